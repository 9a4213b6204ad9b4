// Drop-in for the reference's src/neuralUtils/shape.hh:3-7 (struct Shape{size_t x,y;}).
#pragma once
#include <cstddef>

struct Shape {
    size_t x, y;
    Shape(size_t x = 1, size_t y = 1) : x(x), y(y) {}
};
