// HighFive-subset shim: the reference includes this header but uses nothing from it
// (src/neuralNetwork.cpp:4, src/simpleInfer.cpp:2).
#pragma once
#include "H5File.hpp"
