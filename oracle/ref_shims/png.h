/* Empty stand-in for libpng's <png.h>: the reference includes it (src/neuralUtils/image.hh:9) but uses
 * nothing from it (its PNG codec is the vendored lodepng).  Test infrastructure only (oracle/_ref build). */
