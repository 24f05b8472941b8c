// ORACLE test infrastructure: forwards the OpenCV header name the reference includes to the shim.
#include "cvshim.hpp"
