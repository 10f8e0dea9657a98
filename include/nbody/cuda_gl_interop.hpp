// cuda_gl_interop.hpp -- forwarding header: code written against the reference's include/nbody/cuda_gl_interop.hpp
// compiles unchanged; everything lives in nbody.hpp (thin C++20 wrappers over include/nbody_b200.h).
#pragma once
#include "nbody.hpp"
