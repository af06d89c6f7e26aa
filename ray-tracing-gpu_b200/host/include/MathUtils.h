// MathUtils.h -- drop-in header name of the reference; the implementation lives in src/rtb_math.hpp
#pragma once
#include "../src/rtb_math.hpp"
