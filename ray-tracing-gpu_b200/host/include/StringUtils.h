// StringUtils.h -- drop-in header name of the reference; the classes live in src/rtb_scene.hpp
#pragma once
#include "../src/rtb_scene.hpp"
