// ISurface.h -- drop-in header name of the reference; the classes live in src/rtb_surfaces.hpp
#pragma once
#include "../src/rtb_surfaces.hpp"
