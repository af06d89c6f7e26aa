// Couleur.h -- drop-in header name; see src/rtb_colour.hpp
#pragma once
#include "../src/rtb_colour.hpp"
