// Util.hpp — same include name as the reference tree; everything lives in PathTracing.hpp.
#pragma once
#include "PathTracing.hpp"
