// Name kept so that `#include "hip_helper.hpp"` in reference-era callers still resolves.
#pragma once
#include "tha_helper.hpp"
