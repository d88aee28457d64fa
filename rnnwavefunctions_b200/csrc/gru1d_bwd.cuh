// Reverse sweep of the small 1D models (placeholder header: the kernel is added in the next step).
#pragma once
#include "gru1d.cuh"
