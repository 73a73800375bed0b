// Scan: shared device structures (patterns as kernel parameters, exact verify) live in
// dp_device.cuh, which is also the header of the run-time compiled kernels (dp_jit.cu).
#pragma once
#include "dp_internal.cuh"
