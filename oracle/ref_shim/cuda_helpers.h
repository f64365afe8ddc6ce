// Serial host stand-in for the reference's per-op cuda_helpers.h (wsl/layers/csrc/*/cuda_helpers.h:3-5):
// the grid-stride loop becomes a plain for over all output elements.  TEST INFRASTRUCTURE ONLY.
#pragma once
#include "ref_host_cuda.h"
#define CUDA_1D_KERNEL_LOOP(i, n) for (int i = 0; i < (n); ++i)
