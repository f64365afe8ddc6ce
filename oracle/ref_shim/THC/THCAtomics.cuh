// Empty stand-in (test infrastructure); atomicAdd for the serial host build lives in ref_host_cuda.h.
#pragma once
#include "ref_host_cuda.h"
