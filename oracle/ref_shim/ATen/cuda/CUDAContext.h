// Empty stand-in (test infrastructure): the reference's device kernels are compiled as host C++ by
// oracle/Makefile.ref; only the kernel bodies are used, none of the ATen launcher code. See ref_host_cuda.h.
#pragma once
#include "ref_host_cuda.h"
