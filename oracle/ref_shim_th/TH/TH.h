// oracle/ref_shim_th/TH/TH.h -- the legacy <TH/TH.h> the reference's ROIMerge_cpu.cpp / ROILabel_cpu.cpp include is gone from
// this image's PyTorch; neither file uses anything from it.  TEST INFRASTRUCTURE ONLY (oracle/Makefile.ref).
#pragma once
#include <set>
