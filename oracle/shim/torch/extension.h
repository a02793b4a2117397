// Test infrastructure (oracle/): stand-in for <torch/extension.h> so the reference's CUDA
// kernel bodies can be compiled as host C++ (see oracle/build_ref.sh). Not product code.
#pragma once
#include "../cpu_cuda_shim.h"
