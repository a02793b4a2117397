// Test infrastructure (oracle/): empty stand-in for <cuda_runtime.h> (see oracle/build_ref.sh).
#pragma once
