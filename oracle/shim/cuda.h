// Test infrastructure (oracle/): empty stand-in for <cuda.h> (see oracle/build_ref.sh).
#pragma once
