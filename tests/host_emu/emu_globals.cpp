// TEST INFRASTRUCTURE ONLY: storage for the emulated CUDA built-in variables.
#include "cuda_runtime.h"
thread_local uint3 threadIdx = {0, 0, 0}, blockIdx = {0, 0, 0};
thread_local dim3 blockDim, gridDim;

// Positive control for tests/test_emu_order_independence.py: a deliberately RACY launch -- every thread reads its right
// neighbour's cell of the array the launch is writing (x[i] = x[i + 1] + 1, in place), the pattern of the reference's
// OUTFLOW rows.  Ascending order reads old values, descending order reads new ones.  Single block: no OpenMP involved.
extern "C" void jrl_emu_racy_probe(float *x, int n)
{
    jrl_emu::launch(dim3(1, 1, 1), dim3((unsigned)n, 1, 1), [=]() {
        const int i = (int)threadIdx.x;
        if (i + 1 < n)
            x[i] = x[i + 1] + 1.0f;
    });
}
