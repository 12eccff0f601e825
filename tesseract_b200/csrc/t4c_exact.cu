/*
 * t4c_exact.cu -- the FP64 exact narrowphase kernel, compiled as its own translation unit with -fmad=false.
 *
 * The exact pass follows the reference's FP64 detector step by step (t4_bullet_detector.h); its termination tests
 * compare quantities that differ by rounding noise, so the iteration only tracks the reference (which is built
 * without FMA contraction on x86-64) when a*b+c is evaluated with two roundings.  Every other kernel keeps FMA.
 */
#include <cuda_runtime.h>

#include "t4_kernels2.h"

using namespace t4;

__global__ void __launch_bounds__(128) gjkKernel(KParams P, int64_t n)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    gjkBody(i, P);
}

__global__ void __launch_bounds__(128) epaKernel(KParams P, int64_t n)
{
  Epa<double> epa;  // polytope pools live in thread-local memory
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    epaBody(i, P, epa);
}

// FK feeds the exact pass: same two-rounding arithmetic so that link poses equal the reference's FP64 FK
__global__ void __launch_bounds__(128) fkKernel(KParams P, int64_t n)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    fkBody(i, P);
}

void t4LaunchFk(const KParams& P, int64_t n, int grid, cudaStream_t st)
{
  fkKernel<<<grid, 128, 0, st>>>(P, n);
}

void t4LaunchGjk(const KParams& P, int64_t n, int grid, cudaStream_t st)
{
  gjkKernel<<<grid, 128, 0, st>>>(P, n);
}

void t4LaunchEpa(const KParams& P, int64_t n, int grid, cudaStream_t st)
{
  epaKernel<<<grid, 128, 0, st>>>(P, n);
}
