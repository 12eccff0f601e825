/*
 * t4c_exact.cu -- the FP64 exact narrowphase kernel, compiled as its own translation unit with -fmad=false.
 *
 * The exact pass follows the reference's FP64 detector step by step (t4_bullet_detector.h); its termination tests
 * compare quantities that differ by rounding noise, so the iteration only tracks the reference (which is built
 * without FMA contraction on x86-64) when a*b+c is evaluated with two roundings.  Every other kernel keeps FMA.
 */
#include <cuda_runtime.h>

#include <cstdlib>

#include "t4_kernels2.h"

using namespace t4;

__global__ void __launch_bounds__(128, 2) gjkKernel(KParams P, int64_t n)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    gjkBody(i, P);
}
// same body under a 128-register budget (twice the resident warps): the default (measured faster on B200, r1d);
// T4C_EXACT_REGS=255 selects the unconstrained build
__global__ void __launch_bounds__(128, 4) gjkKernelR128(KParams P, int64_t n)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    gjkBody(i, P);
}

__global__ void __launch_bounds__(128, 2) epaKernel(KParams P, int64_t n)
{
  Epa<double> epa;  // polytope pools live in thread-local memory
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    epaBody(i, P, epa);
}
__global__ void __launch_bounds__(128, 4) epaKernelR128(KParams P, int64_t n)
{
  Epa<double> epa;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    epaBody(i, P, epa);
}

// experiment knobs: tighter register budgets (T4C_GJK_MINB / T4C_EPA_MINB = 5 -> 96 registers, 6 -> 80 registers)
template <int MINB>
__global__ void __launch_bounds__(128, MINB) gjkKernelB(KParams P, int64_t n)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    gjkBody(i, P);
}
template <int MINB>
__global__ void __launch_bounds__(128, MINB) epaKernelB(KParams P, int64_t n)
{
  Epa<double> epa;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    epaBody(i, P, epa);
}
static int envMinB(const char* name)
{
  const char* e = std::getenv(name);
  const int v = e ? std::atoi(e) : 4;
  return (v == 5 || v == 6) ? v : 4;
}

// lane-refill GJK (t4_kernels2.h gjkLaneLoop).  Measured 2x SLOWER than the pair-per-thread kernel on B200 (r1g/r1h:
// the per-lane state machine lives in local memory and every trip reloads it), so it is opt-in: T4C_GJK_LANES=1.
__global__ void __launch_bounds__(128, 4) gjkLaneKernel(KParams P, int64_t n)
{
  gjkLaneLoop(P, n);
}

// lane-refill EPA (t4_kernels2.h epaLaneLoop).  Measured 35-50 % SLOWER than the task-per-thread kernel on B200 (r1m:
// the kernel is bound by local-memory latency, not by idle lanes), so it is opt-in: T4C_EPA_LANES=1.
__global__ void __launch_bounds__(128, 4) epaLaneKernel(KParams P, int64_t n)
{
  Epa<double> epa;
  epaLaneLoop(P, n, epa);
}

static int epaLanes()
{
  static int v = -1;
  if (v < 0)
  {
    const char* e = std::getenv("T4C_EPA_LANES");
    v = (e && std::atoi(e) == 1) ? 1 : 0;
  }
  return v;
}

static int gjkLanes()
{
  static int v = -1;
  if (v < 0)
  {
    const char* e = std::getenv("T4C_GJK_LANES");
    v = (e && std::atoi(e) == 1) ? 1 : 0;
  }
  return v;
}

static int exactRegs()
{
  static int v = 0;
  if (!v)
  {
    const char* e = std::getenv("T4C_EXACT_REGS");
    v = (e && std::atoi(e) == 255) ? 255 : 128;
  }
  return v;
}

// FK feeds the exact pass: same two-rounding arithmetic so that link poses equal the reference's FP64 FK
// states [first, first + n) of the chunk (the host-buffer entries run FK piecewise under the upload of q)
__global__ void __launch_bounds__(128) fkKernel(KParams P, int64_t first, int64_t n)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    fkBody(first + i, P);
}

// robots with at most 16 joints (iiwa: 9, IRB2400: 8) keep the joint frames in a 1.5 KB frame instead of 6 KB
__global__ void __launch_bounds__(128) fkKernel16(KParams P, int64_t first, int64_t n)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    fkBodyT<16>(first + i, P);
}

void t4LaunchFk(const KParams& P, int64_t first, int64_t n, int grid, cudaStream_t st)
{
  if (P.n_joints <= 16)
    fkKernel16<<<grid, 128, 0, st>>>(P, first, n);
  else
    fkKernel<<<grid, 128, 0, st>>>(P, first, n);
}

// LVS subdivision: FP64, two roundings per a*b+c like the host formula it restates (Eigen LinSpaced)
__global__ void __launch_bounds__(256) lvsCountKernel(LvsParams L, int64_t n)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    lvsCountBody(i, L);
}
__global__ void __launch_bounds__(256) lvsExpandKernel(LvsParams L, int64_t n)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    lvsExpandBody(i, L);
}
void t4LaunchLvsCount(const LvsParams& L, int64_t n, int grid, cudaStream_t st) { lvsCountKernel<<<grid, 256, 0, st>>>(L, n); }
void t4LaunchLvsExpand(const LvsParams& L, int64_t n, int grid, cudaStream_t st) { lvsExpandKernel<<<grid, 256, 0, st>>>(L, n); }

void t4LaunchGjk(const KParams& P, int64_t n, int grid, int block, cudaStream_t st)
{
  static const int minb = envMinB("T4C_GJK_MINB");
  if (gjkLanes())
    gjkLaneKernel<<<grid, 128, 0, st>>>(P, n);
  else if (minb == 5)
    gjkKernelB<5><<<grid, block, 0, st>>>(P, n);
  else if (minb == 6)
    gjkKernelB<6><<<grid, block, 0, st>>>(P, n);
  else if (exactRegs() == 128)
    gjkKernelR128<<<grid, block, 0, st>>>(P, n);
  else
    gjkKernel<<<grid, block, 0, st>>>(P, n);
}

void t4LaunchEpa(const KParams& P, int64_t n, int grid, int block, cudaStream_t st)
{
  static const int minb = envMinB("T4C_EPA_MINB");
  if (epaLanes())
    epaLaneKernel<<<grid, 128, 0, st>>>(P, n);
  else if (minb == 5)
    epaKernelB<5><<<grid, block, 0, st>>>(P, n);
  else if (minb == 6)
    epaKernelB<6><<<grid, block, 0, st>>>(P, n);
  else if (exactRegs() == 128)
    epaKernelR128<<<grid, block, 0, st>>>(P, n);
  else
    epaKernel<<<grid, block, 0, st>>>(P, n);
}
