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

// 128-register budget (four 128-thread CTAs per SM): measured faster on B200 than the unconstrained build (r1d); tighter
// budgets (96 / 80 registers) made no difference (r2a sweep)
#ifndef T4_GJK_MIN_CTAS
#define T4_GJK_MIN_CTAS 4
#endif
__global__ void __launch_bounds__(128, T4_GJK_MIN_CTAS) gjkKernelR128(KParams P, int64_t n)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    gjkBody(i, P);
}

// One EPA task per thread, polytope pools in thread-local memory (t4_epa.h).  A group-cooperative version with the pools
// in shared memory was built and measured 3.7x SLOWER (experiments/t4_epa_coop.h): the scalar part
// of a task -- seed GJK, support mapping, phase 4 -- is most of its instructions, a warp instruction costs the same
// issue slot for one active lane as for 32, and shared memory holds 64 tasks per SM where this kernel keeps 512 in
// flight to hide the FP64 dependency latency.
__global__ void __launch_bounds__(128, T4_EPA_MIN_CTAS) epaKernelR128(KParams P, int64_t n)
{
  Epa<double> epa;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    epaBody(i, P, epa);
}

// FK feeds the exact pass: same two-rounding arithmetic (and the host libm's sine / cosine, t4_sincos.h) so that link
// poses equal the reference's FP64 FK bit for bit.  States [first, first + n) of the chunk (the host-buffer entries run
// FK piecewise under the upload of q).
// Stores: the pose arrays are link-major ([alink][state][12]), so the 32 states of a warp own one contiguous run of
// 32 x 96 B (FP64) and 32 x 48 B (FP32) per link.  A thread writing its own record scatters 8- / 16-byte pieces over
// 32 sectors per instruction and the L2 had to merge them: ncu (r2e) counted 4.4 GB of DRAM traffic for 1.06 GB of
// poses.  Here a warp stages the 32 records in shared memory and writes the run as full 128-bit lanes, line by line.
template <int MAXJ>
__global__ void __launch_bounds__(128) fkKernelT(KParams P, int64_t first, int64_t n)
{
  constexpr int STRIDE = 13;  // doubles per staged record (12 + 1: conflict-free column access)
  __shared__ double stage[4][32 * STRIDE];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double* sm = stage[warp];
  for (int64_t base = blockIdx.x * (int64_t)blockDim.x; base < n; base += (int64_t)gridDim.x * blockDim.x)
  {
    const int64_t i = base + threadIdx.x;
    const bool valid = i < n;
    const int64_t s = first + i;
    if (P.fk_out)
    {
      if (valid)
      {
        Xf<double> Tj[MAXJ];
        fkOne(P, P.q0, s, Tj);
        fkStoreAllLinks(P, Tj, s);
      }
      continue;
    }
    const int64_t s0 = first + base + warp * 32;                     // first state of this warp
    const int cnt = (int)t4min<int64_t>(32, n - (base + warp * 32));  // states of this warp inside the range
    // the warp writes the 32 poses of one link as one contiguous run of full 128-bit lanes
    auto writeLink = [&](int a, const Xf<double>& x, float* dst, double* dstd) {
      if (valid)
      {
        double* r = sm + lane * STRIDE;
        r[0] = x.r[0], r[1] = x.r[1], r[2] = x.r[2], r[3] = x.p.x;
        r[4] = x.r[3], r[5] = x.r[4], r[6] = x.r[5], r[7] = x.p.y;
        r[8] = x.r[6], r[9] = x.r[7], r[10] = x.r[8], r[11] = x.p.z;
      }
      __syncwarp();
      if (cnt > 0)
      {
        const size_t off = poseOffset(P, s0, a);  // 96 B (48 B) per state: 16-byte aligned for any s0
        double2* o64 = reinterpret_cast<double2*>(dstd + off);
        for (int e = lane; e < cnt * 6; e += 32)  // 6 x 128 bit per FP64 record
        {
          const int st = e / 6, k = (e - st * 6) * 2;
          o64[e] = make_double2(sm[st * STRIDE + k], sm[st * STRIDE + k + 1]);
        }
        float4* o32 = reinterpret_cast<float4*>(dst + off);
        for (int e = lane; e < cnt * 3; e += 32)  // 3 x 128 bit per FP32 record
        {
          const int st = e / 3, k = (e - st * 3) * 4;
          const double* r = sm + st * STRIDE + k;
          o32[e] = make_float4((float)r[0], (float)r[1], (float)r[2], (float)r[3]);
        }
      }
      __syncwarp();
    };
    for (int pass = 0; pass < (P.cast ? 2 : 1); ++pass)
    {
      const void* q = pass ? P.q1 : P.q0;
      float* dst = pass ? P.poses1 : P.poses0;
      double* dstd = pass ? P.poses1d : P.poses0d;
      // Joint frames are consumed as they are produced: the frame of joint j is the pose of the link it drives and the
      // parent of joint j + 1.  Only frames that a LATER joint hangs on (tree branches, DJoint::keep) go to the
      // thread-local array -- for a serial arm nothing does.  Round 1 kept all MAXJ frames in local memory: 1.5 KB per
      // state written and read back through L2, three times the traffic of the poses themselves (ncu r3e: 0.85 GB of
      // DRAM writes per 256 Ki states for 0.26 GB of poses).
      Xf<double> kept[MAXJ];
      Xf<double> prev;
      for (int j = 0; j < P.n_joints; ++j)  // uniform across the warp: the joint table is the same for every state
      {
        const DJoint& J = P.joints[j];
        Xf<double> cur;
        if (valid)
        {
          const Xf<double> local = fkLocal(P, J, q, s);
          const Xf<double> parent = J.parent_joint < 0 ? Xf<double>::from12(J.pre) :
                                                         (J.parent_joint == j - 1 ? prev : kept[J.parent_joint]);
          cur = parent * local;
          if (J.keep)
            kept[j] = cur;
          prev = cur;
        }
        if (J.alink >= 0)
          writeLink(J.alink, cur, dst, dstd);
      }
      // kinematic links that no joint of the model drives keep their scene pose
      for (int a = 0; a < P.n_alinks; ++a)
        if (P.alink_joint[a] < 0)
          writeLink(a, Xf<double>::from12(P.alink_static + (size_t)a * 12), dst, dstd);
    }
  }
}

void t4LaunchFk(const KParams& P, int64_t first, int64_t n, int grid, cudaStream_t st)
{
  // robots with at most 16 joints (iiwa: 9, IRB2400: 8) keep the joint frames in a 1.5 KB frame instead of 6 KB
  if (P.n_joints <= 16)
    fkKernelT<16><<<grid, 128, 0, st>>>(P, first, n);
  else
    fkKernelT<T4_MAX_JOINTS><<<grid, 128, 0, st>>>(P, first, n);
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
  gjkKernelR128<<<grid, block, 0, st>>>(P, n);
}

void t4LaunchEpa(const KParams& P, int64_t n, int grid, int block, cudaStream_t st)
{
  epaKernelR128<<<grid, block, 0, st>>>(P, n);
}
