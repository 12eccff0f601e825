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
__global__ void __launch_bounds__(128, 4) gjkKernelR128(KParams P, int64_t n)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    gjkBody(i, P);
}

// ---- EPA: groups of G lanes per task, polytope pools in shared memory (t4_epa.h).  Persistent grid: a group takes the
//      next task from a device counter as soon as it is done with the last one.
//      tier 1 (FORWARD): small pools, small groups -- flat-faced pairs (median 7 iterations, ~20 faces);
//      tier 2: Bullet's full caps (128 vertices / 256 faces), one warp per task -- curved cores and whatever overflowed
//      tier 1, read from the retry queue (or every task directly when there is no retry queue).
#ifndef T4_EPA_SMALL_G
#define T4_EPA_SMALL_G 4
#endif
#define T4_EPA_SMALL_V 20
#define T4_EPA_SMALL_F 36
#define T4_EPA_SMALL_THREADS 64
#define T4_EPA_BIG_THREADS 128

template <int G, int MAXV, int MAXF, bool FORWARD, int THREADS>
__global__ void __launch_bounds__(THREADS) epaCoopKernel(KParams P, int64_t n)
{
  extern __shared__ __align__(16) unsigned char epa_smem[];
  typedef EpaPool<MAXV, MAXF> Pool;
  constexpr int GROUPS = THREADS / G;
  const int grp = threadIdx.x / G;
  Pool& S = reinterpret_cast<Pool*>(epa_smem)[grp];
  V3<double>* Pw = reinterpret_cast<V3<double>*>(P.epa_scratch) + ((size_t)blockIdx.x * GROUPS + grp) * MAXV;
  Coop<G> g;
  const bool from_retry = !FORWARD && P.epa_retry != nullptr;
  const uint32_t count = from_retry ? P.qcount[Q_RETRY] : (uint32_t)n;
  for (;;)
  {
    uint32_t i = 0;
    if (g.lane == 0)
      i = atomicAdd(&P.qcount[FORWARD ? Q_FETCH : Q_FETCH_EPA], 1u);
    i = g.bcast(i);
    if (i >= count)
      break;
    epaTaskBody<G, MAXV, MAXF, FORWARD>(from_retry ? P.epa_retry[i] : i, P, S, Pw);
  }
}

typedef void (*EpaKernelFn)(KParams, int64_t);
struct EpaLaunchCfg
{
  EpaKernelFn fn;
  int threads, groups, maxv, blocks_per_sm;
  size_t smem;
};
static EpaLaunchCfg epaCfg(bool small)
{
  static EpaLaunchCfg cfg[2];
  static bool ready = false;
  if (!ready)
  {
    cfg[0].fn = epaCoopKernel<T4_EPA_SMALL_G, T4_EPA_SMALL_V, T4_EPA_SMALL_F, true, T4_EPA_SMALL_THREADS>;
    cfg[0].threads = T4_EPA_SMALL_THREADS, cfg[0].groups = T4_EPA_SMALL_THREADS / T4_EPA_SMALL_G;
    cfg[0].maxv = T4_EPA_SMALL_V;
    cfg[0].smem = cfg[0].groups * sizeof(EpaPool<T4_EPA_SMALL_V, T4_EPA_SMALL_F>);
    cfg[1].fn = epaCoopKernel<32, 128, 256, false, T4_EPA_BIG_THREADS>;
    cfg[1].threads = T4_EPA_BIG_THREADS, cfg[1].groups = T4_EPA_BIG_THREADS / 32, cfg[1].maxv = 128;
    cfg[1].smem = cfg[1].groups * sizeof(EpaPool<128, 256>);
    for (auto& c : cfg)
    {
      cudaFuncSetAttribute(c.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c.smem);
      int nb = 0;
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, c.fn, c.threads, c.smem) != cudaSuccess || nb < 1)
        nb = 1;
      c.blocks_per_sm = nb;
    }
    ready = true;
  }
  return cfg[small ? 0 : 1];
}

// bytes of witness scratch the two tiers need (tier 1 first, tier 2 behind it)
size_t t4EpaScratchBytes(int sm_count)
{
  const EpaLaunchCfg a = epaCfg(true), b = epaCfg(false);
  return ((size_t)sm_count * a.blocks_per_sm * a.groups * a.maxv + (size_t)sm_count * b.blocks_per_sm * b.groups * b.maxv) *
         sizeof(V3<double>);
}

// P.qcount[Q_FETCH], [Q_FETCH_EPA] and [Q_RETRY] must be zero; P.epa_retry holds n entries (null: tier 2 only);
// P.epa_scratch holds t4EpaScratchBytes().  Returns the number of launches.
int t4LaunchEpa(const KParams& P, int64_t n, int sm_count, cudaStream_t st)
{
  const EpaLaunchCfg a = epaCfg(true), b = epaCfg(false);
  KParams Q = P;
  int launches = 0;
  if (P.epa_retry)
  {
    const int64_t want = (n + a.groups - 1) / a.groups;
    const int grid = (int)t4max<int64_t>(1, t4min<int64_t>(want, (int64_t)sm_count * a.blocks_per_sm));
    a.fn<<<grid, a.threads, a.smem, st>>>(Q, n);
    ++launches;
  }
  Q.epa_scratch = P.epa_scratch + (size_t)sm_count * a.blocks_per_sm * a.groups * a.maxv * 3;
  // tier 2 does not know its task count on the host (it is the retry counter): a full persistent grid unless the
  // whole launch is smaller
  const int64_t want = (n + b.groups - 1) / b.groups;
  const int grid = (int)t4max<int64_t>(1, t4min<int64_t>(want, (int64_t)sm_count * b.blocks_per_sm));
  b.fn<<<grid, b.threads, b.smem, st>>>(Q, n);
  return launches + 1;
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
  Xf<double> Tj[MAXJ];
  for (int64_t base = blockIdx.x * (int64_t)blockDim.x; base < n; base += (int64_t)gridDim.x * blockDim.x)
  {
    const int64_t i = base + threadIdx.x;
    const bool valid = i < n;
    const int64_t s = first + i;
    if (P.fk_out)
    {
      if (valid)
      {
        fkOne(P, P.q0, s, Tj);
        fkStoreAllLinks(P, Tj, s);
      }
      continue;
    }
    const int64_t s0 = first + base + warp * 32;                     // first state of this warp
    const int cnt = (int)t4min<int64_t>(32, n - (base + warp * 32));  // states of this warp inside the range
    for (int pass = 0; pass < (P.cast ? 2 : 1); ++pass)
    {
      if (valid)
        fkOne(P, pass ? P.q1 : P.q0, s, Tj);
      float* dst = pass ? P.poses1 : P.poses0;
      double* dstd = pass ? P.poses1d : P.poses0d;
      for (int a = 0; a < P.n_alinks; ++a)
      {
        if (valid)
        {
          const Xf<double> x = fkAlinkPose(P, Tj, a);
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
      }
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

