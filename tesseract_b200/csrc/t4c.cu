/*
 * t4c.cu -- the C ABI of include/t4c.h: scene objects, kernel launches, batched discrete / cast checks.
 *
 * Built by nvcc for sm_100a into libt4c.so (the product; no CPU execution path: every compute entry returns
 * T4C_ERR_CUDA when no device is usable).  The same source also compiles with g++ -DT4C_HOSTSIM into
 * tests/hostsim/libt4c_hostsim.so, a TEST-ONLY simulator that runs the identical kernel bodies serially on the host
 * so numerics can be debugged without a GPU; it is never loaded by the tesseract_b200 package.
 */
#include <mutex>
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "t4_host_scene.h"
#include "t4_jacobian.h"
#include "t4_kernels2.h"

#ifndef T4C_HOSTSIM
#include <cuda_runtime.h>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#endif

using namespace t4;

// ===================================================================================================================
// platform layer
// ===================================================================================================================
static thread_local std::string g_err;
static thread_local bool g_alloc_failed = false;  // a device allocation failed since the flag was last cleared
static std::atomic<unsigned long long> g_launches{ 0 };

static t4c_status fail(t4c_status st, const char* fmt, ...)
{
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
  return st;
}

#ifdef T4C_HOSTSIM
typedef void* t4stream;
namespace plat
{
static bool ok() { return true; }
static int deviceCount() { return 1; }
static bool setDevice(int) { return true; }
static void* alloc(size_t n) { return std::malloc(n ? n : 1); }
static void release(void* p) { std::free(p); }
static bool h2d(void* d, const void* s, size_t n, t4stream) { std::memcpy(d, s, n); return true; }
static bool d2h(void* d, const void* s, size_t n, t4stream) { std::memcpy(d, s, n); return true; }
static bool d2d(void* d, const void* s, size_t n, t4stream) { std::memcpy(d, s, n); return true; }
static bool zero(void* d, size_t n, t4stream) { std::memset(d, 0, n); return true; }
static bool fill(void* d, int v, size_t n, t4stream) { std::memset(d, v, n); return true; }
static bool sync(t4stream) { return true; }
static t4stream newStream() { return nullptr; }
static bool isPinned(const void*) { return true; }  // exercises the progressive-copy logic in the host simulator
typedef void* t4event;
static t4event newEvent() { return nullptr; }
static void freeEvent(t4event) {}
static bool record(t4event, t4stream) { return true; }
static bool wait(t4stream, t4event) { return true; }
static void freeStream(t4stream) {}
static const char* lastError() { return "hostsim"; }
}  // namespace plat
#else
typedef cudaStream_t t4stream;
namespace plat
{
static cudaError_t g_last = cudaSuccess;
static bool chk(cudaError_t e)
{
  if (e != cudaSuccess)
  {
    g_last = e;
    return false;
  }
  return true;
}
static int deviceCount()
{
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess)
    return 0;
  return n;
}
static bool setDevice(int d) { return chk(cudaSetDevice(d)); }
static void* alloc(size_t n)
{
  void* p = nullptr;
  if (!chk(cudaMalloc(&p, n ? n : 1)))
  {
    g_alloc_failed = true;
    cudaGetLastError();  // out of memory is not sticky: clear it so that the next call starts clean
    return nullptr;
  }
  return p;
}
static void release(void* p) { cudaFree(p); }
static bool h2d(void* d, const void* s, size_t n, t4stream st)
{
  return n == 0 || chk(cudaMemcpyAsync(d, s, n, cudaMemcpyHostToDevice, st));
}
static bool d2h(void* d, const void* s, size_t n, t4stream st)
{
  return n == 0 || chk(cudaMemcpyAsync(d, s, n, cudaMemcpyDeviceToHost, st));
}
// device to device, also into a peer's memory (an IPC-mapped or peer-enabled pointer: the copy goes over NVLink)
static bool d2d(void* d, const void* s, size_t n, t4stream st)
{
  return n == 0 || chk(cudaMemcpyAsync(d, s, n, cudaMemcpyDeviceToDevice, st));
}
static bool zero(void* d, size_t n, t4stream st) { return n == 0 || chk(cudaMemsetAsync(d, 0, n, st)); }
static bool fill(void* d, int v, size_t n, t4stream st) { return n == 0 || chk(cudaMemsetAsync(d, v, n, st)); }
static bool sync(t4stream st) { return chk(cudaStreamSynchronize(st)); }
static t4stream newStream()
{
  cudaStream_t s = nullptr;
  if (!chk(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking)))
    return nullptr;
  return s;
}
static void freeStream(t4stream s)
{
  if (s)
    cudaStreamDestroy(s);
}
typedef cudaEvent_t t4event;
static t4event newEvent()
{
  cudaEvent_t e = nullptr;
  if (!chk(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)))
    return nullptr;
  return e;
}
static void freeEvent(t4event e)
{
  if (e)
    cudaEventDestroy(e);
}
static bool record(t4event e, t4stream st) { return chk(cudaEventRecord(e, st)); }
static bool wait(t4stream st, t4event e) { return chk(cudaStreamWaitEvent(st, e, 0)); }
// page-locked host memory: the only kind an asynchronous copy does not stage through the launching thread
static bool isPinned(const void* p)
{
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess)
  {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeHost;
}
static const char* lastError() { return cudaGetErrorString(g_last); }
}  // namespace plat
#endif

// growable device buffer
struct DevBuf
{
  void* p = nullptr;
  size_t cap = 0;
  ~DevBuf() { reset(); }
  void reset()
  {
    if (p)
      plat::release(p);
    p = nullptr;
    cap = 0;
  }
  bool ensure(size_t bytes)
  {
    if (bytes <= cap)
      return true;
    reset();
    size_t want = bytes + bytes / 8 + 256;
    p = plat::alloc(want);
    if (!p)
      return false;
    cap = want;
    return true;
  }
  // grow to at least `bytes`, preserving the first `keep` bytes (queues that are appended to across launches)
  bool ensureKeep(size_t bytes, size_t keep, t4stream st);
  template <class T>
  T* as() const
  {
    return reinterpret_cast<T*>(p);
  }
};


bool DevBuf::ensureKeep(size_t bytes, size_t keep, t4stream st)
{
  if (bytes <= cap)
    return true;
  const size_t want = bytes + bytes / 8 + 256;
  void* q = plat::alloc(want);
  if (!q)
    return false;
  bool ok = true;
  if (p && keep)
  {
#ifdef T4C_HOSTSIM
    std::memcpy(q, p, std::min(keep, cap));
#else
    ok = plat::chk(cudaMemcpyAsync(q, p, std::min(keep, cap), cudaMemcpyDeviceToDevice, st)) && plat::sync(st);
#endif
  }
  (void)st;
  if (p)
    plat::release(p);
  p = q;
  cap = want;
  return ok;
}

template <class T>
static bool upload(DevBuf& b, const std::vector<T>& v, t4stream st)
{
  if (!b.ensure(std::max<size_t>(v.size(), 1) * sizeof(T)))
    return false;
  return plat::h2d(b.p, v.data(), v.size() * sizeof(T), st);
}

// ===================================================================================================================
// kernels (sm_100a) / serial loops (host simulator)
// ===================================================================================================================
#ifndef T4C_HOSTSIM
#define T4_SM_COUNT_FALLBACK 148
static int g_sm_count = 0;
static int smCount()
{
  if (!g_sm_count)
  {
    int dev = 0, n = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = T4_SM_COUNT_FALLBACK;
    g_sm_count = n;
  }
  return g_sm_count;
}
// Grid policy of the grid-stride kernels: one CTA per `block` items, so that the hardware hands the queue out as CTAs
// retire; a grid capped at the resident CTAs gives every thread a fixed stride of the queue and the slowest warp sets
// the time (measured r2a: fk -5 %, broad -6 %, cull -15 %, gjk -6 % uncapped; the thread-per-task EPA alone was 2-4 %
// faster capped).  `capped` selects the resident-CTA grid.
static int gridFor(int64_t n, int block, int ctas_per_sm, bool capped = false)
{
  const int64_t need = (n + block - 1) / block;
  if (!capped)
    return (int)std::max<int64_t>(1, std::min<int64_t>(need, 0x7fffffff));
  const int64_t full = (int64_t)smCount() * ctas_per_sm;
  return (int)std::max<int64_t>(1, std::min<int64_t>(need, full));
}

void t4LaunchFk(const KParams& P, int64_t first, int64_t n, int grid, cudaStream_t st);  // t4c_exact.cu (-fmad=false)
__global__ void __launch_bounds__(256) convertPosesKernel(KParams P, int64_t n)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    convertPosesBody(i, P);
}
__global__ void __launch_bounds__(256) broadStaticKernel(KParams P, int64_t n)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    broadStaticBody(i, P);
}
__global__ void __launch_bounds__(256) broadActiveKernel(KParams P, int64_t n)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    broadActiveBody(i, P);
}
#ifndef T4_CULL_MIN_CTAS
#define T4_CULL_MIN_CTAS 4  // 64 registers: measured 9-10 % faster than the unconstrained 128-register build (r1n)
#endif
__global__ void __launch_bounds__(256, T4_CULL_MIN_CTAS) cullKernel(KParams P, int64_t n)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    cullBody(i, P);
}
__global__ void __launch_bounds__(256) liveKeyKernel(KParams P, int64_t n)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    liveKeyBody(i, P);
}
__global__ void __launch_bounds__(256) pruneKernel(KParams P, int64_t n)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    pruneBody(i, P);
}
void t4LaunchGjk(const KParams& P, int64_t n, int grid, int block, cudaStream_t st);  // t4c_exact.cu (-fmad=false)
void t4LaunchLvsCount(const LvsParams& L, int64_t n, int grid, cudaStream_t st);
void t4LaunchLvsExpand(const LvsParams& L, int64_t n, int grid, cudaStream_t st);
void t4LaunchEpa(const KParams& P, int64_t n, int grid, int block, cudaStream_t st);
__global__ void __launch_bounds__(256) closestInsertKernel(KParams P, int64_t n)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    closestInsertBody(i, P);
}
__global__ void __launch_bounds__(256) closestEmitKernel(KParams P, int64_t n)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    closestEmitBody(i, P);
}
__global__ void __launch_bounds__(256) countHitKernel(KParams P, int64_t n)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    statAdd(P.stats, ST_COLLIDING, P.hit[i] ? 1u : 0u);
}
// sort key of a queued pair: (type of shape A, type of shape B).  Sorting the survivor / EPA task queues by it puts
// pairs that take the same branches of the support mapping (and similar iteration counts: EPA needs ~7 iterations
// against boxes and spheres, 15-75 against cylinders) into the same warps.
__global__ void __launch_bounds__(256) typeKeyKernel(KParams P, const uint32_t* queue, int from_tasks, uint16_t* keys,
                                                    uint32_t* idx, int64_t n)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
  {
    const uint32_t c = from_tasks ? P.tasks[i].cand : queue[i];
    // major: type of B (the environment primitive in robot-vs-environment pairs), then type of A, then which shape A is
    // (same hull vertices / support table inside a warp)
    const uint32_t a = P.cand_a[c];
    // minor (survivor queues only): the iteration count of the FP32 cull, a predictor of the exact pass's
    // ... and, above it, whether phase 1 will be skipped (bit 7 of the class byte): a warp of certified pairs runs no
    // boolean GJK at all, a mixed warp pays for it with the certified lanes idle
    const uint32_t cb = (!from_tasks && P.cand_class) ? P.cand_class[c] : 0u;
    const uint32_t cls = ((cb >> 7) << 6) | (cb & 63u);
    keys[i] = (uint16_t)(((((P.shapes[P.cand_b[c]].type & 7) << 5) | ((P.shapes[a].type & 3) << 3) | (a & 7u)) << 7) | cls);
    if (from_tasks)
      idx[i] = (uint32_t)i;
  }
}
__global__ void fmaPeakKernel(float* out, int iters)
{
  float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1.f, a2 = a0 + 2.f, a3 = a0 + 3.f, a4 = a0 + 4.f, a5 = a0 + 5.f,
        a6 = a0 + 6.f, a7 = a0 + 7.f;
  const float b = 1.0000001f, c = 1e-7f;
  for (int i = 0; i < iters; ++i)
  {
    a0 = fmaf(a0, b, c), a1 = fmaf(a1, b, c), a2 = fmaf(a2, b, c), a3 = fmaf(a3, b, c);
    a4 = fmaf(a4, b, c), a5 = fmaf(a5, b, c), a6 = fmaf(a6, b, c), a7 = fmaf(a7, b, c);
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

__global__ void dfmaPeakKernel(double* out, int iters)
{
  double a0 = threadIdx.x * 1e-3, a1 = a0 + 1., a2 = a0 + 2., a3 = a0 + 3., a4 = a0 + 4., a5 = a0 + 5., a6 = a0 + 6.,
         a7 = a0 + 7.;
  const double b = 1.0000001, c = 1e-7;
  for (int i = 0; i < iters; ++i)
  {
    a0 = fma(a0, b, c), a1 = fma(a1, b, c), a2 = fma(a2, b, c), a3 = fma(a3, b, c);
    a4 = fma(a4, b, c), a5 = fma(a5, b, c), a6 = fma(a6, b, c), a7 = fma(a7, b, c);
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}
template <int MAXJ>
__global__ void __launch_bounds__(128) jacobianKernel(JacParams Q, int64_t n)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    jacobianBody<MAXJ>(i, Q);
}
template <int MAXJ>
__global__ void __launch_bounds__(128) contactGradKernel(JacParams Q, int64_t n)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    contactGradBody<MAXJ>(i, Q);
}
__global__ void sincosDebugKernel(const double* x, int64_t n, double* s, double* c)
{
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
  {
    s[i] = t4sin(x[i]);
    c[i] = t4cos(x[i]);
  }
}

// optional per-kernel timing with CUDA events recorded on the launching stream (bench.py roofline figures)
enum
{
  PK_FK = 0,
  PK_CONVERT,
  PK_BROAD_STATIC,
  PK_BROAD_ACTIVE,
  PK_CULL,
  PK_GJK,
  PK_EPA,
  PK_CLOSEST_INSERT,
  PK_CLOSEST_EMIT,
  PK_COUNT_HIT,
  PK_COUNT
};
struct ProfRec
{
  int k;
  cudaEvent_t e0, e1;
};
static bool g_prof_on = false;
static std::vector<ProfRec> g_prof_pending;
static double g_prof_ms[PK_COUNT];
static unsigned long long g_prof_n[PK_COUNT];
static void profBegin(int k, cudaStream_t st)
{
  if (!g_prof_on)
    return;
  ProfRec r;
  r.k = k;
  cudaEventCreate(&r.e0);
  cudaEventCreate(&r.e1);
  cudaEventRecord(r.e0, st);
  g_prof_pending.push_back(r);
}
static void profEnd(cudaStream_t st)
{
  if (g_prof_on && !g_prof_pending.empty())
    cudaEventRecord(g_prof_pending.back().e1, st);
}
static void profResolve()
{
  for (auto& r : g_prof_pending)
  {
    float ms = 0;
    if (cudaEventSynchronize(r.e1) == cudaSuccess && cudaEventElapsedTime(&ms, r.e0, r.e1) == cudaSuccess)
    {
      g_prof_ms[r.k] += ms;
      g_prof_n[r.k] += 1;
    }
    cudaEventDestroy(r.e0);
    cudaEventDestroy(r.e1);
  }
  g_prof_pending.clear();
}

#define T4_LAUNCH(KERNEL, PK, BLOCK, CTAS, P, n, st)                                                                   \
  do                                                                                                                   \
  {                                                                                                                    \
    if ((n) > 0)                                                                                                       \
    {                                                                                                                  \
      profBegin((PK), (st));                                                                                           \
      KERNEL<<<gridFor((n), (BLOCK), (CTAS)), (BLOCK), 0, (st)>>>((P), (n));                                           \
      profEnd((st));                                                                                                   \
      ++g_launches;                                                                                                    \
    }                                                                                                                  \
  } while (0)
static bool launchOk() { return plat::chk(cudaGetLastError()); }
static void launchFk(const KParams& P, int64_t first, int64_t n, t4stream st)
{
  if (n <= 0)
    return;
  profBegin(PK_FK, st);
  t4LaunchFk(P, first, n, gridFor(n, 128, 8), st);
  profEnd(st);
  ++g_launches;
}
static void launchConvert(const KParams& P, int64_t n, t4stream st) { T4_LAUNCH(convertPosesKernel, PK_CONVERT, 256, 8, P, n, st); }
static void launchBroadStatic(const KParams& P, int64_t n, t4stream st) { T4_LAUNCH(broadStaticKernel, PK_BROAD_STATIC, 256, 8, P, n, st); }
static void launchBroadActive(const KParams& P, int64_t n, t4stream st) { T4_LAUNCH(broadActiveKernel, PK_BROAD_ACTIVE, 256, 8, P, n, st); }
static void launchCull(const KParams& P, int64_t n, t4stream st) { T4_LAUNCH(cullKernel, PK_CULL, 256, (T4_CULL_MIN_CTAS > 4 ? T4_CULL_MIN_CTAS : 4), P, n, st); }
static void launchGjk(const KParams& P, int64_t n, t4stream st)
{
  if (n <= 0)
    return;
  profBegin(PK_GJK, st);
  t4LaunchGjk(P, n, gridFor(n, 128, 4), 128, st);
  profEnd(st);
  ++g_launches;
}
static void launchEpa(const KParams& P, int64_t n, t4stream st)
{
  if (n <= 0)
    return;
  profBegin(PK_EPA, st);
  // a launch of a few thousand tasks lasts as long as its slowest task (124 iterations against a cylinder): small
  // launches use small CTAs so that the tasks spread over all SMs and each SM's L1 holds fewer polytope pools
  // (16 k tasks: 1.97 -> 1.47 ms, profiles/r3m).  Keeping the polytopes of such launches in a dense global pool
  // instead of thread-local memory changed nothing (profiles/r3q): the tail is a serial FP64 dependency chain.
  int block = 128;
  if (n <= (int64_t)smCount() * 4 * 32)
    block = 32;
  else if (n <= (int64_t)smCount() * 4 * 64)
    block = 64;
  t4LaunchEpa(P, n, gridFor(n, block, T4_EPA_MIN_CTAS * (128 / block), true), block, st);
  profEnd(st);
  ++g_launches;
}
static void launchPrune(const KParams& P, int64_t n, t4stream st) { T4_LAUNCH(pruneKernel, PK_CULL, 256, 8, P, n, st); }
static void launchLiveKey(const KParams& P, int64_t n, t4stream st) { T4_LAUNCH(liveKeyKernel, PK_BROAD_STATIC, 256, 8, P, n, st); }
static void launchClosestInsert(const KParams& P, int64_t n, t4stream st) { T4_LAUNCH(closestInsertKernel, PK_CLOSEST_INSERT, 256, 8, P, n, st); }
static void launchClosestEmit(const KParams& P, int64_t n, t4stream st) { T4_LAUNCH(closestEmitKernel, PK_CLOSEST_EMIT, 256, 8, P, n, st); }
static void launchCountHit(const KParams& P, int64_t n, t4stream st) { T4_LAUNCH(countHitKernel, PK_COUNT_HIT, 256, 8, P, n, st); }
#else
#define T4_SERIAL(BODY, P, n)                                                                                          \
  do                                                                                                                   \
  {                                                                                                                    \
    for (int64_t _i = 0; _i < (n); ++_i)                                                                               \
      BODY(_i, (P));                                                                                                   \
    if ((n) > 0)                                                                                                       \
      ++g_launches;                                                                                                    \
  } while (0)
static bool launchOk() { return true; }
static void launchFk(const KParams& P, int64_t first, int64_t n, t4stream)
{
  for (int64_t i = 0; i < n; ++i)
    fkBody(first + i, P);
}
static void launchConvert(const KParams& P, int64_t n, t4stream) { T4_SERIAL(convertPosesBody, P, n); }
static void launchBroadStatic(const KParams& P, int64_t n, t4stream) { T4_SERIAL(broadStaticBody, P, n); }
static void launchBroadActive(const KParams& P, int64_t n, t4stream) { T4_SERIAL(broadActiveBody, P, n); }
static void launchCull(const KParams& P, int64_t n, t4stream) { T4_SERIAL(cullBody, P, n); }
static void launchGjk(const KParams& P, int64_t n, t4stream) { T4_SERIAL(gjkBody, P, n); }
static void launchEpa(const KParams& P, int64_t n, t4stream)
{
  std::unique_ptr<Epa<double>> epa(new Epa<double>());
  for (int64_t i = 0; i < n; ++i)
    epaBody(i, P, *epa);
  if (n > 0)
    ++g_launches;
}
static void launchPrune(const KParams& P, int64_t n, t4stream) { T4_SERIAL(pruneBody, P, n); }
static void launchLiveKey(const KParams& P, int64_t n, t4stream) { T4_SERIAL(liveKeyBody, P, n); }
static void launchClosestInsert(const KParams& P, int64_t n, t4stream) { T4_SERIAL(closestInsertBody, P, n); }
static void launchClosestEmit(const KParams& P, int64_t n, t4stream) { T4_SERIAL(closestEmitBody, P, n); }
static void launchCountHit(const KParams& P, int64_t n, t4stream)
{
  for (int64_t i = 0; i < n; ++i)
    statAdd(P.stats, ST_COLLIDING, P.hit[i] ? 1u : 0u);
}
#endif

// ===================================================================================================================
// scene / FK model objects
// ===================================================================================================================
struct t4c_fk_model
{
  int n_links = 0, n_q = 0, root_link = 0;
  std::vector<t4c_joint> joints;  // parent-before-child
};

struct FkCache
{
  unsigned long long scene_rev = ~0ull;
  DevBuf joints, alink_joint, alink_static, link_joint, link_static;
  int n_joints = 0;
};

// Compiled scene + its device copy.  Immutable once built, so the clones of a scene share ONE image by reference count
// (every consumer of an Environment takes a manager clone, tesseract_environment/src/core/environment.cpp:482-488: a clone
// must not cost a compile and an upload); a scene that is modified afterwards builds a fresh image of its own and lets
// go of the shared one.
struct SceneImage
{
  CompiledScene comp;
  DevBuf d_shapes, d_statics, d_verts, d_sup_cells, d_sup_cand, d_bvh, d_active, d_aa, d_acm, d_margin, d_shape_d, d_statics_d;
  DevBuf d_link_bvh, d_link_root, d_link_group;
};

struct t4c_scene
{
  int device = 0;
  HostScene host;
  unsigned long long rev = 1;       // bumped by every mutation
  unsigned long long uploaded = 0;  // rev of the device copy
  std::shared_ptr<SceneImage> img = std::make_shared<SceneImage>();
  t4stream stream = nullptr;
  t4stream copy_stream = nullptr;  // progressive download of finished contact records (hostRun)
  t4stream aux_stream = nullptr;   // long-pair GJK pass, beside the deferred tiers / EPA on the main stream
  t4stream epa_stream = nullptr;   // EPA tasks of the first deferred pass, beside the other deferred passes
  plat::t4event q_events[4] = { nullptr, nullptr, nullptr, nullptr };  // piecewise upload of q (hostRun)
  // per-call workspace
  DevBuf d_q0, d_q1, d_pin0, d_pin1, d_poses0, d_poses1, d_poses0d, d_poses1d;
  DevBuf d_cand_state, d_cand_a, d_cand_b, d_cand_class, d_long, d_tasks_long, d_surv, d_surv2, d_surv3, d_surv4, d_cull_claim, d_tasks;
  DevBuf d_qcount, d_stats, d_hit, d_claim, d_raw, d_raw_local, d_raw_d64, d_ht_rep, d_ht_best;
  DevBuf d_out, d_out_local, d_fk_out, d_hit_full;
  DevBuf d_hit_send, d_shard_out, d_gather_hit, d_gather_out;  // multi-GPU entries (t4_multi.h)
  DevBuf d_traj, d_lvs_counts, d_lvs_offsets, d_lvs_tmp;
  DevBuf d_sort_keys, d_sort_keys2, d_sort_vals, d_sort_idx, d_sort_tmp;
  DevBuf d_key_min, d_cand_lb;
  DevBuf d_epa_order;  // type order of the EPA tasks that start early (private copy: the sort buffers are reused)
  DevBuf d_live_a, d_live_b, d_live_key, d_live_key2, d_live_val;  // FIRST in rounds: ordered state lists
  uint32_t cand_cap = 0;
  int64_t chunk_hint = 0;  // states per chunk that fitted the device memory (0: the default)
  std::unordered_map<const t4c_fk_model*, std::unique_ptr<FkCache>> fk;

  // frees what runBatch allocates per chunk (not the caller-facing staging buffers of the host entries)
  void releaseWorkspace()
  {
    DevBuf* bufs[] = { &d_poses0, &d_poses1, &d_poses0d, &d_poses1d, &d_cand_state, &d_cand_a, &d_cand_b, &d_cand_class,
                       &d_long, &d_tasks_long, &d_surv, &d_surv2, &d_surv3, &d_surv4, &d_cull_claim, &d_tasks, &d_hit,
                       &d_claim, &d_raw, &d_raw_local, &d_raw_d64, &d_ht_rep, &d_ht_best, &d_sort_keys, &d_sort_keys2,
                       &d_sort_vals, &d_sort_idx, &d_sort_tmp, &d_key_min, &d_cand_lb, &d_live_a, &d_live_b,
                       &d_live_key, &d_live_key2, &d_live_val, &d_epa_order };
    for (DevBuf* b : bufs)
      b->reset();
    cand_cap = 0;
  }

  ~t4c_scene()
  {
    fk.clear();
    plat::freeStream(stream);
    plat::freeStream(copy_stream);
    plat::freeStream(aux_stream);
    plat::freeStream(epa_stream);
    for (auto& e : q_events)
      plat::freeEvent(e);
  }
};

static bool bindDevice(const t4c_scene* s)
{
  return plat::setDevice(s->device);
}

// upload of a scene image's device buffers (shared by the full build and the refit)
static bool uploadImage(SceneImage& I, t4stream st)
{
  CompiledScene& C = I.comp;
  return upload(I.d_shapes, C.shapes, st) && upload(I.d_statics, C.statics, st) && upload(I.d_verts, C.verts, st) &&
         upload(I.d_sup_cells, C.sup_cells, st) && upload(I.d_sup_cand, C.sup_cand, st) && upload(I.d_bvh, C.bvh, st) &&
         upload(I.d_active, C.active_shapes, st) && upload(I.d_aa, C.aa_pairs, st) && upload(I.d_acm, C.acm_bits, st) &&
         upload(I.d_margin, C.pair_margin, st) && upload(I.d_shape_d, C.shape_d, st) &&
         upload(I.d_statics_d, C.statics_d, st) && upload(I.d_link_bvh, C.link_bvh, st) &&
         upload(I.d_link_root, C.link_root, st) && upload(I.d_link_group, C.link_group, st) && plat::sync(st);
}

static t4c_status syncScene(t4c_scene* s)
{
  if (!bindDevice(s))
    return fail(T4C_ERR_CUDA, "cudaSetDevice(%d): %s", s->device, plat::lastError());
  if (s->uploaded == s->rev)
    return T4C_OK;
  std::string err;
  // never rebuild in place: other clones may still be reading the old image
  std::shared_ptr<SceneImage> img = std::make_shared<SceneImage>();
  if (!compileScene(s->host, img->comp, err))
    return fail(T4C_ERR_UNSUPPORTED, "%s", err.c_str());
  if (!uploadImage(*img, s->stream))
    return fail(T4C_ERR_CUDA, "scene upload failed: %s", plat::lastError());
  s->img = img;
  s->uploaded = s->rev;
  return T4C_OK;
}

static void fillSceneParams(const t4c_scene* s, KParams& P)
{
  const CompiledScene& C = s->img->comp;
  const SceneImage& I = *s->img;
  std::memset(&P, 0, sizeof(P));
  P.shapes = I.d_shapes.as<DShape>();
  P.statics = I.d_statics.as<DStaticW>();
  P.shape_d = I.d_shape_d.as<double>();
  P.statics_d = I.d_statics_d.as<double>();
  P.verts = I.d_verts.as<f4>();
  P.sup_cells = C.sup_cells.empty() ? nullptr : I.d_sup_cells.as<uint32_t>();
  P.sup_cand = I.d_sup_cand.as<uint16_t>();
  P.bvh = I.d_bvh.as<DBvhNode>();
  P.n_bvh_nodes = (int)C.bvh.size();
  P.active_shapes = I.d_active.as<int32_t>();
  P.n_active_shapes = (int)C.active_shapes.size();
  P.aa_pairs = I.d_aa.as<int32_t>();
  P.n_aa_pairs = (int)(C.aa_pairs.size() / 3);
  P.link_bvh = I.d_link_bvh.as<DBvhNode>();
  P.link_root = I.d_link_root.as<int32_t>();
  P.acm_bits = I.d_acm.as<uint32_t>();
  P.acm_words = C.acm_words;
  P.pair_margin = C.pair_margin.empty() ? nullptr : I.d_margin.as<double>();
  P.n_links = (int)s->host.links.size();
  P.n_alinks = (int)C.alinks.size();
  P.default_margin = C.default_margin;
  P.max_margin = C.max_margin;
  P.default_margin_d = C.default_margin_d;
  P.max_margin_d = C.max_margin_d;
}

// FK tables of `model` against the current scene revision (device arrays cached in the scene)
static t4c_status syncFk(t4c_scene* s, const t4c_fk_model* m, FkCache** out)
{
  auto& slot = s->fk[m];
  if (!slot)
    slot.reset(new FkCache());
  FkCache* fc = slot.get();
  *out = fc;
  if (fc->scene_rev == s->rev)
    return T4C_OK;
  const int nl = (int)s->host.links.size();
  const int nj = (int)m->joints.size();
  if (nj > T4_MAX_JOINTS)
    return fail(T4C_ERR_UNSUPPORTED, "FK model has %d joints (max %d)", nj, T4_MAX_JOINTS);
  std::vector<DJoint> dj((size_t)nj);
  std::vector<int32_t> child_joint((size_t)std::max(nl, m->n_links), -1);
  for (int j = 0; j < nj; ++j)
  {
    const t4c_joint& J = m->joints[j];
    if (J.parent_link < 0 || J.child_link < 0 || J.parent_link >= (int)child_joint.size() ||
        J.child_link >= (int)child_joint.size())
      return fail(T4C_ERR_INVALID, "joint %d refers to an unknown link", j);
    if (J.type != T4C_JOINT_FIXED && (J.q_index < 0 || J.q_index >= m->n_q))
      return fail(T4C_ERR_INVALID, "joint %d has q_index %d outside [0,%d)", j, J.q_index, m->n_q);
    DJoint& D = dj[j];
    std::memset(&D, 0, sizeof(D));
    std::memcpy(D.origin, J.origin, sizeof(D.origin));
    const double al = std::sqrt(J.axis[0] * J.axis[0] + J.axis[1] * J.axis[1] + J.axis[2] * J.axis[2]);
    for (int k = 0; k < 3; ++k)
      D.axis[k] = al > 0 ? J.axis[k] / al : (k == 2 ? 1.0 : 0.0);
    D.type = J.type;
    D.q_index = J.q_index;
    D.parent_joint = child_joint[J.parent_link];
    static const double I12[12] = { 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0 };
    std::memcpy(D.pre, J.parent_link < nl ? s->host.links[J.parent_link].pose : I12, sizeof(D.pre));
    child_joint[J.child_link] = j;
    D.alink = (J.child_link < (int)s->img->comp.link_alink.size()) ? s->img->comp.link_alink[J.child_link] : -1;
    D.keep = 0;
  }
  // a frame has to outlive the next joint only when a LATER joint hangs on it (tree branch): serial chains keep nothing
  for (int j = 0; j < nj; ++j)
    if (dj[j].parent_joint >= 0 && dj[j].parent_joint != j - 1)
      dj[dj[j].parent_joint].keep = 1;
  const CompiledScene& C = s->img->comp;
  std::vector<int32_t> aj(std::max<size_t>(C.alinks.size(), 1), -1);
  std::vector<double> as(std::max<size_t>(C.alinks.size(), 1) * 12, 0.0);
  for (size_t a = 0; a < C.alinks.size(); ++a)
  {
    aj[a] = child_joint[C.alinks[a]];
    std::memcpy(&as[a * 12], s->host.links[C.alinks[a]].pose, 12 * sizeof(double));
  }
  std::vector<int32_t> lj((size_t)std::max(nl, 1), -1);
  std::vector<double> ls((size_t)std::max(nl, 1) * 12, 0.0);
  for (int l = 0; l < nl; ++l)
  {
    lj[l] = child_joint[l];
    std::memcpy(&ls[(size_t)l * 12], s->host.links[l].pose, 12 * sizeof(double));
  }
  t4stream st = s->stream;
  bool ok = upload(fc->joints, dj, st) && upload(fc->alink_joint, aj, st) && upload(fc->alink_static, as, st) &&
            upload(fc->link_joint, lj, st) && upload(fc->link_static, ls, st) && plat::sync(st);
  if (!ok)
    return fail(T4C_ERR_CUDA, "FK model upload failed: %s", plat::lastError());
  fc->n_joints = nj;
  fc->scene_rev = s->rev;
  return T4C_OK;
}

static void fillFkParams(const t4c_fk_model* m, const FkCache* fc, KParams& P)
{
  P.joints = fc->joints.as<DJoint>();
  P.n_joints = fc->n_joints;
  P.n_q = m->n_q;
  P.alink_joint = fc->alink_joint.as<int32_t>();
  P.alink_static = fc->alink_static.as<double>();
  P.link_joint = fc->link_joint.as<int32_t>();
  P.link_static = fc->link_static.as<double>();
}

// ===================================================================================================================
// batched pipeline
// ===================================================================================================================
struct BatchIO
{
  int64_t n = 0;
  bool cast = false;
  // joint-space input (device pointers), or link poses (device, FP64, n x n_alinks x 12) when model == nullptr
  const t4c_fk_model* model = nullptr;
  const void* q0 = nullptr;
  const void* q1 = nullptr;
  int q_layout = T4C_Q_F64_ROWMAJOR;
  const double* poses0 = nullptr;
  const double* poses1 = nullptr;
  t4c_request req{ T4C_TEST_FIRST, 1, 1 };
  int64_t state_offset = 0;  // added to the state index of every record (shard of a larger batch, t4_multi.h)
  // outputs (device pointers, may be null)
  uint8_t* hit = nullptr;
  t4c_contact* contacts = nullptr;
  float* cc_local = nullptr;
  int64_t capacity = 0;
  // host-buffer entries with a page-locked result buffer: records are downloaded on the scene's copy stream as soon
  // as they are final, while later rounds / chunks compute; host_copied = records already on their way
  t4c_contact* host_contacts = nullptr;
  int64_t host_copied = 0;
  // sharded entries (t4_multi.h): records are ALSO pushed, as soon as they are final, into this rank's region of the
  // root's gather slab -- peer memory, the copy engine writes it over NVLink on the copy stream while the later rounds
  // compute; pushed = records already on their way (the final exchange sends only what is left)
  t4c_contact* push_dst = nullptr;
  int64_t push_cap = 0, pushed = 0;
  // q arrives in q_pieces uploads of q_piece_states states each on the copy stream (event k = piece k is resident):
  // FK runs piece by piece under the rest of the upload
  int q_pieces = 0;
  int64_t q_piece_states = 0;
  // results
  int64_t n_contacts = 0;
  t4c_stats stats{};
};

// Pipeline parameters.  The values are the measured choices (profiles/README.md).  The environment variables are read
// ONCE, here: six TEST HOOKS that let small test batches reach the code paths of the million-state runs (rounds,
// budgets, chunking, progressive copies, the phase-1 certificate check), a trace of the exact passes, and three
// A/B switches (peer push, live ordering, slice growth) that the profiles' on / off comparisons use.  Nothing else in
// the library reads the environment.
struct Tuning
{
  // FIRST queries of at least this many states (per chunk) run in rounds over the robot's shapes (runBatch)
  int64_t round_min_states = 16384;  // T4C_ROUND_MIN_STATES
  // support pairs a survivor may spend in phases 1-2 of the exact pass before it is deferred to the long-pair pass
  // (0 = no budget).  One pair in a million cycles towards the detector's 1000-iteration caps; alone in its warp it
  // holds the launch it is in for milliseconds.  -1: 64 for FIRST queries (+8 % on batches that contain such pairs, no
  // change on batches that do not, profiles/r2k), none for CLOSEST / ALL (there the extra pass costs 5 %).
  int gjk_budget = -1;               // T4C_GJK_BUDGET
  int64_t chunk_states = 1ll << 20;  // T4C_CHUNK_STATES
  int64_t drain_min = 4096;          // T4C_DRAIN_MIN: records per progressive download
  int64_t upload_piece_min = 65536;  // T4C_UPLOAD_PIECE_MIN: states per piece of the piecewise q upload
  int max_rounds = 8;
  int verify_skip = 0;               // T4C_VERIFY_SKIP: run phase 1 for certified pairs too, count contradictions
  int trace = 0;                     // T4C_TRACE_PASSES: queue sizes of every exact pass on stderr
  int peer_push = 1;                 // T4C_PEER_PUSH=0: sharded entries move every record in the final exchange
  int epa_early = 1;                 // T4C_EPA_EARLY=0: one EPA launch after all deferred passes (r3o behaviour)
  int live_order = 1;                // T4C_LIVE_ORDER=0: rounds walk all states in index order (r3j behaviour)
  // a round takes unit * (this % * n / live states) shapes; 0 = off, -1 = the measured choice (profiles/r3l): 35 for
  // queries with records, 70 for flags-only queries
  int round_merge = -1;              // T4C_ROUND_MERGE
  Tuning()
  {
    auto env = [](const char* name, long long dflt, long long lo) {
      const char* e = std::getenv(name);
      return e ? std::max<long long>(lo, std::atoll(e)) : dflt;
    };
    round_min_states = env("T4C_ROUND_MIN_STATES", round_min_states, 1);
    gjk_budget = (int)env("T4C_GJK_BUDGET", gjk_budget, 0);
    chunk_states = env("T4C_CHUNK_STATES", chunk_states, 1);
    drain_min = env("T4C_DRAIN_MIN", drain_min, 1);
    upload_piece_min = env("T4C_UPLOAD_PIECE_MIN", upload_piece_min, 1);
    verify_skip = (int)env("T4C_VERIFY_SKIP", 0, 0);
    trace = (int)env("T4C_TRACE_PASSES", 0, 0);
    peer_push = (int)env("T4C_PEER_PUSH", 1, 0);
    live_order = (int)env("T4C_LIVE_ORDER", 1, 0);
    epa_early = (int)env("T4C_EPA_EARLY", 1, 0);
    round_merge = (int)env("T4C_ROUND_MERGE", round_merge, -1);
  }
};
static const Tuning& tuning()
{
  static const Tuning t;
  return t;
}
static int64_t roundMinStates() { return tuning().round_min_states; }
static int gjkBudget() { return tuning().gjk_budget; }
static int maxRounds() { return tuning().max_rounds; }
static int64_t chunkStates() { return tuning().chunk_states; }
// survivor queues are sorted by (types, iteration count of the FP32 cull): the exact pass of a pair needs about as many
// iterations as its cull did, so a warp's lanes finish together (GJK -6 %, r2g)
static bool sortClass() { return true; }
// 0: last active shapes first (distal links of a serial arm), 1: first shapes first.  Distal-first settles most
// states in the first rounds (the end of the arm sweeps the largest volume): 85 M instead of 208 M tree nodes per 1 Mi
// states on configs[1].  Round 1 measured queries WITH records faster proximal-first (fewer EPA runs); since tier 0
// holds only certified pairs EPA hardly runs, and distal-first wins for both (profiles/r3l: 79.9 M vs 75.7 M states/s
// with records, 239.6 M vs 195.6 M flags-only).
static int roundOrder(bool) { return 0; }

static uint32_t nextPow2(uint64_t x)
{
  uint64_t p = 16;
  while (p < x)
    p <<= 1;
  return (uint32_t)std::min<uint64_t>(p, 1ull << 31);
}

#define T4_TRY(expr, what)                                                                                             \
  do                                                                                                                   \
  {                                                                                                                    \
    if (!(expr))                                                                                                       \
      return fail(T4C_ERR_CUDA, "%s: %s", what, plat::lastError());                                                    \
  } while (0)

#ifndef T4C_HOSTSIM
// stable radix sort of a queue of n pairs by shape-type key; returns the sorted values (candidate indices when
// from_tasks == 0, task indices otherwise) or null when sorting is off / not worthwhile
static const uint32_t* sortByType(t4c_scene* s, const KParams& P, const uint32_t* queue, int from_tasks, uint32_t n)
{
  if (n < 4096)
    return nullptr;
  t4stream st = s->stream;
  if (!(s->d_sort_keys.ensure((size_t)n * 2) && s->d_sort_keys2.ensure((size_t)n * 2) && s->d_sort_vals.ensure((size_t)n * 4) &&
        s->d_sort_idx.ensure((size_t)n * 4)))
    return nullptr;
  typeKeyKernel<<<gridFor(n, 256, 8), 256, 0, st>>>(P, queue, from_tasks, s->d_sort_keys.as<uint16_t>(),
                                                    s->d_sort_idx.as<uint32_t>(), (int64_t)n);
  ++g_launches;
  const uint32_t* vals_in = from_tasks ? s->d_sort_idx.as<uint32_t>() : queue;
  size_t tmp = 0;
  // task queues and class-less runs have zero low bits: one 8-bit digit is enough for them
  const int bit0 = (!from_tasks && P.cand_class) ? 0 : 7;
  if (cub::DeviceRadixSort::SortPairs(nullptr, tmp, s->d_sort_keys.as<uint16_t>(), s->d_sort_keys2.as<uint16_t>(), vals_in,
                                      s->d_sort_vals.as<uint32_t>(), (int)n, bit0, 15, st) != cudaSuccess ||
      !s->d_sort_tmp.ensure(tmp) ||
      cub::DeviceRadixSort::SortPairs(s->d_sort_tmp.p, tmp, s->d_sort_keys.as<uint16_t>(), s->d_sort_keys2.as<uint16_t>(),
                                      vals_in, s->d_sort_vals.as<uint32_t>(), (int)n, bit0, 15, st) != cudaSuccess)
    return nullptr;
  return s->d_sort_vals.as<uint32_t>();
}
#else
static const uint32_t* sortByType(t4c_scene*, const KParams&, const uint32_t*, int, uint32_t) { return nullptr; }
#endif

// stable radix sort of the round's state list by (done flag, Morton cell): keys / values of liveKeyBody -> `out`
static bool sortLive(t4c_scene* s, const KParams& P, uint32_t n, uint32_t* out)
{
#ifndef T4C_HOSTSIM
  t4stream st = s->stream;
  size_t tmp = 0;
  const int bits = 3 * T4_LIVE_BITS + 1;
  return s->d_live_key2.ensure((size_t)n * 4) &&
         cub::DeviceRadixSort::SortPairs(nullptr, tmp, P.live_key, s->d_live_key2.as<uint32_t>(), P.live_val, out, (int)n, 0,
                                         bits, st) == cudaSuccess &&
         s->d_sort_tmp.ensure(tmp) &&
         cub::DeviceRadixSort::SortPairs(s->d_sort_tmp.p, tmp, P.live_key, s->d_live_key2.as<uint32_t>(), P.live_val, out,
                                         (int)n, 0, bits, st) == cudaSuccess;
#else
  (void)s;
  std::vector<uint32_t> idx(n);
  for (uint32_t i = 0; i < n; ++i)
    idx[i] = i;
  std::stable_sort(idx.begin(), idx.end(), [&](uint32_t a, uint32_t b) { return P.live_key[a] < P.live_key[b]; });
  for (uint32_t i = 0; i < n; ++i)
    out[i] = P.live_val[idx[i]];
  return true;
#endif
}

static t4c_status runBatchChunked(t4c_scene* s, BatchIO& io, int64_t CH);

// The per-chunk workspace (pose arrays, candidate / survivor queues, raw-contact staging, CLOSEST hash table) scales with
// the states of a chunk times the pairs per state, which the caller cannot know in advance: 1 Mi states against ten
// thousand primitives need 2 GB, against a million voxels at a 0.10 m margin a hundred times that.  When a device
// allocation fails, the workspace is released and the batch is run again in chunks of half the size; the scene remembers
// the size that worked.
static t4c_status runBatch(t4c_scene* s, BatchIO& io)
{
  int64_t ch = s->chunk_hint > 0 ? s->chunk_hint : chunkStates();
  for (;;)
  {
    BatchIO attempt = io;
    g_alloc_failed = false;
    const t4c_status rc = runBatchChunked(s, attempt, ch);
    const bool nomem = rc == T4C_ERR_NOMEM || (rc == T4C_ERR_CUDA && g_alloc_failed);
    if (!nomem || ch <= 2048)
    {
      io = attempt;
      if (nomem)
        return fail(T4C_ERR_NOMEM, "out of device memory even in chunks of %lld states", (long long)ch);
      return rc;
    }
    plat::sync(s->stream);
    if (s->aux_stream)
      plat::sync(s->aux_stream);
    if (s->epa_stream)
      plat::sync(s->epa_stream);
    if (s->copy_stream)
      plat::sync(s->copy_stream);
    s->releaseWorkspace();
    ch = std::max<int64_t>(2048, ch / 2);
    s->chunk_hint = ch;
  }
}

static t4c_status runBatchChunked(t4c_scene* s, BatchIO& io, int64_t CH)
{
  t4c_status rc = syncScene(s);
  if (rc != T4C_OK)
    return rc;
  if (io.req.type < T4C_TEST_FIRST || io.req.type > T4C_TEST_ALL)
    return fail(T4C_ERR_INVALID, "unknown ContactTestType %d (LIMITED is not implemented by the reference either)",
                io.req.type);
  t4stream st = s->stream;
  KParams P;
  fillSceneParams(s, P);
  FkCache* fc = nullptr;
  if (io.model)
  {
    rc = syncFk(s, io.model, &fc);
    if (rc != T4C_OK)
      return rc;
    fillFkParams(io.model, fc, P);
  }
  P.cast = io.cast ? 1 : 0;
  P.test_type = io.req.type;
  P.calc_pen = io.req.calculate_penetration;
  P.calc_dist = io.req.calculate_distance;
  P.hit_only = io.contacts == nullptr ? 1 : 0;
  P.q_layout = io.q_layout;
  P.q_stride_n = io.n;
  T4_TRY(s->d_qcount.ensure(Q_COUNT * sizeof(uint32_t)) && s->d_stats.ensure(ST_COUNT * sizeof(unsigned long long)),
         "workspace");
  P.qcount = s->d_qcount.as<uint32_t>();
  P.stats = s->d_stats.as<unsigned long long>();
  T4_TRY(plat::zero(P.stats, ST_COUNT * sizeof(unsigned long long), st), "memset");
  unsigned long long stats_host[ST_COUNT];
  std::memset(stats_host, 0, sizeof(stats_host));
  unsigned long long stats_pruned = 0;  // survivors of the cull that the CLOSEST pruning kept out of the exact pass
  const int na = P.n_alinks;
  int64_t total = 0;
  const size_t qelem = io.q_layout == T4C_Q_F64_ROWMAJOR ? sizeof(double) : sizeof(float);
  // records [0, upto) are final on the device (the compute stream has just been synchronised): send what is new to
  // the caller's page-locked buffer on the copy stream, under the kernels that follow
  bool aux_busy = false;  // set while kernels of the long-pair pass run beside the main stream (they append records too)
  auto drain = [&](int64_t upto) -> bool {
    if (!io.contacts || aux_busy)
      return true;
    upto = std::min<int64_t>(upto, io.capacity);
    const int64_t min_records = tuning().drain_min;
    bool ok = true;
    if (io.host_contacts && upto - io.host_copied >= min_records)
    {
      ok = plat::d2h(io.host_contacts + io.host_copied, io.contacts + io.host_copied,
                     (size_t)(upto - io.host_copied) * sizeof(t4c_contact), s->copy_stream);
      io.host_copied = upto;
    }
    const int64_t to = std::min<int64_t>(upto, io.push_cap);
    if (io.push_dst && to - io.pushed >= min_records)
    {
      ok = ok && plat::d2d(io.push_dst + io.pushed, io.contacts + io.pushed,
                           (size_t)(to - io.pushed) * sizeof(t4c_contact), s->copy_stream);
      io.pushed = to;
    }
    return ok;
  };
  for (int64_t base = 0; base < io.n; base += CH)
  {
    const int64_t nc = std::min<int64_t>(CH, io.n - base);
    P.n = nc;
    P.state_base = base + io.state_offset;
    // ---- poses of the kinematic links
    const size_t pose_elems = (size_t)nc * (size_t)std::max(na, 1) * 12;
    T4_TRY(s->d_poses0.ensure(pose_elems * sizeof(float)) && s->d_poses0d.ensure(pose_elems * sizeof(double)),
           "workspace");
    P.poses0 = s->d_poses0.as<float>();
    P.poses0d = s->d_poses0d.as<double>();
    if (io.cast)
    {
      T4_TRY(s->d_poses1.ensure(pose_elems * sizeof(float)) && s->d_poses1d.ensure(pose_elems * sizeof(double)),
             "workspace");
      P.poses1 = s->d_poses1.as<float>();
      P.poses1d = s->d_poses1d.as<double>();
    }
    P.fk_out = nullptr;
    if (io.model)
    {
      const size_t qoff = io.q_layout == T4C_Q_F32_SOA ? (size_t)base : (size_t)base * (size_t)io.model->n_q;
      P.q0 = (const char*)io.q0 + qoff * qelem;
      P.q1 = io.q1 ? (const char*)io.q1 + qoff * qelem : nullptr;
      if (io.q_pieces > 0)
      {
        for (int k = 0; k < io.q_pieces; ++k)
        {
          const int64_t a = std::max<int64_t>(base, (int64_t)k * io.q_piece_states);
          const int64_t b = std::min<int64_t>(base + nc, (int64_t)(k + 1) * io.q_piece_states);
          if (a >= b)
            continue;
          T4_TRY(plat::wait(st, s->q_events[k]), "upload q");
          launchFk(P, a - base, b - a, st);
        }
      }
      else
        launchFk(P, 0, nc, st);
    }
    else if (na > 0)
    {
      P.poses_in0 = io.poses0 + (size_t)base * na * 12;
      P.poses_in1 = io.poses1 ? io.poses1 + (size_t)base * na * 12 : nullptr;
      launchConvert(P, nc * na, st);
    }
    // ---- per-state flags
    if (io.hit)
      P.hit = io.hit + base;
    else
    {
      T4_TRY(s->d_hit.ensure((size_t)nc), "workspace");
      P.hit = s->d_hit.as<uint8_t>();
    }
    T4_TRY(plat::zero(P.hit, (size_t)nc, st), "memset");
    if (io.req.type == T4C_TEST_FIRST && io.contacts)
    {
      T4_TRY(s->d_claim.ensure((size_t)nc * sizeof(uint32_t)), "workspace");
      P.first_claim = s->d_claim.as<uint32_t>();
      T4_TRY(plat::zero(P.first_claim, (size_t)nc * sizeof(uint32_t), st), "memset");
    }
    // ---- FIRST queries run in ROUNDS over the robot's shapes: broadphase + cull + exact pass for one slice of the
    //      active shapes at a time, and states that already have a contact take no part in the later rounds.  This
    //      is the reference's early exit (ContactTestData::done, core/common.h:129-133: Bullet stops dispatching
    //      pairs once the first contact is stored) applied per state across the batch.
    uint32_t qc[Q_COUNT] = {};
    const bool first_rec = io.req.type == T4C_TEST_FIRST && io.contacts;
    const bool closest = io.req.type == T4C_TEST_CLOSEST && io.contacts;
    const int64_t room = std::max<int64_t>(0, io.capacity - total);
    const int32_t* all_active = P.active_shapes;
    const int n_all_active = P.n_active_shapes, n_all_aa = P.n_aa_pairs;
    int n_rounds = 1;
    if (io.req.type == T4C_TEST_FIRST && nc >= roundMinStates() && n_all_active > 1 && P.n_bvh_nodes > 0)
      n_rounds = std::min(n_all_active, maxRounds());
    P.skip_hit = n_rounds > 1 ? 1 : 0;
    // rounds walk an ordered list of the states that are still without a contact (Morton grid over the scene box)
    bool use_live = n_rounds > 1 && P.n_bvh_nodes > 0 && tuning().live_order;
    const uint32_t* live_cur = nullptr;
    uint32_t live_n = 0;
    if (use_live)
    {
      const size_t lb = (size_t)nc * sizeof(uint32_t);
      if (!(s->d_live_a.ensure(lb) && s->d_live_b.ensure(lb) && s->d_live_key.ensure(lb) && s->d_live_val.ensure(lb)))
      {
        g_alloc_failed = false;
        use_live = false;
      }
      const DBvhNode& R = s->img->comp.bvh[0];
      for (int k = 0; k < 3; ++k)
      {
        const float mn = std::min(R.lmn[k], R.rmn[k]) - 0.5f, mx = std::max(R.lmx[k], R.rmx[k]) + 0.5f;
        P.live_org[k] = mn;
        P.live_scale[k] = (float)(1 << T4_LIVE_BITS) / std::max(mx - mn, 1e-3f);
      }
    }
    // Candidates APPEND across the rounds (survivor queues hold candidate indices).  Per round only the first tier of
    // survivors is resolved: pairs that are certainly in contact without penetrating -- one cheap GJK each.  The
    // other tiers (one certainly penetrating pair per state, then the rest) accumulate and are resolved after the
    // last round, for the states that are still without a contact: EPA runs at most once per such state.
    if (s->cand_cap == 0)
      s->cand_cap = (uint32_t)std::min<int64_t>(std::max<int64_t>(1 << 16, 24 * nc), 1ll << 30);
    T4_TRY(plat::zero(P.qcount, Q_COUNT * sizeof(uint32_t), st), "memset");
    P.surv2_w = nullptr;
    P.surv3_w = nullptr;
    P.surv4_w = nullptr;
    P.cull_claim = nullptr;
    if (first_rec)
    {
      T4_TRY(s->d_cull_claim.ensure((size_t)nc * sizeof(uint32_t)) &&
                 plat::zero(s->d_cull_claim.p, (size_t)nc * sizeof(uint32_t), st),
             "survivor queue");
      P.cull_claim = s->d_cull_claim.as<uint32_t>();
    }
    P.raw = nullptr;
    P.raw_local = nullptr;
    P.raw_d64 = nullptr;
    P.raw_cap = 0;
    if (io.contacts && !closest)
    {
      P.raw = io.contacts + std::min<int64_t>(total, io.capacity);
      P.raw_cap = (uint32_t)std::min<int64_t>(room, 0xfffffff0ll);
      P.raw_local = (io.cast && io.cc_local) ? io.cc_local + (size_t)std::min<int64_t>(total, io.capacity) * 6 : nullptr;
    }
    uint32_t cand_begin = 0, nsurv2 = 0, nsurv3 = 0, nsurv4 = 0;
    uint32_t long_cap_bound = 0, long_done = 0;
    // CLOSEST against links that are compounds of many shapes: prune the exact queue with the cull's bounds (pruneBody)
    const int n_groups = s->img->comp.n_groups;
    const bool prune = closest && n_groups > 0 && (int64_t)na * n_groups <= 4096;
    P.link_group = nullptr, P.n_groups = 0, P.key_min_ub = nullptr, P.cand_lb = nullptr;
    if (prune)
    {
      const size_t kb = (size_t)nc * (size_t)na * (size_t)n_groups * sizeof(uint32_t);
      T4_TRY(s->d_key_min.ensure(kb) && plat::fill(s->d_key_min.p, 0xff, kb, st), "pruning table");
      P.link_group = s->img->d_link_group.as<int32_t>();
      P.n_groups = n_groups;
      P.key_min_ub = s->d_key_min.as<uint32_t>();
    }
    // kernels of the long-pair pass may be running on the auxiliary stream (see launchLongAsync below): they are joined
    // before any queue they read is reallocated, and on EVERY way out of this function
    struct AuxJoin
    {
      t4stream aux;
      bool inflight = false;
      void join()
      {
        if (inflight && aux)
          plat::sync(aux);
        inflight = false;
      }
      ~AuxJoin() { join(); }
    } aux_join{ s->aux_stream };
    uint32_t bound_cap = 0;
    auto bindQueues = [&]() -> bool {
      if (aux_join.inflight && s->cand_cap != bound_cap)
        aux_join.join();  // the queues are about to grow (and move)
      bound_cap = s->cand_cap;
      const size_t cb = (size_t)s->cand_cap * sizeof(uint32_t), keep = (size_t)cand_begin * sizeof(uint32_t);
      if (!(s->d_cand_state.ensureKeep(cb, keep, st) && s->d_cand_a.ensureKeep(cb, keep, st) &&
            s->d_cand_b.ensureKeep(cb, keep, st) && s->d_surv.ensure(cb)))
        return false;
      if (first_rec && !(s->d_surv2.ensureKeep(cb, keep, st) && s->d_surv3.ensureKeep(cb, keep, st) &&
                         s->d_surv4.ensureKeep(cb, keep, st)))
        return false;
      P.cand_state = s->d_cand_state.as<uint32_t>();
      P.cand_a = s->d_cand_a.as<uint32_t>();
      P.cand_b = s->d_cand_b.as<uint32_t>();
      P.cand_class = nullptr;
      if (sortClass())
      {
        if (!s->d_cand_class.ensureKeep((size_t)s->cand_cap, (size_t)cand_begin, st))
          return false;
        P.cand_class = s->d_cand_class.as<uint8_t>();
      }
      P.cand_cap = s->cand_cap;
      if (prune)
      {
        if (!(s->d_cand_lb.ensureKeep((size_t)s->cand_cap * sizeof(float), (size_t)cand_begin * sizeof(float), st) &&
              s->d_surv2.ensure(cb)))
          return false;
        P.cand_lb = s->d_cand_lb.as<float>();
        P.surv2_w = s->d_surv2.as<uint32_t>();
      }
      P.long_w = nullptr;
      P.long_cap = 0;
      P.task_q = Q_TASK;
      P.gjk_budget = gjkBudget() >= 0 ? gjkBudget() : (io.req.type == T4C_TEST_FIRST ? 64 : 0);
      P.verify_skip = tuning().verify_skip;
      if (P.gjk_budget > 0)
      {
        if (!s->d_long.ensureKeep(cb, (size_t)std::min(qc[Q_LONG], long_cap_bound) * sizeof(uint32_t), st))
          return false;
        long_cap_bound = s->cand_cap;
        P.long_w = s->d_long.as<uint32_t>();
        P.long_cap = s->cand_cap;
      }
      P.surv_w = s->d_surv.as<uint32_t>();
      P.surv_cap = s->cand_cap;
      if (first_rec)
      {
        P.surv2_w = s->d_surv2.as<uint32_t>();
        P.surv3_w = s->d_surv3.as<uint32_t>();
        P.surv4_w = s->d_surv4.as<uint32_t>();
      }
      return true;
    };
    uint32_t n_long_async = 0;  // long pairs whose GJK has been launched on the auxiliary stream
    auto launchLongAsync = [&]() -> t4c_status {
      if (!P.long_w || !s->aux_stream)
        return T4C_OK;
      const uint32_t nlong = std::min(qc[Q_LONG], P.long_cap);  // qc: read back by the exact pass just before
      if (nlong <= long_done)
        return T4C_OK;
      const size_t need = (size_t)nlong * sizeof(EpaTask);
      if (need > s->d_tasks_long.cap)
      {
        aux_join.join();  // the task queue of the running launches moves
        T4_TRY(s->d_tasks_long.ensureKeep(std::max<size_t>(need, 4096 * sizeof(EpaTask)),
                                          (size_t)n_long_async * sizeof(EpaTask), st),
               "EPA task queue");
      }
      KParams PL = P;
      PL.gjk_budget = 0;
      PL.long_w = nullptr;
      PL.surv = P.long_w + long_done;
      PL.tasks = s->d_tasks_long.as<EpaTask>();
      PL.task_cap = (uint32_t)(s->d_tasks_long.cap / sizeof(EpaTask));
      PL.task_q = Q_TASK_LONG;
      PL.task_order = nullptr;
      launchGjk(PL, nlong - long_done, s->aux_stream);
      aux_join.inflight = true;
      aux_busy = true;  // slots below the record counter may still be unwritten: no progressive download meanwhile
      n_long_async += nlong - long_done;
      long_done = nlong;
      return T4C_OK;
    };
    // one exact pass over a survivor queue: GJK kernel -> EPA task queue [-> EPA kernel].  With run_epa == false the
    // tasks stay queued (and the next pass appends to them): a launch of a few hundred EPA tasks lasts as long as its
    // slowest task, so the rounds share ONE EPA launch instead of paying that tail once per round.
    uint32_t ntask_pending = 0;
    size_t task_seg = 0;  // tasks in front of the pending ones that have been handed to an EPA launch already
    struct EpaJoin
    {
      t4stream stream;
      bool inflight = false;
      void join()
      {
        if (inflight && stream)
          plat::sync(stream);
        inflight = false;
      }
      ~EpaJoin() { join(); }
    } epa_join{ s->epa_stream };
    auto runEpa = [&](uint32_t ntask) -> t4c_status {
      launchEpa(P, ntask, st);
      return T4C_OK;
    };
    auto exactPass = [&](uint32_t* queue, uint32_t nq, bool run_epa) -> t4c_status {
      if (nq > 0)
      {
        const size_t need = (task_seg + ntask_pending + nq) * sizeof(EpaTask);
        if (need > s->d_tasks.cap)
          epa_join.join();  // the queue is about to move under the launch that reads its first segment
        T4_TRY(s->d_tasks.ensureKeep(need, (task_seg + ntask_pending) * sizeof(EpaTask), st), "EPA task queue");
        P.tasks = s->d_tasks.as<EpaTask>() + task_seg;
        P.task_cap = ntask_pending + nq;
        P.surv = queue;
        P.task_order = nullptr;
        if (const uint32_t* sorted = sortByType(s, P, P.surv, 0, nq))
          P.surv = const_cast<uint32_t*>(sorted);
        launchGjk(P, nq, st);
        T4_TRY(launchOk() && plat::d2h(qc, P.qcount, sizeof(qc), st) && plat::sync(st), "narrowphase (GJK)");
        if (tuning().trace)
          std::fprintf(stderr, "[t4c] exact pass: %u pairs, EPA tasks %u -> %u, long %u, records %u\n", nq, ntask_pending,
                       std::min(qc[Q_TASK], P.task_cap), qc[Q_LONG], qc[Q_RAW]);
        ntask_pending = std::min(qc[Q_TASK], P.task_cap);
      }
      if (run_epa && ntask_pending > 0)
      {
        P.task_order = sortByType(s, P, nullptr, 1, ntask_pending);
        const t4c_status re = runEpa(ntask_pending);
        if (re != T4C_OK)
          return re;
        ntask_pending = 0;
        T4_TRY(plat::zero(P.qcount + Q_TASK, sizeof(uint32_t), st), "memset");
      }
      return T4C_OK;
    };
    // The pending EPA tasks start NOW on their own stream while the main stream goes on with the next exact passes;
    // tasks queued from here on form a new segment of the queue.  An EPA launch of a few thousand tasks lasts as long
    // as its slowest task (1.5 - 2 ms) with a tenth of the GPU's thread slots busy: it costs the passes beside it
    // next to nothing, and they no longer wait in front of it.
    auto startEpaEarly = [&]() -> t4c_status {
      if (!s->epa_stream || !tuning().epa_early || ntask_pending < 256 || epa_join.inflight)
        return T4C_OK;
      KParams PE = P;
      const uint32_t* order = sortByType(s, P, nullptr, 1, ntask_pending);
      if (order)
      {
        if (!s->d_epa_order.ensure((size_t)ntask_pending * sizeof(uint32_t)))
        {
          g_alloc_failed = false;
          return T4C_OK;  // no room for the private order: the tasks stay pending for the launch at the end
        }
        T4_TRY(plat::d2d(s->d_epa_order.p, order, (size_t)ntask_pending * sizeof(uint32_t), st), "EPA task order");
        PE.task_order = s->d_epa_order.as<uint32_t>();
      }
      else
        PE.task_order = nullptr;
      T4_TRY(plat::zero(P.qcount + Q_TASK, sizeof(uint32_t), st) && plat::sync(st), "EPA task queue");
      launchEpa(PE, ntask_pending, s->epa_stream);
      epa_join.inflight = true;
      task_seg += ntask_pending;
      ntask_pending = 0;
      return T4C_OK;
    };
    // pairs whose GJK ran out of its iteration budget (Q_LONG): all of them together, without a budget
    auto longPass = [&](bool run_epa) -> t4c_status {
      if (!P.long_w)
        return run_epa ? exactPass(nullptr, 0, true) : T4C_OK;
      T4_TRY(launchOk() && plat::d2h(qc, P.qcount, sizeof(qc), st) && plat::sync(st), "narrowphase (GJK)");
      const uint32_t nlong = std::min(qc[Q_LONG], P.long_cap);
      const int budget = P.gjk_budget;
      P.gjk_budget = 0;
      const t4c_status r = exactPass(P.long_w + long_done, nlong - long_done, run_epa);
      P.gjk_budget = budget;
      long_done = nlong;
      return r;
    };
    // Slices of the robot's shapes, one per round.  A round has a fixed cost (five host round trips, a dozen
    // launches: ~0.15 ms) and pays by what the states it settles do not have to do in the later rounds, so the slices
    // grow as the live states get fewer: `unit` shapes while most states are live, then unit * (merge * n / live).
    const int unit = n_rounds > 1 ? (n_all_active + maxRounds() - 1) / maxRounds() : n_all_active;
    const int order = roundOrder(first_rec);
    int cursor = 0;  // shapes taken so far
    for (int round = 0; cursor < n_all_active; ++round)
    {
      const int remaining = n_all_active - cursor;
      int take = std::min(unit, remaining);
      const int merge = tuning().round_merge >= 0 ? tuning().round_merge : (first_rec ? 35 : 70);
      if (n_rounds > 1 && round > 0 && merge > 0)
      {
        int64_t est_live = first_rec ? nc - (int64_t)std::min<uint32_t>(qc[Q_RAW], (uint32_t)nc) : nc;
        if (live_cur)
          est_live = std::min<int64_t>(est_live, live_n);  // counted at the start of the previous round
        const int64_t mult = (int64_t)merge * nc / (100 * std::max<int64_t>(est_live, 1));
        take = (int)std::min<int64_t>(remaining, (int64_t)unit * std::max<int64_t>(mult, 1));
      }
      if (n_rounds > 1 && round >= maxRounds() - 1)
        take = remaining;
      const bool last_round = cursor + take == n_all_active;
      // order 0: from the far end of the kinematic chain (distal links sweep the largest volume), 1: from the base
      const int lo = order ? cursor : n_all_active - cursor - take;
      const int hi = lo + take;
      cursor += take;
      P.active_shapes = all_active + lo;
      P.n_active_shapes = hi - lo;
      P.n_aa_pairs = round == 0 ? n_all_aa : 0;
      // ---- this round's states in Morton order of the round's first shape, the states that are done last
      //      (liveKeyBody).  The list of the previous round is the input: it shrinks as the states settle.
      P.live = nullptr;
      P.n_live = 0;
      if (use_live)
      {
        const uint32_t n_prev = live_cur ? live_n : (uint32_t)nc;
        uint32_t* const out = (live_cur == s->d_live_a.as<uint32_t>()) ? s->d_live_b.as<uint32_t>() : s->d_live_a.as<uint32_t>();
        if (n_prev > 0)
        {
          T4_TRY(plat::zero(P.qcount + Q_LIVE, sizeof(uint32_t), st), "memset");
          P.live_prev = live_cur;
          P.live_key = s->d_live_key.as<uint32_t>();
          P.live_val = s->d_live_val.as<uint32_t>();
          launchLiveKey(P, n_prev, st);
          if (sortLive(s, P, n_prev, out))
          {
            P.live = out;
            P.n_live = n_prev;
          }
          else
          {
            g_alloc_failed = false;  // no room for the sort: this round walks every state in index order
            use_live = false;
          }
        }
        else
        {
          P.live = out;  // every state is done: nothing to traverse
          P.n_live = 0;
        }
      }
      const int64_t per_shape = P.live ? (int64_t)P.n_live : nc;
      // ---- broadphase (retried with a larger queue when it overflows)
      for (int attempt = 0;; ++attempt)
      {
        T4_TRY(bindQueues(), "candidate queue");
        if (P.n_bvh_nodes > 0)
          launchBroadStatic(P, per_shape * (int64_t)P.n_active_shapes, st);
        launchBroadActive(P, nc * (int64_t)P.n_aa_pairs, st);
        T4_TRY(launchOk() && plat::d2h(qc, P.qcount, sizeof(qc), st) && plat::sync(st), "broadphase");
        if (qc[Q_CAND] <= s->cand_cap)
          break;
        if (attempt >= 2)
          return fail(T4C_ERR_NOMEM, "candidate queue overflow (%u pairs)", qc[Q_CAND]);
        // grow (keeping the candidates of the earlier rounds), rewind the append counter to this round's start and
        // run the round's broadphase again; the statistics counters count the repeated traversal twice
        s->cand_cap = (uint32_t)std::min<uint64_t>((uint64_t)qc[Q_CAND] + qc[Q_CAND] / 8 + 1024, 0xfffffff0ull);
        T4_TRY(plat::h2d(P.qcount + Q_CAND, &cand_begin, sizeof(uint32_t), st) && plat::sync(st), "queue rewind");
      }
      const uint32_t ncand = qc[Q_CAND];
      if (P.live)
      {
        live_cur = P.live;
        live_n = std::min(qc[Q_LIVE], P.n_live);
      }
      // ---- FP32 cull of this round's candidates
      T4_TRY(plat::zero(P.qcount + Q_SURV, sizeof(uint32_t), st), "memset");
      P.cull_base = cand_begin;
      launchCull(P, ncand - cand_begin, st);
      T4_TRY(launchOk() && plat::d2h(qc, P.qcount, sizeof(qc), st) && plat::sync(st), "cull");
      if (!closest)
        T4_TRY(drain(total + qc[Q_RAW]), "contact download");  // the records of the earlier rounds
      cand_begin = ncand;
      const uint32_t nsurv = qc[Q_SURV];
      nsurv2 = qc[Q_SURV2], nsurv3 = qc[Q_SURV3], nsurv4 = qc[Q_SURV4];
      // ---- CLOSEST pruning: the survivors that can still be the minimum of their key
      uint32_t* exact_queue = s->d_surv.as<uint32_t>();
      uint32_t n_exact = nsurv;
      if (prune && nsurv > 0)
      {
        T4_TRY(plat::zero(P.qcount + Q_SURV2, sizeof(uint32_t), st), "memset");
        P.surv = s->d_surv.as<uint32_t>();
        launchPrune(P, nsurv, st);
        T4_TRY(launchOk() && plat::d2h(qc, P.qcount, sizeof(qc), st) && plat::sync(st), "prune");
        exact_queue = s->d_surv2.as<uint32_t>();
        n_exact = qc[Q_SURV2];
        stats_pruned += nsurv - n_exact;
      }
      // ---- FP64 exact pass over the first tier
      if (closest)
      {
        T4_TRY(s->d_raw.ensure(std::max<size_t>(nsurv, 1) * sizeof(t4c_contact)), "contact staging");
        P.raw = s->d_raw.as<t4c_contact>();
        P.raw_cap = nsurv;
        T4_TRY(s->d_raw_d64.ensure(std::max<size_t>(nsurv, 1) * sizeof(double)), "contact staging");
        P.raw_d64 = s->d_raw_d64.as<double>();
        if (io.cast && io.cc_local)
        {
          T4_TRY(s->d_raw_local.ensure(std::max<size_t>(nsurv, 1) * 6 * sizeof(float)), "contact staging");
          P.raw_local = s->d_raw_local.as<float>();
        }
      }
      // A launch of the exact kernel lasts at least as long as one pair (~0.15 ms on an otherwise idle GPU): the few
      // tier-0 pairs of the last round (the links next to the base) join the first deferred pass instead of paying
      // that latency for a launch of their own.
      if (first_rec && last_round && n_rounds > 1 && n_exact > 0 && n_exact < 4096 &&
          (uint64_t)nsurv4 + n_exact <= (uint64_t)P.surv_cap)
      {
        T4_TRY(plat::d2d(s->d_surv4.as<uint32_t>() + nsurv4, exact_queue, (size_t)n_exact * sizeof(uint32_t), st),
               "survivor queue");
        nsurv4 += n_exact;
        n_exact = 0;
      }
      // with a long-pair queue the EPA launch waits until the long pairs' GJK has been started beside it
      rc = exactPass(exact_queue, n_exact, last_round && !P.long_w);
      if (rc != T4C_OK)
        return rc;
      // ---- long pairs of this round: a pair that cycles towards the detector's 1000-iteration caps keeps one lane
      //      busy for ~9 ms.  Its GJK starts NOW on the auxiliary stream, under the rounds that follow, the deferred
      //      tiers and the EPA launch of the main stream (r3c: started only after the last round, the pass outlasted
      //      its cover by 3 ms on the ranks whose batch holds such pairs -- the 0.90 scaling efficiency of round 1).
      rc = launchLongAsync();
      if (rc != T4C_OK)
        return rc;
    }
    // ---- deferred tiers, for the states that are still without a contact (no pair of theirs is certainly in contact
    //      with its cores certainly apart): the rest first -- a pair that turns out to be separated and within the
    //      margin settles the state without EPA -- then the certainly penetrating ones.  A launch of a few thousand
    //      EPA tasks lasts as long as its slowest task (1.5 - 2 ms against a cylinder): the tasks of the first pass
    //      start at once on their own stream (startEpaEarly), the others share ONE launch at the end (longPass
    //      below), and tasks of states that were settled meanwhile leave at once.
    rc = exactPass(s->d_surv4.as<uint32_t>(), nsurv4, false);
    if (rc != T4C_OK)
      return rc;
    if (first_rec && n_rounds > 1)
    {
      rc = startEpaEarly();  // most EPA tasks come out of this pass (two thirds on configs[1])
      if (rc != T4C_OK)
        return rc;
    }
    rc = exactPass(s->d_surv2.as<uint32_t>(), nsurv2, false);
    if (rc != T4C_OK)
      return rc;
    rc = exactPass(s->d_surv3.as<uint32_t>(), nsurv3, false);
    if (rc != T4C_OK)
      return rc;
    if (n_long_async > 0)
    {
      // join the long-pair pass; its penetrating pairs go through EPA now
      aux_join.join();
      aux_busy = false;
      T4_TRY(launchOk() && plat::d2h(qc, P.qcount, sizeof(qc), st) && plat::sync(st), "narrowphase (long pairs)");
      const uint32_t ntl = std::min(qc[Q_TASK_LONG], n_long_async);
      if (ntl > 0)
      {
        EpaTask* const keep_tasks = P.tasks;
        P.tasks = s->d_tasks_long.as<EpaTask>();
        P.task_order = nullptr;
        const t4c_status re = runEpa(ntl);
        P.tasks = keep_tasks;
        if (re != T4C_OK)
          return re;
      }
    }
    rc = longPass(true);  // long pairs that came out of the deferred tiers
    if (rc != T4C_OK)
      return rc;
    P.active_shapes = all_active;
    P.n_active_shapes = n_all_active;
    P.n_aa_pairs = n_all_aa;
    epa_join.join();
    T4_TRY(launchOk() && plat::d2h(qc, P.qcount, sizeof(qc), st) && plat::sync(st), "narrowphase (EPA)");
    const uint32_t nraw = qc[Q_RAW];
    if (closest)
    {
      // ---- ContactTestType::CLOSEST: minimum per (state, link pair)
      P.ht_size = nextPow2(2ull * nraw);
      T4_TRY(s->d_ht_rep.ensure((size_t)P.ht_size * sizeof(int32_t)) &&
                 s->d_ht_best.ensure((size_t)P.ht_size * sizeof(unsigned long long)),
             "hash table");
      P.ht_rep = s->d_ht_rep.as<int32_t>();
      P.ht_best = s->d_ht_best.as<unsigned long long>();
      T4_TRY(plat::fill(P.ht_rep, 0xff, (size_t)P.ht_size * sizeof(int32_t), st) &&
                 plat::fill(P.ht_best, 0xff, (size_t)P.ht_size * sizeof(unsigned long long), st),
             "memset");
      P.out = io.contacts + std::min<int64_t>(total, io.capacity);
      P.out_cap = (uint32_t)std::min<int64_t>(room, 0xfffffff0ll);
      P.out_local = (io.cast && io.cc_local) ? io.cc_local + (size_t)std::min<int64_t>(total, io.capacity) * 6 : nullptr;
      launchClosestInsert(P, nraw, st);
      launchClosestEmit(P, nraw, st);
      T4_TRY(launchOk() && plat::d2h(qc, P.qcount, sizeof(qc), st) && plat::sync(st), "closest reduction");
      total += qc[Q_OUT];
    }
    else
      total += nraw;
    T4_TRY(drain(total), "contact download");  // this chunk's records, under the next chunk
    launchCountHit(P, nc, st);
    T4_TRY(launchOk() && plat::d2h(stats_host, P.stats, sizeof(stats_host), st) && plat::sync(st), "stats");
  }
  io.n_contacts = total;
  stats_host[ST_STATES] = (unsigned long long)io.n;
  stats_host[ST_CONTACTS] = (unsigned long long)total;
  io.stats.states = stats_host[ST_STATES];
  io.stats.bvh_nodes = stats_host[ST_BVH_NODES];
  io.stats.candidates = stats_host[ST_CANDIDATES];
  io.stats.culled = stats_host[ST_CULLED];
  io.stats.gjk_pairs = stats_host[ST_GJK_PAIRS];
  io.stats.gjk_iterations = stats_host[ST_CULL_ITERS] + stats_host[ST_P1_ITERS] + stats_host[ST_P2_ITERS] +
                            stats_host[ST_EPA_GJK_ITERS];
  io.stats.epa_pairs = stats_host[ST_EPA_PAIRS];
  io.stats.epa_iterations = stats_host[ST_EPA_ITERS];
  io.stats.contacts = stats_host[ST_CONTACTS];
  io.stats.colliding_states = stats_host[ST_COLLIDING];
  io.stats.cull_iterations = stats_host[ST_CULL_ITERS];
  io.stats.exact_pairs = stats_host[ST_EXACT_PAIRS];
  io.stats.phase1_iterations = stats_host[ST_P1_ITERS];
  io.stats.phase2_iterations = stats_host[ST_P2_ITERS];
  io.stats.epa_gjk_iterations = stats_host[ST_EPA_GJK_ITERS];
  io.stats.support_vertices = stats_host[ST_SUPV] + stats_host[ST_SUPV64];
  io.stats.support_vertices_fp64 = stats_host[ST_SUPV64];
  io.stats.bvh_stack_overflows = stats_host[ST_BVH_OVERFLOW];
  io.stats.phase1_skipped = stats_host[ST_P1_SKIPPED];
  io.stats.support_vertices_epa = stats_host[ST_SUPV_EPA];
  io.stats.phase1_skip_violations = stats_host[ST_P1_VIOLATIONS];
  io.stats.closest_pruned = stats_pruned;
  if (stats_host[ST_BVH_OVERFLOW])
    return fail(T4C_ERR_UNSUPPORTED, "environment BVH traversal overflowed its stack %llu times: results incomplete",
                stats_host[ST_BVH_OVERFLOW]);
  if (io.contacts && total > io.capacity)
    return fail(T4C_ERR_CAPACITY, "contact buffer too small: %lld needed, capacity %lld", (long long)total,
                (long long)io.capacity);
  return T4C_OK;
}

// ===================================================================================================================
// C ABI
// ===================================================================================================================
static t4c_status hostRun(t4c_scene* s, BatchIO& io, const double* hq0, const double* hq1, const double* hp0,
                          const double* hp1, t4c_result* r)
{
  if (!s || !r || io.n < 0)
    return fail(T4C_ERR_INVALID, "null scene / result or negative n");
  if (!bindDevice(s))
    return fail(T4C_ERR_CUDA, "cudaSetDevice(%d): %s", s->device, plat::lastError());
  t4c_status rc = syncScene(s);
  if (rc != T4C_OK)
    return rc;
  t4stream st = s->stream;
  const int na = (int)s->img->comp.alinks.size();
  if (io.model && !hq0 && io.q0)
  {
    // joint vectors already on the device (LVS expansion kernels)
  }
  else if (io.model)
  {
    const size_t row = (size_t)io.model->n_q * sizeof(double);
    const size_t qb = (size_t)io.n * row;
    T4_TRY(s->d_q0.ensure(qb) && (!io.cast || s->d_q1.ensure(qb)), "workspace");
    io.q0 = s->d_q0.p;
    if (io.cast)
      io.q1 = s->d_q1.p;
    const int64_t piece_min = tuning().upload_piece_min;
    const bool pieces = io.n >= 4 * piece_min && ((s->copy_stream && s->q_events[3]) || !s->stream) &&
                        plat::isPinned(hq0) && (!io.cast || plat::isPinned(hq1));
    if (pieces)
    {
      // four uploads on the copy stream; FK of piece k starts when piece k is resident (runBatch)
      io.q_pieces = 4;
      io.q_piece_states = (io.n + 3) / 4;
      for (int k = 0; k < 4; ++k)
      {
        const int64_t a = (int64_t)k * io.q_piece_states, b = std::min<int64_t>(io.n, a + io.q_piece_states);
        if (a >= b)
        {
          io.q_pieces = k;
          break;
        }
        bool ok = plat::h2d((char*)s->d_q0.p + (size_t)a * row, (const char*)hq0 + (size_t)a * row,
                            (size_t)(b - a) * row, s->copy_stream);
        if (io.cast)
          ok = ok && plat::h2d((char*)s->d_q1.p + (size_t)a * row, (const char*)hq1 + (size_t)a * row,
                               (size_t)(b - a) * row, s->copy_stream);
        T4_TRY(ok && plat::record(s->q_events[k], s->copy_stream), "upload q");
      }
    }
    else
    {
      T4_TRY(plat::h2d(s->d_q0.p, hq0, qb, st), "upload q");
      if (io.cast)
        T4_TRY(plat::h2d(s->d_q1.p, hq1, qb, st), "upload q1");
    }
    io.q_layout = T4C_Q_F64_ROWMAJOR;
  }
  else
  {
    const size_t pb = (size_t)io.n * (size_t)na * 12 * sizeof(double);
    T4_TRY(s->d_pin0.ensure(pb) && plat::h2d(s->d_pin0.p, hp0, pb, st), "upload poses");
    io.poses0 = s->d_pin0.as<double>();
    if (io.cast)
    {
      T4_TRY(s->d_pin1.ensure(pb) && plat::h2d(s->d_pin1.p, hp1, pb, st), "upload poses1");
      io.poses1 = s->d_pin1.as<double>();
    }
  }
  if (r->hit)
  {
    T4_TRY(s->d_hit_full.ensure((size_t)std::max<int64_t>(io.n, 1)), "workspace");
    io.hit = s->d_hit_full.as<uint8_t>();
  }
  if (r->contacts)
  {
    T4_TRY(s->d_out.ensure((size_t)std::max<int64_t>(r->capacity, 1) * sizeof(t4c_contact)), "workspace");
    io.contacts = s->d_out.as<t4c_contact>();
    io.capacity = r->capacity;
    if (s->copy_stream || !s->stream)
      io.host_contacts = plat::isPinned(r->contacts) ? r->contacts : nullptr;
    if (r->cc_local && io.cast)
    {
      T4_TRY(s->d_out_local.ensure((size_t)std::max<int64_t>(r->capacity, 1) * 6 * sizeof(float)), "workspace");
      io.cc_local = s->d_out_local.as<float>();
    }
  }
  rc = runBatch(s, io);
  r->n_contacts = io.n_contacts;
  r->stats = io.stats;
  if (rc != T4C_OK && rc != T4C_ERR_CAPACITY)
  {
    // nothing may still be reading from or writing into the caller's buffers after the return: the progressive record
    // download AND the piecewise q upload both run on the copy stream
    if ((io.host_copied > 0 || io.q_pieces > 0) && s->copy_stream)
      plat::sync(s->copy_stream);
    plat::sync(st);
    return rc;
  }
  const int64_t ncopy = std::min<int64_t>(io.n_contacts, r->capacity);
  bool ok = true;
  if (r->hit)
    ok = ok && plat::d2h(r->hit, io.hit, (size_t)io.n, st);
  if (r->contacts && ncopy > io.host_copied)
    ok = ok && plat::d2h(r->contacts + io.host_copied, io.contacts + io.host_copied,
                         (size_t)(ncopy - io.host_copied) * sizeof(t4c_contact), st);
  if (io.host_copied > 0)
    ok = ok && plat::sync(s->copy_stream);
  if (r->contacts && r->cc_local && io.cast)
    ok = ok && plat::d2h(r->cc_local, io.cc_local, (size_t)ncopy * 6 * sizeof(float), st);
  T4_TRY(ok && plat::sync(st), "download results");
  return rc;
}

// Eigen::VectorXd::LinSpaced (core/utils.h:194-200)
static double linspaced(int64_t cnt, double low, double high, int64_t i)
{
  const int64_t size1 = cnt == 1 ? 1 : cnt - 1;
  const double step = cnt == 1 ? 0.0 : (high - low) / double(cnt - 1);
  const bool flip = std::fabs(high) < std::fabs(low);
  if (flip)
    return (i == 0) ? low : (high - double(size1 - i) * step);
  return (i == size1) ? high : (low + double(i) * step);
}

// LVS walk (tesseract_environment/core/utils.h:189-205 continuous, :328-347 discrete): calls
// visit(trajectory, waypoint step, sub-step, cnt, a, b) for every work item in the reference's visiting order;
// cnt == 0 marks an undivided step (the state / segment itself)
template <class F>
static void lvsWalk(const double* traj, int64_t n_traj, int n_wp, int n_q, double lvs, bool cast, F&& visit)
{
  for (int64_t t = 0; t < n_traj; ++t)
  {
    const double* tr = traj + (size_t)t * n_wp * n_q;
    const int last = cast ? n_wp - 1 : n_wp;
    for (int s = 0; s < last; ++s)
    {
      const double* a = tr + (size_t)s * n_q;
      const double* b = tr + (size_t)(s + 1) * n_q;
      double dist = -1;
      if (s < n_wp - 1)
      {
        double acc = 0;
        for (int k = 0; k < n_q; ++k)
          acc += (b[k] - a[k]) * (b[k] - a[k]);
        dist = std::sqrt(acc);
      }
      if (dist > 0 && dist > lvs)
      {
        const int64_t cnt = (int64_t)std::ceil(dist / lvs) + 1;
        for (int64_t i = 0; i < cnt - 1; ++i)
          visit(t, s, i, cnt, a, b);
      }
      else
        visit(t, s, (int64_t)0, (int64_t)0, a, b);
    }
  }
}

static void lvsExpand(const double* traj, int64_t n_traj, int n_wp, int n_q, double lvs, bool cast,
                      std::vector<double>& q0, std::vector<double>& q1)
{
  lvsWalk(traj, n_traj, n_wp, n_q, lvs, cast,
          [&](int64_t, int, int64_t i, int64_t cnt, const double* a, const double* b) {
            for (int k = 0; k < n_q; ++k)
            {
              q0.push_back(cnt ? linspaced(cnt, a[k], b[k], i) : a[k]);
              if (cast)
                q1.push_back(cnt ? linspaced(cnt, a[k], b[k], i + 1) : b[k]);
            }
          });
}

#ifndef T4C_HOSTSIM
template <class T, class K>
static t4c_status measurePeak(int32_t device, double* tflops, K kernel, int iters)
{
  if (device < 0 || device >= plat::deviceCount())
    return fail(T4C_ERR_CUDA, "no usable CUDA device %d", device);
  plat::setDevice(device);
  g_sm_count = 0;
  const int blocks = smCount() * 8, threads = 256;
  T* out = (T*)plat::alloc((size_t)blocks * threads * sizeof(T));
  if (!out)
    return fail(T4C_ERR_NOMEM, "alloc");
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  double best = 0;
  for (int rep = 0; rep < 5; ++rep)
  {
    cudaEventRecord(e0);
    kernel<<<blocks, threads>>>(out, iters);
    ++g_launches;
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess)
      break;
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double fl = 2.0 * 8.0 * (double)iters * (double)blocks * threads;
    if (ms > 0)
      best = std::max(best, fl / (ms * 1e-3) / 1e12);
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  plat::release(out);
  *tflops = best;
  return best > 0 ? T4C_OK : fail(T4C_ERR_CUDA, "FMA peak kernel failed: %s", cudaGetErrorString(cudaGetLastError()));
}
#endif

extern "C" {
const char* t4c_last_error(void) { return g_err.c_str(); }
int32_t t4c_abi_version(void) { return T4C_ABI_VERSION; }
const char* t4c_build_info(void)
{
#ifdef T4C_HOSTSIM
  return "hostsim";
#else
  return "cuda sm_100a";
#endif
}
int32_t t4c_device_count(void) { return plat::deviceCount(); }
uint64_t t4c_kernel_launch_count(void) { return g_launches.load(); }
void t4c_profile_enable(int32_t on)
{
#ifndef T4C_HOSTSIM
  g_prof_on = on != 0;
#else
  (void)on;
#endif
}
void t4c_profile_reset(void)
{
#ifndef T4C_HOSTSIM
  profResolve();
  std::memset(g_prof_ms, 0, sizeof(g_prof_ms));
  std::memset(g_prof_n, 0, sizeof(g_prof_n));
#endif
}
int32_t t4c_profile_get(double* ms, uint64_t* launches, int32_t capacity)
{
#ifndef T4C_HOSTSIM
  profResolve();
  for (int i = 0; i < PK_COUNT && i < capacity; ++i)
  {
    ms[i] = g_prof_ms[i];
    launches[i] = g_prof_n[i];
  }
  return PK_COUNT;
#else
  (void)ms, (void)launches, (void)capacity;
  return 0;
#endif
}

t4c_status t4c_scene_create(const t4c_scene_desc* d, int32_t device, t4c_scene** out)
{
  if (!d || !out)
    return fail(T4C_ERR_INVALID, "null argument");
  if (d->n_links < 0 || d->n_shapes < 0 || d->n_vertices < 0 || d->n_acm_pairs < 0 || d->n_margin_pairs < 0)
    return fail(T4C_ERR_INVALID, "negative count in scene description");
  if (device < 0 || device >= plat::deviceCount())
    return fail(T4C_ERR_CUDA, "no usable CUDA device %d (%d visible); this library has no CPU path", device,
                plat::deviceCount());
  std::unique_ptr<t4c_scene> s(new t4c_scene());
  s->device = device;
  if (!plat::setDevice(device))
    return fail(T4C_ERR_CUDA, "cudaSetDevice(%d): %s", device, plat::lastError());
  HostScene& H = s->host;
  H.links.resize((size_t)d->n_links);
  for (int i = 0; i < d->n_links; ++i)
  {
    H.links[i].name = d->links[i].name ? d->links[i].name : "";
    H.links[i].type_id = d->links[i].type_id;
    H.links[i].enabled = d->links[i].enabled != 0;
    H.links[i].active = d->links[i].active != 0;
    std::memcpy(H.links[i].pose, d->links[i].pose, sizeof(H.links[i].pose));
  }
  H.geom->shapes.assign(d->shapes, d->shapes + d->n_shapes);
  for (auto& sh : H.geom->shapes)
    if (sh.type < T4C_SPHERE || sh.type > T4C_TRIANGLE || sh.link < 0 || sh.link >= d->n_links)
      return fail(T4C_ERR_INVALID, "bad shape record (type %d, link %d)", sh.type, sh.link);
  H.geom->vertices.assign(d->vertices, d->vertices + 3 * (size_t)d->n_vertices);
  for (int i = 0; i < d->n_acm_pairs; ++i)
    H.acm.insert(hostPairKey(d->acm_pairs[2 * i], d->acm_pairs[2 * i + 1]));
  H.default_margin = d->default_margin;
  for (int i = 0; i < d->n_margin_pairs; ++i)
    H.pair_margins[hostPairKey(d->margin_pairs[2 * i], d->margin_pairs[2 * i + 1])] = d->margin_values[i];
  s->stream = plat::newStream();
  s->copy_stream = plat::newStream();
  s->aux_stream = plat::newStream();
  s->epa_stream = plat::newStream();
  for (auto& e : s->q_events)
    e = plat::newEvent();
  t4c_status rc = syncScene(s.get());
  if (rc != T4C_OK)
    return rc;
  *out = s.release();
  return T4C_OK;
}

t4c_status t4c_scene_clone(const t4c_scene* src, t4c_scene** out)
{
  if (!src || !out)
    return fail(T4C_ERR_INVALID, "null argument");
  std::unique_ptr<t4c_scene> s(new t4c_scene());
  s->device = src->device;
  s->host = src->host;  // links, ACM, margins by value; geometry (shapes, vertices) by reference
  if (!plat::setDevice(s->device))
    return fail(T4C_ERR_CUDA, "cudaSetDevice(%d): %s", s->device, plat::lastError());
  s->stream = plat::newStream();
  s->copy_stream = plat::newStream();
  s->aux_stream = plat::newStream();
  s->epa_stream = plat::newStream();
  for (auto& e : s->q_events)
    e = plat::newEvent();
  if (src->uploaded == src->rev)
  {
    // the source's compiled image (host tables + device buffers) is current: share it, no compile, no upload
    s->img = src->img;
    s->rev = s->uploaded = 1;
  }
  else
  {
    t4c_status rc = syncScene(s.get());
    if (rc != T4C_OK)
      return rc;
  }
  s->chunk_hint = src->chunk_hint;
  *out = s.release();
  return T4C_OK;
}

void t4c_scene_destroy(t4c_scene* s)
{
  if (!s)
    return;
  plat::setDevice(s->device);
  delete s;
}

t4c_status t4c_scene_set_active(t4c_scene* s, const int32_t* flags)
{
  if (!s || !flags)
    return fail(T4C_ERR_INVALID, "null argument");
  bool changed = false;
  for (size_t i = 0; i < s->host.links.size(); ++i)
  {
    changed = changed || s->host.links[i].active != (flags[i] != 0);
    s->host.links[i].active = flags[i] != 0;
  }
  if (changed)  // an unchanged setting costs nothing: no recompile at the next check
    ++s->rev;
  return T4C_OK;
}
t4c_status t4c_scene_set_enabled(t4c_scene* s, int32_t link, int32_t enabled)
{
  if (!s || link < 0 || link >= (int)s->host.links.size())
    return fail(T4C_ERR_INVALID, "unknown link %d", link);
  if (s->host.links[link].enabled == (enabled != 0))
    return T4C_OK;
  s->host.links[link].enabled = enabled != 0;
  ++s->rev;
  return T4C_OK;
}
t4c_status t4c_scene_set_acm(t4c_scene* s, int32_t n, const int32_t* pairs)
{
  if (!s || n < 0 || (n > 0 && !pairs))
    return fail(T4C_ERR_INVALID, "bad argument");
  std::unordered_set<uint64_t> acm;
  for (int i = 0; i < n; ++i)
    acm.insert(hostPairKey(pairs[2 * i], pairs[2 * i + 1]));
  if (acm == s->host.acm)
    return T4C_OK;
  s->host.acm.swap(acm);
  ++s->rev;
  return T4C_OK;
}
t4c_status t4c_scene_set_margins(t4c_scene* s, double def, int32_t n, const int32_t* pairs, const double* values)
{
  if (!s || n < 0 || (n > 0 && (!pairs || !values)))
    return fail(T4C_ERR_INVALID, "bad argument");
  std::unordered_map<uint64_t, double> pm;
  for (int i = 0; i < n; ++i)
    pm[hostPairKey(pairs[2 * i], pairs[2 * i + 1])] = values[i];
  if (def == s->host.default_margin && pm == s->host.pair_margins)
    return T4C_OK;
  s->host.default_margin = def;
  s->host.pair_margins.swap(pm);
  ++s->rev;
  return T4C_OK;
}
t4c_status t4c_scene_set_link_poses(t4c_scene* s, int32_t n, const int32_t* links, const double* poses12)
{
  if (!s || n < 0 || (n > 0 && (!links || !poses12)))
    return fail(T4C_ERR_INVALID, "bad argument");
  std::vector<char> moved(s->host.links.size(), 0);
  bool any = false, only_static = true;
  for (int i = 0; i < n; ++i)
  {
    const int link = links[i];
    if (link < 0 || link >= (int)s->host.links.size())
      return fail(T4C_ERR_INVALID, "unknown link %d", link);
    HostLink& L = s->host.links[link];
    if (std::memcmp(L.pose, poses12 + 12 * (size_t)i, 12 * sizeof(double)) == 0)
      continue;  // unchanged: nothing to do
    std::memcpy(L.pose, poses12 + 12 * (size_t)i, 12 * sizeof(double));
    moved[link] = 1;
    any = true;
    only_static = only_static && !L.active && L.enabled;
  }
  if (!any)
    return T4C_OK;
  if (s->uploaded == s->rev && only_static && bindDevice(s))
  {
    // Static objects moved and nothing else changed: REFIT the environment BVH (boxes follow, topology stays) instead of
    // compiling the scene again.  The image may be shared with clones, so the refit goes into a copy.
    std::shared_ptr<SceneImage> img = std::make_shared<SceneImage>();
    img->comp = s->img->comp;
    refitStaticLinks(s->host, img->comp, moved);
    if (!uploadImage(*img, s->stream))
      return fail(T4C_ERR_CUDA, "scene upload failed: %s", plat::lastError());
    s->img = img;
    s->uploaded = ++s->rev;
    return T4C_OK;
  }
  ++s->rev;
  return T4C_OK;
}

t4c_status t4c_scene_set_link_pose(t4c_scene* s, int32_t link, const double* pose12)
{
  if (!s || !pose12)
    return fail(T4C_ERR_INVALID, "null argument");
  return t4c_scene_set_link_poses(s, 1, &link, pose12);
}

// DiscreteContactManager::addCollisionObject / removeCollisionObject on an existing scene
// (discrete_contact_manager.h:75-98; the Environment's add / remove link commands end here, environment.cpp:613-640).
// The geometry may be shared with clones: it is copied before the first edit.  The next check compiles the scene again.
static HostGeometry& ownGeometry(t4c_scene* s)
{
  if (s->host.geom.use_count() > 1)
    s->host.geom = std::make_shared<HostGeometry>(*s->host.geom);
  return *s->host.geom;
}

t4c_status t4c_scene_add_object(t4c_scene* s, const t4c_link* link, int32_t n_shapes, const t4c_shape* shapes,
                                int32_t n_vertices, const double* vertices, int32_t* link_index)
{
  if (!s || !link || n_shapes < 0 || (n_shapes > 0 && !shapes) || n_vertices < 0 || (n_vertices > 0 && !vertices))
    return fail(T4C_ERR_INVALID, "bad argument");
  for (int i = 0; i < n_shapes; ++i)
  {
    const t4c_shape& sh = shapes[i];
    const int need = sh.type == T4C_TRIANGLE ? 3 : sh.vertex_count;
    if (sh.type < T4C_SPHERE || sh.type > T4C_TRIANGLE ||
        ((sh.type == T4C_HULL || sh.type == T4C_TRIANGLE) && (sh.vertex_offset < 0 || sh.vertex_offset + need > n_vertices)))
      return fail(T4C_ERR_INVALID, "bad shape record %d", i);
  }
  HostGeometry& G = ownGeometry(s);
  const std::string name = link->name ? link->name : "";
  int li = -1;
  for (size_t i = 0; i < s->host.links.size(); ++i)
    if (!name.empty() && s->host.links[i].name == name)
      li = (int)i;
  if (li >= 0)
  {
    // re-adding an existing name replaces the object (bullet_discrete_bvh_manager.cpp:113-114)
    G.shapes.erase(std::remove_if(G.shapes.begin(), G.shapes.end(), [li](const t4c_shape& x) { return x.link == li; }),
                   G.shapes.end());
  }
  else
  {
    li = (int)s->host.links.size();
    s->host.links.emplace_back();
  }
  HostLink& L = s->host.links[li];
  L.name = name;
  L.type_id = link->type_id;
  L.enabled = link->enabled != 0;
  L.active = link->active != 0;
  std::memcpy(L.pose, link->pose, sizeof(L.pose));
  const int off = (int)(G.vertices.size() / 3);
  G.vertices.insert(G.vertices.end(), vertices, vertices + 3 * (size_t)n_vertices);
  for (int i = 0; i < n_shapes; ++i)
  {
    t4c_shape sh = shapes[i];
    sh.link = li;
    if (sh.type == T4C_HULL || sh.type == T4C_TRIANGLE)
      sh.vertex_offset += off;
    G.shapes.push_back(sh);
  }
  if (link_index)
    *link_index = li;
  ++s->rev;
  return T4C_OK;
}

t4c_status t4c_scene_remove_object(t4c_scene* s, int32_t link)
{
  if (!s || link < 0 || link >= (int)s->host.links.size())
    return fail(T4C_ERR_INVALID, "unknown link %d", link);
  HostGeometry& G = ownGeometry(s);
  G.shapes.erase(std::remove_if(G.shapes.begin(), G.shapes.end(), [link](const t4c_shape& x) { return x.link == link; }),
                 G.shapes.end());
  // link indices of the other objects must not move (contact records and the ACM refer to them): the slot stays,
  // shapeless and disabled
  s->host.links[link].enabled = false;
  s->host.links[link].active = false;
  ++s->rev;
  return T4C_OK;
}
int32_t t4c_scene_num_active(const t4c_scene* s)
{
  if (!s)
    return 0;
  int n = 0;
  for (auto& l : s->host.links)
    n += l.active ? 1 : 0;
  return n;
}

t4c_status t4c_fk_model_create(const t4c_fk_model_desc* d, const t4c_scene*, t4c_fk_model** out)
{
  if (!d || !out || d->n_joints < 0 || d->n_q < 0)
    return fail(T4C_ERR_INVALID, "bad FK model description");
  if (d->n_joints > T4_MAX_JOINTS)
    return fail(T4C_ERR_UNSUPPORTED, "FK model has %d joints (max %d)", d->n_joints, T4_MAX_JOINTS);
  std::unique_ptr<t4c_fk_model> m(new t4c_fk_model());
  m->n_links = d->n_links;
  m->n_q = d->n_q;
  m->root_link = d->root_link;
  m->joints.assign(d->joints, d->joints + d->n_joints);
  std::vector<char> placed((size_t)std::max(d->n_links, 1), 0);
  for (auto& j : m->joints)
  {
    if (j.parent_link < 0 || j.parent_link >= d->n_links || j.child_link < 0 || j.child_link >= d->n_links)
      return fail(T4C_ERR_INVALID, "joint refers to a link outside [0,%d)", d->n_links);
    if (placed[j.child_link])
      return fail(T4C_ERR_INVALID, "link %d has two parent joints", j.child_link);
    placed[j.child_link] = 1;
  }
  std::fill(placed.begin(), placed.end(), 0);
  for (auto& j : m->joints)
  {
    // parent-before-child: the parent must be a root or already produced
    bool parent_is_child_later = false;
    for (auto& k : m->joints)
      if (&k > &j && k.child_link == j.parent_link)
        parent_is_child_later = true;
    if (parent_is_child_later)
      return fail(T4C_ERR_INVALID, "joints are not ordered parent-before-child");
  }
  *out = m.release();
  return T4C_OK;
}
void t4c_fk_model_destroy(t4c_fk_model* m) { delete m; }
}  // extern "C"

extern "C" {
t4c_status t4c_fk(const t4c_fk_model* m, const t4c_scene* scene_c, const double* q, int64_t n, double* link_poses)
{
  t4c_scene* s = const_cast<t4c_scene*>(scene_c);
  if (!m || !s || !q || !link_poses || n < 0)
    return fail(T4C_ERR_INVALID, "null argument");
  t4c_status rc = syncScene(s);
  if (rc != T4C_OK)
    return rc;
  FkCache* fc = nullptr;
  rc = syncFk(s, m, &fc);
  if (rc != T4C_OK)
    return rc;
  t4stream st = s->stream;
  const int nl = (int)s->host.links.size();
  KParams P;
  fillSceneParams(s, P);
  fillFkParams(m, fc, P);
  const int64_t CH = chunkStates();
  for (int64_t base = 0; base < n; base += CH)
  {
    const int64_t nc = std::min<int64_t>(CH, n - base);
    const size_t qb = (size_t)nc * (size_t)m->n_q * sizeof(double);
    const size_t ob = (size_t)nc * (size_t)nl * 12 * sizeof(double);
    T4_TRY(s->d_q0.ensure(qb) && s->d_fk_out.ensure(ob) && plat::h2d(s->d_q0.p, q + (size_t)base * m->n_q, qb, st),
           "FK workspace");
    P.n = nc;
    P.q0 = s->d_q0.p;
    P.q_layout = T4C_Q_F64_ROWMAJOR;
    P.q_stride_n = nc;
    P.cast = 0;
    P.fk_out = s->d_fk_out.as<double>();
    launchFk(P, 0, nc, st);
    T4_TRY(launchOk() && plat::d2h(link_poses + (size_t)base * nl * 12, P.fk_out, ob, st) && plat::sync(st), "FK");
  }
  return T4C_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// t4c_fk_noalloc: t4c_fk for callers that may not allocate device memory (code under CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// t4c_fk stages q and the poses in growable device buffers (s->d_q0, s->d_fk_out: cudaMalloc when they
// grow); this entry stages them in ONE fixed array of device memory that is part of the library image:
//   - every GPU that loads the library holds T4C_NOALLOC_SCRATCH_BYTES (8 MiB) of it until the process exits;
//   - n * (n_q * 8 + n_links * 96) bytes (+ 256 of padding) have to fit, larger requests are refused (T4C_ERR_CAPACITY);
//   - the scene image and the FK model must already be resident (any earlier t4c_fk / check call after the last edit):
//     this entry never uploads them, it refuses instead (T4C_ERR_INVALID);
//   - every call on a device shares the array and finds the previous call's bytes in it.  Nothing has to start at
//     zero: the upload overwrites the q region and the kernel writes every link of every state (fkStoreAllLinks)
//     before the download reads it.  Calls that overlap on one GPU would race on it; inside one process they cannot
//     (the entry is synchronous and holds a per-device lock from upload to download), two PROCESSES have separate
//     copies of the array.
// In host code `g_noalloc_scratch` is a stand-in for the symbol, not a device address: the pointer the kernel gets
// comes from cudaGetSymbolAddress, taken once per device with that device current.
// ---------------------------------------------------------------------------------------------------------------------
#define T4C_NOALLOC_SCRATCH_BYTES (8u << 20)
#define T4C_NOALLOC_MAX_DEVICES 64
#ifndef T4C_HOSTSIM
static __device__ __align__(256) unsigned char g_noalloc_scratch[T4C_NOALLOC_SCRATCH_BYTES];
#else
alignas(256) static unsigned char g_noalloc_scratch[T4C_NOALLOC_SCRATCH_BYTES];
#endif
static std::mutex g_noalloc_lock[T4C_NOALLOC_MAX_DEVICES];
static void* g_noalloc_addr[T4C_NOALLOC_MAX_DEVICES] = {};

// device address of the scratch array on `device` (current); looked up once per device
static t4c_status noallocScratch(int device, unsigned char** out)
{
  if (device < 0 || device >= T4C_NOALLOC_MAX_DEVICES)
    return fail(T4C_ERR_INVALID, "device %d outside [0, %d)", device, T4C_NOALLOC_MAX_DEVICES);
  if (!g_noalloc_addr[device])
  {
#ifndef T4C_HOSTSIM
    void* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
    if (e != cudaSuccess || !ptr)
    {
      cudaGetLastError();
      return fail(T4C_ERR_CUDA, "cudaGetSymbolAddress(scratch) on device %d: %s", device, cudaGetErrorString(e));
    }
    g_noalloc_addr[device] = ptr;
#else
    g_noalloc_addr[device] = g_noalloc_scratch;
#endif
  }
  *out = static_cast<unsigned char*>(g_noalloc_addr[device]);
  return T4C_OK;
}

t4c_status t4c_fk_noalloc(const t4c_fk_model* m, const t4c_scene* scene_c, const double* q, int64_t n, double* link_poses)
{
  t4c_scene* s = const_cast<t4c_scene*>(scene_c);
  if (!m || !s || !q || !link_poses || n < 0)
    return fail(T4C_ERR_INVALID, "null argument");
  if (!bindDevice(s))
    return fail(T4C_ERR_CUDA, "cudaSetDevice(%d): %s", s->device, plat::lastError());
  const auto it = s->fk.find(m);
  if (s->uploaded != s->rev || !s->img || it == s->fk.end() || !it->second || it->second->scene_rev != s->rev)
    return fail(T4C_ERR_INVALID, "t4c_fk_noalloc never uploads: the scene or the FK model is not resident on the device "
                                 "(run t4c_fk or a check once after creating / editing them)");
  const int nl = (int)s->host.links.size();
  const size_t qb = (size_t)n * (size_t)m->n_q * sizeof(double);
  const size_t q_region = (qb + 255) & ~(size_t)255;
  const size_t ob = (size_t)n * (size_t)nl * 12 * sizeof(double);
  if (q_region + ob > T4C_NOALLOC_SCRATCH_BYTES)
    return fail(T4C_ERR_CAPACITY, "t4c_fk_noalloc: %lld states x (%d joint values + %d link poses) need %zu bytes of "
                                  "scratch, the fixed array holds %u (at most %lld states of this model)",
                (long long)n, m->n_q, nl, q_region + ob, (unsigned)T4C_NOALLOC_SCRATCH_BYTES,
                (long long)((T4C_NOALLOC_SCRATCH_BYTES - 256) / ((size_t)m->n_q * 8 + (size_t)nl * 96)));
  if (n == 0)
    return T4C_OK;
  std::lock_guard<std::mutex> guard(g_noalloc_lock[s->device < T4C_NOALLOC_MAX_DEVICES && s->device >= 0 ? s->device : 0]);
  unsigned char* scratch = nullptr;
  const t4c_status rc = noallocScratch(s->device, &scratch);
  if (rc != T4C_OK)
    return rc;
  t4stream st = s->stream;
  KParams P;
  fillSceneParams(s, P);
  fillFkParams(m, it->second.get(), P);
  T4_TRY(plat::h2d(scratch, q, qb, st), "FK upload");
  P.n = n;
  P.q0 = scratch;
  P.q_layout = T4C_Q_F64_ROWMAJOR;
  P.q_stride_n = n;
  P.cast = 0;
  P.fk_out = reinterpret_cast<double*>(scratch + q_region);
  launchFk(P, 0, n, st);
  T4_TRY(launchOk() && plat::d2h(link_poses, P.fk_out, ob, st) && plat::sync(st), "FK");
  return T4C_OK;
}

static void launchJacobian(const JacParams& Q, int64_t n, bool grad, t4stream st)
{
  if (n <= 0)
    return;
#ifndef T4C_HOSTSIM
  const int grid = gridFor(n, 128, 8);
  if (Q.P.n_joints <= 16)
  {
    if (grad)
      contactGradKernel<16><<<grid, 128, 0, st>>>(Q, n);
    else
      jacobianKernel<16><<<grid, 128, 0, st>>>(Q, n);
  }
  else if (grad)
    contactGradKernel<T4_MAX_JOINTS><<<grid, 128, 0, st>>>(Q, n);
  else
    jacobianKernel<T4_MAX_JOINTS><<<grid, 128, 0, st>>>(Q, n);
#else
  (void)st;
  for (int64_t i = 0; i < n; ++i)
    if (grad)
      contactGradBody<T4_MAX_JOINTS>(i, Q);
    else
      jacobianBody<T4_MAX_JOINTS>(i, Q);
#endif
  ++g_launches;
}

t4c_status t4c_jacobian(const t4c_fk_model* m, const t4c_scene* scene_c, const double* q, int64_t n, int32_t link,
                        const double* point_local, double* J)
{
  t4c_scene* s = const_cast<t4c_scene*>(scene_c);
  if (!m || !s || !q || !J || n < 0 || link < 0 || link >= (int)s->host.links.size())
    return fail(T4C_ERR_INVALID, "bad argument");
  t4c_status rc = syncScene(s);
  if (rc != T4C_OK)
    return rc;
  FkCache* fc = nullptr;
  rc = syncFk(s, m, &fc);
  if (rc != T4C_OK)
    return rc;
  t4stream st = s->stream;
  JacParams Q;
  std::memset(&Q, 0, sizeof(Q));
  fillSceneParams(s, Q.P);
  fillFkParams(m, fc, Q.P);
  int lj = -1;
  for (size_t j = 0; j < m->joints.size(); ++j)
    if (m->joints[j].child_link == link)
      lj = (int)j;
  Q.link_joint = lj;
  Q.has_point = point_local ? 1 : 0;
  if (point_local)
    std::memcpy(Q.point, point_local, 3 * sizeof(double));
  const size_t qb = (size_t)n * (size_t)m->n_q * sizeof(double), jb = (size_t)n * 6 * (size_t)m->n_q * sizeof(double);
  T4_TRY(s->d_q0.ensure(std::max<size_t>(qb, 8)) && s->d_fk_out.ensure(std::max<size_t>(jb, 8)) &&
             plat::h2d(s->d_q0.p, q, qb, st),
         "Jacobian workspace");
  Q.P.n = n, Q.P.q0 = s->d_q0.p, Q.P.q_layout = T4C_Q_F64_ROWMAJOR, Q.P.q_stride_n = n;
  Q.J = s->d_fk_out.as<double>();
  launchJacobian(Q, n, false, st);
  T4_TRY(launchOk() && plat::d2h(J, Q.J, jb, st) && plat::sync(st), "Jacobian");
  return T4C_OK;
}

static t4c_status contactGradients(t4c_scene* s, const t4c_fk_model* m, const void* q_dev, int32_t q_layout, int64_t n,
                                   const t4c_contact* con_dev, int64_t nc, double* grad_dev)
{
  t4c_status rc = syncScene(s);
  if (rc != T4C_OK)
    return rc;
  FkCache* fc = nullptr;
  rc = syncFk(s, m, &fc);
  if (rc != T4C_OK)
    return rc;
  JacParams Q;
  std::memset(&Q, 0, sizeof(Q));
  fillSceneParams(s, Q.P);
  fillFkParams(m, fc, Q.P);
  Q.P.n = n, Q.P.q0 = q_dev, Q.P.q_layout = q_layout, Q.P.q_stride_n = n;
  Q.contacts = con_dev, Q.n_contacts = nc, Q.grad = grad_dev;
  Q.link_to_joint = fc->link_joint.as<int32_t>();
  launchJacobian(Q, nc, true, s->stream);
  T4_TRY(launchOk(), "contact gradients");
  return T4C_OK;
}

t4c_status t4c_contact_gradients_device(t4c_scene* s, const t4c_fk_model* m, const void* q, int32_t q_layout, int64_t n,
                                        const t4c_contact* contacts, int64_t n_contacts, double* grad)
{
  if (!s || !m || !q || n < 0 || n_contacts < 0 || (n_contacts > 0 && (!contacts || !grad)))
    return fail(T4C_ERR_INVALID, "bad argument");
  const t4c_status rc = contactGradients(s, m, q, q_layout, n, contacts, n_contacts, grad);
  if (rc != T4C_OK)
    return rc;
  T4_TRY(plat::sync(s->stream), "contact gradients");
  return T4C_OK;
}

t4c_status t4c_contact_gradients(t4c_scene* s, const t4c_fk_model* m, const double* q, int64_t n,
                                 const t4c_contact* contacts, int64_t n_contacts, double* grad)
{
  if (!s || !m || !q || n < 0 || n_contacts < 0 || (n_contacts > 0 && (!contacts || !grad)))
    return fail(T4C_ERR_INVALID, "bad argument");
  if (!bindDevice(s))
    return fail(T4C_ERR_CUDA, "cudaSetDevice(%d): %s", s->device, plat::lastError());
  for (int64_t i = 0; i < n_contacts; ++i)
    if (contacts[i].state < 0 || contacts[i].state >= n || contacts[i].link[0] < 0 || contacts[i].link[1] < 0 ||
        contacts[i].link[0] >= (int)s->host.links.size() || contacts[i].link[1] >= (int)s->host.links.size())
      return fail(T4C_ERR_INVALID, "contact %lld refers to a state or link outside the batch", (long long)i);
  t4stream st = s->stream;
  const size_t qb = (size_t)n * (size_t)m->n_q * sizeof(double), cb = (size_t)n_contacts * sizeof(t4c_contact);
  const size_t gb = (size_t)n_contacts * (size_t)m->n_q * sizeof(double);
  T4_TRY(s->d_q0.ensure(std::max<size_t>(qb, 8)) && s->d_out.ensure(std::max<size_t>(cb, 8)) &&
             s->d_fk_out.ensure(std::max<size_t>(gb, 8)) && plat::h2d(s->d_q0.p, q, qb, st) &&
             plat::h2d(s->d_out.p, contacts, cb, st),
         "gradient workspace");
  const t4c_status rc = contactGradients(s, m, s->d_q0.p, T4C_Q_F64_ROWMAJOR, n, s->d_out.as<t4c_contact>(), n_contacts,
                                         s->d_fk_out.as<double>());
  if (rc != T4C_OK)
    return rc;
  T4_TRY(plat::d2h(grad, s->d_fk_out.p, gb, st) && plat::sync(st), "contact gradients");
  return T4C_OK;
}

t4c_status t4c_check_discrete(t4c_scene* s, const t4c_fk_model* m, const double* q, int64_t n, const t4c_request* req,
                              t4c_result* r)
{
  if (!s || !m || !req || !r || (n > 0 && !q))
    return fail(T4C_ERR_INVALID, "null argument");
  BatchIO io;
  io.n = n;
  io.model = m;
  io.req = *req;
  return hostRun(s, io, q, nullptr, nullptr, nullptr, r);
}

t4c_status t4c_check_link_poses_discrete(t4c_scene* s, const double* poses, int64_t n, const t4c_request* req,
                                         t4c_result* r)
{
  if (!s || !req || !r)
    return fail(T4C_ERR_INVALID, "null argument");
  BatchIO io;
  io.n = n;
  io.req = *req;
  return hostRun(s, io, nullptr, nullptr, poses, nullptr, r);
}

t4c_status t4c_check_cast(t4c_scene* s, const t4c_fk_model* m, const double* q0, const double* q1, int64_t n,
                          const t4c_request* req, t4c_result* r)
{
  if (!s || !m || !req || !r || (n > 0 && (!q0 || !q1)))
    return fail(T4C_ERR_INVALID, "null argument");
  BatchIO io;
  io.n = n;
  io.cast = true;
  io.model = m;
  io.req = *req;
  return hostRun(s, io, q0, q1, nullptr, nullptr, r);
}

t4c_status t4c_check_link_poses_cast(t4c_scene* s, const double* p0, const double* p1, int64_t n,
                                     const t4c_request* req, t4c_result* r)
{
  if (!s || !req || !r)
    return fail(T4C_ERR_INVALID, "null argument");
  BatchIO io;
  io.n = n;
  io.cast = true;
  io.req = *req;
  return hostRun(s, io, nullptr, nullptr, p0, p1, r);
}

static t4c_status checkTrajectories(t4c_scene* s, const t4c_fk_model* m, const double* traj, int64_t n_traj,
                                    int32_t n_wp, double lvs, const t4c_request* req, t4c_result* r, int64_t* n_items,
                                    bool cast)
{
  if (!s || !m || !req || !r || !traj || n_traj < 0 || n_wp < (cast ? 2 : 1) || !(lvs > 0))
    return fail(T4C_ERR_INVALID, "bad trajectory arguments");
#ifdef T4C_HOSTSIM
  std::vector<double> q0, q1;
  lvsExpand(traj, n_traj, n_wp, m->n_q, lvs, cast, q0, q1);
  const int64_t n = (int64_t)(q0.size() / (size_t)std::max(m->n_q, 1));
  if (n_items)
    *n_items = n;
  BatchIO io;
  io.n = n;
  io.cast = cast;
  io.model = m;
  io.req = *req;
  return hostRun(s, io, q0.data(), cast ? q1.data() : nullptr, nullptr, nullptr, r);
#else
  // LVS subdivision on the device: count per waypoint step -> exclusive scan -> expansion into q0 / q1
  if (!bindDevice(s))
    return fail(T4C_ERR_CUDA, "cudaSetDevice(%d): %s", s->device, plat::lastError());
  t4stream st = s->stream;
  LvsParams L;
  L.n_traj = n_traj, L.n_wp = n_wp, L.n_q = m->n_q, L.cast = cast ? 1 : 0, L.lvs = lvs;
  const int64_t n_steps = n_traj * (int64_t)L.steps();
  int64_t n = 0;
  if (n_steps > 0)
  {
    const size_t tb = (size_t)n_traj * n_wp * m->n_q * sizeof(double);
    T4_TRY(s->d_traj.ensure(tb) && plat::h2d(s->d_traj.p, traj, tb, st) &&
               s->d_lvs_counts.ensure((size_t)n_steps * sizeof(uint64_t)) &&
               s->d_lvs_offsets.ensure((size_t)n_steps * sizeof(uint64_t)),
           "LVS workspace");
    L.traj = s->d_traj.as<double>();
    L.counts = s->d_lvs_counts.as<uint64_t>();
    L.offsets = s->d_lvs_offsets.as<uint64_t>();
    L.q0 = L.q1 = nullptr;
    t4LaunchLvsCount(L, n_steps, gridFor(n_steps, 256, 8), st);
    ++g_launches;
    size_t tmp_bytes = 0;
    T4_TRY(plat::chk(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, L.counts, s->d_lvs_offsets.as<uint64_t>(), n_steps, st)) &&
               s->d_lvs_tmp.ensure(tmp_bytes) &&
               plat::chk(cub::DeviceScan::ExclusiveSum(s->d_lvs_tmp.p, tmp_bytes, L.counts,
                                                       s->d_lvs_offsets.as<uint64_t>(), n_steps, st)),
           "LVS scan");
    uint64_t last[2] = { 0, 0 };
    T4_TRY(plat::d2h(&last[0], L.offsets + (n_steps - 1), sizeof(uint64_t), st) &&
               plat::d2h(&last[1], L.counts + (n_steps - 1), sizeof(uint64_t), st) && plat::sync(st),
           "LVS count");
    n = (int64_t)(last[0] + last[1]);
    const size_t qb = (size_t)n * m->n_q * sizeof(double);
    T4_TRY(s->d_q0.ensure(qb) && (!cast || s->d_q1.ensure(qb)), "LVS workspace");
    L.q0 = s->d_q0.as<double>();
    L.q1 = cast ? s->d_q1.as<double>() : nullptr;
    t4LaunchLvsExpand(L, n_steps, gridFor(n_steps, 256, 8), st);
    ++g_launches;
    T4_TRY(launchOk(), "LVS expansion");
  }
  if (n_items)
    *n_items = n;
  BatchIO io;
  io.n = n;
  io.cast = cast;
  io.model = m;
  io.req = *req;
  if (n == 0)
    return hostRun(s, io, traj, traj, nullptr, nullptr, r);
  io.q0 = s->d_q0.p;
  io.q1 = cast ? s->d_q1.p : nullptr;
  io.q_layout = T4C_Q_F64_ROWMAJOR;
  return hostRun(s, io, nullptr, nullptr, nullptr, nullptr, r);
#endif
}

t4c_status t4c_lvs_count(const double* traj, int64_t n_traj, int32_t n_wp, int32_t n_q, double lvs, int32_t cast,
                         int64_t* n_items)
{
  return t4c_lvs_items(traj, n_traj, n_wp, n_q, lvs, cast, 0, nullptr, nullptr, nullptr, n_items);
}

t4c_status t4c_lvs_items(const double* traj, int64_t n_traj, int32_t n_wp, int32_t n_q, double lvs, int32_t cast,
                         int64_t capacity, int32_t* item_traj, int32_t* item_step, int32_t* item_sub, int64_t* n_items)
{
  if (!traj || !n_items || n_q <= 0 || n_wp <= 0 || n_traj < 0 || !(lvs > 0))
    return fail(T4C_ERR_INVALID, "bad arguments");
  int64_t n = 0;
  lvsWalk(traj, n_traj, n_wp, n_q, lvs, cast != 0,
          [&](int64_t t, int s, int64_t i, int64_t, const double*, const double*) {
            if (n < capacity)
            {
              if (item_traj)
                item_traj[n] = (int32_t)t;
              if (item_step)
                item_step[n] = s;
              if (item_sub)
                item_sub[n] = (int32_t)i;
            }
            ++n;
          });
  *n_items = n;
  return T4C_OK;
}

t4c_status t4c_lvs_item_states(const double* traj, int64_t n_traj, int32_t n_wp, int32_t n_q, double lvs, int32_t cast,
                               int64_t n_sel, const int64_t* items, double* q0, double* q1)
{
  if (!traj || n_q <= 0 || n_wp <= 0 || n_traj < 0 || !(lvs > 0) || n_sel < 0 || (n_sel > 0 && (!items || !q0)) ||
      (cast && n_sel > 0 && !q1))
    return fail(T4C_ERR_INVALID, "bad arguments");
  for (int64_t k = 1; k < n_sel; ++k)
    if (items[k] <= items[k - 1])
      return fail(T4C_ERR_INVALID, "item indices must be strictly ascending");
  int64_t n = 0, k = 0;
  lvsWalk(traj, n_traj, n_wp, n_q, lvs, cast != 0,
          [&](int64_t, int, int64_t i, int64_t cnt, const double* a, const double* b) {
            if (k < n_sel && items[k] == n)
            {
              for (int c = 0; c < n_q; ++c)
              {
                q0[(size_t)k * n_q + c] = cnt ? linspaced(cnt, a[c], b[c], i) : a[c];
                if (cast)
                  q1[(size_t)k * n_q + c] = cnt ? linspaced(cnt, a[c], b[c], i + 1) : b[c];
              }
              ++k;
            }
            ++n;
          });
  if (k != n_sel)
    return fail(T4C_ERR_INVALID, "item index %lld outside [0,%lld)", (long long)items[k], (long long)n);
  return T4C_OK;
}

t4c_status t4c_check_trajectories_discrete(t4c_scene* s, const t4c_fk_model* m, const double* traj, int64_t n_traj,
                                           int32_t n_wp, double lvs, const t4c_request* req, t4c_result* r,
                                           int64_t* n_items)
{
  return checkTrajectories(s, m, traj, n_traj, n_wp, lvs, req, r, n_items, false);
}
t4c_status t4c_check_trajectories_cast(t4c_scene* s, const t4c_fk_model* m, const double* traj, int64_t n_traj,
                                       int32_t n_wp, double lvs, const t4c_request* req, t4c_result* r,
                                       int64_t* n_items)
{
  return checkTrajectories(s, m, traj, n_traj, n_wp, lvs, req, r, n_items, true);
}

t4c_status t4c_check_discrete_device(t4c_scene* s, const t4c_fk_model* m, int64_t n, const t4c_request* req,
                                     t4c_device_io* d)
{
  if (!s || !m || !req || !d || (n > 0 && !d->q))
    return fail(T4C_ERR_INVALID, "null argument");
  BatchIO io;
  io.n = n;
  io.model = m;
  io.req = *req;
  io.q0 = d->q;
  io.q_layout = d->q_layout;
  io.hit = d->hit;
  io.contacts = d->contacts;
  io.capacity = d->capacity;
  const t4c_status rc = runBatch(s, io);
  d->n_contacts = io.n_contacts;
  d->stats = io.stats;
  return rc;
}

t4c_status t4c_check_cast_device(t4c_scene* s, const t4c_fk_model* m, int64_t n, const void* q1,
                                 const t4c_request* req, t4c_device_io* d)
{
  if (!s || !m || !req || !d || (n > 0 && (!d->q || !q1)))
    return fail(T4C_ERR_INVALID, "null argument");
  BatchIO io;
  io.n = n;
  io.cast = true;
  io.model = m;
  io.req = *req;
  io.q0 = d->q;
  io.q1 = q1;
  io.q_layout = d->q_layout;
  io.hit = d->hit;
  io.contacts = d->contacts;
  io.cc_local = d->cc_local;
  io.capacity = d->capacity;
  const t4c_status rc = runBatch(s, io);
  d->n_contacts = io.n_contacts;
  d->stats = io.stats;
  return rc;
}

// Test hook (host only, no device work): evaluates the hull support mapping of shape `shape` for n directions twice,
// through the support table and by the plain scan, in FP64 and FP32.  out[4*i..] = table64, scan64, table32, scan32.
t4c_status t4c_debug_hull_support(t4c_scene* s, int32_t shape, const double* dirs, int64_t n, int32_t* out,
                                  double* mean_candidates)
{
  if (!s || !dirs || !out || shape < 0 || shape >= (int)s->img->comp.shapes.size())
    return fail(T4C_ERR_INVALID, "bad argument");
  const CompiledScene& C = s->img->comp;
  const DShape& D = C.shapes[shape];
  if (D.type != SH_HULL)
    return fail(T4C_ERR_INVALID, "shape %d is not a hull", shape);
  Xf<double> I;
  Xf<float> If;
  ShapeW<double> t64;
  ShapeW<float> t32;
  makeShapeW<double>(t64, D, C.verts.data(), C.sup_cells.data(), C.sup_cand.data(), I, I, false);
  makeShapeW<float>(t32, D, C.verts.data(), C.sup_cells.data(), C.sup_cand.data(), If, If, false);
  ShapeW<double> b64 = t64;
  ShapeW<float> b32 = t32;
  b64.sup_cells = nullptr, b32.sup_cells = nullptr;
  auto indexOf = [&](double x, double y, double z) {
    for (int i = 0; i < D.vert_cnt; ++i)
    {
      const f4& v = C.verts[(size_t)D.vert_off + i];
      if ((double)v.x == x && (double)v.y == y && (double)v.z == z)
        return i;
    }
    return -1;
  };
  for (int64_t i = 0; i < n; ++i)
  {
    const V3<double> d(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]);
    const V3<float> df((float)d.x, (float)d.y, (float)d.z);
    const V3<double> a = coreSupport(t64, d), b = coreSupport(b64, d);
    const V3<float> c = coreSupport(t32, df), e = coreSupport(b32, df);
    out[4 * i + 0] = indexOf(a.x, a.y, a.z), out[4 * i + 1] = indexOf(b.x, b.y, b.z);
    out[4 * i + 2] = indexOf(c.x, c.y, c.z), out[4 * i + 3] = indexOf(e.x, e.y, e.z);
  }
  if (mean_candidates)
  {
    *mean_candidates = -1;
    if (D.sup_off >= 0)
    {
      const int cells = 6 * T4_SUP_R * T4_SUP_R;
      *mean_candidates = double(C.sup_cells[D.sup_off + cells] - C.sup_cells[D.sup_off]) / cells;
    }
  }
  return T4C_OK;
}

t4c_status t4c_measure_fp32_peak(int32_t device, double* tflops)
{
  if (!tflops)
    return fail(T4C_ERR_INVALID, "null argument");
#ifdef T4C_HOSTSIM
  (void)device;
  return fail(T4C_ERR_CUDA, "host simulator has no device");
#else
  return measurePeak<float>(device, tflops, fmaPeakKernel, 1 << 15);
#endif
}

t4c_status t4c_measure_fp64_peak(int32_t device, double* tflops)
{
  if (!tflops)
    return fail(T4C_ERR_INVALID, "null argument");
#ifdef T4C_HOSTSIM
  (void)device;
  return fail(T4C_ERR_CUDA, "host simulator has no device");
#else
  return measurePeak<double>(device, tflops, dfmaPeakKernel, 1 << 13);
#endif
}

t4c_status t4c_debug_sincos(int32_t device, const double* x, int64_t n, double* sin_out, double* cos_out)
{
  if (!x || !sin_out || !cos_out || n < 0)
    return fail(T4C_ERR_INVALID, "null argument");
#ifdef T4C_HOSTSIM
  (void)device;
  for (int64_t i = 0; i < n; ++i)
  {
    sin_out[i] = t4sin(x[i]);
    cos_out[i] = t4cos(x[i]);
  }
  return T4C_OK;
#else
  if (device < 0 || device >= plat::deviceCount())
    return fail(T4C_ERR_CUDA, "no usable CUDA device %d", device);
  plat::setDevice(device);
  DevBuf dx, ds, dc;
  const size_t bytes = (size_t)n * sizeof(double);
  T4_TRY(dx.ensure(bytes) && ds.ensure(bytes) && dc.ensure(bytes) && plat::h2d(dx.p, x, bytes, nullptr), "sincos workspace");
  if (n > 0)
  {
    sincosDebugKernel<<<gridFor(n, 256, 8), 256>>>(dx.as<double>(), n, ds.as<double>(), dc.as<double>());
    ++g_launches;
  }
  T4_TRY(launchOk() && plat::d2h(sin_out, ds.p, bytes, nullptr) && plat::d2h(cos_out, dc.p, bytes, nullptr) &&
             plat::chk(cudaDeviceSynchronize()),
         "sincos");
  return T4C_OK;
#endif
}
}  // extern "C"

#include "t4_multi.h"
