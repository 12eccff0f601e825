/*
 * t4_multi.h -- the multi-GPU entries of the C ABI (included by t4c.cu).
 *
 * States are independent given the read-only scene, so a batch is cut into one contiguous shard per GPU, every GPU runs
 * the single-device pipeline on its shard against its own replica of the scene, and the per-GPU results -- hit flags AND
 * contact records -- are gathered on the root GPU in ONE exchange over NVLink:
 *     ncclGroupStart(); ncclSend (ranks != root) / ncclRecv x (world - 1) (root); ncclGroupEnd();
 * preceded by a 256-byte header per rank (states, contacts, work counters) so that the root can place every shard's
 * records behind the previous one's.  No all-reduce, no all-to-all.
 *
 * Compute and transfer overlap: the root owns a gather slab (one region per rank); at the start of a call it tells the
 * peers where it is (CUDA IPC handle between processes, the plain pointer between threads of one process), and every
 * peer's pipeline copies its contact records into its region AS SOON AS A ROUND HAS MADE THEM FINAL -- peer-memory
 * writes by the copy engine over NVLink, on the copy stream, under the rounds that follow (runBatchChunked `drain`).
 * What the final exchange still has to move is the last round's records (about 1 % of a FIRST batch) and the flags;
 * the root then places the regions behind one another with device-local copies.  Measured on 8 x B200 (profiles/r3f):
 * the exchange at the end cost 1.4 ms of a 23.3 ms step when it moved all 560 MB at once.  The reference does the same thing with one manager
 * clone per thread (tesseract_collision/test_suite/.../collision_multi_threaded_unit.hpp:79-127).
 *
 * Two ways in:
 *   t4c_check_discrete_sharded / t4c_check_cast_sharded   one process per GPU (torchrun, MPI): every rank calls with
 *                                                         its shard; the communicator comes from t4c_comm_create
 *   t4c_check_discrete_multi                              one process, a host thread per GPU: the caller hands over
 *                                                         the whole batch (host buffers), the split is internal
 * NCCL is reached through dlopen("libnccl.so.2"): inside a PyTorch process that is the copy PyTorch already loaded,
 * elsewhere the system's; libt4c.so itself has no link-time dependency on it.
 */
#ifndef T4_MULTI_H_
#define T4_MULTI_H_

#ifndef T4C_HOSTSIM
#include <dlfcn.h>
#include <nccl.h>
#include <unistd.h>

#include <mutex>
#include <thread>

namespace
{
struct NcclApi
{
  void* handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  bool ok = false;
};

const NcclApi& nccl()
{
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    for (const char* name : { "libnccl.so.2", "libnccl.so" })
    {
      api.handle = dlopen(name, RTLD_NOW | RTLD_LOCAL);
      if (api.handle)
        break;
    }
    if (!api.handle)
      return;
    auto sym = [](const char* n) { return dlsym(api.handle, n); };
    api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
    api.CommInitAll = (decltype(api.CommInitAll))sym("ncclCommInitAll");
    api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
    api.Send = (decltype(api.Send))sym("ncclSend");
    api.Recv = (decltype(api.Recv))sym("ncclRecv");
    api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
    api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
    api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
    api.ok = api.GetUniqueId && api.CommInitRank && api.CommInitAll && api.CommDestroy && api.Send && api.Recv &&
             api.GroupStart && api.GroupEnd && api.GetErrorString;
  });
  return api;
}
}  // namespace

// what the root tells the peers at the start of a sharded call: where its gather slab is
struct GatherInfo
{
  uint64_t magic, generation;
  int64_t region_cap;  // records per rank region; 0: no slab this call (nothing to push into)
  int64_t pid;         // process of the root: the same process maps nothing, it uses raw_ptr
  uint64_t raw_ptr;
  int32_t root_device, pad;
  cudaIpcMemHandle_t handle;
};
#define T4_GATHER_MAGIC 0x7434635f736c6162ull

struct t4c_comm
{
  ncclComm_t comm = nullptr;
  int rank = 0, world = 1, device = 0;
  DevBuf d_header;  // [world] headers (root receives into it; every rank stages its own in slot `rank`)
  // root side: the gather slab, [world] regions of slab_region records
  DevBuf d_slab;
  std::vector<std::unique_ptr<DevBuf>> retired;  // earlier, smaller slabs (see pushHandshake)
  int64_t slab_region = 0;
  uint64_t slab_gen = 0;
  cudaIpcMemHandle_t slab_handle;
  bool slab_handle_ok = false;
  // peer side: the root's slab as this process sees it
  void* peer_base = nullptr;
  bool peer_mapped = false;  // peer_base came from cudaIpcOpenMemHandle (to be closed)
  uint64_t peer_gen = 0;
  int peer_root = -1;
  DevBuf d_info;
  GatherInfo info_host;
  ~t4c_comm()
  {
    if (peer_base && peer_mapped)
      cudaIpcCloseMemHandle(peer_base);
  }
};

// what a rank tells the root before the payload
struct ShardHeader
{
  int64_t n_states, n_contacts, state_offset, n_pushed;  // n_pushed: records already in this rank's slab region
  t4c_stats stats;
};
static_assert(sizeof(GatherInfo) <= 256, "info slot");
static_assert(sizeof(ShardHeader) <= 256, "header slot");
#define T4_HEADER_SLOT 256

#define T4_NCCL(expr, what)                                                                                            \
  do                                                                                                                   \
  {                                                                                                                    \
    const ncclResult_t _r = (expr);                                                                                    \
    if (_r != ncclSuccess)                                                                                             \
      return fail(T4C_ERR_CUDA, "%s: NCCL %s", what, nccl().GetErrorString(_r));                                       \
  } while (0)

static void addStats(t4c_stats& a, const t4c_stats& b)
{
  uint64_t* pa = reinterpret_cast<uint64_t*>(&a);
  const uint64_t* pb = reinterpret_cast<const uint64_t*>(&b);
  for (size_t i = 0; i < sizeof(t4c_stats) / sizeof(uint64_t); ++i)
    pa[i] += pb[i];
}

// Start of a sharded call: the root makes sure its gather slab exists and tells every peer where it is; a peer maps it
// (once per slab) and points its pipeline's progressive copies at its own region.  EVERY rank of a communicator with
// more than one rank runs the same NCCL calls here whatever it has to say (a root without a contact buffer announces
// region_cap 0): the ranks cannot know each other's arguments.
static t4c_status pushHandshake(t4c_scene* s, t4c_comm* c, int root, const t4c_gathered* out, BatchIO& io)
{
  if (c->world < 2 || !tuning().peer_push)
    return T4C_OK;
  const NcclApi& N = nccl();
  t4stream st = s->stream;
  T4_TRY(c->d_info.ensure(T4_HEADER_SLOT), "handshake staging");
  if (c->rank == root)
  {
    GatherInfo& gi = c->info_host;
    std::memset(&gi, 0, sizeof(gi));
    gi.magic = T4_GATHER_MAGIC;
    if (out && out->contacts && out->capacity > 0)
    {
      // a region holds a rank's fair share twice over; a shard with more records sends the rest in the final exchange
      const int64_t region = std::min<int64_t>(out->capacity, std::max<int64_t>(1 << 16, 2 * out->capacity / c->world));
      const size_t bytes = (size_t)c->world * (size_t)region * sizeof(t4c_contact);
      bool ok = true;
      if (bytes > c->d_slab.cap)
      {
        // a slab that peers may still have mapped is never freed under them (freeing exported memory before the
        // importers have closed it is undefined): it is retired until the communicator goes away
        if (c->d_slab.p)
        {
          c->retired.emplace_back(new DevBuf());
          std::swap(c->retired.back()->p, c->d_slab.p);
          std::swap(c->retired.back()->cap, c->d_slab.cap);
        }
        ok = c->d_slab.ensure(bytes);
        c->slab_handle_ok = ok && cudaIpcGetMemHandle(&c->slab_handle, c->d_slab.p) == cudaSuccess;
        ++c->slab_gen;
        if (!ok)
          g_alloc_failed = false;  // no slab, no overlap: the final exchange moves everything
        cudaGetLastError();
      }
      if (ok)
      {
        c->slab_region = region;
        gi.generation = c->slab_gen;
        gi.region_cap = region;
        gi.pid = (int64_t)getpid();
        gi.raw_ptr = (uint64_t)(uintptr_t)c->d_slab.p;
        gi.root_device = c->device;
        if (c->slab_handle_ok)
          gi.handle = c->slab_handle;
      }
    }
    if (gi.region_cap == 0)
      c->slab_region = 0;
    T4_TRY(plat::h2d(c->d_info.p, &gi, sizeof(gi), st), "handshake");
    T4_NCCL(N.GroupStart(), "handshake");
    for (int r = 0; r < c->world; ++r)
      if (r != root)
        T4_NCCL(N.Send(c->d_info.p, T4_HEADER_SLOT, ncclUint8, r, c->comm, st), "handshake");
    T4_NCCL(N.GroupEnd(), "handshake");
    return T4C_OK;
  }
  T4_NCCL(N.Recv(c->d_info.p, T4_HEADER_SLOT, ncclUint8, root, c->comm, st), "handshake");
  GatherInfo& gi = c->info_host;
  T4_TRY(plat::d2h(&gi, c->d_info.p, sizeof(gi), st) && plat::sync(st), "handshake");
  if (gi.magic != T4_GATHER_MAGIC)
    return fail(T4C_ERR_CUDA, "gather handshake: unexpected message from rank %d", root);
  if (gi.region_cap <= 0 || !io.contacts || !s->copy_stream)
    return T4C_OK;
  if (!c->peer_base || c->peer_gen != gi.generation || c->peer_root != root)
  {
    if (c->peer_base && c->peer_mapped)
      cudaIpcCloseMemHandle(c->peer_base);
    c->peer_base = nullptr;
    c->peer_mapped = false;
    if (gi.pid == (int64_t)getpid())
    {
      // threads of one process: the root's pointer is valid here once this device may address the root's memory
      const cudaError_t e = cudaDeviceEnablePeerAccess(gi.root_device, 0);
      if (e == cudaSuccess || e == cudaErrorPeerAccessAlreadyEnabled)
        c->peer_base = (void*)(uintptr_t)gi.raw_ptr;
    }
    else
    {
      void* base = nullptr;
      if (cudaIpcOpenMemHandle(&base, gi.handle, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess)
      {
        c->peer_base = base;
        c->peer_mapped = true;
      }
    }
    cudaGetLastError();  // no mapping: this rank pushes nothing and sends everything at the end
    c->peer_gen = gi.generation;
    c->peer_root = root;
  }
  if (c->peer_base)
  {
    io.push_dst = reinterpret_cast<t4c_contact*>(c->peer_base) + (size_t)c->rank * (size_t)gi.region_cap;
    io.push_cap = gi.region_cap;
  }
  return T4C_OK;
}

// The single exchange: local results (device) -> gathered buffers on the root (device).  `stats` = this rank's counters.
static t4c_status gatherShard(t4c_scene* s, t4c_comm* c, int root, int64_t n_local, int64_t state_offset,
                              const uint8_t* hit_local, const t4c_contact* con_local, int64_t n_contacts_local,
                              int64_t n_pushed, const t4c_stats& stats, t4c_gathered* out)
{
  const NcclApi& N = nccl();
  t4stream st = s->stream;
  if (c->rank == root && !out)
    return fail(T4C_ERR_INVALID, "the root rank needs a t4c_gathered");
  T4_TRY(c->d_header.ensure((size_t)c->world * T4_HEADER_SLOT), "header staging");
  ShardHeader mine;
  std::memset(&mine, 0, sizeof(mine));
  mine.n_states = n_local, mine.n_contacts = con_local ? n_contacts_local : 0, mine.state_offset = state_offset;
  mine.n_pushed = con_local ? std::min(n_pushed, n_contacts_local) : 0;
  mine.stats = stats;
  // the header tells the root that this many records ARE in the slab: the copies have to be over before it leaves
  if (mine.n_pushed > 0)
    T4_TRY(plat::sync(s->copy_stream), "record push");
  char* hdr = c->d_header.as<char>();
  T4_TRY(plat::h2d(hdr + (size_t)c->rank * T4_HEADER_SLOT, &mine, sizeof(mine), st), "header upload");
  // ---- headers to the root
  if (c->world > 1)
  {
    T4_NCCL(N.GroupStart(), "gather");
    if (c->rank != root)
      T4_NCCL(N.Send(hdr + (size_t)c->rank * T4_HEADER_SLOT, T4_HEADER_SLOT, ncclUint8, root, c->comm, st), "gather");
    else
      for (int r = 0; r < c->world; ++r)
        if (r != root)
          T4_NCCL(N.Recv(hdr + (size_t)r * T4_HEADER_SLOT, T4_HEADER_SLOT, ncclUint8, r, c->comm, st), "gather");
    T4_NCCL(N.GroupEnd(), "gather");
  }
  std::vector<ShardHeader> all((size_t)c->world);
  bool overflow = false;
  if (c->rank == root)
  {
    std::vector<char> raw((size_t)c->world * T4_HEADER_SLOT);
    T4_TRY(plat::d2h(raw.data(), hdr, raw.size(), st) && plat::sync(st), "header download");
    for (int r = 0; r < c->world; ++r)
      std::memcpy(&all[r], raw.data() + (size_t)r * T4_HEADER_SLOT, sizeof(ShardHeader));
    int64_t ns = 0, ncon = 0;
    std::memset(&out->stats, 0, sizeof(out->stats));
    for (auto& h : all)
    {
      ns += h.n_states, ncon += h.n_contacts;
      addStats(out->stats, h.stats);
    }
    out->n_states = ns;
    out->n_contacts = ncon;
    if (tuning().trace)
      for (int r = 0; r < c->world; ++r)
        if (r != root)
          std::fprintf(stderr, "[t4c] gather: rank %d pushed %lld of %lld records into the slab before the exchange\n", r,
                       (long long)all[r].n_pushed, (long long)all[r].n_contacts);
    overflow = out->contacts && ncon > out->capacity;
    // scratch for what the root does not keep (no buffer given, or the records do not fit): the senders transmit
    // exactly what their headers announced, and every send needs its receive
    int64_t max_states = 0, max_con = 0;
    for (auto& h : all)
      max_states = std::max(max_states, h.n_states), max_con = std::max(max_con, h.n_contacts);
    if (!out->hit)
      T4_TRY(s->d_hit_full.ensure((size_t)std::max<int64_t>(max_states, 1)), "gather scratch");
    if (!out->contacts || overflow)
      T4_TRY(s->d_out.ensure((size_t)std::max<int64_t>(max_con, 1) * sizeof(t4c_contact)), "gather scratch");
  }
  // ---- payload: hit flags at the shard's state offset, contact records behind those of the lower ranks
  if (c->world > 1)
    T4_NCCL(N.GroupStart(), "gather");
  if (c->rank != root)
  {
    if (hit_local && n_local > 0)
      T4_NCCL(N.Send(hit_local, (size_t)n_local, ncclUint8, root, c->comm, st), "gather");
    if (con_local && n_contacts_local > mine.n_pushed)
      T4_NCCL(N.Send(con_local + mine.n_pushed, (size_t)(n_contacts_local - mine.n_pushed) * sizeof(t4c_contact),
                     ncclUint8, root, c->comm, st),
              "gather");
  }
  else
  {
    int64_t con_at = 0;
    for (int r = 0; r < c->world; ++r)
    {
      const ShardHeader& h = all[r];
      uint8_t* hit_dst = out->hit ? out->hit + h.state_offset : nullptr;
      t4c_contact* con_dst = (out->contacts && !overflow) ? out->contacts + con_at : nullptr;
      if (r == root)
      {
        if (hit_dst && hit_local && h.n_states > 0)
          T4_TRY(plat::chk(cudaMemcpyAsync(hit_dst, hit_local, (size_t)h.n_states, cudaMemcpyDeviceToDevice, st)), "gather");
        if (con_dst && con_local && h.n_contacts > 0)
          T4_TRY(plat::chk(cudaMemcpyAsync(con_dst, con_local, (size_t)h.n_contacts * sizeof(t4c_contact),
                                           cudaMemcpyDeviceToDevice, st)),
                 "gather");
      }
      else
      {
        // a sender transmits exactly what its header announced; a buffer the root did not ask for still has to be
        // received (into scratch) to keep the pairs of the group matched
        if (h.n_states > 0)
        {
          if (!hit_dst)
            hit_dst = s->d_hit_full.as<uint8_t>();
          T4_NCCL(N.Recv(hit_dst, (size_t)h.n_states, ncclUint8, r, c->comm, st), "gather");
        }
        // records the peer has already written into its slab region: a device-local copy puts them in place
        if (con_dst && h.n_pushed > 0)
          T4_TRY(plat::d2d(con_dst, c->d_slab.as<t4c_contact>() + (size_t)r * (size_t)c->slab_region,
                           (size_t)h.n_pushed * sizeof(t4c_contact), st),
                 "gather");
        if (h.n_contacts > h.n_pushed)
        {
          t4c_contact* rest = con_dst ? con_dst + h.n_pushed : s->d_out.as<t4c_contact>();
          T4_NCCL(N.Recv(rest, (size_t)(h.n_contacts - h.n_pushed) * sizeof(t4c_contact), ncclUint8, r, c->comm, st),
                  "gather");
        }
      }
      con_at += h.n_contacts;
    }
  }
  if (c->world > 1)
    T4_NCCL(N.GroupEnd(), "gather");
  T4_TRY(plat::sync(st), "gather");
  if (overflow)
    return fail(T4C_ERR_CAPACITY, "gathered contact buffer too small: %lld needed, capacity %lld",
                (long long)out->n_contacts, (long long)out->capacity);
  return T4C_OK;
}

static t4c_status checkSharded(t4c_scene* s, const t4c_fk_model* m, t4c_comm* c, int32_t root, int64_t n_local,
                               int64_t state_offset, const void* q1, bool cast, const t4c_request* req, t4c_device_io* d,
                               t4c_gathered* out)
{
  if (!s || !m || !c || !req || !d || (n_local > 0 && !d->q) || (cast && n_local > 0 && !q1) || root < 0 ||
      root >= c->world)
    return fail(T4C_ERR_INVALID, "bad argument");
  if (!nccl().ok)
    return fail(T4C_ERR_CUDA, "libnccl.so.2 could not be loaded");
  if (s->device != c->device)
    return fail(T4C_ERR_INVALID, "scene is on device %d, communicator on device %d", s->device, c->device);
  if (!bindDevice(s))
    return fail(T4C_ERR_CUDA, "cudaSetDevice(%d): %s", s->device, plat::lastError());
  BatchIO io;
  io.n = n_local;
  io.cast = cast;
  io.model = m;
  io.req = *req;
  io.q0 = d->q;
  io.q1 = q1;
  io.q_layout = d->q_layout;
  io.hit = d->hit;
  io.contacts = d->contacts;
  io.cc_local = d->cc_local;
  io.capacity = d->capacity;
  io.state_offset = state_offset;
  // a rank without a hit buffer of its own still has flags to send
  if (!io.hit)
  {
    T4_TRY(s->d_hit_send.ensure((size_t)std::max<int64_t>(n_local, 1)), "workspace");
    io.hit = s->d_hit_send.as<uint8_t>();
  }
  t4c_status rc = pushHandshake(s, c, root, out, io);
  if (rc != T4C_OK)
    return rc;
  rc = runBatch(s, io);
  d->n_contacts = io.n_contacts;
  d->stats = io.stats;
  if (rc != T4C_OK)
    return rc;  // every rank of a failed step has to be handled by the caller (the reference has no partial results)
  return gatherShard(s, c, root, n_local, state_offset, io.hit, io.contacts, std::min(io.n_contacts, io.capacity),
                     io.pushed, io.stats, out);
}

extern "C" {
t4c_status t4c_comm_unique_id(uint8_t* id128)
{
  if (!id128)
    return fail(T4C_ERR_INVALID, "null argument");
  if (!nccl().ok)
    return fail(T4C_ERR_CUDA, "libnccl.so.2 could not be loaded");
  ncclUniqueId id;
  T4_NCCL(nccl().GetUniqueId(&id), "ncclGetUniqueId");
  static_assert(sizeof(id) == 128, "ncclUniqueId");
  std::memcpy(id128, &id, 128);
  return T4C_OK;
}

t4c_status t4c_comm_create(const uint8_t* id128, int32_t rank, int32_t world, int32_t device, t4c_comm** out)
{
  if (!id128 || !out || world < 1 || rank < 0 || rank >= world)
    return fail(T4C_ERR_INVALID, "bad argument");
  if (!nccl().ok)
    return fail(T4C_ERR_CUDA, "libnccl.so.2 could not be loaded");
  if (device < 0 || device >= plat::deviceCount() || !plat::setDevice(device))
    return fail(T4C_ERR_CUDA, "no usable CUDA device %d", device);
  std::unique_ptr<t4c_comm> c(new t4c_comm());
  c->rank = rank, c->world = world, c->device = device;
  ncclUniqueId id;
  std::memcpy(&id, id128, 128);
  T4_NCCL(nccl().CommInitRank(&c->comm, world, id, rank), "ncclCommInitRank");
  *out = c.release();
  return T4C_OK;
}

t4c_status t4c_comm_create_local(const int32_t* devices, int32_t n, t4c_comm** out)
{
  if (!devices || !out || n < 1)
    return fail(T4C_ERR_INVALID, "bad argument");
  if (!nccl().ok)
    return fail(T4C_ERR_CUDA, "libnccl.so.2 could not be loaded");
  std::vector<ncclComm_t> comms((size_t)n);
  std::vector<int> devs(devices, devices + n);
  T4_NCCL(nccl().CommInitAll(comms.data(), n, devs.data()), "ncclCommInitAll");
  for (int i = 0; i < n; ++i)
  {
    out[i] = new t4c_comm();
    out[i]->comm = comms[i], out[i]->rank = i, out[i]->world = n, out[i]->device = devs[i];
  }
  return T4C_OK;
}

void t4c_comm_destroy(t4c_comm* c)
{
  if (!c)
    return;
  plat::setDevice(c->device);
  if (c->comm && nccl().ok)
    nccl().CommDestroy(c->comm);
  delete c;
}

t4c_status t4c_check_discrete_sharded(t4c_scene* s, const t4c_fk_model* m, t4c_comm* c, int32_t root, int64_t n_local,
                                      int64_t state_offset, const t4c_request* req, t4c_device_io* d, t4c_gathered* out)
{
  return checkSharded(s, m, c, root, n_local, state_offset, nullptr, false, req, d, out);
}

t4c_status t4c_check_cast_sharded(t4c_scene* s, const t4c_fk_model* m, t4c_comm* c, int32_t root, int64_t n_local,
                                  int64_t state_offset, const void* q1, const t4c_request* req, t4c_device_io* d,
                                  t4c_gathered* out)
{
  return checkSharded(s, m, c, root, n_local, state_offset, q1, true, req, d, out);
}

// One process, one host thread per device: contiguous shard per device, gather on device 0, download to the caller.
t4c_status t4c_check_discrete_multi(t4c_scene* const* scenes, const t4c_fk_model* m, t4c_comm* const* comms,
                                    int32_t n_dev, const double* q, int64_t n, const t4c_request* req, t4c_result* r)
{
  if (!scenes || !m || !comms || n_dev < 1 || !req || !r || n < 0 || (n > 0 && !q))
    return fail(T4C_ERR_INVALID, "bad argument");
  const int64_t per = (n + n_dev - 1) / n_dev;
  std::vector<t4c_status> rcs((size_t)n_dev, T4C_OK);
  std::vector<std::string> errs((size_t)n_dev);
  const int nq = m->n_q;
  auto work = [&](int k) {
    t4c_scene* s = scenes[k];
    t4c_comm* c = comms[k];
    auto run = [&]() -> t4c_status {
      if (!s || !c || !bindDevice(s))
        return fail(T4C_ERR_INVALID, "bad scene / communicator for device slot %d", k);
      const int64_t a = std::min<int64_t>(n, (int64_t)k * per), b = std::min<int64_t>(n, a + per);
      const int64_t nl = b - a;
      t4stream st = s->stream;
      const size_t qb = (size_t)nl * nq * sizeof(double);
      // shard-local capacity: no shard can hold more than the caller's whole buffer
      const int64_t cap = r->contacts ? r->capacity : 0;
      T4_TRY(s->d_q0.ensure(std::max<size_t>(qb, 8)) && s->d_hit_send.ensure((size_t)std::max<int64_t>(nl, 1)) &&
                 (!cap || s->d_shard_out.ensure((size_t)cap * sizeof(t4c_contact))) &&
                 plat::h2d(s->d_q0.p, q + (size_t)a * nq, qb, st),
             "shard upload");
      t4c_device_io d;
      std::memset(&d, 0, sizeof(d));
      d.q = s->d_q0.p, d.q_layout = T4C_Q_F64_ROWMAJOR;
      d.hit = s->d_hit_send.as<uint8_t>();
      d.contacts = cap ? s->d_shard_out.as<t4c_contact>() : nullptr;
      d.capacity = cap;
      t4c_gathered g;
      std::memset(&g, 0, sizeof(g));
      if (k == 0)
      {
        T4_TRY(s->d_gather_hit.ensure((size_t)std::max<int64_t>(n, 1)) &&
                   (!cap || s->d_gather_out.ensure((size_t)cap * sizeof(t4c_contact))),
               "gather buffers");
        g.hit = s->d_gather_hit.as<uint8_t>();
        g.contacts = cap ? s->d_gather_out.as<t4c_contact>() : nullptr;
        g.capacity = cap;
      }
      t4c_status rc = t4c_check_discrete_sharded(s, m, c, 0, nl, a, req, &d, k == 0 ? &g : nullptr);
      if (k == 0 && (rc == T4C_OK || rc == T4C_ERR_CAPACITY))
      {
        r->n_contacts = g.n_contacts;
        r->stats = g.stats;
        if (rc == T4C_OK)
        {
          bool ok = true;
          if (r->hit)
            ok = ok && plat::d2h(r->hit, g.hit, (size_t)n, st);
          if (r->contacts && g.n_contacts > 0)
            ok = ok && plat::d2h(r->contacts, g.contacts, (size_t)g.n_contacts * sizeof(t4c_contact), st);
          T4_TRY(ok && plat::sync(st), "download results");
        }
      }
      return rc;
    };
    rcs[k] = run();
    if (rcs[k] != T4C_OK)
      errs[k] = g_err;  // thread-local message of this worker
  };
  std::vector<std::thread> threads;
  for (int k = 1; k < n_dev; ++k)
    threads.emplace_back(work, k);
  work(0);
  for (auto& t : threads)
    t.join();
  for (int k = 0; k < n_dev; ++k)
    if (rcs[k] != T4C_OK)
      return fail(rcs[k], "device slot %d: %s", k, errs[k].c_str());
  return T4C_OK;
}
}  // extern "C"

#else  // T4C_HOSTSIM: no devices, no NCCL
struct t4c_comm
{
};
extern "C" {
t4c_status t4c_comm_unique_id(uint8_t*) { return fail(T4C_ERR_CUDA, "host simulator has no device"); }
t4c_status t4c_comm_create(const uint8_t*, int32_t, int32_t, int32_t, t4c_comm**) { return fail(T4C_ERR_CUDA, "host simulator has no device"); }
t4c_status t4c_comm_create_local(const int32_t*, int32_t, t4c_comm**) { return fail(T4C_ERR_CUDA, "host simulator has no device"); }
void t4c_comm_destroy(t4c_comm*) {}
t4c_status t4c_check_discrete_sharded(t4c_scene*, const t4c_fk_model*, t4c_comm*, int32_t, int64_t, int64_t,
                                      const t4c_request*, t4c_device_io*, t4c_gathered*)
{
  return fail(T4C_ERR_CUDA, "host simulator has no device");
}
t4c_status t4c_check_cast_sharded(t4c_scene*, const t4c_fk_model*, t4c_comm*, int32_t, int64_t, int64_t, const void*,
                                  const t4c_request*, t4c_device_io*, t4c_gathered*)
{
  return fail(T4C_ERR_CUDA, "host simulator has no device");
}
t4c_status t4c_check_discrete_multi(t4c_scene* const*, const t4c_fk_model*, t4c_comm* const*, int32_t, const double*,
                                    int64_t, const t4c_request*, t4c_result*)
{
  return fail(T4C_ERR_CUDA, "host simulator has no device");
}
}
#endif
#endif
