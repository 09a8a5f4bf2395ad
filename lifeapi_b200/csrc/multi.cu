// multi.cu -- the multi-GPU side of the C ABI (SURVEY.md 8b "device management (init with device list)",
// "multi-GPU run + gather"; 8e): a group of contexts, one per GPU, that shards a batch by contiguous id range
// and gathers the hit lists and counters over NCCL.
//
// The path shards with no exchange step (closed 64x64 torus, independent patterns; the reference loop a
// caller shards is LifeAPI.hpp:877-881), so the only collective is one gather per batch:
//   1. every rank runs its shard (one host thread per local GPU) and leaves its hit records, sorted by id, and
//      its four counters on its device (survey_to_device / enumerate_to_device, csrc/match.cu)
//   2. ncclAllGather of the counters (4 x u64 per rank) -- every rank learns every count (a group that lives in
//      one process already has them on the host and skips this)
//   3. grouped ncclSend / ncclRecv of the variable-length hit lists to rank 0's device buffer, at the offsets
//      the counts give; ranks own ascending id ranges, so the concatenation IS the sorted list
//   4. rank 0 downloads the list once
// Two ways to build the group: all GPUs in one process (lifeb200_multi_create: ncclCommInitAll, one host thread
// per device inside the calls) or one process per GPU (lifeb200_multi_create_rank: ncclCommInitRank with a
// unique id the caller distributes -- torchrun / MPI style).  The calls are SPMD either way: every process
// passes the same global arguments.  NCCL is dlopen-ed at group creation (libnccl.so.2, or $LIFEB200_NCCL_LIB),
// so the core library keeps linking only libcudart and a process that already loaded an NCCL (torch) shares it.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "ctx.h"

namespace lifeb200 {

struct Nccl {
  void* handle = nullptr;
  ncclResult_t (*GetVersion)(int*) = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

static Nccl g_nccl;

static int load_nccl() {
  if (g_nccl.handle) return LIFEB200_OK;
  const char* env = std::getenv("LIFEB200_NCCL_LIB");
  const char* names[] = {env && *env ? env : "libnccl.so.2", "libnccl.so.2", "libnccl.so"};
  void* h = nullptr;
  for (const char* nm : names)
    if ((h = dlopen(nm, RTLD_NOW | RTLD_LOCAL))) break;
  if (!h) return set_error(LIFEB200_E_NCCL, "cannot load NCCL (libnccl.so.2; set LIFEB200_NCCL_LIB): %s", dlerror());
  Nccl n;
  n.handle = h;
  bool ok = true;
  auto sym = [&](const char* name) { void* p = dlsym(h, name); if (!p) ok = false; return p; };
  n.GetVersion = reinterpret_cast<decltype(n.GetVersion)>(sym("ncclGetVersion"));
  n.GetUniqueId = reinterpret_cast<decltype(n.GetUniqueId)>(sym("ncclGetUniqueId"));
  n.CommInitAll = reinterpret_cast<decltype(n.CommInitAll)>(sym("ncclCommInitAll"));
  n.CommInitRank = reinterpret_cast<decltype(n.CommInitRank)>(sym("ncclCommInitRank"));
  n.CommDestroy = reinterpret_cast<decltype(n.CommDestroy)>(sym("ncclCommDestroy"));
  n.GroupStart = reinterpret_cast<decltype(n.GroupStart)>(sym("ncclGroupStart"));
  n.GroupEnd = reinterpret_cast<decltype(n.GroupEnd)>(sym("ncclGroupEnd"));
  n.Send = reinterpret_cast<decltype(n.Send)>(sym("ncclSend"));
  n.Recv = reinterpret_cast<decltype(n.Recv)>(sym("ncclRecv"));
  n.AllGather = reinterpret_cast<decltype(n.AllGather)>(sym("ncclAllGather"));
  n.GetErrorString = reinterpret_cast<decltype(n.GetErrorString)>(sym("ncclGetErrorString"));
  if (!ok) { dlclose(h); return set_error(LIFEB200_E_NCCL, "the loaded NCCL library lacks a required symbol"); }
  g_nccl = n;
  return LIFEB200_OK;
}

#define LIFEB200_NCCL(expr)                                                                              \
  do {                                                                                                   \
    ncclResult_t _r = (expr);                                                                            \
    if (_r != ncclSuccess)                                                                               \
      return lifeb200::set_error(LIFEB200_E_NCCL, "NCCL error %d (%s): %s", (int)_r,                      \
                                 lifeb200::g_nccl.GetErrorString ? lifeb200::g_nccl.GetErrorString(_r) : "?", #expr); \
  } while (0)

}  // namespace lifeb200

struct lifeb200_multi {
  int world = 0;       // ranks in the group
  int first_rank = 0;  // rank of local device 0; local device i is rank first_rank + i
  std::vector<lifeb200_ctx*> ctx;   // local contexts
  std::vector<ncclComm_t> comm;     // one communicator per local device (empty when world == 1)
  std::vector<void*> d_allc;        // per local device: world x 4 u64 gathered counters
  unsigned long long* h_allc = nullptr;  // pinned: world x 4 u64
  int gather_mode = 0;              // 0 = NCCL send/recv to rank 0's GPU, one download; 1 = every local GPU downloads
                                    //     its list straight into the caller's buffer (single-process groups only)
  float last_gather_ms = 0.f;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;  // on local device 0: the gather's own timer
  int n_local() const { return (int)ctx.size(); }
  bool owns_root() const { return first_rank == 0; }
};

namespace lifeb200 {

static void shard(uint64_t lo, uint64_t n, int rank, int world, uint64_t* b, uint64_t* e) {
  // contiguous id ranges, the same split as the CPU baseline's thread chunks (128-bit product: n * rank can pass 2^64)
  *b = lo + (uint64_t)(((unsigned __int128)n * (unsigned)rank) / (unsigned)world);
  *e = lo + (uint64_t)(((unsigned __int128)n * (unsigned)(rank + 1)) / (unsigned)world);
}

// run fn(i) for every local device on its own host thread (the calls synchronise their streams)
template <class F>
static int for_each_device(lifeb200_multi* m, F fn) {
  const int k = m->n_local();
  if (k == 1) return fn(0);
  static const bool trace = std::getenv("LIFEB200_TRACE_GATHER") != nullptr;
  const auto t0 = std::chrono::steady_clock::now();
  auto ms = [&] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); };
  std::vector<int> rc(k, LIFEB200_OK);
  std::vector<std::string> msg(k);
  std::vector<double> t_in(k), t_out(k);
  std::vector<std::thread> pool;
  pool.reserve(k);
  for (int i = 0; i < k; i++)
    pool.emplace_back([&, i] {
      t_in[i] = ms();
      rc[i] = fn(i);
      t_out[i] = ms();
      if (rc[i]) msg[i] = lifeb200_last_error();  // the error text is per thread: carry it to the caller's
    });
  for (auto& t : pool) t.join();
  if (trace)
    for (int i = 0; i < k; i++)
      std::fprintf(stderr, "[lifeb200 shard rank %d] thread in +%.3f ms, out +%.3f ms (joined +%.3f ms)\n", m->first_rank + i, t_in[i],
                   t_out[i], ms());
  for (int i = 0; i < k; i++)
    if (rc[i]) return set_error(rc[i], "device %d (rank %d): %s", m->ctx[i]->device, m->first_rank + i, msg[i].c_str());
  return LIFEB200_OK;
}

// Steps 2-4 of the header comment.  Every local context has its sorted hit records in ctx->d_hitbuf and its
// counters in d_counters / h_counters.  On return: counters[4] summed (xor for [3]) over all ranks on every
// rank; on rank 0, hits[0 .. total) = the whole list in id order.
static int gather_body(lifeb200_multi* m, lifeb200_hit* hits, size_t cap, uint64_t* counters) {
  const int world = m->world, k = m->n_local();
  static const bool trace = std::getenv("LIFEB200_TRACE_GATHER") != nullptr;  // host-side timestamps of the steps below
  const auto t_begin = std::chrono::steady_clock::now();
  auto stamp = [&](const char* what) {
    if (trace)
      std::fprintf(stderr, "[lifeb200 gather rank %d] %-28s +%.3f ms\n", m->first_rank, what,
                   std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count());
  };
  std::vector<unsigned long long> all((size_t)world * 4, 0);
  lifeb200_ctx* c0 = m->ctx[0];
  if (k == world) {
    // one process holds every rank: the counts are already on the host (each shard call ended with a download of
    // its counters); the collective is the hit lists
    for (int i = 0; i < k; i++)
      for (int j = 0; j < 4; j++) all[(size_t)i * 4 + j] = m->ctx[i]->h_counters[j];
  } else {
    LIFEB200_NCCL(g_nccl.GroupStart());
    for (int i = 0; i < k; i++)
      LIFEB200_NCCL(g_nccl.AllGather(m->ctx[i]->d_counters, m->d_allc[i], 4, ncclUint64, m->comm[i], m->ctx[i]->stream));
    LIFEB200_NCCL(g_nccl.GroupEnd());
    LIFEB200_CUDA(cudaSetDevice(c0->device));
    LIFEB200_CUDA(cudaMemcpyAsync(m->h_allc, m->d_allc[0], (size_t)world * 4 * 8, cudaMemcpyDeviceToHost, c0->stream));
    LIFEB200_CUDA(cudaStreamSynchronize(c0->stream));
    std::copy(m->h_allc, m->h_allc + (size_t)world * 4, all.begin());
  }
  std::vector<uint64_t> cnt(world), off(world);
  uint64_t total = 0;
  counters[0] = counters[1] = counters[2] = counters[3] = 0;
  for (int r = 0; r < world; r++) {
    cnt[r] = all[(size_t)r * 4];
    off[r] = total;
    total += cnt[r];
    counters[1] += all[(size_t)r * 4 + 1];
    counters[2] += all[(size_t)r * 4 + 2];
    counters[3] ^= all[(size_t)r * 4 + 3];
  }
  counters[0] = total;
  stamp("counts known");
  if (total > cap) return set_error(LIFEB200_E_OVERFLOW, "hit buffer too small: %llu hits, capacity %llu",
                                    (unsigned long long)total, (unsigned long long)cap);
  if (total == 0) return LIFEB200_OK;
  if (world == 1 || (m->gather_mode == 1 && k == world)) {
    // direct: every GPU copies its list to its place in the caller's buffer (parallel PCIe downloads, no NVLink hop)
    for (int i = 0; i < k; i++) {
      if (!cnt[i]) continue;
      LIFEB200_CUDA(cudaSetDevice(m->ctx[i]->device));
      LIFEB200_CUDA(cudaMemcpyAsync(hits + off[i], m->ctx[i]->d_hitbuf, cnt[i] * sizeof(lifeb200_hit), cudaMemcpyDeviceToHost,
                                    m->ctx[i]->stream));
    }
  } else {
    // rank 0's own list already sits at the front of its hit buffer, which the call sized for the whole list before
    // the shards ran (no allocation here: growing a buffer after NCCL has mapped the peers cost 48 ms on the box)
    char* root_buf = m->owns_root() ? static_cast<char*>(c0->d_hitbuf) : nullptr;
    if (m->owns_root() && c0->hitbuf_cap < total * sizeof(lifeb200_hit))
      return set_error(LIFEB200_E_ARG, "internal: gather buffer smaller than the hit list");
    LIFEB200_NCCL(g_nccl.GroupStart());
    for (int i = 0; i < k; i++) {
      const int r = m->first_rank + i;
      if (r == 0) {
        for (int src = 1; src < world; src++)
          if (cnt[src])
            LIFEB200_NCCL(g_nccl.Recv(root_buf + off[src] * sizeof(lifeb200_hit), cnt[src] * sizeof(lifeb200_hit), ncclUint8, src,
                                      m->comm[i], m->ctx[i]->stream));
      } else if (cnt[r]) {
        LIFEB200_NCCL(g_nccl.Send(m->ctx[i]->d_hitbuf, cnt[r] * sizeof(lifeb200_hit), ncclUint8, 0, m->comm[i], m->ctx[i]->stream));
      }
    }
    LIFEB200_NCCL(g_nccl.GroupEnd());
    stamp("send / recv enqueued");
    if (m->owns_root()) {
      LIFEB200_CUDA(cudaSetDevice(c0->device));
      LIFEB200_CUDA(cudaMemcpyAsync(hits, root_buf, total * sizeof(lifeb200_hit), cudaMemcpyDeviceToHost, c0->stream));
    }
  }
  stamp("download enqueued");
  for (int i = 0; i < k; i++) {
    LIFEB200_CUDA(cudaSetDevice(m->ctx[i]->device));
    LIFEB200_CUDA(cudaStreamSynchronize(m->ctx[i]->stream));
  }
  stamp("streams drained");
  return LIFEB200_OK;
}

static int gather(lifeb200_multi* m, lifeb200_hit* hits, size_t cap, uint64_t* counters) {
  lifeb200_ctx* c0 = m->ctx[0];
  LIFEB200_CUDA(cudaSetDevice(c0->device));
  LIFEB200_CUDA(cudaEventRecord(m->ev0, c0->stream));
  if (int rc = gather_body(m, hits, cap, counters)) return rc;
  LIFEB200_CUDA(cudaSetDevice(c0->device));
  LIFEB200_CUDA(cudaEventRecord(m->ev1, c0->stream));
  LIFEB200_CUDA(cudaEventSynchronize(m->ev1));
  LIFEB200_CUDA(cudaEventElapsedTime(&m->last_gather_ms, m->ev0, m->ev1));
  return LIFEB200_OK;
}

// NCCL sets a peer connection up at the first send / recv that uses it (40 ms in one process, 0.4 s across
// processes on the box): do that once at creation with an 8-byte message per rank, not inside the first batch.
static int warm_connections(lifeb200_multi* m) {
  const int world = m->world, k = m->n_local();
  LIFEB200_NCCL(g_nccl.GroupStart());
  for (int i = 0; i < k; i++)
    LIFEB200_NCCL(g_nccl.AllGather(m->ctx[i]->d_counters, m->d_allc[i], 4, ncclUint64, m->comm[i], m->ctx[i]->stream));
  LIFEB200_NCCL(g_nccl.GroupEnd());
  LIFEB200_NCCL(g_nccl.GroupStart());
  for (int i = 0; i < k; i++) {
    const int r = m->first_rank + i;
    if (r == 0) {
      for (int src = 1; src < world; src++)
        LIFEB200_NCCL(g_nccl.Recv(static_cast<char*>(m->d_allc[i]) + 8 * src, 8, ncclUint8, src, m->comm[i], m->ctx[i]->stream));
    } else {
      LIFEB200_NCCL(g_nccl.Send(m->ctx[i]->d_counters, 8, ncclUint8, 0, m->comm[i], m->ctx[i]->stream));
    }
  }
  LIFEB200_NCCL(g_nccl.GroupEnd());
  for (int i = 0; i < k; i++) {
    LIFEB200_CUDA(cudaSetDevice(m->ctx[i]->device));
    LIFEB200_CUDA(cudaStreamSynchronize(m->ctx[i]->stream));
  }
  return LIFEB200_OK;
}

static int finish_create(lifeb200_multi* m) {
  LIFEB200_CUDA(cudaSetDevice(m->ctx[0]->device));
  LIFEB200_CUDA(cudaEventCreate(&m->ev0));
  LIFEB200_CUDA(cudaEventCreate(&m->ev1));
  if (m->world > 1) {
    m->d_allc.assign(m->n_local(), nullptr);
    for (int i = 0; i < m->n_local(); i++) {
      LIFEB200_CUDA(cudaSetDevice(m->ctx[i]->device));
      LIFEB200_CUDA(cudaMalloc(&m->d_allc[i], (size_t)m->world * 4 * 8));
    }
    LIFEB200_CUDA(cudaMallocHost(&m->h_allc, (size_t)m->world * 4 * 8));
    for (int i = 0; i < m->n_local(); i++) {
      LIFEB200_CUDA(cudaSetDevice(m->ctx[i]->device));
      LIFEB200_CUDA(cudaMemsetAsync(m->ctx[i]->d_counters, 0, 8 * sizeof(unsigned long long), m->ctx[i]->stream));
    }
    return warm_connections(m);
  }
  return LIFEB200_OK;
}

}  // namespace lifeb200

using namespace lifeb200;

extern "C" {

int lifeb200_multi_destroy(lifeb200_multi* m) {
  if (!m) return LIFEB200_OK;
  for (size_t i = 0; i < m->comm.size(); i++)
    if (m->comm[i] && g_nccl.CommDestroy) g_nccl.CommDestroy(m->comm[i]);
  for (size_t i = 0; i < m->d_allc.size(); i++)
    if (m->d_allc[i]) { cudaSetDevice(m->ctx[i]->device); cudaFree(m->d_allc[i]); }
  if (m->h_allc) cudaFreeHost(m->h_allc);
  if (!m->ctx.empty()) {
    cudaSetDevice(m->ctx[0]->device);
    if (m->ev0) cudaEventDestroy(m->ev0);
    if (m->ev1) cudaEventDestroy(m->ev1);
  }
  for (lifeb200_ctx* c : m->ctx) lifeb200_destroy(c);
  delete m;
  return LIFEB200_OK;
}

static int lifeb200_multi_create_body(const int* devices, int n_devices, lifeb200_multi** out) {
  if (!out) return set_error(LIFEB200_E_ARG, "null out");
  *out = nullptr;
  int count = 0;
  LIFEB200_CUDA(cudaGetDeviceCount(&count));
  if (count <= 0) return set_error(LIFEB200_E_NODEVICE, "no CUDA device; this library has no CPU fallback");
  if (n_devices <= 0) n_devices = devices ? 0 : count;  // NULL list and n <= 0: every visible device
  if (n_devices <= 0) return set_error(LIFEB200_E_ARG, "empty device list");
  std::vector<int> devs(n_devices);
  for (int i = 0; i < n_devices; i++) {
    devs[i] = devices ? devices[i] : i;
    for (int j = 0; j < i; j++)
      if (devs[j] == devs[i]) return set_error(LIFEB200_E_ARG, "device %d listed twice", devs[i]);
  }
  lifeb200_multi* m = new (std::nothrow) lifeb200_multi();
  if (!m) return set_error(LIFEB200_E_ARG, "out of host memory");
  m->world = n_devices;
  m->first_rank = 0;
  for (int i = 0; i < n_devices; i++) {
    lifeb200_ctx* c = nullptr;
    if (int rc = lifeb200_create(devs[i], &c)) { lifeb200_multi_destroy(m); return rc; }
    m->ctx.push_back(c);
  }
  auto body = [&]() -> int {
    if (n_devices > 1) {
      if (int rc = load_nccl()) return rc;
      m->comm.assign(n_devices, nullptr);
      LIFEB200_NCCL(g_nccl.CommInitAll(m->comm.data(), n_devices, devs.data()));
    }
    return finish_create(m);
  };
  if (int rc = body()) { lifeb200_multi_destroy(m); return rc; }
  *out = m;
  return LIFEB200_OK;
}

int lifeb200_multi_unique_id(char* id128) {
  if (!id128) return set_error(LIFEB200_E_ARG, "null id buffer");
  if (int rc = load_nccl()) return rc;
  ncclUniqueId id;
  LIFEB200_NCCL(g_nccl.GetUniqueId(&id));
  static_assert(sizeof(id) == LIFEB200_UNIQUE_ID_BYTES, "ncclUniqueId size");
  std::memcpy(id128, &id, sizeof id);
  return LIFEB200_OK;
}

static int lifeb200_multi_create_rank_body(int device, int rank, int world, const char* id128, lifeb200_multi** out) {
  if (!out) return set_error(LIFEB200_E_ARG, "null out");
  *out = nullptr;
  if (world < 1 || rank < 0 || rank >= world) return set_error(LIFEB200_E_ARG, "bad rank / world %d / %d", rank, world);
  if (world > 1 && !id128) return set_error(LIFEB200_E_ARG, "null unique id");
  lifeb200_multi* m = new (std::nothrow) lifeb200_multi();
  if (!m) return set_error(LIFEB200_E_ARG, "out of host memory");
  m->world = world;
  m->first_rank = rank;
  lifeb200_ctx* c = nullptr;
  if (int rc = lifeb200_create(device, &c)) { lifeb200_multi_destroy(m); return rc; }
  m->ctx.push_back(c);
  auto body = [&]() -> int {
    if (world > 1) {
      if (int rc = load_nccl()) return rc;
      ncclUniqueId id;
      std::memcpy(&id, id128, sizeof id);
      m->comm.assign(1, nullptr);
      LIFEB200_CUDA(cudaSetDevice(device));
      LIFEB200_NCCL(g_nccl.CommInitRank(&m->comm[0], world, id, rank));
    }
    return finish_create(m);
  };
  if (int rc = body()) { lifeb200_multi_destroy(m); return rc; }
  *out = m;
  return LIFEB200_OK;
}

int lifeb200_multi_info(lifeb200_multi* m, int* world, int* first_rank, int* n_local, int* nccl_version) {
  if (!m) return set_error(LIFEB200_E_ARG, "null group");
  if (world) *world = m->world;
  if (first_rank) *first_rank = m->first_rank;
  if (n_local) *n_local = m->n_local();
  if (nccl_version) {
    *nccl_version = 0;
    if (g_nccl.GetVersion) g_nccl.GetVersion(nccl_version);
  }
  return LIFEB200_OK;
}

int lifeb200_multi_ctx(lifeb200_multi* m, int local_index, lifeb200_ctx** ctx) {
  if (!m || !ctx) return set_error(LIFEB200_E_ARG, "null argument");
  if (local_index < 0 || local_index >= m->n_local()) return set_error(LIFEB200_E_ARG, "local device index %d out of range", local_index);
  *ctx = m->ctx[local_index];
  return LIFEB200_OK;
}

int lifeb200_multi_set_gather(lifeb200_multi* m, int mode) {
  if (!m) return set_error(LIFEB200_E_ARG, "null group");
  if (mode != 0 && mode != 1) return set_error(LIFEB200_E_ARG, "gather mode must be 0 (NCCL to rank 0) or 1 (direct downloads)");
  if (mode == 1 && m->n_local() != m->world) return set_error(LIFEB200_E_ARG, "direct downloads need every rank in this process");
  m->gather_mode = mode;
  return LIFEB200_OK;
}

int lifeb200_multi_last_gather_ms(lifeb200_multi* m, float* ms) {
  if (!m || !ms) return set_error(LIFEB200_E_ARG, "null argument");
  *ms = m->last_gather_ms;
  return LIFEB200_OK;
}

int lifeb200_multi_shard(lifeb200_multi* m, uint64_t first_id, uint64_t n, int rank, uint64_t* begin, uint64_t* end) {
  if (!m || !begin || !end) return set_error(LIFEB200_E_ARG, "null argument");
  if (rank < 0 || rank >= m->world) return set_error(LIFEB200_E_ARG, "rank %d out of range", rank);
  shard(first_id, n, rank, m->world, begin, end);
  return LIFEB200_OK;
}

static int lifeb200_multi_reserve_body(lifeb200_multi* m, uint64_t n, size_t cap) {
  if (!m) return set_error(LIFEB200_E_ARG, "null group");
  return for_each_device(m, [&](int i) -> int {
    lifeb200_ctx* c = m->ctx[i];
    LIFEB200_CUDA(cudaSetDevice(c->device));
    uint64_t b, e;
    shard(0, n, m->first_rank + i, m->world, &b, &e);
    const size_t shard_cap = (size_t)std::min<uint64_t>(e - b, cap);
    const size_t room = (m->first_rank + i == 0) ? (size_t)std::min<uint64_t>(n, cap) : shard_cap;
    if (int r = ensure_scratch(c->d_hitbuf, c->hitbuf_cap, std::max<size_t>(room, 1) * sizeof(lifeb200_hit))) return r;
    return reserve_ids(c, (size_t)std::max<uint64_t>(e - b, 1));
  });
}

static int lifeb200_multi_soup_survey_body(lifeb200_multi* m, uint64_t first_id, uint64_t n, uint64_t seed, int dist, unsigned gens,
                               lifeb200_hit* hits, size_t cap, uint64_t* n_hits, uint64_t* pop_sum, uint64_t* hash_xor) {
  if (!m) return set_error(LIFEB200_E_ARG, "null group");
  if (gens < 2) return set_error(LIFEB200_E_ARG, "survey needs gens >= 2");
  if (dist != 0 && dist != 1) return set_error(LIFEB200_E_ARG, "dist must be 0 or 1");
  if (n_hits) *n_hits = 0;
  if (pop_sum) *pop_sum = 0;
  if (hash_xor) *hash_xor = 0;
  if (m->owns_root() && !hits && cap) return set_error(LIFEB200_E_ARG, "null hits with non-zero capacity");
  int rc = for_each_device(m, [&](int i) -> int {
    lifeb200_ctx* c = m->ctx[i];
    LIFEB200_CUDA(cudaSetDevice(c->device));
    uint64_t b, e;
    shard(first_id, n, m->first_rank + i, m->world, &b, &e);
    std::memset(c->h_counters, 0, 4 * sizeof(unsigned long long));
    LIFEB200_CUDA(cudaMemsetAsync(c->d_counters, 0, 4 * sizeof(unsigned long long), c->stream));
    // a shard can report at most its own ids and never more than the caller can take
    const size_t shard_cap = (size_t)std::min<uint64_t>(e - b, cap);
    // rank 0's buffer doubles as the gather buffer: room for the whole list the caller can take
    const size_t room = (m->first_rank + i == 0) ? (size_t)std::min<uint64_t>(n, cap) : shard_cap;
    if (int r = ensure_scratch(c->d_hitbuf, c->hitbuf_cap, std::max<size_t>(room, 1) * sizeof(lifeb200_hit))) return r;
    if (e == b) { LIFEB200_CUDA(cudaStreamSynchronize(c->stream)); return LIFEB200_OK; }
    return survey_to_device(c, b, (size_t)(e - b), seed, dist, gens, static_cast<lifeb200_hit*>(c->d_hitbuf), shard_cap);
  });
  if (rc) return rc;
  uint64_t counters[4];
  rc = gather(m, hits, cap, counters);
  if (n_hits) *n_hits = counters[0];
  if (pop_sum) *pop_sum = counters[2];
  if (hash_xor) *hash_xor = counters[3];
  return rc;
}

static int lifeb200_multi_enumerate_collisions_body(lifeb200_multi* m, const uint64_t* catalysts, unsigned n_catalysts,
                                        const uint64_t* projectiles, unsigned n_projectiles, uint64_t id_begin,
                                        uint64_t id_end, unsigned gens, unsigned stable_interval, lifeb200_hit* hits,
                                        size_t cap, uint64_t* n_hits, uint64_t* n_skipped) {
  if (!m) return set_error(LIFEB200_E_ARG, "null group");
  if (!catalysts || !projectiles || n_catalysts == 0 || n_projectiles == 0)
    return set_error(LIFEB200_E_ARG, "empty catalyst / projectile table");
  if (stable_interval == 0) return set_error(LIFEB200_E_ARG, "stable_interval must be >= 1");
  const uint64_t total = (uint64_t)n_catalysts * n_projectiles * 4096ull;
  if (id_end > total) id_end = total;
  if (n_hits) *n_hits = 0;
  if (n_skipped) *n_skipped = 0;
  if (id_begin > id_end) id_begin = id_end;
  if (m->owns_root() && !hits && cap) return set_error(LIFEB200_E_ARG, "null hits with non-zero capacity");
  int rc = for_each_device(m, [&](int i) -> int {
    lifeb200_ctx* c = m->ctx[i];
    LIFEB200_CUDA(cudaSetDevice(c->device));
    uint64_t b, e;
    shard(id_begin, id_end - id_begin, m->first_rank + i, m->world, &b, &e);
    std::memset(c->h_counters, 0, 4 * sizeof(unsigned long long));
    LIFEB200_CUDA(cudaMemsetAsync(c->d_counters, 0, 4 * sizeof(unsigned long long), c->stream));
    const size_t shard_cap = (size_t)std::min<uint64_t>(e - b, cap);
    const size_t room = (m->first_rank + i == 0) ? (size_t)std::min<uint64_t>(id_end - id_begin, cap) : shard_cap;
    if (int r = ensure_scratch(c->d_hitbuf, c->hitbuf_cap, std::max<size_t>(room, 1) * sizeof(lifeb200_hit))) return r;
    if (e == b) { LIFEB200_CUDA(cudaStreamSynchronize(c->stream)); return LIFEB200_OK; }
    return enumerate_to_device(c, catalysts, n_catalysts, projectiles, n_projectiles, b, e, gens, stable_interval,
                               static_cast<lifeb200_hit*>(c->d_hitbuf), shard_cap);
  });
  if (rc) return rc;
  uint64_t counters[4];
  rc = gather(m, hits, cap, counters);
  if (n_hits) *n_hits = counters[0];
  if (n_skipped) *n_skipped = counters[1];
  return rc;
}

static int lifeb200_multi_step_batch_host_body(lifeb200_multi* m, const uint64_t* in, uint64_t* out, size_t n, unsigned gens) {
  if (!m) return set_error(LIFEB200_E_ARG, "null group");
  if (n == 0) return LIFEB200_OK;
  if (!in || !out) return set_error(LIFEB200_E_ARG, "null pattern pointer");
  const int k = m->n_local();
  return for_each_device(m, [&](int i) -> int {
    const size_t b = n * (size_t)i / k, e = n * (size_t)(i + 1) / k;  // this process's patterns over its own GPUs
    return lifeb200_step_batch_host(m->ctx[i], in + b * 64, out + b * 64, e - b, gens);
  });
}

int lifeb200_multi_sync(lifeb200_multi* m) {
  if (!m) return set_error(LIFEB200_E_ARG, "null group");
  for (lifeb200_ctx* c : m->ctx)
    if (int rc = lifeb200_sync(c)) return rc;
  return LIFEB200_OK;
}

int lifeb200_multi_create(const int* devices, int n_devices, lifeb200_multi** out) {
  return guarded([&]() -> int { return lifeb200_multi_create_body(devices, n_devices, out); });
}

int lifeb200_multi_create_rank(int device, int rank, int world, const char* id128, lifeb200_multi** out) {
  return guarded([&]() -> int { return lifeb200_multi_create_rank_body(device, rank, world, id128, out); });
}

int lifeb200_multi_reserve(lifeb200_multi* m, uint64_t n, size_t cap) {
  return guarded([&]() -> int { return lifeb200_multi_reserve_body(m, n, cap); });
}

int lifeb200_multi_soup_survey(lifeb200_multi* m, uint64_t first_id, uint64_t n, uint64_t seed, int dist, unsigned gens,
                               lifeb200_hit* hits, size_t cap, uint64_t* n_hits, uint64_t* pop_sum, uint64_t* hash_xor) {
  return guarded([&]() -> int { return lifeb200_multi_soup_survey_body(m, first_id, n, seed, dist, gens, hits, cap, n_hits, pop_sum, hash_xor); });
}

int lifeb200_multi_enumerate_collisions(lifeb200_multi* m, const uint64_t* catalysts, unsigned n_catalysts,
                                        const uint64_t* projectiles, unsigned n_projectiles, uint64_t id_begin,
                                        uint64_t id_end, unsigned gens, unsigned stable_interval, lifeb200_hit* hits,
                                        size_t cap, uint64_t* n_hits, uint64_t* n_skipped) {
  return guarded([&]() -> int { return lifeb200_multi_enumerate_collisions_body(m, catalysts, n_catalysts, projectiles, n_projectiles, id_begin, id_end, gens, stable_interval, hits, cap, n_hits, n_skipped); });
}

int lifeb200_multi_step_batch_host(lifeb200_multi* m, const uint64_t* in, uint64_t* out, size_t n, unsigned gens) {
  return guarded([&]() -> int { return lifeb200_multi_step_batch_host_body(m, in, out, n, gens); });
}

}  // extern "C"
