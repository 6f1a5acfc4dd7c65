// Communicator, peer arena, flags, status word, event handles.
//
// B200-native stand-in for what the reference gets from Horovod's C++ core and c10d/NCCL
// (patch_files/horovod/torch/mpi_ops.py:28-88,420-439): every rank owns one
// cudaMalloc'ed arena, exports it with CUDA IPC, and maps all peers' arenas.  Over NVSwitch
// every peer is one hop away at full bandwidth, so kernels simply store to / load from the
// mapped pointers; completion is signalled by epoch-numbered flags written with
// st.release.sys and polled with ld.acquire.sys.  No negotiation thread is needed because
// every rank issues the same plans in the same order.
#include "engine.cuh"

#include <stdarg.h>
#include <stdlib.h>

#include <mutex>
#include <unordered_map>

static thread_local char g_err[512] = "";

void gb_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* gb_last_error(void) { return g_err; }
extern "C" int gb_version(void) { return 100; }

extern "C" int gb_set_device(int device) {
  GB_CUDA(cudaSetDevice(device));
  return 0;
}

static int comm_init_common(gb_comm* c, size_t arena_bytes) {
  c->bytes = (arena_bytes + 255) & ~size_t(255);
  GB_CHECK(c->bytes > GB_FLAG_BYTES, "arena of %zu bytes is smaller than the flag region", arena_bytes);
  GB_CUDA(cudaGetDevice(&c->device));
  GB_CUDA(cudaMalloc((void**)&c->base, c->bytes));
  GB_CUDA(cudaMemset(c->base, 0, c->bytes));
  GB_CUDA(cudaHostAlloc((void**)&c->status_host, 64, cudaHostAllocMapped));
  memset(c->status_host, 0, 64);
  GB_CUDA(cudaHostGetDevicePointer((void**)&c->status_dev, c->status_host, 0));
  c->cursor = GB_FLAG_BYTES;
  c->next_channel = 1;  // channel 0 is the barrier
  // NCCL-like patience by default (a rank may be saving a checkpoint); GRACE_B200_TIMEOUT_MS overrides
  c->timeout_ns = 120ull * 1000 * 1000 * 1000;
  if (const char* e = getenv("GRACE_B200_TIMEOUT_MS")) {
    const long long ms = atoll(e);
    if (ms > 0) c->timeout_ns = (uint64_t)ms * 1000000ull;
  }
  c->need_bytes = 0;
  c->siblings = nullptr;
  c->guard = getenv("GRACE_B200_GUARD") && atoi(getenv("GRACE_B200_GUARD")) != 0;
  c->barrier_epoch = 0;
  for (int i = 0; i < GB_MAX_WORLD; ++i) { c->peer[i] = nullptr; c->opened[i] = false; }
  c->peer[c->rank] = c->base;
  GB_CUDA(cudaDeviceSynchronize());
  return 0;
}

extern "C" int gb_comm_create(int rank, int world, size_t arena_bytes, gb_comm** out) {
  GB_CHECK(world >= 1 && world <= GB_MAX_WORLD, "world size %d not in [1, %d]", world, GB_MAX_WORLD);
  GB_CHECK(rank >= 0 && rank < world, "rank %d out of range", rank);
  gb_comm* c = new gb_comm();
  c->rank = rank;
  c->world = world;
  c->local_group = false;
  if (comm_init_common(c, arena_bytes)) { delete c; return -1; }
  *out = c;
  return 0;
}

extern "C" int gb_comm_ipc_handle(gb_comm* c, void* out64) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
  cudaIpcMemHandle_t h;
  GB_CUDA(cudaIpcGetMemHandle(&h, c->base));
  memcpy(out64, &h, 64);
  return 0;
}

extern "C" int gb_comm_open_peers(gb_comm* c, const void* handles) {
  for (int r = 0; r < c->world; ++r) {
    if (r == c->rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, (const char*)handles + 64 * r, 64);
    void* p = nullptr;
    GB_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    c->peer[r] = (char*)p;
    c->opened[r] = true;
  }
  return 0;
}

extern "C" int gb_comm_create_local(int world, size_t arena_bytes, gb_comm** out) {
  GB_CHECK(world >= 1 && world <= GB_MAX_WORLD, "world size %d not in [1, %d]", world, GB_MAX_WORLD);
  for (int r = 0; r < world; ++r) {
    gb_comm* c = new gb_comm();
    c->rank = r;
    c->world = world;
    c->local_group = true;
    if (comm_init_common(c, arena_bytes)) return -1;
    out[r] = c;
  }
  for (int r = 0; r < world; ++r)
    for (int q = 0; q < world; ++q) out[r]->peer[q] = out[q]->base;
  gb_comm** sib = new gb_comm*[world];           // shared by the group; released with rank 0
  for (int r = 0; r < world; ++r) sib[r] = out[r];
  for (int r = 0; r < world; ++r) out[r]->siblings = sib;
  return 0;
}

extern "C" int gb_comm_destroy(gb_comm* c) {
  if (!c) return 0;
  cudaDeviceSynchronize();
  for (int r = 0; r < c->world; ++r)
    if (c->opened[r]) cudaIpcCloseMemHandle(c->peer[r]);
  for (gb_region& g : c->extra) {
    for (int r = 0; r < c->world; ++r)
      if (g.opened[r]) cudaIpcCloseMemHandle(g.peer[r]);
    if (!g.adopted) cudaFree(g.base);
  }
  if (c->siblings && c->rank == 0) delete[] c->siblings;
  cudaFree(c->base);
  cudaFreeHost(c->status_host);
  delete c;
  return 0;
}

extern "C" int gb_comm_status(gb_comm* c, uint32_t* out) {
  *out = *(volatile uint32_t*)c->status_host;
  return 0;
}
extern "C" int gb_comm_clear_status(gb_comm* c) {
  *(volatile uint32_t*)c->status_host = 0;
  return 0;
}
extern "C" int gb_comm_set_timeout_ms(gb_comm* c, uint64_t ms) {
  c->timeout_ns = ms * 1000000ull;
  return 0;
}
extern "C" int gb_comm_rank(gb_comm* c) { return c->rank; }
extern "C" int gb_comm_world(gb_comm* c) { return c->world; }
extern "C" size_t gb_comm_arena_free(gb_comm* c) {
  return c->extra.empty() ? c->bytes - c->cursor : c->extra.back().bytes - c->extra.back().cursor;
}
extern "C" size_t gb_comm_arena_need(gb_comm* c) { return c->need_bytes; }
extern "C" int gb_comm_regions(gb_comm* c) { return 1 + (int)c->extra.size(); }

// Channels are carved from the LAST region only, so every rank (same call sequence, same region list) takes the
// same decision.  When the slot does not fit, need_bytes tells the caller how much to add (gb_comm_region_add is
// collective: every rank adds a region of the same size, exchanges its IPC handle, and retries).
int gb_comm_alloc_channel(gb_comm* c, size_t slot_bytes, int* channel, size_t* offset) {
  size_t need = ((slot_bytes + 255) & ~size_t(255)) * 2 * (size_t)c->world;
  GB_CHECK(c->next_channel < GB_MAX_CHANNELS, "out of flag channels (%d)", GB_MAX_CHANNELS);
  size_t& cursor = c->extra.empty() ? c->cursor : c->extra.back().cursor;
  const size_t bytes = c->extra.empty() ? c->bytes : c->extra.back().bytes;
  if (cursor + need > bytes) {
    c->need_bytes = need;
    gb_set_error("peer arena exhausted: need %zu more bytes, %zu free (grow it with gb_comm_region_add)", need, bytes - cursor);
    return -1;
  }
  c->need_bytes = 0;
  *channel = c->next_channel++;
  *offset = cursor | ((size_t)c->extra.size() << GB_REGION_SHIFT);
  cursor += need;
  if (c->guard && slot_bytes > GB_GUARD_BYTES) {
    // poison the last GB_GUARD_BYTES of each of the 2 * W slots of this channel in MY arena (peers write here)
    const size_t sb = (slot_bytes + 255) & ~size_t(255);
    char* first = gb_slot_of(c, c->rank, *offset, sb, 0, 0) + sb - GB_GUARD_BYTES;
    std::vector<uint32_t> pat(GB_GUARD_BYTES / 4, GB_GUARD_WORD);
    for (int s = 0; s < 2 * c->world; ++s)
      GB_CUDA(cudaMemcpy(first + (size_t)s * sb, pat.data(), GB_GUARD_BYTES, cudaMemcpyHostToDevice));
    c->guarded.push_back(gb_comm::GuardRec{*offset, sb});
  }
  return 0;
}

// ---------------------------------------------------------------- guard bands
__global__ void guard_check_kernel(const uint32_t* band, size_t stride_words, int nslots, uint32_t* status) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nslots * (GB_GUARD_BYTES / 4); i += gridDim.x * blockDim.x) {
    const int s = i / (GB_GUARD_BYTES / 4), w = i % (GB_GUARD_BYTES / 4);
    if (band[(size_t)s * stride_words + w] != GB_GUARD_WORD) atomicOr(status, GB_STATUS_GUARD);
  }
}

// test hook: overwrite one word of the first guard band, as a stray peer store would
extern "C" int gb_comm_guard_poke(gb_comm* c) {
  GB_CHECK(!c->guarded.empty(), "no guarded channel");
  const gb_comm::GuardRec& g = c->guarded.front();
  char* first = gb_slot_of(c, c->rank, g.chan_off, g.slot_bytes, 0, 0) + g.slot_bytes - GB_GUARD_BYTES;
  GB_CUDA(cudaMemset(first + 8, 0, 4));
  return 0;
}

// Scans the guard band behind every wire slot of this rank's arena (GRACE_B200_GUARD=1); an overwritten band raises
// GB_STATUS_GUARD.  Synchronises the device; meant for tests and debugging runs.
extern "C" int gb_comm_check_guards(gb_comm* c, int* nchannels) {
  if (nchannels) *nchannels = (int)c->guarded.size();
  for (const gb_comm::GuardRec& g : c->guarded) {
    const char* first = gb_slot_of(c, c->rank, g.chan_off, g.slot_bytes, 0, 0) + g.slot_bytes - GB_GUARD_BYTES;
    guard_check_kernel<<<2, 256>>>((const uint32_t*)first, g.slot_bytes / 4, 2 * c->world, c->status_dev);
    GB_CUDA(cudaGetLastError());
  }
  GB_CUDA(cudaDeviceSynchronize());
  return 0;
}

static int region_alloc(gb_comm* c, size_t bytes) {
  gb_region g;
  memset(&g, 0, sizeof(g));
  g.bytes = (bytes + 255) & ~size_t(255);
  GB_CUDA(cudaMalloc((void**)&g.base, g.bytes));
  GB_CUDA(cudaMemset(g.base, 0, g.bytes));
  GB_CUDA(cudaDeviceSynchronize());
  g.peer[c->rank] = g.base;
  c->extra.push_back(g);
  c->need_bytes = 0;
  return 0;
}

extern "C" int gb_comm_region_add(gb_comm* c, size_t bytes, int* index) {
  if (c->local_group && c->siblings) {            // single-process group: grow every virtual rank at once
    for (int r = 0; r < c->world; ++r)
      if (region_alloc(c->siblings[r], bytes)) return -1;
    for (int r = 0; r < c->world; ++r)
      for (int q = 0; q < c->world; ++q) c->siblings[r]->extra.back().peer[q] = c->siblings[q]->extra.back().base;
  } else if (region_alloc(c, bytes)) {
    return -1;
  }
  if (index) *index = (int)c->extra.size();
  return 0;
}

// Adopt a symmetric allocation made by the caller (torch.distributed._symmetric_memory): `peers` are this process'
// mappings of every rank's copy (rank order, peers[rank] = the local one), `mc` its NVLS multicast address (0: none).
// The memory stays owned by the caller and must outlive the communicator.  Collective like gb_comm_region_add.
extern "C" int gb_comm_region_adopt(gb_comm* c, const uint64_t* peers, uint64_t mc, size_t bytes, int* index) {
  gb_region g;
  memset(&g, 0, sizeof(g));
  g.bytes = bytes & ~size_t(255);
  for (int r = 0; r < c->world; ++r) g.peer[r] = (char*)(uintptr_t)peers[r];
  g.base = g.peer[c->rank];
  g.mc = (char*)(uintptr_t)mc;
  g.adopted = true;
  GB_CHECK(g.base != nullptr && g.bytes > 0, "gb_comm_region_adopt: empty region");
  c->extra.push_back(g);
  c->need_bytes = 0;
  if (index) *index = (int)c->extra.size();
  return 0;
}
extern "C" int gb_comm_has_multicast(gb_comm* c) { return !c->extra.empty() && c->extra.back().mc != nullptr; }

extern "C" int gb_comm_region_ipc_handle(gb_comm* c, int index, void* out64) {
  GB_CHECK(index >= 1 && index <= (int)c->extra.size(), "no region %d", index);
  cudaIpcMemHandle_t h;
  GB_CUDA(cudaIpcGetMemHandle(&h, c->extra[index - 1].base));
  memcpy(out64, &h, 64);
  return 0;
}

extern "C" int gb_comm_region_open_peers(gb_comm* c, int index, const void* handles) {
  GB_CHECK(index >= 1 && index <= (int)c->extra.size(), "no region %d", index);
  gb_region& g = c->extra[index - 1];
  for (int r = 0; r < c->world; ++r) {
    if (r == c->rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, (const char*)handles + 64 * r, 64);
    void* p = nullptr;
    GB_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    g.peer[r] = (char*)p;
    g.opened[r] = true;
  }
  return 0;
}

PeerPtrs gb_comm_peers(const gb_comm* c) {
  PeerPtrs p;
  for (int i = 0; i < GB_MAX_WORLD; ++i) p.base[i] = i < c->world ? c->peer[i] : nullptr;
  p.world = c->world;
  p.rank = c->rank;
  return p;
}

// ---------------------------------------------------------------- flag barrier
__global__ void barrier_kernel(PeerPtrs peers, uint32_t epoch, uint32_t* status, uint64_t timeout_ns) {
  int r = threadIdx.x;
  if (r < peers.world) {
    __threadfence_system();
    st_release_sys(gb_flag(peers.base[r], 0, epoch & 1, peers.rank), epoch);
    wait_flag(gb_flag(peers.base[peers.rank], 0, epoch & 1, r), epoch, status, timeout_ns);
  }
}

extern "C" int gb_comm_barrier(gb_comm* c, void* stream) {
  c->barrier_epoch++;
  barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(gb_comm_peers(c), c->barrier_epoch, c->status_dev, c->timeout_ns);
  GB_CUDA(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------- event handles
static std::mutex g_hmu;
static std::unordered_map<int64_t, cudaEvent_t> g_handles;
static int64_t g_next_handle = 1;

extern "C" int gb_handle_record(void* stream, int64_t* out_handle) {
  cudaEvent_t ev;
  GB_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  GB_CUDA(cudaEventRecord(ev, (cudaStream_t)stream));
  std::lock_guard<std::mutex> lk(g_hmu);
  *out_handle = g_next_handle++;
  g_handles[*out_handle] = ev;
  return 0;
}

static bool find_handle(int64_t h, cudaEvent_t* ev) {
  std::lock_guard<std::mutex> lk(g_hmu);
  auto it = g_handles.find(h);
  if (it == g_handles.end()) return false;
  *ev = it->second;
  return true;
}

extern "C" int gb_handle_poll(int64_t handle) {
  cudaEvent_t ev;
  GB_CHECK(find_handle(handle, &ev), "unknown handle %lld", (long long)handle);
  cudaError_t e = cudaEventQuery(ev);
  if (e == cudaSuccess) return 1;
  if (e == cudaErrorNotReady) return 0;
  gb_set_error("cudaEventQuery: %s", cudaGetErrorString(e));
  return -1;
}

extern "C" int gb_handle_release(int64_t handle) {
  cudaEvent_t ev;
  {
    std::lock_guard<std::mutex> lk(g_hmu);
    auto it = g_handles.find(handle);
    if (it == g_handles.end()) return 0;
    ev = it->second;
    g_handles.erase(it);
  }
  cudaEventDestroy(ev);
  return 0;
}

extern "C" int gb_handle_wait(int64_t handle) {
  cudaEvent_t ev;
  GB_CHECK(find_handle(handle, &ev), "unknown handle %lld", (long long)handle);
  GB_CUDA(cudaEventSynchronize(ev));
  return gb_handle_release(handle);
}

extern "C" int gb_handle_stream_wait(int64_t handle, void* stream) {
  cudaEvent_t ev;
  GB_CHECK(find_handle(handle, &ev), "unknown handle %lld", (long long)handle);
  GB_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, ev, 0));
  return 0;
}
