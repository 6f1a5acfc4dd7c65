// Horovod-style point collectives on the peer-memory engine.
//
// The reference's op layer (patch_files/horovod/torch/mpi_ops.py:157,216,362: allreduce_async_, allgather_async,
// broadcast_async_ -> Horovod's C++ core -> NCCL/MPI) serves everything that is NOT a gradient bucket:
// broadcast_parameters / broadcast_optimizer_state at start-up, metric averaging, user-level hvd.allreduce.  Here the
// same calls for CUDA tensors on one NVLink domain run as two small kernels over the arena:
//   gb_ops_post    every rank stores its bytes into slot [rank] of every destination (16-byte P2P stores) -- or
//                  nothing but its flag when it has no payload for this op (non-roots of a broadcast) -- and the
//                  last CTA releases the op's epoch flag on every peer;
//   gb_ops_collect waits for ALL ranks' flags (also for a broadcast: the full handshake is what makes the
//                  double-buffered slots safe to reuse two ops later), then copies slot [r] for the wanted ranks
//                  into the output, or sums them in rank order (all-reduce of small tensors).
// Large fp32 all-reduces use the dense two-shot plans (K0) instead; see grace_b200/torch/mpi_ops.py.
#include "kernels.cuh"

#include <algorithm>

struct gb_ops {
  gb_comm* comm;
  int channel;
  size_t chan_off, slot_bytes;
  uint32_t epoch;
  uint32_t* d_done;
  int64_t launches;
};

struct OpsPeers { char* slot[GB_MAX_WORLD]; int n; int mc; };

__global__ void __launch_bounds__(GB_THREADS) gb_ops_post(const uint4* __restrict__ src, size_t n16, OpsPeers dst, uint32_t* done,
                                                         PeerPtrs peers, int channel, uint32_t epoch) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) {
    const uint4 v = src[i];
    for (int r = 0; r < dst.n; ++r) wire_st_u4(reinterpret_cast<uint4*>(dst.slot[r]) + i, v, dst.mc);
  }
  release_flags_last_block(done, peers, channel, epoch);
}

// mode 0: out[r * stride16 + i] = slot[r][i] for r in [first, first + count)    (all-gather / broadcast)
// mode 1: out[i] = sum_r slot[r][i] as float, rank order, times scale             (all-reduce of small fp32 tensors)
// mode 2: the same for double, mode 3 for int64, mode 4 for int32
__global__ void __launch_bounds__(GB_THREADS) gb_ops_collect(OpsPeers src, size_t n16, int first, int count, size_t stride16, int mode,
                                                            double scale, uint4* out, PeerPtrs peers, int channel, uint32_t epoch,
                                                            uint32_t* status, uint64_t timeout_ns) {
  wait_all_ranks(peers, channel, epoch, status, timeout_ns);
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) {
    if (mode == 0) {
      for (int r = first; r < first + count; ++r)
        out[(size_t)(r - first) * stride16 + i] = __ldcg(reinterpret_cast<const uint4*>(src.slot[r]) + i);
    } else if (mode == 1) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int r = 0; r < src.n; ++r) {
        const uint4 u = __ldcg(reinterpret_cast<const uint4*>(src.slot[r]) + i);
        acc.x = __fadd_rn(acc.x, __uint_as_float(u.x)); acc.y = __fadd_rn(acc.y, __uint_as_float(u.y));
        acc.z = __fadd_rn(acc.z, __uint_as_float(u.z)); acc.w = __fadd_rn(acc.w, __uint_as_float(u.w));
      }
      const float s = (float)scale;
      if (s != 1.f) { acc.x = __fmul_rn(acc.x, s); acc.y = __fmul_rn(acc.y, s); acc.z = __fmul_rn(acc.z, s); acc.w = __fmul_rn(acc.w, s); }
      out[i] = make_uint4(__float_as_uint(acc.x), __float_as_uint(acc.y), __float_as_uint(acc.z), __float_as_uint(acc.w));
    } else if (mode == 2) {
      double a0 = 0.0, a1 = 0.0;
      for (int r = 0; r < src.n; ++r) {
        const uint4 u = __ldcg(reinterpret_cast<const uint4*>(src.slot[r]) + i);
        a0 += __hiloint2double((int)u.y, (int)u.x);
        a1 += __hiloint2double((int)u.w, (int)u.z);
      }
      a0 *= scale; a1 *= scale;
      out[i] = make_uint4((uint32_t)__double2loint(a0), (uint32_t)__double2hiint(a0), (uint32_t)__double2loint(a1),
                          (uint32_t)__double2hiint(a1));
    } else if (mode == 3) {
      long long a0 = 0, a1 = 0;
      for (int r = 0; r < src.n; ++r) {
        const uint4 u = __ldcg(reinterpret_cast<const uint4*>(src.slot[r]) + i);
        a0 += (long long)(((unsigned long long)u.y << 32) | u.x);
        a1 += (long long)(((unsigned long long)u.w << 32) | u.z);
      }
      out[i] = make_uint4((uint32_t)a0, (uint32_t)((unsigned long long)a0 >> 32), (uint32_t)a1, (uint32_t)((unsigned long long)a1 >> 32));
    } else {
      int4 a = make_int4(0, 0, 0, 0);
      for (int r = 0; r < src.n; ++r) {
        const uint4 u = __ldcg(reinterpret_cast<const uint4*>(src.slot[r]) + i);
        a.x += (int)u.x; a.y += (int)u.y; a.z += (int)u.z; a.w += (int)u.w;
      }
      out[i] = make_uint4((uint32_t)a.x, (uint32_t)a.y, (uint32_t)a.z, (uint32_t)a.w);
    }
  }
}

extern "C" int gb_ops_create(gb_comm* c, size_t slot_bytes, gb_ops** out) {
  gb_ops* o = new gb_ops();
  o->comm = c;
  o->slot_bytes = gb_guarded(c, slot_bytes);
  o->epoch = 0;
  o->launches = 0;
  o->d_done = nullptr;
  if (gb_comm_alloc_channel(c, o->slot_bytes, &o->channel, &o->chan_off)) { delete o; return -1; }
  if (cudaMalloc((void**)&o->d_done, 64) != cudaSuccess || cudaMemset(o->d_done, 0, 64) != cudaSuccess) {
    gb_set_error("gb_ops_create: cudaMalloc failed");
    cudaFree(o->d_done);
    delete o;
    return -1;
  }
  *out = o;
  return 0;
}

extern "C" int gb_ops_destroy(gb_ops* o) {
  if (!o) return 0;
  cudaDeviceSynchronize();
  cudaFree(o->d_done);
  delete o;
  return 0;
}

extern "C" size_t gb_ops_capacity(gb_ops* o) { return o->slot_bytes - (o->comm->guard ? GB_GUARD_BYTES : 0); }
extern "C" int64_t gb_ops_launches(gb_ops* o) { return o->launches; }

static int ops_grid(size_t n16) { return (int)std::min<size_t>((n16 + GB_THREADS - 1) / GB_THREADS + 1, 148 * 4); }

// One op = gb_ops_send + gb_ops_recv on one stream (two calls so that single-stream emulation of several ranks can post
// all ranks first).  d_src: nbytes of payload (multiple of 16, 16-byte aligned) or NULL when this rank contributes
// nothing (broadcast non-root).  Output: mode 0 -> `count` ranks' payloads starting at rank `first`, `stride` bytes
// apart; modes 1-4 -> the reduced payload (float, double, int64, int32).
static void ops_slots(gb_ops* o, uint32_t epoch, OpsPeers* dst, OpsPeers* src) {
  gb_comm* c = o->comm;
  for (int r = 0; r < c->world; ++r) {
    dst->slot[r] = gb_slot_of(c, r, o->chan_off, o->slot_bytes, epoch & 1, c->rank);
    src->slot[r] = gb_slot_of(c, c->rank, o->chan_off, o->slot_bytes, epoch & 1, r);
  }
  dst->n = src->n = c->world;
  dst->mc = src->mc = 0;
  if (char* mc = gb_slot_mc(c, o->chan_off, o->slot_bytes, epoch & 1, c->rank)) { dst->n = 1; dst->mc = 1; dst->slot[0] = mc; }
}

extern "C" int gb_ops_send(gb_ops* o, const void* d_src, size_t nbytes, void* stream) {
  gb_comm* c = o->comm;
  GB_CHECK(nbytes % 16 == 0 && nbytes <= gb_ops_capacity(o), "gb_ops_send: %zu bytes (slot capacity %zu, multiple of 16 required)",
           nbytes, o->slot_bytes);
  GB_CHECK(((uintptr_t)d_src & 15) == 0, "gb_ops_send: 16-byte alignment");
  const uint32_t epoch = ++o->epoch;
  OpsPeers dst, src;
  ops_slots(o, epoch, &dst, &src);
  const size_t n16 = d_src ? nbytes / 16 : 0;
  gb_ops_post<<<d_src ? ops_grid(n16) : 1, GB_THREADS, 0, (cudaStream_t)stream>>>((const uint4*)d_src, n16, dst, o->d_done,
                                                                                 gb_comm_peers(c), o->channel, epoch);
  GB_CUDA(cudaGetLastError());
  o->launches++;
  return 0;
}

extern "C" int gb_ops_recv(gb_ops* o, size_t nbytes, int first, int count, size_t stride, int mode, double scale, void* d_out,
                           void* stream) {
  gb_comm* c = o->comm;
  GB_CHECK(o->epoch > 0, "gb_ops_recv before gb_ops_send");
  GB_CHECK(nbytes % 16 == 0 && nbytes <= gb_ops_capacity(o) && stride % 16 == 0 && ((uintptr_t)d_out & 15) == 0,
           "gb_ops_recv: sizes must be multiples of 16 within the slot capacity");
  GB_CHECK(first >= 0 && count >= 1 && first + count <= c->world, "gb_ops_recv: rank range");
  OpsPeers dst, src;
  ops_slots(o, o->epoch, &dst, &src);
  const size_t n16 = nbytes / 16;
  gb_ops_collect<<<ops_grid(n16), GB_THREADS, 0, (cudaStream_t)stream>>>(src, n16, first, count, stride / 16, mode, scale, (uint4*)d_out,
                                                                        gb_comm_peers(c), o->channel, o->epoch, c->status_dev,
                                                                        c->timeout_ns);
  GB_CUDA(cudaGetLastError());
  o->launches++;
  return 0;
}
