// K0 -- dense all-reduce over peer memory for the None / FP16 codecs at world size > 1.
//
// Reference: communicator/allreduce.py:8-13 (dist.all_reduce + div_) with compressor/none.py or
// fp16.py, i.e. NCCL ring/tree all-reduce.  Behind NVSwitch every peer is one hop away at full
// bandwidth, so the natural schedule is the direct two-shot one:
//   k0_scatter  every rank stores shard d of its gradient (fp32, or fp16 for the FP16 codec)
//               into rank d's arena                        egress (W-1)/W of the payload
//   k0_reduce   rank d adds the W copies of its shard in rank order, scales by 1/W and stores the
//               result into every rank's arena             egress (W-1)/W of the payload (fp32)
//   k0_gather   every rank copies the W reduced shards out of its arena into the gradient views
// Same numbers, bit for bit, as the all-gather formulation (decode each rank, add in rank order,
// multiply by 1/W) at 2(W-1)/W instead of (W-1) payloads of traffic.
#include "kernels.cuh"
#include <cuda_fp16.h>

struct K0Slots { char* slot[GB_MAX_WORLD]; int n; };
__device__ __forceinline__ void shift(K0Slots& o, size_t bytes) {
  for (int r = 0; r < o.n; ++r) o.slot[r] += bytes;
}

// slot layout: [shard copy: cap elements (fp32 or fp16)] [reduced piece: cap fp32]
template <bool HALF>
__global__ void __launch_bounds__(GB_THREADS) k0_scatter(const float* __restrict__ grad, int64_t padded, int64_t cap,
                                                         K0Slots mine_in_peer, uint32_t* done, PeerPtrs peers, int channel,
                                                         const StepCtl* ctl, size_t par_stride) {
  const uint32_t epoch = ctl->epoch;
  if (epoch & 1) shift(mine_in_peer, par_stride);
  const int64_t n4 = padded / 4;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = i * 4;
    const int d = (int)(e / cap);
    const int64_t local = e - (int64_t)d * cap;
    const float4 v = ld_stream_f4(grad + e);
    if (HALF) {
      const __half2 a = __floats2half2_rn(v.x, v.y), b = __floats2half2_rn(v.z, v.w);
      uint2 w;
      w.x = *reinterpret_cast<const uint32_t*>(&a);
      w.y = *reinterpret_cast<const uint32_t*>(&b);
      *reinterpret_cast<uint2*>(mine_in_peer.slot[d] + local * 2) = w;
    } else {
      *reinterpret_cast<float4*>(mine_in_peer.slot[d] + local * 4) = v;
    }
  }
  release_flags_last_block(done, peers, channel, epoch);
}

template <bool HALF>
__global__ void __launch_bounds__(GB_THREADS) k0_reduce(int64_t shard_len, int64_t cap, K0Slots in_mine, K0Slots mine_in_peer,
                                                        size_t result_off, float inv_world, uint32_t* done, PeerPtrs peers,
                                                        int chan_in, int chan_out, const StepCtl* ctl, size_t par_stride,
                                                        uint32_t* status, uint64_t timeout_ns) {
  const uint32_t epoch = ctl->epoch;
  if (epoch & 1) { shift(in_mine, par_stride); shift(mine_in_peer, par_stride); }
  wait_all_ranks(peers, chan_in, epoch, status, timeout_ns);
  const int64_t n4 = shard_len / 4;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = 0; r < in_mine.n; ++r) {
      float4 v;
      if (HALF) {
        const uint2 w = *reinterpret_cast<const uint2*>(in_mine.slot[r] + i * 8);
        const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&w.x));
        const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&w.y));
        v = make_float4(a.x, a.y, b.x, b.y);
      } else {
        v = *reinterpret_cast<const float4*>(in_mine.slot[r] + i * 16);
      }
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    if (inv_world != 1.f) {
      s.x = __fmul_rn(s.x, inv_world); s.y = __fmul_rn(s.y, inv_world);
      s.z = __fmul_rn(s.z, inv_world); s.w = __fmul_rn(s.w, inv_world);
    }
    for (int r = 0; r < mine_in_peer.n; ++r) *reinterpret_cast<float4*>(mine_in_peer.slot[r] + result_off + i * 16) = s;
  }
  release_flags_last_block(done, peers, chan_out, epoch);
}

__global__ void __launch_bounds__(GB_THREADS) k0_gather(float* out, int64_t padded, int64_t cap, K0Slots in_mine, size_t result_off,
                                                        PeerPtrs peers, int channel, const StepCtl* ctl, size_t par_stride,
                                                        uint32_t* status, uint64_t timeout_ns) {
  const uint32_t epoch = ctl->epoch;
  if (epoch & 1) shift(in_mine, par_stride);
  wait_all_ranks(peers, channel, epoch, status, timeout_ns);
  const int64_t n4 = padded / 4;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = i * 4;
    const int d = (int)(e / cap);
    const int64_t local = e - (int64_t)d * cap;
    *reinterpret_cast<float4*>(out + e) = *reinterpret_cast<const float4*>(in_mine.slot[d] + result_off + local * 4);
  }
}

#define LAUNCHED(p) do { (p)->launches++; GB_CUDA(cudaGetLastError()); } while (0)

static K0Slots slots_of(gb_plan* p, bool remote, uint32_t epoch = 0) {   // parity 0; kernels shift by the device epoch
  K0Slots s;
  gb_comm* c = p->comm;
  s.n = c->world;
  for (int r = 0; r < c->world; ++r)
    s.slot[r] = remote ? gb_slot_of(c, r, p->chan_off, p->slot_bytes, epoch & 1, c->rank)
                       : gb_slot_of(c, c->rank, p->chan_off, p->slot_bytes, epoch & 1, r);
  return s;
}

int64_t k0_shard_cap(int64_t padded, int world) { return ((padded + world - 1) / world + 127) / 128 * 128; }

size_t k0_slot_bytes(int64_t padded, int world, int bits) {
  const int64_t cap = k0_shard_cap(padded, world);
  return (size_t)cap * (bits / 8) + (size_t)cap * 4;
}

int k0_send(gb_plan* p, const float* d_grad, cudaStream_t st) {
  gb_comm* c = p->comm;
  GB_CHECK(((uintptr_t)d_grad & 15) == 0, "gradient buffer must be 16-byte aligned");
  ++p->epoch;
  const int64_t cap = k0_shard_cap(p->padded, c->world);
  const K0Slots out = slots_of(p, true);
  const int grid = (int)std::min<int64_t>((p->padded / 4 + GB_THREADS - 1) / GB_THREADS, 148 * 8);
  if (p->bits == 16) k0_scatter<true><<<grid, GB_THREADS, 0, st>>>(d_grad, p->padded, cap, out, p->d_done, gb_comm_peers(c), p->channel, p->d_ctl, p->par_stride);
  else k0_scatter<false><<<grid, GB_THREADS, 0, st>>>(d_grad, p->padded, cap, out, p->d_done, gb_comm_peers(c), p->channel, p->d_ctl, p->par_stride);
  LAUNCHED(p);
  return 0;
}

int k0_mid(gb_plan* p, cudaStream_t st) {
  gb_comm* c = p->comm;
  const uint32_t epoch = p->epoch;
  const int64_t cap = k0_shard_cap(p->padded, c->world);
  const int64_t start = (int64_t)c->rank * cap;
  int64_t len = std::min<int64_t>(cap, p->padded - start);
  if (len < 0) len = 0;
  const size_t result_off = (size_t)cap * (p->bits / 8);
  const float inv = p->prm.average ? 1.0f / (float)c->world : 1.0f;
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((len / 4 + GB_THREADS - 1) / GB_THREADS, 148 * 8));
  if (p->bits == 16)
    k0_reduce<true><<<grid, GB_THREADS, 0, st>>>(len, cap, slots_of(p, false), slots_of(p, true), result_off, inv, p->d_done, gb_comm_peers(c), p->channel, p->channel2, p->d_ctl, p->par_stride, c->status_dev, c->timeout_ns);
  else
    k0_reduce<false><<<grid, GB_THREADS, 0, st>>>(len, cap, slots_of(p, false), slots_of(p, true), result_off, inv, p->d_done, gb_comm_peers(c), p->channel, p->channel2, p->d_ctl, p->par_stride, c->status_dev, c->timeout_ns);
  LAUNCHED(p);
  p->mid_epoch = epoch;
  return 0;
}

int k0_recv(gb_plan* p, float* d_out, cudaStream_t st) {
  gb_comm* c = p->comm;
  GB_CHECK(((uintptr_t)d_out & 15) == 0, "output buffer must be 16-byte aligned");
  if (p->mid_epoch != p->epoch) {           // caller skipped gb_plan_mid: run it now
    if (k0_mid(p, st)) return -1;
  }
  const int64_t cap = k0_shard_cap(p->padded, c->world);
  const size_t result_off = (size_t)cap * (p->bits / 8);
  const int grid = (int)std::min<int64_t>((p->padded / 4 + GB_THREADS - 1) / GB_THREADS, 148 * 8);
  k0_gather<<<grid, GB_THREADS, 0, st>>>(d_out, p->padded, cap, slots_of(p, false), result_off, gb_comm_peers(c), p->channel2, p->d_ctl, p->par_stride, c->status_dev, c->timeout_ns);
  LAUNCHED(p);
  return 0;
}

// ---------------------------------------------------------------- NVLS (in-switch reduction) variant
// When the host exposes NVLink SHARP multicast (NCCL_DEBUG=INFO shows "NVLS multicast support is
// available"), the gradient bucket itself lives in symmetric memory bound to a multicast address.
// One kernel per bucket then does the whole all-reduce in place:
//   wait until every rank has signalled "my gradients are final"          (epoch flags, channel A)
//   for my shard:  v = multimem.ld_reduce.add.f32x4 [mc + i]  -- the switch adds all W copies
//                  multimem.st.f32x4 [mc + i] = v / W         -- and broadcasts the result to all W copies
//   last CTA: release "my shard is stored" to every rank; then wait for all W such flags (channel B)
// Traffic per GPU: N/W in, N/W out -- the NVSwitch does the W-fold fan-in/fan-out.
// Every rank receives the same bits (one rank computes each shard); the switch's summation order is
// its own, so the value can differ in the last ulp from the rank-ordered P2P path.
struct gb_nvls {
  gb_comm* comm;
  int64_t padded;
  int chan_a, chan_b;
  uint32_t epoch;
  uint32_t* d_done;
  StepCtl* d_ctl;
  int64_t launches;
};

__global__ void k0_nvls_signal(PeerPtrs peers, int channel, StepCtl* ctl) {
  __shared__ uint32_t s_epoch;
  if (threadIdx.x == 0) { ctl->epoch += 1; s_epoch = ctl->epoch; }     // device-side epoch: graph-replayable
  __syncthreads();
  const uint32_t epoch = s_epoch;
  if (threadIdx.x < peers.world) {
    __threadfence_system();
    st_release_sys(gb_flag(peers.base[threadIdx.x], channel, epoch & 1, peers.rank), epoch);
  }
}

__global__ void __launch_bounds__(GB_THREADS) k0_nvls_reduce(float* mc, int64_t start, int64_t len, float inv_world, uint32_t* done,
                                                             PeerPtrs peers, int chan_a, int chan_b, const StepCtl* ctl,
                                                             uint32_t* status, uint64_t timeout_ns) {
  const uint32_t epoch = ctl->epoch;
  wait_all_ranks(peers, chan_a, epoch, status, timeout_ns);
  const int64_t n4 = len / 4;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  // four switch reductions in flight per thread: a multimem.ld_reduce round trip through NVSwitch is several
  // microseconds, one outstanding request per thread leaves the links idle
  for (int64_t i0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i0 < n4; i0 += 4 * stride) {
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t i = i0 + u * stride;
      if (i < n4) {
        float* p = mc + start + i * 4;
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(v[u].x), "=f"(v[u].y), "=f"(v[u].z), "=f"(v[u].w) : "l"(p) : "memory");
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t i = i0 + u * stride;
      if (i < n4) {
        float* p = mc + start + i * 4;
        const float4 w = make_float4(v[u].x * inv_world, v[u].y * inv_world, v[u].z * inv_world, v[u].w * inv_world);
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(w.x), "f"(w.y), "f"(w.z), "f"(w.w)
                     : "memory");
      }
    }
  }
  // publish "my shard is stored", then wait until every rank's shard is
  __threadfence_system();
  __syncthreads();
  __shared__ uint32_t last;
  if (threadIdx.x == 0) last = (atomicAdd(done, 1u) == gridDim.x - 1) ? 1u : 0u;
  __syncthreads();
  if (last) {
    if (threadIdx.x == 0) *done = 0;
    if (threadIdx.x < peers.world) {
      __threadfence_system();
      st_release_sys(gb_flag(peers.base[threadIdx.x], chan_b, epoch & 1, peers.rank), epoch);
      wait_flag(gb_flag(peers.base[peers.rank], chan_b, epoch & 1, threadIdx.x), epoch, status, timeout_ns);
    }
  }
}

extern "C" int gb_nvls_create(gb_comm* c, int64_t padded, gb_nvls** out) {
  GB_CHECK(padded % 128 == 0, "bucket length must be a multiple of 128 elements");
  gb_nvls* p = new gb_nvls();
  p->comm = c; p->padded = padded; p->epoch = 0; p->launches = 0;
  size_t unused;
  if (gb_comm_alloc_channel(c, 0, &p->chan_a, &unused)) return -1;
  if (gb_comm_alloc_channel(c, 0, &p->chan_b, &unused)) return -1;
  GB_CUDA(cudaMalloc((void**)&p->d_done, 128));
  GB_CUDA(cudaMemset(p->d_done, 0, 128));
  p->d_ctl = (StepCtl*)(p->d_done + 16);
  *out = p;
  return 0;
}
extern "C" int gb_nvls_destroy(gb_nvls* p) {
  if (!p) return 0;
  cudaDeviceSynchronize();
  cudaFree(p->d_done);
  delete p;
  return 0;
}
// "my copy of the bucket is final" -- enqueue right after the last gradient of the bucket
extern "C" int gb_nvls_signal(gb_nvls* p, void* stream) {
  ++p->epoch;
  k0_nvls_signal<<<1, 32, 0, (cudaStream_t)stream>>>(gb_comm_peers(p->comm), p->chan_a, p->d_ctl);
  LAUNCHED(p);
  return 0;
}
// in-place mean over all ranks of the multicast-bound bucket; returns (in stream order) once every shard is in place
extern "C" int gb_nvls_reduce(gb_nvls* p, void* d_multicast, int average, void* stream) {
  gb_comm* c = p->comm;
  GB_CHECK(p->epoch > 0, "gb_nvls_reduce before gb_nvls_signal");
  const int64_t cap = k0_shard_cap(p->padded, c->world);
  const int64_t start = (int64_t)c->rank * cap;
  int64_t len = std::min<int64_t>(cap, p->padded - start);
  if (len < 0) len = 0;
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((len / 4 + GB_THREADS - 1) / GB_THREADS, 148 * 4));
  k0_nvls_reduce<<<grid, GB_THREADS, 0, (cudaStream_t)stream>>>((float*)d_multicast, start, len, average ? 1.0f / (float)c->world : 1.0f,
                                                               p->d_done, gb_comm_peers(c), p->chan_a, p->chan_b, p->d_ctl,
                                                               c->status_dev, c->timeout_ns);
  LAUNCHED(p);
  return 0;
}
extern "C" int gb_nvls_launches(gb_nvls* p, int64_t* out) { *out = p->launches; return 0; }
