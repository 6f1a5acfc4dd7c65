// K2-sketch -- SketchML quantile-bucket codec (TensorFlow-only in the reference:
// grace_dl/tensorflow/compressor/sketch.py:6-39; restated in grace_b200/compressor/quant.py).
//
// Reference op chain per tensor: tfp.stats.quantiles (a full sort), find_bins, unsorted_segment_mean,
// cast; receive: gather means[bins] per rank + adds.  Here, per bucket:
//   sk_prep    x = compensate(grad, residual) -> keys                      (one pass)
//   cub::DeviceRadixSort::SortKeys on 64-bit keys (tensor id << 32 | order-preserving float bits): ONE sort
//              for all tensors of the bucket; only 32 + log2(#tensors) key bits are sorted
//   sk_edges   q+1 cut points per tensor, linear interpolation between order statistics
//   sk_means   one CTA per (bucket, tensor): the elements of a bucket are a contiguous range of the sorted
//              tensor, so the mean is a fixed-order range sum (deterministic, no float atomics);
//              means go straight into every peer's slot header
//   sk_encode  bin = largest b with edges[b] <= x (7-step search in shared memory), 4 codes per 32-bit
//              word, stores into every peer, residual = x - means[bin]; last CTA releases the flags
//   sk_decode  wait flags; sum_r means_r[code_r] in rank order, x 1/W
#include "kernels.cuh"

#include <cub/device/device_radix_sort.cuh>

#define SK_MAXQ 255

struct SkSlots { char* slot[GB_MAX_WORLD]; int n; };
struct SkIn { const char* slot[GB_MAX_WORLD]; int n; };
__device__ __forceinline__ void shift(SkSlots& o, size_t bytes) { for (int r = 0; r < o.n; ++r) o.slot[r] += bytes; }
__device__ __forceinline__ void shift(SkIn& o, size_t bytes) { for (int r = 0; r < o.n; ++r) o.slot[r] += bytes; }

__device__ __forceinline__ uint32_t flip_f(float x) {            // order-preserving float -> uint32
  const uint32_t b = __float_as_uint(x);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float unflip_f(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k);
}
__device__ __forceinline__ float sorted_at(const uint64_t* s, int64_t i) { return unflip_f((uint32_t)s[i]); }

// x = compensate(grad, residual) -> vals (kept for the encode pass) and 64-bit sort keys; pad elements of a
// tensor sort behind its real elements
__global__ void __launch_bounds__(GB_THREADS) sk_prep(SegTable t, const float* __restrict__ grad, const float* __restrict__ res,
                                                     float* __restrict__ vals, uint64_t* __restrict__ keys, int has_mem, float beta,
                                                     float gamma, const StepCtl* ctl) {
  if (ctl->step == 0) { beta = 0.f; gamma = 1.f; }
  for (int tile = blockIdx.x; tile < t.ntiles; tile += gridDim.x) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, tile, seg, base, local0, len);
    const int padded_len = min(GB_TILE, (int)(t.seg_off[seg + 1] - base));      // tile incl. the tensor's tail padding
    const uint64_t hi = (uint64_t)seg << 32;
    for (int e = threadIdx.x; e < padded_len; e += blockDim.x) {
      float x = 0.f;
      uint32_t k = 0xFFFFFFFFu;
      if (e < len) {
        x = grad[base + e];
        if (has_mem) x = __fadd_rn(__fmul_rn(beta, res[base + e]), __fmul_rn(gamma, x));
        k = flip_f(x);
      }
      vals[base + e] = x;
      keys[base + e] = hi | k;
    }
  }
}

// edges[seg][0..q]: srt[lo] + (srt[hi] - srt[lo]) * frac with pos = j * ((n-1)/q)  (quant.py: linear_quantile_edges)
__global__ void sk_edges(SegTable t, const uint64_t* __restrict__ sorted, int q, float* edges) {
  const int seg = blockIdx.x;
  const int n = t.seg_numel[seg];
  const uint64_t* s = sorted + t.seg_off[seg];
  const double step = (double)(n - 1) / (double)q;
  for (int j = threadIdx.x; j <= q; j += blockDim.x) {
    const double pos = (double)j * step;
    const double lo = floor(pos);
    int hi = (int)ceil(pos);
    if (hi > n - 1) hi = n - 1;
    const float frac = (float)(pos - lo);
    const float a = sorted_at(s, (int)lo), b = sorted_at(s, hi);
    edges[(size_t)seg * (SK_MAXQ + 1) + j] = __fadd_rn(a, __fmul_rn(__fsub_rn(b, a), frac));
  }
}

__device__ __forceinline__ int lower_bound_f(const uint64_t* s, int n, float v) {   // first i with s[i] >= v
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (sorted_at(s, mid) < v) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// grid (q, nseg): mean of bucket b of tensor seg; written to means[seg][b] and every peer's header
__global__ void __launch_bounds__(128) sk_means(SegTable t, const uint64_t* __restrict__ sorted, const float* __restrict__ edges, int q,
                                                float* means, SkSlots out, const StepCtl* ctl, size_t par_stride) {
  __shared__ double sh[4];
  __shared__ int s_lo, s_hi;
  if (ctl->epoch & 1) shift(out, par_stride);
  const int b = blockIdx.x, seg = blockIdx.y;
  const int n = t.seg_numel[seg];
  const uint64_t* s = sorted + t.seg_off[seg];
  const float* e = edges + (size_t)seg * (SK_MAXQ + 1);
  if (threadIdx.x == 0) {
    // bucket b = { x : b is the largest index (< q) with e[b] <= x }
    int bb = b;
    while (bb + 1 < q && e[bb + 1] == e[b]) ++bb;          // equal cut points: the highest index owns the values
    if (bb != b) { s_lo = 0; s_hi = 0; }
    else {
      s_lo = b == 0 ? 0 : lower_bound_f(s, n, e[b]);
      s_hi = b == q - 1 ? n : lower_bound_f(s, n, e[b + 1]);
      // cut points equal to e[b] at lower indices were skipped above; values equal to e[b+1] belong to b+1
    }
  }
  __syncthreads();
  const int lo = s_lo, hi = s_hi;
  double acc = 0;
  for (int i = lo + threadIdx.x; i < hi; i += blockDim.x) acc += (double)sorted_at(s, i);
  acc = warp_sum_d(acc);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    const double tot = sh[0] + sh[1] + sh[2] + sh[3];
    const float m = hi > lo ? (float)(tot / (double)(hi - lo)) : 0.f;
    means[(size_t)seg * (SK_MAXQ + 1) + b] = m;
    for (int r = 0; r < out.n; ++r) reinterpret_cast<float*>(out.slot[r])[(size_t)seg * q + b] = m;
  }
}

__device__ __forceinline__ int sk_bin(const float* e, int q, float x) {
  int lo = 0, hi = q + 1;                       // upper_bound over the q+1 cut points
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (e[mid] <= x) lo = mid + 1; else hi = mid;
  }
  return min(max(lo - 1, 0), q - 1);
}

__global__ void __launch_bounds__(GB_THREADS) sk_encode(SegTable t, const float* __restrict__ keys, float* res, int has_mem,
                                                       const float* __restrict__ edges, const float* __restrict__ means, int q,
                                                       SkSlots out, size_t hdr_bytes, uint32_t* done, PeerPtrs peers, int channel,
                                                       const StepCtl* ctl, size_t par_stride) {
  __shared__ float se[SK_MAXQ + 1], sm[SK_MAXQ + 1];
  const uint32_t epoch = ctl->epoch;
  if (epoch & 1) shift(out, par_stride);
  int cur = -1;
  for (int tile = blockIdx.x; tile < t.ntiles; tile += gridDim.x) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, tile, seg, base, local0, len);
    if (seg != cur) {
      __syncthreads();
      for (int i = threadIdx.x; i <= q; i += blockDim.x) {
        se[i] = edges[(size_t)seg * (SK_MAXQ + 1) + i];
        sm[i] = i < q ? means[(size_t)seg * (SK_MAXQ + 1) + i] : 0.f;
      }
      __syncthreads();
      cur = seg;
    }
#pragma unroll
    for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j) {
      const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
      if (e0 >= len) continue;
      const float4 v = *reinterpret_cast<const float4*>(keys + base + e0);
      const float x[4] = {v.x, v.y, v.z, v.w};
      uint32_t word = 0;
      float nres[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const bool live = e0 + k < len;
        const int b = live ? sk_bin(se, q, x[k]) : 0;
        word |= (uint32_t)b << (8 * k);
        nres[k] = live ? __fsub_rn(x[k], sm[b]) : 0.f;
      }
      if (has_mem) *reinterpret_cast<float4*>(res + base + e0) = make_float4(nres[0], nres[1], nres[2], nres[3]);
      const int64_t w = (base + e0) >> 2;
      for (int r = 0; r < out.n; ++r) reinterpret_cast<uint32_t*>(out.slot[r] + hdr_bytes)[w] = word;
    }
  }
  release_flags_last_block(done, peers, channel, epoch);
}

__global__ void __launch_bounds__(GB_THREADS) sk_decode(SegTable t, SkIn in, size_t hdr_bytes, int q, float* outp, float divisor,
                                                       PeerPtrs peers, int channel, const StepCtl* ctl, size_t par_stride,
                                                       uint32_t* status, uint64_t timeout_ns) {
  __shared__ float sm[GB_MAX_WORLD][SK_MAXQ + 1];
  const uint32_t epoch = ctl->epoch;
  if (epoch & 1) shift(in, par_stride);
  wait_all_ranks(peers, channel, epoch, status, timeout_ns);
  const float inv = divisor != 1.f ? __fdiv_rn(1.0f, divisor) : 1.f;
  int cur = -1;
  for (int tile = blockIdx.x; tile < t.ntiles; tile += gridDim.x) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, tile, seg, base, local0, len);
    if (seg != cur) {
      __syncthreads();
      for (int r = 0; r < in.n; ++r)
        for (int i = threadIdx.x; i < q; i += blockDim.x) sm[r][i] = reinterpret_cast<const float*>(in.slot[r])[(size_t)seg * q + i];
      __syncthreads();
      cur = seg;
    }
#pragma unroll
    for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j) {
      const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
      if (e0 >= len) continue;
      const int64_t w = (base + e0) >> 2;
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
      for (int r = 0; r < in.n; ++r) {
        const uint32_t word = reinterpret_cast<const uint32_t*>(in.slot[r] + hdr_bytes)[w];
#pragma unroll
        for (int k = 0; k < 4; ++k) acc[k] += sm[r][(word >> (8 * k)) & 0xFFu];
      }
      float o[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) o[k] = (e0 + k < len) ? (divisor != 1.f ? __fmul_rn(acc[k], inv) : acc[k]) : 0.f;
      *reinterpret_cast<float4*>(outp + base + e0) = make_float4(o[0], o[1], o[2], o[3]);
    }
  }
}

// ---------------------------------------------------------------- host side
#define LAUNCHED(p) do { (p)->launches++; GB_CUDA(cudaGetLastError()); } while (0)

struct SketchBufs {          // lives in gb_plan::sketch (allocated by sk_init)
  float *vals, *edges, *means;
  uint64_t *keys, *sorted;
  void* temp;
  size_t temp_bytes;
  int end_bit;
};

size_t sk_header_bytes(int nseg, int q) { return ((size_t)nseg * q * 4 + 255) & ~size_t(255); }

int sk_init(gb_plan* p) {
  const int q = p->prm.quantum_num;
  GB_CHECK(q >= 2 && q <= SK_MAXQ, "native SketchML supports 2..%d quantiles (got %d)", SK_MAXQ, q);
  SketchBufs* b = new SketchBufs();
  p->sketch = b;
  GB_CUDA(cudaMalloc((void**)&b->vals, p->padded * 4));
  GB_CUDA(cudaMalloc((void**)&b->keys, p->padded * 8));
  GB_CUDA(cudaMalloc((void**)&b->sorted, p->padded * 8));
  GB_CUDA(cudaMalloc((void**)&b->edges, (size_t)p->nseg * (SK_MAXQ + 1) * 4));
  GB_CUDA(cudaMalloc((void**)&b->means, (size_t)p->nseg * (SK_MAXQ + 1) * 4));
  b->end_bit = 32;
  while ((1ll << (b->end_bit - 32)) < p->nseg) ++b->end_bit;
  b->temp = nullptr;
  b->temp_bytes = 0;
  GB_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, b->temp_bytes, (const uint64_t*)b->keys, b->sorted, (int)p->padded, 0, b->end_bit));
  GB_CUDA(cudaMalloc(&b->temp, b->temp_bytes + 256));
  return 0;
}

void sk_free(gb_plan* p) {
  SketchBufs* b = (SketchBufs*)p->sketch;
  if (!b) return;
  cudaFree(b->vals); cudaFree(b->keys); cudaFree(b->sorted); cudaFree(b->edges); cudaFree(b->means); cudaFree(b->temp);
  delete b;
  p->sketch = nullptr;
}

static int grid_for(int ntiles) { return ntiles < 148 * 8 ? ntiles : 148 * 8; }

int sk_send(gb_plan* p, const float* d_grad, cudaStream_t st) {
  gb_comm* c = p->comm;
  SketchBufs* b = (SketchBufs*)p->sketch;
  const int q = p->prm.quantum_num;
  const int has_mem = p->prm.memory == GB_MEM_RESIDUAL;
  ++p->epoch;
  SkSlots out;
  out.n = c->world;
  for (int r = 0; r < c->world; ++r) out.slot[r] = gb_slot_of(c, r, p->chan_off, p->slot_bytes, 0, c->rank);
  sk_prep<<<grid_for(p->ntiles), GB_THREADS, 0, st>>>(p->tab, d_grad, p->d_state0, b->vals, b->keys, has_mem, p->prm.beta, p->prm.gamma,
                                                     p->d_ctl);
  LAUNCHED(p);
  size_t tb = b->temp_bytes;
  GB_CUDA(cub::DeviceRadixSort::SortKeys(b->temp, tb, (const uint64_t*)b->keys, b->sorted, (int)p->padded, 0, b->end_bit, st));
  p->launches += (b->end_bit + 7) / 8 + 2;          // histogram + one onesweep pass per 8 key bits
  sk_edges<<<p->nseg, 128, 0, st>>>(p->tab, b->sorted, q, b->edges); LAUNCHED(p);
  sk_means<<<dim3(q, p->nseg), 128, 0, st>>>(p->tab, b->sorted, b->edges, q, b->means, out, p->d_ctl, p->par_stride); LAUNCHED(p);
  sk_encode<<<grid_for(p->ntiles), GB_THREADS, 0, st>>>(p->tab, b->vals, p->d_state0, has_mem, b->edges, b->means, q, out, p->hdr_bytes,
                                                       p->d_done, gb_comm_peers(c), p->channel, p->d_ctl, p->par_stride);
  LAUNCHED(p);
  return 0;
}

int sk_recv(gb_plan* p, float* d_out, cudaStream_t st) {
  gb_comm* c = p->comm;
  SkIn in;
  in.n = c->world;
  for (int r = 0; r < c->world; ++r) in.slot[r] = gb_slot_of(c, c->rank, p->chan_off, p->slot_bytes, 0, r);
  sk_decode<<<grid_for(p->ntiles), GB_THREADS, 0, st>>>(p->tab, in, p->hdr_bytes, p->prm.quantum_num, d_out,
                                                       p->prm.average ? (float)c->world : 1.f, gb_comm_peers(c), p->channel, p->d_ctl,
                                                       p->par_stride, c->status_dev, c->timeout_ns);
  LAUNCHED(p);
  return 0;
}
