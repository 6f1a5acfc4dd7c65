// K2-sketch -- SketchML quantile-bucket codec (TensorFlow-only in the reference:
// grace_dl/tensorflow/compressor/sketch.py:6-39; restated in grace_b200/compressor/quant.py).
//
// Reference op chain per tensor: tfp.stats.quantiles (a full sort), find_bins, unsorted_segment_mean,
// cast; receive: gather means[bins] per rank + adds.  Here, per bucket:
//   sk_prep    x = compensate(grad, residual) -> keys                      (one pass)
//   rs_hist / rs_scan / rs_scatter: a hand-written stable LSD radix sort (8-bit digits, 4 passes) of the
//              order-preserving 32-bit float keys, SEGMENTED: every tensor of the bucket is sorted in place inside its
//              own range of the flat buffer -- a sort block is a plan tile, which never straddles two tensors, so the
//              digit prefixes are taken over the tiles of one tensor and no tensor id has to travel in the key (the
//              first version sorted 64-bit (tensor id, key) words of the whole bucket in 5 passes: twice the bytes).
//              No library kernel is on this path (round 1 called cub::DeviceRadixSort here).
//   sk_edges   q+1 cut points per tensor, linear interpolation between order statistics, and the first sorted position
//              at or above every cut point (q+1 binary searches side by side)
//   sk_means   one CTA per (bucket, tensor): the elements of a bucket are a contiguous range of the sorted
//              tensor, so the mean is a fixed-order range sum (deterministic, no float atomics);
//              means go straight into every peer's slot header
//   sk_encode  bin = largest b with edges[b] <= x (binary search in shared memory), 8-bit codes (q < 256) or
//              16-bit codes (q <= 1023; reference: uint16 bins, tensorflow/compressor/sketch.py:26) packed into
//              32-bit words, stores into every peer, residual = x - means[bin]; last CTA releases the flags
//   sk_decode  wait flags; sum_r means_r[code_r] in rank order, x 1/W
#include "kernels.cuh"

#include <utility>

#define SK_MAXQ 1023

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
__device__ __forceinline__ float sorted_at(const uint32_t* s, int64_t i) { return unflip_f(s[i]); }
__device__ __forceinline__ float4 sk_ld4(const float* p, int e0, int len) {      // p + e0 is 16-byte aligned
  if (e0 + 4 <= len) return *reinterpret_cast<const float4*>(p + e0);
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (e0 < len) v.x = p[e0];
  if (e0 + 1 < len) v.y = p[e0 + 1];
  if (e0 + 2 < len) v.z = p[e0 + 2];
  if (e0 + 3 < len) v.w = p[e0 + 3];
  return v;
}

// ---------------------------------------------------------------- stable segmented LSD radix sort, 8 bits per pass
// A sort block is a plan tile (<= 4096 keys of ONE tensor, the tensor's tail padding included in its last tile).
// rs_hist counts the block's digits into counts[tile][digit]; rs_scan (one CTA per tensor, one thread per digit) turns
// them into exclusive prefixes over the tensor's tiles and the tensor's exclusive digit offsets; rs_scatter ranks
// every key among the equal digits that precede it (warp match + per-warp counts, rounds of 256 keys in index order
// => stable) and stores it at  tensor offset + digit offset + prefix[tile][digit] + rank.
__device__ __forceinline__ int sk_tile_keys(const SegTable& t, int tile, int64_t& base) {
  int seg, local0, len;
  tile_bounds(t, tile, seg, base, local0, len);
  return min(GB_TILE, (int)(t.seg_off[seg + 1] - base));
}

__global__ void __launch_bounds__(GB_THREADS) rs_hist(SegTable t, const uint32_t* __restrict__ keys, int shift, uint32_t* counts) {
  __shared__ uint32_t sh[256];
  sh[threadIdx.x] = 0;
  __syncthreads();
  int64_t base;
  const int n = sk_tile_keys(t, blockIdx.x, base);
  for (int i = threadIdx.x * 4; i < n; i += GB_THREADS * 4) {            // tile bases and lengths are multiples of 4
    const uint4 k = *reinterpret_cast<const uint4*>(keys + base + i);
    atomicAdd(&sh[(k.x >> shift) & 255u], 1u);
    atomicAdd(&sh[(k.y >> shift) & 255u], 1u);
    atomicAdd(&sh[(k.z >> shift) & 255u], 1u);
    atomicAdd(&sh[(k.w >> shift) & 255u], 1u);
  }
  __syncthreads();
  counts[(size_t)blockIdx.x * 256 + threadIdx.x] = sh[threadIdx.x];
}

// grid: one CTA per tensor, RS_GROUPS x 256 threads; thread (g, d) owns digit d over the g-th quarter of the tensor's
// tiles: first the quarter's digit totals, then -- started from the totals of the quarters before it -- the exclusive
// prefixes (a tensor of 576 tiles walked by one group kept the launch at 65 us).
// digit_base[seg][d] = keys of the tensor with a smaller digit.
#define RS_GROUPS 4
__global__ void __launch_bounds__(RS_GROUPS * 256) rs_scan(SegTable t, uint32_t* counts, uint32_t* digit_base) {
  __shared__ uint32_t s_tot[RS_GROUPS][256];
  __shared__ uint32_t sh[256];
  const int seg = blockIdx.x, d = threadIdx.x & 255, g = threadIdx.x >> 8;
  const int t0 = t.seg_tile0[seg], t1 = t.seg_tile0[seg + 1];
  const int per = (t1 - t0 + RS_GROUPS - 1) / RS_GROUPS;
  const int a0 = min(t1, t0 + g * per), a1 = min(t1, a0 + per);
  uint32_t sum = 0;
  int tile = a0;
  for (; tile + 8 <= a1; tile += 8) {                                   // eight independent loads in flight
    uint32_t v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) v[u] = counts[(size_t)(tile + u) * 256 + d];
#pragma unroll
    for (int u = 0; u < 8; ++u) sum += v[u];
  }
  for (; tile < a1; ++tile) sum += counts[(size_t)tile * 256 + d];
  s_tot[g][d] = sum;
  __syncthreads();
  uint32_t run = 0, total = 0;
#pragma unroll
  for (int k = 0; k < RS_GROUPS; ++k) {
    const uint32_t v = s_tot[k][d];
    if (k < g) run += v;
    total += v;
  }
  tile = a0;
  for (; tile + 8 <= a1; tile += 8) {
    uint32_t v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) v[u] = counts[(size_t)(tile + u) * 256 + d];
#pragma unroll
    for (int u = 0; u < 8; ++u) { counts[(size_t)(tile + u) * 256 + d] = run; run += v[u]; }
  }
  for (; tile < a1; ++tile) {
    const uint32_t v = counts[(size_t)tile * 256 + d];
    counts[(size_t)tile * 256 + d] = run;
    run += v;
  }
  if (g == 0) sh[d] = total;
  __syncthreads();
  for (int o = 1; o < 256; o <<= 1) {
    const uint32_t a = (g == 0 && d >= o) ? sh[d - o] : 0u;
    __syncthreads();
    if (g == 0) sh[d] += a;
    __syncthreads();
  }
  if (g == 0) digit_base[(size_t)seg * 256 + d] = sh[d] - total;
}

// Ranking without block barriers: warp w owns keys [512 w, 512 w + 512) of the block and walks them 32 at a time;
// a key's rank among the warp's earlier equal digits is the warp's running counter for that digit plus its position in
// the match mask.  One block-wide step then turns the 8 x 256 counters into exclusive offsets in (digit, warp) order,
// the keys are placed in shared memory in that order (= stable), and every thread copies keys i, i + 256, ... to their
// final places: the keys of one digit are consecutive in both, so a warp stores contiguous runs.
// (First version: 16 block-wide rounds of 256 keys with four barriers each and 4-byte stores straight from the round:
// 240 us per pass for 102 MB.)
__global__ void __launch_bounds__(GB_THREADS, 4) rs_scatter(SegTable t, const uint32_t* __restrict__ keys, uint32_t* __restrict__ out, int shift,
                                                        const uint32_t* __restrict__ prefix, const uint32_t* __restrict__ digit_base) {
  __shared__ uint32_t s_key[GB_TILE];            // the block's keys ordered by digit (stable)
  __shared__ uint32_t s_dst[256];                // final position of the block's first key with digit d, minus its offset in s_key
  __shared__ uint32_t s_scan[256];
  __shared__ uint32_t s_cnt[GB_THREADS / 32][256];
  int seg, local0, len;
  int64_t base;
  tile_bounds(t, blockIdx.x, seg, base, local0, len);
  const int n = min(GB_TILE, (int)(t.seg_off[seg + 1] - base));
  uint32_t* dst = out + t.seg_off[seg];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int R = GB_TILE / GB_THREADS;            // 16 keys per thread
  constexpr int WK = GB_TILE / (GB_THREADS / 32);    // 512 keys per warp
  for (int w = 0; w < GB_THREADS / 32; ++w) s_cnt[w][threadIdx.x] = 0;
  uint32_t key[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int i = warp * WK + r * 32 + lane;
    key[r] = i < n ? keys[base + i] : 0u;
  }
  __syncthreads();
  uint32_t rank[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const bool live = warp * WK + r * 32 + lane < n;
    const uint32_t dg = (key[r] >> shift) & 255u;
    // lanes with the same digit: eight ballots, one per digit bit (MATCH.ANY occupies the SM's ADU pipe for ~64 cycles
    // per warp instruction -- ncu: that pipe 73 % busy, 250 us per pass; a ballot is a few cycles)
    uint32_t same = __ballot_sync(0xffffffffu, live);
#pragma unroll
    for (int bit = 0; bit < 8; ++bit) {
      const uint32_t bal = __ballot_sync(0xffffffffu, (dg >> bit) & 1u);
      same &= ((dg >> bit) & 1u) ? bal : ~bal;
    }
    const uint32_t before = live ? s_cnt[warp][dg] : 0u;
    rank[r] = before + (uint32_t)__popc(same & ((1u << lane) - 1u));
    __syncwarp();
    if (live && (same & ((1u << lane) - 1u)) == 0u) s_cnt[warp][dg] = before + (uint32_t)__popc(same);
    __syncwarp();
  }
  __syncthreads();
  {
    // thread d: counters of digit d over the warps -> exclusive offsets in (digit, warp) order
    uint32_t c[GB_THREADS / 32], tot = 0;
#pragma unroll
    for (int w = 0; w < GB_THREADS / 32; ++w) { c[w] = s_cnt[w][threadIdx.x]; tot += c[w]; }
    s_scan[threadIdx.x] = tot;
    __syncthreads();
    for (int o = 1; o < 256; o <<= 1) {
      const uint32_t a = threadIdx.x >= o ? s_scan[threadIdx.x - o] : 0u;
      __syncthreads();
      s_scan[threadIdx.x] += a;
      __syncthreads();
    }
    uint32_t run = s_scan[threadIdx.x] - tot;
    s_dst[threadIdx.x] = digit_base[(size_t)seg * 256 + threadIdx.x] + prefix[(size_t)blockIdx.x * 256 + threadIdx.x] - run;
#pragma unroll
    for (int w = 0; w < GB_THREADS / 32; ++w) { s_cnt[w][threadIdx.x] = run; run += c[w]; }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < R; ++r) {
    if (warp * WK + r * 32 + lane < n) s_key[s_cnt[warp][(key[r] >> shift) & 255u] + rank[r]] = key[r];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += GB_THREADS) {
    const uint32_t k = s_key[i];
    dst[s_dst[(k >> shift) & 255u] + (uint32_t)i] = k;
  }
}

// x = compensate(grad, residual) -> vals (kept for the encode pass) and 32-bit sort keys; pad elements of a
// tensor sort behind its real elements
__global__ void __launch_bounds__(GB_THREADS) sk_prep(SegTable t, const float* __restrict__ grad, const float* __restrict__ res,
                                                     float* __restrict__ vals, uint32_t* __restrict__ keys, int has_mem, float beta,
                                                     float gamma, const StepCtl* ctl) {
  if (ctl->step == 0) { beta = 0.f; gamma = 1.f; }
  for (int tile = blockIdx.x; tile < t.ntiles; tile += gridDim.x) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, tile, seg, base, local0, len);
    const int padded_len = min(GB_TILE, (int)(t.seg_off[seg + 1] - base));      // tile incl. the tensor's tail padding
    for (int e = threadIdx.x * 4; e < padded_len; e += GB_THREADS * 4) {        // bases and padded lengths: multiples of 4
      float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
      if (e < len) {
        x = sk_ld4(grad + base, e, len);
        if (has_mem) {
          const float4 r = sk_ld4(res + base, e, len);
          x.x = __fadd_rn(__fmul_rn(beta, r.x), __fmul_rn(gamma, x.x)); x.y = __fadd_rn(__fmul_rn(beta, r.y), __fmul_rn(gamma, x.y));
          x.z = __fadd_rn(__fmul_rn(beta, r.z), __fmul_rn(gamma, x.z)); x.w = __fadd_rn(__fmul_rn(beta, r.w), __fmul_rn(gamma, x.w));
        }
      }
      const uint4 k = make_uint4(e < len ? flip_f(x.x) : 0xFFFFFFFFu, e + 1 < len ? flip_f(x.y) : 0xFFFFFFFFu,
                                 e + 2 < len ? flip_f(x.z) : 0xFFFFFFFFu, e + 3 < len ? flip_f(x.w) : 0xFFFFFFFFu);
      if (e + 1 >= len) x.y = 0.f;
      if (e + 2 >= len) x.z = 0.f;
      if (e + 3 >= len) x.w = 0.f;
      *reinterpret_cast<float4*>(vals + base + e) = x;
      *reinterpret_cast<uint4*>(keys + base + e) = k;
    }
  }
}

__device__ __forceinline__ int lower_bound_f(const uint32_t* s, int n, float v) {   // first i with s[i] >= v
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (sorted_at(s, mid) < v) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// edges[seg][0..q]: srt[lo] + (srt[hi] - srt[lo]) * frac with pos = j * ((n-1)/q)  (quant.py: linear_quantile_edges),
// and bound[seg][j] = first sorted position with a value >= edges[j]: thread j does both, so the q + 1 binary searches
// of a tensor run side by side (they used to be done one after the other by thread 0 of every sk_means CTA: two
// 21-step chains of dependent loads in front of every bucket sum, 188 us for the launch)
__global__ void sk_edges(SegTable t, const uint32_t* __restrict__ sorted, int q, float* edges, int* bound) {
  const int seg = blockIdx.x;
  const int n = t.seg_numel[seg];
  const uint32_t* s = sorted + t.seg_off[seg];
  const double step = (double)(n - 1) / (double)q;
  for (int j = threadIdx.x; j <= q; j += blockDim.x) {
    const double pos = (double)j * step;
    const double lo = floor(pos);
    int hi = (int)ceil(pos);
    if (hi > n - 1) hi = n - 1;
    const float frac = (float)(pos - lo);
    const float a = sorted_at(s, (int)lo), b = sorted_at(s, hi);
    const float e = __fadd_rn(a, __fmul_rn(__fsub_rn(b, a), frac));
    edges[(size_t)seg * (SK_MAXQ + 1) + j] = e;
    bound[(size_t)seg * (SK_MAXQ + 1) + j] = j == 0 ? 0 : lower_bound_f(s, n, e);
  }
}

// grid (q, nseg): mean of bucket b of tensor seg; written to means[seg][b] and every peer's header
__global__ void __launch_bounds__(128) sk_means(SegTable t, const uint32_t* __restrict__ sorted, const float* __restrict__ edges,
                                                const int* __restrict__ bound, int q, float* means, SkSlots out, const StepCtl* ctl,
                                                size_t par_stride) {
  __shared__ double sh[4];
  if (ctl->epoch & 1) shift(out, par_stride);
  const int b = blockIdx.x, seg = blockIdx.y;
  const int n = t.seg_numel[seg];
  const uint32_t* s = sorted + t.seg_off[seg];
  const float* e = edges + (size_t)seg * (SK_MAXQ + 1);
  const int* bd = bound + (size_t)seg * (SK_MAXQ + 1);
  // bucket b = { x : b is the largest index (< q) with e[b] <= x }; equal cut points: the highest index owns the values
  int lo = 0, hi = 0;
  if (!(b + 1 < q && e[b + 1] == e[b])) {
    lo = bd[b];                          // cut points equal to e[b] at lower indices own nothing: lower_bound covers them
    hi = b == q - 1 ? n : bd[b + 1];     // values equal to e[b+1] belong to b+1
  }
  // 16-byte loads over the aligned middle of [lo, hi) (the tensor's range starts on a 512-byte boundary), four of
  // them in flight per thread; a big tensor's bucket is 150 KB and one 4-byte load per thread and trip kept the
  // launch at 158 us
  double acc = 0;
  const int a0 = min(hi, (lo + 3) & ~3), a1 = max(a0, hi & ~3);
  for (int i = lo + threadIdx.x; i < a0; i += blockDim.x) acc += (double)sorted_at(s, i);
  for (int i = a1 + threadIdx.x; i < hi; i += blockDim.x) acc += (double)sorted_at(s, i);
  const uint4* s4 = reinterpret_cast<const uint4*>(s + a0);
  const int n4 = (a1 - a0) >> 2;
  int i = threadIdx.x;
  for (; i + 3 * (int)blockDim.x < n4; i += 4 * blockDim.x) {
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = s4[i + u * blockDim.x];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      acc += ((double)unflip_f(v[u].x) + (double)unflip_f(v[u].y)) + ((double)unflip_f(v[u].z) + (double)unflip_f(v[u].w));
  }
  for (; i < n4; i += blockDim.x) {
    const uint4 v = s4[i];
    acc += ((double)unflip_f(v.x) + (double)unflip_f(v.y)) + ((double)unflip_f(v.z) + (double)unflip_f(v.w));
  }
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

template <int CB>   // code bits: 8 (q < 256) or 16
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
      uint32_t word[2] = {0u, 0u};
      float nres[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const bool live = e0 + k < len;
        const int b = live ? sk_bin(se, q, x[k]) : 0;
        if (CB == 8) word[0] |= (uint32_t)b << (8 * k);
        else word[k >> 1] |= (uint32_t)b << (16 * (k & 1));
        nres[k] = live ? __fsub_rn(x[k], sm[b]) : 0.f;
      }
      if (has_mem) *reinterpret_cast<float4*>(res + base + e0) = make_float4(nres[0], nres[1], nres[2], nres[3]);
      if (CB == 8) {
        const int64_t w = (base + e0) >> 2;
        for (int r = 0; r < out.n; ++r) reinterpret_cast<uint32_t*>(out.slot[r] + hdr_bytes)[w] = word[0];
      } else {
        const int64_t w = (base + e0) >> 1;
        for (int r = 0; r < out.n; ++r) *reinterpret_cast<uint2*>(reinterpret_cast<uint32_t*>(out.slot[r] + hdr_bytes) + w) = make_uint2(word[0], word[1]);
      }
    }
  }
  release_flags_last_block(done, peers, channel, epoch);
}

template <int CB>
__global__ void __launch_bounds__(GB_THREADS) sk_decode(SegTable t, SkIn in, size_t hdr_bytes, int q, float* outp, float divisor,
                                                       PeerPtrs peers, int channel, const StepCtl* ctl, size_t par_stride,
                                                       uint32_t* status, uint64_t timeout_ns) {
  constexpr int QCAP = CB == 8 ? 256 : SK_MAXQ + 1;
  __shared__ float sm[GB_MAX_WORLD][QCAP];
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
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
      for (int r = 0; r < in.n; ++r) {
        if (CB == 8) {
          const uint32_t word = reinterpret_cast<const uint32_t*>(in.slot[r] + hdr_bytes)[(base + e0) >> 2];
#pragma unroll
          for (int k = 0; k < 4; ++k) acc[k] += sm[r][(word >> (8 * k)) & 0xFFu];
        } else {
          const uint2 word = *reinterpret_cast<const uint2*>(reinterpret_cast<const uint32_t*>(in.slot[r] + hdr_bytes) + ((base + e0) >> 1));
          acc[0] += sm[r][word.x & 0xFFFFu]; acc[1] += sm[r][word.x >> 16];
          acc[2] += sm[r][word.y & 0xFFFFu]; acc[3] += sm[r][word.y >> 16];
        }
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
  int* bound;                  // [nseg][SK_MAXQ + 1] first sorted position >= edge
  uint32_t *keys, *sorted;
  uint32_t *counts, *totals;   // radix sort: [ntiles][256] digit counts -> prefixes over a tensor's tiles, [nseg][256] digit offsets
};

size_t sk_header_bytes(int nseg, int q) { return ((size_t)nseg * q * 4 + 255) & ~size_t(255); }

int sk_init(gb_plan* p) {
  const int q = p->prm.quantum_num;
  GB_CHECK(q >= 2 && q <= SK_MAXQ, "native SketchML supports 2..%d quantiles (got %d)", SK_MAXQ, q);
  SketchBufs* b = new SketchBufs();
  p->sketch = b;
  GB_CUDA(cudaMalloc((void**)&b->vals, p->padded * 4));
  GB_CUDA(cudaMalloc((void**)&b->keys, p->padded * 4));
  GB_CUDA(cudaMalloc((void**)&b->sorted, p->padded * 4));
  GB_CUDA(cudaMalloc((void**)&b->edges, (size_t)p->nseg * (SK_MAXQ + 1) * 4));
  GB_CUDA(cudaMalloc((void**)&b->means, (size_t)p->nseg * (SK_MAXQ + 1) * 4));
  GB_CUDA(cudaMalloc((void**)&b->bound, (size_t)p->nseg * (SK_MAXQ + 1) * 4));
  GB_CUDA(cudaMalloc((void**)&b->counts, (size_t)256 * p->ntiles * 4));
  GB_CUDA(cudaMalloc((void**)&b->totals, (size_t)256 * p->nseg * 4));
  return 0;
}

void sk_free(gb_plan* p) {
  SketchBufs* b = (SketchBufs*)p->sketch;
  if (!b) return;
  cudaFree(b->vals); cudaFree(b->keys); cudaFree(b->sorted); cudaFree(b->edges); cudaFree(b->means); cudaFree(b->bound); cudaFree(b->counts); cudaFree(b->totals);
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
  // stable segmented LSD radix sort of the 32-bit keys, 8 bits per pass, ping-pong between keys and sorted
  uint32_t* src = b->keys;
  uint32_t* dst = b->sorted;
  for (int shift = 0; shift < 32; shift += 8) {
    rs_hist<<<p->ntiles, GB_THREADS, 0, st>>>(p->tab, src, shift, b->counts); LAUNCHED(p);
    rs_scan<<<p->nseg, RS_GROUPS * 256, 0, st>>>(p->tab, b->counts, b->totals); LAUNCHED(p);
    rs_scatter<<<p->ntiles, GB_THREADS, 0, st>>>(p->tab, src, dst, shift, b->counts, b->totals); LAUNCHED(p);
    std::swap(src, dst);
  }
  const uint32_t* sorted = src;                       // after the last swap `src` holds the sorted keys
  sk_edges<<<p->nseg, 128, 0, st>>>(p->tab, sorted, q, b->edges, b->bound); LAUNCHED(p);
  sk_means<<<dim3(q, p->nseg), 128, 0, st>>>(p->tab, sorted, b->edges, b->bound, q, b->means, out, p->d_ctl, p->par_stride); LAUNCHED(p);
  if (q < 256)
    sk_encode<8><<<grid_for(p->ntiles), GB_THREADS, 0, st>>>(p->tab, b->vals, p->d_state0, has_mem, b->edges, b->means, q, out, p->hdr_bytes,
                                                            p->d_done, gb_comm_peers(c), p->channel, p->d_ctl, p->par_stride);
  else
    sk_encode<16><<<grid_for(p->ntiles), GB_THREADS, 0, st>>>(p->tab, b->vals, p->d_state0, has_mem, b->edges, b->means, q, out, p->hdr_bytes,
                                                             p->d_done, gb_comm_peers(c), p->channel, p->d_ctl, p->par_stride);
  LAUNCHED(p);
  return 0;
}

int sk_recv(gb_plan* p, float* d_out, cudaStream_t st) {
  gb_comm* c = p->comm;
  SkIn in;
  in.n = c->world;
  for (int r = 0; r < c->world; ++r) in.slot[r] = gb_slot_of(c, c->rank, p->chan_off, p->slot_bytes, 0, r);
  if (p->prm.quantum_num < 256)
    sk_decode<8><<<grid_for(p->ntiles), GB_THREADS, 0, st>>>(p->tab, in, p->hdr_bytes, p->prm.quantum_num, d_out,
                                                            p->prm.average ? (float)c->world : 1.f, gb_comm_peers(c), p->channel, p->d_ctl,
                                                            p->par_stride, c->status_dev, c->timeout_ns);
  else
    sk_decode<16><<<grid_for(p->ntiles), GB_THREADS, 0, st>>>(p->tab, in, p->hdr_bytes, p->prm.quantum_num, d_out,
                                                             p->prm.average ? (float)c->world : 1.f, gb_comm_peers(c), p->channel, p->d_ctl,
                                                             p->par_stride, c->status_dev, c->timeout_ns);
  LAUNCHED(p);
  return 0;
}
