// K1 -- sparsifier family: TopK, Threshold, RandomK, DGC.
//
// Replaces, per bucket instead of per tensor, the reference op chains
//   memory/residual.py:10-20 (compensate / update), compressor/topk.py:6-18 (abs, topk,
//   gather, zeros + scatter_), compressor/threshold.py:17-18 (where), compressor/randomk.py:6-12
//   (randperm, index), communicator/allgather.py:8-45 (all_gather, W x scatter, sum, div)
// by (SURVEY §2.9-B, rows TopK/Threshold/RandomK) -- TopK itself now runs the two-pass kernels of k1_topk2.cu; the
// four-pass kernels here stay behind GRACE_B200_TOPK_V1 and serve the stand-alone gb_topk_compress; Threshold, DGC
// and the receive kernel k1_merge (all sparse codecs) live here:
//   send:  pass1  compensate (written in place into the residual buffer) + radix histogram
//          select exact per-segment threshold by 3 histogram rounds (11+10+10 bits of |x|)
//          count / scan / emit: index-ORDERED compaction; (value, index32) pairs are stored
//          straight into every peer's slot (P2P st.global over NVLink); selected positions
//          are zeroed in place, which leaves the new residual; last CTA releases the flags.
//   recv:  acquire W flags; each CTA owns one output tile, adds the W ranks' entries for that
//          tile in rank order in shared memory and writes the tile once (no zero-fill pass, no
//          atomics, bit-identical on every rank).
// Ties at the threshold magnitude are broken towards the lowest index (the reference leaves
// the order unspecified: torch.topk(sorted=False)).
#include "kernels.cuh"

struct SparseOut {
  float* vals[GB_MAX_WORLD];
  int32_t* idx[GB_MAX_WORLD];
  int2* hdr[GB_MAX_WORLD];
  int n;
};

struct SparseIn {
  const float* vals[GB_MAX_WORLD];
  const int32_t* idx[GB_MAX_WORLD];
  const int2* hdr[GB_MAX_WORLD];
  int n;
};
// host code passes the parity-0 slot pointers; kernels move them to the parity of the device-side epoch
__device__ __forceinline__ void shift(SparseOut& o, size_t bytes) {
  for (int r = 0; r < o.n; ++r) {
    o.vals[r] = (float*)((char*)o.vals[r] + bytes);
    o.idx[r] = (int32_t*)((char*)o.idx[r] + bytes);
    o.hdr[r] = (int2*)((char*)o.hdr[r] + bytes);
  }
}
__device__ __forceinline__ void shift(SparseIn& o, size_t bytes) {
  for (int r = 0; r < o.n; ++r) {
    o.vals[r] = (const float*)((const char*)o.vals[r] + bytes);
    o.idx[r] = (const int32_t*)((const char*)o.idx[r] + bytes);
    o.hdr[r] = (const int2*)((const char*)o.hdr[r] + bytes);
  }
}

__device__ __forceinline__ float4 load4(const float* p, int e0, int len) {
  if (e0 + 4 <= len) return *reinterpret_cast<const float4*>(p + e0);
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (e0 < len) v.x = p[e0];
  if (e0 + 1 < len) v.y = p[e0 + 1];
  if (e0 + 2 < len) v.z = p[e0 + 2];
  if (e0 + 3 < len) v.w = p[e0 + 3];
  return v;
}
__device__ __forceinline__ void store4(float* p, int e0, int len, float4 v) {
  if (e0 + 4 <= len) { *reinterpret_cast<float4*>(p + e0) = v; return; }
  if (e0 < len) p[e0] = v.x;
  if (e0 + 1 < len) p[e0 + 1] = v.y;
  if (e0 + 2 < len) p[e0 + 2] = v.z;
  if (e0 + 3 < len) p[e0 + 3] = v.w;
}
__device__ __forceinline__ float compensate1(float beta, float r, float gamma, float g) {
  // two roundings like torch's  beta * e + gamma * g  (no fma contraction)
  return __fadd_rn(__fmul_rn(beta, r), __fmul_rn(gamma, g));
}
__device__ __forceinline__ uint32_t absbits(float v) { return __float_as_uint(v) & 0x7FFFFFFFu; }

__device__ __forceinline__ void flush_hist(uint32_t* sh, uint32_t* dst, int nbins) {
  for (int b = threadIdx.x; b < nbins; b += blockDim.x) {
    uint32_t c = sh[b];
    if (c) atomicAdd(dst + b, c);
    sh[b] = 0;
  }
}

// ROUND 0: compensate (+ write) and histogram bits[30:20].  ROUND 1/2: histogram the next 10
// bits of the elements whose leading bits equal the prefix chosen so far.
template <int ROUND, bool WRITE>
__global__ void __launch_bounds__(GB_THREADS, 5) k1_hist(SegTable t, const float* __restrict__ grad, float* work, float beta,
                                                      float gamma, uint32_t* hist, const uint32_t* __restrict__ sel,
                                                      const StepCtl* ctl) {
  constexpr int NB = ROUND == 0 ? 2048 : 1024;
  if (ROUND == 0 && WRITE && ctl->step == 0) { beta = 0.f; gamma = 1.f; }   // empty memory: compensate() is the identity
  __shared__ uint32_t sh[NB];
  const int per = (t.ntiles + gridDim.x - 1) / gridDim.x;
  const int t0 = blockIdx.x * per;
  const int t1 = min(t0 + per, t.ntiles);
  if (t0 >= t1) return;
  for (int b = threadIdx.x; b < NB; b += blockDim.x) sh[b] = 0;
  __syncthreads();
  int cur = t.tile_seg[t0];
  uint32_t prefix = ROUND == 0 ? 0u : sel[cur * 4];
  for (int tile = t0; tile < t1; ++tile) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, tile, seg, base, local0, len);
    if (seg != cur) {
      __syncthreads();
      flush_hist(sh, hist + (size_t)cur * 2048, NB);
      __syncthreads();
      cur = seg;
      if (ROUND != 0) prefix = sel[cur * 4];
    }
    if (len == GB_TILE) {
      // full tile (the common case): all 16-byte loads are issued before the first histogram update; the
      // generic loop below has a bounds branch in front of every load, which keeps one load in flight per thread
      float4 vv4[GB_TILE / (GB_THREADS * 4)];
      if (ROUND == 0 && WRITE) {
        float4 rr4[GB_TILE / (GB_THREADS * 4)];
#pragma unroll
        for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j) {
          const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
          vv4[j] = *reinterpret_cast<const float4*>(grad + base + e0);
          rr4[j] = *reinterpret_cast<const float4*>(work + base + e0);
        }
#pragma unroll
        for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j) {
          const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
          vv4[j].x = compensate1(beta, rr4[j].x, gamma, vv4[j].x);
          vv4[j].y = compensate1(beta, rr4[j].y, gamma, vv4[j].y);
          vv4[j].z = compensate1(beta, rr4[j].z, gamma, vv4[j].z);
          vv4[j].w = compensate1(beta, rr4[j].w, gamma, vv4[j].w);
          *reinterpret_cast<float4*>(work + base + e0) = vv4[j];
        }
      } else {
        const float* src = ROUND == 0 ? grad : work;
#pragma unroll
        for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j)
          vv4[j] = *reinterpret_cast<const float4*>(src + base + (j * GB_THREADS + threadIdx.x) * 4);
      }
#pragma unroll
      for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j) {
        const float vv[4] = {vv4[j].x, vv4[j].y, vv4[j].z, vv4[j].w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint32_t b = absbits(vv[q]);
          if (ROUND == 0) atomicAdd(&sh[b >> 20], 1u);
          else if (ROUND == 1) { if ((b >> 20) == prefix) atomicAdd(&sh[(b >> 10) & 1023u], 1u); }
          else { if ((b >> 10) == prefix) atomicAdd(&sh[b & 1023u], 1u); }
        }
      }
      continue;
    }
#pragma unroll
    for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j) {
      const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
      if (e0 >= len) continue;
      float4 v;
      if (ROUND == 0) {
        v = load4(grad + base, e0, len);
        if (WRITE) {
          float4 r = load4(work + base, e0, len);
          v.x = compensate1(beta, r.x, gamma, v.x);
          v.y = compensate1(beta, r.y, gamma, v.y);
          v.z = compensate1(beta, r.z, gamma, v.z);
          v.w = compensate1(beta, r.w, gamma, v.w);
          store4(work + base, e0, len, v);
        }
      } else {
        v = load4(work + base, e0, len);
      }
      const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (e0 + q >= len) break;
        const uint32_t b = absbits(vv[q]);
        if (ROUND == 0) atomicAdd(&sh[b >> 20], 1u);
        else if (ROUND == 1) { if ((b >> 20) == prefix) atomicAdd(&sh[(b >> 10) & 1023u], 1u); }
        else { if ((b >> 10) == prefix) atomicAdd(&sh[b & 1023u], 1u); }
      }
    }
  }
  __syncthreads();
  flush_hist(sh, hist + (size_t)cur * 2048, NB);
}

// One CTA per segment: walk the histogram from the top bin down to the bin holding the
// k-th largest magnitude; extend the prefix, update k_remaining, re-zero the histogram.
// Round 2 also clears the segment's payload cursor sel[seg][2] and sets sel[seg][3] = 1 when the
// number of elements equal to the threshold differs from the number still to take: only for such a
// segment does the choice among equal magnitudes matter, and only its tiles go through the
// index-ordered (count / scan / emit) compaction; every other segment is done by k1_emit_fast.
__global__ void __launch_bounds__(GB_THREADS) k1_select(SegTable t, uint32_t* hist, uint32_t* sel, int round) {
  __shared__ uint32_t part[GB_THREADS];
  const int seg = blockIdx.x;
  const int nb = round == 0 ? 2048 : 1024;
  const int per = nb / GB_THREADS;
  uint32_t* h = hist + (size_t)seg * 2048;
  const uint32_t k_rem = round == 0 ? (uint32_t)t.seg_k[seg] : sel[seg * 4 + 1];
  const uint32_t prefix = round == 0 ? 0u : sel[seg * 4];
  const int hi = nb - 1 - threadIdx.x * per;       // this thread's bins: hi, hi-1, ..., hi-per+1
  uint32_t c[8];
  uint32_t mine = 0;
  for (int i = 0; i < per; ++i) { c[i] = h[hi - i]; mine += c[i]; h[hi - i] = 0; }
  part[threadIdx.x] = mine;
  __syncthreads();
  for (int o = 1; o < GB_THREADS; o <<= 1) {           // inclusive Hillis-Steele scan
    uint32_t add = threadIdx.x >= o ? part[threadIdx.x - o] : 0u;
    __syncthreads();
    part[threadIdx.x] += add;
    __syncthreads();
  }
  uint32_t running = part[threadIdx.x] - mine;      // elements in strictly higher bins
  for (int i = 0; i < per; ++i) {
    if (running < k_rem && running + c[i] >= k_rem) {
      const uint32_t bin = (uint32_t)(hi - i);
      sel[seg * 4] = round == 0 ? bin : ((prefix << 10) | bin);
      sel[seg * 4 + 1] = k_rem - running;           // still to take inside this bin
      if (round == 2) {
        sel[seg * 4 + 2] = 0;
        sel[seg * 4 + 3] = (c[i] != k_rem - running) ? 1u : 0u;
      }
    }
    running += c[i];
  }
}

__device__ __forceinline__ int block_sum_i(int v, int* sh) {
  v = warp_sum_i(v);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  int tot = 0;
  for (int w = 0; w < GB_THREADS / 32; ++w) tot += sh[w];
  __syncthreads();
  return tot;
}

// per tile: how many magnitudes are above / equal to the segment threshold
__global__ void __launch_bounds__(GB_THREADS) k1_count(SegTable t, const float* __restrict__ work,
                                                       const uint32_t* __restrict__ sel, uint32_t fixed_thr, int use_fixed,
                                                       int ge, int32_t* tilecnt, int only_ambiguous) {
  __shared__ int sh[GB_THREADS / 32];
  for (int tile = blockIdx.x; tile < t.ntiles; tile += gridDim.x) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, tile, seg, base, local0, len);
    if (only_ambiguous && sel[seg * 4 + 3] == 0) continue;      // k1_emit_fast owns this segment
    const uint32_t thr = use_fixed ? fixed_thr : sel[seg * 4];
    int gt = 0, eq = 0;
#pragma unroll
    for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j) {
      const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
      if (e0 >= len) continue;
      float4 v = load4(work + base, e0, len);
      const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (e0 + q >= len) break;
        const uint32_t b = absbits(vv[q]);
        gt += ge ? (b >= thr) : (b > thr);
        eq += ge ? 0 : (b == thr);
      }
    }
    gt = block_sum_i(gt, sh);
    eq = block_sum_i(eq, sh);
    if (threadIdx.x == 0) { tilecnt[tile * 2] = gt; tilecnt[tile * 2 + 1] = eq; }
  }
}

// Exclusive scan of the tile counts.  PER_SEGMENT: one CTA per segment, positions start at the
// segment's fixed payload base (TopK).  Otherwise one CTA scans the whole bucket (variable-size).
// Output per tile: {first payload position, number of threshold ties before this tile}.
template <bool PER_SEGMENT>
__global__ void __launch_bounds__(GB_THREADS) k1_scan(SegTable t, const uint32_t* __restrict__ sel, int32_t* tilecnt,
                                                      int only_ambiguous) {
  __shared__ int sg[GB_THREADS], se[GB_THREADS];
  if (PER_SEGMENT && only_ambiguous && sel[blockIdx.x * 4 + 3] == 0) return;
  const int seg = blockIdx.x;
  const int first = PER_SEGMENT ? t.seg_tile0[seg] : 0;
  const int last = PER_SEGMENT ? t.seg_tile0[seg + 1] : t.ntiles;
  const int take = PER_SEGMENT ? (int)sel[seg * 4 + 1] : 0;
  const int64_t kbase = PER_SEGMENT ? t.seg_kbase[seg] : 0;
  int carry_g = 0, carry_e = 0;
  for (int c0 = first; c0 < last; c0 += GB_THREADS) {
    const int tile = c0 + threadIdx.x;
    const int g = tile < last ? tilecnt[tile * 2] : 0;
    const int e = tile < last ? tilecnt[tile * 2 + 1] : 0;
    sg[threadIdx.x] = g;
    se[threadIdx.x] = e;
    __syncthreads();
    for (int o = 1; o < GB_THREADS; o <<= 1) {
      int ag = threadIdx.x >= o ? sg[threadIdx.x - o] : 0;
      int ae = threadIdx.x >= o ? se[threadIdx.x - o] : 0;
      __syncthreads();
      sg[threadIdx.x] += ag;
      se[threadIdx.x] += ae;
      __syncthreads();
    }
    const int g_before = carry_g + sg[threadIdx.x] - g;
    const int e_before = carry_e + se[threadIdx.x] - e;
    if (tile < last) {
      tilecnt[tile * 2] = (int)kbase + g_before + min(e_before, take);
      tilecnt[tile * 2 + 1] = e_before;
    }
    carry_g += sg[GB_THREADS - 1];
    carry_e += se[GB_THREADS - 1];
    __syncthreads();
  }
}

// block-wide exclusive scan of one int per thread (256 threads); returns total in `total`
__device__ __forceinline__ int block_excl_scan(int v, int* sh, int& total) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int n = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += n;
  }
  if (lane == 31) sh[wid] = inc;
  __syncthreads();
  int woff = 0, tot = 0;
  for (int w = 0; w < GB_THREADS / 32; ++w) {
    int s = sh[w];
    if (w < wid) woff += s;
    tot += s;
  }
  __syncthreads();
  total = tot;
  return woff + inc - v;
}

// Ordered compaction + peer stores + in-place residual update + flag release.
template <bool ZERO>
__global__ void __launch_bounds__(GB_THREADS) k1_emit(SegTable t, float* work, float* work2, const uint32_t* __restrict__ sel,
                                                      uint32_t fixed_thr, int use_fixed, int ge,
                                                      const int32_t* __restrict__ tilecnt,
                                                      SparseOut out, int64_t capacity, uint32_t* status, uint32_t* done,
                                                      PeerPtrs peers, int channel, const StepCtl* ctl, size_t par_stride,
                                                      int release, int only_ambiguous) {
  __shared__ int sh[GB_THREADS / 32];
  const uint32_t epoch = ctl ? ctl->epoch : 0u;
  if (epoch & 1) shift(out, par_stride);
  for (int tile = blockIdx.x; tile < t.ntiles; tile += gridDim.x) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, tile, seg, base, local0, len);
    if (only_ambiguous && sel[seg * 4 + 3] == 0) continue;
    const uint32_t thr = use_fixed ? fixed_thr : sel[seg * 4];
    const int take = use_fixed ? 0 : (int)sel[seg * 4 + 1];
    const int begin = tilecnt[tile * 2];
    int run_g = 0;                                   // selected-above-threshold so far in this tile
    int run_e = tilecnt[tile * 2 + 1];               // ties seen so far in this segment
    const int e_start = run_e;
#pragma unroll 1
    for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j) {
      const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (e0 < len) v = load4(work + base, e0, len);
      const float vv[4] = {v.x, v.y, v.z, v.w};
      int g = 0, e = 0;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (e0 + q < len) {
          const uint32_t b = absbits(vv[q]);
          g += ge ? (b >= thr) : (b > thr);
          e += ge ? 0 : (b == thr);
        }
      }
      int tot;
      const int packed = block_excl_scan(g | (e << 16), sh, tot);
      int g_before = run_g + (packed & 0xFFFF);
      int e_before = run_e + (packed >> 16);
      bool changed = false;
      bool pk[4] = {false, false, false, false};
      float nv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (e0 + q < len) {
          const uint32_t b = absbits(vv[q]);
          const bool is_g = ge ? (b >= thr) : (b > thr), is_e = ge ? false : (b == thr);
          const bool pick = is_g || (is_e && e_before < take);
          if (pick) {
            const int64_t pos = (int64_t)begin + g_before + (min(e_before, take) - min(e_start, take));
            if (pos < capacity) {
              const int32_t li = local0 + e0 + q;
              for (int r = 0; r < out.n; ++r) { out.vals[r][pos] = vv[q]; out.idx[r][pos] = li; }
            } else {
              atomicOr(status, GB_STATUS_OVERFLOW);
            }
            nv[q] = 0.f;
            pk[q] = true;
            changed = true;
          }
          g_before += is_g;
          e_before += is_e;
        }
      }
      if (ZERO && changed) {
        store4(work + base, e0, len, make_float4(nv[0], nv[1], nv[2], nv[3]));
        if (work2) {                                   // DgcMemory.update masks u as well as v
          float4 u = load4(work2 + base, e0, len);
          if (pk[0]) u.x = 0.f;
          if (pk[1]) u.y = 0.f;
          if (pk[2]) u.z = 0.f;
          if (pk[3]) u.w = 0.f;
          store4(work2 + base, e0, len, u);
        }
      }
      run_g += tot & 0xFFFF;
      run_e += tot >> 16;
    }
    if (threadIdx.x == 0) {
      int64_t cnt = run_g + (min(run_e, take) - min(e_start, take));
      if ((int64_t)begin + cnt > capacity) cnt = max((int64_t)0, capacity - begin);
      for (int r = 0; r < out.n; ++r) out.hdr[r][tile] = make_int2(begin, (int)cnt);
    }
  }
  if (release) release_flags_last_block(done, peers, channel, epoch);
}

// Unordered, single-pass compaction (the common case: no ambiguity among equal magnitudes).
// Per tile: ballot-compact the picked (value, index) pairs into shared memory; a range of the payload is claimed with
// one atomicAdd (per-segment cursor for TopK, one global cursor for the variable-size codecs) and the staged pairs go
// to every peer with coalesced stores.  The receiver only needs the per-tile {begin, count} header, so neither the
// order of tiles in the payload nor the order inside a tile matters for the (deterministic) dense result.
// With the global cursor a CTA keeps staging tile after tile and claims once per BATCH (when the 4096-entry staging
// buffer would overflow, or at its last tile): 25 000 tiles claiming one address one by one were the bound of the
// first version -- same-address atomics serialise in L2, 105 us for the 102 MB DGC bucket whatever else overlapped.
#define EF_BATCH 32
template <bool ZERO>
__global__ void __launch_bounds__(GB_THREADS) k1_emit_fast(SegTable t, float* work, float* work2, uint32_t* sel, uint32_t fixed_thr,
                                                           int use_fixed, int per_segment, uint32_t* cursor, SparseOut out,
                                                           int64_t capacity, uint32_t* status, uint32_t* done, PeerPtrs peers,
                                                           int channel, const StepCtl* ctl, size_t par_stride, int release) {
  __shared__ float sval[GB_TILE];
  const uint32_t epoch = ctl->epoch;
  if (epoch & 1) shift(out, par_stride);
  __shared__ int32_t sidx[GB_TILE];
  __shared__ int scount, sbegin;
  __shared__ int b_tile[EF_BATCH], b_off[EF_BATCH], b_cnt[EF_BATCH], b_n;   // tiles staged in the current batch
  __shared__ int s_warp[GB_THREADS / 32];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (threadIdx.x == 0) { scount = 0; b_n = 0; }
  __syncthreads();
  // claim the staged range, write the tiles' headers, copy the pairs out
  auto flush = [&](int seg_for_cursor) {
    if (threadIdx.x == 0) {
      int tc = scount;
      int64_t begin;
      if (per_segment) begin = t.seg_kbase[seg_for_cursor] + atomicAdd(&sel[seg_for_cursor * 4 + 2], (uint32_t)tc);
      else begin = atomicAdd(cursor, (uint32_t)tc);
      int keep = tc;
      if (begin + tc > capacity) {
        atomicOr(status, GB_STATUS_OVERFLOW);
        keep = (int)max((int64_t)0, capacity - begin);
      }
      for (int i = 0; i < b_n; ++i) {
        const int c = max(0, min(b_cnt[i], keep - b_off[i]));
        for (int r = 0; r < out.n; ++r) out.hdr[r][b_tile[i]] = make_int2((int)begin + b_off[i], c);
      }
      scount = keep;
      sbegin = (int)begin;
    }
    __syncthreads();
    const int tcnt = scount, begin = sbegin;
    for (int r = 0; r < out.n; ++r) {
      float* dv = out.vals[r] + begin;
      int32_t* di = out.idx[r] + begin;
      for (int i = threadIdx.x; i < tcnt; i += GB_THREADS) { dv[i] = sval[i]; di[i] = sidx[i]; }
    }
    __syncthreads();
    if (threadIdx.x == 0) { scount = 0; b_n = 0; }
    __syncthreads();
  };
  int last_seg = -1;
  for (int tile = blockIdx.x; tile < t.ntiles; tile += gridDim.x) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, tile, seg, base, local0, len);
    if (per_segment && sel[seg * 4 + 3] != 0) continue;          // ambiguous ties: ordered kernels do this segment
    const uint32_t thr = use_fixed ? fixed_thr : sel[seg * 4];
    // all 16 values of this thread stay in registers; one warp prefix sum per tile places the picked pairs
    float vv[4][4];
    uint32_t picks = 0;
    int cnt = 0;
    if (len == GB_TILE) {
#pragma unroll
      for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j) {       // four independent 16-byte loads per thread
        const float4 v = *reinterpret_cast<const float4*>(work + base + (j * GB_THREADS + threadIdx.x) * 4);
        vv[j][0] = v.x; vv[j][1] = v.y; vv[j][2] = v.z; vv[j][3] = v.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j) {
        const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (e0 < len) v = load4(work + base, e0, len);
        vv[j][0] = v.x; vv[j][1] = v.y; vv[j][2] = v.z; vv[j][3] = v.w;
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint32_t b = absbits(vv[j][q]);
        // TopK: >= exact threshold; Threshold codec: > fixed; DGC: >= per-segment threshold
        const bool pick = (e0 + q < len) && (use_fixed ? (b > thr) : (b >= thr));
        picks |= (pick ? 1u : 0u) << (j * 4 + q);
        cnt += pick;
      }
    }
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int n = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += n;
    }
    if (lane == 31) s_warp[wid] = incl;
    __syncthreads();
    int wbase = 0, tile_total = 0;
#pragma unroll
    for (int w = 0; w < GB_THREADS / 32; ++w) {
      const int c = s_warp[w];
      if (w < wid) wbase += c;
      tile_total += c;
    }
    // the batch is flushed first when this tile would not fit, when the batch table is full, or (per-segment cursors)
    // when the tile belongs to another tensor than the batch
    const bool must_flush = b_n > 0 && (scount + tile_total > GB_TILE || b_n == EF_BATCH || (per_segment && seg != last_seg));
    __syncthreads();
    if (must_flush) flush(last_seg);
    const int off0 = scount;
    __syncthreads();
    if (picks) {
      int pos = off0 + wbase + incl - cnt;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
        if ((picks >> (j * 4)) & 0xFu) {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            if ((picks >> (j * 4 + q)) & 1u) {
              sval[pos] = vv[j][q];
              sidx[pos] = local0 + e0 + q;
              ++pos;
              vv[j][q] = 0.f;
            }
          }
          if (ZERO) {
            store4(work + base, e0, len, make_float4(vv[j][0], vv[j][1], vv[j][2], vv[j][3]));
            if (work2) {
              float4 u = load4(work2 + base, e0, len);
              if ((picks >> (j * 4 + 0)) & 1u) u.x = 0.f;
              if ((picks >> (j * 4 + 1)) & 1u) u.y = 0.f;
              if ((picks >> (j * 4 + 2)) & 1u) u.z = 0.f;
              if ((picks >> (j * 4 + 3)) & 1u) u.w = 0.f;
              store4(work2 + base, e0, len, u);
            }
          }
        }
      }
    }
    if (threadIdx.x == 0) {
      b_tile[b_n] = tile; b_off[b_n] = off0; b_cnt[b_n] = tile_total;
      b_n = b_n + 1;
      scount = off0 + tile_total;
    }
    last_seg = seg;
    __syncthreads();
  }
  if (b_n > 0) flush(last_seg);
  if (release) release_flags_last_block(done, peers, channel, epoch);
}

// Receive: ordered merge of W sparse lists into the dense aggregate.

// NW: compile-time bound of the number of ranks (1, 2, 4, 8): keeps the per-rank prefetch registers to what is needed
template <bool EXTRAS, int NW>
__global__ void __launch_bounds__(GB_THREADS, 5) k1_merge(SegTable t, SparseIn in, float* out, float divisor, PeerPtrs peers,
                                                       int channel, const StepCtl* ctl, size_t par_stride, uint32_t* status,
                                                       uint64_t timeout_ns, int wait) {
  __shared__ __align__(16) float acc[GB_TILE];
  __shared__ int2 s_hdr[2][2 * NW];         // headers of this tile and of the CTA's next tile: [buf][range * NW + rank]
  const uint32_t epoch = ctl->epoch;
  if (epoch & 1) shift(in, par_stride);
  // the accumulator tile starts at zero and is put back to zero by the write-out below
  for (int i = threadIdx.x; i < GB_TILE; i += GB_THREADS) acc[i] = 0.f;
  if (wait) wait_all_ranks(peers, channel, epoch, status, timeout_ns);
  else __syncthreads();
  constexpr int NR = EXTRAS ? 2 : 1;        // TopK: a second range per tile (picks from the threshold bin)
  auto fetch_hdr = [&](int tile, int buf) {
    if (threadIdx.x < NR * NW) {
      const int r = threadIdx.x % NW, which = threadIdx.x / NW;
      int2 v = make_int2(0, 0);
      if (r < in.n && tile < t.ntiles) v = in.hdr[r][which ? t.ntiles + 1 + tile : tile];
      s_hdr[buf][which * NW + r] = v;
    }
  };
  fetch_hdr(blockIdx.x, 0);
  __syncthreads();
  int buf = 0;
  for (int tile = blockIdx.x; tile < t.ntiles; tile += gridDim.x, buf ^= 1) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, tile, seg, base, local0, len);
    // One memory latency per tile: the headers were fetched during the previous tile, so every rank's entries (both
    // ranges, one per thread) are requested at once before anything is added; 16 threads fetch the next tile's
    // headers meanwhile.  (First version: headers, then entries, then the second range rank by rank -- 2 + W
    // dependent latencies per tile, 177 us for ResNet-50 at W = 8.)  The adds keep rank order.
    float pv[NW];
    int32_t pi[NW];
#pragma unroll
    for (int r = 0; r < NW; ++r) {
      pv[r] = 0.f;
      pi[r] = -1;
      if (r < in.n) {
        const int2 h = s_hdr[buf][r];
        const int2 x = EXTRAS ? s_hdr[buf][NW + r] : make_int2(0, 0);
        const int i = threadIdx.x;
        if (i < h.y) {
          pi[r] = in.idx[r][h.x + i];
          pv[r] = in.vals[r][h.x + i];
        } else if (EXTRAS && i < h.y + x.y) {
          pi[r] = in.idx[r][x.x + i - h.y];
          pv[r] = in.vals[r][x.x + i - h.y];
        }
      }
    }
    fetch_hdr(tile + gridDim.x, buf ^ 1);
#pragma unroll
    for (int r = 0; r < NW; ++r) {
      if (r < in.n) {
        const int l0 = pi[r] - local0;
        if (pi[r] >= 0 && (unsigned)l0 < (unsigned)GB_TILE) acc[l0] += pv[r];
        const int2 h = s_hdr[buf][r];
        const int2 x = EXTRAS ? s_hdr[buf][NW + r] : make_int2(0, 0);
        if (h.y + x.y > GB_THREADS) {                  // rare: more than 256 entries of one rank in one tile
          for (int i = threadIdx.x + GB_THREADS; i < h.y + x.y; i += GB_THREADS) {
            const int64_t e = i < h.y ? (int64_t)h.x + i : (int64_t)x.x + i - h.y;
            const int li = in.idx[r][e] - local0;
            if ((unsigned)li < (unsigned)GB_TILE) acc[li] += in.vals[r][e];
          }
        }
        __syncthreads();                               // indices are unique within a rank; ranks add one after another
      }
    }
#pragma unroll
    for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j) {
      const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
      float4 v = *reinterpret_cast<const float4*>(acc + e0);
      *reinterpret_cast<float4*>(acc + e0) = make_float4(0.f, 0.f, 0.f, 0.f);
      if (e0 >= len) continue;
      if (divisor != 1.f) {          // torch's CUDA tensor / python-scalar is a multiply by the reciprocal
        const float inv = __fdiv_rn(1.0f, divisor);
        v.x = __fmul_rn(v.x, inv); v.y = __fmul_rn(v.y, inv);
        v.z = __fmul_rn(v.z, inv); v.w = __fmul_rn(v.w, inv);
      }
      store4(out + base, e0, len, v);
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------- RandomK
__device__ __forceinline__ int find_segment(const int64_t* kbase, int nseg, int64_t p) {
  int lo = 0, hi = nseg;                       // largest s with kbase[s] <= p
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (kbase[mid] <= p) lo = mid; else hi = mid;
  }
  return lo;
}

__device__ __forceinline__ int64_t randomk_index(const SegTable& t, int seg, int64_t j, uint32_t seed, uint32_t step) {
  const Philox4 k = philox4x32(0u, step, GB_STREAM_PERMUTE, 0u, seed, t.seg_hash[seg]);
  const uint32_t keys[4] = {k.x, k.y, k.z, k.w};
  const uint64_t n = (uint64_t)t.seg_numel[seg];
  return (int64_t)feistel_permute((uint64_t)j, n, feistel_half_bits(n), keys);
}

template <bool WRITE>
__global__ void __launch_bounds__(GB_THREADS) k1_axpby(SegTable t, const float* __restrict__ grad, float* work, float beta,
                                                       float gamma, const StepCtl* ctl) {
  if (ctl->step == 0) { beta = 0.f; gamma = 1.f; }
  const int64_t n4 = t.padded / 4;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float4 g = reinterpret_cast<const float4*>(grad)[i];
    float4 r = reinterpret_cast<float4*>(work)[i];
    r.x = compensate1(beta, r.x, gamma, g.x); r.y = compensate1(beta, r.y, gamma, g.y);
    r.z = compensate1(beta, r.z, gamma, g.z); r.w = compensate1(beta, r.w, gamma, g.w);
    reinterpret_cast<float4*>(work)[i] = r;
  }
}

template <bool ZERO>
__global__ void __launch_bounds__(GB_THREADS) k1_randomk_gather(SegTable t, float* work, int64_t ktotal, uint32_t seed,
                                                               SparseOut out, uint32_t* done, PeerPtrs peers, int channel,
                                                               const StepCtl* ctl, size_t par_stride) {
  const uint32_t epoch = ctl->epoch, step = ctl->step;
  if (epoch & 1) shift(out, par_stride);
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < ktotal; p += (int64_t)gridDim.x * blockDim.x) {
    const int seg = find_segment(t.seg_kbase, t.nseg, p);
    const int64_t idx = randomk_index(t, seg, p - t.seg_kbase[seg], seed, step);
    float* src = work + t.seg_off[seg] + idx;
    const float v = *src;
    for (int r = 0; r < out.n; ++r) out.vals[r][p] = v;
    if (ZERO) *src = 0.f;
  }
  release_flags_last_block(done, peers, channel, epoch);
}

__global__ void __launch_bounds__(GB_THREADS) k1_randomk_scatter(SegTable t, int64_t ktotal, uint32_t seed, SparseIn in,
                                                                float* out, float divisor, PeerPtrs peers, int channel,
                                                                const StepCtl* ctl, size_t par_stride, uint32_t* status,
                                                                uint64_t timeout_ns) {
  const uint32_t epoch = ctl->epoch, step = ctl->step;
  if (epoch & 1) shift(in, par_stride);
  wait_all_ranks(peers, channel, epoch, status, timeout_ns);
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < ktotal; p += (int64_t)gridDim.x * blockDim.x) {
    const int seg = find_segment(t.seg_kbase, t.nseg, p);
    const int64_t idx = randomk_index(t, seg, p - t.seg_kbase[seg], seed, step);
    float s = 0.f;
    for (int r = 0; r < in.n; ++r) s += in.vals[r][p];
    out[t.seg_off[seg] + idx] = divisor != 1.f ? __fmul_rn(s, __fdiv_rn(1.0f, divisor)) : s;
  }
}

// ---------------------------------------------------------------- host side
static int grid_for(int ntiles) { return ntiles < 148 * 8 ? ntiles : 148 * 8; }

static void slot_views(gb_plan* p, int parity, bool remote_side, SparseOut* o, SparseIn* in) {
  // remote_side: pointers into every peer's arena at *my* slot (for sending);
  // otherwise pointers into my arena at every source rank's slot (for receiving).
  gb_comm* c = p->comm;
  const size_t per = ((size_t)p->ktotal * 4 + 255) & ~size_t(255);
  for (int r = 0; r < c->world; ++r) {
    char* s = remote_side ? gb_slot_of(c, r, p->chan_off, p->slot_bytes, parity, c->rank)
                          : gb_slot_of(c, c->rank, p->chan_off, p->slot_bytes, parity, r);
    if (o) { o->hdr[r] = (int2*)s; o->vals[r] = (float*)(s + p->hdr_bytes); o->idx[r] = (int32_t*)(s + p->hdr_bytes + per); }
    if (in) { in->hdr[r] = (const int2*)s; in->vals[r] = (const float*)(s + p->hdr_bytes); in->idx[r] = (const int32_t*)(s + p->hdr_bytes + per); }
  }
  if (o) o->n = c->world;
  if (in) in->n = c->world;
}

#define LAUNCHED(p) do { (p)->launches++; GB_CUDA(cudaGetLastError()); } while (0)

static int k1_dgc_send(gb_plan* p, const float* d_grad, cudaStream_t st);

int k1_send(gb_plan* p, const float* d_grad, cudaStream_t st) {
  if (p->prm.codec == GB_CODEC_DGC) return k1_dgc_send(p, d_grad, st);
  if (p->topk2) return tk_send(p, d_grad, st);
  gb_comm* c = p->comm;
  const bool mem = p->prm.memory == GB_MEM_RESIDUAL;
  // first call: the memory is empty, compensate() returns the gradient unchanged
  const float beta = p->prm.beta, gamma = p->prm.gamma;     // kernels apply the first-call rule (device-side step)
  float* work = mem ? p->d_state0 : const_cast<float*>(d_grad);
  ++p->epoch;                                               // host mirror of the device-side epoch
  SparseOut out;
  slot_views(p, 0, true, &out, nullptr);
  StepCtl* ctl = p->d_ctl;
  const size_t ps = p->par_stride;
  const PeerPtrs peers = gb_comm_peers(c);
  const int grid = grid_for(p->ntiles);

  if (p->prm.codec == GB_CODEC_RANDOMK) {
    if (mem) { k1_axpby<true><<<148 * 4, GB_THREADS, 0, st>>>(p->tab, d_grad, work, beta, gamma, ctl); LAUNCHED(p); }
    const int g2 = (int)std::min<int64_t>((p->ktotal + GB_THREADS - 1) / GB_THREADS, 148 * 8);
    if (mem) k1_randomk_gather<true><<<g2, GB_THREADS, 0, st>>>(p->tab, work, p->ktotal, p->prm.seed, out, p->d_done, peers, p->channel, ctl, ps);
    else k1_randomk_gather<false><<<g2, GB_THREADS, 0, st>>>(p->tab, work, p->ktotal, p->prm.seed, out, p->d_done, peers, p->channel, ctl, ps);
    LAUNCHED(p);
    return 0;
  }

  const bool fixed = p->prm.codec == GB_CODEC_THRESHOLD;
  uint32_t thr_bits = 0u;
  if (fixed) { const float th = p->prm.threshold < 0.f ? 0.f : p->prm.threshold; memcpy(&thr_bits, &th, 4); }
  const int hgrid = std::min(p->ntiles, 148 * 8);
  uint32_t* cursor = p->d_done + 2;
  if (fixed) {
    // variable-size payload: one global cursor, a single compaction pass
    if (mem) { k1_axpby<true><<<148 * 4, GB_THREADS, 0, st>>>(p->tab, d_grad, work, beta, gamma, ctl); LAUNCHED(p); }
    GB_CUDA(cudaMemsetAsync(cursor, 0, 4, st));
    if (mem) k1_emit_fast<true><<<grid, GB_THREADS, 0, st>>>(p->tab, work, nullptr, p->d_sel, thr_bits, 1, 0, cursor, out, p->ktotal, c->status_dev, p->d_done, peers, p->channel, ctl, ps, 1);
    else k1_emit_fast<false><<<grid, GB_THREADS, 0, st>>>(p->tab, work, nullptr, p->d_sel, thr_bits, 1, 0, cursor, out, p->ktotal, c->status_dev, p->d_done, peers, p->channel, ctl, ps, 1);
    LAUNCHED(p);
    return 0;
  }
  if (mem) k1_hist<0, true><<<hgrid, GB_THREADS, 0, st>>>(p->tab, d_grad, work, beta, gamma, p->d_hist, p->d_sel, ctl);
  else k1_hist<0, false><<<hgrid, GB_THREADS, 0, st>>>(p->tab, d_grad, work, beta, gamma, p->d_hist, p->d_sel, ctl);
  LAUNCHED(p);
  k1_select<<<p->nseg, GB_THREADS, 0, st>>>(p->tab, p->d_hist, p->d_sel, 0); LAUNCHED(p);
  k1_hist<1, false><<<hgrid, GB_THREADS, 0, st>>>(p->tab, d_grad, work, beta, gamma, p->d_hist, p->d_sel, ctl); LAUNCHED(p);
  k1_select<<<p->nseg, GB_THREADS, 0, st>>>(p->tab, p->d_hist, p->d_sel, 1); LAUNCHED(p);
  k1_hist<2, false><<<hgrid, GB_THREADS, 0, st>>>(p->tab, d_grad, work, beta, gamma, p->d_hist, p->d_sel, ctl); LAUNCHED(p);
  k1_select<<<p->nseg, GB_THREADS, 0, st>>>(p->tab, p->d_hist, p->d_sel, 2); LAUNCHED(p);
  // segments without tie ambiguity (nearly all): one unordered compaction pass; the others: count /
  // scan / index-ordered emit restricted to their tiles.  The last kernel releases the flags.
  if (mem) k1_emit_fast<true><<<grid, GB_THREADS, 0, st>>>(p->tab, work, nullptr, p->d_sel, 0u, 0, 1, cursor, out, p->ktotal, c->status_dev, p->d_done, peers, p->channel, ctl, ps, 0);
  else k1_emit_fast<false><<<grid, GB_THREADS, 0, st>>>(p->tab, work, nullptr, p->d_sel, 0u, 0, 1, cursor, out, p->ktotal, c->status_dev, p->d_done, peers, p->channel, ctl, ps, 0);
  LAUNCHED(p);
  k1_count<<<grid, GB_THREADS, 0, st>>>(p->tab, work, p->d_sel, 0u, 0, 0, p->d_tilecnt, 1); LAUNCHED(p);
  k1_scan<true><<<p->nseg, GB_THREADS, 0, st>>>(p->tab, p->d_sel, p->d_tilecnt, 1); LAUNCHED(p);
  if (mem) k1_emit<true><<<grid, GB_THREADS, 0, st>>>(p->tab, work, nullptr, p->d_sel, 0u, 0, 0, p->d_tilecnt, out, p->ktotal, c->status_dev, p->d_done, peers, p->channel, ctl, ps, 1, 1);
  else k1_emit<false><<<grid, GB_THREADS, 0, st>>>(p->tab, work, nullptr, p->d_sel, 0u, 0, 0, p->d_tilecnt, out, p->ktotal, c->status_dev, p->d_done, peers, p->channel, ctl, ps, 1, 1);
  LAUNCHED(p);
  return 0;
}

int k1_recv(gb_plan* p, float* d_out, cudaStream_t st) {
  gb_comm* c = p->comm;
  SparseIn in;
  slot_views(p, 0, false, nullptr, &in);
  const PeerPtrs peers = gb_comm_peers(c);
  const float divisor = p->prm.average ? (float)c->world : 1.f;
  if (p->prm.codec == GB_CODEC_RANDOMK) {
    GB_CUDA(cudaMemsetAsync(d_out, 0, (size_t)p->padded * 4, st));
    const int g2 = (int)std::min<int64_t>((p->ktotal + GB_THREADS - 1) / GB_THREADS, 148 * 8);
    k1_randomk_scatter<<<g2, GB_THREADS, 0, st>>>(p->tab, p->ktotal, p->prm.seed, in, d_out, divisor, peers, p->channel, p->d_ctl, p->par_stride, c->status_dev, c->timeout_ns);
    LAUNCHED(p);
    return 0;
  }
  // 48 registers, 16.6 KB of shared memory: five CTAs are resident per SM; the tile loop is persistent
  const int mgrid = p->ntiles < 148 * 5 ? p->ntiles : 148 * 5;
#define MERGE(EX, NW) k1_merge<EX, NW><<<mgrid, GB_THREADS, 0, st>>>(p->tab, in, d_out, divisor, peers, p->channel, p->d_ctl, p->par_stride, c->status_dev, c->timeout_ns, 1)
  const bool ex = p->prm.codec == GB_CODEC_TOPK;
  if (c->world == 1) { if (ex) MERGE(true, 1); else MERGE(false, 1); }
  else if (c->world == 2) { if (ex) MERGE(true, 2); else MERGE(false, 2); }
  else if (c->world <= 4) { if (ex) MERGE(true, 4); else MERGE(false, 4); }
  else { if (ex) MERGE(true, 8); else MERGE(false, 8); }
#undef MERGE
  LAUNCHED(p);
  return 0;
}


// ---------------------------------------------------------------- DGC (+ DgcMemory)
// Reference: compressor/dgc.py:12-50 and memory/dgc.py:15-39.  The reference needs up to 12
// host round trips per tensor (.item()-style compares inside the refinement loop); here the
// sampled threshold and the <=10 multiplicative refinements stay on the device.
// sel[seg] = { threshold bits, 0, done flag, selected count }.

__global__ void __launch_bounds__(GB_THREADS) k1_dgc_accumulate(SegTable t, const float* __restrict__ grad, float* v, float* u,
                                                               float momentum) {
  const int64_t n4 = t.padded / 4;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 g = reinterpret_cast<const float4*>(grad)[i];
    float4 uu = reinterpret_cast<float4*>(u)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    uu.x = __fadd_rn(__fmul_rn(momentum, uu.x), g.x); uu.y = __fadd_rn(__fmul_rn(momentum, uu.y), g.y);
    uu.z = __fadd_rn(__fmul_rn(momentum, uu.z), g.z); uu.w = __fadd_rn(__fmul_rn(momentum, uu.w), g.w);
    vv.x = __fadd_rn(vv.x, uu.x); vv.y = __fadd_rn(vv.y, uu.y); vv.z = __fadd_rn(vv.z, uu.z); vv.w = __fadd_rn(vv.w, uu.w);
    reinterpret_cast<float4*>(u)[i] = uu;
    reinterpret_cast<float4*>(v)[i] = vv;
  }
}

// ---- DgcMemory gradient clipping (memory/dgc.py:17-20): clamp g to +-sqrt(sum_r sum(g_r^2) / W) per tensor.
// k1_dgc_sqsum: per-tile partial sums in double; the last CTA adds each tensor's tiles in tile order (deterministic)
// and stores the W-th part of the scalar all-reduce -- this rank's nseg floats -- into every peer, then releases the
// flags of the plan's second channel.  k1_dgc_accumulate_clip waits for all ranks' scalars, sums them in rank order
// (identical on every rank), clamps, and applies the momentum correction + accumulation in the same pass.
struct ClipOut { float* slot[GB_MAX_WORLD]; int n; };
struct ClipIn { const float* slot[GB_MAX_WORLD]; int n; };

__global__ void __launch_bounds__(GB_THREADS) k1_dgc_sqsum(SegTable t, const float* __restrict__ grad, double* tilesum, ClipOut out,
                                                          uint32_t* done, PeerPtrs peers, int channel2, const StepCtl* ctl,
                                                          size_t par_stride2) {
  __shared__ double sh[GB_THREADS / 32];
  __shared__ int s_last;
  for (int tile = blockIdx.x; tile < t.ntiles; tile += gridDim.x) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, tile, seg, base, local0, len);
    double acc = 0.0;
#pragma unroll
    for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j) {
      const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
      if (e0 >= len) continue;
      const float4 g = load4(grad + base, e0, len);
      acc += (double)g.x * g.x + (double)g.y * g.y + (double)g.z * g.z + (double)g.w * g.w;
    }
    acc = warp_sum_d(acc);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
      double s = 0.0;
      for (int w = 0; w < GB_THREADS / 32; ++w) s += sh[w];
      tilesum[tile] = s;
    }
    __syncthreads();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    s_last = atomicAdd(done, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const uint32_t epoch = ctl->epoch;
  const size_t sh2 = (epoch & 1) ? par_stride2 : 0;
  for (int seg = threadIdx.x; seg < t.nseg; seg += GB_THREADS) {
    double s = 0.0;
    for (int tile = t.seg_tile0[seg]; tile < t.seg_tile0[seg + 1]; ++tile) s += __ldcg(tilesum + tile);
    const float f = (float)s;
    for (int r = 0; r < out.n; ++r) ((float*)((char*)out.slot[r] + sh2))[seg] = f;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (peers.world == 1) __threadfence(); else __threadfence_system();
    for (int r = 0; r < peers.world; ++r) st_release_sys(gb_flag(peers.base[r], channel2, epoch & 1, peers.rank), epoch);
    *done = 0;
  }
}

__global__ void __launch_bounds__(GB_THREADS) k1_dgc_accumulate_clip(SegTable t, const float* __restrict__ grad, float* v, float* u,
                                                                    float momentum, ClipIn in, float world_f, PeerPtrs peers,
                                                                    int channel2, const StepCtl* ctl, size_t par_stride2,
                                                                    uint32_t* status, uint64_t timeout_ns) {
  const uint32_t epoch = ctl->epoch;
  const size_t sh2 = (epoch & 1) ? par_stride2 : 0;
  wait_all_ranks(peers, channel2, epoch, status, timeout_ns);
  for (int tile = blockIdx.x; tile < t.ntiles; tile += gridDim.x) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, tile, seg, base, local0, len);
    float sq = 0.f;
    for (int r = 0; r < in.n; ++r) sq = __fadd_rn(sq, __ldcg((const float*)((const char*)in.slot[r] + sh2) + seg));   // rank order
    const float clip = __fsqrt_rn(__fdiv_rn(sq, world_f));
#pragma unroll
    for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j) {
      const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
      if (e0 >= len) continue;
      float4 g = load4(grad + base, e0, len);
      g.x = fmaxf(fminf(g.x, clip), -clip); g.y = fmaxf(fminf(g.y, clip), -clip);
      g.z = fmaxf(fminf(g.z, clip), -clip); g.w = fmaxf(fminf(g.w, clip), -clip);
      float4 uu = load4(u + base, e0, len);
      float4 vv = load4(v + base, e0, len);
      uu.x = __fadd_rn(__fmul_rn(momentum, uu.x), g.x); uu.y = __fadd_rn(__fmul_rn(momentum, uu.y), g.y);
      uu.z = __fadd_rn(__fmul_rn(momentum, uu.z), g.z); uu.w = __fadd_rn(__fmul_rn(momentum, uu.w), g.w);
      vv.x = __fadd_rn(vv.x, uu.x); vv.y = __fadd_rn(vv.y, uu.y); vv.z = __fadd_rn(vv.z, uu.z); vv.w = __fadd_rn(vv.w, uu.w);
      store4(u + base, e0, len, uu);
      store4(v + base, e0, len, vv);
    }
  }
}

// Sampling with replacement (Philox stream 1), one thread per sample on a (chunk, tensor) grid: magnitudes of
// max(1, int(0.01 n)) positions go to a scratch list (seg_k = samples per tensor, seg_kbase = their offsets).
__global__ void __launch_bounds__(GB_THREADS) k1_dgc_gather(SegTable t, const float* __restrict__ v, uint32_t seed, const StepCtl* ctl,
                                                           uint32_t* samp) {
  const int seg = blockIdx.y;
  const int ns = t.seg_k[seg];
  const int i = blockIdx.x * GB_THREADS + threadIdx.x;
  if (i >= ns) return;
  const int n = t.seg_numel[seg];
  const Philox4 r = philox4x32((uint32_t)(i >> 2), ctl->step, GB_STREAM_SAMPLE, 0u, seed, t.seg_hash[seg]);
  const uint32_t bits = (i & 3) == 0 ? r.x : (i & 3) == 1 ? r.y : (i & 3) == 2 ? r.z : r.w;
  int pos = (int)__fmul_rn(u32_to_unit(bits), (float)n);
  pos = min(pos, n - 1);
  samp[t.seg_kbase[seg] + i] = absbits(v[t.seg_off[seg] + pos]);
}

// One CTA per segment: k-th largest of the sampled magnitudes by three histogram rounds over the scratch list.
__global__ void __launch_bounds__(GB_THREADS) k1_dgc_sample(SegTable t, const uint32_t* __restrict__ samp, double ratio, uint32_t* sel) {
  __shared__ uint32_t hist[2048];
  __shared__ uint32_t part[GB_THREADS];
  __shared__ uint32_t s_prefix, s_krem;
  const int seg = blockIdx.x;
  const int n = t.seg_numel[seg];
  const int ns = t.seg_k[seg];
  int k = max(1, (int)((double)n * ratio * 0.01));
  k = min(k, ns);
  const uint32_t* x = samp + t.seg_kbase[seg];
  if (threadIdx.x == 0) { s_prefix = 0; s_krem = (uint32_t)k; }
  for (int round = 0; round < 3; ++round) {
    const int nb = round == 0 ? 2048 : 1024;
    for (int b = threadIdx.x; b < nb; b += blockDim.x) hist[b] = 0;
    __syncthreads();
    const uint32_t prefix = s_prefix;
    for (int i = threadIdx.x; i < ns; i += blockDim.x) {
      const uint32_t b = x[i];
      if (round == 0) atomicAdd(&hist[b >> 20], 1u);
      else if (round == 1) { if ((b >> 20) == prefix) atomicAdd(&hist[(b >> 10) & 1023u], 1u); }
      else { if ((b >> 10) == prefix) atomicAdd(&hist[b & 1023u], 1u); }
    }
    __syncthreads();
    {
      // parallel walk from the top bin down to the one holding the k-th largest sample (a single thread
      // walking 2048 shared-memory bins cost ~20 us per round)
      const uint32_t krem = s_krem;
      const int per = nb / GB_THREADS;
      const int hi = nb - 1 - threadIdx.x * per;
      uint32_t c[8];
      uint32_t mine = 0;
      for (int i = 0; i < per; ++i) { c[i] = hist[hi - i]; mine += c[i]; }
      part[threadIdx.x] = mine;
      __syncthreads();
      for (int o = 1; o < GB_THREADS; o <<= 1) {
        const uint32_t add = threadIdx.x >= o ? part[threadIdx.x - o] : 0u;
        __syncthreads();
        part[threadIdx.x] += add;
        __syncthreads();
      }
      uint32_t running = part[threadIdx.x] - mine;
      for (int i = 0; i < per; ++i) {
        if (running < krem && running + c[i] >= krem) {
          s_prefix = round == 0 ? (uint32_t)(hi - i) : ((prefix << 10) | (uint32_t)(hi - i));
          s_krem = krem - running;
        }
        running += c[i];
      }
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) { sel[seg * 4] = s_prefix; sel[seg * 4 + 1] = 0; sel[seg * 4 + 2] = 0; sel[seg * 4 + 3] = 0; }
}

// Threshold refinement (compressor/dgc.py:28-36: up to ten times, x1.3 if more than 1.3 k values pass, x0.7 if fewer
// than 0.7 k, else stop).  The factors are coarse, so mid-sized tensors bounce around the band and really need their
// rounds -- and a round is a launch-plus-latency chain of ~12 us however little it counts.  Hence:
//   * one launch counts, per element, against the current threshold AND the 14 thresholds the next three decisions
//     can lead to (a binary tree: child 2i+1 = 1.3f * T[i], child 2i+2 = 0.7f * T[i], the reference's own float
//     products), so the CTA that finishes last walks up to four decisions per tensor: 3 launches instead of 10;
//   * a tensor of one tile is refined to the end by the CTA that owns it, values in registers;
//   * the last CTA lists the tensors that are still open (OpenDir): later launches walk only their tiles and return at
//     once when none is open.
// (All ten rounds in one persistent kernel with a grid barrier were tried too: slower than the launches -- 0.49 vs
// 0.41 ms per ResNet-50 exchange; the barrier's serial latency chain costs more than a launch boundary.)
// sel[seg] = {threshold bits, decisions taken, final flag, -};  cnt[seg][16]: passes per tree node.
#define DGC_PREF_CACHE 1024
#define DGC_TREE 15
#define DGC_LAUNCHES 3
__device__ __forceinline__ void dgc_tree(float thr, float (&T)[DGC_TREE]) {
  T[0] = thr;
#pragma unroll
  for (int i = 0; 2 * i + 2 < DGC_TREE; ++i) { T[2 * i + 1] = __fmul_rn(1.3f, T[i]); T[2 * i + 2] = __fmul_rn(0.7f, T[i]); }
}
__global__ void __launch_bounds__(GB_THREADS) k1_dgc_count(SegTable t, const float* __restrict__ v, uint32_t* sel, double ratio,
                                                          uint32_t* done, uint32_t* dir_scratch, int round) {
  __shared__ int sh[GB_THREADS / 32];
  __shared__ int s_cnt[DGC_TREE];
  __shared__ int s_last;
  __shared__ int s_pref[DGC_PREF_CACHE];
  const OpenDir dir = open_dir(dir_scratch, t.nseg);
  uint32_t* cnt = dir_scratch + ((4 + 2 * t.nseg + 15) & ~15);          // [nseg][16]
  auto count_tile = [&](int tile) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, tile, seg, base, local0, len);
    if (sel[seg * 4 + 2]) return;                      // this segment's threshold is final
    const uint32_t thr = sel[seg * 4];
    constexpr int NJ = GB_TILE / (GB_THREADS * 4);
    uint32_t ab[NJ * 4];
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
      float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
      if (e0 < len) x = load4(v + base, e0, len);
      const float xx[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) ab[j * 4 + q] = e0 + q < len ? absbits(xx[q]) : 0u;
    }
    if (round == 0 && t.seg_tile0[seg + 1] - t.seg_tile0[seg] == 1) {
      // single tile: this CTA is the only one that touches the tensor -- all ten decisions here
      const double n = (double)t.seg_numel[seg];
      float th = __uint_as_float(thr);
      for (int it = 0; it < 10; ++it) {
        const uint32_t tb = __float_as_uint(th);
        int c = 0;
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
          const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
#pragma unroll
          for (int q = 0; q < 4; ++q) c += (e0 + q < len) && ab[j * 4 + q] >= tb;
        }
        const double selected = (double)block_sum_i(c, sh);
        if (selected > 1.3 * n * ratio) th = __fmul_rn(1.3f, th);
        else if (selected < 0.7 * n * ratio) th = __fmul_rn(0.7f, th);
        else break;
      }
      if (threadIdx.x == 0) { sel[seg * 4] = __float_as_uint(th); sel[seg * 4 + 2] = 1u; }
      return;
    }
    float T[DGC_TREE];
    dgc_tree(__uint_as_float(thr), T);
    if (threadIdx.x < DGC_TREE) s_cnt[threadIdx.x] = 0;
    __syncthreads();
#pragma unroll
    for (int i = 0; i < DGC_TREE; ++i) {
      const uint32_t tb = __float_as_uint(T[i]);       // thresholds are >= 0: float order = bit order
      int c = 0;
#pragma unroll
      for (int e = 0; e < NJ * 4; ++e) {
        const int e0 = ((e >> 2) * GB_THREADS + threadIdx.x) * 4 + (e & 3);
        c += (e0 < len) && ab[e] >= tb;
      }
      c = warp_sum_i(c);
      if ((threadIdx.x & 31) == 0 && c) atomicAdd(&s_cnt[i], c);
    }
    __syncthreads();
    if (threadIdx.x < DGC_TREE && s_cnt[threadIdx.x]) atomicAdd(&cnt[seg * 16 + threadIdx.x], (uint32_t)s_cnt[threadIdx.x]);
    __syncthreads();
  };
  if (round == 0) {
    for (int tile = blockIdx.x; tile < t.ntiles; tile += gridDim.x) count_tile(tile);
  } else {
    const int nopen = __ldcg(dir.hdr), total = __ldcg(dir.hdr + 1);
    if (nopen == 0) return;                            // every threshold is final: nothing left to do
    const int cached = min(nopen, DGC_PREF_CACHE);
    for (int i = threadIdx.x; i < cached; i += GB_THREADS) s_pref[i] = __ldcg(dir.pref + i);
    __syncthreads();
    for (int vt = blockIdx.x; vt < total; vt += gridDim.x) count_tile(open_dir_tile(t, dir, s_pref, cached, nopen, vt));
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(done, 1u) == gridDim.x - 1;
  __syncthreads();
  if (s_last) {
    __threadfence();
    for (int seg = threadIdx.x; seg < t.nseg; seg += GB_THREADS) {
      if (__ldcg(sel + seg * 4 + 2)) continue;
      const double n = (double)t.seg_numel[seg];
      float T[DGC_TREE];
      dgc_tree(__uint_as_float(__ldcg(sel + seg * 4)), T);
      int taken = (int)__ldcg(sel + seg * 4 + 1), node = 0;
      float thr = T[0];
      bool fin = false;
      while (node < DGC_TREE && taken < 10) {
        const double selected = (double)__ldcg(cnt + seg * 16 + node);
        thr = T[node];
        if (selected > 1.3 * n * ratio) { thr = __fmul_rn(1.3f, thr); node = 2 * node + 1; ++taken; }
        else if (selected < 0.7 * n * ratio) { thr = __fmul_rn(0.7f, thr); node = 2 * node + 2; ++taken; }
        else { fin = true; break; }
      }
      if (fin || taken >= 10) __stcg(sel + seg * 4 + 2, 1u);
      __stcg(sel + seg * 4, __float_as_uint(thr));
      __stcg(sel + seg * 4 + 1, (uint32_t)taken);
#pragma unroll
      for (int i = 0; i < 16; ++i) __stcg(cnt + seg * 16 + i, 0u);
    }
    __syncthreads();
    open_dir_build(t, dir, [&](int seg) { return __ldcg(sel + seg * 4 + 2) == 0u; });
    if (threadIdx.x == 0) *done = 0;
  }
}


int k1_dgc_presend(gb_plan* p, const float* d_grad, cudaStream_t st) {
  gb_comm* c = p->comm;
  ClipOut out;
  for (int r = 0; r < c->world; ++r) out.slot[r] = (float*)gb_slot_of(c, r, p->chan2_off, p->slot2_bytes, 0, c->rank);
  out.n = c->world;
  k1_dgc_sqsum<<<grid_for(p->ntiles), GB_THREADS, 0, st>>>(p->tab, d_grad, (double*)p->d_tilestats, out, p->d_done + 6, gb_comm_peers(c),
                                                          p->channel2, p->d_ctl, (size_t)c->world * p->slot2_bytes);
  LAUNCHED(p);
  return 0;
}

static int k1_dgc_send(gb_plan* p, const float* d_grad, cudaStream_t st) {
  gb_comm* c = p->comm;
  GB_CHECK(p->prm.memory == GB_MEM_DGC || p->prm.memory == GB_MEM_NONE || p->prm.memory == GB_MEM_RESIDUAL,
           "unsupported memory for DGC");
  ++p->epoch;
  SparseOut out;
  slot_views(p, 0, true, &out, nullptr);
  StepCtl* ctl = p->d_ctl;
  const size_t ps = p->par_stride;
  const PeerPtrs peers = gb_comm_peers(c);
  const int grid = grid_for(p->ntiles);
  float* work;
  float* work2 = nullptr;
  if (p->prm.memory == GB_MEM_DGC && p->slot2_bytes) {
    ClipIn in;
    for (int r = 0; r < c->world; ++r) in.slot[r] = (const float*)gb_slot_of(c, c->rank, p->chan2_off, p->slot2_bytes, 0, r);
    in.n = c->world;
    k1_dgc_accumulate_clip<<<grid, GB_THREADS, 0, st>>>(p->tab, d_grad, p->d_state0, p->d_state1, p->prm.momentum, in,
                                                       (float)c->world, peers, p->channel2, ctl,
                                                       (size_t)c->world * p->slot2_bytes, c->status_dev, c->timeout_ns);
    LAUNCHED(p);
    work = p->d_state0;
    work2 = p->d_state1;
  } else if (p->prm.memory == GB_MEM_DGC) {
    k1_dgc_accumulate<<<148 * 4, GB_THREADS, 0, st>>>(p->tab, d_grad, p->d_state0, p->d_state1, p->prm.momentum); LAUNCHED(p);
    work = p->d_state0;
    work2 = p->d_state1;
  } else if (p->prm.memory == GB_MEM_RESIDUAL) {
    k1_axpby<true><<<148 * 4, GB_THREADS, 0, st>>>(p->tab, d_grad, p->d_state0, p->prm.beta, p->prm.gamma, ctl); LAUNCHED(p);
    work = p->d_state0;
  } else {
    work = const_cast<float*>(d_grad);
  }
  int max_ns = 1;
  for (int s = 0; s < p->nseg; ++s) max_ns = std::max(max_ns, (int)p->h_k[s]);
  GB_CHECK(p->nseg <= 65535 && p->d_samp, "DGC plan: sample scratch missing or too many tensors (%d)", p->nseg);
  k1_dgc_gather<<<dim3((max_ns + GB_THREADS - 1) / GB_THREADS, p->nseg), GB_THREADS, 0, st>>>(p->tab, work, p->prm.seed, ctl, p->d_samp);
  LAUNCHED(p);
  k1_dgc_sample<<<p->nseg, GB_THREADS, 0, st>>>(p->tab, p->d_samp, p->prm.ratio, p->d_sel); LAUNCHED(p);
  for (int it = 0; it < DGC_LAUNCHES; ++it) {           // 4 + 4 + 2 decisions of the reference's ten
    k1_dgc_count<<<grid, GB_THREADS, 0, st>>>(p->tab, work, p->d_sel, p->prm.ratio, p->d_done + 4, p->d_hist, it); LAUNCHED(p);
  }
  uint32_t* cursor = p->d_done + 2;
  GB_CUDA(cudaMemsetAsync(cursor, 0, 4, st));
  if (p->prm.memory != GB_MEM_NONE)
    k1_emit_fast<true><<<grid, GB_THREADS, 0, st>>>(p->tab, work, work2, p->d_sel, 0u, 0, 0, cursor, out, p->ktotal, c->status_dev, p->d_done, peers, p->channel, ctl, ps, 1);
  else
    k1_emit_fast<false><<<grid, GB_THREADS, 0, st>>>(p->tab, work, work2, p->d_sel, 0u, 0, 0, cursor, out, p->ktotal, c->status_dev, p->d_done, peers, p->channel, ctl, ps, 1);
  LAUNCHED(p);
  return 0;
}

// ---------------------------------------------------------------- stand-alone single tensor
// Scratch layout (all 256-byte aligned): tables | hist[2048] | sel[4] | tilecnt[2*ntiles+2] | hdr[ntiles+1]
struct SingleLayout {
  size_t off_tab, off_hist, off_sel, off_cnt, off_hdr, total;
  int ntiles;
};
static SingleLayout single_layout(int64_t n) {
  SingleLayout L;
  L.ntiles = (int)((n + GB_TILE - 1) / GB_TILE);
  size_t cur = 0;
  auto add = [&](size_t b) { size_t o = cur; cur += (b + 255) & ~size_t(255); return o; };
  L.off_tab = add(256 + (size_t)L.ntiles * 4);
  L.off_hist = add(2048 * 4);
  L.off_sel = add(16);
  L.off_cnt = add(((size_t)L.ntiles * 2 + 2) * 4);
  L.off_hdr = add(((size_t)L.ntiles + 1) * 8);
  L.total = cur;
  return L;
}

__global__ void k1_single_setup(char* tab, int64_t n, int64_t k, int ntiles, uint32_t* hist) {
  int64_t* off = (int64_t*)tab;             // [0..1] seg_off, [2..3] kbase
  int32_t* i32 = (int32_t*)(tab + 64);      // numel, k, hash, tile0[0..1]
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    off[0] = 0; off[1] = (n + GB_ALIGN - 1) / GB_ALIGN * GB_ALIGN; off[2] = 0; off[3] = k;
    i32[0] = (int32_t)n; i32[1] = (int32_t)k; i32[2] = 0; i32[3] = 0; i32[4] = ntiles;
  }
  int32_t* tile_seg = (int32_t*)(tab + 256);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < ntiles; i += gridDim.x * blockDim.x) tile_seg[i] = 0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < 2048; i += gridDim.x * blockDim.x) hist[i] = 0;
}

static SegTable single_table(char* tab, int64_t n, int ntiles) {
  SegTable t;
  t.seg_off = (const int64_t*)tab;
  t.seg_kbase = (const int64_t*)tab + 2;
  t.seg_numel = (const int32_t*)(tab + 64);
  t.seg_k = (const int32_t*)(tab + 64) + 1;
  t.seg_hash = (const uint32_t*)(tab + 64) + 2;
  t.seg_tile0 = (const int32_t*)(tab + 64) + 3;
  t.tile_seg = (const int32_t*)(tab + 256);
  t.tile_desc = nullptr;
  t.nseg = 1;
  t.ntiles = ntiles;
  t.padded = (n + GB_ALIGN - 1) / GB_ALIGN * GB_ALIGN;
  return t;
}

extern "C" size_t gb_topk_scratch_bytes(int64_t n) { return single_layout(n).total; }

// shared by gb_topk_compress and the allocation-free variant (k1_static.cu)
int k1_topk_single(const float* d_x, int64_t n, int64_t k, float* d_vals, int32_t* d_idx, char* scratch, cudaStream_t st) {
  GB_CHECK(n >= 1 && n < (int64_t(1) << 31), "n = %lld out of range", (long long)n);
  GB_CHECK(k >= 1 && k <= n, "k = %lld out of range for n = %lld", (long long)k, (long long)n);
  GB_CHECK(((uintptr_t)d_x & 15) == 0, "input must be 16-byte aligned");
  const SingleLayout L = single_layout(n);
  char* tab = scratch + L.off_tab;
  uint32_t* hist = (uint32_t*)(scratch + L.off_hist);
  uint32_t* sel = (uint32_t*)(scratch + L.off_sel);
  int32_t* cnt = (int32_t*)(scratch + L.off_cnt);
  k1_single_setup<<<8, GB_THREADS, 0, st>>>(tab, n, k, L.ntiles, hist);
  GB_CUDA(cudaGetLastError());
  const SegTable t = single_table(tab, n, L.ntiles);
  float* work = const_cast<float*>(d_x);
  const int hgrid = std::min(L.ntiles, 148 * 8), grid = grid_for(L.ntiles);
  k1_hist<0, false><<<hgrid, GB_THREADS, 0, st>>>(t, d_x, work, 0.f, 1.f, hist, sel, nullptr);
  k1_select<<<1, GB_THREADS, 0, st>>>(t, hist, sel, 0);
  k1_hist<1, false><<<hgrid, GB_THREADS, 0, st>>>(t, d_x, work, 0.f, 1.f, hist, sel, nullptr);
  k1_select<<<1, GB_THREADS, 0, st>>>(t, hist, sel, 1);
  k1_hist<2, false><<<hgrid, GB_THREADS, 0, st>>>(t, d_x, work, 0.f, 1.f, hist, sel, nullptr);
  k1_select<<<1, GB_THREADS, 0, st>>>(t, hist, sel, 2);
  k1_count<<<grid, GB_THREADS, 0, st>>>(t, work, sel, 0u, 0, 0, cnt, 0);
  k1_scan<true><<<1, GB_THREADS, 0, st>>>(t, sel, cnt, 0);
  SparseOut out;
  out.n = 1;
  out.vals[0] = d_vals;
  out.idx[0] = d_idx;
  out.hdr[0] = (int2*)(scratch + L.off_hdr);
  PeerPtrs none;
  memset(&none, 0, sizeof(none));
  k1_emit<false><<<grid, GB_THREADS, 0, st>>>(t, work, nullptr, sel, 0u, 0, 0, cnt, out, k, (uint32_t*)(scratch + L.off_sel) + 2, nullptr, none, 0, nullptr, 0, 0, 0);
  GB_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int gb_topk_compress(const float* d_x, int64_t n, int64_t k, float* d_vals, int32_t* d_idx, void* d_scratch,
                                size_t scratch_bytes, void* stream) {
  GB_CHECK(d_scratch && scratch_bytes >= gb_topk_scratch_bytes(n), "scratch too small: %zu < %zu", scratch_bytes,
           gb_topk_scratch_bytes(n));
  GB_CHECK(((uintptr_t)d_scratch & 255) == 0, "scratch must be 256-byte aligned");
  return k1_topk_single(d_x, n, k, d_vals, d_idx, (char*)d_scratch, (cudaStream_t)stream);
}

__global__ void __launch_bounds__(GB_THREADS) k1_scatter_rank(const float* __restrict__ v, const int32_t* __restrict__ ix,
                                                             int64_t count, float scale, float* out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x)
    out[ix[i]] += v[i] * scale;
}

extern "C" int gb_sparse_decompress(const float* const* d_vals, const int32_t* const* d_idx, const int64_t* counts, int nranks,
                                    int64_t n, float scale, float* d_out, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  GB_CUDA(cudaMemsetAsync(d_out, 0, (size_t)n * 4, st));
  for (int r = 0; r < nranks; ++r) {           // one launch per rank keeps the summation order fixed
    if (counts[r] == 0) continue;
    const int g = (int)std::min<int64_t>((counts[r] + GB_THREADS - 1) / GB_THREADS, 148 * 8);
    k1_scatter_rank<<<g, GB_THREADS, 0, st>>>(d_vals[r], d_idx[r], counts[r], scale, d_out);
  }
  GB_CUDA(cudaGetLastError());
  return 0;
}
