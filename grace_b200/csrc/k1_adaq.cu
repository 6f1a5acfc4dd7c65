// K1-adaq -- AdaQ / AdaComp-style sparsifier (SURVEY §2.3 row 15, §2.9-B row "AdaQ").
//
// Reference: grace_dl/tensorflow/compressor/adaq.py:16-93 (TF graph: boolean_mask, a sampled top-k,
// a tf.while_loop of <= 20 threshold refinements per sign, two means, one int32 vector on the wire).
// Oracle of the numerics: grace_b200/compressor/sparse.py (_adaq_part / AdaqCompressor).
//
// Per bucket, for every tensor and both signs at once:
//   aq_count     compensate (ResidualMemory) + count positives / negatives per 128-element sub-block
//   aq_scan      exclusive prefix of the per-tile counts inside each tensor
//   aq_sample    one thread per sample: the r-th positive (negative) element is located through the
//                tile prefix, the 32 sub-block counts of that tile and a scan of one 128-element sub-block -- the
//                reference's boolean_mask + gather without ever materialising the masked tensor
//   aq_kth       k-th largest sampled magnitude (3 histogram rounds, one CTA per tensor and sign)
//   6 x aq_refine_tree   the refinement loop (<= 21 evaluations), device-resident: a launch counts against the current
//                thresholds AND the 14 thresholds the next three decisions can lead to, the CTA that finishes last
//                walks that tree with the x1.25 / x0.9 rule for every (tensor, sign): four evaluations per launch,
//                no host round trips; single-tile tensors are refined to the end by their own CTA
//   aq_select<final> + aq_mean    fixed-order sums -> one mean per tensor and sign, stored to every peer
//   aq_emit      unordered compaction of the picked indices (sign in bit 31) straight into every peer's
//                slot; residual = x - mean at the picked positions; the last CTA releases the flags
//   aq_merge     receive: each CTA owns an output tile and adds mean_r[sign] at rank r's indices in
//                rank order (bit-identical on all ranks)
// Wire per rank: 2 floats per tensor + 4 bytes per kept element (the reference sends the same int32
// list; values never travel).
#include "kernels.cuh"

#include <math.h>

#include <algorithm>
#include <vector>

struct AdaqBuf {
  void* d_all;
  uint8_t* sub;       // [ntiles][2][32]  positives / negatives per 128-element sub-block
  int32_t* tcnt;      // [ntiles][2]      per-tile counts, then exclusive prefix inside the tensor
  int32_t* segcnt;    // [nseg][2]
  int64_t* samp_off;  // [nseg+1]         sample capacity prefix: ceil(0.01 n) + 1 per tensor
  float* samp;        // [2][samp_total]  sampled magnitudes
  uint32_t* sel;      // [nseg][2][4]     {threshold bits, done, selected, adjustments made}
  double* tsum;       // [ntiles][2]      per-tile sum of the finally selected values
  int32_t* tsel;      // [ntiles][2]      per-tile count of the finally selected values
  float* means;       // [nseg][2]
  int64_t samp_total;
  int64_t max_samples;  // largest per-tensor sample capacity
};

struct AdaqOut {
  int2* hdr[GB_MAX_WORLD];
  float* means[GB_MAX_WORLD];
  int32_t* idx[GB_MAX_WORLD];
  int n;
};
__device__ __forceinline__ void shift(AdaqOut& o, size_t bytes) {
  for (int r = 0; r < o.n; ++r) {
    o.hdr[r] = (int2*)((char*)o.hdr[r] + bytes);
    o.means[r] = (float*)((char*)o.means[r] + bytes);
    o.idx[r] = (int32_t*)((char*)o.idx[r] + bytes);
  }
}

size_t aq_means_bytes(int nseg) { return ((size_t)nseg * 2 * 4 + 255) & ~size_t(255); }

namespace {

__device__ __forceinline__ float4 ld4(const float* p, int e0, int len) {
  if (e0 + 4 <= len) return *reinterpret_cast<const float4*>(p + e0);
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (e0 < len) v.x = p[e0];
  if (e0 + 1 < len) v.y = p[e0 + 1];
  if (e0 + 2 < len) v.z = p[e0 + 2];
  if (e0 + 3 < len) v.w = p[e0 + 3];
  return v;
}
__device__ __forceinline__ void st4(float* p, int e0, int len, float4 v) {
  if (e0 + 4 <= len) { *reinterpret_cast<float4*>(p + e0) = v; return; }
  if (e0 < len) p[e0] = v.x;
  if (e0 + 1 < len) p[e0 + 1] = v.y;
  if (e0 + 2 < len) p[e0 + 2] = v.z;
  if (e0 + 3 < len) p[e0 + 3] = v.w;
}
__device__ __forceinline__ float comp1(float beta, float r, float gamma, float g) {
  return __fadd_rn(__fmul_rn(beta, r), __fmul_rn(gamma, g));   // two roundings, like torch's beta*e + gamma*g
}
__device__ __forceinline__ int warp_incl_scan(int v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int n = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += n;
  }
  return v;
}

// ---------------------------------------------------------------- counting
template <bool WRITE>
__global__ void __launch_bounds__(GB_THREADS) aq_count(SegTable t, const float* __restrict__ grad, float* work, float beta,
                                                       float gamma, const StepCtl* ctl, uint8_t* sub, int32_t* tcnt) {
  if (ctl->step == 0) { beta = 0.f; gamma = 1.f; }             // first call: compensate() returns the gradient
  __shared__ int scnt[2];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int tile = blockIdx.x; tile < t.ntiles; tile += gridDim.x) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, tile, seg, base, local0, len);
    if (threadIdx.x < 2) scnt[threadIdx.x] = 0;
    __syncthreads();
#pragma unroll
    for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j) {
      const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (e0 < len) {
        if (WRITE) {
          const float4 g = ld4(grad + base, e0, len);
          const float4 r = ld4(work + base, e0, len);
          v = make_float4(comp1(beta, r.x, gamma, g.x), comp1(beta, r.y, gamma, g.y), comp1(beta, r.z, gamma, g.z),
                          comp1(beta, r.w, gamma, g.w));
          st4(work + base, e0, len, v);
        } else {
          v = ld4(work + base, e0, len);
        }
      }
      int cp = (v.x > 0.f) + (v.y > 0.f) + (v.z > 0.f) + (v.w > 0.f);
      int cn = (v.x < 0.f) + (v.y < 0.f) + (v.z < 0.f) + (v.w < 0.f);
      cp = warp_sum_i(cp);
      cn = warp_sum_i(cn);
      if (lane == 0) {
        sub[((size_t)tile * 2 + 0) * 32 + j * 8 + warp] = (uint8_t)cp;
        sub[((size_t)tile * 2 + 1) * 32 + j * 8 + warp] = (uint8_t)cn;
        if (cp) atomicAdd(&scnt[0], cp);
        if (cn) atomicAdd(&scnt[1], cn);
      }
    }
    __syncthreads();
    if (threadIdx.x < 2) tcnt[tile * 2 + threadIdx.x] = scnt[threadIdx.x];
    __syncthreads();
  }
}

// One CTA per tensor: exclusive prefix of the tile counts (both signs), totals, reset of the selection state.
__global__ void __launch_bounds__(GB_THREADS) aq_scan(SegTable t, int32_t* tcnt, int32_t* segcnt, uint32_t* sel) {
  __shared__ int wsum[GB_THREADS / 32];
  __shared__ int carry;
  const int seg = blockIdx.x;
  const int t0 = t.seg_tile0[seg], t1 = t.seg_tile0[seg + 1];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int s = 0; s < 2; ++s) {
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int c0 = t0; c0 < t1; c0 += GB_THREADS) {
      const int tile = c0 + threadIdx.x;
      const int v = tile < t1 ? tcnt[tile * 2 + s] : 0;
      const int incl = warp_incl_scan(v, lane);
      if (lane == 31) wsum[warp] = incl;
      __syncthreads();
      int woff = 0;
      for (int w = 0; w < warp; ++w) woff += wsum[w];
      const int before = carry;
      if (tile < t1) tcnt[tile * 2 + s] = before + woff + incl - v;
      __syncthreads();
      if (threadIdx.x == GB_THREADS - 1) carry = before + woff + incl;
      __syncthreads();
    }
    if (threadIdx.x == 0) {
      segcnt[seg * 2 + s] = carry;
      uint32_t* q = sel + (seg * 2 + s) * 4;
      q[0] = 0; q[1] = 0; q[2] = 0; q[3] = 0;
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------- sampling
// One THREAD per sample, grid = (sample chunks, tensor x sign).  The reference gathers
// `boolean_mask(x, sign)[pos]`; here the pos-th element of that sign is found through three levels of counts
// (binary search over the tile prefix -> the tile's 32 sub-block counts -> a scan of one 128-element
// sub-block, 16 elements at a time).  A first version used one warp per sample and was latency-bound
// (226 us for ResNet-50's 511 K samples: ~20 dependent loads per warp, 54 waves); with a thread per sample the
// same chains run 32x more in parallel.
__global__ void __launch_bounds__(GB_THREADS) aq_sample(SegTable t, const float* __restrict__ work, const int32_t* __restrict__ tcnt,
                                                        const int32_t* __restrict__ segcnt, const uint8_t* __restrict__ sub,
                                                        const int64_t* __restrict__ samp_off, float* samp, int64_t samp_total,
                                                        uint32_t seed, const StepCtl* ctl) {
  const int seg = blockIdx.y >> 1, s = blockIdx.y & 1;
  const int count = segcnt[seg * 2 + s];
  const int ns = (int)ceil((double)count * 0.01);
  const int i = blockIdx.x * GB_THREADS + threadIdx.x;
  if (i >= ns) return;
  const uint32_t step = ctl->step;
  const Philox4 r4 = philox4x32((uint32_t)(i >> 2), 2u * step + (uint32_t)s, GB_STREAM_SAMPLE, 0u, seed, t.seg_hash[seg]);
  const uint32_t bits = (i & 3) == 0 ? r4.x : (i & 3) == 1 ? r4.y : (i & 3) == 2 ? r4.z : r4.w;
  int r = (int)__fmul_rn(u32_to_unit(bits), (float)count);
  r = min(r, count - 1);
  // tile: the last one whose exclusive prefix is <= r
  int lo = t.seg_tile0[seg], hi = t.seg_tile0[seg + 1];
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (tcnt[mid * 2 + s] <= r) lo = mid; else hi = mid;
  }
  int rr = r - tcnt[lo * 2 + s];
  // sub-block: 32 one-byte counts = two 16-byte loads
  const uint4* sc = reinterpret_cast<const uint4*>(sub + ((size_t)lo * 2 + s) * 32);
  const uint4 c0 = sc[0], c1 = sc[1];
  const uint32_t cw[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
  int sb = 31;
  {
    int acc = 0;
    bool found = false;
#pragma unroll
    for (int k = 0; k < 32; ++k) {
      const int c = (int)((cw[k >> 2] >> (8 * (k & 3))) & 0xFFu);
      if (!found && acc + c > rr) { sb = k; rr -= acc; found = true; }
      acc += c;
    }
  }
  // element inside the 128-element sub-block
  int sg2, local0, len;
  int64_t base;
  tile_bounds(t, lo, sg2, base, local0, len);
  const float* x = work + base;
  float val = 0.f;
  int seen = 0;
  bool done = false;
  for (int e0 = sb * 128; e0 < sb * 128 + 128 && !done; e0 += 16) {
    float4 v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] = (e0 + 4 * k < len) ? ld4(x, e0 + 4 * k, len) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float vv[4] = {v[k].x, v[k].y, v[k].z, v[k].w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const bool hit = s ? vv[q] < 0.f : vv[q] > 0.f;
        if (hit) {
          if (!done && seen == rr) { val = fabsf(vv[q]); done = true; }
          ++seen;
        }
      }
    }
  }
  samp[(int64_t)s * samp_total + samp_off[seg] + i] = val;
}

// Parallel version of "walk the histogram from the top bin down to the bin that holds the k-th largest":
// every thread sums nb/GB_THREADS bins (highest first), a block-wide inclusive scan gives the number of elements
// in strictly higher bins.  Thread 0 walking 2048 shared-memory bins serially cost ~20 us per round.
__device__ __forceinline__ void select_bin(const uint32_t* hist, int nb, uint32_t krem, uint32_t* part, uint32_t* out_bin,
                                           uint32_t* out_krem) {
  const int per = nb / GB_THREADS;
  const int hi = nb - 1 - threadIdx.x * per;
  uint32_t c[8];
  uint32_t mine = 0;
  for (int k = 0; k < per; ++k) { c[k] = hist[hi - k]; mine += c[k]; }
  part[threadIdx.x] = mine;
  __syncthreads();
  for (int o = 1; o < GB_THREADS; o <<= 1) {
    const uint32_t add = threadIdx.x >= o ? part[threadIdx.x - o] : 0u;
    __syncthreads();
    part[threadIdx.x] += add;
    __syncthreads();
  }
  uint32_t running = part[threadIdx.x] - mine;
  for (int k = 0; k < per; ++k) {
    if (running < krem && running + c[k] >= krem) { *out_bin = (uint32_t)(hi - k); *out_krem = krem - running; }
    running += c[k];
  }
  __syncthreads();
}

// One CTA per (tensor, sign): k-th largest of the ns sampled magnitudes -> initial threshold.
__global__ void __launch_bounds__(GB_THREADS) aq_kth(SegTable t, const int32_t* __restrict__ segcnt, const int64_t* __restrict__ samp_off,
                                                     const float* __restrict__ samp, int64_t samp_total, double ratio, uint32_t* sel) {
  __shared__ uint32_t hist[2048];
  __shared__ uint32_t part[GB_THREADS];
  __shared__ uint32_t s_prefix, s_krem, s_bin;
  const int seg = blockIdx.x >> 1, s = blockIdx.x & 1;
  const int count = segcnt[seg * 2 + s];
  uint32_t* q = sel + (seg * 2 + s) * 4;
  if (count == 0) {                                   // empty half: nothing selected, mean 0
    if (threadIdx.x == 0) { q[0] = 0x7F800000u; q[1] = 1u; q[2] = 0u; q[3] = 0u; }
    return;
  }
  const int ns = (int)ceil((double)count * 0.01);
  int k = (int)ceil((double)count * 0.01 * ratio);
  k = max(1, min(ns, k));
  const float* x = samp + (int64_t)s * samp_total + samp_off[seg];
  if (threadIdx.x == 0) { s_prefix = 0; s_krem = (uint32_t)k; }
  for (int round = 0; round < 3; ++round) {
    const int nb = round == 0 ? 2048 : 1024;
    for (int b = threadIdx.x; b < nb; b += blockDim.x) hist[b] = 0;
    __syncthreads();
    const uint32_t prefix = s_prefix;
    for (int i = threadIdx.x; i < ns; i += blockDim.x) {
      const uint32_t b = __float_as_uint(x[i]) & 0x7FFFFFFFu;
      if (round == 0) atomicAdd(&hist[b >> 20], 1u);
      else if (round == 1) { if ((b >> 20) == prefix) atomicAdd(&hist[(b >> 10) & 1023u], 1u); }
      else { if ((b >> 10) == prefix) atomicAdd(&hist[b & 1023u], 1u); }
    }
    __syncthreads();
    const uint32_t krem = s_krem;
    __syncthreads();
    select_bin(hist, nb, krem, part, &s_bin, &s_krem);
    if (threadIdx.x == 0) s_prefix = round == 0 ? s_bin : ((prefix << 10) | s_bin);
    __syncthreads();
  }
  if (threadIdx.x == 0) { q[0] = s_prefix; q[1] = 0u; q[2] = 0u; q[3] = 0u; }
}

// ---------------------------------------------------------------- threshold refinement
// Oracle (sparse.py:_adaq_part): `masked` is |x| on this sign's elements and 0 elsewhere; the first count
// uses `>`, every recount after an adjustment `>=`, the final mask `>` again.
__device__ __forceinline__ double block_sum_d(double v, double* sh) {
  v = warp_sum_d(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  double tot = 0.0;
  for (int w = 0; w < GB_THREADS / 32; ++w) tot += sh[w];          // fixed order
  return tot;
}

__device__ __forceinline__ void aq_adjust_one(uint32_t* q, int count, double ratio) {
  if (q[1]) return;
  const double target = ceil((double)count * ratio);
  const double selected = (double)q[2];
  float thr = __uint_as_float(q[0]);
  const bool over = selected > 1.25 * target, under = selected < 0.8 * target;
  if (q[3] < 20u && (over || under)) {
    thr = over ? __fmul_rn(1.25f, thr) : __fmul_rn(0.9f, thr);
    q[3] += 1u;
    q[2] = 0u;
  } else {
    if (selected < 1.0) thr = __fmul_rn(0.8f, thr);
    q[1] = 1u;
  }
  q[0] = __float_as_uint(thr);
}

// One round of the threshold adjustment (FINAL = false) or the final selection pass (FINAL = true: per-tile counts and
// sums of the selected values).  Adjustment rounds: round 0 counts every tile; the CTA that finishes last applies the
// rule to every (tensor, sign) and lists the tensors with an open sign (OpenDir); later rounds walk only those
// tensors' tiles and return at once when none is open (first version: 21 launches that each walked every tile,
// 14-24 us apiece; one persistent kernel with a grid barrier measured slower than the launches, see k1_dgc_count).
#define AQ_PREF_CACHE 1024
template <bool FINAL>
__global__ void __launch_bounds__(GB_THREADS) aq_select(SegTable t, const float* __restrict__ work, uint32_t* sel, double* tsum,
                                                        int32_t* tsel, const int32_t* __restrict__ segcnt, double ratio,
                                                        uint32_t* done, uint32_t* dir_scratch, int round) {
  __shared__ double shd[GB_THREADS / 32];
  __shared__ int shc[2];
  __shared__ int s_last;
  __shared__ int s_pref[FINAL ? 1 : AQ_PREF_CACHE];
  const OpenDir dir = open_dir(dir_scratch, t.nseg);
  auto do_tile = [&](int tile) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, tile, seg, base, local0, len);
    const uint32_t* q0 = sel + (seg * 2) * 4;
    const uint32_t* q1 = q0 + 4;
    const bool d0 = q0[1] != 0, d1 = q1[1] != 0;
    if (!FINAL && d0 && d1) return;
    if (!FINAL && round == 0 && t.seg_tile0[seg + 1] - t.seg_tile0[seg] == 1) {
      // A tensor of one tile belongs to this CTA alone: all rounds of both signs run here on the values in registers
      // (these small tensors are the ones that keep adjusting; a global round costs ~12 us however little it counts)
      constexpr int NJ = GB_TILE / (GB_THREADS * 4);
      float vals[NJ * 4];
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (e0 < len) v = ld4(work + base, e0, len);
        vals[j * 4] = v.x; vals[j * 4 + 1] = v.y; vals[j * 4 + 2] = v.z; vals[j * 4 + 3] = v.w;
      }
      uint32_t qa[4] = {q0[0], q0[1], q0[2], q0[3]}, qb[4] = {q1[0], q1[1], q1[2], q1[3]};
      for (int it = 0; it <= 20 && !(qa[1] && qb[1]); ++it) {
        const float ta = __uint_as_float(qa[0]), tb = __uint_as_float(qb[0]);
        const bool ga = qa[3] != 0, gb = qb[3] != 0, da = qa[1] != 0, db = qb[1] != 0;
        int c0 = 0, c1 = 0;
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
          const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            if (e0 + q >= len) continue;
            const float x = vals[j * 4 + q];
            const float mp = x > 0.f ? x : 0.f;
            const float mn = x < 0.f ? -x : 0.f;
            c0 += !da && (ga ? mp >= ta : mp > ta);
            c1 += !db && (gb ? mn >= tb : mn > tb);
          }
        }
        c0 = warp_sum_i(c0);
        c1 = warp_sum_i(c1);
        if (threadIdx.x < 2) shc[threadIdx.x] = 0;
        __syncthreads();
        if ((threadIdx.x & 31) == 0) {
          if (c0) atomicAdd(&shc[0], c0);
          if (c1) atomicAdd(&shc[1], c1);
        }
        __syncthreads();
        qa[2] += (uint32_t)shc[0];
        qb[2] += (uint32_t)shc[1];
        __syncthreads();
        aq_adjust_one(qa, segcnt[seg * 2], ratio);
        aq_adjust_one(qb, segcnt[seg * 2 + 1], ratio);
      }
      if (threadIdx.x == 0) {
        uint32_t* w0 = sel + (seg * 2) * 4;
#pragma unroll
        for (int i = 0; i < 4; ++i) { w0[i] = qa[i]; w0[4 + i] = qb[i]; }
      }
      return;
    }
    const float thr0 = __uint_as_float(q0[0]), thr1 = __uint_as_float(q1[0]);
    const bool ge0 = q0[3] != 0, ge1 = q1[3] != 0;
    if (threadIdx.x < 2) shc[threadIdx.x] = 0;
    __syncthreads();
    int c0 = 0, c1 = 0;
    double s0 = 0.0, s1 = 0.0;
#pragma unroll
    for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j) {
      const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
      if (e0 >= len) continue;
      const float4 v = ld4(work + base, e0, len);
      const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (e0 + q >= len) continue;
        const float mp = vv[q] > 0.f ? vv[q] : 0.f;
        const float mn = vv[q] < 0.f ? -vv[q] : 0.f;
        if (FINAL) {
          if (mp > thr0) { ++c0; s0 += (double)vv[q]; }
          if (mn > thr1) { ++c1; s1 += (double)vv[q]; }
        } else {
          c0 += !d0 && (ge0 ? mp >= thr0 : mp > thr0);
          c1 += !d1 && (ge1 ? mn >= thr1 : mn > thr1);
        }
      }
    }
    c0 = warp_sum_i(c0);
    c1 = warp_sum_i(c1);
    if ((threadIdx.x & 31) == 0) {
      if (c0) atomicAdd(&shc[0], c0);
      if (c1) atomicAdd(&shc[1], c1);
    }
    if (FINAL) {
      const double a = block_sum_d(s0, shd);
      const double b = block_sum_d(s1, shd);
      if (threadIdx.x == 0) { tsum[tile * 2] = a; tsum[tile * 2 + 1] = b; }
    }
    __syncthreads();
    if (threadIdx.x < 2) {
      const int c = shc[threadIdx.x];
      if (FINAL) tsel[tile * 2 + threadIdx.x] = c;
      else if (c) atomicAdd(&sel[(seg * 2 + threadIdx.x) * 4 + 2], (uint32_t)c);
    }
    __syncthreads();
  };
  if (FINAL || round == 0) {
    for (int tile = blockIdx.x; tile < t.ntiles; tile += gridDim.x) do_tile(tile);
  } else {
    const int nopen = __ldcg(dir.hdr), total = __ldcg(dir.hdr + 1);
    if (nopen == 0) return;                            // every threshold is final: nothing left to do in this round
    const int cached = min(nopen, AQ_PREF_CACHE);
    for (int i = threadIdx.x; i < cached; i += GB_THREADS) s_pref[i] = __ldcg(dir.pref + i);
    __syncthreads();
    for (int vt = blockIdx.x; vt < total; vt += gridDim.x) do_tile(open_dir_tile(t, dir, s_pref, cached, nopen, vt));
  }
  if (!FINAL) {
    // the CTA that finishes last applies the refinement rule to every (tensor, sign) and rebuilds the directory
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(done, 1u) == gridDim.x - 1;
    __syncthreads();
    if (s_last) {
      __threadfence();
      for (int i = threadIdx.x; i < t.nseg * 2; i += GB_THREADS) {
        const uint4 qv = __ldcg(reinterpret_cast<const uint4*>(sel + i * 4));
        uint32_t q[4] = {qv.x, qv.y, qv.z, qv.w};
        aq_adjust_one(q, segcnt[i], ratio);
        __stcg(reinterpret_cast<uint4*>(sel + i * 4), make_uint4(q[0], q[1], q[2], q[3]));
      }
      __syncthreads();
      open_dir_build(t, dir, [&](int seg) { return __ldcg(sel + (seg * 2) * 4 + 1) == 0u || __ldcg(sel + (seg * 2 + 1) * 4 + 1) == 0u; });
      if (threadIdx.x == 0) *done = 0;
    }
  }
}

// The threshold adjustment with FOUR decisions per launch: a launch counts every element against the current threshold
// of its sign AND the 14 thresholds the next three decisions can lead to (binary tree: child 2i+1 = 1.25f * T[i] after
// "too many", child 2i+2 = 0.9f * T[i] after "too few" -- the products aq_adjust_one would form), and the CTA that
// finishes last walks the tree with the rule of aq_adjust_one.  21 evaluations = 6 launches instead of 21; single-tile
// tensors are refined to the end by their CTA in launch 0; later launches walk the open tensors' tiles only (OpenDir).
// cnt[(seg * 2 + sign) * 16 + node]: values of that sign that pass node's threshold.
#define AQ_TREE 15
#define AQ_TREE_LAUNCHES 6
__device__ __forceinline__ void aq_tree(float thr, float (&T)[AQ_TREE]) {
  T[0] = thr;
#pragma unroll
  for (int i = 0; 2 * i + 2 < AQ_TREE; ++i) { T[2 * i + 1] = __fmul_rn(1.25f, T[i]); T[2 * i + 2] = __fmul_rn(0.9f, T[i]); }
}
__global__ void __launch_bounds__(GB_THREADS) aq_refine_tree(SegTable t, const float* __restrict__ work, uint32_t* sel,
                                                            const int32_t* __restrict__ segcnt, double ratio, uint32_t* done,
                                                            uint32_t* dir_scratch, int round) {
  __shared__ int shc[2];
  __shared__ int s_cnt[2][16];
  __shared__ int s_last;
  __shared__ int s_pref[AQ_PREF_CACHE];
  const OpenDir dir = open_dir(dir_scratch, t.nseg);
  uint32_t* cnt = dir_scratch + ((4 + 2 * t.nseg + 15) & ~15);          // [nseg * 2][16]
  auto do_tile = [&](int tile) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, tile, seg, base, local0, len);
    const uint32_t* q0 = sel + (seg * 2) * 4;
    const uint32_t* q1 = q0 + 4;
    const bool d0 = q0[1] != 0, d1 = q1[1] != 0;
    if (d0 && d1) return;
    constexpr int NJ = GB_TILE / (GB_THREADS * 4);
    float vals[NJ * 4];
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (e0 < len) v = ld4(work + base, e0, len);
      vals[j * 4] = v.x; vals[j * 4 + 1] = v.y; vals[j * 4 + 2] = v.z; vals[j * 4 + 3] = v.w;
    }
    if (round == 0 && t.seg_tile0[seg + 1] - t.seg_tile0[seg] == 1) {
      // a tensor of one tile belongs to this CTA alone: all evaluations of both signs here, values in registers
      uint32_t qa[4] = {q0[0], q0[1], q0[2], q0[3]}, qb[4] = {q1[0], q1[1], q1[2], q1[3]};
      for (int it = 0; it <= 20 && !(qa[1] && qb[1]); ++it) {
        const float ta = __uint_as_float(qa[0]), tb = __uint_as_float(qb[0]);
        const bool ga = qa[3] != 0, gb = qb[3] != 0, da = qa[1] != 0, db = qb[1] != 0;
        int c0 = 0, c1 = 0;
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
          const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            if (e0 + q >= len) continue;
            const float x = vals[j * 4 + q];
            const float mp = x > 0.f ? x : 0.f;
            const float mn = x < 0.f ? -x : 0.f;
            c0 += !da && (ga ? mp >= ta : mp > ta);
            c1 += !db && (gb ? mn >= tb : mn > tb);
          }
        }
        c0 = warp_sum_i(c0);
        c1 = warp_sum_i(c1);
        if (threadIdx.x < 2) shc[threadIdx.x] = 0;
        __syncthreads();
        if ((threadIdx.x & 31) == 0) {
          if (c0) atomicAdd(&shc[0], c0);
          if (c1) atomicAdd(&shc[1], c1);
        }
        __syncthreads();
        qa[2] += (uint32_t)shc[0];
        qb[2] += (uint32_t)shc[1];
        __syncthreads();
        aq_adjust_one(qa, segcnt[seg * 2], ratio);
        aq_adjust_one(qb, segcnt[seg * 2 + 1], ratio);
      }
      if (threadIdx.x == 0) {
        uint32_t* w0 = sel + (seg * 2) * 4;
#pragma unroll
        for (int i = 0; i < 4; ++i) { w0[i] = qa[i]; w0[4 + i] = qb[i]; }
      }
      return;
    }
    if (threadIdx.x < 32) s_cnt[threadIdx.x >> 4][threadIdx.x & 15] = 0;
    __syncthreads();
#pragma unroll
    for (int sgn = 0; sgn < 2; ++sgn) {
      const uint32_t* q = sgn ? q1 : q0;
      if (q[1] != 0) continue;                         // this sign is final (uniform over the CTA)
      float T[AQ_TREE];
      aq_tree(__uint_as_float(q[0]), T);
      const bool ge_root = q[3] != 0;                  // the root compares like aq_select: > before the first adjustment
#pragma unroll
      for (int i = 0; i < AQ_TREE; ++i) {
        int c = 0;
#pragma unroll
        for (int e = 0; e < NJ * 4; ++e) {
          const int e0 = ((e >> 2) * GB_THREADS + threadIdx.x) * 4 + (e & 3);
          const float x = vals[e];
          const float m = sgn ? (x < 0.f ? -x : 0.f) : (x > 0.f ? x : 0.f);
          c += (e0 < len) && ((i == 0 && !ge_root) ? m > T[i] : m >= T[i]);
        }
        c = warp_sum_i(c);
        if ((threadIdx.x & 31) == 0 && c) atomicAdd(&s_cnt[sgn][i], c);
      }
    }
    __syncthreads();
    if (threadIdx.x < 32) {
      const int c = s_cnt[threadIdx.x >> 4][threadIdx.x & 15];
      if (c) atomicAdd(&cnt[(size_t)(seg * 2 + (threadIdx.x >> 4)) * 16 + (threadIdx.x & 15)], (uint32_t)c);
    }
    __syncthreads();
  };
  if (round == 0) {
    for (int tile = blockIdx.x; tile < t.ntiles; tile += gridDim.x) do_tile(tile);
  } else {
    const int nopen = __ldcg(dir.hdr), total = __ldcg(dir.hdr + 1);
    if (nopen == 0) return;                            // every threshold is final: nothing left to do
    const int cached = min(nopen, AQ_PREF_CACHE);
    for (int i = threadIdx.x; i < cached; i += GB_THREADS) s_pref[i] = __ldcg(dir.pref + i);
    __syncthreads();
    for (int vt = blockIdx.x; vt < total; vt += gridDim.x) do_tile(open_dir_tile(t, dir, s_pref, cached, nopen, vt));
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(done, 1u) == gridDim.x - 1;
  __syncthreads();
  if (s_last) {
    __threadfence();
    for (int i = threadIdx.x; i < t.nseg * 2; i += GB_THREADS) {
      const uint4 qv = __ldcg(reinterpret_cast<const uint4*>(sel + i * 4));
      if (qv.y) continue;                              // final already (or refined by its own CTA)
      float T[AQ_TREE];
      aq_tree(__uint_as_float(qv.x), T);
      const double target = ceil((double)segcnt[i] * ratio);
      uint32_t adj = qv.w, fin = 0;
      float thr = T[0];
      int node = 0;
      while (node < AQ_TREE) {
        const double selected = (double)__ldcg(cnt + (size_t)i * 16 + node);
        thr = T[node];
        const bool over = selected > 1.25 * target, under = selected < 0.8 * target;
        if (adj < 20u && (over || under)) {
          thr = over ? __fmul_rn(1.25f, thr) : __fmul_rn(0.9f, thr);
          node = over ? 2 * node + 1 : 2 * node + 2;
          ++adj;
        } else {
          if (selected < 1.0) thr = __fmul_rn(0.8f, thr);
          fin = 1u;
          break;
        }
      }
      __stcg(reinterpret_cast<uint4*>(sel + i * 4), make_uint4(__float_as_uint(thr), fin, 0u, adj));
#pragma unroll
      for (int k = 0; k < 16; ++k) __stcg(cnt + (size_t)i * 16 + k, 0u);
    }
    __syncthreads();
    open_dir_build(t, dir, [&](int seg) { return __ldcg(sel + (seg * 2) * 4 + 1) == 0u || __ldcg(sel + (seg * 2 + 1) * 4 + 1) == 0u; });
    if (threadIdx.x == 0) *done = 0;
  }
}

// One CTA per tensor: fixed-order sum of the tile partials -> mean per sign, stored locally and to every peer.
__global__ void __launch_bounds__(GB_THREADS) aq_mean(SegTable t, const double* __restrict__ tsum, const int32_t* __restrict__ tsel,
                                                      float* means, AdaqOut out, const StepCtl* ctl, size_t par_stride) {
  __shared__ double shd[GB_THREADS / 32];
  __shared__ int shi[GB_THREADS / 32];
  if (ctl->epoch & 1) shift(out, par_stride);
  const int seg = blockIdx.x;
  const int t0 = t.seg_tile0[seg], t1 = t.seg_tile0[seg + 1];
  for (int s = 0; s < 2; ++s) {
    double a = 0.0;
    int c = 0;
    for (int tile = t0 + threadIdx.x; tile < t1; tile += GB_THREADS) { a += tsum[tile * 2 + s]; c += tsel[tile * 2 + s]; }
    a = block_sum_d(a, shd);
    c = warp_sum_i(c);
    if ((threadIdx.x & 31) == 0) shi[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
      int n = 0;
      for (int w = 0; w < GB_THREADS / 32; ++w) n += shi[w];
      const float m = n ? (float)(a / (double)n) : 0.f;
      means[seg * 2 + s] = m;
      for (int r = 0; r < out.n; ++r) out.means[r][seg * 2 + s] = m;
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------- emit / merge
template <bool WRITE>
__global__ void __launch_bounds__(GB_THREADS) aq_emit(SegTable t, float* work, const uint32_t* __restrict__ sel,
                                                      const float* __restrict__ means, uint32_t* cursor, AdaqOut out, int64_t capacity,
                                                      uint32_t* status, uint32_t* done, PeerPtrs peers, int channel,
                                                      const StepCtl* ctl, size_t par_stride) {
  __shared__ int32_t sidx[GB_TILE];
  __shared__ int scount, sbegin;
  // a CTA stages tile after tile and claims the payload cursor once per batch: same-address atomics serialise in L2
  // (see k1_emit_fast)
  constexpr int BATCH = 32;
  __shared__ int b_tile[BATCH], b_off[BATCH], b_cnt[BATCH], b_n;
  __shared__ int s_warp[GB_THREADS / 32];
  const uint32_t epoch = ctl->epoch;
  if (epoch & 1) shift(out, par_stride);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (threadIdx.x == 0) { scount = 0; b_n = 0; }
  __syncthreads();
  auto flush = [&]() {
    if (threadIdx.x == 0) {
      const int tc = scount;
      const int64_t begin = atomicAdd(cursor, (uint32_t)tc);
      int keep = tc;
      if (begin + tc > capacity) {
        atomicOr(status, GB_STATUS_OVERFLOW);
        keep = (int)max((int64_t)0, capacity - begin);
      }
      sbegin = (int)min(begin, capacity);
      for (int i = 0; i < b_n; ++i) {
        const int c = max(0, min(b_cnt[i], keep - b_off[i]));
        for (int r = 0; r < out.n; ++r) out.hdr[r][b_tile[i]] = make_int2(sbegin + (c ? b_off[i] : 0), c);
      }
      scount = keep;
    }
    __syncthreads();
    const int tcnt = scount, begin = sbegin;
    for (int r = 0; r < out.n; ++r) {
      int32_t* di = out.idx[r] + begin;
      for (int i = threadIdx.x; i < tcnt; i += GB_THREADS) di[i] = sidx[i];
    }
    __syncthreads();
    if (threadIdx.x == 0) { scount = 0; b_n = 0; }
    __syncthreads();
  };
  for (int tile = blockIdx.x; tile < t.ntiles; tile += gridDim.x) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, tile, seg, base, local0, len);
    const float thr0 = __uint_as_float(sel[(seg * 2) * 4]), thr1 = __uint_as_float(sel[(seg * 2 + 1) * 4]);
    const float m0 = means[seg * 2], m1 = means[seg * 2 + 1];
    float vv[4][4];
    uint32_t picks = 0, negs = 0;
    int cnt = 0;
#pragma unroll
    for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j) {
      const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (e0 < len) v = ld4(work + base, e0, len);
      vv[j][0] = v.x; vv[j][1] = v.y; vv[j][2] = v.z; vv[j][3] = v.w;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float x = vv[j][q];
        const bool in = e0 + q < len;
        const bool pp = in && x > 0.f && x > thr0;
        const bool pn = in && x < 0.f && -x > thr1;
        picks |= ((pp || pn) ? 1u : 0u) << (j * 4 + q);
        negs |= (pn ? 1u : 0u) << (j * 4 + q);
        cnt += pp || pn;
      }
    }
    const int incl = warp_incl_scan(cnt, lane);
    if (lane == 31) s_warp[wid] = incl;
    __syncthreads();
    int wbase = 0, tile_total = 0;
#pragma unroll
    for (int w = 0; w < GB_THREADS / 32; ++w) {
      const int c = s_warp[w];
      if (w < wid) wbase += c;
      tile_total += c;
    }
    const bool must_flush = b_n > 0 && (scount + tile_total > GB_TILE || b_n == BATCH);
    __syncthreads();
    if (must_flush) flush();
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
              const bool neg = (negs >> (j * 4 + q)) & 1u;
              sidx[pos++] = (local0 + e0 + q) | (neg ? (int32_t)0x80000000 : 0);
              vv[j][q] = __fsub_rn(vv[j][q], neg ? m1 : m0);        // residual = x - decompress(compress(x))
            }
          }
          if (WRITE) st4(work + base, e0, len, make_float4(vv[j][0], vv[j][1], vv[j][2], vv[j][3]));
        }
      }
    }
    if (threadIdx.x == 0) {
      b_tile[b_n] = tile; b_off[b_n] = off0; b_cnt[b_n] = tile_total;
      b_n = b_n + 1;
      scount = off0 + tile_total;
    }
    __syncthreads();
  }
  if (b_n > 0) flush();
  release_flags_last_block(done, peers, channel, epoch);
}

struct AdaqIn {
  const int2* hdr[GB_MAX_WORLD];
  const float* means[GB_MAX_WORLD];
  const int32_t* idx[GB_MAX_WORLD];
  int n;
};

__global__ void __launch_bounds__(GB_THREADS) aq_merge(SegTable t, AdaqIn in, float* outp, float divisor, PeerPtrs peers, int channel,
                                                       const StepCtl* ctl, size_t par_stride, uint32_t* status, uint64_t timeout_ns) {
  __shared__ float acc[GB_TILE];
  const uint32_t epoch = ctl->epoch;
  if (epoch & 1) {
    for (int r = 0; r < in.n; ++r) {
      in.hdr[r] = (const int2*)((const char*)in.hdr[r] + par_stride);
      in.means[r] = (const float*)((const char*)in.means[r] + par_stride);
      in.idx[r] = (const int32_t*)((const char*)in.idx[r] + par_stride);
    }
  }
  wait_all_ranks(peers, channel, epoch, status, timeout_ns);
  for (int tile = blockIdx.x; tile < t.ntiles; tile += gridDim.x) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, tile, seg, base, local0, len);
    // headers and means of all ranks, then all ranks' first GB_THREADS indices, are loaded before the adds
    int2 h[GB_MAX_WORLD];
    float m0[GB_MAX_WORLD], m1[GB_MAX_WORLD];
#pragma unroll
    for (int r = 0; r < GB_MAX_WORLD; ++r) {
      const int rr = r < in.n ? r : 0;
      h[r] = in.hdr[rr][tile];
      m0[r] = in.means[rr][seg * 2];
      m1[r] = in.means[rr][seg * 2 + 1];
    }
    int32_t pw[GB_MAX_WORLD];
    bool have[GB_MAX_WORLD];
#pragma unroll
    for (int r = 0; r < GB_MAX_WORLD; ++r) {
      have[r] = r < in.n && threadIdx.x < h[r].y;
      pw[r] = have[r] ? in.idx[r][h[r].x + threadIdx.x] : 0;
    }
    for (int i = threadIdx.x; i < GB_TILE; i += GB_THREADS) acc[i] = 0.f;
    __syncthreads();
#pragma unroll
    for (int r = 0; r < GB_MAX_WORLD; ++r) {
      if (r < in.n) {
        if (have[r]) {
          const int l0 = (pw[r] & 0x7FFFFFFF) - local0;
          if ((unsigned)l0 < (unsigned)GB_TILE) acc[l0] += pw[r] < 0 ? m1[r] : m0[r];   // indices are unique per rank
        }
        const int32_t* ix = in.idx[r] + h[r].x;
        for (int i = threadIdx.x + GB_THREADS; i < h[r].y; i += GB_THREADS) {
          const int32_t w = ix[i];
          const int li = (w & 0x7FFFFFFF) - local0;
          if ((unsigned)li < (unsigned)GB_TILE) acc[li] += w < 0 ? m1[r] : m0[r];
        }
        __syncthreads();
      }
    }
#pragma unroll
    for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j) {
      const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
      if (e0 >= len) continue;
      float4 v = make_float4(acc[e0], acc[e0 + 1], acc[e0 + 2], acc[e0 + 3]);
      if (divisor != 1.f) {          // torch's CUDA tensor / python-scalar is a multiply by the reciprocal
        const float inv = __fdiv_rn(1.0f, divisor);
        v.x = __fmul_rn(v.x, inv); v.y = __fmul_rn(v.y, inv); v.z = __fmul_rn(v.z, inv); v.w = __fmul_rn(v.w, inv);
      }
      st4(outp + base, e0, len, v);
    }
    __syncthreads();
  }
}

}  // namespace

// ---------------------------------------------------------------- host side
#define AQ_LAUNCHED(p) do { (p)->launches++; GB_CUDA(cudaGetLastError()); } while (0)

int aq_init(gb_plan* p) {
  GB_CHECK(p->nseg * 2 <= 65535, "AdaQ plan: too many tensors in one bucket (%d)", p->nseg);
  AdaqBuf* b = new AdaqBuf();
  b->max_samples = 1;
  p->adaq = b;
  std::vector<int64_t> off(p->nseg + 1);
  int64_t tot = 0;
  for (int s = 0; s < p->nseg; ++s) {
    off[s] = tot;
    const int64_t cap = (int64_t)ceil((double)p->h_numel[s] * 0.01) + 1;
    tot += cap;
    b->max_samples = std::max(b->max_samples, cap);
  }
  off[p->nseg] = tot;
  b->samp_total = tot;
  size_t bytes = 0;
  auto add = [&](size_t n) { size_t o = bytes; bytes += (n + 255) & ~size_t(255); return o; };
  const size_t o_sub = add((size_t)p->ntiles * 64), o_tcnt = add((size_t)p->ntiles * 8), o_seg = add((size_t)p->nseg * 8);
  const size_t o_off = add((size_t)(p->nseg + 1) * 8), o_samp = add((size_t)tot * 2 * 4), o_sel = add((size_t)p->nseg * 32);
  const size_t o_tsum = add((size_t)p->ntiles * 16), o_tsel = add((size_t)p->ntiles * 8), o_means = add((size_t)p->nseg * 8);
  GB_CUDA(cudaMalloc(&b->d_all, bytes));
  GB_CUDA(cudaMemset(b->d_all, 0, bytes));
  char* base = (char*)b->d_all;
  b->sub = (uint8_t*)(base + o_sub); b->tcnt = (int32_t*)(base + o_tcnt); b->segcnt = (int32_t*)(base + o_seg);
  b->samp_off = (int64_t*)(base + o_off); b->samp = (float*)(base + o_samp); b->sel = (uint32_t*)(base + o_sel);
  b->tsum = (double*)(base + o_tsum); b->tsel = (int32_t*)(base + o_tsel); b->means = (float*)(base + o_means);
  GB_CUDA(cudaMemcpy(b->samp_off, off.data(), (size_t)(p->nseg + 1) * 8, cudaMemcpyHostToDevice));
  return 0;
}

void aq_free(gb_plan* p) {
  AdaqBuf* b = (AdaqBuf*)p->adaq;
  if (!b) return;
  cudaFree(b->d_all);
  delete b;
  p->adaq = nullptr;
}

static int aq_grid(int ntiles) { return ntiles < 148 * 8 ? ntiles : 148 * 8; }

int aq_send(gb_plan* p, const float* d_grad, cudaStream_t st) {
  gb_comm* c = p->comm;
  AdaqBuf* b = (AdaqBuf*)p->adaq;
  GB_CHECK(p->prm.memory == GB_MEM_NONE || p->prm.memory == GB_MEM_RESIDUAL, "unsupported memory for AdaQ");
  const bool mem = p->prm.memory == GB_MEM_RESIDUAL;
  float* work = mem ? p->d_state0 : const_cast<float*>(d_grad);
  ++p->epoch;
  AdaqOut out;
  const size_t mb = aq_means_bytes(p->nseg);
  for (int r = 0; r < c->world; ++r) {
    char* s = gb_slot_of(c, r, p->chan_off, p->slot_bytes, 0, c->rank);
    out.hdr[r] = (int2*)s; out.means[r] = (float*)(s + p->hdr_bytes); out.idx[r] = (int32_t*)(s + p->hdr_bytes + mb);
  }
  out.n = c->world;
  StepCtl* ctl = p->d_ctl;
  const PeerPtrs peers = gb_comm_peers(c);
  const int grid = aq_grid(p->ntiles);
  if (mem) aq_count<true><<<grid, GB_THREADS, 0, st>>>(p->tab, d_grad, work, p->prm.beta, p->prm.gamma, ctl, b->sub, b->tcnt);
  else aq_count<false><<<grid, GB_THREADS, 0, st>>>(p->tab, d_grad, work, 0.f, 1.f, ctl, b->sub, b->tcnt);
  AQ_LAUNCHED(p);
  aq_scan<<<p->nseg, GB_THREADS, 0, st>>>(p->tab, b->tcnt, b->segcnt, b->sel); AQ_LAUNCHED(p);
  const dim3 sgrid((unsigned)((b->max_samples + GB_THREADS - 1) / GB_THREADS), (unsigned)(p->nseg * 2));
  aq_sample<<<sgrid, GB_THREADS, 0, st>>>(p->tab, work, b->tcnt, b->segcnt, b->sub, b->samp_off, b->samp, b->samp_total,
                                           p->prm.seed, ctl); AQ_LAUNCHED(p);
  aq_kth<<<p->nseg * 2, GB_THREADS, 0, st>>>(p->tab, b->segcnt, b->samp_off, b->samp, b->samp_total, p->prm.ratio, b->sel);
  AQ_LAUNCHED(p);
  uint32_t* adj_done = p->d_done + 4;
  if (getenv("GRACE_B200_ADAQ_ROUND_PER_LAUNCH")) {      // the earlier shape (one evaluation per launch), kept for A/B timing
    for (int it = 0; it <= 20; ++it) {
      aq_select<false><<<grid, GB_THREADS, 0, st>>>(p->tab, work, b->sel, b->tsum, b->tsel, b->segcnt, p->prm.ratio, adj_done, p->d_hist, it);
      AQ_LAUNCHED(p);
    }
  } else {
    for (int it = 0; it < AQ_TREE_LAUNCHES; ++it) {       // four evaluations per launch: 6 x 4 >= 21
      aq_refine_tree<<<grid, GB_THREADS, 0, st>>>(p->tab, work, b->sel, b->segcnt, p->prm.ratio, adj_done, p->d_hist, it);
      AQ_LAUNCHED(p);
    }
  }
  aq_select<true><<<grid, GB_THREADS, 0, st>>>(p->tab, work, b->sel, b->tsum, b->tsel, b->segcnt, p->prm.ratio, adj_done, p->d_hist, 0);
  AQ_LAUNCHED(p);
  aq_mean<<<p->nseg, GB_THREADS, 0, st>>>(p->tab, b->tsum, b->tsel, b->means, out, ctl, p->par_stride); AQ_LAUNCHED(p);
  uint32_t* cursor = p->d_done + 2;
  GB_CUDA(cudaMemsetAsync(cursor, 0, 4, st));
  if (mem) aq_emit<true><<<grid, GB_THREADS, 0, st>>>(p->tab, work, b->sel, b->means, cursor, out, p->ktotal, c->status_dev, p->d_done, peers, p->channel, ctl, p->par_stride);
  else aq_emit<false><<<grid, GB_THREADS, 0, st>>>(p->tab, work, b->sel, b->means, cursor, out, p->ktotal, c->status_dev, p->d_done, peers, p->channel, ctl, p->par_stride);
  AQ_LAUNCHED(p);
  return 0;
}

int aq_recv(gb_plan* p, float* d_out, cudaStream_t st) {
  gb_comm* c = p->comm;
  AdaqIn in;
  const size_t mb = aq_means_bytes(p->nseg);
  for (int r = 0; r < c->world; ++r) {
    const char* s = gb_slot_of(c, c->rank, p->chan_off, p->slot_bytes, 0, r);
    in.hdr[r] = (const int2*)s; in.means[r] = (const float*)(s + p->hdr_bytes); in.idx[r] = (const int32_t*)(s + p->hdr_bytes + mb);
  }
  in.n = c->world;
  const float divisor = p->prm.average ? (float)c->world : 1.f;
  aq_merge<<<aq_grid(p->ntiles), GB_THREADS, 0, st>>>(p->tab, in, d_out, divisor, gb_comm_peers(c), p->channel, p->d_ctl,
                                                      p->par_stride, c->status_dev, c->timeout_ns);
  AQ_LAUNCHED(p);
  return 0;
}
