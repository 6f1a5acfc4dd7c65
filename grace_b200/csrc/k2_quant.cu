// K2 -- quantiser family: None, FP16, QSGD, TernGrad, Natural, 8-bit, SignSGD, Signum,
// EFSignSGD, OneBit.
//
// Per bucket instead of per tensor, replaces the reference op chains of
//   compressor/{qsgd.py:12-38, terngrad.py:8-32, natural.py:13-40, signsgd.py:11-30,
//   signum.py:13-37, efsignsgd.py:12-33}, torch/compressor/onebit.py:8-32,
//   tensorflow/compressor/u8bit.py:10-110, memory/{residual.py,efsignsgd.py} and
//   communicator/allgather.py:8-45
// by (SURVEY §2.9-B quantiser rows):
//   send:  k2_stats     one pass, per-tile partial reductions of the compensated gradient
//          k2_finalize  per-segment scalars (norm / scalar / means), summed in a fixed order,
//                       stored into every peer's slot header
//          k2_encode    Philox + quantise + pack BITS-per-element words in registers + P2P
//                       stores into every peer + residual / momentum update; last CTA
//                       releases the flags
//   recv:  k2_decode    acquire W flags; unpack the W ranks' codes, decode with the sender's
//                       scalars, add in rank order (majority vote for sign codecs), write once
// Wire formats are true bit-packs (1 bit sign codecs, 2-bit TernGrad, 4/5/6/7/8/16-bit QSGD) instead
// of the reference's one byte per element.
#include "kernels.cuh"
#include <cuda_fp16.h>

__device__ const float g_u8_table[128] = {
    1.5000001e-06f, 2.75e-06f, 7.2499997e-06f, 1.8750001e-05f, 3.625e-05f, 5.8749996e-05f, 8.6249995e-05f, 1.4375e-04f,
    2.3125e-04f, 3.1875001e-04f, 4.0625001e-04f, 5.1874999e-04f, 6.5624999e-04f, 7.9374999e-04f, 9.3124999e-04f,
    1.21875e-03f, 1.65625e-03f, 2.0937501e-03f, 2.53125e-03f, 2.96875e-03f, 3.4062499e-03f, 3.8437501e-03f,
    4.2812498e-03f, 4.84375e-03f, 5.53125e-03f, 6.21875e-03f, 6.90625e-03f, 7.59375e-03f, 8.2812496e-03f, 8.96875e-03f,
    9.6562495e-03f, 1.109375e-02f, 1.328125e-02f, 1.546875e-02f, 1.765625e-02f, 1.984375e-02f, 2.2031249e-02f,
    2.4218749e-02f, 2.6406251e-02f, 2.8593751e-02f, 3.078125e-02f, 3.2968748e-02f, 3.515625e-02f, 3.7343752e-02f,
    3.953125e-02f, 4.1718751e-02f, 4.3906249e-02f, 4.671875e-02f, 5.0156251e-02f, 5.3593751e-02f, 5.7031251e-02f,
    6.0468748e-02f, 6.3906237e-02f, 6.7343749e-02f, 7.0781253e-02f, 7.4218743e-02f, 7.7656247e-02f, 8.1093743e-02f,
    8.453124e-02f, 8.7968737e-02f, 9.1406241e-02f, 9.4843738e-02f, 9.8281242e-02f, 1.0546875e-01f, 1.1640625e-01f,
    1.2734374e-01f, 1.3828126e-01f, 1.4921875e-01f, 1.6015625e-01f, 1.7109375e-01f, 1.8203124e-01f, 1.9296876e-01f,
    2.0390625e-01f, 2.1484375e-01f, 2.2578125e-01f, 2.3671874e-01f, 2.4765626e-01f, 2.5859374e-01f, 2.6953125e-01f,
    2.8046876e-01f, 2.9140624e-01f, 3.0234376e-01f, 3.1328124e-01f, 3.2421875e-01f, 3.3515626e-01f, 3.4609374e-01f,
    3.5703126e-01f, 3.6796874e-01f, 3.7890625e-01f, 3.8984376e-01f, 4.0078124e-01f, 4.1171876e-01f, 4.2265624e-01f,
    4.3359375e-01f, 4.4453126e-01f, 4.5859376e-01f, 4.7578123e-01f, 4.9296874e-01f, 5.1015621e-01f, 5.2734375e-01f,
    5.4453123e-01f, 5.617187e-01f, 5.7890624e-01f, 5.9609371e-01f, 6.1328125e-01f, 6.3046873e-01f, 6.476562e-01f,
    6.6484374e-01f, 6.8203121e-01f, 6.9921869e-01f, 7.1640623e-01f, 7.335937e-01f, 7.5078118e-01f, 7.6796871e-01f,
    7.8515619e-01f, 8.0234367e-01f, 8.195312e-01f, 8.3671868e-01f, 8.5390615e-01f, 8.7109369e-01f, 8.8828117e-01f,
    9.0546864e-01f, 9.2265618e-01f, 9.3984365e-01f, 9.5703113e-01f, 9.7421867e-01f, 9.9140614e-01f, 9.9570298e-01f};

struct QuantArgs {
  int codec;
  int has_mem;        // residual buffer present
  float beta, gamma;  // compensate
  float mom;          // signum momentum (0 on the first call)
  float om;           // (float)(1 - momentum), rounded from double like python's scalar
  int quantum;        // qsgd levels
  uint32_t seed, step;
};

struct WireOut { char* slot[GB_MAX_WORLD]; int n; int mc; };   // my slot inside every peer; mc: slot[0] is the NVLS multicast view, n == 1
struct WireIn { const char* slot[GB_MAX_WORLD]; int n; };  // every rank's slot inside my arena

// first-call rules (empty memory: compensate() is the identity; Signum: m = x) and the Philox step
// come from the device-side control block
__device__ __forceinline__ QuantArgs resolve(QuantArgs a, const StepCtl* ctl) {
  const uint32_t s = ctl->step;
  if (s == 0) { a.beta = 0.f; a.gamma = 1.f; a.mom = 0.f; }
  a.step = s;
  return a;
}
__device__ __forceinline__ void shift(WireOut& o, size_t bytes) {
  for (int r = 0; r < o.n; ++r) o.slot[r] += bytes;
}
__device__ __forceinline__ void shift(WireIn& o, size_t bytes) {
  for (int r = 0; r < o.n; ++r) o.slot[r] += bytes;
}

__device__ __forceinline__ float comp1(const QuantArgs& a, float r, float g) {
  return a.has_mem ? __fadd_rn(__fmul_rn(a.beta, r), __fmul_rn(a.gamma, g)) : g;
}

// ---------------------------------------------------------------- pass 1: per-tile partials
// tilestats[tile][8] = { sum x, sum x^2, max|x|, sum|x|, count(x<0), sum(x<0), M2 about the tile mean, - }
// Only the reductions the codec's scalars need are computed (QSGD: sum x^2; TernGrad: sum, max, M2;
// EFSignSGD: sum|x|; OneBit: sum, count/sum of negatives; 8-bit: max) -- the generic version spent more
// issue slots on six shuffle reductions per tile than on the loads.  Unused tilestats fields stay zero.
template <int CODEC>
__global__ void __launch_bounds__(GB_THREADS, 4) k2_stats(SegTable t, const float* __restrict__ grad,
                                                                                          const float* __restrict__ res, QuantArgs a_in,
                                                                                          const StepCtl* ctl, float* tilestats) {
  constexpr bool NEED_M2 = CODEC == GB_CODEC_TERNGRAD;
  constexpr bool W_S = CODEC == GB_CODEC_TERNGRAD || CODEC == GB_CODEC_ONEBIT;
  constexpr bool W_S2 = CODEC == GB_CODEC_QSGD;
  constexpr bool W_MX = CODEC == GB_CODEC_TERNGRAD || CODEC == GB_CODEC_U8BIT;
  constexpr bool W_SA = CODEC == GB_CODEC_EFSIGNSGD;
  constexpr bool W_NEG = CODEC == GB_CODEC_ONEBIT;
  __shared__ float sh[6][GB_THREADS / 32];
  const QuantArgs a = resolve(a_in, ctl);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int tile = blockIdx.x; tile < t.ntiles; tile += gridDim.x) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, tile, seg, base, local0, len);
    float s = 0.f, s2 = 0.f, mx = 0.f, sa = 0.f, cn = 0.f, sn = 0.f;
    float keep[GB_TILE / GB_THREADS];                 // this thread's 16 values stay in registers
    float4 gg[GB_TILE / (GB_THREADS * 4)], rr[GB_TILE / (GB_THREADS * 4)];
#pragma unroll
    for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j) {       // all loads of the tile in flight first
      const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
      gg[j] = make_float4(0.f, 0.f, 0.f, 0.f);
      rr[j] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (e0 < len) {
        gg[j] = *reinterpret_cast<const float4*>(grad + base + e0);
        if (a.has_mem) rr[j] = *reinterpret_cast<const float4*>(res + base + e0);
      }
    }
#pragma unroll
    for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j) {
      const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
      const float x[4] = {comp1(a, rr[j].x, gg[j].x), comp1(a, rr[j].y, gg[j].y), comp1(a, rr[j].z, gg[j].z), comp1(a, rr[j].w, gg[j].w)};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (NEED_M2) keep[j * 4 + q] = x[q];
        if (e0 + q < len) {
          const float v = x[q];
          if (W_S) s += v;
          if (W_S2) s2 += v * v;
          if (W_MX) mx = fmaxf(mx, fabsf(v));
          if (W_SA) sa += fabsf(v);
          if (W_NEG && v < 0.f) { cn += 1.f; sn += v; }
        }
      }
    }
    if (W_S) { s = warp_sum(s); if (lane == 0) sh[0][wid] = s; }
    if (W_S2) { s2 = warp_sum(s2); if (lane == 0) sh[1][wid] = s2; }
    if (W_MX) { mx = warp_max(mx); if (lane == 0) sh[2][wid] = mx; }
    if (W_SA) { sa = warp_sum(sa); if (lane == 0) sh[3][wid] = sa; }
    if (W_NEG) { cn = warp_sum(cn); sn = warp_sum(sn); if (lane == 0) { sh[4][wid] = cn; sh[5][wid] = sn; } }
    __syncthreads();
    float tile_sum = 0.f;
    if (NEED_M2)
      for (int w = 0; w < GB_THREADS / 32; ++w) tile_sum += sh[0][w];
    if (threadIdx.x < 6) {
      const int f = threadIdx.x;
      const bool used = (f == 0 && W_S) || (f == 1 && W_S2) || (f == 2 && W_MX) || (f == 3 && W_SA) || (f >= 4 && W_NEG);
      if (used) {
        float acc = 0.f;
        for (int w = 0; w < GB_THREADS / 32; ++w) acc = f == 2 ? fmaxf(acc, sh[2][w]) : acc + sh[f][w];
        tilestats[(size_t)tile * 8 + f] = acc;
      }
    }
    __syncthreads();
    if (NEED_M2) {
      // centred second moment about the tile mean (combined across tiles by Chan's formula in
      // k2_finalize): sum x^2 - n mean^2 would cancel catastrophically
      const float tmean = tile_sum / (float)len;
      float m2 = 0.f;
#pragma unroll
      for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j) {
        const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (e0 + q < len) { const float d = keep[j * 4 + q] - tmean; m2 += d * d; }
      }
      m2 = warp_sum(m2);
      if (lane == 0) sh[0][wid] = m2;
      __syncthreads();
      if (threadIdx.x == 0) {
        float acc = 0.f;
        for (int w = 0; w < GB_THREADS / 32; ++w) acc += sh[0][w];
        tilestats[(size_t)tile * 8 + 6] = acc;
      }
      __syncthreads();
    }
  }
}

// ---------------------------------------------------------------- pass 1b: per-segment scalars
// One warp per segment; tiles are combined in a fixed order in double precision, so the
// scalars are reproducible run to run.  segstats[seg][4] is codec specific:
//   QSGD {norm}  TERNGRAD {scalar}  EFSIGNSGD {mean|x|}  ONEBIT {mean_neg, mean_pos}  U8BIT {max|x|}
__global__ void k2_finalize(SegTable t, const float* __restrict__ tilestats, int codec, float* segstats, WireOut out,
                            const StepCtl* ctl, size_t par_stride) {
  const size_t poff = (ctl->epoch & 1) ? par_stride : 0;      // parity slot; the parameter struct stays in the constant bank
  const int seg = blockIdx.x;
  const int lane = threadIdx.x;
  const int first = t.seg_tile0[seg], last = t.seg_tile0[seg + 1];
  double s = 0, s2 = 0, sa = 0, cn = 0, sn = 0, m2 = 0;
  float mx = 0.f;
  for (int tile = first + lane; tile < last; tile += 32) {
    const float* p = tilestats + (size_t)tile * 8;
    s += p[0]; s2 += p[1]; mx = fmaxf(mx, p[2]); sa += p[3]; cn += p[4]; sn += p[5]; m2 += p[6];
  }
  s = warp_sum_d(s); s2 = warp_sum_d(s2); sa = warp_sum_d(sa); cn = warp_sum_d(cn); sn = warp_sum_d(sn);
  m2 = warp_sum_d(m2);
  mx = warp_max(mx);
  if (codec == GB_CODEC_TERNGRAD) {               // + sum_t n_t (mean_t - mean)^2
    const double mean = s / (double)t.seg_numel[seg];
    double between = 0;
    for (int tile = first + lane; tile < last; tile += 32) {
      const int nt = min(GB_TILE, t.seg_numel[seg] - (tile - first) * GB_TILE);
      const double d = (double)tilestats[(size_t)tile * 8] / nt - mean;
      between += nt * d * d;
    }
    m2 += warp_sum_d(between);
  }
  if (lane == 0) {
    const double n = (double)t.seg_numel[seg];
    float o[4] = {0.f, 0.f, 0.f, 0.f};
    if (codec == GB_CODEC_QSGD) {
      o[0] = (float)sqrt(s2);
    } else if (codec == GB_CODEC_TERNGRAD) {
      o[0] = fminf(mx, (float)(2.5 * sqrt(m2 / n)));
    } else if (codec == GB_CODEC_EFSIGNSGD) {
      o[0] = (float)(sa / n);
    } else if (codec == GB_CODEC_ONEBIT) {
      const double sp = s - sn, np = n - cn;
      o[0] = (float)(cn > 0 ? sn / cn : sn);
      o[1] = (float)(np > 0 ? sp / np : sp);
    } else if (codec == GB_CODEC_U8BIT) {
      o[0] = mx;
    }
    const float4 v = make_float4(o[0], o[1], o[2], o[3]);
    reinterpret_cast<float4*>(segstats)[seg] = v;
    for (int r = 0; r < out.n; ++r)
      wire_st_u4(reinterpret_cast<float4*>(out.slot[r] + poff) + seg,
                 make_uint4(__float_as_uint(v.x), __float_as_uint(v.y), __float_as_uint(v.z), __float_as_uint(v.w)), out.mc);
  }
}

// ---------------------------------------------------------------- codec arithmetic
// encode: value (+ uniform draw) -> BITS-bit code; decode: code + sender's scalars -> float.
// Per-(segment, sender) constants that the element loops must not recompute: IEEE divisions run on the
// XU pipe and were the top cost of the first version (ncu: 44 % XU utilisation in k2_encode<QSGD>).
struct CodecConst { float enc, dec; };
template <int CODEC>
__device__ __forceinline__ CodecConst codec_const(const float4& st, int quantum) {
  CodecConst c = {0.f, 0.f};
  if (CODEC == GB_CODEC_QSGD) {
    c.enc = st.x > 0.f ? __fdiv_rn((float)quantum, st.x) : 0.f;     // s / norm
    c.dec = __fdiv_rn(st.x, (float)quantum);                        // norm / s
  } else if (CODEC == GB_CODEC_U8BIT) {
    c.enc = st.x > 0.f ? __fdiv_rn(1.0f, st.x) : 0.f;
  }
  return c;
}

template <int CODEC>
__device__ __forceinline__ uint32_t encode1(float x, const float4& st, uint32_t rnd, const CodecConst& cc, const float* tab) {
  if (CODEC == GB_CODEC_NONE) return __float_as_uint(x);
  if (CODEC == GB_CODEC_FP16) return (uint32_t)__half_as_ushort(__float2half_rn(x));
  if (CODEC == GB_CODEC_SIGNSGD || CODEC == GB_CODEC_SIGNUM || CODEC == GB_CODEC_EFSIGNSGD) return x >= 0.f ? 1u : 0u;
  if (CODEC == GB_CODEC_ONEBIT) return x < 0.f ? 1u : 0u;
  if (CODEC == GB_CODEC_TERNGRAD) {
    const float scalar = st.x;
    const float mag = fminf(fabsf(x), scalar);
    const bool keep = __fmul_rn(u32_to_unit(rnd), scalar) < mag;
    if (!keep || x == 0.f) return 0u;
    return x > 0.f ? 1u : 3u;                       // 2-bit two's complement of +1 / -1
  }
  if (CODEC == GB_CODEC_QSGD) {
    const float level = __fmul_rn(cc.enc, fabsf(x));
    const float low = floorf(level);
    const float q = low + (u32_to_unit(rnd) < (level - low) ? 1.f : 0.f);
    const int sq = x > 0.f ? (int)q : (x < 0.f ? -(int)q : 0);
    return (uint32_t)sq;                            // caller masks to BITS
  }
  if (CODEC == GB_CODEC_NATURAL) {
    const uint32_t b = __float_as_uint(x);
    const uint32_t sign = b >> 31, mant = b & 0x7FFFFFu;
    int e = (int)((b >> 23) & 0xFFu) + (mant > (rnd >> 9) ? 1 : 0);
    e = min(max(e, 18), 145);
    return (sign << 7) | (uint32_t)(e - 18);
  }
  if (CODEC == GB_CODEC_U8BIT) {
    const float y = __fmul_rn(fabsf(x), cc.enc);
    int lo = 0, hi = 128;                           // upper_bound(table, y)
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (tab[mid] <= y) lo = mid + 1; else hi = mid;
      }
    }
    int bin = min(max(lo - 1, 0), 127);
    const int code = x > 0.f ? bin : (x < 0.f ? -bin : 0);
    return (uint32_t)code;
  }
  return 0u;
}

template <int CODEC, int BITS>
__device__ __forceinline__ float decode1(uint32_t code, const float4& st, const CodecConst& cc, const float* tab) {
  if (CODEC == GB_CODEC_NONE) return __uint_as_float(code);
  if (CODEC == GB_CODEC_FP16) return __half2float(__ushort_as_half((unsigned short)code));
  if (CODEC == GB_CODEC_SIGNSGD || CODEC == GB_CODEC_SIGNUM) return code ? 1.f : -1.f;
  if (CODEC == GB_CODEC_EFSIGNSGD) return __fmul_rn(st.x, code ? 1.f : -1.f);
  if (CODEC == GB_CODEC_ONEBIT) return code ? st.x : st.y;
  const int sv = (int)(code << (32 - BITS)) >> (32 - BITS);     // sign-extend BITS
  if (CODEC == GB_CODEC_TERNGRAD) return __fmul_rn((float)sv, st.x);
  if (CODEC == GB_CODEC_QSGD) return __fmul_rn(cc.dec, (float)sv);
  if (CODEC == GB_CODEC_NATURAL) {
    const uint32_t e = code & 0x7Fu;
    const float mag = __uint_as_float((e + 18u) << 23);
    return e == 0u ? 0.f : ((code & 0x80u) ? -mag : mag);
  }
  if (CODEC == GB_CODEC_U8BIT) {
    const int a = sv < 0 ? -sv : sv;
    const float mag = __fmul_rn(tab[min(a, 127)], st.x);
    return sv > 0 ? mag : (sv < 0 ? -mag : 0.f);
  }
  return 0.f;
}

template <int CODEC>
struct NeedsRng { static constexpr bool value = CODEC == GB_CODEC_QSGD || CODEC == GB_CODEC_TERNGRAD || CODEC == GB_CODEC_NATURAL; };

// ---------------------------------------------------------------- pass 2: encode + pack + push
// Each thread owns 4 consecutive elements (one float4); their 4*BITS code bits are merged with
// the neighbouring lanes' into full 32-bit words by a shuffle butterfly, so loads stay 16-byte
// coalesced and each word is stored once per peer.
template <int CODEC, int BITS>
__global__ void __launch_bounds__(GB_THREADS) k2_encode(SegTable t, const float* __restrict__ grad, float* res, float* mom,
                                                        const float* __restrict__ segstats, QuantArgs a_in, WireOut out,
                                                        size_t hdr_bytes, uint32_t* done, PeerPtrs peers, int channel,
                                                        const StepCtl* ctl, size_t par_stride) {
  const QuantArgs a = resolve(a_in, ctl);
  const uint32_t epoch = ctl->epoch;
  const size_t poff = (epoch & 1) ? par_stride : 0;           // parity slot; the parameter struct stays in the constant bank
  constexpr int SUB = 4 * BITS;                         // bits produced per thread
  constexpr int LANES = SUB >= 32 ? 1 : 32 / SUB;       // lanes sharing one word
  constexpr int WORDS = SUB >= 32 ? SUB / 32 : 1;       // words per thread
  __shared__ float s_tab[128];
  // Sub-word codes (1/2/4/8 bit): the tile's packed words are staged in shared memory and go to every peer
  // as 16-byte stores of one contiguous run (512 B .. 4 KB per tile and peer).  Storing each word straight
  // from the lane that finished it produced 16-byte NVLink writes for the 1-bit codecs.
  constexpr bool ODD = BITS == 5 || BITS == 6 || BITS == 7;     // widths that do not divide 32
  constexpr int STAGE_WORDS = WORDS == 1 ? GB_TILE * BITS / 32 + (ODD ? 4 : 0) : 4;
  __shared__ __align__(16) uint32_t s_words[STAGE_WORDS];
  // 5/6/7-bit lanes (QSGD with 8..63 levels): 20/24/28 bits per thread do not tile a word, so the codes are OR-ed into
  // the zeroed staging words at their bit position (a tile starts on a word boundary: 128 elements = 4 BITS words)
  if (ODD) {
    for (int i = threadIdx.x; i < STAGE_WORDS; i += GB_THREADS) s_words[i] = 0u;
    __syncthreads();
  }
  if (CODEC == GB_CODEC_U8BIT) {
    if (threadIdx.x < 128) s_tab[threadIdx.x] = g_u8_table[threadIdx.x];
    __syncthreads();
  }
  for (int tile = blockIdx.x; tile < t.ntiles; tile += gridDim.x) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, tile, seg, base, local0, len);
    const float4 st = reinterpret_cast<const float4*>(segstats)[seg];
    const CodecConst cc = codec_const<CODEC>(st, a.quantum);
    const uint32_t hash = t.seg_hash[seg];
    // full tile (the common case): the gradient (and residual) loads of all four element groups are issued up
    // front; behind the per-group bounds branches only one 16-byte load per thread was in flight (ncu: the 1-bit
    // encode read its 102 MB at 1.6 TB/s with the SM 28 % busy)
    const bool full = len == GB_TILE;
    float4 gq[GB_TILE / (GB_THREADS * 4)], rq[GB_TILE / (GB_THREADS * 4)];
#pragma unroll
    for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j) {
      gq[j] = make_float4(0.f, 0.f, 0.f, 0.f);
      rq[j] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (full) {
        const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
        gq[j] = *reinterpret_cast<const float4*>(grad + base + e0);
        if (a.has_mem) rq[j] = *reinterpret_cast<const float4*>(res + base + e0);
      }
    }
#pragma unroll
    for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j) {
      const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
      // whole warps leave together (len is checked on the warp's first element), so the
      // shuffles below always see full participation
      if ((e0 & ~127) >= len) continue;
      uint32_t code[4] = {0u, 0u, 0u, 0u};
      if (e0 < len) {
        float4 g = gq[j];
        float4 r = rq[j];
        if (!full) {
          g = *reinterpret_cast<const float4*>(grad + base + e0);
          if (a.has_mem) r = *reinterpret_cast<const float4*>(res + base + e0);
        }
        float x[4] = {comp1(a, r.x, g.x), comp1(a, r.y, g.y), comp1(a, r.z, g.z), comp1(a, r.w, g.w)};
        float src[4] = {x[0], x[1], x[2], x[3]};
        if (CODEC == GB_CODEC_SIGNUM) {
          const float4 m = *reinterpret_cast<const float4*>(mom + base + e0);
          const float mm[4] = {m.x, m.y, m.z, m.w};
          const float om = a.om;
#pragma unroll
          for (int q = 0; q < 4; ++q) src[q] = a.mom == 0.f ? x[q] : __fadd_rn(__fmul_rn(om, x[q]), __fmul_rn(a.mom, mm[q]));
          float4 w = make_float4(src[0], src[1], src[2], src[3]);
          if (e0 + 4 > len) {                             // keep pad lanes at zero
            if (e0 + 1 >= len) w.y = 0.f;
            if (e0 + 2 >= len) w.z = 0.f;
            if (e0 + 3 >= len) w.w = 0.f;
          }
          *reinterpret_cast<float4*>(mom + base + e0) = w;
        }
        Philox4 rnd = {0u, 0u, 0u, 0u};
        if (NeedsRng<CODEC>::value) rnd = philox4x32((uint32_t)((local0 + e0) >> 2), a.step, GB_STREAM_UNIFORM, 0u, a.seed, hash);
        const uint32_t rr[4] = {rnd.x, rnd.y, rnd.z, rnd.w};
        float nres[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const bool live = e0 + q < len;
          const uint32_t c = live ? (encode1<CODEC>(src[q], st, rr[q], cc, s_tab) & (0xFFFFFFFFu >> (32 - BITS))) : 0u;
          code[q] = c;
          nres[q] = live ? __fsub_rn(x[q], decode1<CODEC, BITS>(c, st, cc, s_tab)) : 0.f;
        }
        if (a.has_mem) *reinterpret_cast<float4*>(res + base + e0) = make_float4(nres[0], nres[1], nres[2], nres[3]);
      }
      const int64_t flat = base + e0;                   // flat element index of code[0]
      if (WORDS > 1) {
        // 16 or 32 bits per element: this thread owns WORDS whole words
        uint32_t w[WORDS];
        if (BITS == 32) { w[0] = code[0]; w[1 % WORDS] = code[1]; w[2 % WORDS] = code[2]; w[3 % WORDS] = code[3]; }
        else { w[0] = code[0] | (code[1] << 16); w[1 % WORDS] = code[2] | (code[3] << 16); }
        if (e0 < len) {
          for (int r = 0; r < out.n; ++r) {
            uint32_t* dst = reinterpret_cast<uint32_t*>(out.slot[r] + poff + hdr_bytes) + flat * BITS / 32;
            if (WORDS == 4) wire_st_u4(dst, make_uint4(w[0], w[1 % WORDS], w[2 % WORDS], w[3 % WORDS]), out.mc);
            else wire_st_u2(dst, make_uint2(w[0], w[1 % WORDS]), out.mc);
          }
        }
      } else {
        uint32_t sub = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) sub |= code[q] << (q * BITS);
        if (ODD) {
          if (e0 < len && sub) {
            const int bit = e0 * BITS, wi = bit >> 5, sh = bit & 31;
            atomicOr(&s_words[wi], sub << sh);
            if (sh > 32 - SUB) atomicOr(&s_words[wi + 1], sub >> (32 - sh));
          }
        } else {
          const int lane = threadIdx.x & 31;
          uint32_t word = sub << ((lane % LANES) * SUB);
#pragma unroll
          for (int o = 1; o < LANES; o <<= 1) word |= __shfl_xor_sync(0xffffffffu, word, o);
          if ((lane % LANES) == 0 && e0 < len) s_words[e0 * BITS / 32] = word;
        }
      }
    }
    if (WORDS == 1) {
      __syncthreads();
      const int nwords = (len * BITS + 31) / 32;
      const int n4 = nwords >> 2;
      const int64_t w0 = base * BITS / 32;              // base is a multiple of 128 elements -> 16-byte aligned
      for (int r = 0; r < out.n; ++r) {
        uint32_t* dst = reinterpret_cast<uint32_t*>(out.slot[r] + poff + hdr_bytes) + w0;
        for (int i = threadIdx.x; i < n4; i += GB_THREADS)
          wire_st_u4(dst + 4 * i, *reinterpret_cast<const uint4*>(s_words + 4 * i), out.mc);
        if (threadIdx.x < (nwords & 3)) wire_st_u32(dst + 4 * n4 + threadIdx.x, s_words[4 * n4 + threadIdx.x], out.mc);
      }
      __syncthreads();
      if (ODD) {
        for (int i = threadIdx.x; i < STAGE_WORDS; i += GB_THREADS) s_words[i] = 0u;
        __syncthreads();
      }
    }
  }
  release_flags_last_block(done, peers, channel, epoch);
}

// ---------------------------------------------------------------- receive: unpack, decode, add
// final: 0 -> sum / divisor;  1 -> majority vote (sum >= 0 -> +1 else -1);  2 -> sum / lr
template <int CODEC, int BITS>
__global__ void __launch_bounds__(GB_THREADS, 4) k2_decode(SegTable t, WireIn in, size_t hdr_bytes, float* outp, int quantum,
                                                        int final_mode, float divisor, PeerPtrs peers, int channel,
                                                        const StepCtl* ctl, size_t par_stride, uint32_t* status,
                                                        uint64_t timeout_ns) {
  const uint32_t epoch = ctl->epoch;
  const size_t poff = (epoch & 1) ? par_stride : 0;           // parity slot; the parameter struct stays in the constant bank
  __shared__ float s_tab[128];
  __shared__ float4 sts[GB_MAX_WORLD];
  __shared__ CodecConst ccs[GB_MAX_WORLD];
  constexpr bool HAS_STATS = CODEC == GB_CODEC_QSGD || CODEC == GB_CODEC_TERNGRAD || CODEC == GB_CODEC_EFSIGNSGD ||
                             CODEC == GB_CODEC_ONEBIT || CODEC == GB_CODEC_U8BIT;
  if (threadIdx.x < GB_MAX_WORLD) { sts[threadIdx.x] = make_float4(0.f, 0.f, 0.f, 0.f); ccs[threadIdx.x] = CodecConst{0.f, 0.f}; }
  if (CODEC == GB_CODEC_U8BIT && threadIdx.x < 128) s_tab[threadIdx.x] = g_u8_table[threadIdx.x];
  wait_all_ranks(peers, channel, epoch, status, timeout_ns);
  for (int tile = blockIdx.x; tile < t.ntiles; tile += gridDim.x) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, tile, seg, base, local0, len);
    // the senders' per-tensor scalars live in shared memory (8 ranks x 24 bytes in registers halved the
    // occupancy of this bandwidth-bound kernel); codecs without scalars skip the loads altogether
    if (HAS_STATS) {
      __syncthreads();
      if (threadIdx.x < in.n) {
        const float4 s4 = reinterpret_cast<const float4*>(in.slot[threadIdx.x] + poff)[seg];
        sts[threadIdx.x] = s4;
        ccs[threadIdx.x] = codec_const<CODEC>(s4, quantum);
      }
      __syncthreads();
    }
    const float inv_div = divisor != 1.f ? __fdiv_rn(1.0f, divisor) : 1.f;
    // Loads first, arithmetic second: every rank's words of JCH element groups are requested before any of them
    // is decoded (rank indices beyond the world size re-read rank 0 and are ignored).  The earlier loop left the
    // rank loop through a `break`, which serialised W x 4 four-byte loads per thread and tile and capped the
    // kernel at ~1 TB/s at W = 8.
    constexpr int NJ = GB_TILE / (GB_THREADS * 4);
    constexpr bool ODD = BITS == 5 || BITS == 6 || BITS == 7;                       // widths that do not divide 32
    constexpr int WPT = BITS == 32 ? 4 : ((BITS == 16 || ODD) ? 2 : 1);            // 32-bit words per thread, rank and group
    constexpr int JCH = BITS == 32 ? 1 : ((BITS == 16 || ODD) ? 2 : 4);            // groups whose loads are in flight together
#pragma unroll
    for (int j0 = 0; j0 < NJ; j0 += JCH) {
      uint32_t raw[JCH][GB_MAX_WORLD][WPT];
#pragma unroll
      for (int jj = 0; jj < JCH; ++jj) {
        const int e0 = ((j0 + jj) * GB_THREADS + threadIdx.x) * 4;
        const int64_t flat = base + (e0 < len ? e0 : 0);
#pragma unroll
        for (int r = 0; r < GB_MAX_WORLD; ++r) {
          const uint32_t* words = reinterpret_cast<const uint32_t*>(in.slot[r < in.n ? r : 0] + poff + hdr_bytes);
          if (BITS == 32) {
            const uint4 w = *reinterpret_cast<const uint4*>(words + flat);
            raw[jj][r][0] = w.x; raw[jj][r][1 % WPT] = w.y; raw[jj][r][2 % WPT] = w.z; raw[jj][r][3 % WPT] = w.w;
          } else if (BITS == 16) {
            const uint2 w = *reinterpret_cast<const uint2*>(words + flat / 2);
            raw[jj][r][0] = w.x; raw[jj][r][1 % WPT] = w.y;
          } else if (ODD) {
            raw[jj][r][0] = words[(flat * BITS) >> 5];                   // 4 BITS bits from bit (BITS flat) & 31: may straddle
            raw[jj][r][1 % WPT] = words[((flat * BITS) >> 5) + 1];
          } else {
            raw[jj][r][0] = words[flat * BITS / 32];
          }
        }
      }
#pragma unroll
      for (int jj = 0; jj < JCH; ++jj) {
        const int e0 = ((j0 + jj) * GB_THREADS + threadIdx.x) * 4;
        const int64_t flat = base + (e0 < len ? e0 : 0);
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int r = 0; r < GB_MAX_WORLD; ++r) {
          uint32_t code[4];
          if (BITS == 32) {
            code[0] = raw[jj][r][0]; code[1] = raw[jj][r][1 % WPT]; code[2] = raw[jj][r][2 % WPT]; code[3] = raw[jj][r][3 % WPT];
          } else if (BITS == 16) {
            code[0] = raw[jj][r][0] & 0xFFFFu; code[1] = raw[jj][r][0] >> 16;
            code[2] = raw[jj][r][1 % WPT] & 0xFFFFu; code[3] = raw[jj][r][1 % WPT] >> 16;
          } else if (ODD) {
            const uint64_t v = ((uint64_t)raw[jj][r][1 % WPT] << 32) | raw[jj][r][0];
            const int sh = (int)((flat * BITS) & 31);
#pragma unroll
            for (int q = 0; q < 4; ++q) code[q] = (uint32_t)(v >> (sh + q * BITS)) & (0xFFFFFFFFu >> (32 - BITS));
          } else {
            const int sh = (int)((flat * BITS) & 31);
#pragma unroll
            for (int q = 0; q < 4; ++q) code[q] = (raw[jj][r][0] >> (sh + q * BITS)) & (0xFFFFFFFFu >> (32 - BITS));
          }
          if (r < in.n) {
#pragma unroll
            for (int q = 0; q < 4; ++q) acc[q] += decode1<CODEC, BITS>(code[q], sts[r], ccs[r], s_tab);
          }
        }
        if (e0 < len) {
          float o[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            if (final_mode == 1) o[q] = acc[q] >= 0.f ? 1.f : -1.f;
            else o[q] = divisor != 1.f ? __fmul_rn(acc[q], inv_div) : acc[q];
            if (e0 + q >= len) o[q] = 0.f;
          }
          *reinterpret_cast<float4*>(outp + flat) = make_float4(o[0], o[1], o[2], o[3]);
        }
      }
    }
  }
}

// ---------------------------------------------------------------- receive for SignSGD / Signum: majority vote
// sum_r (2 b_r - 1) >= 0  <=>  popcount over ranks >= ceil(W / 2)  (tie -> +1, signsgd.py:25-30).  The W bitmaps
// of a tile are added bit-sliced (a 4-bit carry-save counter per element position, ~7 bitwise ops per rank
// for 32 elements) instead of decoding W x 32 floats; the vote word is then expanded to +-1 with coalesced
// 16-byte stores.
__global__ void __launch_bounds__(GB_THREADS) k2_vote(SegTable t, WireIn in, size_t hdr_bytes, float* outp, PeerPtrs peers, int channel,
                                                      const StepCtl* ctl, size_t par_stride, uint32_t* status, uint64_t timeout_ns) {
  __shared__ uint32_t s_vote[GB_TILE / 32];
  const uint32_t epoch = ctl->epoch;
  const size_t poff = (epoch & 1) ? par_stride : 0;           // parity slot; the parameter struct stays in the constant bank
  wait_all_ranks(peers, channel, epoch, status, timeout_ns);
  const uint32_t need = (uint32_t)(in.n + 1) / 2;
  for (int tile = blockIdx.x; tile < t.ntiles; tile += gridDim.x) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, tile, seg, base, local0, len);
    if (threadIdx.x < GB_TILE / 32 && threadIdx.x * 32 < len) {
      uint32_t c0 = 0, c1 = 0, c2 = 0, c3 = 0;
#pragma unroll
      for (int r = 0; r < GB_MAX_WORLD; ++r) {      // no early exit: the W loads stay independent
        uint32_t w = reinterpret_cast<const uint32_t*>(in.slot[r < in.n ? r : 0] + poff + hdr_bytes)[base / 32 + threadIdx.x];
        if (r >= in.n) w = 0u;
        const uint32_t k0 = c0 & w; c0 ^= w;
        const uint32_t k1 = c1 & k0; c1 ^= k0;
        const uint32_t k2 = c2 & k1; c2 ^= k1;
        c3 ^= k2;
      }
      // count >= need, bit-sliced compare from the most significant plane down
      const uint32_t cs[4] = {c0, c1, c2, c3};
      uint32_t gt = 0u, eq = 0xFFFFFFFFu;
#pragma unroll
      for (int b = 3; b >= 0; --b) {
        const uint32_t tb = (need >> b) & 1u ? 0xFFFFFFFFu : 0u;
        gt |= eq & cs[b] & ~tb;
        eq &= ~(cs[b] ^ tb);
      }
      s_vote[threadIdx.x] = gt | eq;
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < GB_TILE / (GB_THREADS * 4); ++j) {
      const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
      if (e0 >= len) continue;
      const uint32_t w = s_vote[e0 >> 5] >> (e0 & 31);
      float o[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) o[q] = (e0 + q < len) ? (((w >> q) & 1u) ? 1.f : -1.f) : 0.f;
      *reinterpret_cast<float4*>(outp + base + e0) = make_float4(o[0], o[1], o[2], o[3]);
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------- host side
#define LAUNCHED(p) do { (p)->launches++; GB_CUDA(cudaGetLastError()); } while (0)

static int grid_for(int ntiles) { return ntiles < 148 * 8 ? ntiles : 148 * 8; }

static bool needs_stats(int codec) {
  return codec == GB_CODEC_QSGD || codec == GB_CODEC_TERNGRAD || codec == GB_CODEC_EFSIGNSGD || codec == GB_CODEC_ONEBIT ||
         codec == GB_CODEC_U8BIT;
}

template <int CODEC, int BITS>
static int launch_encode(gb_plan* p, const float* g, const QuantArgs& a, const WireOut& out, const PeerPtrs& peers,
                         cudaStream_t st) {
  k2_encode<CODEC, BITS><<<grid_for(p->ntiles), GB_THREADS, 0, st>>>(p->tab, g, p->d_state0, p->d_state1, p->d_segstats, a, out,
                                                                    p->hdr_bytes, p->d_done, peers, p->channel, p->d_ctl,
                                                                    p->par_stride);
  LAUNCHED(p);
  return 0;
}

template <int CODEC, int BITS>
static int launch_decode(gb_plan* p, float* o, const WireIn& in, const PeerPtrs& peers, int mode, float divisor, cudaStream_t st) {
  k2_decode<CODEC, BITS><<<grid_for(p->ntiles), GB_THREADS, 0, st>>>(p->tab, in, p->hdr_bytes, o, p->prm.quantum_num, mode, divisor,
                                                                    peers, p->channel, p->d_ctl, p->par_stride,
                                                                    p->comm->status_dev, p->comm->timeout_ns);
  LAUNCHED(p);
  return 0;
}

#define DISPATCH(FN, ...)                                                                         \
  switch (p->prm.codec) {                                                                         \
    case GB_CODEC_NONE: return FN<GB_CODEC_NONE, 32>(__VA_ARGS__);                                \
    case GB_CODEC_FP16: return FN<GB_CODEC_FP16, 16>(__VA_ARGS__);                                \
    case GB_CODEC_QSGD:                                                                           \
      if (p->bits == 4) return FN<GB_CODEC_QSGD, 4>(__VA_ARGS__);                                 \
      if (p->bits == 5) return FN<GB_CODEC_QSGD, 5>(__VA_ARGS__);                                 \
      if (p->bits == 6) return FN<GB_CODEC_QSGD, 6>(__VA_ARGS__);                                 \
      if (p->bits == 7) return FN<GB_CODEC_QSGD, 7>(__VA_ARGS__);                                 \
      if (p->bits == 8) return FN<GB_CODEC_QSGD, 8>(__VA_ARGS__);                                 \
      return FN<GB_CODEC_QSGD, 16>(__VA_ARGS__);                                                  \
    case GB_CODEC_TERNGRAD: return FN<GB_CODEC_TERNGRAD, 2>(__VA_ARGS__);                         \
    case GB_CODEC_NATURAL: return FN<GB_CODEC_NATURAL, 8>(__VA_ARGS__);                           \
    case GB_CODEC_U8BIT: return FN<GB_CODEC_U8BIT, 8>(__VA_ARGS__);                               \
    case GB_CODEC_SIGNSGD: return FN<GB_CODEC_SIGNSGD, 1>(__VA_ARGS__);                           \
    case GB_CODEC_SIGNUM: return FN<GB_CODEC_SIGNUM, 1>(__VA_ARGS__);                             \
    case GB_CODEC_EFSIGNSGD: return FN<GB_CODEC_EFSIGNSGD, 1>(__VA_ARGS__);                       \
    case GB_CODEC_ONEBIT: return FN<GB_CODEC_ONEBIT, 1>(__VA_ARGS__);                             \
    default: gb_set_error("codec %d is not a quantiser", p->prm.codec); return -1;                \
  }

int k2_send(gb_plan* p, const float* d_grad, cudaStream_t st) {
  gb_comm* c = p->comm;
  GB_CHECK(((uintptr_t)d_grad & 15) == 0, "gradient buffer must be 16-byte aligned");
  QuantArgs a;
  a.codec = p->prm.codec;
  a.has_mem = p->prm.memory == GB_MEM_RESIDUAL;
  a.beta = p->prm.beta;                 // steady-state values; kernels apply the first-call rules
  a.gamma = p->prm.gamma;
  a.mom = p->prm.momentum;
  a.om = p->prm.one_minus_momentum != 0.f ? p->prm.one_minus_momentum : 1.0f - p->prm.momentum;
  a.quantum = p->prm.quantum_num;
  a.seed = p->prm.seed;
  a.step = 0;
  ++p->epoch;                           // host mirror of the device-side epoch
  WireOut out;
  out.n = c->world;
  out.mc = 0;
  for (int r = 0; r < c->world; ++r) out.slot[r] = gb_slot_of(c, r, p->chan_off, p->slot_bytes, 0, c->rank);
  if (char* mc = gb_slot_mc(c, p->chan_off, p->slot_bytes, 0, c->rank)) {   // one multimem.st instead of W peer stores
    out.n = 1; out.mc = 1; out.slot[0] = mc;
  }
  const PeerPtrs peers = gb_comm_peers(c);
  if (needs_stats(a.codec)) {
    const int g = grid_for(p->ntiles);
    switch (a.codec) {
      case GB_CODEC_QSGD: k2_stats<GB_CODEC_QSGD><<<g, GB_THREADS, 0, st>>>(p->tab, d_grad, p->d_state0, a, p->d_ctl, p->d_tilestats); break;
      case GB_CODEC_TERNGRAD: k2_stats<GB_CODEC_TERNGRAD><<<g, GB_THREADS, 0, st>>>(p->tab, d_grad, p->d_state0, a, p->d_ctl, p->d_tilestats); break;
      case GB_CODEC_EFSIGNSGD: k2_stats<GB_CODEC_EFSIGNSGD><<<g, GB_THREADS, 0, st>>>(p->tab, d_grad, p->d_state0, a, p->d_ctl, p->d_tilestats); break;
      case GB_CODEC_ONEBIT: k2_stats<GB_CODEC_ONEBIT><<<g, GB_THREADS, 0, st>>>(p->tab, d_grad, p->d_state0, a, p->d_ctl, p->d_tilestats); break;
      default: k2_stats<GB_CODEC_U8BIT><<<g, GB_THREADS, 0, st>>>(p->tab, d_grad, p->d_state0, a, p->d_ctl, p->d_tilestats); break;
    }
    LAUNCHED(p);
    k2_finalize<<<p->nseg, 32, 0, st>>>(p->tab, p->d_tilestats, a.codec, p->d_segstats, out, p->d_ctl, p->par_stride); LAUNCHED(p);
  }
  DISPATCH(launch_encode, p, d_grad, a, out, peers, st);
}

int k2_recv(gb_plan* p, float* d_out, cudaStream_t st) {
  gb_comm* c = p->comm;
  GB_CHECK(((uintptr_t)d_out & 15) == 0, "output buffer must be 16-byte aligned");
  WireIn in;
  in.n = c->world;
  for (int r = 0; r < c->world; ++r) in.slot[r] = gb_slot_of(c, c->rank, p->chan_off, p->slot_bytes, 0, r);
  const PeerPtrs peers = gb_comm_peers(c);
  int mode = 0;
  float divisor = p->prm.average ? (float)c->world : 1.f;
  if (p->prm.codec == GB_CODEC_SIGNSGD || p->prm.codec == GB_CODEC_SIGNUM) {
    k2_vote<<<grid_for(p->ntiles), GB_THREADS, 0, st>>>(p->tab, in, p->hdr_bytes, d_out, peers, p->channel, p->d_ctl, p->par_stride,
                                                        c->status_dev, c->timeout_ns);
    LAUNCHED(p);
    return 0;
  }
  if (p->prm.codec == GB_CODEC_EFSIGNSGD) { mode = 2; divisor = p->prm.lr; }
  DISPATCH(launch_decode, p, d_out, in, peers, mode, divisor, st);
}
