// K2-inceptionn -- INCEPTIONN variable-width codec.
//
// Reference: grace_dl/tensorflow/compressor/inceptionn.py:8-188 (TF graph: per element a 32/16/8/0-bit
// encoding chosen by the binary exponent, four separate boolean_mask gathers, 2-bit tags packed by a chain of
// reshape/bitwise ops, decode through find_bins and scatter_nd into freshly created Variables).
// Oracle of the numerics: grace_b200/compressor/inceptionn.py.
//
// One pass per direction, no reductions, no host round trips:
//   in_send   compensate (ResidualMemory) -> classify by exponent -> tag byte per 4 elements stored straight
//             into every peer -> block-wide ordered scan of the three class counts (one 64-bit packed word) ->
//             the tile's values are staged in shared memory as [fp32 ...][16-bit ...][8-bit ...], a byte range
//             of the stream is claimed with one atomic, the staged words go to every peer with coalesced
//             stores; residual = x - decode(x) in place; the last CTA releases the flags.
//   in_recv   each CTA owns an output tile: re-derives the per-element offsets from the sender's tags with the
//             same scan, decodes from the sender's stream, adds the W ranks in rank order in registers.
// Wire per rank: 2 bits per element + 4/2/1/0 bytes per element + 16 bytes per tile.
// quantum_num carries the two exponent cut points: low_exp | mid_exp << 8.
#include "kernels.cuh"

#include <algorithm>

struct IncOut {
  int4* hdr[GB_MAX_WORLD];
  uint8_t* tags[GB_MAX_WORLD];
  uint32_t* stream[GB_MAX_WORLD];
  int n;
};

// per tile: int4 {first stream word, #32-bit, #16-bit, #8-bit} + the exclusive class offsets of its 32 warp chunks
// (128 elements each, packed like class_word) -- the receiver needs no block-wide scan
#define IN_HDR_STRIDE 272
size_t in_hdr_bytes(int ntiles) { return ((size_t)ntiles * IN_HDR_STRIDE + 255) & ~size_t(255); }
size_t in_tag_bytes(int64_t padded) { return ((size_t)padded / 4 + 255) & ~size_t(255); }

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

__device__ __forceinline__ int classify(float x, int low_exp, int mid_exp) {
  const int e = (int)((__float_as_uint(x) >> 23) & 0xFFu);
  if (e >= 127) return 3;
  if (e >= mid_exp) return 2;
  if (e >= low_exp) return 1;            // low_exp <= e < mid_exp
  return 0;
}
// class counts of one thread packed as 16-bit fields: [15:0] 32-bit values, [31:16] 16-bit, [47:32] 8-bit
__device__ __forceinline__ uint64_t class_word(int cls) { return cls == 3 ? 1ull : cls == 2 ? (1ull << 16) : cls == 1 ? (1ull << 32) : 0ull; }

__device__ __forceinline__ uint64_t warp_incl_scan64(uint64_t v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint64_t n = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += n;
  }
  return v;
}

// Exclusive offsets of this thread's four 4-element groups in (j, thread) order; returns the tile totals.
__device__ __forceinline__ uint64_t ordered_offsets(const uint64_t pc[4], uint64_t offs[4], uint64_t (*wtot)[GB_THREADS / 32],
                                                    uint64_t* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint64_t incl[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    incl[j] = warp_incl_scan64(pc[j], lane);
    if (lane == 31) wtot[j][warp] = incl[j];
  }
  __syncthreads();
  if (warp == 0) {                                     // the 4 x 8 warp totals are exactly one warp's worth of lanes
    uint64_t* flat = &wtot[0][0];
    const uint64_t c = flat[lane];
    const uint64_t inc = warp_incl_scan64(c, lane);
    flat[lane] = inc - c;
    if (lane == 31) *total = inc;
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 4; ++j) offs[j] = wtot[j][warp] + incl[j] - pc[j];
  return *total;
}

__device__ __forceinline__ float dec16(uint32_t w) {
  const float m = __fmul_rn((float)(w & 0x7FFFu), 1.0f / 16384.0f);
  return (w >> 15) & 1u ? -m : m;
}
__device__ __forceinline__ float dec8(uint32_t w) {
  const float m = __fmul_rn((float)(w & 0x7Fu), 1.0f / 64.0f);
  return (w >> 7) & 1u ? -m : m;
}

template <bool WRITE>
__global__ void __launch_bounds__(GB_THREADS) in_send_k(SegTable t, const float* __restrict__ grad, float* work, float beta, float gamma,
                                                        int low_exp, int mid_exp, uint32_t* cursor, IncOut out, int64_t cap_words,
                                                        uint32_t* status, uint32_t* done, PeerPtrs peers, int channel,
                                                        const StepCtl* ctl, size_t par_stride) {
  __shared__ uint32_t stage[GB_TILE];
  __shared__ uint64_t wtot[4][GB_THREADS / 32];
  __shared__ uint64_t s_total;
  __shared__ int s_begin;
  const uint32_t epoch = ctl->epoch;
  if (ctl->step == 0) { beta = 0.f; gamma = 1.f; }
  if (epoch & 1) {
    for (int r = 0; r < out.n; ++r) {
      out.hdr[r] = (int4*)((char*)out.hdr[r] + par_stride);
      out.tags[r] = out.tags[r] + par_stride;
      out.stream[r] = (uint32_t*)((char*)out.stream[r] + par_stride);
    }
  }
  for (int tile = blockIdx.x; tile < t.ntiles; tile += gridDim.x) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, tile, seg, base, local0, len);
    float vv[4][4];
    uint64_t pc[4], offs[4];
    uint32_t tagw = 0;                                  // this thread's 4 tag bytes (one per j)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (e0 < len) {
        if (WRITE) {
          const float4 g = ld4(grad + base, e0, len);
          const float4 r = ld4(work + base, e0, len);
          v.x = __fadd_rn(__fmul_rn(beta, r.x), __fmul_rn(gamma, g.x)); v.y = __fadd_rn(__fmul_rn(beta, r.y), __fmul_rn(gamma, g.y));
          v.z = __fadd_rn(__fmul_rn(beta, r.z), __fmul_rn(gamma, g.z)); v.w = __fadd_rn(__fmul_rn(beta, r.w), __fmul_rn(gamma, g.w));
        } else {
          v = ld4(work + base, e0, len);
        }
      }
      vv[j][0] = v.x; vv[j][1] = v.y; vv[j][2] = v.z; vv[j][3] = v.w;
      pc[j] = 0;
      uint32_t tb = 0;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int cls = (e0 + q < len) ? classify(vv[j][q], low_exp, mid_exp) : 0;
        tb |= (uint32_t)cls << (2 * q);
        pc[j] += class_word(cls);
      }
      tagw |= tb << (8 * j);
      if (e0 < len)
        for (int r = 0; r < out.n; ++r) out.tags[r][(base + e0) >> 2] = (uint8_t)tb;
    }
    const uint64_t total = ordered_offsets(pc, offs, wtot, &s_total);
    if (threadIdx.x < 32) {                             // exclusive offsets of the tile's 32 warp chunks -> header
      const uint64_t c = (&wtot[0][0])[threadIdx.x];
      for (int r = 0; r < out.n; ++r)
        reinterpret_cast<uint64_t*>(reinterpret_cast<char*>(out.hdr[r]) + (size_t)tile * IN_HDR_STRIDE + 16)[threadIdx.x] = c;
    }
    const int n32 = (int)(total & 0xFFFFu), n16 = (int)((total >> 16) & 0xFFFFu), n8 = (int)((total >> 32) & 0xFFFFu);
    uint16_t* st16 = reinterpret_cast<uint16_t*>(stage + n32);
    uint8_t* st8 = reinterpret_cast<uint8_t*>(stage + n32) + 2 * n16;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
      int o32 = (int)(offs[j] & 0xFFFFu), o16 = (int)((offs[j] >> 16) & 0xFFFFu), o8 = (int)((offs[j] >> 32) & 0xFFFFu);
      if (((tagw >> (8 * j)) & 0xFFu) == 0) continue;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int cls = (tagw >> (8 * j + 2 * q)) & 3u;
        const float x = vv[j][q];
        const uint32_t neg = x < 0.f ? 1u : 0u;
        if (cls == 3) {
          stage[o32++] = __float_as_uint(x);
          vv[j][q] = 0.f;
        } else if (cls == 2) {
          const uint32_t w = (uint32_t)floorf(__fmul_rn(fabsf(x), 16384.0f)) | (neg << 15);
          st16[o16++] = (uint16_t)w;
          vv[j][q] = __fsub_rn(x, dec16(w));
        } else if (cls == 1) {
          const uint32_t w = (uint32_t)floorf(__fmul_rn(fabsf(x), 64.0f)) | (neg << 7);
          st8[o8++] = (uint8_t)w;
          vv[j][q] = __fsub_rn(x, dec8(w));
        }
      }
      if (WRITE) st4(work + base, e0, len, make_float4(vv[j][0], vv[j][1], vv[j][2], vv[j][3]));
    }
    if (WRITE) {
      // groups without any coded element still need the compensated value written back
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
        if (e0 < len && ((tagw >> (8 * j)) & 0xFFu) == 0) st4(work + base, e0, len, make_float4(vv[j][0], vv[j][1], vv[j][2], vv[j][3]));
      }
    }
    const int words = n32 + (2 * n16 + n8 + 3) / 4;
    __syncthreads();
    if (threadIdx.x == 0) {
      // A tile never needs more stream words than it has elements, so with a slot sized for the whole bucket (the
      // default capacity of 1.0) the tile's own element offset is a free place for its words: no claim at all.  The
      // shared cursor compacts the stream only when the slot is smaller -- one same-address atomic per tile, which
      // serialises in L2 (~100 us for the 25 000 tiles of ResNet-50).
      int64_t begin = 0;
      if (cap_words >= t.padded) begin = base;
      else if (words) begin = (int64_t)atomicAdd(cursor, (uint32_t)words);
      int4 h = make_int4((int)begin, n32, n16, n8);
      if (begin + words > cap_words) {
        atomicOr(status, GB_STATUS_OVERFLOW);
        h = make_int4(-1, 0, 0, 0);                     // receivers skip this tile; the host raises on the status word
      }
      s_begin = h.x;
      for (int r = 0; r < out.n; ++r) *reinterpret_cast<int4*>(reinterpret_cast<char*>(out.hdr[r]) + (size_t)tile * IN_HDR_STRIDE) = h;
    }
    __syncthreads();
    const int begin = s_begin;
    if (begin >= 0) {
      for (int r = 0; r < out.n; ++r) {
        uint32_t* d = out.stream[r] + begin;
        for (int i = threadIdx.x; i < words; i += GB_THREADS) d[i] = stage[i];
      }
    }
    __syncthreads();
  }
  release_flags_last_block(done, peers, channel, epoch);
}

struct IncIn {
  const int4* hdr[GB_MAX_WORLD];
  const uint8_t* tags[GB_MAX_WORLD];
  const uint32_t* stream[GB_MAX_WORLD];
  int n;
};

__global__ void __launch_bounds__(GB_THREADS) in_recv_k(SegTable t, IncIn in, float* outp, float divisor, PeerPtrs peers, int channel,
                                                        const StepCtl* ctl, size_t par_stride, uint32_t* status, uint64_t timeout_ns) {
  const uint32_t epoch = ctl->epoch;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (epoch & 1) {
    for (int r = 0; r < in.n; ++r) {
      in.hdr[r] = (const int4*)((const char*)in.hdr[r] + par_stride);
      in.tags[r] = in.tags[r] + par_stride;
      in.stream[r] = (const uint32_t*)((const char*)in.stream[r] + par_stride);
    }
  }
  wait_all_ranks(peers, channel, epoch, status, timeout_ns);
  for (int tile = blockIdx.x; tile < t.ntiles; tile += gridDim.x) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, tile, seg, base, local0, len);
    float acc[4][4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int q = 0; q < 4; ++q) acc[j][q] = 0.f;
    // No block-wide synchronisation in this kernel: the sender's header carries the class offsets of every
    // 128-element warp chunk, a warp scan of the tags gives the rest.  Headers and tag bytes of every rank are
    // requested up front (independent loads); only the value fetches depend on them.
    int4 hh[GB_MAX_WORLD];
    uint32_t tg[GB_MAX_WORLD];
#pragma unroll
    for (int r = 0; r < GB_MAX_WORLD; ++r) {
      const int rr = r < in.n ? r : 0;
      hh[r] = *reinterpret_cast<const int4*>(reinterpret_cast<const char*>(in.hdr[rr]) + (size_t)tile * IN_HDR_STRIDE);
      tg[r] = 0u;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
        const uint32_t tb = e0 < len ? in.tags[rr][(base + e0) >> 2] : 0u;
        tg[r] |= tb << (8 * j);
      }
    }
#pragma unroll
    for (int r = 0; r < GB_MAX_WORLD; ++r) {
      if (r >= in.n) continue;
      const int4 h = hh[r];
      if (h.x < 0) continue;                            // sender overflowed its slot (status already raised)
      int o32s[4], o16s[4], o8s[4];
      const uint32_t tagw = tg[r];
      const uint64_t* chunk = reinterpret_cast<const uint64_t*>(reinterpret_cast<const char*>(in.hdr[r]) + (size_t)tile * IN_HDR_STRIDE + 16);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint32_t tb = (tagw >> (8 * j)) & 0xFFu;
        // inside a 128-element chunk every class count fits 8 bits: one 32-bit warp scan for all three
        uint32_t pc = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint32_t cls = (tb >> (2 * q)) & 3u;
          pc += cls == 3u ? 1u : (cls == 2u ? (1u << 8) : (cls == 1u ? (1u << 16) : 0u));
        }
        uint32_t inc = pc;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t nb = __shfl_up_sync(0xffffffffu, inc, o);
          if (lane >= o) inc += nb;
        }
        const uint32_t ex = inc - pc;
        const uint64_t cb = chunk[j * (GB_THREADS / 32) + warp];
        o32s[j] = (int)(cb & 0xFFFFu) + (int)(ex & 0xFFu);
        o16s[j] = (int)((cb >> 16) & 0xFFFFu) + (int)((ex >> 8) & 0xFFu);
        o8s[j] = (int)((cb >> 32) & 0xFFFFu) + (int)((ex >> 16) & 0xFFu);
      }
      // Branch-free fetch: every coded element reads the aligned 32-bit word that holds its 4/2/1-byte value
      // (predicated, all 16 loads independent), the class only selects the extraction.  Loads inside the
      // three-way class branch serialised up to 3 x 16 memory latencies per rank.
      const uint32_t* sp = in.stream[r] + h.x;
      const int b16 = 4 * h.y, b8 = 4 * h.y + 2 * h.z;           // byte offsets of the 16-bit and 8-bit runs
      uint32_t wv[4][4];
      uint32_t shf[4];                                             // 4 x 4 shift amounts (bits), 8 bits each
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        int o32 = o32s[j], o16 = o16s[j], o8 = o8s[j];
        shf[j] = 0u;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int cls = (tagw >> (8 * j + 2 * q)) & 3u;
          const bool c3 = cls == 3, c2 = cls == 2, c1 = cls == 1;
          const int bo = c3 ? 4 * o32 : (c2 ? b16 + 2 * o16 : b8 + o8);
          o32 += c3; o16 += c2; o8 += c1;
          wv[j][q] = cls ? sp[bo >> 2] : 0u;
          shf[j] |= (uint32_t)((bo & 3) * 8) << (8 * q);
        }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int cls = (tagw >> (8 * j + 2 * q)) & 3u;
          const uint32_t w = wv[j][q] >> ((shf[j] >> (8 * q)) & 0xFFu);
          // 16- and 8-bit forms share one path: sign + magnitude / 2^14 (or 2^6); exact in fp32
          const bool wide = cls == 2;
          const uint32_t mag = w & (wide ? 0x7FFFu : 0x7Fu);
          const uint32_t sgn = (w >> (wide ? 15 : 7)) & 1u;
          const float m = __fmul_rn((float)mag, wide ? (1.0f / 16384.0f) : (1.0f / 64.0f));
          const float x = cls == 3 ? __uint_as_float(wv[j][q]) : __uint_as_float(__float_as_uint(m) | (sgn << 31));
          if (cls) acc[j][q] += x;
        }
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
      if (e0 >= len) continue;
      float4 v = make_float4(acc[j][0], acc[j][1], acc[j][2], acc[j][3]);
      if (divisor != 1.f) {          // torch's CUDA tensor / python-scalar is a multiply by the reciprocal
        const float inv = __fdiv_rn(1.0f, divisor);
        v.x = __fmul_rn(v.x, inv); v.y = __fmul_rn(v.y, inv); v.z = __fmul_rn(v.z, inv); v.w = __fmul_rn(v.w, inv);
      }
      st4(outp + base, e0, len, v);
    }
  }
}

}  // namespace

// ---------------------------------------------------------------- host side
#define IN_LAUNCHED(p) do { (p)->launches++; GB_CUDA(cudaGetLastError()); } while (0)

static int in_grid(int ntiles) { return ntiles < 148 * 8 ? ntiles : 148 * 8; }

int in_send(gb_plan* p, const float* d_grad, cudaStream_t st) {
  gb_comm* c = p->comm;
  GB_CHECK(p->prm.memory == GB_MEM_NONE || p->prm.memory == GB_MEM_RESIDUAL, "unsupported memory for INCEPTIONN");
  const bool mem = p->prm.memory == GB_MEM_RESIDUAL;
  float* work = mem ? p->d_state0 : const_cast<float*>(d_grad);
  ++p->epoch;
  IncOut out;
  const size_t tb = in_tag_bytes(p->padded);
  for (int r = 0; r < c->world; ++r) {
    char* s = gb_slot_of(c, r, p->chan_off, p->slot_bytes, 0, c->rank);
    out.hdr[r] = (int4*)s; out.tags[r] = (uint8_t*)(s + p->hdr_bytes); out.stream[r] = (uint32_t*)(s + p->hdr_bytes + tb);
  }
  out.n = c->world;
  const int low_exp = p->prm.quantum_num & 0xFF, mid_exp = (p->prm.quantum_num >> 8) & 0x1FF;
  uint32_t* cursor = p->d_done + 2;
  GB_CUDA(cudaMemsetAsync(cursor, 0, 4, st));
  const PeerPtrs peers = gb_comm_peers(c);
  if (mem) in_send_k<true><<<in_grid(p->ntiles), GB_THREADS, 0, st>>>(p->tab, d_grad, work, p->prm.beta, p->prm.gamma, low_exp, mid_exp, cursor, out, p->ktotal, c->status_dev, p->d_done, peers, p->channel, p->d_ctl, p->par_stride);
  else in_send_k<false><<<in_grid(p->ntiles), GB_THREADS, 0, st>>>(p->tab, d_grad, work, 0.f, 1.f, low_exp, mid_exp, cursor, out, p->ktotal, c->status_dev, p->d_done, peers, p->channel, p->d_ctl, p->par_stride);
  IN_LAUNCHED(p);
  return 0;
}

int in_recv(gb_plan* p, float* d_out, cudaStream_t st) {
  gb_comm* c = p->comm;
  IncIn in;
  const size_t tb = in_tag_bytes(p->padded);
  for (int r = 0; r < c->world; ++r) {
    const char* s = gb_slot_of(c, c->rank, p->chan_off, p->slot_bytes, 0, r);
    in.hdr[r] = (const int4*)s; in.tags[r] = (const uint8_t*)(s + p->hdr_bytes); in.stream[r] = (const uint32_t*)(s + p->hdr_bytes + tb);
  }
  in.n = c->world;
  const float divisor = p->prm.average ? (float)c->world : 1.f;
  in_recv_k<<<in_grid(p->ntiles), GB_THREADS, 0, st>>>(p->tab, in, d_out, divisor, gb_comm_peers(c), p->channel, p->d_ctl, p->par_stride,
                                                      c->status_dev, c->timeout_ns);
  IN_LAUNCHED(p);
  return 0;
}
