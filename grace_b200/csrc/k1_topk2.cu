// K1-topk, two streaming passes.
//
// Replaces the reference chain  memory/residual.py:10-20 (compensate / update) + compressor/topk.py:6-18
// (abs, topk, gather, zeros + scatter_) + communicator/allgather.py:8-45 per tensor by five launches per
// BUCKET (all tensors of the bucket, per-tensor k and residual preserved):
//
//   tk_pass1   reads gradient + residual once, writes the compensated values, 11-bit histogram of |x| per tensor;
//              the CTA that completes a tensor's last tile picks the bin that holds the k-th largest magnitude
//              (no separate select launch); the last CTA of the grid advances the device-side step block.
//   tk_pass2   reads the compensated values once: everything in a HIGHER bin is certain to be selected -> staged in
//              shared memory and stored straight into every peer's payload slot, zeroed in place (new residual);
//              elements IN the threshold bin (~1 % of the data) are only marked in a 4096-bit mask per tile and
//              histogrammed by their next 10 bits; the CTA finishing a tensor extends the threshold prefix to 21 bits.
//   tk_refine  one warp per tile walks the mask, histograms the last 10 bits of the few values matching the 21-bit
//              prefix; the warp finishing a tensor fixes the exact 31-bit threshold and how many equal values to take.
//   tk_ties    only for tensors where more values EQUAL the threshold than may be taken (common with bf16-valued
//              gradients): per-tile tie counts + exclusive scan, so ties are taken in index order (lowest first).
//   tk_finish  one warp per tile: the chosen threshold-bin elements go to the peers as a second (value, index) range
//              of the tile, are zeroed in place; the last CTA releases the epoch flags.
//
// The big data is therefore streamed twice (pass 1: 2 reads + 1 write, pass 2: 1 read) instead of four times; the
// three tail kernels touch the masks (0.125 B / element) and gather ~1 % of the values from L2.
#include "kernels.cuh"

#define TK_SEL 8
enum { S_BIN = 0, S_KREM = 1, S_CURA = 2, S_FLAGS = 3, S_THR = 4, S_TAKE = 5 };
#define TKF_TAKEALL 1u   // fewer non-zero candidates than still needed: every candidate is selected
#define TKF_AMBIG 2u     // more values equal to the threshold than may be taken: ties go by index order
#define TK_WPT 8         // warps (tiles) per CTA in the tail kernels

struct TkDev {
  uint32_t* hist0;    // [nseg][2048]
  uint32_t* hist1;    // [nseg][1024]
  uint32_t* hist2;    // [nseg][1024]
  uint32_t* sel;      // [nseg][TK_SEL]
  uint32_t* segdone;  // [nseg][4]   tiles finished per pass (last-done detection), self-resetting
  uint32_t* mask;     // [ntiles][128]  bit e of a tile = element e lies in the tensor's threshold bin
  int32_t* tiecnt;    // [ntiles]
  uint32_t* done;     // [16] grid-wide last-block counters
  uint32_t* status;   // engine status word (debug-build bounds assertions)
};

struct TkHost {
  TkDev d;
  void* block;
};

__device__ __forceinline__ uint32_t tk_absbits(float v) { return __float_as_uint(v) & 0x7FFFFFFFu; }
__device__ __forceinline__ float tk_comp(float beta, float r, float gamma, float g) {
  return __fadd_rn(__fmul_rn(beta, r), __fmul_rn(gamma, g));     // two roundings, like torch's beta * e + gamma * g
}
__device__ __forceinline__ uint32_t ldcg_u32(const uint32_t* p) { return __ldcg(p); }

// Block-wide search of the bin that holds the krem-th largest element: bins are walked from NB-1 downwards.
// Every thread must call it.  Zeroes the histogram.  Returns false when the histogram holds fewer than krem entries.
template <int NB, bool GLOBAL>
__device__ __forceinline__ bool cta_select(uint32_t* h, uint32_t krem, uint32_t& bin, uint32_t& rem, uint32_t& cnt) {
  constexpr int PER = NB / GB_THREADS;
  __shared__ uint32_t wsum[GB_THREADS / 32];
  __shared__ uint32_t res[4];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int hi = NB - 1 - threadIdx.x * PER;
  uint32_t c[PER];
  uint32_t mine = 0;
#pragma unroll
  for (int i = 0; i < PER; ++i) { c[i] = GLOBAL ? ldcg_u32(h + hi - i) : h[hi - i]; mine += c[i]; }
#pragma unroll
  for (int i = 0; i < PER; ++i) if (c[i]) h[hi - i] = 0;
  uint32_t inc = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += n;
  }
  if (threadIdx.x == 0) res[3] = 0;
  if (lane == 31) wsum[wid] = inc;
  __syncthreads();
  uint32_t before = 0;
  for (int w = 0; w < wid; ++w) before += wsum[w];
  uint32_t running = before + inc - mine;                    // elements in strictly higher bins
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    if (running < krem && running + c[i] >= krem) { res[0] = (uint32_t)(hi - i); res[1] = krem - running; res[2] = c[i]; res[3] = 1; }
    running += c[i];
  }
  __syncthreads();
  bin = res[0]; rem = res[1]; cnt = res[2];
  const bool found = res[3] != 0;
  __syncthreads();
  return found;
}

// Tile descriptors of the CTA's next TK_DCACHE tiles in shared memory: the per-tile table look-up (a global load
// at the head of every iteration, and again in the thread that issues the copies) leaves the critical path.
#define TK_DCACHE 128
struct TkTile { int64_t base; int seg, local0, len; };
__device__ __forceinline__ void tk_fill_descs(const SegTable& t, int4* sdesc, int first, int t1) {
  for (int i = threadIdx.x; i < TK_DCACHE && first + i < t1; i += GB_THREADS) {
    int seg, local0, len;
    int64_t base;
    tile_bounds(t, first + i, seg, base, local0, len);
    sdesc[i] = make_int4((int)(uint32_t)(base & 0xFFFFFFFFll), (int)(uint32_t)((uint64_t)base >> 32), seg,
                         (int)((uint32_t)len | ((uint32_t)(local0 / GB_TILE) << 13)));
  }
}
__device__ __forceinline__ TkTile tk_desc(const int4* sdesc, int i) {
  const int4 d = sdesc[i];
  TkTile r;
  r.base = (int64_t)(((uint64_t)(uint32_t)d.y << 32) | (uint32_t)d.x);
  r.seg = d.z;
  r.len = d.w & 0x1FFF;
  r.local0 = (int)((uint32_t)d.w >> 13) * GB_TILE;
  return r;
}

// the same for one warp and 1024 bins: a lane owns 32 consecutive bins = eight 16-byte loads in flight at once
__device__ __forceinline__ bool warp_select1024(uint32_t* h, uint32_t krem, uint32_t& bin, uint32_t& rem, uint32_t& cnt) {
  const int lane = threadIdx.x & 31;
  const int lo = 992 - lane * 32;                            // lane 0 owns the top bins 992..1023
  uint4* h4 = reinterpret_cast<uint4*>(h + lo);
  uint4 c4[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) c4[i] = __ldcg(h4 + i);
  uint32_t mine = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) mine += c4[i].x + c4[i].y + c4[i].z + c4[i].w;
  if (mine) {
#pragma unroll
    for (int i = 0; i < 8; ++i) h4[i] = make_uint4(0u, 0u, 0u, 0u);
  }
  uint32_t inc = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += n;
  }
  uint32_t running = inc - mine;                             // entries in strictly higher bins
  uint32_t fb = 0, fr = 0, fc = 0;
  bool found = false;
  if (running < krem && running + mine >= krem) {
#pragma unroll
    for (int i = 7; i >= 0; --i) {
      const uint32_t c[4] = {c4[i].w, c4[i].z, c4[i].y, c4[i].x};          // descending bin order
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (!found && running + c[q] >= krem) { fb = (uint32_t)(lo + i * 4 + 3 - q); fr = krem - running; fc = c[q]; found = true; }
        running += c[q];
      }
    }
  }
  const uint32_t who = __ballot_sync(0xffffffffu, found);
  if (!who) return false;
  const int src = __ffs(who) - 1;
  bin = __shfl_sync(0xffffffffu, fb, src);
  rem = __shfl_sync(0xffffffffu, fr, src);
  cnt = __shfl_sync(0xffffffffu, fc, src);
  return true;
}

// ---------------------------------------------------------------- bulk-copy (TMA) tile pipeline
// Tiles are contiguous 16 KB runs, so the one-dimensional bulk copy (cp.async.bulk, SASS UBLKCP) moves them:
// one thread issues a tile's loads TK_STAGES tiles ahead, completion is counted in bytes on an mbarrier, nobody
// holds the data in registers while it is in flight.
#define TK_STAGES 3
__device__ __forceinline__ uint32_t tk_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void tk_mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(tk_smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void tk_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tk_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tk_mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(tk_smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// this GPU's memory only (gradient / residual buffers are local); no multicast
__device__ __forceinline__ void tk_bulk_load(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(tk_smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(tk_smem_u32(bar))
               : "memory");
}

// ---------------------------------------------------------------- pass 1
template <bool WRITE>
__global__ void __launch_bounds__(GB_THREADS, 2) tk_pass1(SegTable t, const float* __restrict__ grad, float* work, float beta,
                                                       float gamma, TkDev d, StepCtl* ctl) {
  extern __shared__ __align__(128) unsigned char tk_dyn[];
  float* sg = reinterpret_cast<float*>(tk_dyn);                       // [TK_STAGES][GB_TILE] gradient tiles
  float* sr = sg + TK_STAGES * GB_TILE;                               // [TK_STAGES][GB_TILE] residual tiles (WRITE)
  __shared__ uint32_t sh[2048];
  __shared__ __align__(16) int4 sdesc[TK_DCACHE];
  __shared__ __align__(8) uint64_t full[TK_STAGES];
  __shared__ int s_last;
  if (WRITE && ctl->next_step == 0) { beta = 0.f; gamma = 1.f; }   // empty memory: compensate() is the identity
  const int t0 = (int)((int64_t)blockIdx.x * t.ntiles / gridDim.x);          // even split: the grid is a multiple of the SM count
  const int t1 = (int)((int64_t)(blockIdx.x + 1) * t.ntiles / gridDim.x);
  for (int b = threadIdx.x; b < 2048; b += GB_THREADS) sh[b] = 0;
  if (threadIdx.x == 0) {
    for (int s = 0; s < TK_STAGES; ++s) tk_mbar_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  int dbase = t0;                                                   // first tile held by the descriptor cache
  tk_fill_descs(t, sdesc, dbase, t1);
  __syncthreads();
  auto issue = [&](int tile, int s) {                               // one thread
    const TkTile x = tk_desc(sdesc, tile - dbase);
    const uint32_t bytes = (uint32_t)((x.len + 3) & ~3) * 4u;       // the pad up to the next 128 elements belongs to the segment
    tk_mbar_expect_tx(&full[s], WRITE ? 2 * bytes : bytes);
    tk_bulk_load(sg + s * GB_TILE, grad + x.base, bytes, &full[s]);
    if (WRITE) tk_bulk_load(sr + s * GB_TILE, work + x.base, bytes, &full[s]);
  };
  if (threadIdx.x == 0)
    for (int s = 0; s < TK_STAGES && t0 + s < t1; ++s) issue(t0 + s, s);
  int cur = t0 < t1 ? tk_desc(sdesc, 0).seg : -1;
  int mine = 0;                                   // tiles of `cur` done by this CTA
  auto write_sel = [&](uint32_t bin, uint32_t rem) {
    uint32_t* s = d.sel + cur * TK_SEL;
    s[S_BIN] = bin; s[S_KREM] = rem; s[S_CURA] = 0; s[S_FLAGS] = 0; s[S_THR] = 0; s[S_TAKE] = 0;
  };
  auto close_segment = [&]() {
    __syncthreads();
    uint32_t bin, rem, cnt;
    if (t.seg_tile0[cur] >= t0 && t.seg_tile0[cur + 1] <= t1) {
      // the whole tensor was histogrammed by this CTA (all small tensors): select straight from shared memory
      cta_select<2048, false>(sh, (uint32_t)t.seg_k[cur], bin, rem, cnt);
      if (threadIdx.x == 0) write_sel(bin, rem);
      mine = 0;
      return;
    }
    uint32_t* hg = d.hist0 + (size_t)cur * 2048;
    for (int b = threadIdx.x; b < 2048; b += GB_THREADS) {
      const uint32_t c = sh[b];
      if (c) atomicAdd(hg + b, c);
      sh[b] = 0;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();                             // cumulative: orders the CTA's histogram atomics (seen through the barrier)
      const uint32_t total = (uint32_t)(t.seg_tile0[cur + 1] - t.seg_tile0[cur]);
      const uint32_t old = atomicAdd(d.segdone + cur * 4, (uint32_t)mine);
      s_last = (old + (uint32_t)mine == total);
    }
    __syncthreads();
    if (s_last) {                                  // this CTA finished the tensor: choose its threshold bin
      __threadfence();
      cta_select<2048, true>(hg, (uint32_t)t.seg_k[cur], bin, rem, cnt);      // always found: the histogram holds numel >= k
      if (threadIdx.x == 0) { write_sel(bin, rem); d.segdone[cur * 4] = 0; }
    }
    mine = 0;
  };
  for (int tile = t0, it = 0; tile < t1; ++tile, ++it) {
    const int st = it % TK_STAGES;
    const uint32_t par = (uint32_t)(it / TK_STAGES) & 1u;
    if (tile + TK_STAGES >= dbase + TK_DCACHE && dbase + TK_DCACHE < t1) {      // refill the descriptor cache (rare)
      __syncthreads();
      dbase = tile;
      tk_fill_descs(t, sdesc, dbase, t1);
      __syncthreads();
    }
    const TkTile x = tk_desc(sdesc, tile - dbase);
    const int seg = x.seg, len = x.len;
    const int64_t base = x.base;
    if (seg != cur) { close_segment(); cur = seg; }
    ++mine;
    tk_mbar_wait(&full[st], par);
    const float4* g4 = reinterpret_cast<const float4*>(sg + st * GB_TILE);
    const float4* r4 = reinterpret_cast<const float4*>(sr + st * GB_TILE);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int e0 = (j * GB_THREADS + threadIdx.x) * 4;
      if (e0 < len) {
        float4 v = g4[j * GB_THREADS + threadIdx.x];
        if (WRITE) {
          const float4 r = r4[j * GB_THREADS + threadIdx.x];
          v.x = tk_comp(beta, r.x, gamma, v.x);
          v.y = tk_comp(beta, r.y, gamma, v.y);
          v.z = tk_comp(beta, r.z, gamma, v.z);
          v.w = tk_comp(beta, r.w, gamma, v.w);
        }
        if (e0 + 4 <= len) {
          if (WRITE) *reinterpret_cast<float4*>(work + base + e0) = v;
          atomicAdd(&sh[tk_absbits(v.x) >> 20], 1u);
          atomicAdd(&sh[tk_absbits(v.y) >> 20], 1u);
          atomicAdd(&sh[tk_absbits(v.z) >> 20], 1u);
          atomicAdd(&sh[tk_absbits(v.w) >> 20], 1u);
        } else {
          const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            if (e0 + q < len) {
              if (WRITE) work[base + e0 + q] = vv[q];
              atomicAdd(&sh[tk_absbits(vv[q]) >> 20], 1u);
            }
          }
        }
      }
    }
    __syncthreads();                               // every thread has read stage st: refill it
    if (threadIdx.x == 0 && tile + TK_STAGES < t1) issue(tile + TK_STAGES, st);
  }
  if (cur >= 0) close_segment();
  // grid-wide last block: advance the device-side step block (what gb_tick_kernel does for the other codecs)
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(d.done, 1u) == gridDim.x - 1) {
      d.done[0] = 0;
      ctl->epoch += 1;
      ctl->step = ctl->next_step;
      ctl->next_step += 1;
    }
  }
}

// ---------------------------------------------------------------- pass 2
struct TkOut {
  float* vals[GB_MAX_WORLD];
  int32_t* idx[GB_MAX_WORLD];
  int2* hdr[GB_MAX_WORLD];      // [ntiles + 1] certain picks, then [ntiles + 1] threshold-bin picks
  int n;
  int mc;                       // entry 0 is the NVLS multicast view of my slot (n == 1): one multimem.st serves all ranks
};
__device__ __forceinline__ void tk_put(const TkOut& out, int r, size_t psh, int pos, float v, int32_t ix) {
  wire_st_u32((char*)out.vals[r] + psh + (size_t)pos * 4, __float_as_uint(v), out.mc);
  wire_st_u32((char*)out.idx[r] + psh + (size_t)pos * 4, (uint32_t)ix, out.mc);
}
__device__ __forceinline__ void tk_put_hdr(const TkOut& out, int r, size_t psh, int slot, int begin, int count) {
  wire_st_u2((char*)out.hdr[r] + psh + (size_t)slot * 8, make_uint2((uint32_t)begin, (uint32_t)count), out.mc);
}
// peers are addressed with compile-time indices (the kernel parameter stays in the constant bank, no local copy)
#define TK_PEERS(r) _Pragma("unroll") for (int r = 0; r < GB_MAX_WORLD; ++r) if (r < out.n)
#define TK_VALS(o, r, sh) ((float*)((char*)(o).vals[r] + (sh)))
#define TK_IDX(o, r, sh) ((int32_t*)((char*)(o).idx[r] + (sh)))
#define TK_HDR(o, r, sh) ((int2*)((char*)(o).hdr[r] + (sh)))

#define TK_STAGES2 2            // pass 2: two tiles in flight per CTA, four CTAs per SM (the pass is issue-bound, not copy-bound)
#define TK_STAGE_CAP 512        // certain picks of one tile staged in shared memory (more: the tile is re-read)

// Mask layout: word 4 * g + q of a tile, bit l  <->  element 128 * g + 4 * l + q  (g = 0..31 groups of 128 elements,
// one warp-wide ballot per q).  A lane of the tail kernels loads the uint4 of group `lane`, i.e. 128 consecutive elements.
template <bool ZERO>
__global__ void __launch_bounds__(GB_THREADS, 4) tk_pass2(SegTable t, float* work, TkDev d, TkOut out, const StepCtl* ctl,
                                                       size_t par_stride) {
  extern __shared__ __align__(128) unsigned char tk_dyn[];
  float* sw = reinterpret_cast<float*>(tk_dyn);                                 // [TK_STAGES2][GB_TILE] tiles of `work`
  float* sval = sw + TK_STAGES2 * GB_TILE;                                       // [3][TK_STAGE_CAP]
  int32_t* sidx = reinterpret_cast<int32_t*>(sval + 3 * TK_STAGE_CAP);          // [3][TK_STAGE_CAP]
  __shared__ uint32_t sh1[1024];
  __shared__ __align__(16) int4 sdesc[TK_DCACHE];
  __shared__ __align__(8) uint64_t full[TK_STAGES2];
  __shared__ int scnt[3], s_tc[3], s_begin[3], s_len[3], s_local0[3], s_ovcnt[3], s_kb[3];
  __shared__ uint32_t s_bin[3];
  __shared__ long long s_base[3];
  __shared__ int s_last;
  const size_t psh = (ctl->epoch & 1) ? par_stride : 0;       // parity of the device-side epoch selects the slot
  const int t0 = (int)((int64_t)blockIdx.x * t.ntiles / gridDim.x);
  const int t1 = (int)((int64_t)(blockIdx.x + 1) * t.ntiles / gridDim.x);
  if (t0 >= t1) return;
  for (int b = threadIdx.x; b < 1024; b += GB_THREADS) sh1[b] = 0;
  if (threadIdx.x < 3) scnt[threadIdx.x] = 0;
  if (threadIdx.x == 0) {
    for (int s = 0; s < TK_STAGES2; ++s) tk_mbar_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  int dbase = t0;
  tk_fill_descs(t, sdesc, dbase, t1);
  __syncthreads();
  auto issue = [&](int tile, int s) {                               // one thread
    const TkTile x = tk_desc(sdesc, tile - dbase);
    const uint32_t bytes = (uint32_t)((x.len + 3) & ~3) * 4u;
    tk_mbar_expect_tx(&full[s], bytes);
    tk_bulk_load(sw + s * GB_TILE, work + x.base, bytes, &full[s]);
  };
  if (threadIdx.x == 0)
    for (int s = 0; s < TK_STAGES2 && t0 + s < t1; ++s) issue(t0 + s, s);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  int cur = tk_desc(sdesc, 0).seg;
  uint32_t bin = d.sel[cur * TK_SEL + S_BIN];
  int mine = 0;
  auto extend = [&](bool found, uint32_t b1, uint32_t rem) {          // thread 0: threshold prefix 11 -> 21 bits
    uint32_t* s = d.sel + cur * TK_SEL;
    if (found) { s[S_BIN] = (bin << 10) | b1; s[S_KREM] = rem; }
    else s[S_FLAGS] = TKF_TAKEALL;                 // only zeros are left in the bin: they would add nothing
  };
  auto close_segment = [&]() {
    __syncthreads();
    uint32_t b1, rem, cnt;
    const uint32_t krem = d.sel[cur * TK_SEL + S_KREM];
    if (t.seg_tile0[cur] >= t0 && t.seg_tile0[cur + 1] <= t1) {     // tensor local to this CTA: no global histogram
      const bool found = cta_select<1024, false>(sh1, krem, b1, rem, cnt);
      if (threadIdx.x == 0) extend(found, b1, rem);
      mine = 0;
      return;
    }
    uint32_t* hg = d.hist1 + (size_t)cur * 1024;
    for (int b = threadIdx.x; b < 1024; b += GB_THREADS) {
      const uint32_t c = sh1[b];
      if (c) atomicAdd(hg + b, c);
      sh1[b] = 0;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      const uint32_t total = (uint32_t)(t.seg_tile0[cur + 1] - t.seg_tile0[cur]);
      const uint32_t old = atomicAdd(d.segdone + cur * 4 + 1, (uint32_t)mine);
      s_last = (old + (uint32_t)mine == total);
    }
    __syncthreads();
    if (s_last) {                                  // tensor complete: extend the threshold prefix by 10 bits
      __threadfence();
      const bool found = cta_select<1024, true>(hg, krem, b1, rem, cnt);
      if (threadIdx.x == 0) { extend(found, b1, rem); d.segdone[cur * 4 + 1] = 0; }
    }
    mine = 0;
  };
  // stores the staged certain picks of the tile in buffer b into every peer's payload and zeroes them in place;
  // its range was claimed by thread 0 one barrier earlier
  auto copy_out = [&](int b) {
    const int tc = s_tc[b], begin = s_begin[b];
    if (tc == 0) return;
    const int64_t tb = (int64_t)s_base[b];
    const int l0 = s_local0[b];
    if (tc <= TK_STAGE_CAP) {
      const float* sv = sval + b * TK_STAGE_CAP;
      const int32_t* si = sidx + b * TK_STAGE_CAP;
      for (int i = threadIdx.x; i < tc; i += GB_THREADS) {
        const float v = sv[i];
        const int32_t ix = si[i];
        TK_PEERS(r) tk_put(out, r, psh, begin + i, v, ix);
        if (ZERO) work[tb + (ix - l0)] = 0.f;
      }
    } else {
      // crowded tile: more certain picks than the staging buffer holds -> re-read the tile from memory
      const int len = s_len[b];
      const uint32_t tbin = s_bin[b];
      for (int e = threadIdx.x; e < len; e += GB_THREADS) {
        const float v = __ldcg(work + tb + e);
        if ((tk_absbits(v) >> 20) > tbin) {
          const int pos = begin + atomicAdd(&s_ovcnt[b], 1);
          TK_PEERS(r) tk_put(out, r, psh, pos, v, l0 + e);
          if (ZERO) work[tb + e] = 0.f;
        }
      }
    }
  };
  // Per element: |x| bits against the two bin bounds.  Elements of the threshold bin vote into the tile mask (one
  // ballot per element slot); anything at or above the bin only sets a bit in the thread's 16-bit "rare" word and is
  // handled after the streaming part by a short data-driven loop (~1.7 % of the elements), so the common path has no
  // divergent branches.
#define TK_CLASSIFY(VALID)                                                                                         \
  _Pragma("unroll") for (int j = 0; j < 4; ++j) {                                                                  \
    const float4 v = w4[j * GB_THREADS + threadIdx.x];                                                             \
    const uint32_t bb[4] = {tk_absbits(v.x), tk_absbits(v.y), tk_absbits(v.z), tk_absbits(v.w)};                   \
    bool pc[4];                                                                                                    \
    _Pragma("unroll") for (int q = 0; q < 4; ++q) {                                                                \
      const bool ge = (VALID) && bb[q] >= lo_cand;                                                                 \
      if (ge) rm |= 1u << (j * 4 + q);                                                                             \
      pc[q] = ge && bb[q] < lo_def;                                                                                \
    }                                                                                                              \
    const uint32_t m0 = __ballot_sync(0xffffffffu, pc[0]), m1 = __ballot_sync(0xffffffffu, pc[1]),                 \
                   m2 = __ballot_sync(0xffffffffu, pc[2]), m3 = __ballot_sync(0xffffffffu, pc[3]);                 \
    if (lane == 0) *reinterpret_cast<uint4*>(mrow + (j * 8 + wid) * 4) = make_uint4(m0, m1, m2, m3);               \
  }
  int it = 0, b = 0;                               // b = it % 3: staging buffer of this tile
  int pend_tc = 0, pend_b = -1, pend_k = 0;        // thread 0: claim of the previous tile, not yet published
  uint32_t pend_raw = 0;
  for (int tile = t0; tile < t1; ++tile, ++it) {
    const int st = it & 1;
    const uint32_t par = (uint32_t)(it >> 1) & 1u;
    if (tile + TK_STAGES2 >= dbase + TK_DCACHE && dbase + TK_DCACHE < t1) {      // refill the descriptor cache (rare)
      __syncthreads();
      dbase = tile;
      tk_fill_descs(t, sdesc, dbase, t1);
      __syncthreads();
    }
    const TkTile x = tk_desc(sdesc, tile - dbase);
    const int seg = x.seg, len = x.len, local0 = x.local0;
    const int64_t base = x.base;
    if (seg != cur) { close_segment(); cur = seg; bin = d.sel[cur * TK_SEL + S_BIN]; }
    ++mine;
    tk_mbar_wait(&full[st], par);
    const float* swt = sw + st * GB_TILE;
    const float4* w4 = reinterpret_cast<const float4*>(swt);
    const uint32_t lo_def = bin < 2047u ? (bin + 1u) << 20 : 0x80000000u;      // |x| bits >= lo_def: a higher bin, certain pick
    const uint32_t lo_cand = bin ? (bin << 20) : 1u;         // threshold bin; exact zeros never matter
    uint32_t* mrow = d.mask + (size_t)tile * 128;
    uint32_t rm = 0;
    if (len == GB_TILE) {
      TK_CLASSIFY(true)
    } else {
      TK_CLASSIFY((j * GB_THREADS + (int)threadIdx.x) * 4 + q < len)
    }
    if (rm) {
      float* sv = sval + b * TK_STAGE_CAP;
      int32_t* si = sidx + b * TK_STAGE_CAP;
      do {
        const int i = __ffs(rm) - 1;
        rm &= rm - 1;
        const int e = ((i >> 2) * GB_THREADS + threadIdx.x) * 4 + (i & 3);
        const float v = swt[e];
        const uint32_t bb = tk_absbits(v);
        if (bb >= lo_def) {
          const int pos = atomicAdd(&scnt[b], 1);
          if (pos < TK_STAGE_CAP) { sv[pos] = v; si[pos] = local0 + e; }
        } else {
          atomicAdd(&sh1[(bb >> 10) & 1023u], 1u);
        }
      } while (rm);
    }
    // thread 0 claimed the previous tile's payload range right after the last barrier; the atomic's result is only
    // consumed here, so its round trip to L2 overlapped this tile's classification instead of holding up the barrier
    if (threadIdx.x == 0 && pend_b >= 0) {
      const int pb = *(volatile int*)&s_kb[pend_b] + (int)pend_raw;      // first use of the atomic's return value
      // debug build: a tile's certain picks must stay inside its tensor's k payload entries
      if (!GB_ASSERT_STORE(pend_raw + (uint32_t)pend_tc <= (uint32_t)pend_k, d.status)) s_tc[pend_b] = 0;
      s_begin[pend_b] = pb;
      TK_PEERS(r) tk_put_hdr(out, r, psh, tile - 1, pb, pend_tc);
    }
    __syncthreads();      // tile staged, stage st consumed, and every thread has stored the tile before the previous one
    const int bprev = b == 0 ? 2 : b - 1;          // (it - 1) % 3 == (it + 2) % 3
    if (threadIdx.x == 0) {
      if (tile + TK_STAGES2 < t1) issue(tile + TK_STAGES2, st);
      const int tc = scnt[b];
      s_kb[b] = (int)t.seg_kbase[seg];
      pend_raw = tc ? atomicAdd(d.sel + seg * TK_SEL + S_CURA, (uint32_t)tc) : 0u;
      pend_tc = tc;
      pend_k = t.seg_k[seg];
      pend_b = b;
      s_tc[b] = tc; s_len[b] = len; s_local0[b] = local0; s_bin[b] = bin; s_base[b] = (long long)base;
      scnt[bprev] = 0;                             // buffer of the tile after the next one (its last user is stored by now)
      s_ovcnt[b] = 0;
    }
    if (it > 0) copy_out(bprev);
    if (++b == 3) b = 0;
  }
  if (threadIdx.x == 0 && pend_b >= 0) {
    const int pb = s_kb[pend_b] + (int)pend_raw;
    if (!GB_ASSERT_STORE(pend_raw + (uint32_t)pend_tc <= (uint32_t)pend_k, d.status)) s_tc[pend_b] = 0;
    s_begin[pend_b] = pb;
    TK_PEERS(r) tk_put_hdr(out, r, psh, t1 - 1, pb, pend_tc);
  }
  __syncthreads();
  copy_out(b == 0 ? 2 : b - 1);
  close_segment();
#undef TK_CLASSIFY
}

// ---------------------------------------------------------------- tail kernels: one warp per tile
__device__ __forceinline__ bool tk_tile(const SegTable& t, int& tile, int& seg, int64_t& base, int& local0, int& len) {
  tile = blockIdx.x * TK_WPT + (threadIdx.x >> 5);
  if (tile >= t.ntiles) return false;
  tile_bounds(t, tile, seg, base, local0, len);
  return true;
}

// Marked elements of a tile as a list in shared memory (index order), so that the gathers of their values are
// independent loads spread over the 32 lanes instead of a dependent chain per lane.  Returns the number of marked
// elements; the list is only built when it fits TK_LIST entries (else the callers walk the mask lane by lane).
#define TK_LIST 256
__device__ __forceinline__ int tk_build_list(const uint32_t mw[4], int lane, uint16_t* pos) {
  const int mine = __popc(mw[0]) + __popc(mw[1]) + __popc(mw[2]) + __popc(mw[3]);
  int inc = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int n = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += n;
  }
  const int total = __shfl_sync(0xffffffffu, inc, 31);
  if (total <= TK_LIST) {
    int o = inc - mine;
    uint32_t u = mw[0] | mw[1] | mw[2] | mw[3];
    while (u) {                                   // index order: l ascending, then q
      const int l = __ffs(u) - 1;
      u &= u - 1;
#pragma unroll
      for (int q = 0; q < 4; ++q)
        if ((mw[q] >> l) & 1u) pos[o++] = (uint16_t)(lane * 128 + 4 * l + q);
    }
  }
  __syncwarp();
  return total;
}

__global__ void __launch_bounds__(TK_WPT * 32) tk_refine(SegTable t, const float* __restrict__ work, TkDev d) {
  __shared__ uint16_t s_pos[TK_WPT][TK_LIST];
  int tile, seg, local0, len;
  int64_t base;
  const int lane = threadIdx.x & 31;
  // the mask row does not depend on the tile descriptor: both loads are in flight together, then one 16-byte load
  // fetches the tensor's selection state (three dependent memory latencies per warp instead of six)
  const int tile_guess = min(blockIdx.x * TK_WPT + (threadIdx.x >> 5), t.ntiles - 1);
  const uint4 m = *reinterpret_cast<const uint4*>(d.mask + (size_t)tile_guess * 128 + lane * 4);
  if (!tk_tile(t, tile, seg, base, local0, len)) return;
  uint32_t* s = d.sel + seg * TK_SEL;
  const uint4 s0 = *reinterpret_cast<const uint4*>(s);         // BIN, KREM, CURA, FLAGS
  if (s0.w & TKF_TAKEALL) return;
  uint16_t* pos = s_pos[threadIdx.x >> 5];
  const uint32_t prefix = s0.x;                                // 21 bits
  const uint32_t mw[4] = {m.x, m.y, m.z, m.w};
  uint32_t* h2 = d.hist2 + (size_t)seg * 1024;
  const int total = tk_build_list(mw, lane, pos);
  if (total <= TK_LIST) {
    for (int i0 = 0; i0 < total; i0 += 128) {                  // four independent gathers per lane in flight
      uint32_t key[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * 32 + lane;
        key[u] = i < total ? tk_absbits(__ldcg(work + base + pos[i])) : 0xFFFFFFFFu;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if ((key[u] >> 10) == prefix) atomicAdd(h2 + (key[u] & 1023u), 1u);
    }
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint32_t w = mw[i];
      while (w) {
        const int bit = __ffs(w) - 1;
        w &= w - 1;
        const uint32_t key = tk_absbits(__ldcg(work + base + lane * 128 + 4 * bit + i));
        if ((key >> 10) == prefix) atomicAdd(h2 + (key & 1023u), 1u);
      }
    }
  }
  __syncwarp();                                                // orders the lanes' atomics before lane 0's (cumulative) fence
  const uint32_t total_tiles = (uint32_t)(t.seg_tile0[seg + 1] - t.seg_tile0[seg]);
  int last = total_tiles == 1;                                 // a one-tile tensor needs no counter
  if (!last) {
    if (lane == 0) {
      __threadfence();
      last = atomicAdd(d.segdone + seg * 4 + 2, 1u) == total_tiles - 1;
    }
    last = __shfl_sync(0xffffffffu, last, 0);
  }
  if (last) {                                                  // this warp finished the tensor: exact threshold
    __threadfence();
    uint32_t b2 = 0, take = 0, cnt = 0;
    const bool found = warp_select1024(h2, s0.y, b2, take, cnt);
    if (lane == 0) {
      if (found) {
        s[S_THR] = (prefix << 10) | b2;
        s[S_TAKE] = take;
        if (cnt != take) s[S_FLAGS] = TKF_AMBIG;
      } else {
        s[S_FLAGS] = TKF_TAKEALL;                              // cannot happen (the prefix bin holds >= KREM values); safe
      }
      if (total_tiles != 1) d.segdone[seg * 4 + 2] = 0;
    }
  }
}

__global__ void __launch_bounds__(TK_WPT * 32) tk_ties(SegTable t, const float* __restrict__ work, TkDev d) {
  __shared__ uint16_t s_pos[TK_WPT][TK_LIST];
  int tile, seg, local0, len;
  int64_t base;
  if (!tk_tile(t, tile, seg, base, local0, len)) return;
  const uint32_t* s = d.sel + seg * TK_SEL;
  const uint4 s0 = *reinterpret_cast<const uint4*>(s);
  if (!(s0.w & TKF_AMBIG)) return;
  const int lane = threadIdx.x & 31;
  uint16_t* pos = s_pos[threadIdx.x >> 5];
  const uint32_t thr = s[S_THR];
  const uint4 m = *reinterpret_cast<const uint4*>(d.mask + (size_t)tile * 128 + lane * 4);
  const uint32_t mw[4] = {m.x, m.y, m.z, m.w};
  int c = 0;
  const int total = tk_build_list(mw, lane, pos);
  if (total <= TK_LIST) {
    for (int i0 = 0; i0 < total; i0 += 128) {
      uint32_t key[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * 32 + lane;
        key[u] = i < total ? tk_absbits(__ldcg(work + base + pos[i])) : 0xFFFFFFFFu;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) c += key[u] == thr;
    }
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint32_t w = mw[i];
      while (w) {
        const int bit = __ffs(w) - 1;
        w &= w - 1;
        c += tk_absbits(__ldcg(work + base + lane * 128 + 4 * bit + i)) == thr;
      }
    }
  }
  c = warp_sum_i(c);
  const int first = t.seg_tile0[seg], ntile = t.seg_tile0[seg + 1] - first;
  int last = 0;
  if (lane == 0) {
    d.tiecnt[tile] = c;
    __threadfence();
    last = atomicAdd(d.segdone + seg * 4 + 3, 1u) == (uint32_t)ntile - 1;
  }
  last = __shfl_sync(0xffffffffu, last, 0);
  if (last) {                                                  // exclusive scan of the tensor's tile counts, in tile order
    __threadfence();
    int carry = 0;
    for (int c0 = 0; c0 < ntile; c0 += 32) {
      const int i = c0 + lane;
      const int v = i < ntile ? __ldcg(d.tiecnt + first + i) : 0;
      int inc = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
      }
      if (i < ntile) d.tiecnt[first + i] = carry + inc - v;
      carry += __shfl_sync(0xffffffffu, inc, 31);
    }
    if (lane == 0) d.segdone[seg * 4 + 3] = 0;
  }
}

template <bool ZERO>
__global__ void __launch_bounds__(TK_WPT * 32) tk_finish(SegTable t, float* work, TkDev d, TkOut out, PeerPtrs peers, int channel,
                                                        const StepCtl* ctl, size_t par_stride) {
  __shared__ uint16_t s_pos[TK_WPT][TK_LIST];
  __shared__ float s_val[TK_WPT][TK_LIST];
  const uint32_t epoch = ctl->epoch;
  const size_t psh = (epoch & 1) ? par_stride : 0;
  int tile, seg, local0, len;
  int64_t base;
  const int tile_guess = min(blockIdx.x * TK_WPT + (threadIdx.x >> 5), t.ntiles - 1);
  const uint4 m = *reinterpret_cast<const uint4*>(d.mask + (size_t)tile_guess * 128 + (threadIdx.x & 31) * 4);
  if (tk_tile(t, tile, seg, base, local0, len)) {
    uint32_t* s = d.sel + seg * TK_SEL;
    const int lane = threadIdx.x & 31;
    uint16_t* pos = s_pos[threadIdx.x >> 5];
    float* val = s_val[threadIdx.x >> 5];
    const uint4 s0 = *reinterpret_cast<const uint4*>(s);       // BIN, KREM, CURA, FLAGS
    const uint2 s1 = *reinterpret_cast<const uint2*>(s + 4);   // THR, TAKE
    const uint32_t flags = s0.w, thr = s1.x;
    const int take = (int)s1.y;
    const bool takeall = flags & TKF_TAKEALL, ambig = flags & TKF_AMBIG;
    const uint32_t mw[4] = {m.x, m.y, m.z, m.w};
    const int xoff = t.ntiles + 1;
    const int nmarked = tk_build_list(mw, lane, pos);
    if (nmarked <= TK_LIST) {
      // common case: gather all marked values at once (independent loads), decide in index order with ballots,
      // compact the selected ones in place in shared memory, claim one payload range, store
      for (int i0 = 0; i0 < nmarked; i0 += 128) {
        float v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int i = i0 + u * 32 + lane;
          v[u] = i < nmarked ? __ldcg(work + base + pos[i]) : 0.f;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int i = i0 + u * 32 + lane;
          if (i < nmarked) val[i] = v[u];
        }
      }
      __syncwarp();
      int nsel = 0;
      int tie_run = ambig ? d.tiecnt[tile] : 0;
      const uint32_t lt = (1u << lane) - 1u;
      for (int i0 = 0; i0 < nmarked; i0 += 32) {
        const int i = i0 + lane;
        const bool in = i < nmarked;
        const float v = in ? val[i] : 0.f;
        const uint16_t e = in ? pos[i] : (uint16_t)0;
        const uint32_t key = tk_absbits(v);
        const bool gt = in && (takeall || key > thr);
        const bool eq = in && !gt && key == thr;
        bool chosen = gt || eq;
        if (ambig) {
          const uint32_t eb = __ballot_sync(0xffffffffu, eq);
          chosen = gt || (eq && tie_run + __popc(eb & lt) < take);
          tie_run += __popc(eb);
        }
        const uint32_t cb = __ballot_sync(0xffffffffu, chosen);
        __syncwarp();
        if (chosen) { const int p = nsel + __popc(cb & lt); pos[p] = e; val[p] = v; }
        nsel += __popc(cb);
        __syncwarp();
      }
      int begin = 0;
      if (lane == 0) {
        if (nsel) {
          const uint32_t at = atomicAdd(s + S_CURA, (uint32_t)nsel);
          begin = (int)t.seg_kbase[seg] + (int)at;
          if (!GB_ASSERT_STORE(at + (uint32_t)nsel <= (uint32_t)t.seg_k[seg], d.status)) nsel = 0;     // debug build
        }
        TK_PEERS(r) tk_put_hdr(out, r, psh, xoff + tile, begin, nsel);
      }
      begin = __shfl_sync(0xffffffffu, begin, 0);
#ifdef GB_DEBUG
      nsel = __shfl_sync(0xffffffffu, nsel, 0);
#endif
      for (int i = lane; i < nsel; i += 32) {
        const int e = pos[i];
        const float v = val[i];
        TK_PEERS(r) tk_put(out, r, psh, begin + i, v, local0 + e);
        if (ZERO) work[base + e] = 0.f;
      }
    } else {
      // crowded tile (many values share the threshold bin): walk the mask lane by lane.  A lane owns 128 consecutive
      // elements; word q, bit l is element 4 l + q, so index order is l ascending, then q.
      uint32_t gt[4] = {0, 0, 0, 0}, eq[4] = {0, 0, 0, 0};
      int neq = 0;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint32_t w = mw[q];
        while (w) {
          const int l = __ffs(w) - 1;
          w &= w - 1;
          const uint32_t key = tk_absbits(__ldcg(work + base + lane * 128 + 4 * l + q));
          if (takeall || key > thr) gt[q] |= 1u << l;
          else if (key == thr) { eq[q] |= 1u << l; ++neq; }
        }
      }
      uint32_t ch[4] = {gt[0], gt[1], gt[2], gt[3]};
      if (!ambig) {
#pragma unroll
        for (int q = 0; q < 4; ++q) ch[q] |= eq[q];
      } else {
        int inc = neq;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int n = __shfl_up_sync(0xffffffffu, inc, o);
          if (lane >= o) inc += n;
        }
        int rank = d.tiecnt[tile] + inc - neq;                 // ties before this lane's first element
        uint32_t u = eq[0] | eq[1] | eq[2] | eq[3];
        while (u) {
          const int l = __ffs(u) - 1;
          u &= u - 1;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            if ((eq[q] >> l) & 1u) {
              if (rank < take) ch[q] |= 1u << l;
              ++rank;
            }
          }
        }
      }
      const int mine = __popc(ch[0]) + __popc(ch[1]) + __popc(ch[2]) + __popc(ch[3]);
      int inc = mine;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
      }
      const int total = __shfl_sync(0xffffffffu, inc, 31);
      int begin = 0;
      if (lane == 0) {
        if (total) begin = (int)t.seg_kbase[seg] + (int)atomicAdd(s + S_CURA, (uint32_t)total);
        TK_PEERS(r) tk_put_hdr(out, r, psh, xoff + tile, begin, total);
      }
      begin = __shfl_sync(0xffffffffu, begin, 0);
      int p = begin + inc - mine;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint32_t w = ch[q];
        while (w) {
          const int l = __ffs(w) - 1;
          w &= w - 1;
          const int e = lane * 128 + 4 * l + q;
          const float v = __ldcg(work + base + e);
          TK_PEERS(r) tk_put(out, r, psh, p, v, local0 + e);
          if (ZERO) work[base + e] = 0.f;
          ++p;
        }
      }
    }
  }
  release_flags_last_block(d.done + 4, peers, channel, epoch);
}

// ---------------------------------------------------------------- host side
#define TK_LAUNCHED(p) do { (p)->launches++; GB_CUDA(cudaGetLastError()); } while (0)

#define TK_SMEM_P1 (TK_STAGES * GB_TILE * 4)
#define TK_SMEM_P1W (2 * TK_STAGES * GB_TILE * 4)
#define TK_SMEM_P2 (TK_STAGES2 * GB_TILE * 4 + 3 * TK_STAGE_CAP * 8)

int tk_init(gb_plan* p) {
  GB_CUDA(cudaFuncSetAttribute(tk_pass1<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, TK_SMEM_P1W));
  GB_CUDA(cudaFuncSetAttribute(tk_pass1<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, TK_SMEM_P1));
  GB_CUDA(cudaFuncSetAttribute(tk_pass2<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, TK_SMEM_P2));
  GB_CUDA(cudaFuncSetAttribute(tk_pass2<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, TK_SMEM_P2));
  TkHost* h = new TkHost();
  memset(h, 0, sizeof(*h));
  p->topk2 = h;
  size_t bytes = 0;
  auto add = [&](size_t n) { size_t o = bytes; bytes += (n + 255) & ~size_t(255); return o; };
  const size_t o_h1 = add((size_t)p->nseg * 1024 * 4), o_h2 = add((size_t)p->nseg * 1024 * 4);
  const size_t o_sel = add((size_t)p->nseg * TK_SEL * 4), o_sd = add((size_t)p->nseg * 4 * 4);
  const size_t o_mask = add((size_t)p->ntiles * 512), o_tc = add((size_t)p->ntiles * 4), o_done = add(64);
  GB_CUDA(cudaMalloc(&h->block, bytes));
  GB_CUDA(cudaMemset(h->block, 0, bytes));
  char* b = (char*)h->block;
  h->d.hist0 = p->d_hist;
  h->d.hist1 = (uint32_t*)(b + o_h1);
  h->d.hist2 = (uint32_t*)(b + o_h2);
  h->d.sel = (uint32_t*)(b + o_sel);
  h->d.segdone = (uint32_t*)(b + o_sd);
  h->d.mask = (uint32_t*)(b + o_mask);
  h->d.tiecnt = (int32_t*)(b + o_tc);
  h->d.done = (uint32_t*)(b + o_done);
  h->d.status = p->comm->status_dev;
  return 0;
}

void tk_free(gb_plan* p) {
  TkHost* h = (TkHost*)p->topk2;
  if (!h) return;
  cudaFree(h->block);
  delete h;
  p->topk2 = nullptr;
}

int tk_send(gb_plan* p, const float* d_grad, cudaStream_t st) {
  gb_comm* c = p->comm;
  TkHost* h = (TkHost*)p->topk2;
  const bool mem = p->prm.memory == GB_MEM_RESIDUAL;
  float* work = mem ? p->d_state0 : const_cast<float*>(d_grad);
  ++p->epoch;                                               // host mirror of the device-side epoch
  TkOut out;
  const size_t per = ((size_t)p->ktotal * 4 + 255) & ~size_t(255);
  for (int r = 0; r < c->world; ++r) {
    char* s = gb_slot_of(c, r, p->chan_off, p->slot_bytes, 0, c->rank);
    out.hdr[r] = (int2*)s;
    out.vals[r] = (float*)(s + p->hdr_bytes);
    out.idx[r] = (int32_t*)(s + p->hdr_bytes + per);
  }
  out.n = c->world;
  out.mc = 0;
  if (char* mc = gb_slot_mc(c, p->chan_off, p->slot_bytes, 0, c->rank)) {   // NVLS: the switch replicates every store
    out.n = 1; out.mc = 1;
    out.hdr[0] = (int2*)mc;
    out.vals[0] = (float*)(mc + p->hdr_bytes);
    out.idx[0] = (int32_t*)(mc + p->hdr_bytes + per);
  }
  const int grid = std::min(p->ntiles, 148 * 2);            // pass 1: two resident CTAs per SM, each with a 3-stage tile pipeline
  const int grid2 = std::min(p->ntiles, 148 * 4);
  const int wgrid = (p->ntiles + TK_WPT - 1) / TK_WPT;
  if (mem) tk_pass1<true><<<grid, GB_THREADS, TK_SMEM_P1W, st>>>(p->tab, d_grad, work, p->prm.beta, p->prm.gamma, h->d, p->d_ctl);
  else tk_pass1<false><<<grid, GB_THREADS, TK_SMEM_P1, st>>>(p->tab, d_grad, work, p->prm.beta, p->prm.gamma, h->d, p->d_ctl);
  TK_LAUNCHED(p);
  if (mem) tk_pass2<true><<<grid2, GB_THREADS, TK_SMEM_P2, st>>>(p->tab, work, h->d, out, p->d_ctl, p->par_stride);
  else tk_pass2<false><<<grid2, GB_THREADS, TK_SMEM_P2, st>>>(p->tab, work, h->d, out, p->d_ctl, p->par_stride);
  TK_LAUNCHED(p);
  tk_refine<<<wgrid, TK_WPT * 32, 0, st>>>(p->tab, work, h->d);
  TK_LAUNCHED(p);
  tk_ties<<<wgrid, TK_WPT * 32, 0, st>>>(p->tab, work, h->d);
  TK_LAUNCHED(p);
  if (mem) tk_finish<true><<<wgrid, TK_WPT * 32, 0, st>>>(p->tab, work, h->d, out, gb_comm_peers(c), p->channel, p->d_ctl, p->par_stride);
  else tk_finish<false><<<wgrid, TK_WPT * 32, 0, st>>>(p->tab, work, h->d, out, gb_comm_peers(c), p->channel, p->d_ctl, p->par_stride);
  TK_LAUNCHED(p);
  return 0;
}
