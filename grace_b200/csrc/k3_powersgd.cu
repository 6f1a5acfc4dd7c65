// K3 -- PowerSGD (rank <= 16) as tcgen05/TMEM GEMMs fed by TMA, with the low-rank exchange
// issued tile by tile from the GEMM epilogue.
//
// Reference: grace_dl/dist/compressor/powersgd.py:7-65 + memory/powersgd.py:12-37: per matrix
//   M += e;  Q0 = orth(N(0,1));  P = M Q0;  all_reduce(P)/W;  P^ = orth(P);  Q = M^T P^;
//   all_reduce(Q)/W;  out = P^ Q^T;  e = M - out
// i.e. ~20 + 12r library launches, three cuBLAS SGEMMs and two blocking NCCL all-reduces per
// matrix, x 54 matrices for ResNet-50 (SURVEY §2.9-B PowerSGD row, §3.4).
//
// Here, for ALL matrices of a bucket at once (grouped through a work-item table), five launches after the tick:
//   phase 1 (gb_psgd_send):  k3_prep   CTAs [0, nmat): Philox normal Q0 (or the previous Q on a warm start),
//                                      CholeskyQR2, Q0^T (K-major B operand); band CTAs: M = grad + (M - P^ Q^T) -- the
//                                      error-feedback memory is implicit, see the band kernels -- and clear the
//                                      split-K tile counters
//                            k3_gemm<1> P_r = M Q0 : persistent CTAs walk the work items; TMA -> smem (128B swizzle)
//                                      -> tcgen05.mma kind::tf32 -> TMEM (two alternating accumulators) ->
//                                      tcgen05.ld epilogue.  Split-K slices store their partial tiles; the CTA that
//                                      completes a 128-row tile adds them in slice order (no floating-point atomics:
//                                      the same bits on every run) and stores that tile straight into EVERY peer's
//                                      slot (P2P st.global over NVLink, or one multimem.st) -- the all-reduce is an
//                                      all-gather of tiny P_r tiles + rank-ordered sum on the receiver.  The CTA that
//                                      finishes last releases the epoch flags.
//                            (k3_gemm1_simt: matrices smaller than 8 x 32 on CUDA cores, pushes + releases itself)
//   phase 2 (gb_psgd_mid):   k3_orth   wait flags; P = mean_r P_r (rank order); CholeskyQR2; P^ and P^^T
//                            k3_gemm<2> Q_r = M^T P^ : A operand is MN-major straight from the row-major M tile
//   phase 3 (gb_psgd_recv):  k3_final  wait flags; Q = mean_r Q_r (kept for the next prep); out = P^ Q^T (write only)
// Results are bit-identical on all ranks (every rank sums the same W tiles in the same order).
// Orthogonalisation: two rounds of Cholesky-QR (Gram matrix, Cholesky, triangular solve) -- the
// same Q factor Gram-Schmidt produces, with the reference formula's missing epsilon added
// (SURVEY §9 item 3), in 2 passes per round instead of ~6r launches.
#include "kernels.cuh"

#include <cuda.h>

#include <stdlib.h>

#include <algorithm>
#include <vector>

#define PS_RMAX 16       // max rank handled natively = UMMA N; rows of P / Q are stored with a pitch RP of 4, 8 or 16 floats
#define PS_N 16          // UMMA N (rank padded to the minimum N for M=128)
#define PS_BM 128        // UMMA M / rows of a tile in the GEMM's M dimension
#define PS_BOXK 32       // K elements of one TMA box (= one 128-byte swizzle row of tf32)
#define PS_HALVES 2      // boxes along K per pipeline stage: the two 128-byte pieces of a row are requested together
#define PS_BK (PS_BOXK * PS_HALVES)      // K elements per pipeline stage
#define PS_STAGES 3
#define PS_A_HALF (PS_BM * PS_BOXK * 4)  // 16 KiB
#define PS_B_HALF (PS_N * PS_BOXK * 4)   //  2 KiB
#define PS_A_BYTES (PS_A_HALF * PS_HALVES)
#define PS_B_BYTES (PS_B_HALF * PS_HALVES)
#define PS_SMEM (PS_STAGES * (PS_A_BYTES + PS_B_BYTES) + 1024 + 256)
#define PS_GRID (148 * 2)   // persistent GEMM: two resident CTAs per SM

struct PsMat {          // one matrix of the bucket (device copy)
  int64_t off;          // element offset of M in the flat gradient / output buffers (row pitch m)
  int64_t woff;         // element offset of M in `work`, whose rows have a pitch of ldw floats
  int32_t n, m, r;      // rows, cols, effective rank
  int32_t ldw;          // m rounded up to 4: every row of `work` starts on 16 bytes, so TMA can address any matrix
  int32_t prow0, qrow0; // first row of this matrix in the P / Q row spaces
  uint32_t hash;
  int32_t tc;           // 1: tcgen05 path, 0: CUDA-core fallback (TMA constraints not met)
};

struct PsBand { int32_t mat, i0, j0, j1; };   // one CTA of the band kernels: rows [i0, i0 + 32) x columns [j0, j1) of one matrix

struct PsItem {         // one CTA of a grouped GEMM launch
  int32_t mat, tile, m0, k0, k1, nchunks, chunk;   // chunk: index of this K slice inside its tile (items of a tile are consecutive)
};

struct gb_psgd {
  gb_comm* comm;
  int nmat, rank_r, has_mem, use_tc;
  int rp;                   // row pitch of the stored P / Q rows in floats: 4, 8 or 16 (rank rounded up)
  float *d_part1, *d_part2; // split-K partial tiles of the two grouped GEMMs [items][128][rp]
  float* d_qprev;           // the previous exchange's averaged Q [qrows][rp]: implicit residual, and Q0 of a warm start
  int warm;                 // 1: Q0 = previous Q (compressor/powersgd.py:37-39), 0: fresh normal draw every step
  int* d_simt;              // indices of the matrices on the CUDA-core fallback
  int nsimt, simt_rows, simt_cblocks;
  uint32_t seed, step, epoch;
  int64_t padded, wpadded, prows, qrows;   // wpadded: floats of `work` (rows padded to 16 bytes)
  std::vector<PsMat> h_mat;
  std::vector<int64_t> h_off;
  PsMat* d_mat;
  PsItem *d_items1, *d_items2;
  int nitems1, nitems2, ntiles1, ntiles2;
  CUtensorMap* d_maps;      // [nmat][4]: A1 (M, box 32x128), B1 (Q0^T), A2 (M, box 32x32), B2 (P^^T)
  float* d_work;            // compensated gradient M of the last exchange (padded floats); residual = M - P^ Q^T is implicit
  float *d_q0t, *d_pht;     // Q0^T [16][qrows_pad], P^^T [16][prows_pad]  (per-matrix column windows)
  float *d_pacc, *d_qacc;   // local accumulators [prows][4], [qrows][4]
  float *d_ph;              // P^ [prows][4]
  uint32_t *d_cnt1, *d_cnt2, *d_done;
  StepCtl* d_ctl;           // device-side epoch / step (graph-replayable)
  struct PsBand* d_bands;
  int nbands;
  int64_t qpad, ppad;       // row pitch of d_q0t / d_pht
  int chan_p, chan_q;
  size_t off_p, off_q, slot_p, slot_q;
  int64_t launches;
};

// ---------------------------------------------------------------- PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
                   smem_u32(dst)),
               "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
        "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// shared-memory matrix descriptor, 128-byte swizzle, Blackwell version field = 1
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout = 2) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;      // descriptor version
  d |= (uint64_t)layout << 61; // 2 = SWIZZLE_128B, 1 = SWIZZLE_128B with 32-byte atoms (MN-major 32-bit operands)
  return d;
}
// instruction descriptor: D=f32, A=B=tf32, M=128, N=16, A major selectable (0 K-major, 1 MN-major)
__host__ __device__ constexpr uint32_t idesc_tf32(int a_mn_major) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_mn_major << 15) | ((uint32_t)(PS_N >> 3) << 17) |
         ((uint32_t)(PS_BM >> 4) << 24);
}

// ---------------------------------------------------------------- grouped tcgen05 GEMM
// MODE 1:  D[128 x 16] = M[m0:m0+128, k0:k1] * Q0[k0:k1, :]         (A K-major,  K = columns of M)
// MODE 2:  D[128 x 16] = M[k0:k1, m0:m0+128]^T * P^[k0:k1, :]       (A MN-major, K = rows of M)
// warp 0: TMA producer; warp 1: TMEM alloc + MMA issuer; warps 2-5: epilogue (TMEM lane quarter = warp % 4)
struct PsPeers { float* slot[GB_MAX_WORLD]; int n; int mc; };   // mc: slot[0] is the NVLS multicast view, n == 1
__device__ __forceinline__ void shift(PsPeers& o, size_t bytes) {
  for (int r = 0; r < o.n; ++r) o.slot[r] = (float*)((char*)o.slot[r] + bytes);
}


// PERSISTENT: the grid is two CTAs per SM and CTA b walks work items b, b + grid, b + 2 grid, ...  The barriers, the
// TMEM allocation and the pipeline live across items: the producer runs ahead into the next item's tiles while the
// epilogue of the previous one drains, and the two 16-column TMEM accumulators alternate so the MMA of item i+1
// starts before item i has been read back.  (First version: one CTA per item -- ncu showed the SMs active 68 % of the
// launch: 2.24 waves of ~10-stage items, each paying its own barrier init, TMEM alloc, pipeline fill and epilogue.)
template <int MODE, int RP>
__global__ void __launch_bounds__(192, 2) k3_gemm(const PsItem* __restrict__ items, int nitems, const PsMat* __restrict__ mats,
                                                  const CUtensorMap* __restrict__ maps, float* part, uint32_t* tile_cnt,
                                                  PsPeers peers, const StepCtl* ctl, size_t par_stride, uint32_t* done,
                                                  PeerPtrs flag_peers, int channel, int release) {
  extern __shared__ uint8_t smem_raw[];
  if (ctl->epoch & 1) shift(peers, par_stride);
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;
  uint8_t* sB = smem + PS_STAGES * PS_A_BYTES;
  uint64_t* full = (uint64_t*)(smem + PS_STAGES * (PS_A_BYTES + PS_B_BYTES));
  uint64_t* empty = full + PS_STAGES;
  uint64_t* accum_full = empty + PS_STAGES;       // [2] MMAs of an item have landed in accumulator b
  uint64_t* accum_free = accum_full + 2;          // [2] the epilogue has read accumulator b back
  uint32_t* tmem_slot = (uint32_t*)(accum_free + 2);
  uint32_t* last_flag = tmem_slot + 1;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < PS_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&accum_full[b], 1); mbar_init(&accum_free[b], 128); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(32u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0 && lane == 0) {
    // ---- TMA producer: one running stage counter over all items of this CTA
    uint32_t g = 0;
    for (int ii = blockIdx.x; ii < nitems; ii += gridDim.x) {
      const PsItem it = items[ii];
      const CUtensorMap* mapA = maps + it.mat * 4 + (MODE == 1 ? 0 : 2);
      const CUtensorMap* mapB = maps + it.mat * 4 + (MODE == 1 ? 1 : 3);
      const int nkb = (it.k1 - it.k0 + PS_BK - 1) / PS_BK;
      for (int kb = 0; kb < nkb; ++kb, ++g) {
        const int s = g % PS_STAGES;
        if (g >= PS_STAGES) mbar_wait(&empty[s], ((g / PS_STAGES) - 1) & 1);
        mbar_expect_tx(&full[s], PS_A_BYTES + PS_B_BYTES);
        const int k = it.k0 + kb * PS_BK;
#pragma unroll
        for (int h = 0; h < PS_HALVES; ++h) {
          const int kh = k + h * PS_BOXK;
          if (MODE == 1) {
            tma_load_2d(sA + s * PS_A_BYTES + h * PS_A_HALF, mapA, &full[s], kh, it.m0);   // box 32 (K) x 128 (rows)
          } else {
#pragma unroll
            for (int b = 0; b < 4; ++b)                                              // 4 boxes 32 (cols) x 32 (K rows)
              tma_load_2d(sA + s * PS_A_BYTES + h * PS_A_HALF + b * 4096, mapA, &full[s], it.m0 + b * 32, kh);
          }
          tma_load_2d(sB + s * PS_B_BYTES + h * PS_B_HALF, mapB, &full[s], kh, 0);   // box 32 (K) x 16
        }
      }
    }
  } else if (warp == 1 && lane == 0) {
    // ---- MMA issuer (one thread)
    constexpr uint32_t idesc = idesc_tf32(MODE == 2);
    uint32_t g = 0, n = 0;
    for (int ii = blockIdx.x; ii < nitems; ii += gridDim.x, ++n) {
      const PsItem it = items[ii];
      const int nkb = (it.k1 - it.k0 + PS_BK - 1) / PS_BK;
      const uint32_t buf = n & 1u, use = n >> 1;
      if (use > 0) {                                      // accumulator `buf` was used by item n - 2: wait for its read-back
        mbar_wait(&accum_free[buf], (use - 1) & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      }
      const uint32_t tmem_d = tmem_base + buf * PS_N;
      for (int kb = 0; kb < nkb; ++kb, ++g) {
        const int s = g % PS_STAGES;
        mbar_wait(&full[s], (g / PS_STAGES) & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t a0 = smem_u32(sA + s * PS_A_BYTES), b0 = smem_u32(sB + s * PS_B_BYTES);
#pragma unroll
        for (int j = 0; j < PS_BK / 8; ++j) {                                        // UMMA_K = 8 for tf32
          const uint32_t ah = a0 + (j / 4) * PS_A_HALF, bh = b0 + (j / 4) * PS_B_HALF;  // which 32-wide box of the stage
          const int jj = j & 3;
          uint64_t ad, bd;
          if (MODE == 1) ad = smem_desc_sw128(ah + jj * 32, 16, 1024);              // K-major: +32 B per K step
          // MN-major tf32 must use the 128B swizzle with 32-byte atoms (atom = 32 MN x 4 K): blocks of 32
          // columns are 4096 B apart (LBO), groups of 4 K rows 512 B apart (SBO), 8 K rows per MMA
          else ad = smem_desc_sw128(ah + jj * 1024, 4096, 512, 1);
          bd = smem_desc_sw128(bh + jj * 32, 16, 1024);
          umma_tf32(tmem_d, ad, bd, idesc, (kb | j) != 0);
        }
        umma_commit(&empty[s]);                 // frees the smem stage once these MMAs have read it
      }
      umma_commit(&accum_full[buf]);
    }
  } else if (warp >= 2) {
    // ---- epilogue: TMEM -> registers -> accumulate / push
    uint32_t n = 0;
    for (int ii = blockIdx.x; ii < nitems; ii += gridDim.x, ++n) {
      const PsItem it = items[ii];
      const PsMat mt = mats[it.mat];
      const uint32_t buf = n & 1u, use = n >> 1;
      mbar_wait(&accum_full[buf], use & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const int q = warp & 3;
      uint32_t v[16];
      tmem_ld16(tmem_base + buf * PS_N + ((uint32_t)(q * 32) << 16), v);
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&accum_free[buf])) : "memory");
      const int row = it.m0 + q * 32 + lane;
      const int nrows = MODE == 1 ? mt.n : mt.m;
      const int row0 = MODE == 1 ? mt.prow0 : mt.qrow0;
      const bool live = row < nrows;
      if (it.nchunks == 1) {
        if (live) {
#pragma unroll
          for (int c4 = 0; c4 < RP / 4; ++c4) {
            const float4 o = make_float4(__uint_as_float(v[c4 * 4]), __uint_as_float(v[c4 * 4 + 1]), __uint_as_float(v[c4 * 4 + 2]),
                                         __uint_as_float(v[c4 * 4 + 3]));
            for (int r = 0; r < peers.n; ++r)
              wire_st_u4(reinterpret_cast<float4*>(peers.slot[r] + (size_t)(row0 + row) * RP) + c4,
                         make_uint4(__float_as_uint(o.x), __float_as_uint(o.y), __float_as_uint(o.z), __float_as_uint(o.w)), peers.mc);
          }
        }
      } else {
        // split-K without floating-point atomics: every K slice stores its partial tile, the CTA that completes the
        // tile adds the slices in slice order (same bits on every run and every rank) and ships the sum
        const int rl = q * 32 + lane;                                 // row inside the 128-row tile
        float4* mine = reinterpret_cast<float4*>(part + ((size_t)ii * PS_BM + rl) * RP);
#pragma unroll
        for (int c4 = 0; c4 < RP / 4; ++c4)
          mine[c4] = make_float4(__uint_as_float(v[c4 * 4]), __uint_as_float(v[c4 * 4 + 1]), __uint_as_float(v[c4 * 4 + 2]),
                                 __uint_as_float(v[c4 * 4 + 3]));
        __threadfence();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (threadIdx.x == 64) *last_flag = (atomicAdd(tile_cnt + it.tile, 1u) == (uint32_t)it.nchunks - 1) ? 1u : 0u;
        asm volatile("bar.sync 1, 128;" ::: "memory");
        const bool last = *last_flag != 0u;
        asm volatile("bar.sync 1, 128;" ::: "memory");               // last_flag is reused by the next item
        if (last && live) {                       // this CTA completed the tile: ship it to every peer
          __threadfence();
          const size_t item0 = (size_t)ii - it.chunk;
#pragma unroll
          for (int c4 = 0; c4 < RP / 4; ++c4) {
            float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int ch = 0; ch < it.nchunks; ++ch) {
              const float4 x = __ldcg(reinterpret_cast<const float4*>(part + ((item0 + ch) * PS_BM + rl) * RP) + c4);
              o.x += x.x; o.y += x.y; o.z += x.z; o.w += x.w;
            }
            for (int r = 0; r < peers.n; ++r)
              wire_st_u4(reinterpret_cast<float4*>(peers.slot[r] + (size_t)(row0 + row) * RP) + c4,
                         make_uint4(__float_as_uint(o.x), __float_as_uint(o.y), __float_as_uint(o.z), __float_as_uint(o.w)), peers.mc);
          }
        }
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(32u) : "memory");
  }
  // every matrix of the bucket went through this launch: the CTA that finishes last publishes the epoch (no separate
  // flag kernel); with CUDA-core fallback matrices in the bucket their launch, which follows on the stream, does it
  if (release) release_flags_last_block(done, flag_peers, channel, ctl->epoch);
}

// ---------------------------------------------------------------- CUDA-core fallbacks (shapes TMA cannot address)
// Matrices whose row pitch is not a multiple of 16 bytes (ResNet-50: the 64 x 147 stem convolution) or that are
// smaller than one MMA tile.  `simt` lists their indices, so nobody walks the whole matrix table.  Both kernels store
// their rows straight into every peer's slot and -- running last on the stream in their phase -- release the flags.
template <int RP>
__device__ __forceinline__ void push_row(const PsPeers& peers, size_t row, const float (&s)[RP]) {
#pragma unroll
  for (int c4 = 0; c4 < RP / 4; ++c4)
    for (int r = 0; r < peers.n; ++r)
      wire_st_u4(reinterpret_cast<float4*>(peers.slot[r] + row * RP) + c4,
                 make_uint4(__float_as_uint(s[c4 * 4]), __float_as_uint(s[c4 * 4 + 1]), __float_as_uint(s[c4 * 4 + 2]),
                            __float_as_uint(s[c4 * 4 + 3])), peers.mc);
}

// one warp per row: P[i][c] = sum_j M[i][j] Q0[j][c]
template <int RP>
__global__ void __launch_bounds__(GB_THREADS) k3_gemm1_simt(const PsMat* __restrict__ mats, const int* __restrict__ simt, int nsimt,
                                                           const float* __restrict__ work, const float* __restrict__ q0t,
                                                           int64_t qpad, PsPeers out, uint32_t* done, PeerPtrs peers, int channel,
                                                           const StepCtl* ctl, size_t par_stride) {
  const uint32_t epoch = ctl->epoch;
  if (epoch & 1) shift(out, par_stride);
  const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  int base = 0;
  for (int si = 0; si < nsimt; ++si) {
    const PsMat mt = mats[simt[si]];
    if (gw >= base && gw < base + mt.n) {
      const int i = gw - base;
      const float* rowp = work + mt.woff + (size_t)i * mt.ldw;
      float s[RP];
#pragma unroll
      for (int c = 0; c < RP; ++c) s[c] = 0.f;
      for (int j = lane; j < mt.m; j += 32) {
        const float x = rowp[j];
#pragma unroll
        for (int c = 0; c < RP; ++c) s[c] += x * q0t[(size_t)c * qpad + mt.qrow0 + j];
      }
#pragma unroll
      for (int c = 0; c < RP; ++c) s[c] = warp_sum(s[c]);
      if (lane == 0) push_row<RP>(out, (size_t)(mt.prow0 + i), s);
      break;
    }
    base += mt.n;
  }
  release_flags_last_block(done, peers, channel, epoch);
}

// Q[j][c] = sum_i M[i][j] P^[i][c]: a CTA owns 32 columns, its 8 warps take every 8th row, the slices are added in
// slice order (first version: one thread per column walking all rows, 27 us for the 64 x 147 matrix alone)
template <int RP>
__global__ void __launch_bounds__(GB_THREADS) k3_gemm2_simt(const PsMat* __restrict__ mats, const int* __restrict__ simt, int nsimt,
                                                           const float* __restrict__ work, const float* __restrict__ ph, PsPeers out,
                                                           uint32_t* done, PeerPtrs peers, int channel, const StepCtl* ctl,
                                                           size_t par_stride) {
  const uint32_t epoch = ctl->epoch;
  if (epoch & 1) shift(out, par_stride);
  __shared__ float part[8][32][RP + 1];
  const int lane = threadIdx.x & 31, slice = threadIdx.x >> 5;
  int base = 0;
  for (int si = 0; si < nsimt; ++si) {
    const PsMat mt = mats[simt[si]];
    const int nb = (mt.m + 31) / 32;
    if ((int)blockIdx.x >= base && (int)blockIdx.x < base + nb) {
      const int j = ((int)blockIdx.x - base) * 32 + lane;
      float s[RP];
#pragma unroll
      for (int c = 0; c < RP; ++c) s[c] = 0.f;
      if (j < mt.m) {
#pragma unroll 4
        for (int i = slice; i < mt.n; i += 8) {
          const float x = work[mt.woff + (size_t)i * mt.ldw + j];
          const float* pr = ph + (size_t)(mt.prow0 + i) * RP;
#pragma unroll
          for (int c = 0; c < RP; ++c) s[c] += x * pr[c];
        }
      }
#pragma unroll
      for (int c = 0; c < RP; ++c) part[slice][lane][c] = s[c];
      __syncthreads();
      if (slice == 0 && j < mt.m) {
#pragma unroll
        for (int c = 0; c < RP; ++c) {
          float t = part[0][lane][c];
          for (int k = 1; k < 8; ++k) t += part[k][lane][c];
          s[c] = t;
        }
        push_row<RP>(out, (size_t)(mt.qrow0 + j), s);
      }
      break;
    }
    base += nb;
  }
  release_flags_last_block(done, peers, channel, epoch);
}

// ---------------------------------------------------------------- Cholesky-QR on a tall [len][4] matrix
// One CTA per matrix.  `x` rows are read through `load(i)`; two rounds.  Returns via `store`.
__device__ __forceinline__ void block_reduce10(float (&g)[10], float* sh) {
#pragma unroll
  for (int e = 0; e < 10; ++e) g[e] = warp_sum(g[e]);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
  if (lane == 0)
    for (int e = 0; e < 10; ++e) sh[wid * 10 + e] = g[e];
  __syncthreads();
  if (threadIdx.x < 10) {
    float s = 0.f;
    for (int w = 0; w < nw; ++w) s += sh[w * 10 + threadIdx.x];
    sh[320 + threadIdx.x] = s;
  }
  __syncthreads();
#pragma unroll
  for (int e = 0; e < 10; ++e) g[e] = sh[320 + e];
  __syncthreads();
}

// upper-triangular R with G = R^T R (eps-guarded), then Rinv = R^{-1}; columns >= r are dropped
__device__ void chol_inv4(const float (&g)[10], int r, float (&inv)[4][4]) {
  // g packs the symmetric Gram matrix: (0,0)(0,1)(0,2)(0,3)(1,1)(1,2)(1,3)(2,2)(2,3)(3,3)
  float G[4][4] = {{g[0], g[1], g[2], g[3]}, {g[1], g[4], g[5], g[6]}, {g[2], g[5], g[7], g[8]}, {g[3], g[6], g[8], g[9]}};
  float R[4][4] = {};
  for (int i = 0; i < 4; ++i) {
    if (i >= r) { R[i][i] = 1.f; continue; }
    float d = G[i][i];
    for (int k = 0; k < i; ++k) d -= R[k][i] * R[k][i];
    const float rii = sqrtf(fmaxf(d, 0.f)) + 1e-8f;
    R[i][i] = rii;
    for (int j = i + 1; j < r; ++j) {
      float s = G[i][j];
      for (int k = 0; k < i; ++k) s -= R[k][i] * R[k][j];
      R[i][j] = s / rii;
    }
  }
  for (int j = 0; j < 4; ++j) {                     // back substitution: R * inv = I
    for (int i = 3; i >= 0; --i) {
      float s = (i == j) ? 1.f : 0.f;
      for (int k = i + 1; k < 4; ++k) s -= R[i][k] * inv[k][j];
      inv[i][j] = s / R[i][i];
    }
  }
}

__device__ __forceinline__ float4 apply_inv(const float4 x, const float (&inv)[4][4], int r) {
  const float v[4] = {x.x, x.y, x.z, x.w};
  float o[4] = {0.f, 0.f, 0.f, 0.f};
  for (int j = 0; j < 4; ++j) {
    if (j >= r) break;
    float s = 0.f;
    for (int k = 0; k <= j; ++k) s += v[k] * inv[k][j];
    o[j] = s;
  }
  return make_float4(o[0], o[1], o[2], o[3]);
}

__device__ __forceinline__ void gram_add(float (&g)[10], const float4 x) {
  g[0] += x.x * x.x; g[1] += x.x * x.y; g[2] += x.x * x.z; g[3] += x.x * x.w;
  g[4] += x.y * x.y; g[5] += x.y * x.z; g[6] += x.y * x.w;
  g[7] += x.z * x.z; g[8] += x.z * x.w; g[9] += x.w * x.w;
}

// in-place CholeskyQR2 of buf[len][4] (row-major, first r columns live), by the whole CTA
__device__ void cholqr2_inplace(float* buf, int len, int r, float* sh) {
  for (int round = 0; round < 2; ++round) {
    float g[10] = {};
    for (int i = threadIdx.x; i < len; i += blockDim.x) gram_add(g, *reinterpret_cast<const float4*>(buf + (size_t)i * 4));
    block_reduce10(g, sh);
    float inv[4][4];
    chol_inv4(g, r, inv);
    for (int i = threadIdx.x; i < len; i += blockDim.x) {
      float4* p = reinterpret_cast<float4*>(buf + (size_t)i * 4);
      *p = apply_inv(*p, inv, r);
    }
    __syncthreads();
  }
}

// CholeskyQR2 for row pitches 8 and 16 (ranks 5..16): the Gram matrix is accumulated by RP x RP threads over row chunks
// staged in shared memory, one warp factorises it (same eps-guarded formulas as chol_inv4) and inverts the triangular
// factor, every thread then applies R^-1 to its rows.  sh: at least 128 * RP + 2 * RP * RP floats.
template <int RP>
__device__ void cholqr2_wide(float* buf, int len, int r, float* sh) {
  constexpr int CH = 128;
  float* tile = sh;                 // [CH][RP]
  float* G = sh + CH * RP;          // [RP][RP] Gram matrix, then the upper-triangular factor R in place
  float* INV = G + RP * RP;         // [RP][RP] R^-1
  const int a = threadIdx.x / RP, b = threadIdx.x % RP;
  const bool owner = threadIdx.x < RP * RP;
  for (int round = 0; round < 2; ++round) {
    float accv = 0.f;
    for (int c0 = 0; c0 < len; c0 += CH) {
      const int n = min(CH, len - c0);
      for (int i = threadIdx.x; i < n * RP; i += blockDim.x) tile[i] = buf[(size_t)c0 * RP + i];
      __syncthreads();
      if (owner && a <= b) {
#pragma unroll 4
        for (int i = 0; i < n; ++i) accv += tile[i * RP + a] * tile[i * RP + b];
      }
      __syncthreads();
    }
    if (owner) G[a * RP + b] = accv;            // upper triangle valid
    __syncthreads();
    if (threadIdx.x < 32) {
      const int lane = threadIdx.x;
      for (int i = 0; i < RP; ++i) {
        if (i >= r) {
          if (lane < RP) G[i * RP + lane] = lane == i ? 1.f : 0.f;
          __syncwarp();
          continue;
        }
        float d = G[i * RP + i];
        for (int k = 0; k < i; ++k) d -= G[k * RP + i] * G[k * RP + i];
        const float rii = sqrtf(fmaxf(d, 0.f)) + 1e-8f;
        float sv = 0.f;
        if (lane > i && lane < r) {
          sv = G[i * RP + lane];
          for (int k = 0; k < i; ++k) sv -= G[k * RP + i] * G[k * RP + lane];
          sv /= rii;
        }
        __syncwarp();
        if (lane == i) G[i * RP + i] = rii;
        else if (lane > i && lane < RP) G[i * RP + lane] = lane < r ? sv : 0.f;
        __syncwarp();
      }
      if (lane < RP) {                            // column `lane` of R^-1 by back substitution
        for (int i = RP - 1; i >= 0; --i) {
          float sv = i == lane ? 1.f : 0.f;
          for (int k = i + 1; k < RP; ++k) sv -= G[i * RP + k] * INV[k * RP + lane];
          INV[i * RP + lane] = sv / G[i * RP + i];
        }
      }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < len; i += blockDim.x) {
      float x[RP], o[RP];
      float4* row = reinterpret_cast<float4*>(buf + (size_t)i * RP);
#pragma unroll
      for (int c4 = 0; c4 < RP / 4; ++c4) { const float4 v = row[c4]; x[c4 * 4] = v.x; x[c4 * 4 + 1] = v.y; x[c4 * 4 + 2] = v.z; x[c4 * 4 + 3] = v.w; }
#pragma unroll
      for (int j = 0; j < RP; ++j) {
        float sv = 0.f;
#pragma unroll
        for (int k = 0; k <= j; ++k) sv += x[k] * INV[k * RP + j];
        o[j] = j < r ? sv : 0.f;
      }
#pragma unroll
      for (int c4 = 0; c4 < RP / 4; ++c4) row[c4] = make_float4(o[c4 * 4], o[c4 * 4 + 1], o[c4 * 4 + 2], o[c4 * 4 + 3]);
    }
    __syncthreads();
  }
}

template <int RP>
__device__ __forceinline__ void cholqr2_any(float* buf, int len, int r, float* sh) {
  if (RP == 4) cholqr2_inplace(buf, len, r, sh);
  else cholqr2_wide<RP>(buf, len, r, sh);
}
template <int RP>
struct PsShared { static constexpr int floats = RP == 4 ? 512 : 128 * RP + 2 * RP * RP; };

// Q0 per matrix (one CTA of k3_prep): a fresh Philox normal draw -- or, with a warm start and from the second exchange
// on, the previous exchange's Q (reference: compressor/powersgd.py:37-39 use_memory) -- orthogonalised; writes Q0^T [16][qpad]
template <int RP>
__device__ __forceinline__ void q0_body(const PsMat& mt, float* q0, const float* __restrict__ qprev, float* q0t, int64_t qpad,
                                        uint32_t seed, uint32_t step, float* sh) {
  float* buf = q0 + (size_t)mt.qrow0 * RP;
  const bool warm = qprev != nullptr && step > 0;
  for (int j = threadIdx.x; j < mt.m; j += blockDim.x) {
    float v[RP];
#pragma unroll
    for (int c = 0; c < RP; ++c) v[c] = 0.f;
    if (warm) {
#pragma unroll
      for (int c = 0; c < RP; ++c) v[c] = c < mt.r ? qprev[(size_t)(mt.qrow0 + j) * RP + c] : 0.f;
    } else {
#pragma unroll
      for (int c = 0; c < RP; ++c) {
        if (c < mt.r) {
          // element (j, c) of the (m, r) row-major normal draw: one Box-Muller pair per element index
          const uint32_t e = (uint32_t)(j * mt.r + c);
          const Philox4 p = philox4x32(e, step, GB_STREAM_NORMAL, 0u, seed, mt.hash);
          const float u1 = (float)((p.x >> 8) + 1u) * (1.0f / 16777216.0f);
          const float u2 = (float)(p.y >> 8) * (1.0f / 16777216.0f);
          v[c] = sqrtf(-2.0f * logf(u1)) * cosf(6.283185307179586f * u2);
        }
      }
    }
#pragma unroll
    for (int c = 0; c < RP; ++c) buf[(size_t)j * RP + c] = v[c];
  }
  __syncthreads();
  cholqr2_any<RP>(buf, mt.m, mt.r, sh);
  for (int j = threadIdx.x; j < mt.m; j += blockDim.x) {
#pragma unroll
    for (int c = 0; c < RP; ++c) q0t[(size_t)c * qpad + mt.qrow0 + j] = c < mt.r ? buf[(size_t)j * RP + c] : 0.f;
  }
}

struct PsIn { const float* slot[GB_MAX_WORLD]; int n; };
__device__ __forceinline__ void shift(PsIn& o, size_t bytes) {
  for (int r = 0; r < o.n; ++r) o.slot[r] = (const float*)((const char*)o.slot[r] + bytes);
}

// wait flags; P = mean_r P_r; orthogonalise; write P^ [prows][RP] and P^^T [16][ppad]
template <int RP>
__global__ void __launch_bounds__(512) k3_orth(const PsMat* __restrict__ mats, PsIn in, float inv_world, float* ph, float* pht,
                                               int64_t ppad, PeerPtrs peers, int channel, const StepCtl* ctl, size_t par_stride,
                                               uint32_t* status, uint64_t timeout_ns) {
  const uint32_t epoch = ctl->epoch;
  if (epoch & 1) shift(in, par_stride);
  __shared__ float sh[PsShared<RP>::floats];
  wait_all_ranks(peers, channel, epoch, status, timeout_ns);
  const PsMat mt = mats[blockIdx.x];
  float* buf = ph + (size_t)mt.prow0 * RP;
  for (int i = threadIdx.x; i < mt.n * (RP / 4); i += blockDim.x) {
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = 0; r < in.n; ++r) {
      const float4 x = reinterpret_cast<const float4*>(in.slot[r] + (size_t)mt.prow0 * RP)[i];
      s.x += x.x; s.y += x.y; s.z += x.z; s.w += x.w;
    }
    if (in.n > 1) { s.x *= inv_world; s.y *= inv_world; s.z *= inv_world; s.w *= inv_world; }
    reinterpret_cast<float4*>(buf)[i] = s;
  }
  __syncthreads();
  cholqr2_any<RP>(buf, mt.n, mt.r, sh);
  for (int i = threadIdx.x; i < mt.n; i += blockDim.x) {
#pragma unroll
    for (int c = 0; c < RP; ++c) pht[(size_t)c * ppad + mt.prow0 + i] = c < mt.r ? buf[(size_t)i * RP + c] : 0.f;
  }
}

// wait flags; out = P^ (mean_r Q_r)^T; residual = work - out.  One CTA per (matrix, 32-row band).
template <int RP>
__device__ __forceinline__ void mean_q(const PsIn& in, size_t row, float inv_world, float (&q)[RP]) {
#pragma unroll
  for (int c = 0; c < RP; ++c) q[c] = 0.f;
  for (int r = 0; r < in.n; ++r) {
#pragma unroll
    for (int c4 = 0; c4 < RP / 4; ++c4) {
      const float4 x = reinterpret_cast<const float4*>(in.slot[r] + row * RP)[c4];
      q[c4 * 4] += x.x; q[c4 * 4 + 1] += x.y; q[c4 * 4 + 2] += x.z; q[c4 * 4 + 3] += x.w;
    }
  }
  if (in.n > 1) {
#pragma unroll
    for (int c = 0; c < RP; ++c) q[c] *= inv_world;
  }
}
template <int RP>
__device__ __forceinline__ float dot_r(const float* p, const float (&q)[RP]) {
  float s = 0.f;
#pragma unroll
  for (int c = 0; c < RP; ++c) s += p[c] * q[c];
  return s;
}

// ---------------------------------------------------------------- band kernels: prep, final, residual read-out
// One CTA per (matrix, 32-row band).  A thread owns four consecutive columns (16-byte accesses) and every `rl`-th row
// of the band, eight rows at a time with all loads issued before the dependent stores; narrow matrices spread their
// 32 rows over the threads the columns leave idle.
//
// The error-feedback memory is IMPLICIT: `work` keeps the compensated gradient M of the last exchange, `ph` and
// `qkeep` its P^ and averaged Q, and the residual  e = M - P^ Q^T  (memory/powersgd.py:24-37) is recomputed where it
// is consumed -- in the next exchange's prep:  M' = grad + (M - P^ Q^T)  -- by the same dot product that produced the
// output, so the bits are those a stored residual would have had.  A write-only k3_final (102 MB instead of reading
// 102 MB and writing 204 MB for ResNet-50; streaming writes run at about half the read bandwidth here) and no
// residual buffer at all are what that buys.
//   MODE 0 (prep)   work = grad [+ work - P^ Q^T]          Q from qkeep
//   MODE 1 (final)  out  = P^ Q^T, qkeep = Q               Q = mean over the ranks' slots
//   MODE 2 (resid)  out  = work - P^ Q^T                   Q from qkeep (state read-out)
//   MODE 3 (fold)   work = work - P^ Q^T                   the residual made explicit in place
// grad / out are flat (row pitch m); `work` is private to the plan and pads its rows to 16 bytes (PsMat::ldw).
template <int RP, int MODE>
__device__ __forceinline__ void band_body(const PsMat& mt, const PsBand& b, const float* __restrict__ ph, const float* qsrc,
                                          const PsIn& in, float inv_world, const float* __restrict__ grad, float* work,
                                          float* out, float* qkeep, bool use_prev, float* sp) {
  const int nrow = min(32, mt.n - b.i0);
  const bool need_pq = MODE != 0 || use_prev;
  if (need_pq)
    for (int i = threadIdx.x; i < nrow * RP; i += blockDim.x) sp[i] = ph[(size_t)(mt.prow0 + b.i0) * RP + i];
  __syncthreads();
  const size_t band0 = mt.off + (size_t)b.i0 * mt.m;            // in grad / out
  const size_t wband0 = mt.woff + (size_t)b.i0 * mt.ldw;        // in work
  const bool vec = (mt.m & 3) == 0;                             // (then j0 and j1 are multiples of 4 as well)
  const int ncg = vec ? (b.j1 - b.j0) >> 2 : b.j1 - b.j0;       // column groups: 4 columns (or 1 on the scalar path)
  const int cgs = min(ncg, (int)blockDim.x), rl = (int)blockDim.x / cgs;
  const int my_cg = threadIdx.x % cgs, my_r = threadIdx.x / cgs;
  if (my_r >= rl) return;
  const bool keep = MODE == 1 && qkeep != nullptr && b.i0 == 0 && my_r == 0;
  if (vec) {
    for (int cg = my_cg; cg < ncg; cg += cgs) {
      const int j = b.j0 + cg * 4;
      float q[4][RP];
      if (need_pq) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          if (MODE == 1) {
            mean_q<RP>(in, (size_t)(mt.qrow0 + j + c), inv_world, q[c]);
          } else {
#pragma unroll
            for (int c4 = 0; c4 < RP / 4; ++c4) {
              const float4 x = reinterpret_cast<const float4*>(qsrc + (size_t)(mt.qrow0 + j + c) * RP)[c4];
              q[c][c4 * 4] = x.x; q[c][c4 * 4 + 1] = x.y; q[c][c4 * 4 + 2] = x.z; q[c][c4 * 4 + 3] = x.w;
            }
          }
          if (keep) {
#pragma unroll
            for (int c4 = 0; c4 < RP / 4; ++c4)
              reinterpret_cast<float4*>(qkeep + (size_t)(mt.qrow0 + j + c) * RP)[c4] =
                  make_float4(q[c][c4 * 4], q[c][c4 * 4 + 1], q[c][c4 * 4 + 2], q[c][c4 * 4 + 3]);
          }
        }
      }
      constexpr int U = MODE == 0 ? 4 : 8;       // rows in flight per thread (prep loads two streams: same bytes in flight)
      for (int i0 = my_r; i0 < nrow; i0 += U * rl) {
        float4 w[U], g[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int i = i0 + u * rl;
          if (i < nrow) {
            const size_t e = band0 + (size_t)i * mt.m + j, ew = wband0 + (size_t)i * mt.ldw + j;
            if (MODE == 0) g[u] = *reinterpret_cast<const float4*>(grad + e);
            if (MODE >= 2 || (MODE == 0 && use_prev)) w[u] = *reinterpret_cast<const float4*>(work + ew);
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int i = i0 + u * rl;
          if (i < nrow) {
            const size_t e = band0 + (size_t)i * mt.m + j, ew = wband0 + (size_t)i * mt.ldw + j;
            float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
            if (need_pq) {
              const float* pr = sp + i * RP;
              o = make_float4(dot_r<RP>(pr, q[0]), dot_r<RP>(pr, q[1]), dot_r<RP>(pr, q[2]), dot_r<RP>(pr, q[3]));
            }
            if (MODE == 1) {
              *reinterpret_cast<float4*>(out + e) = o;
            } else if (MODE == 2) {
              *reinterpret_cast<float4*>(out + e) = make_float4(w[u].x - o.x, w[u].y - o.y, w[u].z - o.z, w[u].w - o.w);
            } else if (MODE == 3) {
              *reinterpret_cast<float4*>(work + ew) = make_float4(w[u].x - o.x, w[u].y - o.y, w[u].z - o.z, w[u].w - o.w);
            } else if (use_prev) {
              *reinterpret_cast<float4*>(work + ew) =
                  make_float4(__fadd_rn(g[u].x, w[u].x - o.x), __fadd_rn(g[u].y, w[u].y - o.y), __fadd_rn(g[u].z, w[u].z - o.z),
                              __fadd_rn(g[u].w, w[u].w - o.w));
            } else {
              *reinterpret_cast<float4*>(work + ew) = g[u];
            }
          }
        }
      }
    }
  } else {
    for (int jj = my_cg; jj < ncg; jj += cgs) {
      const int j = b.j0 + jj;
      float q[RP];
#pragma unroll
      for (int c = 0; c < RP; ++c) q[c] = 0.f;
      if (need_pq) {
        if (MODE == 1) {
          mean_q<RP>(in, (size_t)(mt.qrow0 + j), inv_world, q);
        } else {
#pragma unroll
          for (int c = 0; c < RP; ++c) q[c] = qsrc[(size_t)(mt.qrow0 + j) * RP + c];
        }
        if (keep) {
#pragma unroll
          for (int c = 0; c < RP; ++c) qkeep[(size_t)(mt.qrow0 + j) * RP + c] = q[c];
        }
      }
      for (int i = my_r; i < nrow; i += rl) {
        const size_t e = band0 + (size_t)i * mt.m + j, ew = wband0 + (size_t)i * mt.ldw + j;
        const float o = need_pq ? dot_r<RP>(sp + i * RP, q) : 0.f;
        if (MODE == 1) out[e] = o;
        else if (MODE == 2) out[e] = work[ew] - o;
        else if (MODE == 3) work[ew] = work[ew] - o;
        else if (use_prev) work[ew] = __fadd_rn(grad[e], work[ew] - o);
        else work[ew] = grad[e];
      }
    }
  }
}

#define PS_BAND_THREADS 512

// phase 1 prologue, one launch: CTAs [0, nmat) draw and orthogonalise Q0 (latency-bound, one matrix each), the band
// CTAs behind them stream  work = grad + residual  and clear the split-K tile counters
template <int RP>
__global__ void __launch_bounds__(PS_BAND_THREADS, RP == 16 ? 1 : 2) k3_prep(const PsBand* __restrict__ bands, const PsMat* __restrict__ mats, int nmat,
                                                           const float* __restrict__ grad, float* work, const float* __restrict__ ph,
                                                           float* qkeep, int has_mem, int warm, float* q0, float* q0t, int64_t qpad,
                                                           uint32_t seed, uint32_t* cnt, int ncnt, const StepCtl* ctl) {
  __shared__ __align__(16) float sh[PsShared<RP>::floats > 32 * RP ? PsShared<RP>::floats : 32 * RP];
  const uint32_t step = ctl->step;
  if ((int)blockIdx.x < nmat) {
    q0_body<RP>(mats[blockIdx.x], q0, warm ? qkeep : nullptr, q0t, qpad, seed, step, sh);
    return;
  }
  const int bi = (int)blockIdx.x - nmat;
  for (int i = bi * (int)blockDim.x + (int)threadIdx.x; i < ncnt; i += ((int)gridDim.x - nmat) * (int)blockDim.x) cnt[i] = 0u;
  const PsBand b = bands[bi];
  const PsMat mt = mats[b.mat];
  PsIn none; none.n = 0;
  band_body<RP, 0>(mt, b, ph, qkeep, none, 1.f, grad, work, nullptr, nullptr, has_mem && step > 0, sh);
}

// phase 3: wait flags; Q = mean_r Q_r; out = P^ Q^T (write only; the band that starts at row 0 keeps Q)
template <int RP>
__global__ void __launch_bounds__(PS_BAND_THREADS) k3_final(const PsBand* __restrict__ bands, const PsMat* __restrict__ mats, PsIn in,
                                                            float inv_world, const float* __restrict__ ph, float* __restrict__ out,
                                                            float* qkeep, PeerPtrs peers, int channel, const StepCtl* ctl,
                                                            size_t par_stride, uint32_t* status, uint64_t timeout_ns) {
  const uint32_t epoch = ctl->epoch;
  if (epoch & 1) shift(in, par_stride);
  __shared__ __align__(16) float sp[32 * RP];
  wait_all_ranks(peers, channel, epoch, status, timeout_ns);
  const PsBand b = bands[blockIdx.x];
  const PsMat mt = mats[b.mat];
  band_body<RP, 1>(mt, b, ph, nullptr, in, inv_world, nullptr, nullptr, out, qkeep, true, sp);
}

// state read-out: dst = work - P^ Q^T, the residual a PowerSGDMemory would hold after the last exchange (FOLD: in place)
template <int RP, int FOLD>
__global__ void __launch_bounds__(PS_BAND_THREADS) k3_resid(const PsBand* __restrict__ bands, const PsMat* __restrict__ mats,
                                                            const float* __restrict__ ph, const float* __restrict__ qkeep,
                                                            float* work, float* dst) {
  __shared__ __align__(16) float sp[32 * RP];
  const PsBand b = bands[blockIdx.x];
  const PsMat mt = mats[b.mat];
  PsIn none; none.n = 0;
  band_body<RP, FOLD ? 3 : 2>(mt, b, ph, qkeep, none, 1.f, nullptr, work, dst, nullptr, true, sp);
}

// state write: work = src (flat layout -> padded rows)
__global__ void __launch_bounds__(PS_BAND_THREADS) k3_load(const PsBand* __restrict__ bands, const PsMat* __restrict__ mats,
                                                           const float* __restrict__ src, float* work) {
  __shared__ __align__(16) float sp[32 * 4];
  const PsBand b = bands[blockIdx.x];
  const PsMat mt = mats[b.mat];
  PsIn none; none.n = 0;
  band_body<4, 0>(mt, b, nullptr, nullptr, none, 1.f, src, work, nullptr, nullptr, false, sp);
}

// ---------------------------------------------------------------- host side
#define LAUNCHED(p) do { (p)->launches++; GB_CUDA(cudaGetLastError()); } while (0)
// cuTensorMapEncodeTiled is fetched through the runtime (cudaGetDriverEntryPoint) so that the library
// has no link-time dependency on libcuda.so and can be loaded on a machine without a driver.
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)sym;
  }
  return fn;
}

static int make_map(CUtensorMap* out, const float* base, uint64_t inner, uint64_t outer, uint64_t pitch_elems, uint32_t box_inner,
                    uint32_t box_outer, CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B) {
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {pitch_elems * 4};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  EncodeTiledFn enc = encode_tiled_fn();
  GB_CHECK(enc != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
  const CUresult rc = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)base, dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  GB_CHECK(rc == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d (shape %llu x %llu, pitch %llu)", (int)rc,
           (unsigned long long)outer, (unsigned long long)inner, (unsigned long long)pitch_elems);
  return 0;
}

static void build_items(const std::vector<PsMat>& mats, int mode, std::vector<PsItem>& items, int& ntiles) {
  // aim for >= ~3 CTAs per SM overall; never split K finer than 8 pipeline stages
  int64_t total = 0;
  for (auto& m : mats) if (m.tc) total += (int64_t)m.n * m.m;
  // two CTAs are resident per SM (2 x 110 KB of pipeline stages): one streams while the other is in its prologue or
  // epilogue; four items per SM = two full waves
  const char* ips = getenv("GB_PSGD_ITEMS_PER_SM");
  const int items_per_sm = ips ? std::max(1, atoi(ips)) : 4;
  int64_t per_item = std::max<int64_t>((int64_t)PS_BM * PS_BK * 8, total / (148 * items_per_sm) + 1);
  if (getenv("GB_PSGD_NOSPLIT")) per_item = int64_t(1) << 40;      // debugging: one K chunk per tile
  ntiles = 0;
  for (int mi = 0; mi < (int)mats.size(); ++mi) {
    const PsMat& m = mats[mi];
    if (!m.tc) continue;
    const int mdim = mode == 1 ? m.n : m.m, kdim = mode == 1 ? m.m : m.n;
    const int mtiles = (mdim + PS_BM - 1) / PS_BM;
    int chunk = (int)std::min<int64_t>(kdim, std::max<int64_t>(PS_BK * 8, per_item / PS_BM));
    chunk = (chunk + PS_BK - 1) / PS_BK * PS_BK;
    const int nch = (kdim + chunk - 1) / chunk;
    for (int t = 0; t < mtiles; ++t) {
      for (int c = 0; c < nch; ++c)
        items.push_back(PsItem{mi, ntiles, t * PS_BM, c * chunk, std::min(kdim, (c + 1) * chunk), nch, c});
      ++ntiles;
    }
  }
}

extern "C" int gb_psgd_destroy(gb_psgd* p);

static int psgd_build(gb_psgd* p, gb_comm* c, int nmat, const int64_t* rows, const int64_t* cols, const uint32_t* name_hash,
                      int rank_r, uint32_t seed, int has_memory, int use_tcgen05, int warm_start);

extern "C" int gb_psgd_create2(gb_comm* c, int nmat, const int64_t* rows, const int64_t* cols, const uint32_t* name_hash, int rank_r,
                               uint32_t seed, int has_memory, int use_tcgen05, int warm_start, gb_psgd** out) {
  GB_CHECK(nmat >= 1, "need at least one matrix");
  GB_CHECK(rank_r >= 1 && rank_r <= PS_RMAX, "native PowerSGD supports rank 1..%d (got %d)", PS_RMAX, rank_r);
  gb_psgd* p = new gb_psgd();            // value-initialised: every pointer starts as nullptr
  if (psgd_build(p, c, nmat, rows, cols, name_hash, rank_r, seed, has_memory, use_tcgen05, warm_start)) {
    gb_psgd_destroy(p);                  // one failure exit: releases whatever the builder had allocated
    return -1;
  }
  *out = p;
  return 0;
}
extern "C" int gb_psgd_create(gb_comm* c, int nmat, const int64_t* rows, const int64_t* cols, const uint32_t* name_hash, int rank_r,
                              uint32_t seed, int has_memory, int use_tcgen05, gb_psgd** out) {
  return gb_psgd_create2(c, nmat, rows, cols, name_hash, rank_r, seed, has_memory, use_tcgen05, 0, out);
}

static int psgd_build(gb_psgd* p, gb_comm* c, int nmat, const int64_t* rows, const int64_t* cols, const uint32_t* name_hash,
                      int rank_r, uint32_t seed, int has_memory, int use_tcgen05, int warm_start) {
  p->rp = rank_r <= 4 ? 4 : rank_r <= 8 ? 8 : 16;
  p->d_qprev = nullptr; p->d_simt = nullptr; p->warm = warm_start ? 1 : 0;
  p->comm = c; p->nmat = nmat; p->rank_r = rank_r; p->has_mem = has_memory; p->use_tc = use_tcgen05;
  p->seed = seed; p->step = 0; p->epoch = 0; p->launches = 0;
  int64_t off = 0, woff = 0, prow = 0, qrow = 0;
  for (int i = 0; i < nmat; ++i) {
    GB_CHECK(rows[i] >= 1 && cols[i] >= 1 && rows[i] * cols[i] < (int64_t(1) << 31), "matrix %d has a bad shape", i);
    PsMat m;
    m.off = off; m.n = (int)rows[i]; m.m = (int)cols[i];
    m.ldw = (m.m + 3) / 4 * 4; m.woff = woff;
    m.r = (int)std::min<int64_t>(rank_r, std::min(rows[i], cols[i]));
    m.prow0 = (int)prow; m.qrow0 = (int)qrow; m.hash = name_hash[i];
    m.tc = use_tcgen05 && m.n >= 8 && m.m >= 32;      // smaller than a pipeline stage: CUDA-core fallback
    p->h_mat.push_back(m);
    p->h_off.push_back(off);
    off += (rows[i] * cols[i] + GB_ALIGN - 1) / GB_ALIGN * GB_ALIGN;
    woff += (rows[i] * m.ldw + GB_ALIGN - 1) / GB_ALIGN * GB_ALIGN;
    prow += (rows[i] + 3) / 4 * 4;        // keep every matrix's row window 16-byte aligned in the transposed buffers
    qrow += (cols[i] + 3) / 4 * 4;
  }
  p->padded = off; p->wpadded = woff; p->prows = prow; p->qrows = qrow;
  p->ppad = (prow + 31) / 32 * 32; p->qpad = (qrow + 31) / 32 * 32;
  std::vector<PsItem> it1, it2;
  build_items(p->h_mat, 1, it1, p->ntiles1);
  build_items(p->h_mat, 2, it2, p->ntiles2);
  p->nitems1 = (int)it1.size(); p->nitems2 = (int)it2.size();
  // band items: 32 rows x at most 2048 columns (one 4-column group per thread), largest first so that the tail of the
  // launch is made of small CTAs (the matrices arrive in network order: the widest ones last)
  std::vector<PsBand> bands;
  for (int i = 0; i < nmat; ++i) {
    const int m = p->h_mat[i].m;
    const int nchunk = (m + 4 * PS_BAND_THREADS - 1) / (4 * PS_BAND_THREADS);
    const int chunk = ((m + nchunk - 1) / nchunk + 3) / 4 * 4;
    for (int r0 = 0; r0 < p->h_mat[i].n; r0 += 32)
      for (int j0 = 0; j0 < m; j0 += chunk) bands.push_back(PsBand{i, r0, j0, std::min(m, j0 + chunk)});
  }
  std::stable_sort(bands.begin(), bands.end(), [&](const PsBand& a, const PsBand& b) {
    const int64_t sa = (int64_t)std::min(32, p->h_mat[a.mat].n - a.i0) * (a.j1 - a.j0);
    const int64_t sb = (int64_t)std::min(32, p->h_mat[b.mat].n - b.i0) * (b.j1 - b.j0);
    return sa > sb;
  });

  GB_CUDA(cudaMalloc((void**)&p->d_mat, sizeof(PsMat) * nmat));
  GB_CUDA(cudaMemcpy(p->d_mat, p->h_mat.data(), sizeof(PsMat) * nmat, cudaMemcpyHostToDevice));
  GB_CUDA(cudaMalloc((void**)&p->d_items1, sizeof(PsItem) * (it1.size() + 1)));
  GB_CUDA(cudaMalloc((void**)&p->d_items2, sizeof(PsItem) * (it2.size() + 1)));
  if (!it1.empty()) GB_CUDA(cudaMemcpy(p->d_items1, it1.data(), sizeof(PsItem) * it1.size(), cudaMemcpyHostToDevice));
  if (!it2.empty()) GB_CUDA(cudaMemcpy(p->d_items2, it2.data(), sizeof(PsItem) * it2.size(), cudaMemcpyHostToDevice));
  GB_CUDA(cudaMalloc((void**)&p->d_work, p->wpadded * 4));
  GB_CUDA(cudaMemset(p->d_work, 0, p->wpadded * 4));
  GB_CUDA(cudaMalloc((void**)&p->d_q0t, PS_N * p->qpad * 4)); GB_CUDA(cudaMemset(p->d_q0t, 0, PS_N * p->qpad * 4));
  GB_CUDA(cudaMalloc((void**)&p->d_pht, PS_N * p->ppad * 4)); GB_CUDA(cudaMemset(p->d_pht, 0, PS_N * p->ppad * 4));
  const size_t RP = (size_t)p->rp;
  GB_CUDA(cudaMalloc((void**)&p->d_pacc, (p->prows + 8) * RP * 4));
  GB_CUDA(cudaMalloc((void**)&p->d_qacc, (p->qrows + 8) * RP * 4 * 2));   // second half: Q0 scratch for k3_q0
  GB_CUDA(cudaMalloc((void**)&p->d_ph, (p->prows + 8) * RP * 4));
  GB_CUDA(cudaMemset(p->d_ph, 0, (p->prows + 8) * RP * 4));          // P^ = 0: no residual before the first exchange
  GB_CUDA(cudaMalloc((void**)&p->d_part1, (it1.size() + 1) * PS_BM * RP * 4));
  GB_CUDA(cudaMalloc((void**)&p->d_part2, (it2.size() + 1) * PS_BM * RP * 4));
  GB_CUDA(cudaMalloc((void**)&p->d_qprev, (p->qrows + 8) * RP * 4));
  GB_CUDA(cudaMemset(p->d_qprev, 0, (p->qrows + 8) * RP * 4));
  {
    std::vector<int> simt;
    p->simt_rows = p->simt_cblocks = 0;
    for (int i = 0; i < nmat; ++i)
      if (!p->h_mat[i].tc) { simt.push_back(i); p->simt_rows += p->h_mat[i].n; p->simt_cblocks += (p->h_mat[i].m + 31) / 32; }
    p->nsimt = (int)simt.size();
    GB_CUDA(cudaMalloc((void**)&p->d_simt, sizeof(int) * (simt.size() + 1)));
    if (!simt.empty()) GB_CUDA(cudaMemcpy(p->d_simt, simt.data(), sizeof(int) * simt.size(), cudaMemcpyHostToDevice));
  }
  GB_CUDA(cudaMalloc((void**)&p->d_cnt1, (p->ntiles1 + p->ntiles2 + 64) * 4));
  p->d_cnt2 = p->d_cnt1 + p->ntiles1;
  p->d_done = p->d_cnt1 + p->ntiles1 + p->ntiles2;
  p->d_ctl = (StepCtl*)(p->d_done + 16);
  GB_CUDA(cudaMemset(p->d_cnt1, 0, (p->ntiles1 + p->ntiles2 + 64) * 4));
  p->nbands = (int)bands.size();
  GB_CUDA(cudaMalloc((void**)&p->d_bands, sizeof(PsBand) * bands.size()));
  GB_CUDA(cudaMemcpy(p->d_bands, bands.data(), sizeof(PsBand) * bands.size(), cudaMemcpyHostToDevice));
  p->d_maps = nullptr;
  // tensor maps
  std::vector<CUtensorMap> maps(nmat * 4);
  for (int i = 0; i < nmat; ++i) {
    const PsMat& m = p->h_mat[i];
    if (!m.tc) continue;
    const float* M = p->d_work + m.woff;
    if (make_map(&maps[i * 4 + 0], M, m.m, m.n, m.ldw, PS_BOXK, PS_BM)) return -1;
    if (make_map(&maps[i * 4 + 1], p->d_q0t + m.qrow0, m.m, PS_N, p->qpad, PS_BOXK, PS_N)) return -1;
    if (make_map(&maps[i * 4 + 2], M, m.m, m.n, m.ldw, 32, PS_BOXK,
                 getenv("GB_PSGD_SWZ128") ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B)) return -1;
    if (make_map(&maps[i * 4 + 3], p->d_pht + m.prow0, m.n, PS_N, p->ppad, PS_BOXK, PS_N)) return -1;
  }
  GB_CUDA(cudaMalloc((void**)&p->d_maps, sizeof(CUtensorMap) * nmat * 4));
  GB_CUDA(cudaMemcpy(p->d_maps, maps.data(), sizeof(CUtensorMap) * nmat * 4, cudaMemcpyHostToDevice));
  GB_CUDA(cudaFuncSetAttribute(k3_gemm<1, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, PS_SMEM));
  GB_CUDA(cudaFuncSetAttribute(k3_gemm<2, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, PS_SMEM));
  GB_CUDA(cudaFuncSetAttribute(k3_gemm<1, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, PS_SMEM));
  GB_CUDA(cudaFuncSetAttribute(k3_gemm<2, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, PS_SMEM));
  GB_CUDA(cudaFuncSetAttribute(k3_gemm<1, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, PS_SMEM));
  GB_CUDA(cudaFuncSetAttribute(k3_gemm<2, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, PS_SMEM));
  p->slot_p = gb_guarded(c, (size_t)(p->prows + 8) * RP * 4);
  p->slot_q = gb_guarded(c, (size_t)(p->qrows + 8) * RP * 4);
  // both channels in one request, so that a retry after the arena has grown does not leave a half-allocated pair behind
  {
    size_t need2 = (p->slot_p + p->slot_q) * 2 * (size_t)c->world;
    if (gb_comm_arena_free(c) < need2) {
      int dummy; size_t off;
      return gb_comm_alloc_channel(c, p->slot_p + p->slot_q, &dummy, &off);     // fails and reports the bytes needed for both
    }
  }
  if (gb_comm_alloc_channel(c, p->slot_p, &p->chan_p, &p->off_p)) return -1;
  if (gb_comm_alloc_channel(c, p->slot_q, &p->chan_q, &p->off_q)) return -1;
  GB_CUDA(cudaDeviceSynchronize());
  return 0;
}

extern "C" int gb_psgd_destroy(gb_psgd* p) {
  if (!p) return 0;
  cudaDeviceSynchronize();
  cudaFree(p->d_mat); cudaFree(p->d_items1); cudaFree(p->d_items2); cudaFree(p->d_work); cudaFree(p->d_simt);
  cudaFree(p->d_q0t); cudaFree(p->d_pht); cudaFree(p->d_pacc); cudaFree(p->d_qacc); cudaFree(p->d_ph); cudaFree(p->d_cnt1);
  cudaFree(p->d_maps);
  cudaFree(p->d_bands);
  cudaFree(p->d_qprev);
  cudaFree(p->d_part1);
  cudaFree(p->d_part2);
  delete p;
  return 0;
}

extern "C" int64_t gb_psgd_padded_numel(gb_psgd* p) { return p->padded; }
extern "C" int gb_psgd_offsets(gb_psgd* p, int64_t* out) { for (int i = 0; i < p->nmat; ++i) out[i] = p->h_off[i]; return 0; }
extern "C" int gb_psgd_launches(gb_psgd* p, int64_t* out) { *out = p->launches; return 0; }
extern "C" int gb_psgd_tc_matrices(gb_psgd* p) { int n = 0; for (auto& m : p->h_mat) n += m.tc; return n; }
extern "C" int gb_psgd_get_step(gb_psgd* p, uint32_t* s) {
  GB_CUDA(cudaMemcpy(s, &p->d_ctl->next_step, 4, cudaMemcpyDeviceToHost));
  p->step = *s;
  return 0;
}
extern "C" int gb_psgd_set_step(gb_psgd* p, uint32_t s) {
  p->step = s;
  GB_CUDA(cudaMemcpy(&p->d_ctl->next_step, &s, 4, cudaMemcpyHostToDevice));
  return 0;
}
// non-NULL when the plan keeps error-feedback state.  The residual itself is implicit (work - P^ Q^T, see the band
// kernels): read it with gb_psgd_state_read, which materialises it into the caller's buffer.
extern "C" int gb_psgd_residual(gb_psgd* p, void** ptr, size_t* bytes) {
  *ptr = p->has_mem ? p->d_work : nullptr;
  *bytes = p->has_mem ? (size_t)p->padded * 4 : 0;
  return 0;
}
// debugging / tests: copy an internal buffer out.  what: 0 P accumulator, 1 Q accumulator, 2 Q0^T, 3 P^^T, 4 P^, 5 work
extern "C" int gb_psgd_peek(gb_psgd* p, int what, float* d_dst, int64_t nfloats, void* stream) {
  const float* src = what == 0 ? p->d_pacc : what == 1 ? p->d_qacc : what == 2 ? p->d_q0t : what == 3 ? p->d_pht : what == 4 ? p->d_ph : p->d_work;
  GB_CUDA(cudaMemcpyAsync(d_dst, src, (size_t)nfloats * 4, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return 0;
}
#define PS_DISPATCH(rp, CALL)                    \
  do {                                           \
    if ((rp) == 4) { constexpr int RP = 4; CALL; }        \
    else if ((rp) == 8) { constexpr int RP = 8; CALL; }   \
    else { constexpr int RP = 16; CALL; }                 \
  } while (0)

extern "C" int gb_psgd_state_read(gb_psgd* p, void* d_dst, void* stream) {
  GB_CHECK(p->has_mem, "this PowerSGD plan has no residual memory");
  cudaStream_t st = (cudaStream_t)stream;
  GB_CUDA(cudaMemsetAsync(d_dst, 0, (size_t)p->padded * 4, st));          // alignment gaps between the matrices
  PS_DISPATCH(p->rp, (k3_resid<RP, 0><<<p->nbands, PS_BAND_THREADS, 0, st>>>(p->d_bands, p->d_mat, p->d_ph, p->d_qprev, p->d_work, (float*)d_dst)));
  GB_CUDA(cudaGetLastError());
  return 0;
}
// makes the implicit residual explicit in `work` (work -= P^ Q^T, P^ = 0) so that Q may be replaced underneath it
static int psgd_fold_residual(gb_psgd* p, cudaStream_t st) {
  PS_DISPATCH(p->rp, (k3_resid<RP, 1><<<p->nbands, PS_BAND_THREADS, 0, st>>>(p->d_bands, p->d_mat, p->d_ph, p->d_qprev, p->d_work, nullptr)));
  GB_CUDA(cudaGetLastError());
  GB_CUDA(cudaMemsetAsync(p->d_ph, 0, (size_t)(p->prows + 8) * p->rp * 4, st));
  return 0;
}
extern "C" int gb_psgd_state_write(gb_psgd* p, const void* d_src, void* stream) {
  GB_CHECK(p->has_mem, "this PowerSGD plan has no residual memory");
  cudaStream_t st = (cudaStream_t)stream;
  k3_load<<<p->nbands, PS_BAND_THREADS, 0, st>>>(p->d_bands, p->d_mat, (const float*)d_src, p->d_work);
  GB_CUDA(cudaGetLastError());
  GB_CUDA(cudaMemsetAsync(p->d_ph, 0, (size_t)(p->prows + 8) * p->rp * 4, st));      // residual = work - 0
  return 0;
}
// warm-start Q (state index 1 of the Python plan): copy out / in, [qrows + 8][rp] floats
extern "C" int gb_psgd_qprev(gb_psgd* p, void** ptr, size_t* bytes) {
  *ptr = p->warm ? p->d_qprev : nullptr;
  *bytes = p->warm ? (size_t)(p->qrows + 8) * p->rp * 4 : 0;
  return 0;
}
extern "C" int gb_psgd_qprev_copy(gb_psgd* p, void* d_buf, int to_plan, void* stream) {
  GB_CHECK(p->warm, "this PowerSGD plan has no warm-start memory");
  const size_t bytes = (size_t)(p->qrows + 8) * p->rp * 4;
  if (to_plan && p->has_mem && psgd_fold_residual(p, (cudaStream_t)stream)) return -1;   // the residual refers to the old Q
  GB_CUDA(cudaMemcpyAsync(to_plan ? (void*)p->d_qprev : d_buf, to_plan ? d_buf : (void*)p->d_qprev, bytes, cudaMemcpyDeviceToDevice,
                          (cudaStream_t)stream));
  return 0;
}
extern "C" int64_t gb_psgd_wire_bytes(gb_psgd* p) { return (int64_t)(p->prows + p->qrows) * p->rp * 4; }

static PsPeers peer_slots(gb_psgd* p, size_t chan_off, size_t slot_bytes, uint32_t epoch = 0) {   // parity 0
  PsPeers s; gb_comm* c = p->comm; s.n = c->world; s.mc = 0;
  for (int r = 0; r < c->world; ++r) s.slot[r] = (float*)gb_slot_of(c, r, chan_off, slot_bytes, epoch & 1, c->rank);
  if (char* mc = gb_slot_mc(c, chan_off, slot_bytes, epoch & 1, c->rank)) { s.n = 1; s.mc = 1; s.slot[0] = (float*)mc; }
  return s;
}
static PsIn my_slots(gb_psgd* p, size_t chan_off, size_t slot_bytes, uint32_t epoch = 0) {        // parity 0
  PsIn s; gb_comm* c = p->comm; s.n = c->world;
  for (int r = 0; r < c->world; ++r) s.slot[r] = (const float*)gb_slot_of(c, c->rank, chan_off, slot_bytes, epoch & 1, r);
  return s;
}

// phase 1: compensate + Q0 (one launch), P_r = M Q0 with the P tiles pushed from the epilogue; the last CTA of the
// last launch releases the flags
extern "C" int gb_psgd_send(gb_psgd* p, const float* d_grad, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  gb_comm* c = p->comm;
  GB_CHECK(((uintptr_t)d_grad & 15) == 0, "gradient buffer must be 16-byte aligned");
  ++p->epoch;
  gb_launch_tick(p->d_ctl, st);
  const size_t psp = (size_t)c->world * p->slot_p;
  PS_DISPATCH(p->rp, (k3_prep<RP><<<p->nmat + p->nbands, PS_BAND_THREADS, 0, st>>>(
                         p->d_bands, p->d_mat, p->nmat, d_grad, p->d_work, p->d_ph, p->d_qprev, p->has_mem, p->warm,
                         p->d_qacc + (size_t)(p->qrows + 8) * RP, p->d_q0t, p->qpad, p->seed, p->d_cnt1, p->ntiles1 + p->ntiles2,
                         p->d_ctl)));
  LAUNCHED(p);
  const PsPeers ps = peer_slots(p, p->off_p, p->slot_p);
  if (p->nitems1) {
    PS_DISPATCH(p->rp, (k3_gemm<1, RP><<<std::min(p->nitems1, PS_GRID), 192, PS_SMEM, st>>>(p->d_items1, p->nitems1, p->d_mat, p->d_maps, p->d_part1, p->d_cnt1, ps,
                                                                        p->d_ctl, psp, p->d_done, gb_comm_peers(c), p->chan_p,
                                                                        p->nsimt == 0)));
    LAUNCHED(p);
  }
  if (p->nsimt) {
    PS_DISPATCH(p->rp, (k3_gemm1_simt<RP><<<(p->simt_rows * 32 + GB_THREADS - 1) / GB_THREADS, GB_THREADS, 0, st>>>(
                           p->d_mat, p->d_simt, p->nsimt, p->d_work, p->d_q0t, p->qpad, ps, p->d_done, gb_comm_peers(c), p->chan_p,
                           p->d_ctl, psp)));
    LAUNCHED(p);
  }
  return 0;
}

// one-warp wait for every rank's P tiles (which = 0) or Q tiles (which = 1) of the current exchange (see gb_plan_wait)
extern "C" int gb_psgd_wait(gb_psgd* p, int which, void* stream) {
  if (p->comm->world == 1) return 0;
  gb_launch_wait(p->comm, which == 0 ? p->chan_p : p->chan_q, p->d_ctl, (cudaStream_t)stream);
  LAUNCHED(p);
  return 0;
}

// phase 2: reduce + orthogonalise P, Q_r = M^T P^, push Q tiles, release flags
extern "C" int gb_psgd_mid(gb_psgd* p, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  gb_comm* c = p->comm;
  const size_t psp = (size_t)c->world * p->slot_p, psq = (size_t)c->world * p->slot_q;
  PS_DISPATCH(p->rp, (k3_orth<RP><<<p->nmat, 512, 0, st>>>(p->d_mat, my_slots(p, p->off_p, p->slot_p), 1.0f / (float)c->world, p->d_ph,
                                                         p->d_pht, p->ppad, gb_comm_peers(c), p->chan_p, p->d_ctl, psp,
                                                         c->status_dev, c->timeout_ns)));
  LAUNCHED(p);
  const PsPeers qs = peer_slots(p, p->off_q, p->slot_q);
  if (p->nitems2) {
    PS_DISPATCH(p->rp, (k3_gemm<2, RP><<<std::min(p->nitems2, PS_GRID), 192, PS_SMEM, st>>>(p->d_items2, p->nitems2, p->d_mat, p->d_maps, p->d_part2, p->d_cnt2, qs,
                                                                        p->d_ctl, psq, p->d_done, gb_comm_peers(c), p->chan_q,
                                                                        p->nsimt == 0)));
    LAUNCHED(p);
  }
  if (p->nsimt) {
    PS_DISPATCH(p->rp, (k3_gemm2_simt<RP><<<p->simt_cblocks, GB_THREADS, 0, st>>>(p->d_mat, p->d_simt, p->nsimt, p->d_work, p->d_ph, qs,
                                                                                 p->d_done, gb_comm_peers(c), p->chan_q, p->d_ctl,
                                                                                 psq)));
    LAUNCHED(p);
  }
  return 0;
}

// phase 3: reduce Q, out = P^ Q^T (the residual stays implicit: see the band kernels)
extern "C" int gb_psgd_recv(gb_psgd* p, float* d_out, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  gb_comm* c = p->comm;
  PS_DISPATCH(p->rp, (k3_final<RP><<<p->nbands, PS_BAND_THREADS, 0, st>>>(p->d_bands, p->d_mat, my_slots(p, p->off_q, p->slot_q),
                                                                         1.0f / (float)c->world, p->d_ph, d_out, p->d_qprev,
                                                                         gb_comm_peers(c), p->chan_q, p->d_ctl,
                                                                         (size_t)c->world * p->slot_q, c->status_dev, c->timeout_ns)));
  LAUNCHED(p);
  p->step++;
  return 0;
}
