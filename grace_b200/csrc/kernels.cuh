// Launch entry points of the kernel families (one translation unit each).
#pragma once
#include "engine.cuh"

bool gb_codec_is_sparse(int codec);

// K1: sparsifiers -- compensate + select + ordered compaction + peer stores / ordered merge
int k1_send(gb_plan* p, const float* d_grad, cudaStream_t st);
int k1_recv(gb_plan* p, float* d_out, cudaStream_t st);
// K1-topk, two-pass pipeline (k1_topk2.cu); GRACE_B200_TOPK_V1=1 keeps the four-pass kernels of k1_sparse.cu
int tk_init(gb_plan* p);
void tk_free(gb_plan* p);
int tk_send(gb_plan* p, const float* d_grad, cudaStream_t st);
// K2: quantisers -- stats + encode/pack + peer stores / unpack-and-sum (+ majority vote)
int k2_send(gb_plan* p, const float* d_grad, cudaStream_t st);
int k2_recv(gb_plan* p, float* d_out, cudaStream_t st);

// K2-sketch: SketchML (quantile buckets) -- k2_sketch.cu
int sk_init(gb_plan* p);
void sk_free(gb_plan* p);
int sk_send(gb_plan* p, const float* d_grad, cudaStream_t st);
int sk_recv(gb_plan* p, float* d_out, cudaStream_t st);
size_t sk_header_bytes(int nseg, int q);

// K1-adaq: AdaQ (two sampled thresholds, indices + 2 means per tensor) -- k1_adaq.cu
int aq_init(gb_plan* p);
void aq_free(gb_plan* p);
int aq_send(gb_plan* p, const float* d_grad, cudaStream_t st);
int aq_recv(gb_plan* p, float* d_out, cudaStream_t st);
size_t aq_means_bytes(int nseg);

// K2-inceptionn: INCEPTIONN (2-bit tag + 32/16/8/0-bit value per element) -- k2_inceptionn.cu
int in_send(gb_plan* p, const float* d_grad, cudaStream_t st);
int in_recv(gb_plan* p, float* d_out, cudaStream_t st);
size_t in_hdr_bytes(int ntiles);
size_t in_tag_bytes(int64_t padded);

// K0: dense two-shot all-reduce over peer memory (None / FP16 codecs, world > 1)
int k0_send(gb_plan* p, const float* d_grad, cudaStream_t st);
int k0_mid(gb_plan* p, cudaStream_t st);
int k0_recv(gb_plan* p, float* d_out, cudaStream_t st);
size_t k0_slot_bytes(int64_t padded, int world, int bits);

// Per-plan step control block in DEVICE memory.  Kernels take the epoch (flag value, slot parity)
// and the codec step (Philox counter, first-call rules) from here instead of from kernel
// arguments, so a captured CUDA graph of a whole training step can be replayed: the tick kernel
// at the head of every send advances the block on the device.
struct StepCtl { uint32_t epoch, step, next_step, pad; };
void gb_launch_tick(StepCtl* ctl, cudaStream_t st);       // plan.cu
void gb_launch_wait(gb_comm* c, int channel, const StepCtl* ctl, cudaStream_t st);   // plan.cu: one-warp flag wait

// Last-block-done flag release shared by the send kernels: every CTA calls this after its
// final peer store; the CTA that arrives last publishes `epoch` to all peers.
__device__ __forceinline__ void release_flags_last_block(uint32_t* done, PeerPtrs peers, int channel, uint32_t epoch) {
  // One thread fences for the CTA: the barrier orders every thread's payload stores before thread 0's
  // system-scope fence (fences are cumulative), so 255 threads do not each wait for a membar.sys.
  __syncthreads();
  if (threadIdx.x == 0) {
    if (peers.world == 1) __threadfence();         // nothing left this GPU: device scope is enough
    else __threadfence_system();
    uint32_t ticket = atomicAdd(done, 1u);
    if (ticket == gridDim.x - 1) {
      __threadfence_system();
      for (int r = 0; r < peers.world; ++r) st_release_sys(gb_flag(peers.base[r], channel, epoch & 1, peers.rank), epoch);
      *done = 0;
    }
  }
}

// Directory of the tiles that still have work in a multi-round refinement (DGC / AdaQ thresholds): the CTA that
// finishes a round last lists the open tensors with a running tile count; the next round's launch walks
// 0 .. #open tiles  instead of every tile of the bucket and returns at once when nothing is open.  Lives in global
// scratch (4 + 2 nseg ints); L2 accesses (__ldcg / __stcg) so the scheme also holds inside one persistent kernel.
struct OpenDir { int* hdr; int* seg; int* pref; };      // hdr[0] = open tensors, hdr[1] = their tiles
__device__ __forceinline__ OpenDir open_dir(uint32_t* scratch, int nseg) {
  OpenDir d;
  d.hdr = reinterpret_cast<int*>(scratch);
  d.seg = d.hdr + 4;
  d.pref = d.seg + nseg;
  return d;
}
template <typename F>
__device__ __forceinline__ void open_dir_build(const SegTable& t, const OpenDir& d, F is_open) {      // whole CTA
  __shared__ unsigned long long s_w[GB_THREADS / 32];
  __shared__ unsigned long long s_base;
  constexpr unsigned long long LOW = (1ull << 40) - 1ull;
  if (threadIdx.x == 0) s_base = 0ull;
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int b0 = 0; b0 < t.nseg; b0 += GB_THREADS) {
    const int seg = b0 + (int)threadIdx.x;
    const bool open = seg < t.nseg && is_open(seg);
    const unsigned long long val = open ? ((1ull << 40) | (unsigned long long)(t.seg_tile0[seg + 1] - t.seg_tile0[seg])) : 0ull;
    unsigned long long inc = val;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned long long nb = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += nb;
    }
    if (lane == 31) s_w[wid] = inc;
    __syncthreads();
    unsigned long long wbase = s_base;
    for (int w = 0; w < wid; ++w) wbase += s_w[w];
    const unsigned long long ex = wbase + inc - val;
    if (open) {
      const int k = (int)(ex >> 40);
      __stcg(d.seg + k, seg);
      __stcg(d.pref + k, (int)(ex & LOW));
    }
    __syncthreads();
    if (threadIdx.x == GB_THREADS - 1) s_base = wbase + inc;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    __stcg(d.hdr, (int)(s_base >> 40));
    __stcg(d.hdr + 1, (int)(s_base & LOW));
  }
}
// the vt-th open tile; s_pref caches the first `cached` running tile counts in shared memory
__device__ __forceinline__ int open_dir_tile(const SegTable& t, const OpenDir& d, const int* s_pref, int cached, int nopen, int vt) {
  int lo = 0, hi = nopen;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    const int pm = mid < cached ? s_pref[mid] : __ldcg(d.pref + mid);
    if (pm <= vt) lo = mid; else hi = mid;
  }
  const int pl = lo < cached ? s_pref[lo] : __ldcg(d.pref + lo);
  return t.seg_tile0[__ldcg(d.seg + lo)] + vt - pl;
}

// All threads of the CTA return once every rank's flag for `epoch` has arrived.
__device__ __forceinline__ void wait_all_ranks(PeerPtrs peers, int channel, uint32_t epoch, uint32_t* status,
                                               uint64_t timeout_ns) {
  if (threadIdx.x < peers.world)
    wait_flag(gb_flag(peers.base[peers.rank], channel, epoch & 1, threadIdx.x), epoch, status, timeout_ns);
  __syncthreads();
}
