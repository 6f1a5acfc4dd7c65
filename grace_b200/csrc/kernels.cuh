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

// All threads of the CTA return once every rank's flag for `epoch` has arrived.
__device__ __forceinline__ void wait_all_ranks(PeerPtrs peers, int channel, uint32_t epoch, uint32_t* status,
                                               uint64_t timeout_ns) {
  if (threadIdx.x < peers.world)
    wait_flag(gb_flag(peers.base[peers.rank], channel, epoch & 1, threadIdx.x), epoch, status, timeout_ns);
  __syncthreads();
}
