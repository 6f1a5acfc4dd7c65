/* libgrace_b200 -- C ABI of the Blackwell-native gradient-compression engine.
 *
 * The reference (Crystal-wxy/grace) ships no native code; the native pieces it relies on are
 * Horovod's C++ core (handles, poll/synchronize), c10d/NCCL collectives and ATen/CuPy kernels
 * (SURVEY.md §2.7).  This library replaces them with:
 *   - a communicator owning a peer-mapped arena, epoch flags and a host-visible status word;
 *   - "plans": one flattened gradient bucket + its error-feedback state + one compressed
 *     exchange, executed as  gb_plan_send (compensate + compress + stores into every peer)
 *     and gb_plan_recv (wait flags + decode + aggregate, written once);
 *   - stand-alone single-tensor kernels used by the synchronous Communicator.step API.
 * Every function returns 0 on success and -1 on failure (message in gb_last_error()).
 * All pointers named d_* are device pointers; `stream` is a cudaStream_t passed as void*.
 */
#ifndef GRACE_B200_H
#define GRACE_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct gb_comm gb_comm;
typedef struct gb_plan gb_plan;

/* codecs (reference: grace_dl/dist/compressor/<name>.py) */
enum {
  GB_CODEC_NONE = 0,      /* none.py:4-12      fp32 on the wire            */
  GB_CODEC_FP16 = 1,      /* fp16.py:6-22      fp16 on the wire            */
  GB_CODEC_TOPK = 2,      /* topk.py:6-36      (fp32, int32) pairs         */
  GB_CODEC_RANDOMK = 3,   /* randomk.py:6-40   fp32 values only            */
  GB_CODEC_THRESHOLD = 4, /* threshold.py:6-27 variable-size pairs         */
  GB_CODEC_DGC = 5,       /* dgc.py:6-50       variable-size pairs         */
  GB_CODEC_QSGD = 6,      /* qsgd.py:6-38      4/8/16-bit levels + norm    */
  GB_CODEC_TERNGRAD = 7,  /* terngrad.py:6-32  2-bit codes + scalar        */
  GB_CODEC_NATURAL = 8,   /* natural.py:9-40   8-bit sign+exponent         */
  GB_CODEC_U8BIT = 9,     /* tensorflow/compressor/u8bit.py:6-110          */
  GB_CODEC_SIGNSGD = 10,  /* signsgd.py:6-30   1 bit, majority vote        */
  GB_CODEC_SIGNUM = 11,   /* signum.py:6-37    1 bit of the momentum       */
  GB_CODEC_EFSIGNSGD = 12,/* efsignsgd.py:6-33 1 bit + mean|x|             */
  GB_CODEC_ONEBIT = 13,   /* torch/compressor/onebit.py:6-32 1 bit + 2 means */
  GB_CODEC_SKETCH = 14,   /* tensorflow/compressor/sketch.py:6-39  8-bit quantile bucket + q means (quantum_num = q) */
  GB_CODEC_ADAQ = 15,     /* tensorflow/compressor/adaq.py:6-93    int32 indices (sign in bit 31) + 2 means per tensor */
  GB_CODEC_INCEPTIONN = 16 /* tensorflow/compressor/inceptionn.py:8-188  2-bit tag + 32/16/8/0-bit value per element
                              (quantum_num = low_exp | mid_exp << 8) */
};

/* memories (reference: grace_dl/dist/memory/<name>.py) */
enum {
  GB_MEM_NONE = 0,     /* none.py:4-11                                         */
  GB_MEM_RESIDUAL = 1, /* residual.py:4-20 and efsignsgd.py:4-19 (beta=1, gamma=lr) */
  GB_MEM_DGC = 2       /* dgc.py:7-39  momentum correction + accumulation       */
};

typedef struct gb_codec_params {
  double ratio;            /* topk / randomk / dgc compress_ratio (double: k = int(numel*ratio) as in python) */
  int32_t codec;
  int32_t memory;
  float beta, gamma;       /* residual memory: x = beta*e + gamma*g                    */
  float threshold;         /* threshold codec                                          */
  int32_t quantum_num;     /* qsgd                                                     */
  float momentum;          /* signum (codec) or dgc (memory)                           */
  float lr;                /* efsignsgd: aggregate = sum / lr                          */
  uint32_t seed;           /* Philox key word 0 (shared by all ranks)                  */
  int32_t average;         /* divide the aggregate by world size                       */
  float capacity;          /* variable-size codecs: slot capacity as a fraction of numel (0 -> 1.0) */
  float one_minus_momentum;/* signum: (float)(1.0 - momentum) evaluated in double like python; 0 -> 1.0f - momentum */
  int32_t clip;            /* dgc memory: clamp the gradient to +-sqrt(sum over ranks of sum(g^2) / W) per tensor
                              (memory/dgc.py:17-20, the dist helper's default) before the momentum correction      */
} gb_codec_params;

const char* gb_last_error(void);
int gb_version(void);
/* make `device` current for this thread inside the library's CUDA runtime instance */
int gb_set_device(int device);

/* ---- communicator: arena + peers (replaces Horovod core / c10d bootstrap) */
int gb_comm_create(int rank, int world, size_t arena_bytes, gb_comm** out);
int gb_comm_ipc_handle(gb_comm* c, void* out64);                 /* 64-byte cudaIpcMemHandle_t   */
int gb_comm_open_peers(gb_comm* c, const void* handles);         /* world * 64 bytes, rank order */
/* `world` virtual ranks on the current device (tests, single-GPU emulation of the protocol) */
int gb_comm_create_local(int world, size_t arena_bytes, gb_comm** out /* [world] */);
int gb_comm_destroy(gb_comm* c);
int gb_comm_status(gb_comm* c, uint32_t* out);                   /* 0 ok; bit0 timeout; bit1 overflow; bit2 guard band
                                                                    overwritten; bit3 debug-build bounds assertion */
/* With GRACE_B200_GUARD=1 (set before the communicator is created, on every rank) each wire slot ends in a 256-byte
 * poisoned band that no kernel may touch; this scans them (device synchronisation) and raises status bit 2. */
int gb_comm_check_guards(gb_comm* c, int* nchannels);
int gb_comm_clear_status(gb_comm* c);
int gb_comm_set_timeout_ms(gb_comm* c, uint64_t ms);
int gb_comm_rank(gb_comm* c);
int gb_comm_world(gb_comm* c);
size_t gb_comm_arena_free(gb_comm* c);
/* The arena grows by whole regions.  When a plan's wire slot does not fit, gb_plan_create / gb_psgd_create fail and
 * gb_comm_arena_need() reports the missing bytes; every rank then adds a region of the same size (collective:
 * gb_comm_region_add, exchange the 64-byte IPC handles, gb_comm_region_open_peers) and retries.  The Python Engine
 * does this automatically (Horovod's fusion buffer and NCCL's buffers also grow on demand). */
size_t gb_comm_arena_need(gb_comm* c);
int gb_comm_regions(gb_comm* c);
int gb_comm_region_add(gb_comm* c, size_t bytes, int* index);
/* adopt caller-owned symmetric memory (all ranks' mappings + its NVLS multicast address) as the next region: wire slots
 * carved from it are written with ONE multimem.st per 16 bytes for all ranks instead of W - 1 peer stores */
int gb_comm_region_adopt(gb_comm* c, const uint64_t* peers, uint64_t multicast, size_t bytes, int* index);
int gb_comm_has_multicast(gb_comm* c);
int gb_comm_region_ipc_handle(gb_comm* c, int index, void* out64);
int gb_comm_region_open_peers(gb_comm* c, int index, const void* handles);
/* all ranks reach this point (flag barrier over peer memory; used by tests and K0) */
int gb_comm_barrier(gb_comm* c, void* stream);

/* ---- events with integer handles (Horovod poll/synchronize, mpi_ops.py:420-439) */
int gb_handle_record(void* stream, int64_t* out_handle);
int gb_handle_poll(int64_t handle);            /* 1 done, 0 pending, -1 error */
int gb_handle_wait(int64_t handle);            /* blocks the host; releases the handle */
int gb_handle_stream_wait(int64_t handle, void* stream);  /* make `stream` wait; keeps the handle */
int gb_handle_release(int64_t handle);

/* ---- point collectives for everything that is not a gradient bucket (reference: patch_files/horovod/torch/mpi_ops.py
 * :157 allreduce_async_, :216 allgather_async, :362 broadcast_async_ over Horovod's core).  One op = gb_ops_send
 * (store this rank's bytes into every peer, release the op's flag; NULL source = flag only) + gb_ops_recv (wait for
 * all ranks, then mode 0: copy `count` ranks' payloads starting at `first` into d_out, `stride` bytes apart;
 * mode 1/2/3/4: sum all ranks' payloads as float / double / int64 / int32 in rank order, times `scale`).
 * Sizes are multiples of 16 bytes up to the capacity given at creation.  Issue the ops of one gb_ops on ONE stream. */
typedef struct gb_ops gb_ops;
int gb_ops_create(gb_comm* c, size_t slot_bytes, gb_ops** out);
int gb_ops_destroy(gb_ops* o);
size_t gb_ops_capacity(gb_ops* o);
int64_t gb_ops_launches(gb_ops* o);
int gb_ops_send(gb_ops* o, const void* d_src, size_t nbytes, void* stream);
int gb_ops_recv(gb_ops* o, size_t nbytes, int first, int count, size_t stride, int mode, double scale, void* d_out,
                void* stream);

/* ---- plans: one gradient bucket, one compressed exchange */
int gb_plan_create(gb_comm* c, int nseg, const int64_t* numels, const uint32_t* name_hash,
                   const gb_codec_params* params, gb_plan** out);
int gb_plan_destroy(gb_plan* p);
int64_t gb_plan_padded_numel(gb_plan* p);
int gb_plan_offsets(gb_plan* p, int64_t* out /* [nseg] element offsets into the flat buffer */);
int64_t gb_plan_wire_bytes(gb_plan* p);        /* payload bytes one rank sends to one peer */
/* compensate + compress + store into every peer's slot + release flags */
int gb_plan_send(gb_plan* p, const float* d_grad, void* stream);
/* Optional first half of gb_plan_send for plans whose send itself waits for the peers (DGC with gradient clipping:
 * every rank needs all ranks' sum(g^2) before it can clamp).  It advances the step and pushes this rank's scalars;
 * the following gb_plan_send then continues from there.  Only needed when several ranks are driven from one
 * stream (single-GPU emulation): all ranks' presend first, then all ranks' send.  A no-op for every other plan. */
int gb_plan_presend(gb_plan* p, const float* d_grad, void* stream);
/* optional middle phase: the dense two-shot all-reduce (None/FP16 codec, world > 1) reduces its shard
 * here; a no-op for every other codec; gb_plan_recv runs it itself if the caller did not */
int gb_plan_mid(gb_plan* p, void* stream);
/* one warp polls every rank's flag of the current exchange (no-op at world size 1).  Enqueued before gb_plan_mid /
 * gb_plan_recv it lets a caller queue the receive side right behind the send, overlapping the rest of backward,
 * without the receive kernels' grids occupying SMs while a slower rank is still sending.  The reference receives
 * everything after backward (patch_files/horovod/torch/__init__.py:143-161). */
int gb_plan_wait(gb_plan* p, void* stream);
/* wait for all ranks' flags, decode, aggregate; writes every element of d_out exactly once */
int gb_plan_recv(gb_plan* p, float* d_out, void* stream);
/* state access for state_dict(): which = 0 residual / dgc v, 1 momentum / dgc u */
int gb_plan_state(gb_plan* p, int which, void** d_ptr, size_t* bytes);
/* copy the whole padded state buffer out of / into the plan (device to device, on `stream`) */
int gb_plan_state_read(gb_plan* p, int which, void* d_dst, void* stream);
int gb_plan_state_write(gb_plan* p, int which, const void* d_src, void* stream);
int gb_plan_get_step(gb_plan* p, uint32_t* step);
int gb_plan_set_step(gb_plan* p, uint32_t step);
int gb_plan_launches(gb_plan* p, int64_t* out);  /* kernels launched by this plan so far */

/* ---- NVLS (NVLink SHARP multicast) dense all-reduce, in place on a multicast-bound bucket.
 * Reference: communicator/allreduce.py:8-13 with compressor/none.py (NCCL all-reduce + div_).
 * The caller provides the multicast address of a symmetric allocation that holds the bucket on
 * every rank (e.g. torch.distributed._symmetric_memory); flags travel through the communicator. */
typedef struct gb_nvls gb_nvls;
int gb_nvls_create(gb_comm* c, int64_t padded_numel, gb_nvls** out);
int gb_nvls_destroy(gb_nvls* p);
int gb_nvls_signal(gb_nvls* p, void* stream);                 /* this rank's copy is final */
int gb_nvls_reduce(gb_nvls* p, void* d_multicast, int average, void* stream);
int gb_nvls_launches(gb_nvls* p, int64_t* out);

/* ---- PowerSGD plans (reference: compressor/powersgd.py:7-65, memory/powersgd.py:6-37), rank <= 16 (rows of P and Q
 * travel with a pitch of 4, 8 or 16 floats).  gb_psgd_create2 adds warm_start: the averaged Q of one exchange seeds the
 * next one instead of a fresh normal draw (compressor/powersgd.py:37-39 use_memory); gb_psgd_qprev* expose that state.
 * All >=2-D gradients of a bucket, flattened like a plan (128-element aligned).  One exchange is
 *   gb_psgd_send  (M = grad + e; Q0; P_r = M Q0 on tcgen05/TMEM via TMA; P tiles stored into every peer)
 *   gb_psgd_mid   (wait; P = mean; orthogonalise; Q_r = M^T P^ on tcgen05; Q tiles to every peer)
 *   gb_psgd_recv  (wait; Q = mean; out = P^ Q^T)
 * The error-feedback memory e = M - P^ Q^T is implicit: the plan keeps M, P^ and Q of the last exchange and the next
 * gb_psgd_send recomputes e where it is consumed (same bits as a stored residual); gb_psgd_state_read materialises it
 * into a caller buffer of gb_psgd_padded_numel floats, gb_psgd_state_write loads one (checkpoint / resume).
 * use_tcgen05 = 0 forces the CUDA-core GEMMs (also taken per matrix smaller than 8 x 32). */
typedef struct gb_psgd gb_psgd;
int gb_psgd_create(gb_comm* c, int nmat, const int64_t* rows, const int64_t* cols, const uint32_t* name_hash,
                   int rank, uint32_t seed, int has_memory, int use_tcgen05, gb_psgd** out);
int gb_psgd_create2(gb_comm* c, int nmat, const int64_t* rows, const int64_t* cols, const uint32_t* name_hash,
                    int rank, uint32_t seed, int has_memory, int use_tcgen05, int warm_start, gb_psgd** out);
int gb_psgd_qprev(gb_psgd* p, void** d_ptr, size_t* bytes);
int gb_psgd_qprev_copy(gb_psgd* p, void* d_buf, int to_plan, void* stream);
int gb_psgd_destroy(gb_psgd* p);
int64_t gb_psgd_padded_numel(gb_psgd* p);
int gb_psgd_offsets(gb_psgd* p, int64_t* out);
int gb_psgd_send(gb_psgd* p, const float* d_grad, void* stream);
int gb_psgd_wait(gb_psgd* p, int which /* 0: P tiles, 1: Q tiles */, void* stream);
int gb_psgd_mid(gb_psgd* p, void* stream);
int gb_psgd_recv(gb_psgd* p, float* d_out, void* stream);
int gb_psgd_launches(gb_psgd* p, int64_t* out);
int gb_psgd_tc_matrices(gb_psgd* p);          /* how many matrices take the tcgen05 path */
int gb_psgd_get_step(gb_psgd* p, uint32_t* step);
int gb_psgd_set_step(gb_psgd* p, uint32_t step);
int gb_psgd_residual(gb_psgd* p, void** d_ptr, size_t* bytes);    /* non-NULL iff the plan keeps error-feedback state */
int gb_psgd_state_read(gb_psgd* p, void* d_dst, void* stream);    /* residual, padded floats (computed on demand) */
int gb_psgd_state_write(gb_psgd* p, const void* d_src, void* stream);
int64_t gb_psgd_wire_bytes(gb_psgd* p);

/* ---- stand-alone single-tensor kernels (synchronous API, microbenchmarks) */
/* exact top-k by magnitude: k (value, index) pairs in ascending index order.
 * d_scratch: at least gb_topk_scratch_bytes(n) bytes of device memory, any contents. */
size_t gb_topk_scratch_bytes(int64_t n);
int gb_topk_compress(const float* d_x, int64_t n, int64_t k, float* d_vals, int32_t* d_idx,
                     void* d_scratch, size_t scratch_bytes, void* stream);
/* Same result as gb_topk_compress, for callers that cannot hand over a scratch buffer because
 * they cannot allocate one (no allocator of their own and all device memory already reserved by
 * another pool; or a hook / communication thread where cudaMalloc's device-wide synchronisation
 * is not acceptable).  This entry point has no counterpart in the reference.  The scratch is a
 * 96 KiB __device__ array inside the library, resident on every GPU that loads it; therefore
 * n <= gb_topk_static_max_n() (2^24), larger inputs are refused with an error.  All calls on a
 * GPU share that array: calls whose kernels overlap on the GPU race.  The first call on a device
 * resolves the array's address (once); make it before starting any stream capture.  Never calls
 * cudaMalloc/cudaFree. */
int64_t gb_topk_static_max_n(void);
int gb_topk_compress_static(const float* d_x, int64_t n, int64_t k, float* d_vals, int32_t* d_idx, void* stream);
/* dense = sum_r scatter(vals_r, idx_r) * scale, ranks added in order; lists sorted by index */
int gb_sparse_decompress(const float* const* d_vals, const int32_t* const* d_idx, const int64_t* counts,
                         int nranks, int64_t n, float scale, float* d_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GRACE_B200_H */
