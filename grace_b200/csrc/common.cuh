// Shared device/host helpers for libgrace_b200 (sm_100a only).
//
// The reference has no native code at all; what it borrows from Horovod's
// C++ core, c10d/NCCL, ATen and CuPy is re-designed here around three
// primitives: peer-mapped arenas, epoch flags with system-scope release/acquire, and
// counter-based random numbers shared bit-for-bit with grace_b200/utils/philox.py.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#define GB_MAX_WORLD 8
#define GB_TILE 4096          // elements per work tile (one CTA iteration)
#define GB_ALIGN 128          // segment starts are multiples of this many elements
#define GB_THREADS 256

// ---------------------------------------------------------------- error plumbing
extern "C" const char* gb_last_error(void);
void gb_set_error(const char* fmt, ...);

#define GB_CUDA(expr)                                                                     \
  do {                                                                                    \
    cudaError_t _e = (expr);                                                              \
    if (_e != cudaSuccess) {                                                              \
      gb_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e));  \
      return -1;                                                                          \
    }                                                                                     \
  } while (0)

#define GB_CHECK(cond, ...)            \
  do {                                 \
    if (!(cond)) {                     \
      gb_set_error(__VA_ARGS__);       \
      return -1;                       \
    }                                  \
  } while (0)

// status word bits written by kernels (host-visible, see engine.cu)
#define GB_STATUS_TIMEOUT 1u
#define GB_STATUS_OVERFLOW 2u
#define GB_STATUS_GUARD 4u      // a guard band behind a wire slot was overwritten (GRACE_B200_GUARD=1)
#define GB_STATUS_ASSERT 8u     // a bounds assertion of a -DGB_DEBUG build failed (the store was skipped)
#define GB_GUARD_BYTES 256
#define GB_GUARD_WORD 0x7FC0DEADu   // a NaN pattern no codec produces

// Bounds assertions of the debug build (bash csrc/build.sh with GB_NVCC_EXTRA=-DGB_DEBUG): a failed check raises a
// status bit that Engine.check_status() turns into an exception, and skips the offending store.  compute-sanitizer
// is not available on the GPU pool, so this and the guard bands are what watches the peer stores.
#ifdef GB_DEBUG
#define GB_ASSERT_STORE(cond, status) ((cond) ? true : (atomicOr((status), GB_STATUS_ASSERT), false))
#else
#define GB_ASSERT_STORE(cond, status) (true)
#endif

// ---------------------------------------------------------------- memory-model helpers
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// streaming 16-byte accesses: gradients are read once per pass, payloads written once
__device__ __forceinline__ float4 ld_stream_f4(const float* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ void st_stream_u4(void* p, uint4 v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// Wire stores.  `mc` != 0: the pointer is an NVLS multicast address of a symmetric allocation -- ONE multimem.st reaches the
// slot in every rank's arena (the NVSwitch replicates it: one egress copy instead of W - 1 P2P stores); else a plain
// (peer-mapped) global store.  multimem.st moves bits, the .f32 type is only a width.
__device__ __forceinline__ void wire_st_u4(void* p, uint4 v, int mc) {
  if (mc) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(__uint_as_float(v.x)),
                 "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
                 : "memory");
  } else {
    st_stream_u4(p, v);
  }
}
__device__ __forceinline__ void wire_st_u2(void* p, uint2 v, int mc) {
  if (mc) {
    asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1,%2};" ::"l"(p), "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y))
                 : "memory");
  } else {
    *reinterpret_cast<uint2*>(p) = v;
  }
}
__device__ __forceinline__ void wire_st_u32(void* p, uint32_t v, int mc) {
  if (mc) asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(p), "f"(__uint_as_float(v)) : "memory");
  else *reinterpret_cast<uint32_t*>(p) = v;
}

// Wait until *flag has reached `epoch` (wrap-safe).  Bounded: a missing peer sets the
// TIMEOUT status bit instead of hanging the GPU (SURVEY §5 failure-detection row).
__device__ __forceinline__ bool wait_flag(const uint32_t* flag, uint32_t epoch, uint32_t* status, uint64_t timeout_ns) {
  uint64_t t0 = 0;
  uint32_t spins = 0;
  while ((int32_t)(ld_acquire_sys(flag) - epoch) < 0) {
    if ((++spins & 1023u) == 0) {
      uint64_t now = globaltimer_ns();
      if (t0 == 0) t0 = now;
      if (now - t0 > timeout_ns) {
        atomicOr(status, GB_STATUS_TIMEOUT);
        return false;
      }
    }
    __nanosleep(64);
  }
  return true;
}

// ---------------------------------------------------------------- Philox4x32-10
// Must match grace_b200/utils/philox.py bit for bit.
struct Philox4 { uint32_t x, y, z, w; };

__host__ __device__ __forceinline__ void philox_mulhilo(uint32_t a, uint32_t b, uint32_t& hi, uint32_t& lo) {
  uint64_t p = (uint64_t)a * (uint64_t)b;
  hi = (uint32_t)(p >> 32);
  lo = (uint32_t)p;
}

__host__ __device__ __forceinline__ Philox4 philox4x32(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                                       uint32_t k1) {
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    uint32_t hi0, lo0, hi1, lo1;
    philox_mulhilo(0xD2511F53u, c0, hi0, lo0);
    philox_mulhilo(0xCD9E8D57u, c2, hi1, lo1);
    uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return Philox4{c0, c1, c2, c3};
}

#define GB_STREAM_UNIFORM 0u
#define GB_STREAM_SAMPLE 1u
#define GB_STREAM_NORMAL 2u
#define GB_STREAM_PERMUTE 3u

__host__ __device__ __forceinline__ float u32_to_unit(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }

// ---------------------------------------------------------------- keyed permutation (RandomK)
__host__ __device__ __forceinline__ uint32_t lowbias32(uint32_t x) {
  x ^= x >> 16; x *= 0x7FEB352Du; x ^= x >> 15; x *= 0x846CA68Bu; x ^= x >> 16;
  return x;
}
__host__ __device__ __forceinline__ int feistel_half_bits(uint64_t n) {
  int b = 2;
  uint64_t m = (n < 2 ? 2 : n) - 1;
  int len = 0;
  while (m) { ++len; m >>= 1; }
  if (len > b) b = len;
  if (b & 1) ++b;
  return b / 2;
}
__host__ __device__ __forceinline__ uint64_t feistel_permute(uint64_t j, uint64_t n, int half, const uint32_t keys[4]) {
  const uint64_t mask = (1ull << half) - 1;
  uint64_t v = j;
  do {
    uint64_t left = (v >> half) & mask, right = v & mask;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      uint64_t nl = right;
      right = left ^ ((uint64_t)lowbias32((uint32_t)right ^ keys[r]) & mask);
      left = nl;
    }
    v = (left << half) | right;
  } while (v >= n);
  return v;
}

// ---------------------------------------------------------------- warp / block reductions
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ int warp_sum_i(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---------------------------------------------------------------- bucket geometry
// A "plan" flattens the gradient tensors of one bucket into one padded buffer.  Segment s
// occupies [seg_off[s], seg_off[s] + seg_numel[s]); seg_off is a multiple of GB_ALIGN; the
// pad elements are kept at zero.  Work is cut into tiles of GB_TILE elements that never
// straddle a segment, so one launch serves every tensor of the bucket while per-tensor
// semantics (k, norm, mean, residual) are preserved (SURVEY §7.1 decision 2).
struct SegTable {
  const int64_t* seg_off;     // [nseg + 1] padded element offsets
  const int32_t* seg_numel;   // [nseg]
  const int32_t* seg_k;       // [nseg]   payload entries per segment (sparse codecs)
  const int64_t* seg_kbase;   // [nseg+1] prefix sum of seg_k
  const uint32_t* seg_hash;   // [nseg]   fnv1a(name)
  const int32_t* seg_tile0;   // [nseg+1] first tile of each segment
  const int32_t* tile_seg;    // [ntiles]
  const int4* tile_desc;      // [ntiles] {base lo, base hi, segment, valid length}: one 16-byte load per tile (may be null)
  int32_t nseg;
  int32_t ntiles;
  int64_t padded;             // total padded elements
};

__device__ __forceinline__ void tile_bounds(const SegTable& t, int tile, int& seg, int64_t& base, int& local0, int& len) {
  if (t.tile_desc) {
    // precomputed descriptor: a single load instead of a chain of four dependent table look-ups
    const int4 d = __ldg(t.tile_desc + tile);
    base = (int64_t)(((uint64_t)(uint32_t)d.y << 32) | (uint32_t)d.x);
    seg = d.z;
    len = d.w & 0x1FFF;                       // 1..4096 fits 13 bits
    local0 = (int)((uint32_t)d.w >> 13) * GB_TILE;   // tile index inside the segment: 19 bits cover 2^31 elements
    return;
  }
  seg = t.tile_seg[tile];
  int within = tile - t.seg_tile0[seg];
  local0 = within * GB_TILE;
  int rem = t.seg_numel[seg] - local0;
  len = rem < GB_TILE ? rem : GB_TILE;
  base = t.seg_off[seg] + local0;
}

// pointers into every rank's arena, passed by value to kernels
struct PeerPtrs {
  char* base[GB_MAX_WORLD];
  int world;
  int rank;
};
