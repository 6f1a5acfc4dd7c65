// gb_topk_compress_static: exact top-k selection for callers that must not allocate.
//
// Who needs this.  gb_topk_compress() asks its caller for a scratch buffer (histograms, tile
// counters, the single-segment tables).  Some callers have nowhere to get one from: a plain
// C/C++ host that drives the library without a framework allocator and whose device memory has
// already been handed out in full to someone else's pool (a late cudaMalloc then fails); and
// code on a latency-critical path -- a gradient hook, a communication thread next to peer-memory
// kernels -- where cudaMalloc/cudaFree are off limits because they synchronise the whole device
// and would stall, or dead-lock against, kernels that are spinning on another rank's flag.
// For them the scratch lives inside the library image itself.
//
// Cost and limits.
//  * The scratch is one file-scope __device__ array of GB_TOPK_STATIC_SCRATCH_BYTES (96 KiB).
//    Every GPU that loads libgrace_b200.so carries it from module load until process exit,
//    whether or not this function is ever called.
//  * That bounds the input: n <= GB_TOPK_STATIC_MAX_N (2^24 elements); larger n is rejected
//    with an error message, nothing is launched.
//  * There is ONE array per GPU and every call uses it.  It keeps whatever the previous call
//    left behind; each call re-initialises the parts that must start from zero (the histogram
//    and the segment tables) on the stream before using them.  Two calls whose kernels overlap
//    on the same GPU (different streams, or different host threads) race on the array and
//    produce garbage: serialise them.
//  * The first call on each device looks up the array's device address (cudaGetSymbolAddress,
//    once per device, with that device current).  Make that first call outside any stream
//    capture; later calls only launch kernels.
//  * No cudaMalloc / cudaFree / allocating library routine is called here or below.
#include "kernels.cuh"

#include <mutex>

#define GB_TOPK_STATIC_MAX_N (int64_t(1) << 24)
#define GB_TOPK_STATIC_SCRATCH_BYTES (96 * 1024)
#define GB_MAX_DEVICES 64

// In host code this name is only a stand-in for the symbol; it is NOT a device address.
static __device__ __align__(256) unsigned char g_topk_static_scratch[GB_TOPK_STATIC_SCRATCH_BYTES];

int k1_topk_single(const float* d_x, int64_t n, int64_t k, float* d_vals, int32_t* d_idx, char* scratch, cudaStream_t st);

static std::mutex g_mu;
static char* g_scratch_addr[GB_MAX_DEVICES];   // device address of the array, per device

static int scratch_address(char** out) {
  int dev = -1;
  GB_CUDA(cudaGetDevice(&dev));
  GB_CHECK(dev >= 0 && dev < GB_MAX_DEVICES, "device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_scratch_addr[dev]) {
    void* p = nullptr;
    GB_CUDA(cudaGetSymbolAddress(&p, g_topk_static_scratch));   // `dev` is current: its copy of the array
    GB_CHECK(p != nullptr, "cudaGetSymbolAddress returned a null address");
    g_scratch_addr[dev] = (char*)p;
  }
  *out = g_scratch_addr[dev];
  return 0;
}

extern "C" int64_t gb_topk_static_max_n(void) { return GB_TOPK_STATIC_MAX_N; }

extern "C" int gb_topk_compress_static(const float* d_x, int64_t n, int64_t k, float* d_vals, int32_t* d_idx, void* stream) {
  GB_CHECK(n >= 1, "n = %lld must be positive", (long long)n);
  GB_CHECK(n <= GB_TOPK_STATIC_MAX_N,
           "gb_topk_compress_static: n = %lld exceeds the built-in scratch (max %lld elements); "
           "use gb_topk_compress with a caller-provided scratch buffer",
           (long long)n, (long long)GB_TOPK_STATIC_MAX_N);
  GB_CHECK(gb_topk_scratch_bytes(n) <= GB_TOPK_STATIC_SCRATCH_BYTES, "internal: scratch layout for n = %lld needs %zu bytes",
           (long long)n, gb_topk_scratch_bytes(n));
  char* scratch = nullptr;
  if (scratch_address(&scratch)) return -1;
  return k1_topk_single(d_x, n, k, d_vals, d_idx, scratch, (cudaStream_t)stream);
}
