// Internal structures shared by the translation units of libgrace_b200.
#pragma once
#include "../../include/grace_b200.h"
#include "common.cuh"

#include <vector>

#define GB_MAX_CHANNELS 4096
// flags[channel][parity][src_rank] : uint32 epochs, at the start of every arena
#define GB_FLAG_BYTES (GB_MAX_CHANNELS * 2 * GB_MAX_WORLD * 4)

__host__ __device__ __forceinline__ uint32_t* gb_flag(char* arena, int channel, int parity, int src) {
  return (uint32_t*)arena + ((size_t)channel * 2 + parity) * GB_MAX_WORLD + src;
}

// Additional arena regions (index >= 1).  Region 0 is gb_comm::base and also holds the flags; further regions are
// added collectively when a plan does not fit (gb_comm_region_add), so the arena is no longer sized by the first caller.
struct gb_region {
  char* base;
  size_t bytes, cursor;
  char* peer[GB_MAX_WORLD];
  bool opened[GB_MAX_WORLD];
  char* mc;                  // NVLS multicast address of the same allocation on all ranks (adopted symmetric memory), or null
  bool adopted;              // memory owned by the caller (torch symmetric memory): not freed here
};
#define GB_REGION_SHIFT 48   // a channel offset carries its region index in the bits above this

struct gb_comm {
  int rank, world, device;
  bool local_group;          // all ranks live in this process on one device (emulation)
  char* base;                // local arena
  size_t bytes;
  char* peer[GB_MAX_WORLD];  // mapped arenas, peer[rank] == base
  bool opened[GB_MAX_WORLD];
  size_t cursor;             // symmetric bump allocator (same call sequence on every rank)
  int next_channel;
  uint32_t* status_host;     // pinned + mapped: kernels report timeouts/overflow here
  uint32_t* status_dev;
  uint64_t timeout_ns;
  uint32_t barrier_epoch;
  std::vector<gb_region> extra;   // regions 1.. (same order and sizes on every rank)
  size_t need_bytes;              // bytes the last failed channel allocation asked for (0 = none pending)
  gb_comm** siblings;             // local (single-process) groups: all virtual ranks, for collective growth
  bool guard;                     // GRACE_B200_GUARD=1: every wire slot ends in a poisoned guard band
  struct GuardRec { size_t chan_off, slot_bytes; };
  std::vector<GuardRec> guarded;  // channels whose slots carry a guard band
};

// slot size as allocated: rounded to 256 bytes, plus the guard band when the communicator has guards on
static inline size_t gb_guarded(const gb_comm* c, size_t bytes) {
  return bytes ? ((bytes + 255) & ~size_t(255)) + (c->guard ? GB_GUARD_BYTES : 0) : 0;
}

int gb_comm_alloc_channel(gb_comm* c, size_t slot_bytes, int* channel, size_t* offset);
PeerPtrs gb_comm_peers(const gb_comm* c);

// Slot `src` of parity `parity` of a channel inside rank `arena_rank`'s arena (host side; region aware).
static inline char* gb_slot_of(const gb_comm* c, int arena_rank, size_t chan_off, size_t slot_bytes, int parity, int src) {
  const size_t reg = chan_off >> GB_REGION_SHIFT, off = chan_off & ((size_t(1) << GB_REGION_SHIFT) - 1);
  char* base = reg == 0 ? c->peer[arena_rank] : c->extra[reg - 1].peer[arena_rank];
  return base + off + ((size_t)parity * c->world + src) * slot_bytes;
}

// The same slot through the region's multicast window (one multimem.st reaches this slot in EVERY rank's arena), or
// null when the channel's region has no multicast mapping.
static inline char* gb_slot_mc(const gb_comm* c, size_t chan_off, size_t slot_bytes, int parity, int src) {
  const size_t reg = chan_off >> GB_REGION_SHIFT, off = chan_off & ((size_t(1) << GB_REGION_SHIFT) - 1);
  if (reg == 0 || !c->extra[reg - 1].mc) return nullptr;
  return c->extra[reg - 1].mc + off + ((size_t)parity * c->world + src) * slot_bytes;
}

// Slot r of parity q of a channel inside one arena.
__host__ __device__ __forceinline__ char* gb_slot(char* arena, size_t chan_off, size_t slot_bytes, int world, int parity,
                                                   int src) {
  return arena + chan_off + ((size_t)parity * world + src) * slot_bytes;
}

struct StepCtl;
struct gb_plan {
  gb_comm* comm;
  gb_codec_params prm;
  int nseg, ntiles;
  int64_t padded, ktotal;
  // host copies of the geometry
  std::vector<int64_t> h_off, h_kbase;
  std::vector<int32_t> h_numel, h_k, h_tile0;
  // device geometry
  SegTable tab;
  void* d_tables;
  // persistent state (padded floats, zero-initialised)
  float* d_state0;  // residual (ResidualMemory) or DGC accumulated gradient v
  float* d_state1;  // Signum momentum or DGC velocity u
  // scratch
  uint32_t* d_hist;      // [nseg][2048]
  uint32_t* d_sel;       // [nseg][4]: prefix, k_remaining, threshold bits, ties to take
  int32_t* d_tilecnt;    // [ntiles][2] then scanned in place
  float* d_tilestats;    // [ntiles][4]
  float* d_segstats;     // [nseg][4]
  uint32_t* d_done;      // last-block-done counters
  struct StepCtl* d_ctl; // device-side epoch / step (graph-replayable)
  size_t par_stride;     // bytes between the parity-0 and parity-1 slot of one source rank
  // wire
  int channel;
  int channel2;          // K0 only: flags of the second (gather) phase
  uint32_t mid_epoch;    // K0 only: epoch whose reduce phase has been enqueued
  bool dense2;           // K0 two-shot path active (None/FP16 codec, world > 1)
  size_t chan_off, slot_bytes;
  size_t hdr_bytes;      // per-slot header (counts / stats)
  int bits;              // quantiser bits per element
  uint32_t epoch;        // exchanges completed (flag value of the next send = epoch + 1)
  uint32_t step;         // codec call counter (Philox counter word 1)
  int64_t launches;
  void* sketch;          // SketchML buffers (k2_sketch.cu)
  void* adaq;            // AdaQ buffers (k1_adaq.cu)
  void* topk2;           // two-pass TopK buffers (k1_topk2.cu)
  uint32_t* d_samp;      // DGC: sampled magnitudes (k1_sparse.cu)
  size_t chan2_off, slot2_bytes;   // DGC gradient clipping: per-tensor sum(g^2) of every rank (channel2's slots)
  bool presend_done;     // gb_plan_presend already ran for the exchange being sent
};
