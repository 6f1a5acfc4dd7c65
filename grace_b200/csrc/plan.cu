// Plans: geometry of one gradient bucket, its persistent error-feedback state, its wire slot.
//
// The reference keeps per-name python dicts of full-size tensors (grace_dl/dist/memory/
// residual.py:6, memory/dgc.py:12-13, compressor/signum.py:11) and issues one collective per
// payload tensor per parameter (SURVEY §2.8 "no bucketing").  A plan flattens a whole bucket:
// state lives in two flat device buffers, and one send + one recv launch sequence serves every
// tensor of the bucket through the segment/tile tables built here.
#include "engine.cuh"
#include "kernels.cuh"

#include <stdlib.h>

#include <algorithm>

static int quant_bits(const gb_codec_params& p) {
  switch (p.codec) {
    case GB_CODEC_NONE: return 32;
    case GB_CODEC_FP16: return 16;
    // signed levels -q..q in two's complement: b bits hold q <= 2^(b-1) - 1 ("QSGD 8 levels" of BASELINE config 4: 5 bits)
    case GB_CODEC_QSGD: {
      const int q = p.quantum_num;
      return q <= 7 ? 4 : q <= 15 ? 5 : q <= 31 ? 6 : q <= 63 ? 7 : q <= 127 ? 8 : 16;
    }
    case GB_CODEC_TERNGRAD: return 2;
    case GB_CODEC_SKETCH: return p.quantum_num < 256 ? 8 : 16;     // uint16 bins for >= 256 quantiles (sketch.py:26)
    case GB_CODEC_NATURAL:
    case GB_CODEC_U8BIT: return 8;
    case GB_CODEC_SIGNSGD:
    case GB_CODEC_SIGNUM:
    case GB_CODEC_EFSIGNSGD:
    case GB_CODEC_ONEBIT: return 1;
    default: return 0;
  }
}

bool gb_codec_is_sparse(int codec) {
  return codec == GB_CODEC_TOPK || codec == GB_CODEC_RANDOMK || codec == GB_CODEC_THRESHOLD || codec == GB_CODEC_DGC ||
         codec == GB_CODEC_ADAQ;
}

template <typename T>
static T* carve(char*& cur, size_t count) {
  T* p = (T*)cur;
  cur += (count * sizeof(T) + 255) & ~size_t(255);
  return p;
}

extern "C" int gb_plan_create(gb_comm* c, int nseg, const int64_t* numels, const uint32_t* name_hash,
                              const gb_codec_params* params, gb_plan** out) {
  GB_CHECK(nseg >= 1, "a plan needs at least one segment");
  GB_CHECK(params->codec >= GB_CODEC_NONE && params->codec <= GB_CODEC_INCEPTIONN, "unknown codec %d", params->codec);
  GB_CHECK(params->memory != GB_MEM_DGC || params->codec == GB_CODEC_DGC, "DGC memory needs the DGC codec");
  gb_plan* p = new gb_plan();            // value-initialised: every pointer member starts as nullptr
  // any failure below leaves through this lambda; the partially built plan is then released in one place
  auto build = [&]() -> int {
  p->comm = c;
  p->prm = *params;
  p->nseg = nseg;
  p->epoch = 0;
  p->step = 0;
  p->launches = 0;
  p->sketch = nullptr;
  p->adaq = nullptr;
  p->topk2 = nullptr;
  p->d_samp = nullptr;
  p->bits = quant_bits(*params);
  const bool sparse = gb_codec_is_sparse(params->codec);

  // ---- geometry (host)
  p->h_off.resize(nseg + 1);
  p->h_kbase.resize(nseg + 1);
  p->h_numel.resize(nseg);
  p->h_k.resize(nseg);
  p->h_tile0.resize(nseg + 1);
  std::vector<int32_t> tile_seg;
  int64_t off = 0, kbase = 0;
  for (int s = 0; s < nseg; ++s) {
    GB_CHECK(numels[s] >= 1 && numels[s] < (int64_t(1) << 31), "segment %d has %lld elements (must be in [1, 2^31))", s,
             (long long)numels[s]);
    p->h_off[s] = off;
    p->h_numel[s] = (int32_t)numels[s];
    int64_t k = 0;
    if (params->codec == GB_CODEC_TOPK || params->codec == GB_CODEC_RANDOMK)
      k = std::max<int64_t>(1, (int64_t)((double)numels[s] * (double)params->ratio));
    if (params->codec == GB_CODEC_DGC)          // samples per tensor (dgc.py:17); seg_kbase = offsets into the sample scratch
      k = std::max<int64_t>(1, (int64_t)((double)numels[s] * 0.01));
    k = std::min<int64_t>(k, numels[s]);
    p->h_k[s] = (int32_t)k;
    p->h_kbase[s] = kbase;
    kbase += k;
    p->h_tile0[s] = (int32_t)tile_seg.size();
    int nt = (int)((numels[s] + GB_TILE - 1) / GB_TILE);
    for (int t = 0; t < nt; ++t) tile_seg.push_back(s);
    off += (numels[s] + GB_ALIGN - 1) / GB_ALIGN * GB_ALIGN;
  }
  p->h_off[nseg] = off;
  p->h_kbase[nseg] = kbase;
  p->h_tile0[nseg] = (int32_t)tile_seg.size();
  p->padded = off;
  p->ntiles = (int)tile_seg.size();
  if (params->codec == GB_CODEC_THRESHOLD || params->codec == GB_CODEC_DGC || params->codec == GB_CODEC_ADAQ ||
      params->codec == GB_CODEC_INCEPTIONN) {            // INCEPTIONN: capacity of the value stream in 32-bit words
    float cap = params->capacity > 0.f ? params->capacity : 1.0f;
    kbase = std::max<int64_t>(1, (int64_t)((double)cap * (double)off));
  }
  p->ktotal = kbase;

  // ---- device tables, state and scratch in one allocation
  size_t bytes = 0;
  auto add = [&](size_t n) { bytes += (n + 255) & ~size_t(255); };
  add((nseg + 1) * 8); add(nseg * 4); add(nseg * 4); add((nseg + 1) * 8); add(nseg * 4); add((nseg + 1) * 4);
  add(p->ntiles * 4); add((size_t)p->ntiles * 16);
  add((size_t)nseg * 2048 * 4); add((size_t)nseg * 4 * 4); add((size_t)p->ntiles * 2 * 4 + 8);
  add((size_t)p->ntiles * 8 * 4); add((size_t)nseg * 4 * 4); add(64); add(64);
  GB_CUDA(cudaMalloc(&p->d_tables, bytes));
  GB_CUDA(cudaMemset(p->d_tables, 0, bytes));
  char* cur = (char*)p->d_tables;
  int64_t* d_off = carve<int64_t>(cur, nseg + 1);
  int32_t* d_numel = carve<int32_t>(cur, nseg);
  int32_t* d_k = carve<int32_t>(cur, nseg);
  int64_t* d_kbase = carve<int64_t>(cur, nseg + 1);
  uint32_t* d_hash = carve<uint32_t>(cur, nseg);
  int32_t* d_tile0 = carve<int32_t>(cur, nseg + 1);
  int32_t* d_tileseg = carve<int32_t>(cur, p->ntiles);
  int4* d_tiledesc = carve<int4>(cur, p->ntiles);
  p->d_hist = carve<uint32_t>(cur, (size_t)nseg * 2048);
  p->d_sel = carve<uint32_t>(cur, (size_t)nseg * 4);
  p->d_tilecnt = carve<int32_t>(cur, (size_t)p->ntiles * 2 + 2);
  p->d_tilestats = carve<float>(cur, (size_t)p->ntiles * 8);
  p->d_segstats = carve<float>(cur, (size_t)nseg * 4);
  p->d_done = carve<uint32_t>(cur, 16);
  p->d_ctl = (StepCtl*)carve<uint32_t>(cur, 16);
  GB_CUDA(cudaMemcpy(d_off, p->h_off.data(), (nseg + 1) * 8, cudaMemcpyHostToDevice));
  GB_CUDA(cudaMemcpy(d_numel, p->h_numel.data(), nseg * 4, cudaMemcpyHostToDevice));
  GB_CUDA(cudaMemcpy(d_k, p->h_k.data(), nseg * 4, cudaMemcpyHostToDevice));
  GB_CUDA(cudaMemcpy(d_kbase, p->h_kbase.data(), (nseg + 1) * 8, cudaMemcpyHostToDevice));
  GB_CUDA(cudaMemcpy(d_hash, name_hash, nseg * 4, cudaMemcpyHostToDevice));
  GB_CUDA(cudaMemcpy(d_tile0, p->h_tile0.data(), (nseg + 1) * 4, cudaMemcpyHostToDevice));
  GB_CUDA(cudaMemcpy(d_tileseg, tile_seg.data(), p->ntiles * 4, cudaMemcpyHostToDevice));
  bool desc_ok = true;
  {
    std::vector<int4> desc(p->ntiles);
    for (int tile = 0; tile < p->ntiles; ++tile) {
      const int sg = tile_seg[tile];
      const int within = tile - p->h_tile0[sg];
      const int64_t base = p->h_off[sg] + (int64_t)within * GB_TILE;
      const int len = (int)std::min<int64_t>(GB_TILE, p->h_numel[sg] - (int64_t)within * GB_TILE);
      if (within >= (1 << 19)) desc_ok = false;          // cannot happen below 2^31 elements; kept as a guard
      desc[tile] = make_int4((int)(uint32_t)(base & 0xFFFFFFFFll), (int)(uint32_t)((uint64_t)base >> 32), sg,
                             (int)((uint32_t)len | ((uint32_t)within << 13)));
    }
    GB_CUDA(cudaMemcpy(d_tiledesc, desc.data(), (size_t)p->ntiles * 16, cudaMemcpyHostToDevice));
  }
  p->tab = SegTable{d_off, d_numel, d_k, d_kbase, d_hash, d_tile0, d_tileseg, desc_ok ? d_tiledesc : nullptr, nseg, p->ntiles,
                    p->padded};

  p->d_state0 = p->d_state1 = nullptr;
  const bool need0 = params->memory != GB_MEM_NONE;
  const bool need1 = params->codec == GB_CODEC_SIGNUM || params->memory == GB_MEM_DGC;
  if (need0) {
    GB_CUDA(cudaMalloc((void**)&p->d_state0, p->padded * 4));
    GB_CUDA(cudaMemset(p->d_state0, 0, p->padded * 4));
  }
  if (need1) {
    GB_CUDA(cudaMalloc((void**)&p->d_state1, p->padded * 4));
    GB_CUDA(cudaMemset(p->d_state1, 0, p->padded * 4));
  }

  p->d_samp = nullptr;
  if (params->codec == GB_CODEC_DGC) GB_CUDA(cudaMalloc((void**)&p->d_samp, (size_t)(p->h_kbase[nseg] + 1) * 4));

  // ---- wire slot
  if (sparse) {
    // per-tile {begin, count}; TopK carries a second range per tile (threshold-bin picks, k1_topk2.cu)
    p->hdr_bytes = ((size_t)(p->ntiles + 1) * 2 * 4 * (params->codec == GB_CODEC_TOPK ? 2 : 1) + 255) & ~size_t(255);
    size_t per = ((size_t)p->ktotal * 4 + 255) & ~size_t(255);
    p->slot_bytes = p->hdr_bytes + (params->codec == GB_CODEC_RANDOMK ? per : 2 * per);
    if (params->codec == GB_CODEC_ADAQ) p->slot_bytes = p->hdr_bytes + aq_means_bytes(nseg) + per;   // indices only
  } else if (params->codec == GB_CODEC_INCEPTIONN) {
    p->hdr_bytes = in_hdr_bytes(p->ntiles);
    p->slot_bytes = p->hdr_bytes + in_tag_bytes(p->padded) + (((size_t)p->ktotal * 4 + 255) & ~size_t(255));
  } else {
    p->hdr_bytes = params->codec == GB_CODEC_SKETCH ? sk_header_bytes(nseg, params->quantum_num)
                                                    : (((size_t)nseg * 4 * 4 + 255) & ~size_t(255));
    p->slot_bytes = p->hdr_bytes + (((size_t)p->padded * p->bits / 8 + 16 + 255) & ~size_t(255));   // + 16: the 5-bit decoder reads one word ahead
  }
  p->channel2 = 0;
  p->mid_epoch = 0;
  p->dense2 = (params->codec == GB_CODEC_NONE || params->codec == GB_CODEC_FP16) && params->memory == GB_MEM_NONE &&
              c->world > 1 && !getenv("GRACE_B200_NO_K0");   // with a residual memory the K2 path keeps the state
  if (p->dense2) p->slot_bytes = k0_slot_bytes(p->padded, c->world, p->bits);
  p->slot_bytes = gb_guarded(c, p->slot_bytes);
  if (gb_comm_alloc_channel(c, p->slot_bytes, &p->channel, &p->chan_off)) return -1;
  if (p->dense2) {
    // second flag channel for the gather phase (no payload of its own)
    size_t unused;
    if (gb_comm_alloc_channel(c, 0, &p->channel2, &unused)) return -1;
  }
  p->presend_done = false;
  p->chan2_off = 0;
  p->slot2_bytes = 0;
  if (params->memory == GB_MEM_DGC && params->clip) {
    // every rank's per-tensor sum(g^2) travels through a small channel of its own
    p->slot2_bytes = gb_guarded(c, (size_t)nseg * 4);
    if (gb_comm_alloc_channel(c, p->slot2_bytes, &p->channel2, &p->chan2_off)) return -1;
  }
  p->slot_bytes = (p->slot_bytes + 255) & ~size_t(255);
  p->par_stride = (size_t)c->world * p->slot_bytes;
  if (params->codec == GB_CODEC_SKETCH && sk_init(p)) return -1;
  if (params->codec == GB_CODEC_ADAQ && aq_init(p)) return -1;
  if (params->codec == GB_CODEC_TOPK && !getenv("GRACE_B200_TOPK_V1") && tk_init(p)) return -1;
  GB_CUDA(cudaDeviceSynchronize());
  return 0;
  };
  if (build()) {
    gb_plan_destroy(p);
    return -1;
  }
  *out = p;
  return 0;
}

extern "C" int gb_plan_destroy(gb_plan* p) {
  if (!p) return 0;
  cudaDeviceSynchronize();
  sk_free(p);
  aq_free(p);
  tk_free(p);
  cudaFree(p->d_tables);
  cudaFree(p->d_state0);
  cudaFree(p->d_state1);
  cudaFree(p->d_samp);
  delete p;
  return 0;
}

extern "C" int64_t gb_plan_padded_numel(gb_plan* p) { return p->padded; }

extern "C" int gb_plan_offsets(gb_plan* p, int64_t* out) {
  for (int s = 0; s < p->nseg; ++s) out[s] = p->h_off[s];
  return 0;
}

extern "C" int64_t gb_plan_wire_bytes(gb_plan* p) { return (int64_t)p->slot_bytes; }

extern "C" int gb_plan_state(gb_plan* p, int which, void** d_ptr, size_t* bytes) {
  float* s = which == 0 ? p->d_state0 : p->d_state1;
  *d_ptr = s;
  *bytes = s ? (size_t)p->padded * 4 : 0;
  return 0;
}
extern "C" int gb_plan_state_read(gb_plan* p, int which, void* d_dst, void* stream) {
  float* s = which == 0 ? p->d_state0 : p->d_state1;
  GB_CHECK(s, "plan has no state buffer %d", which);
  GB_CUDA(cudaMemcpyAsync(d_dst, s, (size_t)p->padded * 4, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return 0;
}
extern "C" int gb_plan_state_write(gb_plan* p, int which, const void* d_src, void* stream) {
  float* s = which == 0 ? p->d_state0 : p->d_state1;
  GB_CHECK(s, "plan has no state buffer %d", which);
  GB_CUDA(cudaMemcpyAsync(s, d_src, (size_t)p->padded * 4, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return 0;
}
// the device copy is authoritative: replays of a captured graph advance it without the host
extern "C" int gb_plan_get_step(gb_plan* p, uint32_t* step) {
  GB_CUDA(cudaMemcpy(step, &p->d_ctl->next_step, 4, cudaMemcpyDeviceToHost));
  p->step = *step;
  return 0;
}
extern "C" int gb_plan_set_step(gb_plan* p, uint32_t step) {
  p->step = step;
  GB_CUDA(cudaMemcpy(&p->d_ctl->next_step, &step, 4, cudaMemcpyHostToDevice));
  return 0;
}
extern "C" int gb_plan_launches(gb_plan* p, int64_t* out) { *out = p->launches; return 0; }

__global__ void gb_tick_kernel(StepCtl* c) {
  c->epoch += 1;
  c->step = c->next_step;
  c->next_step += 1;
}
void gb_launch_tick(StepCtl* ctl, cudaStream_t st) { gb_tick_kernel<<<1, 1, 0, st>>>(ctl); }

int k1_dgc_presend(gb_plan* p, const float* d_grad, cudaStream_t st);   // k1_sparse.cu

extern "C" int gb_plan_presend(gb_plan* p, const float* d_grad, void* stream) {
  if (!p->slot2_bytes || p->presend_done) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  gb_launch_tick(p->d_ctl, st);
  p->launches++;
  if (k1_dgc_presend(p, d_grad, st)) return -1;
  p->presend_done = true;
  return 0;
}

extern "C" int gb_plan_send(gb_plan* p, const float* d_grad, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (p->slot2_bytes && !p->presend_done && gb_plan_presend(p, d_grad, stream)) return -1;
  if (!p->topk2 && !p->presend_done) {   // the two-pass TopK advances the step block in its first kernel
    gb_launch_tick(p->d_ctl, st);
    p->launches++;
  }
  p->presend_done = false;
  int rc = p->prm.codec == GB_CODEC_SKETCH ? sk_send(p, d_grad, st) : p->prm.codec == GB_CODEC_ADAQ ? aq_send(p, d_grad, st) : p->prm.codec == GB_CODEC_INCEPTIONN ? in_send(p, d_grad, st) : p->dense2 ? k0_send(p, d_grad, st) : gb_codec_is_sparse(p->prm.codec) ? k1_send(p, d_grad, st) : k2_send(p, d_grad, st);
  if (rc) return rc;
  p->step++;
  return 0;
}

extern "C" int gb_plan_recv(gb_plan* p, float* d_out, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  GB_CHECK(p->epoch > 0, "gb_plan_recv before any gb_plan_send");   // host mirror; counts captured sends too
  if (p->prm.codec == GB_CODEC_SKETCH) return sk_recv(p, d_out, st);
  if (p->prm.codec == GB_CODEC_ADAQ) return aq_recv(p, d_out, st);
  if (p->prm.codec == GB_CODEC_INCEPTIONN) return in_recv(p, d_out, st);
  if (p->dense2) return k0_recv(p, d_out, st);
  return gb_codec_is_sparse(p->prm.codec) ? k1_recv(p, d_out, st) : k2_recv(p, d_out, st);
}

// One warp polls the peers' flags of the current exchange.  Enqueued in front of the receive kernels, it keeps their
// (large) grids from occupying SM slots while a slower rank is still sending -- the receive side can then be queued
// right behind its send and overlap the rest of backward (the receive kernels' own waits pass immediately).
__global__ void gb_wait_kernel(PeerPtrs peers, int channel, const StepCtl* ctl, uint32_t* status, uint64_t timeout_ns) {
  const uint32_t epoch = ctl->epoch;
  if (threadIdx.x < peers.world)
    wait_flag(gb_flag(peers.base[peers.rank], channel, epoch & 1, threadIdx.x), epoch, status, timeout_ns);
}
void gb_launch_wait(gb_comm* c, int channel, const StepCtl* ctl, cudaStream_t st) {
  gb_wait_kernel<<<1, 32, 0, st>>>(gb_comm_peers(c), channel, ctl, c->status_dev, c->timeout_ns);
}

extern "C" int gb_plan_wait(gb_plan* p, void* stream) {
  if (p->comm->world == 1) return 0;
  GB_CHECK(p->epoch > 0, "gb_plan_wait before any gb_plan_send");
  gb_launch_wait(p->comm, p->channel, p->d_ctl, (cudaStream_t)stream);
  p->launches++;
  GB_CUDA(cudaGetLastError());
  return 0;
}

// Optional middle phase (only the dense two-shot path has one): lets a caller enqueue every
// bucket's reduce before any bucket's gather.  gb_plan_recv runs it itself when skipped.
extern "C" int gb_plan_mid(gb_plan* p, void* stream) {
  if (!p->dense2) return 0;
  GB_CHECK(p->epoch > 0, "gb_plan_mid before any gb_plan_send");
  return k0_mid(p, (cudaStream_t)stream);
}
