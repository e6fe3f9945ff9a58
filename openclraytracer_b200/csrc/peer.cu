// Frame exchange over peer memory (the reference is single-device, OpenCLRayTracerDemo.cpp:271-326 renders one
// frame on one OpenCL device).  One process per GPU; every process exports its raster-ordered framebuffers and a block of flag
// words (cudaIpc), opens those of its peers, and then a frame needs no collective: each GPU stores the pixels of its own tiles
// straight into the framebuffer of EVERY GPU over NVLink — from the shading kernel itself (clrt_render_stripe_peers) or by one
// push pass over its stripe buffer (clrt_frame_push) — raises a per-frame-slot "ready" word on every peer and waits until the
// other GPUs' words arrived.  "ack" words give the back-pressure before a slot's framebuffer is overwritten.
#include "common.cuh"

namespace {

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v; asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t;
}

// own stripe (tile-major, 256 px per tile, local tile lt = global tile rank + lt*nranks) -> raster of every GPU
__global__ void __launch_bounds__(256)
push_kernel(const uint32_t* __restrict__ stripe, uint32_t resX, uint32_t resY, uint32_t rank, uint32_t nranks, const PeerTargets peers) {
  const TileGrid g = make_tile_grid(resX, resY);
  const uint32_t gt = rank + blockIdx.x * nranks, t = threadIdx.x;
  if (gt >= g.tiles) return;
  const uint32_t x = (gt % g.tilesX) * CLRT_TILE_W + (t & 31u), y = (gt / g.tilesX) * CLRT_TILE_H + (t >> 5);
  if (x >= resX || y >= resY) return;
  const uint32_t v = __ldcs(stripe + (size_t)blockIdx.x * CLRT_TILE_PIXELS + t);
#pragma unroll
  for (uint32_t p = 0; p < 8; p++) if (p < peers.n) peers.p[p][y * resX + x] = v;   // a warp writes one 128-byte row per peer
}

// Runs after the kernels that wrote the pixels (stream order): their stores were performed when they completed; the fence +
// release store orders this word behind them for an observer on any GPU.
__global__ void signal_kernel(const PeerTargets flags, uint32_t word, uint32_t value) {
  __threadfence_system();
  if (threadIdx.x < flags.n) st_release_sys(flags.p[threadIdx.x] + word, value);
}

// Spins until `n` consecutive words reached `value`.  A peer that died must not hang the box: after timeout_ns the kernel traps,
// which surfaces as a CUDA error at the caller's next synchronisation.
__global__ void wait_kernel(const uint32_t* __restrict__ words, uint32_t n, uint32_t value, unsigned long long timeout_ns) {
  if (threadIdx.x >= n) return;
  const unsigned long long t0 = global_ns();
  while ((int32_t)(ld_acquire_sys(words + threadIdx.x) - value) < 0) {
    __nanosleep(200);
    if (global_ns() - t0 > timeout_ns) __trap();
  }
}

// ready[rank] = seq on every peer, wait for everybody's ready word here, then (ack_base != ~0u) ack[rank] = seq on every peer.
__global__ void exchange_kernel(const PeerTargets flags, const uint32_t* __restrict__ own, uint32_t ready_base, uint32_t ack_base,
                                uint32_t rank, uint32_t seq, unsigned long long timeout_ns) {
  const uint32_t t = threadIdx.x;
  __threadfence_system();
  if (t < flags.n) {
    st_release_sys(flags.p[t] + ready_base + rank, seq);
    const unsigned long long t0 = global_ns();
    while ((int32_t)(ld_acquire_sys(own + ready_base + t) - seq) < 0) {
      __nanosleep(100);
      if (global_ns() - t0 > timeout_ns) __trap();
    }
  }
  __syncthreads();
  if (ack_base != ~0u && t < flags.n) st_release_sys(flags.p[t] + ack_base + rank, seq);
}

int fill_targets(PeerTargets& t, void* const* ptrs, uint32_t n, const char* what) {
  if (!ptrs || !n || n > 8) CLRT_FAIL("%s: %u peers (1..8)", what, n);
  t.n = n;
  for (uint32_t i = 0; i < 8; i++) t.p[i] = i < n ? (uint32_t*)ptrs[i] : nullptr;
  for (uint32_t i = 0; i < n; i++) if (!t.p[i]) CLRT_FAIL("%s: peer %u is NULL", what, i);
  return 0;
}

} // namespace

// internal (accel.cu: clrt_render_frame_peers)
int peer_launch_wait(cudaStream_t stream, const uint32_t* words, uint32_t n, uint32_t value) {
  wait_kernel<<<1, 32, 0, stream>>>(words, n, value, 30000000000ull);
  CLRT_CHECK_LAUNCH();
  return 0;
}
int peer_launch_exchange(cudaStream_t stream, const PeerTargets& flags, const uint32_t* own, uint32_t ready_base, uint32_t ack_base,
                         uint32_t rank, uint32_t seq) {
  exchange_kernel<<<1, 32, 0, stream>>>(flags, own, ready_base, ack_base, rank, seq, 30000000000ull);
  CLRT_CHECK_LAUNCH();
  return 0;
}
int peer_launch_signal(cudaStream_t stream, const PeerTargets& flags, uint32_t word, uint32_t value) {
  signal_kernel<<<1, 32, 0, stream>>>(flags, word, value);
  CLRT_CHECK_LAUNCH();
  return 0;
}

extern "C" int clrt_peer_alloc(clrt_ctx* ctx, uint64_t bytes, void** out_dptr) {
  if (!ctx || !out_dptr || !bytes) CLRT_FAIL("clrt_peer_alloc: bad argument");
  CLRT_CUDA(cudaSetDevice(ctx->device));
  void* p = nullptr;
  CLRT_CUDA(cudaMalloc(&p, bytes));                      // a plain allocation: pool / VMM memory cannot be exported with cudaIpc
  CLRT_CUDA(cudaMemset(p, 0, bytes));
  CLRT_CUDA(cudaDeviceSynchronize());
  *out_dptr = p;
  return 0;
}
extern "C" int clrt_peer_free(clrt_ctx* ctx, void* dptr) {
  if (!ctx) CLRT_FAIL("clrt_peer_free: bad argument");
  CLRT_CUDA(cudaSetDevice(ctx->device));
  if (dptr) CLRT_CUDA(cudaFree(dptr));
  return 0;
}
extern "C" int clrt_peer_export(clrt_ctx* ctx, void* dptr, uint8_t* out_handle64) {
  if (!ctx || !dptr || !out_handle64) CLRT_FAIL("clrt_peer_export: bad argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == CLRT_PEER_HANDLE_BYTES, "handle size");
  CLRT_CUDA(cudaSetDevice(ctx->device));
  cudaIpcMemHandle_t h;
  CLRT_CUDA(cudaIpcGetMemHandle(&h, dptr));
  memcpy(out_handle64, &h, sizeof h);
  return 0;
}
extern "C" int clrt_peer_open(clrt_ctx* ctx, const uint8_t* handle64, void** out_dptr) {
  if (!ctx || !handle64 || !out_dptr) CLRT_FAIL("clrt_peer_open: bad argument");
  CLRT_CUDA(cudaSetDevice(ctx->device));
  cudaIpcMemHandle_t h; memcpy(&h, handle64, sizeof h);
  void* p = nullptr;
  CLRT_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *out_dptr = p;
  return 0;
}
extern "C" int clrt_peer_close(clrt_ctx* ctx, void* dptr) {
  if (!ctx) CLRT_FAIL("clrt_peer_close: bad argument");
  CLRT_CUDA(cudaSetDevice(ctx->device));
  if (dptr) CLRT_CUDA(cudaIpcCloseMemHandle(dptr));
  return 0;
}

extern "C" int clrt_frame_push(clrt_ctx* ctx, void* stream, const void* dptr_stripe, uint32_t resX, uint32_t resY, uint32_t rank,
                               uint32_t nranks, void* const* peer_rasters) {
  if (!ctx || !dptr_stripe || rank >= nranks) CLRT_FAIL("clrt_frame_push: bad argument");
  CLRT_CUDA(cudaSetDevice(ctx->device));
  PeerTargets t; CLRT_TRY(fill_targets(t, peer_rasters, nranks, "clrt_frame_push"));
  const TileGrid g = make_tile_grid(resX, resY);
  const uint32_t tiles = stripe_tile_count(g.tiles, rank, nranks);
  if (!tiles) return 0;
  push_kernel<<<tiles, 256, 0, (cudaStream_t)stream>>>((const uint32_t*)dptr_stripe, resX, resY, rank, nranks, t);
  CLRT_CHECK_LAUNCH();
  return 0;
}
extern "C" int clrt_peer_signal(clrt_ctx* ctx, void* stream, void* const* peer_words, uint32_t npeers, uint32_t word, uint32_t value) {
  if (!ctx) CLRT_FAIL("clrt_peer_signal: bad argument");
  CLRT_CUDA(cudaSetDevice(ctx->device));
  PeerTargets t; CLRT_TRY(fill_targets(t, peer_words, npeers, "clrt_peer_signal"));
  signal_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(t, word, value);
  CLRT_CHECK_LAUNCH();
  return 0;
}
extern "C" int clrt_peer_wait(clrt_ctx* ctx, void* stream, const void* dptr_words, uint32_t first_word, uint32_t n, uint32_t value,
                              uint32_t timeout_ms) {
  if (!ctx || !dptr_words || !n || n > 32) CLRT_FAIL("clrt_peer_wait: bad argument");
  CLRT_CUDA(cudaSetDevice(ctx->device));
  wait_kernel<<<1, 32, 0, (cudaStream_t)stream>>>((const uint32_t*)dptr_words + first_word, n, value,
                                                  (unsigned long long)(timeout_ms ? timeout_ms : 30000u) * 1000000ull);
  CLRT_CHECK_LAUNCH();
  return 0;
}
