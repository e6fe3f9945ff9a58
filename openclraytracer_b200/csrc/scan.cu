// scan.cu — single-pass decoupled-lookback INCLUSIVE scan of uint32.
// Replaces PrefixSum::computePrefixSum (R/src/CLRayTracerEngine/PrefixSum.cpp:94-133; kernels
// R/src/CLRayTracerEngine/PrefixSumKernels.cl:29-106 — a multi-level Blelloch scan + uniform add, one
// synchronous launch pair per level).  Semantics kept: inclusive, out[i] = in[0]+...+in[i]; in may alias out.
//
// HBM roofline: 8 B per element (4 read + 4 written), one pass.  Layout: each CTA owns a tile of
// THREADS*IPT consecutive elements, loaded as 16-byte vectors (blocked arrangement: thread t owns elements
// [t*IPT, t*IPT+IPT) so the serial part needs no transposition), warp shuffles for the intra-CTA scan, and a
// 64-bit {flag,value} descriptor per tile for the look-back (one relaxed 8-byte store publishes both).
// Tile ids come from an atomic ticket so a CTA only ever waits on CTAs that already started.
#include "common.cuh"

namespace {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_IPT = 16;                 // 4 x uint4 per thread
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_IPT;

constexpr unsigned long long FLAG_AGG = 1ull << 32, FLAG_PREFIX = 2ull << 32;

__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
  unsigned long long v; asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}

__global__ void __launch_bounds__(SCAN_THREADS)
scan_lookback_kernel(const uint32_t* in, uint32_t* out, uint32_t n,   /* in may alias out */
                     unsigned long long* __restrict__ status, uint32_t* __restrict__ ticket) {
  __shared__ uint32_t s_tile;
  __shared__ uint32_t s_warp[SCAN_THREADS / 32];
  __shared__ uint32_t s_prefix;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_tile = atomicAdd(ticket, 1u);
  __syncthreads();
  const uint32_t tile = s_tile;
  const size_t base = (size_t)tile * SCAN_TILE + (size_t)tid * SCAN_IPT;

  uint32_t v[SCAN_IPT];
  if (base + SCAN_IPT <= n && (((uintptr_t)in & 15) == 0)) {
    const uint4* p = reinterpret_cast<const uint4*>(in + base);
#pragma unroll
    for (int k = 0; k < SCAN_IPT / 4; k++) { uint4 q = p[k]; v[4*k] = q.x; v[4*k+1] = q.y; v[4*k+2] = q.z; v[4*k+3] = q.w; }
  } else {
#pragma unroll
    for (int k = 0; k < SCAN_IPT; k++) v[k] = (base + k < n) ? in[base + k] : 0u;
  }
  // thread-serial inclusive scan
#pragma unroll
  for (int k = 1; k < SCAN_IPT; k++) v[k] += v[k - 1];
  // warp scan of thread totals
  uint32_t tot = v[SCAN_IPT - 1], incl = tot;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += t; }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  uint32_t warp_excl = 0, tile_total = 0;
#pragma unroll
  for (int w = 0; w < SCAN_THREADS / 32; w++) { uint32_t t = s_warp[w]; if (w < warp) warp_excl += t; tile_total += t; }
  const uint32_t thread_excl = warp_excl + incl - tot;

  // decoupled look-back by warp 0: lanes inspect 32 predecessors at a time
  if (warp == 0) {
    if (lane == 0) st_relaxed_u64(status + tile, (tile == 0 ? FLAG_PREFIX : FLAG_AGG) | tile_total);
    uint32_t excl = 0;
    if (tile > 0) {
      int look = (int)tile - 1 - lane;
      while (true) {
        unsigned long long s = 0;
        bool have;
        do {
          s = look >= 0 ? ld_relaxed_u64(status + look) : FLAG_PREFIX;   // before tile 0: prefix 0
          have = (s >> 32) != 0;
        } while (!__all_sync(0xffffffffu, have));
        const unsigned pmask = __ballot_sync(0xffffffffu, (s >> 32) == 2);
        // sum the values of lanes up to and including the first PREFIX lane (lanes are ordered newest -> oldest)
        const int first = pmask ? __ffs(pmask) - 1 : 32;
        uint32_t val = (lane <= first) ? (uint32_t)s : 0u;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) val += __shfl_xor_sync(0xffffffffu, val, d);
        excl += val;
        if (pmask) break;
        look -= 32;
      }
      if (lane == 0) st_relaxed_u64(status + tile, FLAG_PREFIX | (unsigned long long)(uint32_t)(excl + tile_total));
    }
    if (lane == 0) s_prefix = excl;
  }
  __syncthreads();
  const uint32_t add = s_prefix + thread_excl;
  if (base + SCAN_IPT <= n && (((uintptr_t)out & 15) == 0)) {
    uint4* p = reinterpret_cast<uint4*>(out + base);
#pragma unroll
    for (int k = 0; k < SCAN_IPT / 4; k++) p[k] = make_uint4(v[4*k] + add, v[4*k+1] + add, v[4*k+2] + add, v[4*k+3] + add);
  } else {
#pragma unroll
    for (int k = 0; k < SCAN_IPT; k++) if (base + k < n) out[base + k] = v[k] + add;
  }
}

} // namespace

int scan_inclusive(clrt_ctx* ctx, const uint32_t* in, uint32_t* out, size_t n) {
  if (n == 0) return 0;
  if (n > 0xFFFFFFFFull) CLRT_FAIL("scan: n = %zu exceeds 2^32-1", n);
  const size_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
  void* scratch = nullptr;
  const size_t bytes = tiles * 8 + 16;
  CLRT_TRY(clrt_ctx_scratch(ctx, bytes, &scratch));
  CLRT_CUDA(cudaMemsetAsync(scratch, 0, bytes, ctx->stream));
  uint32_t* ticket = (uint32_t*)scratch;
  unsigned long long* status = (unsigned long long*)((uint8_t*)scratch + 16);
  scan_lookback_kernel<<<(unsigned)tiles, SCAN_THREADS, 0, ctx->stream>>>(in, out, (uint32_t)n, status, ticket);
  CLRT_CHECK_LAUNCH();
  return 0;
}

extern "C" int clrt_scan_inclusive_u32(clrt_ctx* ctx, const void* dptr_in, void* dptr_out, size_t n) {
  if (!ctx) CLRT_FAIL("clrt_scan_inclusive_u32: ctx is NULL");
  if (n && (!dptr_in || !dptr_out)) CLRT_FAIL("clrt_scan_inclusive_u32: NULL buffer");
  if (((uintptr_t)dptr_in & 3) || ((uintptr_t)dptr_out & 3)) CLRT_FAIL("clrt_scan_inclusive_u32: buffers must be 4-byte aligned");
  CLRT_CUDA(cudaSetDevice(ctx->device));
  return scan_inclusive(ctx, (const uint32_t*)dptr_in, (uint32_t*)dptr_out, n);
}
