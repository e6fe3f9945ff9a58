// scan.cu — single-pass decoupled-lookback INCLUSIVE scan of uint32.
// Replaces PrefixSum::computePrefixSum (R/src/CLRayTracerEngine/PrefixSum.cpp:94-133; kernels
// R/src/CLRayTracerEngine/PrefixSumKernels.cl:29-106 — a multi-level Blelloch scan + uniform add, one
// synchronous launch pair per level).  Semantics kept: inclusive, out[i] = in[0]+...+in[i]; in may alias out.
//
// HBM roofline: 8 B per element (4 read + 4 written), one pass.  Each CTA owns a tile of 16384 consecutive elements as
// SCAN_VEC sub-tiles of THREADS x 4: thread t holds the uint4 at vector index k*THREADS + t of every sub-tile k, so each warp-wide
// 16-byte load / store covers 512 contiguous bytes (round 1 loaded 4 x uint4 per thread at a 64-byte lane stride: every
// instruction touched 32 half-used sectors, 0.33-0.45 of the HBM peak).  The CTA scan is two shuffle levels: a warp scan of the uint4
// sums per sub-tile, then every warp scans the SCAN_VEC x WARPS warp totals from shared memory.  One 64-bit {generation|flag,
// value} descriptor per tile drives the look-back (a single relaxed 8-byte store publishes both).  The grid is persistent (every CTA
// resident, tiles dealt round-robin), so a CTA only ever waits on CTAs that are running.  No reset launch: descriptors carry the
// call's GENERATION number, so a call is exactly one kernel.
#include "common.cuh"

namespace {

#ifndef CLRT_SCAN_THREADS
#define CLRT_SCAN_THREADS 512
#endif
#ifndef CLRT_SCAN_DEBUG_NOLOOK
#define CLRT_SCAN_DEBUG_NOLOOK 0     /* experiment only: skip the look-back (wrong results) to see what the rest of the kernel costs */
#endif
#ifndef CLRT_SCAN_SLEEP_NS
#define CLRT_SCAN_SLEEP_NS 200
#endif
#ifndef CLRT_SCAN_LOOK
#define CLRT_SCAN_LOOK 1
#endif
constexpr int SCAN_THREADS = CLRT_SCAN_THREADS;
constexpr int SCAN_WARPS = SCAN_THREADS / 32;
constexpr int SCAN_VEC = 4;                                  // uint4 per thread
constexpr int SCAN_SUB = SCAN_THREADS * 4;                   // elements per sub-tile
constexpr int SCAN_TILE = SCAN_SUB * SCAN_VEC;               // 4096 at 256 threads
constexpr int SCAN_LOOK = CLRT_SCAN_LOOK;                    // descriptors per lane and look-back round
constexpr int SCAN_EPL = SCAN_VEC * SCAN_WARPS / 32;         // (sub-tile, warp) totals per lane in the second-level scan
static_assert(SCAN_EPL == 1 || SCAN_EPL == 2 || SCAN_EPL == 4, "256, 512 or 1024 threads");
constexpr unsigned FLAG_AGG = 1u, FLAG_PREFIX = 2u;          // descriptor: hi = generation << 2 | flag, lo = value

__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
  unsigned long long v; asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}

template <bool ALIGNED>
__device__ __forceinline__ void scan_load_tile(const uint32_t* in, uint32_t n, uint32_t tile, int tid, uint4 (&v)[SCAN_VEC]) {
  const size_t tile_base = (size_t)tile * SCAN_TILE;
  const bool full = tile_base + SCAN_TILE <= n;
#pragma unroll
  for (int k = 0; k < SCAN_VEC; k++) {
    const size_t e = tile_base + (size_t)k * SCAN_SUB + 4 * (size_t)tid;
    if (ALIGNED && full) v[k] = __ldcs(reinterpret_cast<const uint4*>(in + e));
    else {
      v[k].x = e     < n ? in[e]     : 0u; v[k].y = e + 1 < n ? in[e + 1] : 0u;
      v[k].z = e + 2 < n ? in[e + 2] : 0u; v[k].w = e + 3 < n ? in[e + 3] : 0u;
    }
  }
}

// Persistent CTAs (grid = resident capacity, CTA b owns tiles b, b + grid, ...), software pipelined: the loads of the CTA's NEXT tile
// are issued before the look-back of the current one, so reads stay in flight while a tile waits for its predecessors and while
// its results are stored.  (One tile per CTA — rounds 1 and 2a — alternated a read phase, a waiting phase and a write phase per
// CTA and reached half of the copy bandwidth whatever the tile size, ticketing or look-back width: profiles/r02_scan_variants.md.)
// Every tile's predecessors belong to CTAs that are resident, so the look-back cannot deadlock.
template <bool ALIGNED>
__global__ void __launch_bounds__(SCAN_THREADS)
scan_lookback_kernel(const uint32_t* in, uint32_t* out, uint32_t n,   /* in may alias out */
                     unsigned long long* __restrict__ status, uint32_t tiles, uint32_t gen) {
  __shared__ __align__(16) uint32_t s_part[2][SCAN_VEC * SCAN_WARPS];    // [sub-tile][warp] totals, double buffered over tiles
  __shared__ uint32_t s_prefix[2];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  uint32_t tile = blockIdx.x;
  uint4 v[SCAN_VEC], vn[SCAN_VEC];
  scan_load_tile<ALIGNED>(in, n, tile, tid, v);
  for (int it = 0; tile < tiles; it ^= 1) {
    const uint32_t next = tile + gridDim.x;
    // In place (in == out) the next tile's loads precede this tile's stores in program order, and they touch other addresses.
    if (next < tiles) scan_load_tile<ALIGNED>(in, n, next, tid, vn);
    const size_t tile_base = (size_t)tile * SCAN_TILE;
    const bool full = tile_base + SCAN_TILE <= n;
    // inclusive scan inside each uint4, then a warp scan of the uint4 totals per sub-tile
    uint32_t wexcl[SCAN_VEC];
#pragma unroll
    for (int k = 0; k < SCAN_VEC; k++) {
      v[k].y += v[k].x; v[k].z += v[k].y; v[k].w += v[k].z;
      uint32_t incl = v[k].w;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += t; }
      wexcl[k] = incl - v[k].w;
      if (lane == 31) s_part[it][k * SCAN_WARPS + warp] = incl;
    }
    __syncthreads();
    // every warp scans the SCAN_VEC x WARPS (sub-tile, warp) totals: lane holds entries EPL*lane .. EPL*lane+EPL-1 (entry = k * WARPS + warp, tile order)
    uint32_t tile_total;
    {
      uint32_t pr[SCAN_EPL];
      uint32_t sum = 0;
#pragma unroll
      for (int q = 0; q < SCAN_EPL; q++) { pr[q] = s_part[it][SCAN_EPL * lane + q]; sum += pr[q]; }
      uint32_t incl = sum;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += t; }
      tile_total = __shfl_sync(0xffffffffu, incl, 31);
      // exclusive prefix of the lane's entry number (warp % EPL): entry k*WARPS + warp sits in lane (k*WARPS + warp)/EPL at that sub-index
      uint32_t mine = incl - sum;
#pragma unroll
      for (int q = 0; q < SCAN_EPL - 1; q++) if (q < (warp % SCAN_EPL)) mine += pr[q];
#pragma unroll
      for (int k = 0; k < SCAN_VEC; k++) wexcl[k] += __shfl_sync(0xffffffffu, mine, (k * SCAN_WARPS + warp) / SCAN_EPL);
    }
    // Decoupled look-back by warp 0.  A round inspects 32 x SCAN_LOOK predecessors.
    if (warp == 0) {
      const unsigned long long tag = (unsigned long long)(gen << 2) << 32;
      if (lane == 0) st_relaxed_u64(status + tile, tag | ((unsigned long long)(tile == 0 ? FLAG_PREFIX : FLAG_AGG) << 32) | tile_total);
      uint32_t excl = 0;
      if (tile > 0 && !CLRT_SCAN_DEBUG_NOLOOK) {
        long long look = (long long)tile - 1 - (long long)lane * SCAN_LOOK;       // newest descriptor of this lane; lanes and j run newest -> oldest
        while (true) {
          uint32_t val[SCAN_LOOK]; unsigned flag[SCAN_LOOK];
          bool all;
          bool first_poll = true;
          do {
            // Back-off between polls: hundreds of resident CTAs spinning on neighbouring descriptors (16 to a cache line) hammer one
            // L2 slice and slow the publishing stores as well as the streaming traffic — without any waiting the kernel ran at half the
            // rate it reaches with the look-back switched off (profiles/r02_scan_variants.md).  CUB's delay constructors do the same.
            if (!first_poll && CLRT_SCAN_SLEEP_NS) __nanosleep(CLRT_SCAN_SLEEP_NS);
            first_poll = false;
            all = true;
#pragma unroll
            for (int j = 0; j < SCAN_LOOK; j++) {
              if (look - j >= 0) {
                const unsigned long long sd = ld_relaxed_u64(status + (look - j));
                const uint32_t hi = (uint32_t)(sd >> 32);
                flag[j] = (hi >> 2) == gen ? (hi & 3u) : 0u; val[j] = (uint32_t)sd;
              } else { flag[j] = FLAG_PREFIX; val[j] = 0u; }                        // before tile 0: prefix 0
              all = all && flag[j] != 0u;
            }
          } while (!__all_sync(0xffffffffu, all));
          // the lane's sum up to and including its first PREFIX descriptor
          uint32_t sum = 0; bool closed = false;
#pragma unroll
          for (int j = 0; j < SCAN_LOOK; j++) { if (!closed) sum += val[j]; closed = closed || flag[j] == FLAG_PREFIX; }
          const unsigned pmask = __ballot_sync(0xffffffffu, closed);
          const int first = pmask ? __ffs(pmask) - 1 : 32;
          uint32_t v2 = (lane <= first) ? sum : 0u;
#pragma unroll
          for (int d = 16; d > 0; d >>= 1) v2 += __shfl_xor_sync(0xffffffffu, v2, d);
          excl += v2;
          if (pmask) break;
          look -= 32 * SCAN_LOOK;
        }
        if (lane == 0) st_relaxed_u64(status + tile, tag | ((unsigned long long)FLAG_PREFIX << 32) | (unsigned long long)(uint32_t)(excl + tile_total));
      }
      if (lane == 0) s_prefix[it] = excl;
    }
    __syncthreads();
    const uint32_t base = s_prefix[it];
#pragma unroll
    for (int k = 0; k < SCAN_VEC; k++) {
      const size_t e = tile_base + (size_t)k * SCAN_SUB + 4 * (size_t)tid;
      const uint32_t add = base + wexcl[k];
      const uint4 r = make_uint4(v[k].x + add, v[k].y + add, v[k].z + add, v[k].w + add);
      if (ALIGNED && full) __stcs(reinterpret_cast<uint4*>(out + e), r);
      else {
        if (e < n) out[e] = r.x;
        if (e + 1 < n) out[e + 1] = r.y;
        if (e + 2 < n) out[e + 2] = r.z;
        if (e + 3 < n) out[e + 3] = r.w;
      }
    }
#pragma unroll
    for (int k = 0; k < SCAN_VEC; k++) v[k] = vn[k];
    tile = next;
  }
}

} // namespace

// Look-back state of the context: [16 B pad | status u64 x tiles], zeroed when (re)allocated and never reset afterwards.
// A descriptor is valid only when it carries the generation of the running call.
int scan_inclusive(clrt_ctx* ctx, const uint32_t* in, uint32_t* out, size_t n) {
  if (n == 0) return 0;
  if (n > 0xFFFFFFFFull) CLRT_FAIL("scan: n = %zu exceeds 2^32-1", n);
  const size_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
  const size_t bytes = tiles * 8 + 16;
  if (bytes > ctx->scan_state_bytes || ctx->scan_gen >= (1u << 30) - 1u) {
    CLRT_CUDA(cudaStreamSynchronize(ctx->stream));
    if (bytes > ctx->scan_state_bytes) CLRT_TRY(clrt_grow(&ctx->scan_state, &ctx->scan_state_bytes, bytes + bytes / 2));
    CLRT_CUDA(cudaMemsetAsync(ctx->scan_state, 0, ctx->scan_state_bytes, ctx->stream));
    ctx->scan_gen = 0;
  }
  const uint32_t gen = ++ctx->scan_gen;
  unsigned long long* status = (unsigned long long*)((uint8_t*)ctx->scan_state + 16);
  const bool aligned = (((uintptr_t)in | (uintptr_t)out) & 15) == 0;
  static int per_sm[2] = {0, 0};
  if (!per_sm[aligned]) {
    if (aligned) CLRT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm[1], scan_lookback_kernel<true>, SCAN_THREADS, 0));
    else CLRT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm[0], scan_lookback_kernel<false>, SCAN_THREADS, 0));
    if (per_sm[aligned] < 1) per_sm[aligned] = 1;
  }
  const unsigned grid = (unsigned)min(tiles, (size_t)ctx->sm_count * per_sm[aligned]);     // all CTAs resident: the look-back never waits on an unscheduled tile
  if (aligned) scan_lookback_kernel<true><<<grid, SCAN_THREADS, 0, ctx->stream>>>(in, out, (uint32_t)n, status, (uint32_t)tiles, gen);
  else scan_lookback_kernel<false><<<grid, SCAN_THREADS, 0, ctx->stream>>>(in, out, (uint32_t)n, status, (uint32_t)tiles, gen);
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
