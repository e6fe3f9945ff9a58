// scan.cu — single-pass INCLUSIVE scan of uint32.
// Replaces PrefixSum::computePrefixSum (R/src/CLRayTracerEngine/PrefixSum.cpp:94-133; kernels
// R/src/CLRayTracerEngine/PrefixSumKernels.cl:29-106 — a multi-level Blelloch scan + uniform add, one
// synchronous launch pair per level).  Semantics kept: inclusive, out[i] = in[0]+...+in[i]; in may alias out.
//
// HBM roofline: 8 B per element (4 read + 4 written), one pass, one launch.  A CTA owns tiles of THREADS x 4 x VEC consecutive
// elements as VEC sub-tiles of THREADS x 4: thread t holds the uint4 at vector index k*THREADS + t of every sub-tile k, so each
// warp-wide 16-byte load / store covers 512 contiguous bytes.  The CTA scan is two shuffle levels: a warp scan of the uint4 sums per
// sub-tile, then every warp scans the VEC x WARPS warp totals from shared memory.  The grid is persistent (every CTA resident, tiles
// dealt round-robin), so a CTA only ever waits on CTAs that are running.  One 64-bit {generation|flag, value} descriptor per tile
// carries the tile's aggregate between CTAs (a single relaxed 8-byte store publishes both).  No reset launch: descriptors carry the
// call's GENERATION number, so a call is exactly one kernel.
//
// Two kernels:
//   scan_window_kernel  (16-byte aligned buffers — every caller in this library): tiles staged through shared memory by bulk-async
//                       copies, prefix resolved one iteration late from a WINDOW of aggregates (below).  0.86 of the measured copy
//                       bandwidth at 2^28 elements (profiles/r02_experiments.md section 4).
//   scan_lookback_kernel (buffers that are only 4-byte aligned): register prefetch + classic decoupled look-back (round 2a's kernel).
#include "common.cuh"
#include <mutex>

namespace {

#ifndef CLRT_SCAN_THREADS
#define CLRT_SCAN_THREADS 1024
#endif
#ifndef CLRT_SCAN_DEBUG_NOLOOK
#define CLRT_SCAN_DEBUG_NOLOOK 0     /* experiment only: skip the look-back (wrong results) to see what the rest of the kernel costs */
#endif
#ifndef CLRT_SCAN_SLEEP_NS
#define CLRT_SCAN_SLEEP_NS 200
#endif
#ifndef CLRT_SCAN_STAGES
#define CLRT_SCAN_STAGES 3
#endif
#ifndef CLRT_SCAN_FORCE_VEC
#define CLRT_SCAN_FORCE_VEC 0        /* experiment only: 2 / 4 = fixed tile size whatever n */
#endif
constexpr int SCAN_THREADS = CLRT_SCAN_THREADS;
constexpr int SCAN_WARPS = SCAN_THREADS / 32;
constexpr int SCAN_SUB = SCAN_THREADS * 4;                   // elements per sub-tile
constexpr int SCAN_STAGES = CLRT_SCAN_STAGES;
constexpr unsigned FLAG_AGG = 1u, FLAG_PREFIX = 2u;          // descriptor: hi = generation << 2 | flag, lo = value
static_assert(SCAN_THREADS == 256 || SCAN_THREADS == 512 || SCAN_THREADS == 1024, "256, 512 or 1024 threads");

__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
  unsigned long long v; asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}

// ---- the steps of a tile, shared by both kernels ------------------------------------------------------------------------------
template <int VEC>
__device__ __forceinline__ void scan_load_tile(const uint32_t* in, uint32_t n, uint32_t tile, int tid, uint4 (&v)[VEC]) {
  const size_t tile_base = (size_t)tile * (SCAN_SUB * VEC);
#pragma unroll
  for (int k = 0; k < VEC; k++) {
    const size_t e = tile_base + (size_t)k * SCAN_SUB + 4 * (size_t)tid;
    v[k].x = e     < n ? in[e]     : 0u; v[k].y = e + 1 < n ? in[e + 1] : 0u;
    v[k].z = e + 2 < n ? in[e + 2] : 0u; v[k].w = e + 3 < n ? in[e + 3] : 0u;
  }
}

// CTA scan: inclusive scan inside each uint4, a warp scan of the uint4 totals per sub-tile, then every warp scans the VEC x WARPS
// (sub-tile, warp) totals.  Leaves v[] scanned inside each uint4, wexcl[k] = what precedes the thread's uint4 k inside the tile, and
// returns the tile's total.  One __syncthreads (after it, every thread has consumed its inputs of this tile).
template <int VEC>
__device__ __forceinline__ uint32_t scan_cta_tile(uint4 (&v)[VEC], uint32_t (&wexcl)[VEC], uint32_t* s_part, int lane, int warp) {
  constexpr int ENTRIES = VEC * SCAN_WARPS;                  // (sub-tile, warp) totals, tile order: entry = k * WARPS + warp
  constexpr int EPL = (ENTRIES + 31) / 32;                   // entries per lane in the second-level scan (lane holds EPL*lane ..)
#pragma unroll
  for (int k = 0; k < VEC; k++) {
    v[k].y += v[k].x; v[k].z += v[k].y; v[k].w += v[k].z;
    uint32_t incl = v[k].w;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += t; }
    wexcl[k] = incl - v[k].w;
    if (lane == 31) s_part[k * SCAN_WARPS + warp] = incl;
  }
  __syncthreads();
  uint32_t pr[EPL];
  uint32_t sum = 0;
#pragma unroll
  for (int q = 0; q < EPL; q++) { const int e = EPL * lane + q; pr[q] = e < ENTRIES ? s_part[e] : 0u; sum += pr[q]; }
  uint32_t incl = sum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += t; }
  const uint32_t tile_total = __shfl_sync(0xffffffffu, incl, 31);
  // This warp's entry of sub-tile k sits in lane e / EPL at sub-index e % EPL = warp % EPL (WARPS is a multiple of EPL): every lane
  // prepares the exclusive prefix of ITS entry number warp % EPL, one shuffle per sub-tile fetches it.
  static_assert(SCAN_WARPS % EPL == 0, "entry sub-index must not depend on the sub-tile");
  uint32_t mine = incl - sum;
#pragma unroll
  for (int q = 0; q < EPL - 1; q++) if (q < (warp % EPL)) mine += pr[q];
#pragma unroll
  for (int k = 0; k < VEC; k++) wexcl[k] += __shfl_sync(0xffffffffu, mine, (k * SCAN_WARPS + warp) / EPL);
  return tile_total;
}

template <bool ALIGNED, int VEC>
__device__ __forceinline__ void scan_store_tile(uint32_t* out, uint32_t n, uint32_t tile, int tid, const uint4 (&v)[VEC],
                                                const uint32_t (&wexcl)[VEC], uint32_t base) {
  const size_t tile_base = (size_t)tile * (SCAN_SUB * VEC);
  const bool full = tile_base + SCAN_SUB * VEC <= n;
#pragma unroll
  for (int k = 0; k < VEC; k++) {
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
}

// ---- kernel 1: buffers that are only 4-byte aligned ----------------------------------------------------------------------------------
// Decoupled look-back by ONE warp: publishes the tile's aggregate, sums the predecessors' descriptors (32 per round) back to the
// first inclusive prefix, publishes the tile's inclusive prefix; returns the tile's exclusive prefix.
__device__ __forceinline__ uint32_t scan_lookback(unsigned long long* __restrict__ status, uint32_t tile, uint32_t tile_total, uint32_t gen, int lane) {
  const unsigned long long tag = (unsigned long long)(gen << 2) << 32;
  if (lane == 0) st_relaxed_u64(status + tile, tag | ((unsigned long long)(tile == 0 ? FLAG_PREFIX : FLAG_AGG) << 32) | tile_total);
  uint32_t excl = 0;
  if (tile > 0 && !CLRT_SCAN_DEBUG_NOLOOK) {
    long long look = (long long)tile - 1 - (long long)lane;              // lanes run newest -> oldest
    while (true) {
      uint32_t val; unsigned flag;
      bool first_poll = true;
      do {
        // Back-off between polls: hundreds of resident CTAs spinning on neighbouring descriptors (16 to a cache line) hammer one
        // L2 slice and slow the publishing stores as well as the streaming traffic (profiles/r02_experiments.md section 4).
        if (!first_poll && CLRT_SCAN_SLEEP_NS) __nanosleep(CLRT_SCAN_SLEEP_NS);
        first_poll = false;
        if (look >= 0) {
          const unsigned long long sd = ld_relaxed_u64(status + look);
          const uint32_t hi = (uint32_t)(sd >> 32);
          flag = (hi >> 2) == gen ? (hi & 3u) : 0u; val = (uint32_t)sd;
        } else { flag = FLAG_PREFIX; val = 0u; }                          // before tile 0: prefix 0
      } while (!__all_sync(0xffffffffu, flag != 0u));
      const unsigned pmask = __ballot_sync(0xffffffffu, flag == FLAG_PREFIX);
      const int first = pmask ? __ffs(pmask) - 1 : 32;
      uint32_t v2 = (lane <= first) ? val : 0u;
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) v2 += __shfl_xor_sync(0xffffffffu, v2, d);
      excl += v2;
      if (pmask) break;
      look -= 32;
    }
    if (lane == 0) st_relaxed_u64(status + tile, tag | ((unsigned long long)FLAG_PREFIX << 32) | (unsigned long long)(uint32_t)(excl + tile_total));
  }
  return excl;
}

// Persistent CTAs (grid = resident capacity, CTA b owns tiles b, b + grid, ...), software pipelined through REGISTERS: the loads of
// the CTA's NEXT tile are issued before the look-back of the current one.  Every tile's predecessors belong to CTAs that are resident,
// so the look-back cannot deadlock.
constexpr int SCAN_VEC_UNALIGNED = 4;
__global__ void __launch_bounds__(SCAN_THREADS)
scan_lookback_kernel(const uint32_t* in, uint32_t* out, uint32_t n,   /* in may alias out */
                     unsigned long long* __restrict__ status, uint32_t tiles, uint32_t gen) {
  constexpr int VEC = SCAN_VEC_UNALIGNED;
  __shared__ __align__(16) uint32_t s_part[2][VEC * SCAN_WARPS];        // [sub-tile][warp] totals, double buffered over tiles
  __shared__ uint32_t s_prefix[2];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  uint32_t tile = blockIdx.x;
  uint4 v[VEC], vn[VEC];
  scan_load_tile<VEC>(in, n, tile, tid, v);
  for (int it = 0; tile < tiles; it ^= 1) {
    const uint32_t next = tile + gridDim.x;
    // In place (in == out) the next tile's loads precede this tile's stores in program order, and they touch other addresses.
    if (next < tiles) scan_load_tile<VEC>(in, n, next, tid, vn);
    uint32_t wexcl[VEC];
    const uint32_t tile_total = scan_cta_tile<VEC>(v, wexcl, s_part[it], lane, warp);
    if (warp == 0) {
      const uint32_t excl = scan_lookback(status, tile, tile_total, gen, lane);
      if (lane == 0) s_prefix[it] = excl;
    }
    __syncthreads();
    scan_store_tile<false, VEC>(out, n, tile, tid, v, wexcl, s_prefix[it]);
#pragma unroll
    for (int k = 0; k < VEC; k++) v[k] = vn[k];
    tile = next;
  }
}

// ---- kernel 2: 16-byte aligned buffers — staged tiles + DEFERRED window look-back -----------------------------------------------------
// A classic look-back resolves a tile's prefix between reading it and writing it: a round trip to L2 (or several, along the chain of
// predecessors) sits on every tile's critical path while the tile occupies the registers of a CTA of waiting threads, and reads stop
// while the CTA waits.  Every variant of that scheme — tile sizes, ticketing, look-back widths, back-off, register prefetch, staging —
// sat at 0.40-0.62 of the copy bandwidth; switching the look-back off gave 0.88 (profiles/r02_experiments.md section 4).  Here a
// tile's life is split over two iterations of its CTA, and it waits in SHARED MEMORY:
//   R(i):   the tile arrives in its stage (one cp.async.bulk per tile, completion on the stage's mbarrier; thread 0 keeps
//           SCAN_STAGES - 1 tiles ahead), the CTA sums it and publishes the aggregate A(i);
//   S(i-1): one iteration later the CTA resolves the prefix of its PREVIOUS tile k, scans it from its stage and stores it.
// Tiles are dealt round-robin, so k's predecessors back to this CTA's own tile k - G (whose inclusive prefix the CTA holds in a
// register) are the G - 1 tiles k-G+1 .. k-1 of the other CTAs: every thread loads ONE of their descriptors at the top of the
// iteration — published about an iteration ago, so the loads return valid values without polling (a late one is polled) — and the
// CTA sums them.  No inclusive-prefix descriptors, no chain, no spinning in the steady state; the round trip overlaps R(i).
// Deadlock-free: a CTA publishes A(i) after waiting only for aggregates of EARLIER iterations, and the whole grid is resident.
constexpr int SCAN_WIN = (148 * 8 + SCAN_THREADS - 1) / SCAN_THREADS;      // descriptors per thread: G - 1 <= SCAN_WIN * SCAN_THREADS (host clamps G)

template <int VEC>
__global__ void __launch_bounds__(SCAN_THREADS)
scan_window_kernel(const uint32_t* in, uint32_t* out, uint32_t n,     /* both 16-byte aligned; in may alias out */
                   unsigned long long* __restrict__ status, uint32_t tiles, uint32_t gen) {
  constexpr uint32_t TILE = SCAN_SUB * VEC, STAGE_BYTES = TILE * 4u;
  extern __shared__ __align__(128) unsigned char scan_smem[];          // SCAN_STAGES tiles, then one mbarrier per stage
  __shared__ __align__(16) uint32_t s_part[VEC * SCAN_WARPS];
  __shared__ uint2 s_red[2][SCAN_WARPS];                                // per warp {sum of the tile being reduced, sum of the window's aggregates}
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t G = gridDim.x, full_tiles = n / TILE;                  // tiles [0, full_tiles) are whole: staged by bulk copies
  const uint32_t smem_a = (uint32_t)__cvta_generic_to_shared(scan_smem);
  const uint32_t bars_a = smem_a + SCAN_STAGES * STAGE_BYTES;
  const uint32_t want = (gen << 2) | FLAG_AGG;
  auto issue = [&](int s, uint64_t tile) {
    const uint32_t bar_a = bars_a + 8u * s;
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar_a), "r"(STAGE_BYTES) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_a + s * STAGE_BYTES), "l"(in + tile * TILE), "r"(STAGE_BYTES), "r"(bar_a) : "memory");
  };
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < SCAN_STAGES; s++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(bars_a + 8u * s) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
#pragma unroll
    for (int s = 0; s < SCAN_STAGES; s++) {
      const uint64_t t = (uint64_t)blockIdx.x + (uint64_t)s * G;
      if (t < full_tiles) issue(s, t);
    }
  }
  __syncthreads();
  uint32_t prev_incl = 0;          // inclusive prefix of the CTA's last resolved tile
  uint32_t total_k = 0;            // aggregate of the tile awaiting S
  for (uint32_t i = 0;; i++) {
    const uint64_t tile64 = (uint64_t)blockIdx.x + (uint64_t)i * G;
    const bool has_r = tile64 < tiles, has_s = i > 0 && tile64 - G < tiles;
    if (!has_r && !has_s) break;
    const uint32_t k = (uint32_t)(tile64 - G);                          // the tile awaiting S (meaningful when has_s)
    // (a) the window's descriptors: in flight while R(i) runs.  Slot p = tile k - 1 - p.
    unsigned long long sd[SCAN_WIN];
    if (has_s && !CLRT_SCAN_DEBUG_NOLOOK) {
#pragma unroll
      for (int q = 0; q < SCAN_WIN; q++) {
        const uint32_t p = (uint32_t)tid + (uint32_t)q * SCAN_THREADS;
        sd[q] = (p < G - 1u && p < k) ? ld_relaxed_u64(status + (k - 1u - p)) : ((unsigned long long)want << 32);
      }
    }
    // (b) R(i): sum the arriving tile
    uint32_t mine = 0;
    if (has_r) {
      const int s = i % SCAN_STAGES;
      uint4* st = reinterpret_cast<uint4*>(scan_smem + (size_t)s * STAGE_BYTES);
      if (tile64 < full_tiles) {
        const uint32_t parity = (i / SCAN_STAGES) & 1u, bar_a = bars_a + 8u * s;
        uint32_t done = 0;
        while (!done)
          asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                       : "=r"(done) : "r"(bar_a), "r"(parity) : "memory");
#pragma unroll
        for (int q = 0; q < VEC; q++) { const uint4 x = st[q * SCAN_THREADS + tid]; mine += x.x + x.y + x.z + x.w; }
      } else {                                   // the ragged last tile: loaded directly and parked in the stage (each thread re-reads its own words)
        uint4 v[VEC];
        scan_load_tile<VEC>(in, n, (uint32_t)tile64, tid, v);
#pragma unroll
        for (int q = 0; q < VEC; q++) { st[q * SCAN_THREADS + tid] = v[q]; mine += v[q].x + v[q].y + v[q].z + v[q].w; }
      }
    }
    // (c) the window: every descriptor must carry this call's generation
    uint32_t win = 0;
    if (has_s && !CLRT_SCAN_DEBUG_NOLOOK) {
#pragma unroll
      for (int q = 0; q < SCAN_WIN; q++) {
        const uint32_t p = (uint32_t)tid + (uint32_t)q * SCAN_THREADS;
        while ((uint32_t)(sd[q] >> 32) != want) { if (CLRT_SCAN_SLEEP_NS) __nanosleep(CLRT_SCAN_SLEEP_NS); sd[q] = ld_relaxed_u64(status + (k - 1u - p)); }
        win += (uint32_t)sd[q];
      }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) { mine += __shfl_xor_sync(0xffffffffu, mine, d); win += __shfl_xor_sync(0xffffffffu, win, d); }
    if (lane == 0) s_red[i & 1][warp] = make_uint2(mine, win);
    __syncthreads();
    uint2 w = lane < SCAN_WARPS ? s_red[i & 1][lane] : make_uint2(0u, 0u);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) { w.x += __shfl_xor_sync(0xffffffffu, w.x, d); w.y += __shfl_xor_sync(0xffffffffu, w.y, d); }
    if (has_r && tid == 0) st_relaxed_u64(status + tile64, ((unsigned long long)want << 32) | w.x);
    // (d) S(i-1): scan the previous tile from its stage and store it
    if (has_s) {
      const uint32_t base = prev_incl + w.y;                             // the CTA's first tile: prev_incl = 0 and the window reaches back to tile 0
      prev_incl = base + total_k;
      const int s = (i - 1) % SCAN_STAGES;
      const uint4* st = reinterpret_cast<const uint4*>(scan_smem + (size_t)s * STAGE_BYTES);
      uint4 v[VEC];
#pragma unroll
      for (int q = 0; q < VEC; q++) v[q] = st[q * SCAN_THREADS + tid];
      uint32_t wexcl[VEC];
      scan_cta_tile<VEC>(v, wexcl, s_part, lane, warp);                  // its __syncthreads: the stage has been read by everyone
      const uint64_t refill = (uint64_t)k + (uint64_t)SCAN_STAGES * G;
      if (tid == 0 && refill < full_tiles) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");    // generic-proxy reads of the stage before the async-proxy write
        issue(s, refill);
      }
      scan_store_tile<true, VEC>(out, n, k, tid, v, wexcl, base);
    }
    total_k = w.x;
  }
}

template <int VEC>
int scan_window_launch(clrt_ctx* ctx, const uint32_t* in, uint32_t* out, size_t n, unsigned long long* status, uint32_t gen) {
  constexpr size_t TILE = (size_t)SCAN_SUB * VEC, SMEM = SCAN_STAGES * TILE * 4 + 8 * SCAN_STAGES;
  static int per_sm = 0;
  if (!per_sm) {
    CLRT_CUDA(cudaFuncSetAttribute(scan_window_kernel<VEC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM));
    CLRT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, scan_window_kernel<VEC>, SCAN_THREADS, SMEM));
    if (per_sm < 1) CLRT_FAIL("scan: the staged kernel does not fit an SM (%zu bytes of shared memory)", SMEM);
  }
  const size_t tiles = (n + TILE - 1) / TILE;
  // all CTAs resident (a CTA only waits on running CTAs); the window (grid - 1 descriptors) fits the CTA's threads
  const unsigned grid = (unsigned)min(min(tiles, (size_t)ctx->sm_count * per_sm), (size_t)SCAN_WIN * SCAN_THREADS);
  scan_window_kernel<VEC><<<grid, SCAN_THREADS, SMEM, ctx->stream>>>(in, out, (uint32_t)n, status, (uint32_t)tiles, gen);
  CLRT_CHECK_LAUNCH();
  return 0;
}

} // namespace

// Descriptor state of the context: [16 B pad | status u64 x tiles], zeroed when (re)allocated and never reset afterwards.
// A descriptor is valid only when it carries the generation of the running call.
int scan_inclusive(clrt_ctx* ctx, const uint32_t* in, uint32_t* out, size_t n) {
  if (n == 0) return 0;
  if (n > 0xFFFFFFFFull) CLRT_FAIL("scan: n = %zu exceeds 2^32-1", n);
  const size_t tiles_max = (n + SCAN_SUB - 1) / SCAN_SUB;                      // the smallest tile any kernel uses
  const size_t bytes = tiles_max * 8 + 16;
  if (bytes > ctx->scan_state_bytes || ctx->scan_gen >= (1u << 30) - 1u) {
    CLRT_CUDA(cudaStreamSynchronize(ctx->stream));
    if (bytes > ctx->scan_state_bytes) CLRT_TRY(clrt_grow(&ctx->scan_state, &ctx->scan_state_bytes, bytes + bytes / 2));
    CLRT_CUDA(cudaMemsetAsync(ctx->scan_state, 0, ctx->scan_state_bytes, ctx->stream));
    ctx->scan_gen = 0;
  }
  const uint32_t gen = ++ctx->scan_gen;
  unsigned long long* status = (unsigned long long*)((uint8_t*)ctx->scan_state + 16);
  const bool aligned = (((uintptr_t)in | (uintptr_t)out) & 15) == 0;
  if (aligned) {
    // Tile size by n (measured, profiles/r02_experiments.md section 4): 16K-element tiles (one 192 KB CTA per SM) from 2^21 elements up,
    // 8K-element tiles (two CTAs per SM) below, where a wave of big tiles would leave most SMs without work.
    const int vec = CLRT_SCAN_FORCE_VEC ? CLRT_SCAN_FORCE_VEC : (n >= (1u << 21) ? 4 : 2);
    if (vec == 4) return scan_window_launch<4>(ctx, in, out, n, status, gen);
    return scan_window_launch<2>(ctx, in, out, n, status, gen);
  }
  constexpr size_t TILE = (size_t)SCAN_SUB * SCAN_VEC_UNALIGNED;
  const size_t tiles = (n + TILE - 1) / TILE;
  static int per_sm = 0;
  if (!per_sm) {
    CLRT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, scan_lookback_kernel, SCAN_THREADS, 0));
    if (per_sm < 1) per_sm = 1;
  }
  const unsigned grid = (unsigned)min(tiles, (size_t)ctx->sm_count * per_sm);     // all CTAs resident: the look-back never waits on an unscheduled tile
  scan_lookback_kernel<<<grid, SCAN_THREADS, 0, ctx->stream>>>(in, out, (uint32_t)n, status, (uint32_t)tiles, gen);
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

// ---- scan without any allocation ----------------------------------------------------------------------------------------------------
// clrt_scan_inclusive_u32_noalloc: the same kernels as above for callers that cannot call cudaMalloc (code under CUDA graph capture,
// tight serving loops, harnesses that forbid it).  It has no counterpart in the reference.
// Differences from clrt_scan_inclusive_u32:
//   * no context: the caller passes a stream; the device is the calling thread's current device;
//   * the tile descriptors — the only scratch memory a scan needs, 8 bytes per tile — live in the file-scope array below instead of
//     the context's grown buffer: 128 KB that every GPU which loads this library holds until the process exits.  That bounds n to
//     CLRT_SCAN_NOALLOC_MAX_ELEMENTS = 2^28; larger sizes are rejected;
//   * the call makes no cudaMalloc / cudaFree / allocating library call.  Per device, the FIRST call does the set-up (address of the
//     array on that device, shared-memory attribute and occupancy of the kernels): make that call outside stream capture;
//   * the array is shared by every call on the device and keeps the previous call's contents, so each call clears the descriptors it is
//     going to use with a cudaMemsetAsync on the caller's stream (a fixed generation number: a captured graph can be replayed).
//     Calls that OVERLAP on one GPU (different streams, no ordering between them) race on the descriptors — undefined results; calls
//     ordered on one stream, or by events, are fine.
constexpr size_t SCAN_NOALLOC_MAX = (size_t)CLRT_SCAN_NOALLOC_MAX_ELEMENTS;
// descriptors needed: 16K-element tiles from 2^21 elements up (and for buffers that are only 4-byte aligned), 8K-element tiles below
constexpr size_t SCAN_NOALLOC_TILES_BIG = (SCAN_NOALLOC_MAX + (size_t)SCAN_SUB * 4 - 1) / ((size_t)SCAN_SUB * 4);
constexpr size_t SCAN_NOALLOC_TILES_SMALL = ((size_t)(1u << 21) + (size_t)SCAN_SUB * 2 - 1) / ((size_t)SCAN_SUB * 2);
constexpr size_t SCAN_NOALLOC_TILES = SCAN_NOALLOC_TILES_BIG > SCAN_NOALLOC_TILES_SMALL ? SCAN_NOALLOC_TILES_BIG : SCAN_NOALLOC_TILES_SMALL;   // 16384 = 128 KB
static_assert(SCAN_NOALLOC_TILES * (size_t)SCAN_SUB * 4 >= SCAN_NOALLOC_MAX, "descriptor array too small for the stated maximum");
static __device__ __align__(256) unsigned long long g_scan_noalloc_status[SCAN_NOALLOC_TILES];

namespace {
struct ScanNoallocDevice {              // per device, filled once with that device current
  bool ready = false;
  unsigned long long* status = nullptr; // DEVICE address of g_scan_noalloc_status on this device (the name itself is a host-side stand-in)
  int sm_count = 0, per_sm_w4 = 0, per_sm_w2 = 0, per_sm_lb = 0;
};
constexpr int SCAN_NOALLOC_MAX_DEVICES = 64;
ScanNoallocDevice g_scan_noalloc_dev[SCAN_NOALLOC_MAX_DEVICES];
std::mutex g_scan_noalloc_mutex;

template <int VEC> constexpr size_t scan_window_smem() { return SCAN_STAGES * (size_t)SCAN_SUB * VEC * 4 + 8 * SCAN_STAGES; }

int scan_noalloc_setup(int device, ScanNoallocDevice** out) {
  if (device < 0 || device >= SCAN_NOALLOC_MAX_DEVICES) CLRT_FAIL("clrt_scan_inclusive_u32_noalloc: device %d out of range", device);
  ScanNoallocDevice& d = g_scan_noalloc_dev[device];
  std::lock_guard<std::mutex> lock(g_scan_noalloc_mutex);
  if (!d.ready) {
    void* p = nullptr;
    CLRT_CUDA(cudaGetSymbolAddress(&p, g_scan_noalloc_status));          // the array's address ON THE CURRENT DEVICE
    if (!p) CLRT_FAIL("clrt_scan_inclusive_u32_noalloc: cudaGetSymbolAddress returned NULL on device %d", device);
    CLRT_CUDA(cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, device));
    CLRT_CUDA(cudaFuncSetAttribute(scan_window_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)scan_window_smem<4>()));
    CLRT_CUDA(cudaFuncSetAttribute(scan_window_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)scan_window_smem<2>()));
    CLRT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.per_sm_w4, scan_window_kernel<4>, SCAN_THREADS, scan_window_smem<4>()));
    CLRT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.per_sm_w2, scan_window_kernel<2>, SCAN_THREADS, scan_window_smem<2>()));
    CLRT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.per_sm_lb, scan_lookback_kernel, SCAN_THREADS, 0));
    if (d.per_sm_w4 < 1 || d.per_sm_w2 < 1 || d.per_sm_lb < 1 || d.sm_count < 1)
      CLRT_FAIL("clrt_scan_inclusive_u32_noalloc: the scan kernels do not fit an SM of device %d", device);
    d.status = (unsigned long long*)p;
    d.ready = true;
  }
  *out = &d;
  return 0;
}
} // namespace

extern "C" int clrt_scan_inclusive_u32_noalloc(void* stream_, const void* dptr_in, void* dptr_out, size_t n) {
  if (n == 0) return 0;
  if (!dptr_in || !dptr_out) CLRT_FAIL("clrt_scan_inclusive_u32_noalloc: NULL buffer");
  if (((uintptr_t)dptr_in & 3) || ((uintptr_t)dptr_out & 3)) CLRT_FAIL("clrt_scan_inclusive_u32_noalloc: buffers must be 4-byte aligned");
  if (n > SCAN_NOALLOC_MAX)
    CLRT_FAIL("clrt_scan_inclusive_u32_noalloc: n = %zu exceeds CLRT_SCAN_NOALLOC_MAX_ELEMENTS = %zu (the static descriptor array); "
              "use clrt_scan_inclusive_u32", n, SCAN_NOALLOC_MAX);
  int device = -1;
  CLRT_CUDA(cudaGetDevice(&device));
  ScanNoallocDevice* d = nullptr;
  CLRT_TRY(scan_noalloc_setup(device, &d));
  cudaStream_t stream = (cudaStream_t)stream_;
  const uint32_t* in = (const uint32_t*)dptr_in;
  uint32_t* out = (uint32_t*)dptr_out;
  const uint32_t gen = 1u;                                                // descriptors are cleared below: any fixed non-zero generation
  const bool aligned = (((uintptr_t)in | (uintptr_t)out) & 15) == 0;
  const int vec = aligned ? (n >= (1u << 21) ? 4 : 2) : SCAN_VEC_UNALIGNED;
  const size_t tile = (size_t)SCAN_SUB * vec, tiles = (n + tile - 1) / tile;
  if (tiles > SCAN_NOALLOC_TILES) CLRT_FAIL("clrt_scan_inclusive_u32_noalloc: %zu tiles exceed the descriptor array (%zu)", tiles, SCAN_NOALLOC_TILES);
  CLRT_CUDA(cudaMemsetAsync(d->status, 0, tiles * sizeof(unsigned long long), stream));      // what must start at zero
  if (aligned) {
    const int per_sm = vec == 4 ? d->per_sm_w4 : d->per_sm_w2;
    const unsigned grid = (unsigned)min(min(tiles, (size_t)d->sm_count * per_sm), (size_t)SCAN_WIN * SCAN_THREADS);
    if (vec == 4) scan_window_kernel<4><<<grid, SCAN_THREADS, scan_window_smem<4>(), stream>>>(in, out, (uint32_t)n, d->status, (uint32_t)tiles, gen);
    else          scan_window_kernel<2><<<grid, SCAN_THREADS, scan_window_smem<2>(), stream>>>(in, out, (uint32_t)n, d->status, (uint32_t)tiles, gen);
  } else {
    const unsigned grid = (unsigned)min(tiles, (size_t)d->sm_count * d->per_sm_lb);
    scan_lookback_kernel<<<grid, SCAN_THREADS, 0, stream>>>(in, out, (uint32_t)n, d->status, (uint32_t)tiles, gen);
  }
  CLRT_CHECK_LAUNCH();
  return 0;
}
