// sort.cu — onesweep LSD radix sort of (key,value) uint32 pairs, by key, stable.
// Replaces BitonicSort::sort (R/src/CLRayTracerEngine/BitonicSort.cpp:118-189; kernels
// R/src/CLRayTracerEngine/SortKernels.cl:266-345): k(k+1)/2 bitonic stages, each launch a full pass over the
// 8*2^k-byte array, each followed by a host wait.  Here: ONE histogram pre-pass (all digits at once) and one
// chained-scan scatter pass per 8-bit digit, no host synchronisation.
//
// Per pass and tile (CTA): keys are loaded warp-striped as 8-byte pairs (a warp reads 256 contiguous bytes per
// instruction), ranked inside the warp with one ballot per digit bit + popc (multi-split), warp counts are combined
// across the CTA in shared memory, the tile's per-digit counts are published for the decoupled look-back (32-bit
// descriptors: 2 flag bits + 30 bit count), the tile is staged in shared memory in digit order and written out with
// consecutive threads writing consecutive addresses inside every digit run.
// Stability: rank order inside a tile == input order, tiles are chained in input order.
//
// HBM roofline: algorithmic 16 B per pair; a P-pass onesweep moves 16*P + 8 B per pair
// (the histogram pre-pass reads the 8-byte pairs once), so the ceiling of `achieved/peak` by that definition is
// 16/(16P+8): 22% at P=4, 28.6% at P=3.
//
// Also here: CLRT_SORT_BITONIC_COMPAT, the reference network replayed stage by stage (tie order identical to
// the reference).
#include "common.cuh"

namespace {

#ifndef CLRT_SORT_BIG_TILES_FROM
#define CLRT_SORT_BIG_TILES_FROM (1u << 18)     /* 1024-pair tiles (more CTAs in flight) up to this size */
#endif
#ifndef CLRT_SORT_DEBUG
#define CLRT_SORT_DEBUG 0            /* experiments only (wrong results): 1 = no look-back, 2 = no look-back and no ranking */
#endif
#ifndef CLRT_SORT_MID_TILES_TO
#define CLRT_SORT_MID_TILES_TO (1u << 19)    /* 2048-pair tiles between BIG_TILES_FROM and this size, 4096-pair tiles above (measured: profiles/r02_experiments.md section 6) */
#endif
#ifndef CLRT_SORT_HUGE_TILES_FROM
#define CLRT_SORT_HUGE_TILES_FROM (1u << 22)   /* 8192-pair tiles (512 threads) above this size */
#endif
#ifndef CLRT_SORT_MIN_CTAS
#define CLRT_SORT_MIN_CTAS 4
#endif
#ifndef CLRT_SORT_LOOK
#define CLRT_SORT_LOOK 4
#endif
constexpr int RADIX_BITS = 8;
constexpr int BINS = 1 << RADIX_BITS;
constexpr int HIST_THREADS = 256;
constexpr int MAX_PASSES = 4;

constexpr uint32_t ST_AGG = 1u << 30, ST_PREFIX = 2u << 30, ST_MASK = (1u << 30) - 1;

__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t* p) {
  uint32_t v; asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v;
}
__device__ __forceinline__ void st_relaxed_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}

// ---- pre-pass: histograms of every digit in one read of the keys ---------------------------------------
__global__ void __launch_bounds__(HIST_THREADS)
radix_histogram_kernel(const uint2* __restrict__ pairs, uint32_t n, uint32_t* __restrict__ hist /*[passes][BINS]*/, int passes) {
  __shared__ uint32_t s_hist[MAX_PASSES * BINS];
  for (int i = threadIdx.x; i < passes * BINS; i += HIST_THREADS) s_hist[i] = 0;
  __syncthreads();
  const size_t stride = (size_t)gridDim.x * HIST_THREADS;
  // two pairs (16 B) per load where possible; a buffer that is only 8-byte aligned (a slice at an odd pair offset) gives its first
  // pair to thread 0 of CTA 0 and is read as 16-byte vectors from the second one
  const uint32_t head = (uint32_t)(((uintptr_t)pairs >> 3) & 1u) & (n ? 1u : 0u);
  if (head && blockIdx.x == 0 && threadIdx.x == 0) {
    const uint32_t k = pairs[0].x;
    for (int p = 0; p < passes; p++) atomicAdd(&s_hist[p * BINS + ((k >> (p * RADIX_BITS)) & (BINS - 1))], 1u);
  }
  pairs += head; n -= head;
  const size_t n2 = n / 2;
  const uint4* p4 = reinterpret_cast<const uint4*>(pairs);
  for (size_t i = blockIdx.x * (size_t)HIST_THREADS + threadIdx.x; i < n2; i += stride) {
    const uint4 q = __ldg(p4 + i);
#pragma unroll
    for (int p = 0; p < MAX_PASSES; p++) if (p < passes) {
      atomicAdd(&s_hist[p * BINS + ((q.x >> (p * RADIX_BITS)) & (BINS - 1))], 1u);
      atomicAdd(&s_hist[p * BINS + ((q.z >> (p * RADIX_BITS)) & (BINS - 1))], 1u);
    }
  }
  if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
    const uint32_t k = pairs[n - 1].x;
    for (int p = 0; p < passes; p++) atomicAdd(&s_hist[p * BINS + ((k >> (p * RADIX_BITS)) & (BINS - 1))], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < passes * BINS; i += HIST_THREADS) { const uint32_t c = s_hist[i]; if (c) atomicAdd(&hist[i], c); }
}

// exclusive scan of each pass's 256 bins (one CTA, one warp-scan chain per pass)
__global__ void __launch_bounds__(BINS)
radix_scan_hist_kernel(uint32_t* __restrict__ hist, int passes) {
  __shared__ uint32_t s_warp[BINS / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int p = 0; p < passes; p++) {
    const uint32_t c = hist[p * BINS + tid];
    uint32_t incl = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += t; }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t off = 0;
    for (int w = 0; w < warp; w++) off += s_warp[w];
    hist[p * BINS + tid] = off + incl - c;
    __syncthreads();
  }
}

// ---- one onesweep pass ---------------------------------------------------------------------------------
// THREADS x IPT pairs per tile (CTA): 256 x 16 = 4096 (4 CTAs per SM), 512 x 16 = 8192 (2 CTAs per SM: longer digit runs in the
// write-out, half as many descriptors and look-back chains), 256 x 4 for small inputs.  Threads 0..255 own one digit each.
template <int THREADS, int IPT, bool FULL>
__global__ void __launch_bounds__(THREADS, (THREADS * IPT >= 8192 ? 2 : CLRT_SORT_MIN_CTAS))
onesweep_pass_kernel(const uint2* __restrict__ in, uint2* __restrict__ out, uint32_t n, int shift,
                     const uint32_t* __restrict__ digit_base,   // [BINS] exclusive global offsets of this pass
                     uint32_t* __restrict__ status,             // [tiles][BINS] look-back descriptors (zeroed)
                     uint32_t* __restrict__ ticket) {
  constexpr int TILE = THREADS * IPT;
  constexpr int WARPS = THREADS / 32, DIGIT_WARPS = BINS / 32;
  static_assert(THREADS >= BINS, "one thread per digit");
  __shared__ uint32_t s_warp_hist[WARPS][BINS];   // per-warp digit counts -> per-warp exclusive offsets
  __shared__ uint32_t s_tile_excl[BINS];               // exclusive scan of the tile's digit counts
  __shared__ uint32_t s_delta[BINS];                   // global position = s_delta[digit] + local position
  __shared__ uint32_t s_scan_warp[DIGIT_WARPS];
  __shared__ uint32_t s_tile;
  extern __shared__ __align__(16) uint2 s_staged[];     // TILE pairs

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_tile = atomicAdd(ticket, 1u);
#pragma unroll
  for (int i = lane; i < BINS; i += 32) s_warp_hist[warp][i] = 0;
  __syncthreads();
  const uint32_t tile = s_tile;
  const size_t tile_base = (size_t)tile * TILE;
  const uint32_t tile_items = (uint32_t)min((size_t)TILE, (size_t)n - tile_base);

  // 1. warp-striped load
  uint2 item[IPT];
  const uint32_t warp_base = warp * (32 * IPT);
#pragma unroll
  for (int j = 0; j < IPT; j++) {
    const uint32_t local = warp_base + j * 32 + lane;
    item[j] = (FULL || local < tile_items) ? in[tile_base + local] : make_uint2(0xFFFFFFFFu, 0xFFFFFFFFu);
  }
  // 2. rank inside the warp (input order = j major, lane minor).  Lanes holding the same digit are found with one ballot per
  //    digit bit: MATCH.ANY costs ~100 cycles per warp when the 32 lanes hold ~30 distinct digits (45 % of this kernel's
  //    stall samples in the first ncu capture); eight VOTE + LOP3 are an order of magnitude cheaper.  The group leader
  //    bumps the warp's digit counter with a shared-memory atomic; same-digit updates of one warp are applied in issue
  //    order (item j before item j+1), which keeps the ranking stable.
  uint32_t rank[IPT];
  const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
  for (int j = 0; j < IPT; j++) {
    const uint32_t local = warp_base + j * 32 + lane;
    const bool valid = FULL || local < tile_items;
    const uint32_t digit = (item[j].x >> shift) & (BINS - 1);
    if (CLRT_SORT_DEBUG == 2) { rank[j] = 0; if (lane == 0 && j == 0) s_warp_hist[warp][digit] = 32 * IPT; continue; }
    uint32_t m = FULL ? 0xffffffffu : __ballot_sync(0xffffffffu, valid);
#pragma unroll
    for (int b = 0; b < RADIX_BITS; b++) {
      const bool bit = (digit >> b) & 1u;
      const uint32_t v = __ballot_sync(0xffffffffu, bit);
      m &= bit ? v : ~v;
    }
    if (!FULL && !valid) m = 1u << lane;
    const int leader = __ffs(m) - 1;
    uint32_t prior = 0;
    if (valid && lane == leader) prior = atomicAdd(&s_warp_hist[warp][digit], (uint32_t)__popc(m));
    prior = __shfl_sync(0xffffffffu, prior, leader);
    rank[j] = prior + __popc(m & lt_mask);
  }
  __syncthreads();
  // 3. per digit: exclusive offsets of the warps, tile count (thread d < 256 owns digit d)
  const bool owner = tid < BINS;
  uint32_t count = 0;
  uint32_t* my_status = status + (size_t)tile * BINS + (owner ? tid : 0);
  if (owner) {
#pragma unroll
    for (int w = 0; w < WARPS; w++) { const uint32_t t = s_warp_hist[w][tid]; s_warp_hist[w][tid] = count; count += t; }
    // 4. publish the tile aggregate, then look back (one thread per digit)
    st_relaxed_u32(my_status, (tile == 0 ? ST_PREFIX : ST_AGG) | count);
  }
  // 5. exclusive scan of the tile's digit counts (for the shared-memory staging order)
  uint32_t incl = count;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += t; }
  if (owner && lane == 31) s_scan_warp[warp] = incl;
  __syncthreads();
  if (owner) {
    uint32_t off = 0;
#pragma unroll
    for (int w = 0; w < DIGIT_WARPS; w++) if (w < warp) off += s_scan_warp[w];
    s_tile_excl[tid] = off + incl - count;
    uint32_t excl = 0;
    if (tile > 0 && CLRT_SORT_DEBUG == 0) {
      // decoupled look-back, one thread per digit; LOOK descriptors are fetched together (independent L2 round trips) and
      // consumed in order — the walk was 18 % of the stall samples with one dependent load per predecessor.
      constexpr int LOOK = CLRT_SORT_LOOK;
      int t = (int)tile - 1;
      bool done = false;
      while (!done) {
        uint32_t sv[LOOK];
#pragma unroll
        for (int w = 0; w < LOOK; w++) sv[w] = (t - w >= 0) ? ld_relaxed_u32(my_status - (size_t)(tile - (uint32_t)(t - w)) * BINS) : ST_PREFIX;
#pragma unroll
        for (int w = 0; w < LOOK; w++) {
          if (!done) {
            uint32_t sx = sv[w];
            if (t - w >= 0) {
              const uint32_t* look = my_status - (size_t)(tile - (uint32_t)(t - w)) * BINS;
              while ((sx >> 30) == 0) sx = ld_relaxed_u32(look);
            }
            excl += sx & ST_MASK;
            done = (sx >> 30) == 2;
          }
        }
        t -= LOOK;
      }
      st_relaxed_u32(my_status, ST_PREFIX | (excl + count));
    }
    s_delta[tid] = digit_base[tid] + excl - s_tile_excl[tid];
  }
  __syncthreads();
  // 6. stage in digit order
#pragma unroll
  for (int j = 0; j < IPT; j++) {
    const uint32_t local = warp_base + j * 32 + lane;
    if (FULL || local < tile_items) {
      const uint32_t digit = (item[j].x >> shift) & (BINS - 1);
      s_staged[s_tile_excl[digit] + s_warp_hist[warp][digit] + rank[j]] = item[j];
    }
  }
  __syncthreads();
  // 7. coalesced write-out
#pragma unroll
  for (int j = 0; j < IPT; j++) {
    const uint32_t k = j * THREADS + tid;
    if (FULL || k < tile_items) {
      const uint2 it = s_staged[k];
      out[s_delta[(it.x >> shift) & (BINS - 1)] + k] = it;
    }
  }
}

// ---- reference bitonic network (compat mode) -----------------------------------------------------------
// One compare-exchange stage: ParallelBitonic_B2 (SortKernels.cl:269-287): i = low bits + 2*high bits,
// reverse = ((dir & i) == 0), swap = reverse ^ (key(x0) < key(x1)), comparison on keys only.
// The reference fuses up to three stages per launch (B8/B4/B2, BitonicSort.cpp:134-153); fusing does not
// change the result, so the replay runs them one by one.
__global__ void bitonic_stage_kernel(uint2* __restrict__ data, uint32_t half_n, int inc, int dir) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= half_n) return;
  const uint32_t low = t & (inc - 1);
  const uint32_t i = (t << 1) - low;
  const bool reverse = ((dir & i) == 0);
  uint2 x0 = data[i], x1 = data[i + inc];
  const bool swap = reverse ^ (x0.x < x1.x);
  if (swap) { data[i] = x1; data[i + inc] = x0; }
}

} // namespace

template <int THREADS, int IPT>
static int launch_pass(clrt_ctx* ctx, const uint2* in, uint2* out, uint32_t n, int shift, const uint32_t* base,
                       uint32_t* status, uint32_t* ticket, unsigned tiles) {
  constexpr size_t SMEM = (size_t)THREADS * IPT * sizeof(uint2);
  static bool attr_set = false;
  if (!attr_set) {
    CLRT_CUDA(cudaFuncSetAttribute(onesweep_pass_kernel<THREADS, IPT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM));
    CLRT_CUDA(cudaFuncSetAttribute(onesweep_pass_kernel<THREADS, IPT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM));
    attr_set = true;
  }
  // FULL: every tile is complete (n is a multiple of the tile: every power-of-two size the reference's callers use) -> no bounds checks
  if (n % (THREADS * IPT) == 0) onesweep_pass_kernel<THREADS, IPT, true><<<tiles, THREADS, SMEM, ctx->stream>>>(in, out, n, shift, base, status, ticket);
  else onesweep_pass_kernel<THREADS, IPT, false><<<tiles, THREADS, SMEM, ctx->stream>>>(in, out, n, shift, base, status, ticket);
  CLRT_CHECK_LAUNCH();
  return 0;
}

int sort_pairs_radix(clrt_ctx* ctx, clrt_uint2* pairs, size_t n, int key_bits) {
  if (n <= 1) return 0;
  if (n >= (1ull << 30)) CLRT_FAIL("sort: n = %zu exceeds 2^30-1", n);
  if (key_bits <= 0 || key_bits > 32) key_bits = 32;
  int passes = (key_bits + RADIX_BITS - 1) / RADIX_BITS;
  // tile size by n: 1024 pairs up to 2^19, 4096 up to CLRT_SORT_HUGE_TILES_FROM, 8192 above
  const int cls = n > (size_t)CLRT_SORT_HUGE_TILES_FROM ? 2 : n > (size_t)CLRT_SORT_MID_TILES_TO ? 1 : n > (size_t)CLRT_SORT_BIG_TILES_FROM ? 3 : 0;
  const size_t tile = cls == 2 ? 8192 : cls == 1 ? 4096 : cls == 3 ? 2048 : 1024;
  const unsigned tiles = (unsigned)((n + tile - 1) / tile);
  // scratch: [hist passes*BINS][tickets 4 (+pad)][status passes*tiles*BINS]
  const size_t hist_bytes = (size_t)MAX_PASSES * BINS * 4, ticket_bytes = 64;
  const size_t status_bytes = (size_t)passes * tiles * BINS * 4;
  void* scratch = nullptr;
  CLRT_TRY(clrt_ctx_scratch(ctx, hist_bytes + ticket_bytes + status_bytes, &scratch));
  void* alt = nullptr;
  CLRT_TRY(clrt_ctx_sort_alt(ctx, n * sizeof(uint2), &alt));
  CLRT_CUDA(cudaMemsetAsync(scratch, 0, hist_bytes + ticket_bytes + status_bytes, ctx->stream));
  uint32_t* hist = (uint32_t*)scratch;
  uint32_t* tickets = (uint32_t*)((uint8_t*)scratch + hist_bytes);
  uint32_t* status = (uint32_t*)((uint8_t*)scratch + hist_bytes + ticket_bytes);

  unsigned hblocks = (unsigned)min((size_t)ctx->sm_count * 8, (n / 2 + HIST_THREADS - 1) / HIST_THREADS);
  if (hblocks == 0) hblocks = 1;
  radix_histogram_kernel<<<hblocks, HIST_THREADS, 0, ctx->stream>>>((const uint2*)pairs, (uint32_t)n, hist, passes);
  CLRT_CHECK_LAUNCH();
  radix_scan_hist_kernel<<<1, BINS, 0, ctx->stream>>>(hist, passes);
  CLRT_CHECK_LAUNCH();

  uint2* src = (uint2*)pairs; uint2* dst = (uint2*)alt;
  for (int p = 0; p < passes; p++) {
    uint32_t* st = status + (size_t)p * tiles * BINS;
    if (cls == 2)      CLRT_TRY((launch_pass<512, 16>(ctx, src, dst, (uint32_t)n, p * RADIX_BITS, hist + p * BINS, st, tickets + p, tiles)));
    else if (cls == 3) CLRT_TRY((launch_pass<256, 8>(ctx, src, dst, (uint32_t)n, p * RADIX_BITS, hist + p * BINS, st, tickets + p, tiles)));
    else if (cls == 1) CLRT_TRY((launch_pass<256, 16>(ctx, src, dst, (uint32_t)n, p * RADIX_BITS, hist + p * BINS, st, tickets + p, tiles)));
    else               CLRT_TRY((launch_pass<256, 4>(ctx, src, dst, (uint32_t)n, p * RADIX_BITS, hist + p * BINS, st, tickets + p, tiles)));
    uint2* t = src; src = dst; dst = t;
  }
  if (src != (uint2*)pairs)
    CLRT_CUDA(cudaMemcpyAsync(pairs, src, n * sizeof(uint2), cudaMemcpyDeviceToDevice, ctx->stream));
  return 0;
}

static int sort_pairs_bitonic_compat(clrt_ctx* ctx, clrt_uint2* pairs, size_t n) {
  if (n <= 1) return 0;
  if (n & (n - 1)) CLRT_FAIL("sort (bitonic compat): n = %zu is not a power of two (BitonicSort.cpp:118-126)", n);
  if (n > (1ull << 31)) CLRT_FAIL("sort (bitonic compat): n too large");
  const uint32_t half = (uint32_t)(n / 2);
  const int threads = 256;
  const unsigned blocks = (half + threads - 1) / threads;
  for (size_t length = 1; length < n; length <<= 1)            // BitonicSort.cpp:127-153
    for (size_t inc = length; inc > 0; inc >>= 1) {
      bitonic_stage_kernel<<<blocks, threads, 0, ctx->stream>>>((uint2*)pairs, half, (int)inc, (int)(length << 1));
      CLRT_CHECK_LAUNCH();
    }
  return 0;
}

extern "C" int clrt_sort_pairs_u32(clrt_ctx* ctx, void* dptr_pairs, size_t n, int mode) {
  if (!ctx) CLRT_FAIL("clrt_sort_pairs_u32: ctx is NULL");
  if (n && !dptr_pairs) CLRT_FAIL("clrt_sort_pairs_u32: NULL buffer");
  if ((uintptr_t)dptr_pairs & 7) CLRT_FAIL("clrt_sort_pairs_u32: buffer must be 8-byte aligned");
  CLRT_CUDA(cudaSetDevice(ctx->device));
  if (mode == CLRT_SORT_RADIX) return sort_pairs_radix(ctx, (clrt_uint2*)dptr_pairs, n, 32);
  if (mode == CLRT_SORT_BITONIC_COMPAT) return sort_pairs_bitonic_compat(ctx, (clrt_uint2*)dptr_pairs, n);
  CLRT_FAIL("clrt_sort_pairs_u32: unknown mode %d", mode);
}
