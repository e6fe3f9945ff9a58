// common.cuh — internal declarations shared by the translation units of libclrt_cuda.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <vector>
#include "../../include/clrt.h"

// ---- error plumbing (Result + Errata convention, R/Include/Common/Errata.h:47-110) -------------------
void clrt_set_error(const char* fmt, ...);
#define CLRT_FAIL(...) do { clrt_set_error(__VA_ARGS__); return 1; } while (0)
#define CLRT_CUDA(expr) do { cudaError_t e__ = (expr); if (e__ != cudaSuccess) { \
  clrt_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e__)); return 1; } } while (0)
#define CLRT_TRY(expr) do { int r__ = (expr); if (r__) return r__; } while (0)
#define CLRT_CHECK_LAUNCH() CLRT_CUDA(cudaGetLastError())

// ---- handles -------------------------------------------------------------------------------------------
struct clrt_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  bool owns_stream = false;
  int sm_count = 0;
  cudaDeviceProp prop;
  // scratch reused by the building blocks (grow-only)
  void* scratch = nullptr; size_t scratch_bytes = 0;
  void* sort_alt = nullptr; size_t sort_alt_bytes = 0;
  void* pinned = nullptr; size_t pinned_bytes = 0;     // staging for small host<->device transfers
  // look-back state of the scan (scan.cu): never reset between calls — descriptors carry a generation
  void* scan_state = nullptr; size_t scan_state_bytes = 0; uint32_t scan_gen = 0;
};
int clrt_ctx_scratch(clrt_ctx* ctx, size_t bytes, void** out);      // grow-only device scratch
int clrt_ctx_sort_alt(clrt_ctx* ctx, size_t bytes, void** out);     // ping-pong buffer of the radix sort
int clrt_ctx_pinned(clrt_ctx* ctx, size_t bytes, void** out);

struct clrt_buf {
  clrt_ctx* ctx = nullptr;
  void* dptr = nullptr;
  size_t size = 0, capacity = 0;
};

struct MeshEntry {           // one per sub-mesh of the packed scene, built on the host at upload
  uint64_t byteOffset;       // offset of the MeshHeader inside the scene buffer
  uint32_t firstTriangle;    // global index of the sub-mesh's first triangle
  uint32_t numTriangles;
  uint32_t numVertices;
  uint32_t materialIndex;
  uint32_t modelIndex;
  uint32_t submeshIndex;
};

struct clrt_scene {
  clrt_ctx* ctx = nullptr;
  uint8_t* d_scene = nullptr; size_t bytes = 0;
  clrt_scene_header header;
  std::vector<MeshEntry> meshes;
  MeshEntry* d_meshes = nullptr;
  clrt_triangle* d_tris = nullptr;      // N x 48 B
  uint32_t* d_triInMesh = nullptr;      // N x 4 B : triangle index inside its sub-mesh (leaf data[1])
  clrt_light* d_lights = nullptr;       // numberOfLights (copy, 16 B aligned)
  clrt_material* d_materials = nullptr; // numberOfMaterials
  uint32_t n_tris = 0, n_lights = 0, n_materials = 0;
};

// 64-byte traversal node written by the refit kernel next to the reference-layout nodes:
// both child boxes + child ids in one aligned record (4 x 16 B loads).
struct __align__(16) TraceNode {
  float ax0, ay0, az0, ax1;   // child A min.xyz, max.x
  float ay1, az1, bx0, by0;   // child A max.yz, child B min.xy
  float bz0, bx1, by1, bz1;   // child B min.z, max.xyz
  uint32_t childA, childB, runStartA, runCounts;  // child id < N : leaf (= SORTED POSITION of the leaf), else N + inner node index
};

// 128-byte 4-wide traversal node (FAST policy): the 2-level collapse of a binary TraceNode tree, written by widen_kernel
// (trace_tree.cu).  Wide node i stands for binary inner node i; only the nodes reachable from the root through wide links
// are ever visited.  child < N: leaf (position in the tree-ordered triangle array); >= N: N + inner index; 0xFFFFFFFF: empty.
#ifndef CLRT_WIDE_Q16
#define CLRT_WIDE_Q16 1
#endif
#if CLRT_WIDE_Q16
// 64-byte variant: the four child boxes as 16-bit fixed point on ONE grid for the whole structure (QFrame below: bound = lo + q * step
// per axis; lo bounds rounded down, hi bounds up, checked against their decoding, so the stored box contains the exact one).  Halves
// of word w of every q array belong to children 2w (low) and 2w+1 (high).  An empty slot holds an inverted box (qlo = 65535, qhi = 0).
// Why: the traversal kernels are bound by the L1 data pipe (ncu r02a: l1tex__data_pipe_lsu_wavefronts at 82 % of peak) — every lane
// fetches its own node, so each 16-byte load instruction costs one wavefront per active lane: 7 per visit with the 128-byte node,
// 4 with this one.  The grid is per structure, not per node, so decoding needs no per-node set-up: (bound - origin) / d =
// A + q * B with A = (lo - origin) / d and B = step / d computed once per RAY (trace.cu: lane_begin).
struct __align__(16) WideNode {
  uint32_t qlx[2], qly[2];
  uint32_t qlz[2], qhx[2];
  uint32_t qhy[2], qhz[2];
  uint32_t child[4];
};
static_assert(sizeof(WideNode) == 64, "WideNode layout");
#else
struct __align__(16) WideNode {
  float lox[4], loy[4], loz[4], hix[4], hiy[4], hiz[4];
  uint32_t child[4];
  uint32_t pad[4];
};
#endif
// Quantisation grid of the 64-byte nodes, written by wide_frame_kernel behind the top table (record n + WIDE_TOP_NODES): the box of
// the structure's root, slightly enlarged, divided into 65535 steps per axis.
struct __align__(16) QFrame { float lox, loy, loz, pad0, stepx, stepy, stepz, pad1; };

// Top table of the wide tree (trace_tree.cu: wide_top_kernel): WIDE_TOP_LEVELS levels in breadth-first order, stored behind the n
// per-node records (index n + slot).  A child id with WIDE_TOP_BIT set is a slot of that table.  n < 2^30, so inner ids stay below 2^31.
#define WIDE_TOP_LEVELS 5
#define WIDE_TOP_NODES  341            /* (4^5 - 1) / 3 */
#define WIDE_TOP_BIT    0x80000000u

// Peer framebuffers of the image-space exchange (accel.cu: clrt_render_stripe_p2p): up to 8 raster-ordered RGBA8 buffers, one per
// GPU of the box (the local one included), that the shading kernel writes its pixels into directly over NVLink.
struct PeerTargets { uint32_t* p[8]; uint32_t n; };

enum AccelKind { ACCEL_BVH = 0, ACCEL_TLG = 1, ACCEL_KD = 2 };

struct clrt_accel {
  AccelKind kind = ACCEL_BVH;
  clrt_ctx* ctx = nullptr;
  const clrt_scene* scene = nullptr;
  int trace_mode = CLRT_TRACE_FAST;
  bool timing = false;
  bool unfused_shadows = false;     // frame entry points: force the three-kernel shadow pipeline (parity checks)
  bool one_kernel_frame = false;    // frame entry points: the whole stripe in one launch (frame_kernel) instead of primary + shadow/shade
  bool stats = false;               // traversal statistics mode (rays, inner visits, triangle tests)
  float stage_ms[CLRT_STAGE_COUNT] = {0};
  cudaEvent_t ev[CLRT_STAGE_COUNT + 1] = {nullptr};
  // --- BVH ---
  uint32_t n = 0;                 // leaves
  clrt_bvh_node* d_nodes = nullptr; size_t nodes_cap = 0;     // (2N-1) x 48 B, reference layout
  clrt_uint2* d_pairs = nullptr;  size_t pairs_cap = 0;       // pow2ceil(N) x 8 B (padding = UINT_MAX pairs)
  void* d_pyr = nullptr; size_t pyr_cap = 0;                  // box pyramid (boxpyr.cuh): 2 x pow2ceil(N) x 32 B
  TraceNode* d_tnodes = nullptr;  size_t tnodes_cap = 0;      // (N-1) x 64 B
  clrt_triangle* d_tris_sorted = nullptr; size_t tris_sorted_cap = 0;   // N x 48 B, sorted-leaf order, v1.w = triangle id
  uint32_t* d_refpos = nullptr; size_t refpos_cap = 0;                  // N x 4 B: triangle -> position in the reference's sorted order
  uint32_t pairs_items = 0;
  uint32_t max_run = 0;           // longest collapsed equal-key run (read back after construct; decides whether a trace tree is built)
  bool built = false;
  // --- separate trace tree (trace_tree.cu): built by construct() when the reference tree is a poor traversal structure ---
  WideNode* d_wnodes = nullptr; size_t wnodes_cap = 0;                  // (N-1) x 128 B: what CLRT_TRACE_FAST walks
  int tt_policy = 0;              // 0 auto, 1 always, 2 never
  bool tt_built = false;
  TraceNode* d_tt_nodes = nullptr; size_t tt_nodes_cap = 0;             // (N-1) x 64 B
  clrt_triangle* d_tt_tris = nullptr; size_t tt_tris_cap = 0;           // N x 48 B in trace-tree order; v1.w = triangle id, v2.w = reference sorted position
  clrt_uint2* d_tt_pairs = nullptr; size_t tt_pairs_cap = 0;            // N x 8 B (fine Morton code, triangle)
  // --- primary contacts / frame scratch ---
  clrt_contact* d_contacts = nullptr; size_t contacts_cap = 0;
  void* d_frame_scratch = nullptr; size_t frame_scratch_cap = 0;
  void* d_fb = nullptr; size_t fb_cap = 0;
  uint32_t* d_tile_counter = nullptr;
  // frame slots: frames in flight on caller-chosen streams (clrt_render_stripe_on), each with its own work counter + scratch
  static constexpr int FRAME_SLOTS = 16;
  void* d_slot_scratch[FRAME_SLOTS] = {nullptr}; size_t slot_scratch_cap[FRAME_SLOTS] = {0};
  uint32_t* d_slot_counters = nullptr;            // FRAME_SLOTS x 64 B
  cudaEvent_t slot_event[FRAME_SLOTS] = {nullptr}; // "this GPU's stripe of the slot's frame is stored everywhere" (clrt_render_frame_peers)
  int slots_seen = 0;                             // highest frame slot used so far + 1 = the caller's frames in flight (grid sizing)
  cudaStream_t launch_stream = nullptr;           // set only while a slot's launches are being enqueued
  PeerTargets launch_targets = {{nullptr}, 0};     // ditto: peer framebuffers of a p2p frame (n == 0: write the local stripe buffer)
  uint32_t* launch_counter = nullptr;
  int launch_batch = 1;                           // frames per launch of the call being enqueued (grid sizing: slots_seen / launch_batch launches run side by side)
  // --- TLG --- (tlg.cu)
  struct TlgState* tlg = nullptr;
  // --- KD-tree --- (kd.cu)
  struct KdState* kd = nullptr;
  float top_density = 2.0f, leaf_density = 2.0f;
};

// ---- internal entry points between translation units ---------------------------------------------------
int clrt_grow(void** p, size_t* cap, size_t bytes);               // cudaFree+cudaMalloc when bytes > *cap
int bvh_init_frame(clrt_accel* a);
int bvh_construct(clrt_accel* a);
int trace_tree_build(clrt_accel* a);                              // trace_tree.cu (ends with the widen pass)
void trace_tree_destroy(clrt_accel* a);
// nranks == 0: whole frame with raster-ordered output; otherwise stripe `rank` of `nranks` in stripe order
int bvh_trace_primary(clrt_accel* a, const clrt_camera& cam, clrt_contact* out, uint32_t rank, uint32_t nranks);
int bvh_trace_rays(clrt_accel* a, const clrt_ray* rays, clrt_contact* out, uint32_t n);
int tlg_init_frame(clrt_accel* a);
int tlg_construct(clrt_accel* a);
int tlg_trace_primary(clrt_accel* a, const clrt_camera& cam, clrt_contact* out, uint32_t rank, uint32_t nranks);
int tlg_trace_rays(clrt_accel* a, const clrt_ray* rays, clrt_contact* out, uint32_t n);
void tlg_destroy(clrt_accel* a);
int kd_init_frame(clrt_accel* a);
int kd_construct(clrt_accel* a);
int kd_trace_primary(clrt_accel* a, const clrt_camera& cam, clrt_contact* out, uint32_t rank, uint32_t nranks);
int kd_trace_rays(clrt_accel* a, const clrt_ray* rays, clrt_contact* out, uint32_t n);
void kd_destroy(clrt_accel* a);
int sort_pairs_radix(clrt_ctx* ctx, clrt_uint2* pairs, size_t n, int key_bits);
int scan_inclusive(clrt_ctx* ctx, const uint32_t* in, uint32_t* out, size_t n);

static inline uint32_t pow2ceil_u32(uint32_t x) { uint32_t p = 1; while (p < x) p <<= 1; return p; }

// ---- image-space tiling: 32x8-pixel tiles, tile t owned by rank t % nranks --------------
#define CLRT_TILE_W 32
#define CLRT_TILE_H 8
#define CLRT_TILE_PIXELS 256
struct TileGrid { uint32_t tilesX, tilesY, tiles; };
__host__ __device__ inline TileGrid make_tile_grid(uint32_t resX, uint32_t resY) {
  TileGrid g; g.tilesX = (resX + CLRT_TILE_W - 1) / CLRT_TILE_W; g.tilesY = (resY + CLRT_TILE_H - 1) / CLRT_TILE_H;
  g.tiles = g.tilesX * g.tilesY; return g;
}
__host__ __device__ inline uint32_t stripe_tile_count(uint32_t tiles, uint32_t rank, uint32_t nranks) {
  return tiles / nranks + ((tiles % nranks) > rank ? 1u : 0u);
}
