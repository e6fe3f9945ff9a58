// tlg.cu — Two-Level Grid builder + traversal (placeholder until the BVH path is parity-green on the GPU).
#include "common.cuh"
int tlg_init_frame(clrt_accel*) { CLRT_FAIL("Two-Level Grid: not built in this revision"); }
int tlg_construct(clrt_accel*) { CLRT_FAIL("Two-Level Grid: not built in this revision"); }
int tlg_trace_primary(clrt_accel*, const clrt_camera&, clrt_contact*, uint32_t, uint32_t) { CLRT_FAIL("Two-Level Grid: not built in this revision"); }
int tlg_trace_rays(clrt_accel*, const clrt_ray*, clrt_contact*, uint32_t) { CLRT_FAIL("Two-Level Grid: not built in this revision"); }
void tlg_destroy(clrt_accel*) {}
extern "C" int clrt_tlg_sizes(clrt_accel*, uint32_t*, uint32_t*, uint32_t*, uint32_t*) { CLRT_FAIL("Two-Level Grid: not built in this revision"); }
extern "C" int clrt_tlg_export(clrt_accel*, clrt_grid_data*, void*, void*, void*, void*, void*) { CLRT_FAIL("Two-Level Grid: not built in this revision"); }
