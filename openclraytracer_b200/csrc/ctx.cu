// ctx.cu — context + buffers: the thin C-ABI CUDA layer that replaces the reference's OpenCL wrapper
// (R/Include/OpenCLUtils/CLExecutionContext.h, CLBuffer.h; R/src/OpenCLUtils/CLExecutionContext.cpp:264-293,386-403,
//  R/src/OpenCLUtils/CLBuffer.cpp:42-122).  One in-order stream per context, like the reference's single queue.
#include "common.cuh"
#include <stdarg.h>

static thread_local char g_err[1024] = "";
void clrt_set_error(const char* fmt, ...) {
  va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof g_err, fmt, ap); va_end(ap);
}

extern "C" const char* clrt_last_error(void) { return g_err; }
extern "C" int clrt_abi_version(void) { return CLRT_ABI_VERSION; }

int clrt_grow(void** p, size_t* cap, size_t bytes) {
  if (bytes <= *cap && *p) return 0;
  if (*p) { CLRT_CUDA(cudaFree(*p)); *p = nullptr; *cap = 0; }
  if (bytes == 0) bytes = 16;
  CLRT_CUDA(cudaMalloc(p, bytes));
  *cap = bytes;
  return 0;
}

extern "C" int clrt_ctx_create(int device, void* stream, clrt_ctx** out) {
  if (!out) CLRT_FAIL("clrt_ctx_create: out is NULL");
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0)
    CLRT_FAIL("clrt_ctx_create: no CUDA device (%s) - this library has no CPU fallback", cudaGetErrorString(e));
  if (device < 0 || device >= count) CLRT_FAIL("clrt_ctx_create: device %d out of range (0..%d)", device, count - 1);
  CLRT_CUDA(cudaSetDevice(device));
  clrt_ctx* c = new clrt_ctx();
  c->device = device;
  CLRT_CUDA(cudaGetDeviceProperties(&c->prop, device));
  if (c->prop.major < 10) {
    int major = c->prop.major, minor = c->prop.minor; delete c;
    CLRT_FAIL("clrt_ctx_create: device is sm_%d%d; this library is built for sm_100a (B200) only", major, minor);
  }
  c->sm_count = c->prop.multiProcessorCount;
  if (stream) { c->stream = (cudaStream_t)stream; c->owns_stream = false; }
  else { CLRT_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)); c->owns_stream = true; }
  *out = c;
  return 0;
}

extern "C" int clrt_ctx_destroy(clrt_ctx* c) {
  if (!c) return 0;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  if (c->scratch) cudaFree(c->scratch);
  if (c->sort_alt) cudaFree(c->sort_alt);
  if (c->scan_state) cudaFree(c->scan_state);
  if (c->pinned) cudaFreeHost(c->pinned);
  if (c->owns_stream) cudaStreamDestroy(c->stream);
  delete c;
  return 0;
}

extern "C" int clrt_ctx_sync(clrt_ctx* c) {
  if (!c) CLRT_FAIL("clrt_ctx_sync: ctx is NULL");
  CLRT_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

extern "C" void* clrt_ctx_stream(clrt_ctx* c) { return c ? (void*)c->stream : nullptr; }

extern "C" int clrt_ctx_device_info(clrt_ctx* c, int what, uint64_t* out) {
  if (!c || !out) CLRT_FAIL("clrt_ctx_device_info: NULL argument");
  switch (what) {
    case CLRT_DEV_LOCAL_MEM_SIZE:      *out = c->prop.sharedMemPerBlockOptin; break;
    case CLRT_DEV_MAX_WORK_GROUP_SIZE: *out = c->prop.maxThreadsPerBlock; break;
    case CLRT_DEV_MAX_COMPUTE_UNITS:   *out = c->prop.multiProcessorCount; break;
    case CLRT_DEV_GLOBAL_MEM_SIZE:     *out = c->prop.totalGlobalMem; break;
    default: CLRT_FAIL("clrt_ctx_device_info: unknown query %d", what);
  }
  return 0;
}

int clrt_ctx_scratch(clrt_ctx* c, size_t bytes, void** out) {
  if (bytes > c->scratch_bytes) {
    CLRT_CUDA(cudaStreamSynchronize(c->stream));   // the old block may still be in use by queued work
    CLRT_TRY(clrt_grow(&c->scratch, &c->scratch_bytes, bytes + bytes / 4));
  }
  *out = c->scratch; return 0;
}
int clrt_ctx_sort_alt(clrt_ctx* c, size_t bytes, void** out) {
  if (bytes > c->sort_alt_bytes) {
    CLRT_CUDA(cudaStreamSynchronize(c->stream));
    CLRT_TRY(clrt_grow(&c->sort_alt, &c->sort_alt_bytes, bytes));
  }
  *out = c->sort_alt; return 0;
}
int clrt_ctx_pinned(clrt_ctx* c, size_t bytes, void** out) {
  if (bytes > c->pinned_bytes) {
    CLRT_CUDA(cudaStreamSynchronize(c->stream));
    if (c->pinned) { CLRT_CUDA(cudaFreeHost(c->pinned)); c->pinned = nullptr; c->pinned_bytes = 0; }
    CLRT_CUDA(cudaMallocHost(&c->pinned, bytes));
    c->pinned_bytes = bytes;
  }
  *out = c->pinned; return 0;
}

// ---- buffers -------------------------------------------------------------------------------------------
extern "C" int clrt_buf_alloc(clrt_ctx* c, size_t bytes, const void* h_src, clrt_buf** out) {
  if (!c || !out) CLRT_FAIL("clrt_buf_alloc: NULL argument");
  CLRT_CUDA(cudaSetDevice(c->device));
  clrt_buf* b = new clrt_buf(); b->ctx = c;
  if (clrt_grow(&b->dptr, &b->capacity, bytes)) { delete b; return 1; }
  b->size = bytes;
  if (h_src && bytes) {
    CLRT_CUDA(cudaMemcpyAsync(b->dptr, h_src, bytes, cudaMemcpyHostToDevice, c->stream));
    CLRT_CUDA(cudaStreamSynchronize(c->stream));
  }
  *out = b; return 0;
}
extern "C" int clrt_buf_free(clrt_buf* b) {
  if (!b) return 0;
  cudaSetDevice(b->ctx->device);
  cudaStreamSynchronize(b->ctx->stream);
  if (b->dptr) cudaFree(b->dptr);
  delete b; return 0;
}
// CLBuffer::resize (CLBuffer.cpp:70-94): no-op when the allocation is large enough, else reallocate (contents lost).
extern "C" int clrt_buf_resize(clrt_buf* b, size_t bytes) {
  if (!b) CLRT_FAIL("clrt_buf_resize: NULL buffer");
  CLRT_CUDA(cudaSetDevice(b->ctx->device));
  if (bytes > b->capacity) {
    CLRT_CUDA(cudaStreamSynchronize(b->ctx->stream));
    CLRT_TRY(clrt_grow(&b->dptr, &b->capacity, bytes));
  }
  b->size = bytes; return 0;
}
extern "C" size_t clrt_buf_size(const clrt_buf* b) { return b ? b->size : 0; }
extern "C" size_t clrt_buf_capacity(const clrt_buf* b) { return b ? b->capacity : 0; }
extern "C" void* clrt_buf_dptr(const clrt_buf* b) { return b ? b->dptr : nullptr; }

__global__ void fill_pattern_kernel(uint32_t* __restrict__ dst, size_t words, uint4 pat, uint32_t pat_words) {
  const uint32_t p[4] = {pat.x, pat.y, pat.z, pat.w};
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < words; i += (size_t)gridDim.x * blockDim.x)
    dst[i] = p[i % pat_words];
}

extern "C" int clrt_buf_fill(clrt_buf* b, const void* h_pattern, size_t pattern_size, size_t bytes) {
  if (!b || !h_pattern) CLRT_FAIL("clrt_buf_fill: NULL argument");
  if (bytes > b->capacity) CLRT_FAIL("clrt_buf_fill: %zu bytes exceed the allocation (%zu)", bytes, b->capacity);
  if (pattern_size != 1 && pattern_size != 2 && pattern_size != 4 && pattern_size != 8 && pattern_size != 16)
    CLRT_FAIL("clrt_buf_fill: pattern size %zu not in {1,2,4,8,16}", pattern_size);
  if (bytes % pattern_size) CLRT_FAIL("clrt_buf_fill: size %zu is not a multiple of the pattern size %zu", bytes, pattern_size);
  if (bytes == 0) return 0;
  CLRT_CUDA(cudaSetDevice(b->ctx->device));
  uint8_t rep[16];
  for (int i = 0; i < 16; i++) rep[i] = ((const uint8_t*)h_pattern)[i % pattern_size];
  if (pattern_size < 4 && (bytes % 4)) {           // byte/short patterns over odd sizes: memset paths
    if (pattern_size == 1) { CLRT_CUDA(cudaMemsetAsync(b->dptr, rep[0], bytes, b->ctx->stream)); return 0; }
    std::vector<uint8_t> tmp(bytes);
    for (size_t i = 0; i < bytes; i++) tmp[i] = rep[i % pattern_size];
    CLRT_CUDA(cudaMemcpyAsync(b->dptr, tmp.data(), bytes, cudaMemcpyHostToDevice, b->ctx->stream));
    CLRT_CUDA(cudaStreamSynchronize(b->ctx->stream));
    return 0;
  }
  uint4 pat; memcpy(&pat, rep, 16);
  const uint32_t pat_words = pattern_size <= 4 ? 1u : (uint32_t)(pattern_size / 4);
  const size_t words = bytes / 4;
  const int threads = 256;
  size_t blocks = (words + threads - 1) / threads;
  if (blocks > (size_t)b->ctx->sm_count * 16) blocks = (size_t)b->ctx->sm_count * 16;
  fill_pattern_kernel<<<(unsigned)blocks, threads, 0, b->ctx->stream>>>((uint32_t*)b->dptr, words, pat, pat_words);
  CLRT_CHECK_LAUNCH();
  return 0;
}

extern "C" int clrt_buf_upload(clrt_buf* b, const void* h_src, size_t offset, size_t bytes) {
  if (!b || (!h_src && bytes)) CLRT_FAIL("clrt_buf_upload: NULL argument");
  if (offset + bytes > b->capacity) CLRT_FAIL("clrt_buf_upload: range [%zu,%zu) exceeds the allocation (%zu)", offset, offset + bytes, b->capacity);
  if (!bytes) return 0;
  CLRT_CUDA(cudaSetDevice(b->ctx->device));
  CLRT_CUDA(cudaMemcpyAsync((uint8_t*)b->dptr + offset, h_src, bytes, cudaMemcpyHostToDevice, b->ctx->stream));
  CLRT_CUDA(cudaStreamSynchronize(b->ctx->stream));
  return 0;
}
extern "C" int clrt_buf_download(const clrt_buf* b, void* h_dst, size_t offset, size_t bytes) {
  if (!b || (!h_dst && bytes)) CLRT_FAIL("clrt_buf_download: NULL argument");
  if (offset + bytes > b->capacity) CLRT_FAIL("clrt_buf_download: range [%zu,%zu) exceeds the allocation (%zu)", offset, offset + bytes, b->capacity);
  if (!bytes) return 0;
  CLRT_CUDA(cudaSetDevice(b->ctx->device));
  CLRT_CUDA(cudaMemcpyAsync(h_dst, (const uint8_t*)b->dptr + offset, bytes, cudaMemcpyDeviceToHost, b->ctx->stream));
  CLRT_CUDA(cudaStreamSynchronize(b->ctx->stream));
  return 0;
}

extern "C" int clrt_memcpy_htod(clrt_ctx* c, void* dptr_dst, const void* h_src, size_t bytes) {
  if (!c || (bytes && (!dptr_dst || !h_src))) CLRT_FAIL("clrt_memcpy_htod: NULL argument");
  if (!bytes) return 0;
  CLRT_CUDA(cudaSetDevice(c->device));
  CLRT_CUDA(cudaMemcpyAsync(dptr_dst, h_src, bytes, cudaMemcpyHostToDevice, c->stream));
  CLRT_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}
extern "C" int clrt_memcpy_dtoh(clrt_ctx* c, void* h_dst, const void* dptr_src, size_t bytes) {
  if (!c || (bytes && (!h_dst || !dptr_src))) CLRT_FAIL("clrt_memcpy_dtoh: NULL argument");
  if (!bytes) return 0;
  CLRT_CUDA(cudaSetDevice(c->device));
  CLRT_CUDA(cudaMemcpyAsync(h_dst, dptr_src, bytes, cudaMemcpyDeviceToHost, c->stream));
  CLRT_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}
