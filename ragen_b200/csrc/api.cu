// Context, error reporting and version for libragen_b200.
#include "common.cuh"

#include <stdlib.h>
#include <string.h>

static thread_local char g_err[512] = "";

void ragen_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* ragen_last_error(void) { return g_err; }

extern "C" int ragen_version(void) { return 100; }  // 0.1.0

extern "C" int ragen_create(int device, ragen_ctx** out) {
  RAGEN_REQUIRE(out, "out is NULL");
  *out = nullptr;
  int count = 0;
  RAGEN_CUDA(cudaGetDeviceCount(&count));
  RAGEN_REQUIRE(device >= 0 && device < count, "device %d out of range (found %d)", device, count);
  cudaDeviceProp prop;
  RAGEN_CUDA(cudaGetDeviceProperties(&prop, device));
  RAGEN_REQUIRE(prop.major == 10, "libragen_b200 needs an sm_100a GPU (B200); device %d is sm_%d%d", device, prop.major,
                prop.minor);
  ragen_ctx* ctx = (ragen_ctx*)calloc(1, sizeof(ragen_ctx));
  if (!ctx) {
    ragen_set_error("out of host memory");
    return RAGEN_ENOMEM;
  }
  ctx->device = device;
  ctx->sm_count = prop.multiProcessorCount;
  *out = ctx;
  return RAGEN_OK;
}

extern "C" int ragen_destroy(ragen_ctx* ctx) {
  if (!ctx) return RAGEN_OK;
  if (ctx->gen_ws || ctx->scratch) {
    DeviceGuard guard(ctx->device);
    if (ctx->gen_ws) cudaFree(ctx->gen_ws);
    if (ctx->scratch) cudaFree(ctx->scratch);
  }
  free(ctx);
  return RAGEN_OK;
}

static void* grow(ragen_ctx* ctx, void** buf, size_t* have, size_t bytes) {
  if (*have >= bytes) return *buf;
  DeviceGuard guard(ctx->device);
  if (*buf) {
    cudaDeviceSynchronize();  // earlier launches may still use the old buffer
    cudaFree(*buf);
    *buf = nullptr;
    *have = 0;
  }
  size_t want = bytes < 4096 ? 4096 : bytes + bytes / 4;
  cudaError_t e = cudaMalloc(buf, want);
  if (e != cudaSuccess) {
    *buf = nullptr;
    ragen_set_error("cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e));
    return nullptr;
  }
  *have = want;
  return *buf;
}

void* ragen_scratch(ragen_ctx* ctx, size_t bytes) { return grow(ctx, &ctx->scratch, &ctx->scratch_bytes, bytes); }
void* ragen_gen_workspace(ragen_ctx* ctx, size_t bytes) { return grow(ctx, &ctx->gen_ws, &ctx->gen_ws_bytes, bytes); }

extern "C" uint64_t ragen_launch_count(const ragen_ctx* ctx) { return ctx ? ctx->launches : 0; }
extern "C" int ragen_sm_count(const ragen_ctx* ctx) { return ctx ? ctx->sm_count : 0; }
