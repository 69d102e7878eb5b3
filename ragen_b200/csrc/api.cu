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
  if (ctx->gen_ws) {
    DeviceGuard guard(ctx->device);
    cudaFree(ctx->gen_ws);
  }
  free(ctx);
  return RAGEN_OK;
}

extern "C" uint64_t ragen_launch_count(const ragen_ctx* ctx) { return ctx ? ctx->launches : 0; }
extern "C" int ragen_sm_count(const ragen_ctx* ctx) { return ctx ? ctx->sm_count : 0; }
