/* pm_ctx.cu — context, error reporting, scratch memory and host/device staging. */
#include <cstdarg>
#include "pm_internal.h"

namespace pm {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

bool is_device_ptr(const void* p) {
  if (!p) return false;
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

int stage_in(pm_ctx* ctx, int slot, const void* p, size_t bytes, const void** dev) {
  if (!p || bytes == 0 || is_device_ptr(p)) { *dev = p; return PM_OK; }
  void* d = nullptr;
  PM_CHECK(ctx->get_scratch(slot, bytes, &d));
  PM_CUDA(cudaMemcpyAsync(d, p, bytes, cudaMemcpyHostToDevice, ctx->stream));
  *dev = d;
  return PM_OK;
}

int stage_out(pm_ctx* ctx, int slot, void* p, size_t bytes, OutStage* st) {
  st->host = nullptr; st->dev = p; st->bytes = bytes;
  if (!p || bytes == 0 || is_device_ptr(p)) return PM_OK;
  void* d = nullptr;
  PM_CHECK(ctx->get_scratch(slot, bytes, &d));
  st->host = p; st->dev = d;
  return PM_OK;
}

int stage_out_finish(pm_ctx* ctx, const OutStage& st) {
  if (!st.host) return PM_OK;
  PM_CUDA(cudaMemcpyAsync(st.host, st.dev, st.bytes, cudaMemcpyDeviceToHost, ctx->stream));
  return PM_OK;
}

}  // namespace pm

int pm_ctx::get_scratch(int slot, size_t bytes, void** out) {
  pm::Scratch& s = scratch[slot];
  if (bytes > s.cap) {
    if (s.ptr) {
      PM_CUDA(cudaStreamSynchronize(stream));
      PM_CUDA(cudaFree(s.ptr));
      s.ptr = nullptr; s.cap = 0;
    }
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&s.ptr, want);
    if (e != cudaSuccess) {
      cudaGetLastError();
      want = bytes;
      e = cudaMalloc(&s.ptr, want);
    }
    if (e != cudaSuccess) {
      cudaGetLastError();
      pm::set_error("out of device memory allocating %zu bytes (scratch slot %d)", bytes, slot);
      return PM_ERR_NOMEM;
    }
    s.cap = want;
  }
  *out = s.ptr;
  return PM_OK;
}

extern "C" {

const char* pm_last_error(void) { return pm::g_err; }
int pm_version(void) { return 100; }

int pm_ctx_create(int device, pm_ctx** out) {
  if (!out) { pm::set_error("pm_ctx_create: out is NULL"); return PM_ERR_INVALID; }
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    pm::set_error("pm_ctx_create: no usable CUDA device (%s); libpmhot has no CPU fallback",
                  e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    return PM_ERR_CUDA;
  }
  if (device < 0 || device >= ndev) { pm::set_error("pm_ctx_create: device %d out of range [0,%d)", device, ndev); return PM_ERR_INVALID; }
  PM_CUDA(cudaSetDevice(device));
  pm_ctx* c = new pm_ctx();
  c->device = device;
  cudaDeviceProp prop;
  PM_CUDA(cudaGetDeviceProperties(&prop, device));
  c->sm_count = prop.multiProcessorCount;
  PM_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  {
    /* photon-map buffers come from the stream-ordered pool (pm_internal.h: map_alloc); keep up to a third of the
     * device's memory cached (a 200 M-photon map owns 11 GB: with less, every rebuild pays the page mapping again) */
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
      uint64_t keep = (uint64_t)prop.totalGlobalMem / 3, cur = 0;
      if (keep < (8ull << 30)) keep = 8ull << 30;
      cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &cur);
      if (cur < keep) cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
    cudaGetLastError();
  }
  PM_CUDA(cudaEventCreate(&c->ev0));
  PM_CUDA(cudaEventCreate(&c->ev1));
  PM_CUDA(cudaEventCreateWithFlags(&c->ev_order, cudaEventDisableTiming));
  PM_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&c->host_counts), 4 * sizeof(int), cudaHostAllocDefault));
  *out = c;
  return PM_OK;
}

void pm_ctx_destroy(pm_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  for (auto& s : ctx->scratch)
    if (s.ptr) cudaFree(s.ptr);
  if (ctx->copy_stream) {
    cudaStreamSynchronize(ctx->copy_stream);
    cudaStreamDestroy(ctx->copy_stream);
    for (int i = 0; i < 2; i++) { cudaEventDestroy(ctx->ev_kernel[i]); cudaEventDestroy(ctx->ev_copy[i]); }
  }
  if (ctx->in_stream) {
    cudaStreamSynchronize(ctx->in_stream);
    cudaStreamDestroy(ctx->in_stream);
    cudaEventDestroy(ctx->ev_in);
  }
  for (cudaEvent_t e : ctx->pipe_events) cudaEventDestroy(e);
  cudaEventDestroy(ctx->ev0);
  cudaEventDestroy(ctx->ev1);
  cudaEventDestroy(ctx->ev_order);
  if (ctx->host_counts) cudaFreeHost(ctx->host_counts);
  cudaStreamDestroy(ctx->stream);
  delete ctx;
}

void* pm_ctx_stream(pm_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

int pm_ctx_sync(pm_ctx* ctx) {
  if (!ctx) return PM_ERR_INVALID;
  PM_CUDA(cudaStreamSynchronize(ctx->stream));
  return PM_OK;
}

int pm_ctx_wait_stream(pm_ctx* ctx, void* other_stream) {
  if (!ctx) return PM_ERR_INVALID;
  PM_CUDA(cudaSetDevice(ctx->device));
  PM_CUDA(cudaEventRecord(ctx->ev_order, (cudaStream_t)other_stream));
  PM_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev_order, 0));
  return PM_OK;
}

int pm_dev_alloc(pm_ctx* ctx, int64_t bytes, void** out) {
  if (!ctx || !out || bytes < 0) { pm::set_error("pm_dev_alloc: bad arguments"); return PM_ERR_INVALID; }
  *out = nullptr;
  if (bytes == 0) return PM_OK;
  PM_CUDA(cudaSetDevice(ctx->device));
  if (cudaMalloc(out, (size_t)bytes) != cudaSuccess) {
    cudaGetLastError();
    *out = nullptr;
    pm::set_error("pm_dev_alloc: out of device memory allocating %lld bytes", (long long)bytes);
    return PM_ERR_NOMEM;
  }
  PM_CUDA(cudaMemsetAsync(*out, 0, (size_t)bytes, ctx->stream));
  PM_CUDA(cudaStreamSynchronize(ctx->stream));
  return PM_OK;
}

void pm_dev_free(pm_ctx* ctx, void* ptr) {
  if (!ctx || !ptr) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  cudaFree(ptr);
}

int pm_dev_copy(pm_ctx* ctx, void* dst, const void* src, int64_t bytes) {
  if (!ctx || bytes < 0 || (bytes > 0 && (!dst || !src))) { pm::set_error("pm_dev_copy: bad arguments"); return PM_ERR_INVALID; }
  if (bytes == 0) return PM_OK;
  PM_CUDA(cudaSetDevice(ctx->device));
  PM_CUDA(cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDefault, ctx->stream));
  PM_CUDA(cudaStreamSynchronize(ctx->stream));
  return PM_OK;
}

int pm_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

int64_t pm_ctx_launch_count(pm_ctx* ctx) { return ctx ? ctx->launches : 0; }
void pm_ctx_reset_launch_count(pm_ctx* ctx) { if (ctx) ctx->launches = 0; }
float pm_ctx_last_kernel_ms(pm_ctx* ctx) { return ctx ? ctx->last_kernel_ms : 0.f; }
int64_t pm_ctx_last_fallback_count(pm_ctx* ctx) { return ctx ? ctx->last_fallbacks : 0; }
int64_t pm_ctx_last_handover_count(pm_ctx* ctx) { return ctx ? ctx->last_handovers : 0; }

}  // extern "C"
