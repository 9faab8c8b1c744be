// ctx.cu -- context, error reporting, memory and event helpers of libpcs_b200 (include/pcs_b200.h).
#include <stdarg.h>
#include <string.h>

#include <new>

#include "common.cuh"

namespace pcs
{
static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...)
{
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what, const char* file, int line)
{
  set_error("CUDA error %d (%s) at %s:%d in %s", (int)e, cudaGetErrorString(e), file, line, what);
  return PCS_ERR_CUDA;
}

int check_launch(const char* what)
{
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess)
  {
    set_error("kernel launch failed: %s: %s", what, cudaGetErrorString(e));
    return PCS_ERR_CUDA;
  }
  return PCS_OK;
}

int ctx_bind(pcs_ctx* ctx)
{
  PCS_CHECK_ARG(ctx != nullptr, "ctx is NULL");
  PCS_CUDA(cudaSetDevice(ctx->device));
  return PCS_OK;
}

int ctx_scratch(pcs_ctx* ctx, size_t bytes, void** out)
{
  if (bytes > ctx->scratch_bytes)
  {
    if (ctx->scratch)
    {
      PCS_CUDA(cudaStreamSynchronize(ctx->stream));
      PCS_CUDA(cudaFree(ctx->scratch));
      ctx->scratch = nullptr;
      ctx->scratch_bytes = 0;
    }
    size_t want = bytes + (bytes >> 2) + 4096;
    PCS_CUDA(cudaMalloc(&ctx->scratch, want));
    ctx->scratch_bytes = want;
  }
  *out = ctx->scratch;
  return PCS_OK;
}

int ctx_staging(pcs_ctx* ctx, int which, size_t bytes, void** out)
{
  if (bytes > ctx->staging_bytes[which])
  {
    if (ctx->staging[which])
    {
      PCS_CUDA(cudaDeviceSynchronize());
      PCS_CUDA(cudaFree(ctx->staging[which]));
      ctx->staging[which] = nullptr;
      ctx->staging_bytes[which] = 0;
    }
    PCS_CUDA(cudaMalloc(&ctx->staging[which], bytes));
    ctx->staging_bytes[which] = bytes;
  }
  *out = ctx->staging[which];
  return PCS_OK;
}

int ctx_pinned(pcs_ctx* ctx, size_t bytes, void** out)
{
  if (bytes > ctx->pinned_bytes)
  {
    if (ctx->pinned)
    {
      PCS_CUDA(cudaStreamSynchronize(ctx->stream));
      PCS_CUDA(cudaFreeHost(ctx->pinned));
      ctx->pinned = nullptr;
      ctx->pinned_bytes = 0;
    }
    size_t want = bytes < 4096 ? 4096 : bytes;
    PCS_CUDA(cudaMallocHost(&ctx->pinned, want));
    ctx->pinned_bytes = want;
  }
  *out = ctx->pinned;
  return PCS_OK;
}
}  // namespace pcs

extern "C" {

int pcs_version(void) { return PCS_VERSION; }

const char* pcs_last_error(void) { return pcs::g_err; }

int pcs_ctx_create(int device, void* cuda_stream, pcs_ctx** out)
{
  PCS_CHECK_ARG(out != nullptr, "ctx out pointer is NULL");
  *out = nullptr;
  int n_dev = 0;
  cudaError_t e = cudaGetDeviceCount(&n_dev);
  if (e != cudaSuccess || n_dev == 0)
  {
    pcs::set_error("no CUDA device available (%s); libpcs_b200 has no CPU fallback",
                   e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    return PCS_ERR_CUDA;
  }
  PCS_CHECK_ARG(device >= 0 && device < n_dev, "device index out of range");
  PCS_CUDA(cudaSetDevice(device));
  pcs_ctx* ctx = new (std::nothrow) pcs_ctx();
  PCS_CHECK_ARG(ctx != nullptr, "out of host memory");
  memset(ctx, 0, sizeof(*ctx));
  ctx->device = device;
  if (cuda_stream)
  {
    ctx->stream = (cudaStream_t)cuda_stream;
    ctx->owns_stream = false;
  }
  else
  {
    PCS_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    ctx->owns_stream = true;
  }
  for (int i = 0; i < pcs::kCopyStreams; i++)
    PCS_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream[i], cudaStreamNonBlocking));
  for (int i = 0; i < 6; i++)
    PCS_CUDA(cudaEventCreateWithFlags(&ctx->pipe_event[i], cudaEventDisableTiming));
  PCS_CUDA(cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device));
  *out = ctx;
  return PCS_OK;
}

int pcs_ctx_destroy(pcs_ctx* ctx)
{
  if (!ctx)
    return PCS_OK;
  PCS_TRY(pcs::ctx_bind(ctx));
  cudaDeviceSynchronize();
  if (ctx->scratch)
    cudaFree(ctx->scratch);
  for (int i = 0; i < 8; i++)
    if (ctx->staging[i])
      cudaFree(ctx->staging[i]);
  if (ctx->pinned)
    cudaFreeHost(ctx->pinned);
  if (ctx->hsv_tables_dev)
    cudaFree(ctx->hsv_tables_dev);
  if (ctx->mesh_counters)
    cudaFree(ctx->mesh_counters);
  for (int i = 0; i < 6; i++)
    if (ctx->pipe_event[i])
      cudaEventDestroy(ctx->pipe_event[i]);
  for (int i = 0; i < pcs::kCopyStreams; i++)
    cudaStreamDestroy(ctx->copy_stream[i]);
  if (ctx->owns_stream)
    cudaStreamDestroy(ctx->stream);
  delete ctx;
  return PCS_OK;
}

int pcs_ctx_sync(pcs_ctx* ctx)
{
  PCS_TRY(pcs::ctx_bind(ctx));
  PCS_CUDA(cudaStreamSynchronize(ctx->stream));
  return PCS_OK;
}

int pcs_get_stats(pcs_ctx* ctx, pcs_stats* out)
{
  PCS_CHECK_ARG(ctx && out, "NULL argument");
  *out = ctx->stats;
  return PCS_OK;
}

int pcs_hsv_default_params(pcs_hsv_params* p)
{
  PCS_CHECK_ARG(p != nullptr, "params is NULL");
  // pcs_detection/include/pcs_detection/hsv_thresholding.h:47-52
  p->h_lo = 95;
  p->h_hi = 115;
  p->s_lo = 95;
  p->s_hi = 220;
  p->v_lo = 40;
  p->v_hi = 150;
  p->inverted = 0;
  p->morph_open = 0;
  return PCS_OK;
}

int pcs_annotate_default_params(pcs_annotate_params* p)
{
  PCS_CHECK_ARG(p != nullptr, "params is NULL");
  pcs_hsv_default_params(&p->hsv);
  // setOctree defaults, octomap_mesh_masking.h:107-109
  p->lower = 0;
  p->upper = 255;
  p->limit_negative = 0;
  p->world_frame_output = 0;
  return PCS_OK;
}

void* pcs_host_alloc(size_t bytes)
{
  void* p = nullptr;
  cudaError_t e = cudaMallocHost(&p, bytes ? bytes : 1);
  if (e != cudaSuccess)
  {
    pcs::cuda_fail(e, "cudaMallocHost", __FILE__, __LINE__);
    return nullptr;
  }
  return p;
}

void pcs_host_free(void* p)
{
  if (p)
    cudaFreeHost(p);
}

int pcs_dev_alloc(pcs_ctx* ctx, size_t bytes, void** dptr)
{
  PCS_CHECK_ARG(dptr != nullptr, "dptr is NULL");
  PCS_TRY(pcs::ctx_bind(ctx));
  PCS_CUDA(cudaMalloc(dptr, bytes ? bytes : 1));
  return PCS_OK;
}

int pcs_dev_free(pcs_ctx* ctx, void* dptr)
{
  PCS_TRY(pcs::ctx_bind(ctx));
  if (dptr)
    PCS_CUDA(cudaFree(dptr));
  return PCS_OK;
}

int pcs_memcpy_h2d(pcs_ctx* ctx, void* dst_dev, const void* src_host, size_t bytes)
{
  PCS_TRY(pcs::ctx_bind(ctx));
  PCS_CUDA(cudaMemcpyAsync(dst_dev, src_host, bytes, cudaMemcpyHostToDevice, ctx->stream));
  PCS_CUDA(cudaStreamSynchronize(ctx->stream));
  ctx->stats.bytes_h2d += bytes;
  return PCS_OK;
}

int pcs_memcpy_d2h(pcs_ctx* ctx, void* dst_host, const void* src_dev, size_t bytes)
{
  PCS_TRY(pcs::ctx_bind(ctx));
  PCS_CUDA(cudaMemcpyAsync(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  PCS_CUDA(cudaStreamSynchronize(ctx->stream));
  ctx->stats.bytes_d2h += bytes;
  return PCS_OK;
}

int pcs_memset_dev(pcs_ctx* ctx, void* dst_dev, int value, size_t bytes)
{
  PCS_TRY(pcs::ctx_bind(ctx));
  PCS_CUDA(cudaMemsetAsync(dst_dev, value, bytes, ctx->stream));
  return PCS_OK;
}

int pcs_event_create(pcs_ctx* ctx, void** event)
{
  PCS_CHECK_ARG(event != nullptr, "event is NULL");
  PCS_TRY(pcs::ctx_bind(ctx));
  cudaEvent_t ev;
  PCS_CUDA(cudaEventCreate(&ev));
  *event = (void*)ev;
  return PCS_OK;
}

int pcs_event_record(pcs_ctx* ctx, void* event)
{
  PCS_TRY(pcs::ctx_bind(ctx));
  PCS_CUDA(cudaEventRecord((cudaEvent_t)event, ctx->stream));
  return PCS_OK;
}

int pcs_stream_wait_event(pcs_ctx* ctx, void* event)
{
  PCS_TRY(pcs::ctx_bind(ctx));
  PCS_CHECK_ARG(event != nullptr, "event is NULL");
  PCS_CUDA(cudaStreamWaitEvent(ctx->stream, (cudaEvent_t)event, 0));
  return PCS_OK;
}

int pcs_event_elapsed_ms(void* start, void* stop, float* ms)
{
  PCS_CHECK_ARG(start && stop && ms, "NULL argument");
  PCS_CUDA(cudaEventSynchronize((cudaEvent_t)stop));
  PCS_CUDA(cudaEventElapsedTime(ms, (cudaEvent_t)start, (cudaEvent_t)stop));
  return PCS_OK;
}

int pcs_event_destroy(void* event)
{
  if (event)
    PCS_CUDA(cudaEventDestroy((cudaEvent_t)event));
  return PCS_OK;
}

}  // extern "C"
