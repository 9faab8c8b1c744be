// passthrough_noalloc.cu -- pcs_color_passthrough_noalloc_dev: the label filter for callers that must not allocate.
//
// Who needs this.  Two kinds of host cannot let this library call cudaMalloc on their behalf:
//   * per-frame callbacks with a latency budget (the annotator node's subscriber callback runs for every camera frame):
//     cudaMalloc / cudaFree serialise against all work on the device, so an allocation in the middle of a scan stalls
//     the detector and every other stream for its duration;
//   * cells where the GPU is shared with a process that reserved almost all device memory when it started (a
//     TensorFlow detector node does that by default): a later allocation here may simply fail, at a moment nobody can
//     do anything about it.
// pcs_color_passthrough (annotate.cu) grows context-owned staging and scratch buffers on demand.  This variant takes
// device pointers for the cloud, the filtered cloud and the count, and keeps the only extra memory the compaction
// needs -- one 32-bit count and one 64-bit offset per block of 2048 points -- in the fixed array below, which is part
// of the library image: every GPU that loads libpcs_b200.so carries these 48.25 KiB until the process exits.
//
// Rules that follow from the fixed array (also in include/pcs_b200.h):
//   * at most PCS_NOALLOC_MAX_POINTS points per call; larger inputs are refused with PCS_ERR_CAPACITY;
//   * all calls on a device share the array, and it still holds the previous call's block counts and offsets when the
//     next call starts.  Nothing in it has to start at zero -- every word a call reads was written earlier in that
//     same call -- so there is nothing to clear, but two calls whose kernels overlap on the GPU (different streams or
//     different contexts on one device) overwrite each other's counts: keep such calls ordered;
//   * the input and output point ranges must not overlap (no in-place filtering): the call refuses them;
//   * the first call on a device looks up the array's device address (cudaGetSymbolAddress, once per device, with that
//     device current).  Under CUDA's lazy module loading that first call is also where the driver loads this library's
//     device code onto the GPU, which is the one moment the driver itself may reserve memory on the call's behalf.
//     Make that first call early, before the latency-critical phase and before starting any stream capture.
// Host code must never hand the array's name to a kernel: on the host the name is only a handle for the runtime, not
// a device address.  Everything below goes through the looked-up pointer.
#include <mutex>

#include "common.cuh"
#include "compact.cuh"

namespace
{
using namespace pcs;

constexpr unsigned long long kMaxPoints = PCS_NOALLOC_MAX_POINTS;
constexpr size_t kMaxBlocks = (size_t)((kMaxPoints + kCompactItems - 1) / kCompactItems);
// same layout as compact_scratch_bytes(): u32 counts padded to 16 bytes, then (blocks + 1) u64 offsets
constexpr size_t kScratchBytes =
    ((((kMaxBlocks * sizeof(unsigned int) + 15) & ~(size_t)15) + (kMaxBlocks + 1) * sizeof(unsigned long long)) + 255) &
    ~(size_t)255;

static __device__ __align__(256) unsigned char g_passthrough_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_addr_mutex;
void* g_scratch_addr[kMaxDevices] = {};

// device address of the array on the CURRENT device; looked up once per device
int scratch_address(int device, void** out)
{
  std::lock_guard<std::mutex> lock(g_addr_mutex);
  if (device < 0 || device >= kMaxDevices)
  {
    set_error("device index %d outside the %d devices this entry point tracks", device, kMaxDevices);
    return PCS_ERR_BAD_ARG;
  }
  if (!g_scratch_addr[device])
  {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_passthrough_scratch);
    if (e != cudaSuccess)
      return cuda_fail(e, "cudaGetSymbolAddress(g_passthrough_scratch)", __FILE__, __LINE__);
    g_scratch_addr[device] = p;
  }
  *out = g_scratch_addr[device];
  return PCS_OK;
}

struct FilterPred
{
  const Point8* pts;
  int lower, upper, neg;
  __device__ bool operator()(unsigned long long i) const
  {
    const unsigned int c = reinterpret_cast<const unsigned int*>(pts + i)[4];  // b | g << 8 | r << 16 | a << 24
    return color_pass((c >> 16) & 0xff, (c >> 8) & 0xff, c & 0xff, lower, upper, neg);
  }
};

struct FilterEmit
{
  const Point8* in;
  Point8* out;
  __device__ void operator()(unsigned long long i, unsigned long long dst) const
  {
    const uint4* s = reinterpret_cast<const uint4*>(in + i);
    uint4* d = reinterpret_cast<uint4*>(out + dst);
    d[0] = s[0];
    d[1] = s[1];
  }
};
}  // namespace

extern "C" int pcs_color_passthrough_noalloc_dev(pcs_ctx* ctx, const pcs_point_xyzrgb* in_dev, uint64_t n, int lower,
                                                 int upper, int limit_negative, pcs_point_xyzrgb* out_dev,
                                                 uint64_t* n_out_dev)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(n_out_dev != nullptr, "n_out_dev is NULL");
  PCS_CHECK_ARG(lower <= upper, "colorPassthrough: lower limit greater than upper limit");
  if (n > kMaxPoints)
  {
    set_error("pcs_color_passthrough_noalloc_dev: %llu points exceed PCS_NOALLOC_MAX_POINTS (%llu); split the cloud or "
              "use pcs_color_passthrough",
              (unsigned long long)n, (unsigned long long)kMaxPoints);
    return PCS_ERR_CAPACITY;
  }
  if (n == 0)
  {
    PCS_CUDA(cudaMemsetAsync(n_out_dev, 0, sizeof(uint64_t), ctx->stream));
    return PCS_OK;
  }
  PCS_CHECK_ARG(in_dev && out_dev, "NULL point buffer");
  PCS_CHECK_ARG(((uintptr_t)in_dev & 15u) == 0 && ((uintptr_t)out_dev & 15u) == 0, "point buffers must be 16-byte aligned");
  PCS_CHECK_ARG(((uintptr_t)n_out_dev & 7u) == 0, "n_out_dev must be 8-byte aligned");
  {
    // The scatter pass writes out[dst <= i] while other blocks still read `in` and re-evaluates the filter on it, so an
    // in-place call would corrupt the result silently: refuse any overlap of the two point ranges.
    const uintptr_t a0 = (uintptr_t)in_dev, a1 = a0 + (uintptr_t)n * sizeof(pcs_point_xyzrgb);
    const uintptr_t b0 = (uintptr_t)out_dev, b1 = b0 + (uintptr_t)n * sizeof(pcs_point_xyzrgb);
    PCS_CHECK_ARG(a1 <= b0 || b1 <= a0, "in_dev and out_dev overlap: the filter cannot run in place");
  }
  if (compact_scratch_bytes(n) > kScratchBytes)  // cannot happen for n <= kMaxPoints; guards a future layout change
  {
    set_error("internal: fixed scratch array too small for %llu points", (unsigned long long)n);
    return PCS_ERR_CAPACITY;
  }
  void* scratch = nullptr;
  PCS_TRY(scratch_address(ctx->device, &scratch));
  FilterPred pred{ (const Point8*)in_dev, lower, upper, limit_negative ? 1 : 0 };
  FilterEmit emit{ (const Point8*)in_dev, (Point8*)out_dev };
  const unsigned long long* d_total = nullptr;
  PCS_TRY(compact_run(ctx, pred, emit, n, scratch, &d_total));
  PCS_CUDA(cudaMemcpyAsync(n_out_dev, d_total, sizeof(uint64_t), cudaMemcpyDeviceToDevice, ctx->stream));
  return PCS_OK;
}
