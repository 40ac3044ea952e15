// umt_model_encode_static: the backbone for at most one frame's worth of crops, with every byte of
// scratch taken from a fixed array that is part of this library's device image.
//
// Why it exists.  umt_model_encode needs an activation workspace that somebody has to allocate
// and bind first.  Some hosts of this library are not in a position to do that at the moment
// they want features: a per-frame tracking loop that runs after the application has frozen its
// device-memory budget, a plug-in living inside another runtime that owns the allocator and
// forbids cudaMalloc on its streams, or code that is called while device allocation is simply
// not permitted (cudaMalloc synchronises the whole device, which a latency-bound loop may not
// tolerate).  Those callers can create the model once at start-up (weights are uploaded there)
// and afterwards call this entry point, which performs no device allocation at all: no
// cudaMalloc/cudaFree, no pool, no torch tensor.
//
// The price: the scratch is a 16 MiB `static __device__` array.  Every GPU on which the CUDA
// runtime loads this part of the library keeps those 16 MiB resident until the process exits
// (with eager module loading that is every GPU the process touches; with lazy loading, the
// default since CUDA 12, it is every GPU on which this entry point has been called once -- the
// driver materialises the array during that first call, which is one more reason to make it at
// start-up), and the batch is capped at kStaticMaxCrops = 4 crops (two hands x two views, i.e.
// one frame of the tracker).
//
// Rules for callers
//   * The first call on each device looks up the device address of the array
//     (cudaGetSymbolAddress; the array's name in host code is only a handle, never a pointer a
//     kernel may dereference).  Make that first call during start-up, with the intended device
//     current, and before beginning any stream capture on that device.
//   * All calls on a device share the one array.  A call first clears it on `stream` (the
//     kernels rely on zero halo rows and the array still holds whatever the previous call left),
//     then runs the layers on `stream`.  Two calls whose GPU work can overlap -- different
//     streams, or different host threads -- would trample each other's activations: serialise
//     them on one stream or with events.
//   * n_img > 4 is rejected with UMT_ERR_INVALID; use umt_model_encode with a bound workspace.
#include <mutex>

#include "model.cuh"

namespace {

constexpr int kStaticMaxCrops = 4;
constexpr size_t kStaticScratchBytes = 16u << 20;   // >= encoder_workspace_bytes(4) = 16,174,080

static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

constexpr int kMaxDevices = 64;
uint8_t* g_scratch_ptr[kMaxDevices] = {nullptr};     // device address of the array, per device
std::mutex g_setup_mutex;

// once per device, with that device current
int static_scratch_for_current_device(uint8_t** out) {
  int dev = -1;
  UMT_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return umt::fail(UMT_ERR_UNSUPPORTED, "device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_setup_mutex);
  if (!g_scratch_ptr[dev]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || !p)
      return umt::fail(UMT_ERR_CUDA, "cudaGetSymbolAddress(static scratch) on device %d: %s", dev,
                       cudaGetErrorString(e));
    if (umt::encoder_workspace_bytes(kStaticMaxCrops) > kStaticScratchBytes)
      return umt::fail(UMT_ERR_UNSUPPORTED, "static scratch (%zu bytes) is smaller than the %d-crop workspace (%zu)",
                       kStaticScratchBytes, kStaticMaxCrops, umt::encoder_workspace_bytes(kStaticMaxCrops));
    g_scratch_ptr[dev] = static_cast<uint8_t*>(p);
  }
  *out = g_scratch_ptr[dev];
  return UMT_OK;
}

}  // namespace

extern "C" int umt_model_encode_static(umt_model* m, int mode, const float* images, int n_img,
                                       float* features, void* stream) {
  if (!m) return umt::fail(UMT_ERR_INVALID, "umt_model_encode_static: model is null");
  if (mode < UMT_MODE_FP32_EQUIV || mode > UMT_MODE_FP16)
    return umt::fail(UMT_ERR_INVALID, "umt_model_encode_static: bad mode %d", mode);
  if (n_img < 0) return umt::fail(UMT_ERR_INVALID, "umt_model_encode_static: n_img=%d", n_img);
  if (n_img == 0) return UMT_OK;   // nothing to do, pointers may be null
  if (!images || !features) return umt::fail(UMT_ERR_INVALID, "umt_model_encode_static: null argument");
  if (n_img > kStaticMaxCrops)
    return umt::fail(UMT_ERR_INVALID,
                     "umt_model_encode_static: %d crops exceed the static scratch capacity of %d crops; "
                     "use umt_model_encode with a bound workspace for larger batches", n_img, kStaticMaxCrops);
  // The weights live on one device; the scratch address below is resolved for the CURRENT device.
  // A mismatch would hand kernels pointers of two different devices.
  UMT_TRY(umt::check_device(m, "umt_model_encode_static"));
  UMT_TRY(umt::check_mode(m, mode));
  uint8_t* scratch = nullptr;
  UMT_TRY(static_scratch_for_current_device(&scratch));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // The array keeps the previous call's activations; halo and guard rows must read as zero.
  UMT_CUDA_TRY(cudaMemsetAsync(scratch, 0, umt::encoder_workspace_bytes(kStaticMaxCrops), st));
  // A private view of the model whose activation tensors live in the static array; the caller's
  // handle (and any workspace bound to it) is left untouched.  The view is a host object built on
  // the first call and kept with the model: layer programs + device weight pointers, no device memory.
  if (!m->static_view) {
    umt_model* view = new umt_model();
    view->convs = m->convs;
    view->tw = m->tw;
    view->proj_conv = m->proj_conv;
    view->device = m->device;
    view->d_status = m->d_status;
    view->chunk_images = kStaticMaxCrops;
    view->ws = scratch;
    view->ws_bytes = kStaticScratchBytes;
    view->ws_max_images = kStaticMaxCrops;
    umt::encoder_bind(view, scratch);
    m->static_view = view;
  }
  m->static_view->launches = 0;
  const int rc = umt::encoder_run_chunk(m->static_view, mode, images, n_img, features, -1, st, nullptr);
  m->launches = m->static_view->launches;   // umt_model_last_launch_count(m) reports this call
  return rc;
}
