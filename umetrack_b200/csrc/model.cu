// Model handle, weight-tensor registry, workspace binding and the C ABI entry points of stage (b).
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "model.cuh"

namespace umt {

char* last_error_buffer() {
  static thread_local char buf[512] = {0};
  return buf;
}

// ---- weight tensor registry: the reference's state_dict (lib/models/model_loader.py:84-87) ----
struct TensorInfo {
  std::string name;
  int64_t numel;
};

static void add_bn(std::vector<TensorInfo>& v, const std::string& p, int c) {
  v.push_back({p + ".weight", c});
  v.push_back({p + ".bias", c});
  v.push_back({p + ".running_mean", c});
  v.push_back({p + ".running_var", c});
}

static void add_block(std::vector<TensorInfo>& v, const std::string& p, int cin, int cout, bool ds) {
  v.push_back({p + ".conv1.weight", (int64_t)cout * cin * 9});
  add_bn(v, p + ".bn1", cout);
  v.push_back({p + ".conv2.weight", (int64_t)cout * cout * 9});
  add_bn(v, p + ".bn2", cout);
  if (ds) {
    v.push_back({p + ".downsample.0.weight", (int64_t)cout * cin});
    add_bn(v, p + ".downsample.1", cout);
  }
}

static const std::vector<TensorInfo>& registry() {
  static std::vector<TensorInfo> v;
  if (!v.empty()) return v;
  const std::string b = "_feature_extractor._image_backbone.0._layers.";
  v.push_back({b + "0.0.weight", 32 * 9});
  v.push_back({b + "0.0.bias", 32});
  add_bn(v, b + "0.1", 32);
  const int blocks[4] = {2, 3, 5, 2}, planes[4] = {32, 64, 128, 256};
  int cin = 32;
  for (int g = 0; g < 4; ++g)
    for (int k = 0; k < blocks[g]; ++k) {
      add_block(v, b + std::to_string(g + 1) + "." + std::to_string(k), cin, planes[g], k == 0 && g > 0);
      cin = planes[g];
    }
  v.push_back({"_feature_extractor._image_backbone.1.weight", 72 * 256});
  v.push_back({"_feature_extractor._image_backbone.1.bias", 72});
  const std::string f = "_feature_extractor._multi_view_fusion.";
  v.push_back({f + "0.weight", 108 * 144});
  v.push_back({f + "0.bias", 108});
  add_bn(v, f + "1", 108);
  v.push_back({f + "3.weight", 72 * 108});
  v.push_back({f + "3.bias", 72});
  add_bn(v, f + "4", 72);
  v.push_back({f + "6.weight", 72 * 72});
  v.push_back({f + "6.bias", 72});
  for (int i = 0; i < 3; ++i) {
    const std::string t = "_temporal._temporal_module." + std::to_string(2 * i);
    v.push_back({t + ".weight", 90 * 90});
    v.push_back({t + ".bias", 90});
  }
  v.push_back({"_skeleton_enc._layers.0.weight", 144 * 132});
  v.push_back({"_skeleton_enc._layers.0.bias", 144});
  add_bn(v, "_skeleton_enc._layers.2", 4);
  const char* heads[2] = {"_regressor_k", "_regressor_u"};
  const int hc[2] = {76, 72}, ho[2] = {62, 63};
  for (int h = 0; h < 2; ++h) {
    const std::string p = std::string(heads[h]) + "._pose_regression_layers.";
    add_block(v, p + "0", hc[h], hc[h], false);
    add_block(v, p + "1", hc[h], hc[h], false);
    v.push_back({p + "2.weight", (int64_t)ho[h] * hc[h]});
    v.push_back({p + "2.bias", ho[h]});
  }
  return v;
}

int tensor_index(const char* name) {
  static std::map<std::string, int> idx;
  if (idx.empty()) {
    const auto& r = registry();
    for (size_t i = 0; i < r.size(); ++i) idx[r[i].name] = (int)i;
  }
  auto it = idx.find(name);
  return it == idx.end() ? -1 : it->second;
}

// UMT_MODE_FP16 needs every BatchNorm-folded weight inside the half range
int check_mode(const umt_model* m, int mode) {
  if (mode == UMT_MODE_FP16 && !m->fp16_weights_ok)
    return fail(UMT_ERR_UNSUPPORTED, "UMT_MODE_FP16: a BatchNorm-folded weight exceeds the half range (65504); "
                                     "use UMT_MODE_FP32_EQUIV or UMT_MODE_BF16 for these weights");
  return UMT_OK;
}

float* upload(umt_model* m, const std::vector<float>& v) {
  float* d = nullptr;
  if (cudaMalloc(&d, v.size() * sizeof(float)) != cudaSuccess) {
    fail(UMT_ERR_CUDA, "cudaMalloc of %zu bytes failed", v.size() * sizeof(float));
    return nullptr;
  }
  m->owned.push_back(d);
  if (cudaMemcpy(d, v.data(), v.size() * sizeof(float), cudaMemcpyHostToDevice) != cudaSuccess) {
    fail(UMT_ERR_CUDA, "weight upload failed");
    return nullptr;
  }
  return d;
}

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

int check_device(const umt_model* m, const char* what) {
  int dev = -1;
  UMT_CUDA_TRY(cudaGetDevice(&dev));
  if (dev != m->device)
    return fail(UMT_ERR_INVALID, "%s: the model lives on device %d but device %d is current", what, m->device, dev);
  return UMT_OK;
}

}  // namespace umt

using namespace umt;

extern "C" const char* umt_last_error(void) { return last_error_buffer(); }
extern "C" int umt_abi_version(void) { return 1; }
extern "C" int umt_model_num_tensors(void) { return (int)registry().size(); }
extern "C" const char* umt_model_tensor_name(int i) {
  const auto& r = registry();
  return (i >= 0 && i < (int)r.size()) ? r[i].name.c_str() : nullptr;
}
extern "C" int64_t umt_model_tensor_numel(int i) {
  const auto& r = registry();
  return (i >= 0 && i < (int)r.size()) ? r[i].numel : -1;
}

extern "C" int umt_model_create(const float* const* host_tensors, int n_tensors, umt_model** out) {
  if (!out) return fail(UMT_ERR_INVALID, "umt_model_create: out is null");
  *out = nullptr;
  if (!host_tensors || n_tensors != (int)registry().size())
    return fail(UMT_ERR_INVALID, "umt_model_create: expected %d tensors, got %d",
                (int)registry().size(), n_tensors);
  for (int i = 0; i < n_tensors; ++i)
    if (!host_tensors[i])
      return fail(UMT_ERR_INVALID, "umt_model_create: tensor %d (%s) is null", i, registry()[i].name.c_str());
  int dev = 0, major = 0;
  UMT_CUDA_TRY(cudaGetDevice(&dev));
  UMT_CUDA_TRY(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  if (major != 10)
    return fail(UMT_ERR_UNSUPPORTED, "umetrack_b200 needs an sm_100 GPU (tcgen05); device %d is sm_%d", dev, major * 10);
  umt_model* m = new umt_model();
  m->device = dev;
  if (cudaMalloc(&m->d_status, sizeof(unsigned)) != cudaSuccess || cudaMemset(m->d_status, 0, sizeof(unsigned)) != cudaSuccess) {
    delete m;
    return fail(UMT_ERR_CUDA, "umt_model_create: status word allocation failed");
  }
  m->owned.push_back(m->d_status);
  int rc = encoder_build(m, host_tensors);
  if (rc == UMT_OK) rc = tail_build(m, host_tensors);
  if (rc != UMT_OK) {
    umt_model_destroy(m);
    return rc;
  }
  *out = m;
  return UMT_OK;
}

extern "C" void umt_model_destroy(umt_model* m) {
  if (!m) return;
  for (void* p : m->owned) cudaFree(p);
  delete m->static_view;   // host object only: it borrows the weights and owns no device memory
  delete m;
}

// tail scratch: dense features, three 6x6 planar flat tensors of 144 channels, skeleton maps
static size_t tail_bytes(int max_images, int max_samples, size_t* feats_off, size_t* a_off,
                         size_t* b_off, size_t* c_off, size_t* s_off, int chunk_images = 0, size_t* u8_off = nullptr) {
  size_t o = 0;
  const size_t flat = align256(flat_tensor_bytes(6, 144, max_samples));
  *feats_off = o; o += align256((size_t)max_images * 72 * 36 * 4);
  *a_off = o; o += flat;
  *b_off = o; o += flat;
  *c_off = o; o += flat;
  *s_off = o; o += align256((size_t)max_samples * 4 * 36 * 4);
  if (u8_off) *u8_off = o;
  o += align256((size_t)chunk_images * 9216 * 4);   // float32 staging of one chunk of uint8 crops
  return o;
}

extern "C" int umt_model_workspace_bytes(const umt_model* m, int chunk_images, int max_images,
                                         int max_samples, size_t* bytes) {
  if (!m || !bytes || chunk_images < 1 || max_images <= 0 || max_samples <= 0)
    return fail(UMT_ERR_INVALID, "umt_model_workspace_bytes: bad arguments");
  size_t f, a, b, c, s;
  *bytes = encoder_workspace_bytes(chunk_images) + tail_bytes(max_images, max_samples, &f, &a, &b, &c, &s, chunk_images);
  return UMT_OK;
}

extern "C" int umt_model_bind_workspace(umt_model* m, void* workspace, size_t bytes, int chunk_images,
                                        int max_images, int max_samples, void* stream) {
  if (!m || !workspace) return fail(UMT_ERR_INVALID, "umt_model_bind_workspace: null argument");
  size_t need = 0;
  UMT_TRY(umt_model_workspace_bytes(m, chunk_images, max_images, max_samples, &need));
  if (bytes < need)
    return fail(UMT_ERR_INVALID, "umt_model_bind_workspace: %zu bytes given, %zu needed", bytes, need);
  if (((uintptr_t)workspace & 255) != 0)
    return fail(UMT_ERR_INVALID, "umt_model_bind_workspace: workspace must be 256-byte aligned");
  UMT_TRY(check_device(m, "umt_model_bind_workspace"));
  {  // the workspace must be memory of the model's own GPU (the kernels' tensor maps are built over it)
    cudaPointerAttributes attr{};
    if (cudaPointerGetAttributes(&attr, workspace) != cudaSuccess || attr.type != cudaMemoryTypeDevice ||
        attr.device != m->device) {
      cudaGetLastError();
      return fail(UMT_ERR_INVALID, "umt_model_bind_workspace: workspace is not device memory of GPU %d", m->device);
    }
  }
  // halo rows of the activation tensors must read as zero for the whole life of the binding
  UMT_CUDA_TRY(cudaMemsetAsync(workspace, 0, need, (cudaStream_t)stream));
  m->ws = (uint8_t*)workspace;
  m->chunk_images = chunk_images;
  m->ws_bytes = bytes;
  m->ws_max_images = max_images;
  m->ws_max_samples = max_samples;
  encoder_bind(m, m->ws);
  size_t f, a, b, c, s, u8;
  tail_bytes(max_images, max_samples, &f, &a, &b, &c, &s, chunk_images, &u8);
  uint8_t* t = m->ws + encoder_workspace_bytes(chunk_images);
  m->img_stage = (float*)(t + u8);
  m->feats = (float*)(t + f);
  m->tail_flat[0] = make_flat_tensor(t + a, 6, 144, max_samples);
  m->tail_flat[1] = make_flat_tensor(t + b, 6, 144, max_samples);
  m->tail_flat[2] = make_flat_tensor(t + c, 6, 144, max_samples);
  m->tail_skel = (float*)(t + s);
  return UMT_OK;
}

// uint8 crop -> float32 in [0,1]: `crop.astype(np.float32) / 255.0` (lib/tracker/tracker.py:211), a true
// IEEE division like numpy's.  16 pixels per thread.
__global__ void u8_to_unit_float_kernel(const uint8_t* __restrict__ src, float* __restrict__ dst, long long n16) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n16) return;
  const uint4 v = __ldg(reinterpret_cast<const uint4*>(src) + i);
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
  float4* o = reinterpret_cast<float4*>(dst) + 4 * i;
#pragma unroll
  for (int k = 0; k < 4; ++k)
    o[k] = make_float4(__fdiv_rn((float)(w[k] & 0xFF), 255.0f), __fdiv_rn((float)((w[k] >> 8) & 0xFF), 255.0f),
                       __fdiv_rn((float)((w[k] >> 16) & 0xFF), 255.0f), __fdiv_rn((float)(w[k] >> 24), 255.0f));
}

static int encode_all(umt_model* m, int mode, const float* images, const uint8_t* images_u8, int n_img,
                      float* features, cudaStream_t st) {
  for (int i0 = 0; i0 < n_img; i0 += m->chunk_images) {
    const int n = (n_img - i0 < m->chunk_images) ? n_img - i0 : m->chunk_images;
    const float* src = images ? images + (size_t)i0 * 9216 : m->img_stage;
    int extra = 0;
    if (!images) {
      const long long n16 = (long long)n * 9216 / 16;
      u8_to_unit_float_kernel<<<(unsigned)((n16 + 255) / 256), 256, 0, st>>>(images_u8 + (size_t)i0 * 9216, m->img_stage, n16);
      UMT_TRY(check_launch("u8_to_unit_float_kernel"));
      extra = 1;
    }
    const int before = m->launches;
    UMT_TRY(encoder_run_chunk(m, mode, src, n, features + (size_t)i0 * 72 * 36, -1, st));
    m->launches = before + (m->launches - before) + extra;
  }
  return UMT_OK;
}

extern "C" int umt_model_encode(umt_model* m, int mode, const float* images, int n_img,
                                float* features, void* stream) {
  if (!m || !images || !features) return fail(UMT_ERR_INVALID, "umt_model_encode: null argument");
  if (mode < UMT_MODE_FP32_EQUIV || mode > UMT_MODE_FP16) return fail(UMT_ERR_INVALID, "bad mode %d", mode);
  if (!m->ws) return fail(UMT_ERR_INVALID, "umt_model_encode: no workspace bound");
  UMT_TRY(check_device(m, "umt_model_encode"));
  UMT_TRY(check_mode(m, mode));
  if (n_img < 0) return fail(UMT_ERR_INVALID, "umt_model_encode: n_img=%d", n_img);
  m->launches = 0;
  if (n_img == 0) return UMT_OK;
  return encode_all(m, mode, images, nullptr, n_img, features, (cudaStream_t)stream);
}

extern "C" int umt_model_encode_u8(umt_model* m, int mode, const uint8_t* images_u8, int n_img, float* features,
                                   void* stream) {
  if (!m || !images_u8 || !features) return fail(UMT_ERR_INVALID, "umt_model_encode_u8: null argument");
  if (mode < UMT_MODE_FP32_EQUIV || mode > UMT_MODE_FP16) return fail(UMT_ERR_INVALID, "bad mode %d", mode);
  if (!m->ws) return fail(UMT_ERR_INVALID, "umt_model_encode_u8: no workspace bound");
  UMT_TRY(check_device(m, "umt_model_encode_u8"));
  UMT_TRY(check_mode(m, mode));
  if (((uintptr_t)images_u8 & 15) != 0) return fail(UMT_ERR_INVALID, "umt_model_encode_u8: images must be 16-byte aligned");
  if (n_img < 0) return fail(UMT_ERR_INVALID, "umt_model_encode_u8: n_img=%d", n_img);
  m->launches = 0;
  if (n_img == 0) return UMT_OK;
  return encode_all(m, mode, nullptr, images_u8, n_img, features, (cudaStream_t)stream);
}

static int model_forward(umt_model* m, int mode, int head, const float* images, const uint8_t* images_u8,
                                 const float* intrinsics, const float* extrinsics,
                                 const int64_t* sample_range, const int64_t* memory_idx,
                                 const uint8_t* use_memory, const int64_t* hand_idx, int n_img,
                                 int n_samples, const float* skel_axes, const float* skel_rest,
                                 int n_skel, float* mem_state, float* prev_extr, int n_slots,
                                 float* out_joint_angles, float* out_wrist, float* out_scale,
                                 float* out_sigma, void* stream) {
  if (!m) return fail(UMT_ERR_INVALID, "umt_model_forward: model is null");
  if (mode < UMT_MODE_FP32_EQUIV || mode > UMT_MODE_FP16) return fail(UMT_ERR_INVALID, "bad mode %d", mode);
  if (head != UMT_HEAD_KNOWN_SKELETON && head != UMT_HEAD_UNKNOWN_SKELETON)
    return fail(UMT_ERR_INVALID, "bad head %d", head);
  if (!m->ws) return fail(UMT_ERR_INVALID, "umt_model_forward: no workspace bound");
  UMT_TRY(check_device(m, "umt_model_forward"));
  UMT_TRY(check_mode(m, mode));
  if (n_img <= 0 || n_samples <= 0 || n_img > m->ws_max_images || n_samples > m->ws_max_samples)
    return fail(UMT_ERR_INVALID, "umt_model_forward: n_img=%d n_samples=%d exceed the bound workspace (%d, %d)",
                n_img, n_samples, m->ws_max_images, m->ws_max_samples);
  if ((!images && !images_u8) || !intrinsics || !extrinsics || !sample_range || !memory_idx || !use_memory ||
      !hand_idx || !mem_state || !prev_extr || !out_joint_angles || !out_wrist || !out_sigma)
    return fail(UMT_ERR_INVALID, "umt_model_forward: null pointer argument");
  if (images_u8 && ((uintptr_t)images_u8 & 15) != 0)
    return fail(UMT_ERR_INVALID, "umt_model_forward_u8: images must be 16-byte aligned");
  if (head == UMT_HEAD_KNOWN_SKELETON) {
    if (!skel_axes || !skel_rest || !(n_skel == 1 || n_skel == n_samples))
      return fail(UMT_ERR_INVALID, "umt_model_forward: known-skeleton head needs 1 or n_samples skeletons");
  } else if (!out_scale) {
    return fail(UMT_ERR_INVALID, "umt_model_forward: unknown-skeleton head needs out_scale");
  }
  if (n_slots <= 0) return fail(UMT_ERR_INVALID, "umt_model_forward: n_slots=%d", n_slots);
  cudaStream_t st = (cudaStream_t)stream;
  m->launches = 0;
  UMT_TRY(encode_all(m, mode, images, images_u8, n_img, m->feats, st));
  return tail_forward(m, head, m->feats, intrinsics, extrinsics, sample_range, memory_idx, use_memory,
                      hand_idx, n_img, n_samples, skel_axes, skel_rest, n_skel, mem_state, prev_extr,
                      n_slots, out_joint_angles, out_wrist, out_scale, out_sigma, st);
}

extern "C" int umt_model_forward(umt_model* m, int mode, int head, const float* images,
                                 const float* intrinsics, const float* extrinsics,
                                 const int64_t* sample_range, const int64_t* memory_idx,
                                 const uint8_t* use_memory, const int64_t* hand_idx, int n_img,
                                 int n_samples, const float* skel_axes, const float* skel_rest,
                                 int n_skel, float* mem_state, float* prev_extr, int n_slots,
                                 float* out_joint_angles, float* out_wrist, float* out_scale,
                                 float* out_sigma, void* stream) {
  if (!images) return fail(UMT_ERR_INVALID, "umt_model_forward: null pointer argument");
  return model_forward(m, mode, head, images, nullptr, intrinsics, extrinsics, sample_range, memory_idx, use_memory,
                       hand_idx, n_img, n_samples, skel_axes, skel_rest, n_skel, mem_state, prev_extr, n_slots,
                       out_joint_angles, out_wrist, out_scale, out_sigma, stream);
}

extern "C" int umt_model_forward_u8(umt_model* m, int mode, int head, const uint8_t* images_u8,
                                    const float* intrinsics, const float* extrinsics,
                                    const int64_t* sample_range, const int64_t* memory_idx,
                                    const uint8_t* use_memory, const int64_t* hand_idx, int n_img,
                                    int n_samples, const float* skel_axes, const float* skel_rest,
                                    int n_skel, float* mem_state, float* prev_extr, int n_slots,
                                    float* out_joint_angles, float* out_wrist, float* out_scale,
                                    float* out_sigma, void* stream) {
  if (!images_u8) return fail(UMT_ERR_INVALID, "umt_model_forward_u8: null pointer argument");
  return model_forward(m, mode, head, nullptr, images_u8, intrinsics, extrinsics, sample_range, memory_idx, use_memory,
                       hand_idx, n_img, n_samples, skel_axes, skel_rest, n_skel, mem_state, prev_extr, n_slots,
                       out_joint_angles, out_wrist, out_scale, out_sigma, stream);
}

// Tools only (not declared in the public header): device buffer [grid][16] of int64 that every
// later conv launch fills with per-role wait/total cycle counters; nullptr switches it off.
extern "C" void umt_debug_set_conv_timeline(long long* device_buffer) { set_conv_debug_buffer(device_buffer); }

extern "C" int umt_model_tail(umt_model* m, int head, const float* features, const float* intrinsics,
                              const float* extrinsics, const int64_t* sample_range,
                              const int64_t* memory_idx, const uint8_t* use_memory,
                              const int64_t* hand_idx, int n_img, int n_samples, const float* skel_axes,
                              const float* skel_rest, int n_skel, float* mem_state, float* prev_extr,
                              int n_slots, float* out_joint_angles, float* out_wrist, float* out_scale,
                              float* out_sigma, void* stream) {
  if (!m) return fail(UMT_ERR_INVALID, "umt_model_tail: model is null");
  if (head != UMT_HEAD_KNOWN_SKELETON && head != UMT_HEAD_UNKNOWN_SKELETON)
    return fail(UMT_ERR_INVALID, "bad head %d", head);
  if (!m->ws) return fail(UMT_ERR_INVALID, "umt_model_tail: no workspace bound");
  UMT_TRY(check_device(m, "umt_model_tail"));
  if (n_img <= 0 || n_samples <= 0 || n_samples > m->ws_max_samples)
    return fail(UMT_ERR_INVALID, "umt_model_tail: n_samples=%d exceeds the bound workspace (%d)", n_samples,
                m->ws_max_samples);
  if (!features || !intrinsics || !extrinsics || !sample_range || !memory_idx || !use_memory || !hand_idx ||
      !mem_state || !prev_extr || !out_joint_angles || !out_wrist || !out_sigma)
    return fail(UMT_ERR_INVALID, "umt_model_tail: null pointer argument");
  if (head == UMT_HEAD_KNOWN_SKELETON) {
    if (!skel_axes || !skel_rest || !(n_skel == 1 || n_skel == n_samples))
      return fail(UMT_ERR_INVALID, "umt_model_tail: known-skeleton head needs 1 or n_samples skeletons");
  } else if (!out_scale) {
    return fail(UMT_ERR_INVALID, "umt_model_tail: unknown-skeleton head needs out_scale");
  }
  if (n_slots <= 0) return fail(UMT_ERR_INVALID, "umt_model_tail: n_slots=%d", n_slots);
  m->launches = 0;
  return tail_forward(m, head, features, intrinsics, extrinsics, sample_range, memory_idx, use_memory,
                      hand_idx, n_img, n_samples, skel_axes, skel_rest, n_skel, mem_state, prev_extr, n_slots,
                      out_joint_angles, out_wrist, out_scale, out_sigma, (cudaStream_t)stream);
}

extern "C" int umt_model_tail_scan(umt_model* m, int head, const float* features, const float* intrinsics,
                                   const float* extrinsics, const int64_t* sample_range,
                                   const int64_t* memory_idx, const uint8_t* use_memory,
                                   const int64_t* hand_idx, int n_img, const int* step_first,
                                   const int* step_rows, int n_steps, const float* skel_axes,
                                   const float* skel_rest, int per_row_skeletons, float* mem_state,
                                   float* prev_extr, int n_slots, float* out_joint_angles, float* out_wrist,
                                   float* out_scale, float* out_sigma, void* stream) {
  if (n_steps < 0 || (n_steps > 0 && (!step_first || !step_rows)))
    return fail(UMT_ERR_INVALID, "umt_model_tail_scan: bad step list");
  int launches = 0;
  for (int s = 0; s < n_steps; ++s) {
    const long long f = step_first[s];
    const int n = step_rows[s];
    if (n == 0) continue;
    if (f < 0 || n < 0) return fail(UMT_ERR_INVALID, "umt_model_tail_scan: step %d = (%lld, %d)", s, f, n);
    const float* ax = skel_axes ? skel_axes + (per_row_skeletons ? f * 66 : 0) : nullptr;
    const float* rs = skel_rest ? skel_rest + (per_row_skeletons ? f * 66 : 0) : nullptr;
    UMT_TRY(umt_model_tail(m, head, features, intrinsics, extrinsics, sample_range + 2 * f, memory_idx + f,
                           use_memory + f, hand_idx + f, n_img, n, ax, rs, per_row_skeletons ? n : 1, mem_state,
                           prev_extr, n_slots, out_joint_angles + 22 * f, out_wrist + 16 * f,
                           out_scale ? out_scale + f : nullptr, out_sigma + 21 * f, stream));
    launches += m->launches;
  }
  if (m) m->launches = launches;
  return UMT_OK;
}

extern "C" int umt_model_last_launch_count(const umt_model* m) { return m ? m->launches : 0; }

extern "C" int umt_model_status(umt_model* m, unsigned* host_status, void* stream) {
  if (!m || !host_status) return fail(UMT_ERR_INVALID, "umt_model_status: null argument");
  UMT_TRY(check_device(m, "umt_model_status"));
  cudaStream_t st = (cudaStream_t)stream;
  UMT_CUDA_TRY(cudaMemcpyAsync(host_status, m->d_status, sizeof(unsigned), cudaMemcpyDeviceToHost, st));
  UMT_CUDA_TRY(cudaMemsetAsync(m->d_status, 0, sizeof(unsigned), st));
  UMT_CUDA_TRY(cudaStreamSynchronize(st));
  return UMT_OK;
}

extern "C" int umt_model_debug_layer(umt_model* m, int mode, int layer, const float* images, int n_img,
                                     float* out_simt, float* out_umma, int* out_shape, void* stream) {
  if (!m || !images) return fail(UMT_ERR_INVALID, "umt_model_debug_layer: null argument");
  if (!m->ws) return fail(UMT_ERR_INVALID, "umt_model_debug_layer: no workspace bound");
  return encoder_debug_layer(m, mode, layer, images, n_img, out_simt, out_umma, out_shape,
                             (cudaStream_t)stream);
}

extern "C" int umt_model_debug_halo_check(umt_model* m, long long* host_nonzero_rows, void* stream) {
  if (!m || !host_nonzero_rows) return fail(UMT_ERR_INVALID, "umt_model_debug_halo_check: null argument");
  UMT_TRY(check_device(m, "umt_model_debug_halo_check"));
  return encoder_halo_check(m, host_nonzero_rows, (cudaStream_t)stream);
}

extern "C" int umt_model_profile_layers(umt_model* m, int mode, const float* images, int n_img,
                                        float* out_ms, void* stream) {
  if (!m || !images || !out_ms) return fail(UMT_ERR_INVALID, "umt_model_profile_layers: null argument");
  if (!m->ws) return fail(UMT_ERR_INVALID, "umt_model_profile_layers: no workspace bound");
  if (n_img <= 0 || n_img > m->chunk_images || n_img > m->ws_max_images)
    return fail(UMT_ERR_INVALID, "umt_model_profile_layers: n_img=%d must fit one chunk (%d)", n_img, m->chunk_images);
  cudaStream_t st = (cudaStream_t)stream;
  constexpr int kEv = 27;  // stem, 24 convolutions, projection, end
  if ((int)m->convs.size() != 24) return fail(UMT_ERR_INVALID, "unexpected layer count");
  cudaEvent_t ev[kEv];
  for (int i = 0; i < kEv; ++i) UMT_CUDA_TRY(cudaEventCreate(&ev[i]));
  m->launches = 0;
  int rc = encoder_run_chunk(m, mode, images, n_img, m->feats, -1, st, ev);
  if (rc == UMT_OK) {
    cudaError_t e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) rc = fail(UMT_ERR_CUDA, "profile sync: %s", cudaGetErrorString(e));
  }
  if (rc == UMT_OK)
    for (int i = 0; i < kEv - 1; ++i) cudaEventElapsedTime(&out_ms[i], ev[i], ev[i + 1]);
  for (int i = 0; i < kEv; ++i) cudaEventDestroy(ev[i]);
  return rc;
}
