// Internal declarations shared by encoder.cu (backbone), tail.cu (fusion / temporal / heads)
// and model.cu (handle, workspace, C ABI).  Not part of the public interface.
#pragma once
#include <vector>

#include "common.cuh"

namespace umt {

// ----------------------------------------------------------------------------------------------
// Activation layout of the backbone ("planar flat"):
//   An activation tensor for N images of H x W pixels and C channels is stored as C/8 planes.
//   A plane is a 1-D array of 16-byte rows, one row = 8 consecutive bf16 channels of one
//   position.  Positions enumerate a halo-padded raster: image i owns (H+1)*(W+1) positions,
//   q = i*(H+1)*(W+1) + y*(W+1) + x, where column x == W and row y == H are halo positions that
//   always hold zeros.  Because the right halo of one row is the left halo of the next and the
//   bottom halo row of one image is the top halo of the next, a 3x3 tap (dy,dx) of position q
//   is position q + dy*(W+1) + dx -- a constant row shift, for every position of every image.
//   `guard` zero rows precede position 0 and follow the last tile.
//   Values are split bf16: x ~= hi + lo, the lo planes live `lo_offset` bytes after the hi
//   planes (UMT_MODE_BF16 reads and writes only hi).
// ----------------------------------------------------------------------------------------------
struct FlatTensor {
  uint8_t* base;            // hi plane 0, first guard row
  long long plane_stride;   // bytes between planes
  long long lo_offset;      // bytes from a hi plane to the matching lo plane
  int guard;                // guard rows before position 0
  int pad_;
};

constexpr int kMaxStages = 40;
constexpr int kMaxTaps = 16;

struct ConvParams {
  FlatTensor src[5];
  FlatTensor out[4];        // [0] for flat output; four parity-phase tensors when phase_out
  const uint8_t* weights;   // packed operand slabs in program order
  const uint8_t* pair_weights;   // non-null: run through the CTA-pair kernel (conv_pair.cu) with this blob
  int pair_G;                    // taps per weight stage of that blob
  float* dense_out;         // if set: write float32 NCHW [n, dense_c, H, W] instead of the planar layout
  int dense_c;
  long long* dbg;           // optional [grid][16] cycle counters (tools/conv_timeline.py); nullptr in production
  const float* bias;        // [COUT]
  unsigned* status;         // device status word (UMT_STATUS_*), may be null
  int H, W;                 // output-domain geometry
  int n_pos;                // valid positions: n_img * (H+1) * (W+1)
  int S;                    // 128-row sub-tiles per CTA
  int lo_shift;             // most negative tap shift (rows)
  int patch_rows;           // S*128 + hi_shift - lo_shift
  int halo_rows;            // hi_shift - lo_shift
  int n_a, n_b;             // pipeline depths (A patches, B slabs)
  int bgroup;               // weight slabs (taps) per B pipeline stage
  int b_resident;           // 1: the layer's weights fit: every CTA loads them once and keeps them (n_b = stages per tile)
  int overlap_launch;       // 1: programmatic dependent launch (set-up and weight stream overlap the previous layer)
  // Output-channel slices (few positions, wide layers): CTA b works on slice b % n_slices of tile
  // b / n_slices; a slice is COUT channels of the layer with its own weight blob and bias, and writes /
  // adds the planes [slice * COUT / 8, ...) of the output / residual tensor.  0 or 1: no slicing.
  int n_slices;
  int force_S;              // >0: at most this many sub-tiles per CTA (sliced launches want many small tiles)
  const uint8_t* slice_weights[4];
  const float* slice_bias[4];
  int n_buf;                // TMEM accumulator buffers (1 or 2)
  int cluster;              // CTAs per cluster sharing (multicasting) the weight slabs
  int occ;                  // CTAs per SM the grid is sized for
  int n_tiles;              // tiles of S*128 positions, padded to a multiple of `cluster`
  int n_stages;
  int phase_out;            // 1: scatter the output into four (H/2, W/2) parity-phase tensors
  int relu;                 // 1: ReLU in the epilogue
  int res_epi;              // 1: the epilogue adds the identity residual `res` (same raster as the output)
  int res_stages;           // trailing program stages that feed the same residual through identity slabs
  FlatTensor res;
  uint8_t st_src[kMaxStages], st_chunk[kMaxStages], st_ntaps[kMaxStages], st_tap0[kMaxStages];
  uint8_t st_flags[kMaxStages];   // bit 0: weight slab has no lo part (identity slab of a residual)
  short tap_shift[kMaxTaps];
  uint8_t tap_src[kMaxTaps];   // source of a tap relative to its stage's first source (merged stages)
  int a_nsrc;                  // consecutive source tensors whose patches share one A stage (1, 2 or 4)
};

struct ConvLayer {
  int cout = 0, cin = 0, stride = 1, geom = 0;   // geom: index into {48,24,12,6}
  int ds_cin = 0;                                 // >0: fused 1x1 stride-2 downsample branch
  int src_kind = 0;                               // 0: flat buffer, 1: phase tensors
  int src_buf = 0, out_buf = 0, res_buf = -1;     // buffer ids within the geometry
  int phase_out = 0;
  int relu = 1;
  int n_slabs = 0;
  uint8_t* d_weights = nullptr;                   // packed [slab][hi|lo][2][cout][8] bf16
  uint8_t* d_weights_hi = nullptr;                // packed [slab][2][cout][8] bf16 (hi only)
  uint8_t* d_weights_h16 = nullptr;               // fp16 weights, one-CTA packing [slab][2][cout][8]
  uint8_t* d_pair[3] = {nullptr, nullptr, nullptr};   // pair-kernel blobs: split bf16, bf16, fp16 (conv_pair.cu)
  int pair_G[3] = {0, 0, 0};
  float* d_bias = nullptr;
  ConvParams prog{};                              // program template (tensors filled per call)
  ConvParams prog16{};                            // stride-2 layers: stage table used by the 16-bit modes
  int has_prog16 = 0;
  int no_pair = 0;                                // 1: no pair-kernel blobs (the slices below)
  std::vector<ConvLayer> slices;                  // 128 / 256 channels: the same layer as 2 / 4 layers of 64 output
                                                  // channels each, for launches on few positions (ConvParams::n_slices)
};

struct Folded {           // conv weight with BatchNorm folded in: w[co][ci][ks*ks], b[co]
  std::vector<float> w, b;
  int cout, cin, ks;
};

// Cluster variant of the fused small-batch tail (tail_fused.cu): a sample is spread over kClusterCtas CTAs,
// CTA r owns ceil(cout / kClusterCtas) <= kSliceStride output channels of every layer; its packed slice is
// [rows][kSliceStride] weights + [kSliceStride] biases (rows = cin, x 9 taps for a 3x3 layer; a constant row
// stride keeps the kernel's shared-memory offsets compile-time constants).
constexpr int kClusterCtas = 8;
constexpr int kSliceStride = 16;
__host__ __device__ constexpr int cluster_slice_floats(int rows) { return (rows + 1) * kSliceStride; }

struct TailWeights {
  // BatchNorm-folded float32 weights, device pointers
  float *proj_w, *proj_b;                 // [256,72] (transposed), [72]: CUDA-core projection (UMT_PROJ_SIMT=1)
  float *skel_w, *skel_b;                 // [144,132] with BN(4) folded, [144]
  float *head_w[2], *head_b[2];           // [n_out, 80] (zero columns for the channel padding), [n_out]
  float *stem_w, *stem_b;                 // [32,9] folded, [32]
  // float32 weights of the one-CTA-per-sample tail kernel (tail_fused.cu), input-channel major: [ci][tap][co]
  float *fus_w[3], *fus_b[3];             // fusion 144->108->72->72 (BN folded)
  float *tmp_w[3], *tmp_b[3];             // temporal 90->90 x3
  float *reg_w[2][4], *reg_b[2][4];       // regressor_k (C = 76) / regressor_u (C = 72): four 3x3 convs (BN folded)
  float *fus_cl[3], *tmp_cl[3], *reg_cl[2][4];   // the same, packed per CTA of the cluster variant (tail_fused.cu)
  uint8_t *stem_pack_bf, *stem_pack_h16;  // tensor-core stem operand [hi|lo][2][128][8] bf16 / [2][128][8] fp16
};

}  // namespace umt

struct umt_model {
  int chunk_images = 0;
  std::vector<umt::ConvLayer> convs;      // 24 backbone convolutions in execution order
  umt::ConvLayer proj_conv;               // 256 -> 72 projection as a tensor-core layer
  umt::TailWeights tw{};
  std::vector<void*> owned;               // device allocations to free
  // workspace binding
  uint8_t* ws = nullptr;
  size_t ws_bytes = 0;
  int ws_max_images = 0, ws_max_samples = 0;
  umt::FlatTensor buf[4][3]{};            // per geometry: ping, pong, mid
  umt::FlatTensor phase[4][4]{};          // phase[g]: input of geometry g (g=1..3), 4 parities
  float* feats = nullptr;                 // [max_images,72,36]
  float* img_stage = nullptr;             // [chunk_images,96,96] float32 staging of uint8 crops
  float* tail_skel = nullptr;             // [max_samples,4,36]
  std::vector<umt::ConvLayer> tail_convs; // fusion(3), temporal(3), regressor_k(4), regressor_u(4)
  umt::FlatTensor tail_flat[3]{};         // 6x6 planar flat scratch, 144 channels capacity
  int launches = 0;
  bool fp16_weights_ok = true;            // no BN-folded weight saturates in half precision (UMT_MODE_FP16 allowed)
  int device = -1;                        // device the weights were uploaded to (umt_model_create)
  unsigned* d_status = nullptr;           // sticky UMT_STATUS_* bits set by kernels, read by umt_model_status
  umt_model* static_view = nullptr;       // host-side view used by umt_model_encode_static (no device memory)
};

namespace umt {
// encoder.cu
int encoder_build(umt_model* m, const float* const* t);
int encoder_run_chunk(umt_model* m, int mode, const float* images, int n_img, float* features,
                      int stop_after_layer, cudaStream_t st, cudaEvent_t* ev = nullptr);
int encoder_debug_layer(umt_model* m, int mode, int layer, const float* images, int n_img,
                        float* out_simt, float* out_umma, int* shape, cudaStream_t st);
int encoder_halo_check(umt_model* m, long long* host_nonzero, cudaStream_t st);
size_t encoder_workspace_bytes(int chunk_images);
void encoder_bind(umt_model* m, uint8_t* base);
void set_conv_debug_buffer(long long* p);
long long* next_conv_debug_slot();
int build_conv(umt_model* m, ConvLayer& L, const Folded& main, const Folded* ds);
int launch_conv(const ConvParams& p, int cout, int mode, cudaStream_t st);
void finalize_conv_params(ConvParams& p, const ConvLayer& L, int mode);
size_t flat_tensor_bytes(int H, int C, int n_img);
FlatTensor make_flat_tensor(uint8_t* base, int H, int C, int n_img);
// conv_pair.cu
int launch_conv_pair(const ConvParams& p, int cout, int mode, cudaStream_t st);
int pair_kernel_enabled(int cout, int mode);
void pack_pair_blob(const std::vector<uint16_t>& blob, int cout, const std::vector<int>& stage_ntaps, int G, int kind,
                    uint16_t (*to_half)(float), float (*from_bf)(uint16_t), std::vector<uint16_t>& out);
// tail.cu
int tail_build(umt_model* m, const float* const* t);
int tail_forward(umt_model* m, int head, const float* feats, const float* intrinsics,
                 const float* extrinsics, const int64_t* sample_range, const int64_t* memory_idx,
                 const uint8_t* use_memory, const int64_t* hand_idx, int n_img, int n_samples,
                 const float* skel_axes, const float* skel_rest, int n_skel, float* mem_state,
                 float* prev_extr, int n_slots, float* out_joint_angles, float* out_wrist,
                 float* out_scale, float* out_sigma, cudaStream_t st);
// tail_fused.cu
int tail_fused_max_samples();
int tail_fused_forward(umt_model* m, int head, const float* feats, const float* intrinsics,
                       const float* extrinsics, const int64_t* sample_range, const int64_t* memory_idx,
                       const uint8_t* use_memory, const int64_t* hand_idx, int n_samples, const float* skel,
                       int n_skel, float* mem_state, float* prev_extr, int n_slots, float* out_joint_angles,
                       float* out_wrist, float* out_scale, float* out_sigma, cudaStream_t st);
// model.cu
int check_device(const umt_model* m, const char* what);
int check_mode(const umt_model* m, int mode);      // UMT_MODE_FP16 only with weights inside the half range   // current device must be the model's
int tensor_index(const char* name);   // index into the umt_model_create tensor list, -1 if absent
float* upload(umt_model* m, const std::vector<float>& v);
}  // namespace umt
