// Stage (b), part 2: everything after the per-view encoder (1 % of the FLOPs).  The ten
// convolutions (fusion 1x1 x3, temporal 1x1 x3, regressor 3x3 x4) run through the same tensor-core
// kernel as the backbone, always in the fp32-equivalent split-bf16 mode, on 6x6 planar flat
// tensors whose channel counts are padded to multiples of 16 with zero weights; the glue between
// them (FTL transforms, memory bank, skeleton encoder, decoders, Procrustes) is float32/float64
// CUDA-core code.
//   feature-transform layer + two-view fusion   lib/models/feature_extractor.py:61-141,
//                                                lib/models/model_utils.py:57-104,141-192
//   temporal memory (SimpleConvRNN)              lib/models/temporal.py:51-139
//   skeleton encoder                             lib/models/skeleton_encoder.py:28-53
//   pose regressor + decoders + Procrustes       lib/models/regressor.py:76-186,
//                                                lib/models/model_utils.py:17-54
//   world-space recovery / right-hand un-mirror  lib/models/umetrack_model.py:77-97
// The glue kernels work on one sample per CTA with the sample's [C][36] tile in shared memory
// (channel-major, the layout the reference's FTL reshape "C channels -> 3 x C/3*36 points" is
// defined on) and convert from/to the planar flat layout with 16-byte rows.
#include <cmath>
#include <string>

#include "model.cuh"
#include "flat.cuh"
#include "tail_math.cuh"

namespace umt {

// ------------------------------------------------------------------------------------------
// sample tile <-> planar flat (6x6 geometry, split bf16)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ long long flat_row6(int b, int p) { return (long long)b * kPos + (p / 6) * 7 + (p % 6); }

// tile[(plane*8+e)*36 + p] <- tensor
__device__ void load_sample_tile(const FlatTensor& T, int b, int nplanes, float* tile) {
  for (int i = threadIdx.x; i < nplanes * kPix; i += blockDim.x) {
    const int pl = i / kPix, p = i % kPix;
    float v[8];
    load_row8<0>(T, pl, flat_row6(b, p), v);
#pragma unroll
    for (int e = 0; e < 8; ++e) tile[(pl * 8 + e) * kPix + p] = v[e];
  }
}

__device__ void store_sample_tile(const FlatTensor& T, int b, int nplanes, const float* tile) {
  for (int i = threadIdx.x; i < nplanes * kPix; i += blockDim.x) {
    const int pl = i / kPix, p = i % kPix;
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = tile[(pl * 8 + e) * kPix + p];
    store_row8<0>(T, pl, flat_row6(b, p), v);
  }
}

// ------------------------------------------------------------------------------------------
// 1. views -> canonical space, channel-concatenated for the fusion convs
//    (feature_extractor.py:61-87,121-133).  One CTA per sample; output 144 channels.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
fuse_prepare_kernel(const float* __restrict__ feats, const float* __restrict__ intrinsics,
                    const float* __restrict__ extrinsics, const int64_t* __restrict__ sample_range,
                    FlatTensor canon) {
  __shared__ float sxf[2][12];
  __shared__ float tile[144 * kPix];
  const int b = blockIdx.x;
  const int r0 = (int)sample_range[2 * b], r1 = (int)sample_range[2 * b + 1];
  if (r1 - r0 != 2) {  // single-view samples skip the fusion convs (feature_extractor.py:89-94)
    for (int i = threadIdx.x; i < 144 * kPix; i += blockDim.x) tile[i] = 0.f;
    __syncthreads();
    store_sample_tile(canon, b, 18, tile);
    return;
  }
  if (threadIdx.x < 2) {
    const int v = threadIdx.x;
    double E0[16], Ev[16], Einv[16], S[16] = {0}, S0inv[16] = {0}, t1[16], t2[16], t3[16];
    load_mat4(extrinsics + (size_t)r0 * 16, E0);
    load_mat4(extrinsics + (size_t)(r0 + v) * 16, Ev);
    const float f0 = intrinsics[(size_t)r0 * 9] / 200.0f;          // S = diag(1,1,fx/200,1)
    const float fv = intrinsics[(size_t)(r0 + v) * 9] / 200.0f;    // (model_utils.py:166-192)
    S[0] = S[5] = S[15] = 1.0; S[10] = (double)fv;
    S0inv[0] = S0inv[5] = S0inv[15] = 1.0; S0inv[10] = (double)(1.0f / f0);
    bool ok = mat4_inv(Ev, Einv);
    mat4_mul(Einv, S, t1);       // view -> world
    mat4_mul(E0, t1, t2);        // -> cam0
    mat4_mul(S0inv, t2, t3);     // -> canonical
    for (int i = 0; i < 12; ++i) sxf[v][i] = ok ? (float)t3[i] : nanf("");
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * 864; i += blockDim.x) {
    const int v = i / 864, j = i % 864;
    const float* src = feats + (size_t)(r0 + v) * 72 * kPix;
    float ox, oy, oz;
    ftl_point(sxf[v], src[j], src[864 + j], src[1728 + j], ox, oy, oz);
    float* d = tile + v * 72 * kPix;
    d[j] = ox;
    d[864 + j] = oy;
    d[1728 + j] = oz;
  }
  __syncthreads();
  store_sample_tile(canon, b, 18, tile);
}

// ------------------------------------------------------------------------------------------
// 2. canonical -> cam0 (feature_extractor.py:135-139) or single-view FTL (:89-94), memory warp
//    and bookkeeping of SimpleConvRNN.transform_memory_features (temporal.py:51-78), and the
//    channel concat [memory(18) | image features(72) | 6 zero channels] (temporal.py:83-85).
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
temporal_prepare_kernel(FlatTensor fused_raw /* 80 ch, 72 valid */, const float* __restrict__ feats,
                        const float* __restrict__ intrinsics, const float* __restrict__ extrinsics,
                        const int64_t* __restrict__ sample_range, const int64_t* __restrict__ memory_idx,
                        const uint8_t* __restrict__ use_memory, float* __restrict__ mem_state,
                        float* __restrict__ prev_extr, int n_slots, FlatTensor tin /* 96 ch */,
                        unsigned* __restrict__ status) {
  __shared__ float sS[12];
  __shared__ float sM[12];
  __shared__ int sUse;
  __shared__ float sin_[72 * kPix];
  __shared__ float sout[96 * kPix];
  const int b = blockIdx.x;
  const int r0 = (int)sample_range[2 * b], r1 = (int)sample_range[2 * b + 1];
  const int slot = (int)memory_idx[b];
  const bool slot_ok = slot >= 0 && slot < n_slots;
  if (threadIdx.x == 0) {
    // a memory index outside the bank runs without memory (the reference would have grown the bank,
    // temporal.py:101-123): report it instead of hiding it
    if (!slot_ok) atomicOr(status, (unsigned)UMT_STATUS_MEMORY_INDEX);
    for (int i = 0; i < 12; ++i) sS[i] = 0.f;
    sS[0] = sS[5] = 1.f;
    sS[10] = intrinsics[(size_t)r0 * 9] / 200.0f;
    int use = (slot_ok && use_memory[b]) ? 1 : 0;
    if (use) {
      double P[16], Pinv[16], C[16], X[16];
      load_mat4(prev_extr + (size_t)slot * 16, P);
      load_mat4(extrinsics + (size_t)r0 * 16, C);
      if (mat4_inv(P, Pinv)) {
        mat4_mul(C, Pinv, X);   // previous cam0 -> current cam0 (temporal.py:66-69)
        for (int i = 0; i < 12; ++i) sM[i] = (float)X[i];
      } else {
        use = 0;                // the reference cannot reach this state (torch.inverse raises)
      }
    }
    sUse = use;
  }
  const bool multi = (r1 - r0) == 2;
  if (multi) {
    load_sample_tile(fused_raw, b, 9, sin_);
  } else {
    const float* src = feats + (size_t)r0 * 72 * kPix;
    for (int i = threadIdx.x; i < 72 * kPix; i += blockDim.x) sin_[i] = src[i];
  }
  __syncthreads();
  const float* mem = slot_ok ? mem_state + (size_t)slot * 18 * kPix : nullptr;
  for (int j = threadIdx.x; j < 216; j += blockDim.x) {   // memory: 18 channels = 216 points
    float ox = 0.f, oy = 0.f, oz = 0.f;
    if (sUse) ftl_point(sM, mem[j], mem[216 + j], mem[432 + j], ox, oy, oz);
    sout[j] = ox;
    sout[216 + j] = oy;
    sout[432 + j] = oz;
  }
  float* o2 = sout + 18 * kPix;
  for (int j = threadIdx.x; j < 864; j += blockDim.x) {   // image features: 72 channels = 864 points
    float ox, oy, oz;
    ftl_point(sS, sin_[j], sin_[864 + j], sin_[1728 + j], ox, oy, oz);
    o2[j] = ox;
    o2[864 + j] = oy;
    o2[1728 + j] = oz;
  }
  for (int j = threadIdx.x; j < 6 * kPix; j += blockDim.x) sout[90 * kPix + j] = 0.f;
  // prev_extrinsics[memory_idx] = cur_extrinsics (temporal.py:77)
  if (slot_ok && threadIdx.x < 16)
    prev_extr[(size_t)slot * 16 + threadIdx.x] = extrinsics[(size_t)r0 * 16 + threadIdx.x];
  __syncthreads();
  store_sample_tile(tin, b, 12, sout);
}

// ------------------------------------------------------------------------------------------
// 3. skeleton encoder: Linear(132->144) -> view(4,6,6) -> BN(4) -> ReLU, BN folded on the host.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(144)
skeleton_kernel(const float* __restrict__ axes, const float* __restrict__ rest,
                const float* __restrict__ w /* [144][132] */, const float* __restrict__ b,
                float* __restrict__ out /* [n,4,36] */) {
  __shared__ float sx[132];
  const int n = blockIdx.x;
  for (int i = threadIdx.x; i < 132; i += blockDim.x) {
    const int j = i / 6, k = i % 6;  // cat((axes, rest), dim=-1) per joint
    sx[i] = (k < 3) ? axes[((size_t)n * 22 + j) * 3 + k] : rest[((size_t)n * 22 + j) * 3 + k - 3];
  }
  __syncthreads();
  const int o = threadIdx.x;
  float a = b[o];
  for (int i = 0; i < 132; ++i) a = fmaf(sx[i], w[(size_t)o * 132 + i], a);
  out[(size_t)n * 144 + o] = fmaxf(a, 0.f);
}

// ------------------------------------------------------------------------------------------
// 4. write back the memory channels and assemble the regressor input
//    (temporal.py:87-91,134; umetrack_model.py:202-210): [72 features | 4 skeleton | zeros] = 80 ch
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
temporal_finish_kernel(FlatTensor tout /* 96 ch */, const int64_t* __restrict__ memory_idx,
                       float* __restrict__ mem_state, int n_slots, const float* __restrict__ skel,
                       int n_skel, FlatTensor rin /* 80 ch */) {
  __shared__ float t[96 * kPix];
  __shared__ float r[80 * kPix];
  const int b = blockIdx.x;
  const int slot = (int)memory_idx[b];
  load_sample_tile(tout, b, 12, t);
  __syncthreads();
  if (slot >= 0 && slot < n_slots) {
    float* mem = mem_state + (size_t)slot * 18 * kPix;
    for (int i = threadIdx.x; i < 18 * kPix; i += blockDim.x) mem[i] = t[i];
  }
  for (int i = threadIdx.x; i < 72 * kPix; i += blockDim.x) r[i] = t[18 * kPix + i];
  const float* sk = skel ? skel + (size_t)(n_skel == 1 ? 0 : b) * 4 * kPix : nullptr;
  for (int i = threadIdx.x; i < 8 * kPix; i += blockDim.x) r[72 * kPix + i] = (sk && i < 4 * kPix) ? sk[i] : 0.f;
  __syncthreads();
  store_sample_tile(rin, b, 10, r);
}

// ------------------------------------------------------------------------------------------
// 5. head: conv1x1(C -> n_out) + global average pool, decoders, Procrustes, world recovery
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
head_kernel(FlatTensor x /* 80 ch */, int C, const float* __restrict__ w /* [n_out][C] */,
            const float* __restrict__ bias, int n_out, int with_scale,
            const float* __restrict__ extrinsics, const int64_t* __restrict__ sample_range,
            const int64_t* __restrict__ hand_idx, float* __restrict__ out_angles,
            float* __restrict__ out_wrist, float* __restrict__ out_scale,
            float* __restrict__ out_sigma) {
  __shared__ float sMean[128];
  __shared__ float sRaw[64];
  const int b = blockIdx.x;
  // channel means over the 36 pixels: thread = (plane, e) reads its 16-byte rows
  if (threadIdx.x < 80) {
    const int pl = threadIdx.x >> 3, e = threadIdx.x & 7;
    float s = 0.f;
    for (int p = 0; p < kPix; ++p) {
      float v[8];
      load_row8<0>(x, pl, flat_row6(b, p), v);
      s += v[e];
    }
    sMean[threadIdx.x] = s * (1.0f / 36.0f);
  }
  __syncthreads();
  // conv1x1 and the average pool commute (both linear); pooling first saves 36x the work
  for (int o = threadIdx.x; o < n_out; o += blockDim.x) {
    float a = bias[o];
    for (int c = 0; c < C; ++c) a = fmaf(sMean[c], w[(size_t)o * C + c], a);
    sRaw[o] = a;
  }
  __syncthreads();
  head_decode(threadIdx.x, sRaw, b, with_scale, extrinsics, sample_range, hand_idx, out_angles, out_wrist, out_scale,
              out_sigma);
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
static const float* tp(const float* const* t, const std::string& name) {
  const int i = tensor_index(name.c_str());
  return i < 0 ? nullptr : t[i];
}

// Fold BN (or nothing) into conv `conv` ([cout][cin][ks][ks]) and zero-pad to [cout_p][cin_p].
static Folded folded_padded(const float* const* t, const std::string& conv, const std::string& bn,
                            int cout, int cin, int ks, bool has_bias, int cout_p, int cin_p) {
  const float* cw = tp(t, conv + ".weight");
  const float* cb = has_bias ? tp(t, conv + ".bias") : nullptr;
  Folded f;
  f.cout = cout_p; f.cin = cin_p; f.ks = ks;
  f.w.assign((size_t)cout_p * cin_p * ks * ks, 0.f);
  f.b.assign(cout_p, 0.f);
  for (int co = 0; co < cout; ++co) {
    float scale = 1.f, shift = 0.f;
    if (!bn.empty()) {
      scale = tp(t, bn + ".weight")[co] / std::sqrt(tp(t, bn + ".running_var")[co] + 1e-5f);
      shift = tp(t, bn + ".bias")[co] - tp(t, bn + ".running_mean")[co] * scale;
    }
    for (int ci = 0; ci < cin; ++ci)
      for (int k = 0; k < ks * ks; ++k)
        f.w[((size_t)co * cin_p + ci) * ks * ks + k] = cw[((size_t)co * cin + ci) * ks * ks + k] * scale;
    f.b[co] = (cb ? cb[co] * scale : 0.f) + shift;
  }
  return f;
}

static int add_tail_conv(umt_model* m, const Folded& f, bool relu, bool residual) {
  ConvLayer L;
  L.cout = f.cout; L.cin = f.cin; L.stride = 1; L.geom = 3; L.relu = relu ? 1 : 0;
  L.res_buf = residual ? 0 : -1;   // only "has a residual" matters here; tensors are bound per call
  UMT_TRY(build_conv(m, L, f, nullptr));
  m->tail_convs.push_back(L);
  return UMT_OK;
}

// [co][ci][k] -> [ci][k][co] on the device (tail_fused.cu reads consecutive output channels per lane)
static float* upload_ci_major(umt_model* m, const Folded& f) {
  const int kk = f.ks * f.ks;
  std::vector<float> w((size_t)f.cin * kk * f.cout);
  for (int co = 0; co < f.cout; ++co)
    for (int ci = 0; ci < f.cin; ++ci)
      for (int k = 0; k < kk; ++k) w[((size_t)ci * kk + k) * f.cout + co] = f.w[((size_t)co * f.cin + ci) * kk + k];
  return upload(m, w);
}

// The same weights packed per CTA of the cluster variant: slice r = [cin*kk][16] (zero beyond its channels), [16] biases.
static float* upload_cluster_slices(umt_model* m, const Folded& f) {
  const int kk = f.ks * f.ks, rows = f.cin * kk, per = (f.cout + kClusterCtas - 1) / kClusterCtas;
  const int n = cluster_slice_floats(rows);
  std::vector<float> w((size_t)kClusterCtas * n, 0.f);
  for (int r = 0; r < kClusterCtas; ++r)
    for (int j = 0; j < per; ++j) {
      const int co = r * per + j;
      if (co >= f.cout) break;
      float* dst = w.data() + (size_t)r * n;
      for (int ci = 0; ci < f.cin; ++ci)
        for (int k = 0; k < kk; ++k) dst[(size_t)(ci * kk + k) * kSliceStride + j] = f.w[((size_t)co * f.cin + ci) * kk + k];
      dst[(size_t)rows * kSliceStride + j] = f.b[co];
    }
  return upload(m, w);
}

int tail_build(umt_model* m, const float* const* t) {
  m->tail_convs.clear();
  {  // unpadded float32 copies for the fused small-batch kernel
    const std::string fu = "_feature_extractor._multi_view_fusion.", te = "_temporal._temporal_module.";
    const Folded f[3] = {folded_padded(t, fu + "0", fu + "1", 108, 144, 1, true, 108, 144),
                         folded_padded(t, fu + "3", fu + "4", 72, 108, 1, true, 72, 108),
                         folded_padded(t, fu + "6", "", 72, 72, 1, true, 72, 72)};
    for (int i = 0; i < 3; ++i) {
      m->tw.fus_w[i] = upload_ci_major(m, f[i]);
      m->tw.fus_b[i] = upload(m, f[i].b);
      m->tw.fus_cl[i] = upload_cluster_slices(m, f[i]);
      const Folded g = folded_padded(t, te + std::to_string(2 * i), "", 90, 90, 1, true, 90, 90);
      m->tw.tmp_w[i] = upload_ci_major(m, g);
      m->tw.tmp_b[i] = upload(m, g.b);
      m->tw.tmp_cl[i] = upload_cluster_slices(m, g);
      if (!m->tw.fus_w[i] || !m->tw.fus_b[i] || !m->tw.tmp_w[i] || !m->tw.tmp_b[i] || !m->tw.fus_cl[i] ||
          !m->tw.tmp_cl[i])
        return UMT_ERR_CUDA;
    }
    const char* heads[2] = {"_regressor_k", "_regressor_u"};
    const int hc[2] = {76, 72};
    for (int h = 0; h < 2; ++h)
      for (int c = 0; c < 4; ++c) {
        const std::string pre = std::string(heads[h]) + "._pose_regression_layers." + std::to_string(c / 2);
        const std::string cv = (c & 1) ? ".conv2" : ".conv1", bn = (c & 1) ? ".bn2" : ".bn1";
        const Folded g = folded_padded(t, pre + cv, pre + bn, hc[h], hc[h], 3, false, hc[h], hc[h]);
        m->tw.reg_w[h][c] = upload_ci_major(m, g);
        m->tw.reg_b[h][c] = upload(m, g.b);
        m->tw.reg_cl[h][c] = upload_cluster_slices(m, g);
        if (!m->tw.reg_w[h][c] || !m->tw.reg_b[h][c] || !m->tw.reg_cl[h][c]) return UMT_ERR_CUDA;
      }
  }
  const std::string fu = "_feature_extractor._multi_view_fusion.";
  // fusion: 144 -> 108 (pad 112) -> 72 (pad 80) -> 72 (pad 80)   (model_utils.py:141-163)
  UMT_TRY(add_tail_conv(m, folded_padded(t, fu + "0", fu + "1", 108, 144, 1, true, 112, 144), true, false));
  UMT_TRY(add_tail_conv(m, folded_padded(t, fu + "3", fu + "4", 72, 108, 1, true, 80, 112), true, false));
  UMT_TRY(add_tail_conv(m, folded_padded(t, fu + "6", "", 72, 72, 1, true, 80, 80), false, false));
  // temporal: three 90 -> 90 (pad 96), ReLU after the first two (temporal.py:31-38)
  const std::string te = "_temporal._temporal_module.";
  UMT_TRY(add_tail_conv(m, folded_padded(t, te + "0", "", 90, 90, 1, true, 96, 96), true, false));
  UMT_TRY(add_tail_conv(m, folded_padded(t, te + "2", "", 90, 90, 1, true, 96, 96), true, false));
  UMT_TRY(add_tail_conv(m, folded_padded(t, te + "4", "", 90, 90, 1, true, 96, 96), false, false));
  // regressors: two BasicBlocks each (model_utils.py:195-208), C = 76 / 72 padded to 80
  const char* heads[2] = {"_regressor_k", "_regressor_u"};
  const int hc[2] = {76, 72}, hout[2] = {62, 63};
  std::vector<float> w, b;
  for (int h = 0; h < 2; ++h) {
    const std::string p = std::string(heads[h]) + "._pose_regression_layers.";
    for (int blk = 0; blk < 2; ++blk) {
      const std::string pre = p + std::to_string(blk);
      UMT_TRY(add_tail_conv(m, folded_padded(t, pre + ".conv1", pre + ".bn1", hc[h], hc[h], 3, false, 80, 80), true, false));
      UMT_TRY(add_tail_conv(m, folded_padded(t, pre + ".conv2", pre + ".bn2", hc[h], hc[h], 3, false, 80, 80), true, true));
    }
    // 1x1 head conv applied after the average pool: [n_out][80] with zero columns for the padding
    const float* cw = tp(t, p + "2.weight");
    const float* cb = tp(t, p + "2.bias");
    w.assign((size_t)hout[h] * 80, 0.f);
    b.assign(hout[h], 0.f);
    for (int o = 0; o < hout[h]; ++o) {
      for (int c = 0; c < hc[h]; ++c) w[(size_t)o * 80 + c] = cw[(size_t)o * hc[h] + c];
      b[o] = cb[o];
    }
    m->tw.head_w[h] = upload(m, w);
    m->tw.head_b[h] = upload(m, b);
  }
  {  // Linear(132->144) then BN2d(4) over view(4,6,6): channel of output o is o / 36
    const float* lw = tp(t, "_skeleton_enc._layers.0.weight");
    const float* lb = tp(t, "_skeleton_enc._layers.0.bias");
    const std::string bn = "_skeleton_enc._layers.2";
    w.resize(144 * 132);
    b.resize(144);
    for (int o = 0; o < 144; ++o) {
      const int c = o / 36;
      const float scale = tp(t, bn + ".weight")[c] / std::sqrt(tp(t, bn + ".running_var")[c] + 1e-5f);
      const float shift = tp(t, bn + ".bias")[c] - tp(t, bn + ".running_mean")[c] * scale;
      for (int i = 0; i < 132; ++i) w[(size_t)o * 132 + i] = lw[(size_t)o * 132 + i] * scale;
      b[o] = lb[o] * scale + shift;
    }
    m->tw.skel_w = upload(m, w);
    m->tw.skel_b = upload(m, b);
  }
  if (!m->tw.skel_w || !m->tw.skel_b || !m->tw.head_w[0] || !m->tw.head_w[1]) return UMT_ERR_CUDA;
  return UMT_OK;
}

// run tail convolution `idx` from `src` to `dst` (residual `res` if the layer has one)
static int run_tail_conv(umt_model* m, int idx, const FlatTensor& src, const FlatTensor& dst,
                         const FlatTensor* res, int n_samples, cudaStream_t st) {
  const ConvLayer& L = m->tail_convs[idx];
  ConvParams p = L.prog;
  p.src[0] = src;
  if (res) { p.src[2] = *res; p.res = *res; }
  p.out[0] = dst;
  p.weights = L.d_weights;
  p.bias = L.d_bias;
  p.n_pos = n_samples * kPos;
  finalize_conv_params(p, L, UMT_MODE_FP32_EQUIV);
  UMT_TRY(launch_conv(p, L.cout, UMT_MODE_FP32_EQUIV, st));
  ++m->launches;
  return UMT_OK;
}

int tail_forward(umt_model* m, int head, const float* feats, const float* intrinsics,
                 const float* extrinsics, const int64_t* sample_range, const int64_t* memory_idx,
                 const uint8_t* use_memory, const int64_t* hand_idx, int n_img, int n_samples,
                 const float* skel_axes, const float* skel_rest, int n_skel, float* mem_state,
                 float* prev_extr, int n_slots, float* out_joint_angles, float* out_wrist,
                 float* out_scale, float* out_sigma, cudaStream_t st) {
  (void)n_img;
  const int B = n_samples;
  if (B <= tail_fused_max_samples()) {   // small batches (per-frame tracker, time scan): one CTA per sample
    const float* skel = nullptr;
    if (head == UMT_HEAD_KNOWN_SKELETON) {
      skeleton_kernel<<<n_skel, 144, 0, st>>>(skel_axes, skel_rest, m->tw.skel_w, m->tw.skel_b, m->tail_skel);
      UMT_TRY(check_launch("skeleton_kernel"));
      ++m->launches;
      skel = m->tail_skel;
    }
    return tail_fused_forward(m, head, feats, intrinsics, extrinsics, sample_range, memory_idx, use_memory, hand_idx, B,
                              skel, n_skel, mem_state, prev_extr, n_slots, out_joint_angles, out_wrist, out_scale,
                              out_sigma, st);
  }
  const FlatTensor &T0 = m->tail_flat[0], &T1 = m->tail_flat[1], &T2 = m->tail_flat[2];
  const TailWeights& tw = m->tw;
  fuse_prepare_kernel<<<B, 128, 0, st>>>(feats, intrinsics, extrinsics, sample_range, T0);
  UMT_TRY(check_launch("fuse_prepare_kernel"));
  UMT_TRY(run_tail_conv(m, 0, T0, T1, nullptr, B, st));
  UMT_TRY(run_tail_conv(m, 1, T1, T2, nullptr, B, st));
  UMT_TRY(run_tail_conv(m, 2, T2, T0, nullptr, B, st));
  temporal_prepare_kernel<<<B, 128, 0, st>>>(T0, feats, intrinsics, extrinsics, sample_range, memory_idx,
                                             use_memory, mem_state, prev_extr, n_slots, T1, m->d_status);
  UMT_TRY(check_launch("temporal_prepare_kernel"));
  UMT_TRY(run_tail_conv(m, 3, T1, T2, nullptr, B, st));
  UMT_TRY(run_tail_conv(m, 4, T2, T0, nullptr, B, st));
  UMT_TRY(run_tail_conv(m, 5, T0, T1, nullptr, B, st));
  m->launches += 2;
  const float* skel = nullptr;
  if (head == UMT_HEAD_KNOWN_SKELETON) {
    skeleton_kernel<<<n_skel, 144, 0, st>>>(skel_axes, skel_rest, tw.skel_w, tw.skel_b, m->tail_skel);
    UMT_TRY(check_launch("skeleton_kernel"));
    ++m->launches;
    skel = m->tail_skel;
  }
  temporal_finish_kernel<<<B, 128, 0, st>>>(T1, memory_idx, mem_state, n_slots, skel, n_skel, T2);
  UMT_TRY(check_launch("temporal_finish_kernel"));
  // two BasicBlocks (backbone_resnet.py:56-72): x=T2 -> T0 -> T1(+T2) -> T0 -> T2(+T1)
  const int c0 = 6 + 4 * head;
  UMT_TRY(run_tail_conv(m, c0 + 0, T2, T0, nullptr, B, st));
  UMT_TRY(run_tail_conv(m, c0 + 1, T0, T1, &T2, B, st));
  UMT_TRY(run_tail_conv(m, c0 + 2, T1, T0, nullptr, B, st));
  UMT_TRY(run_tail_conv(m, c0 + 3, T0, T2, &T1, B, st));
  head_kernel<<<B, 128, 0, st>>>(T2, 80, tw.head_w[head], tw.head_b[head], head == UMT_HEAD_KNOWN_SKELETON ? 62 : 63,
                                 head == UMT_HEAD_UNKNOWN_SKELETON ? 1 : 0, extrinsics, sample_range,
                                 hand_idx, out_joint_angles, out_wrist, out_scale, out_sigma);
  UMT_TRY(check_launch("head_kernel"));
  m->launches += 2;
  return UMT_OK;
}

}  // namespace umt
