// Stage (b), part 2: everything after the per-view encoder, in float32 on CUDA cores (1 % of the
// FLOPs, and the part whose values must stay float32: FTL transforms, memory bank, Procrustes).
//   feature-transform layer + two-view fusion   lib/models/feature_extractor.py:61-141,
//                                                lib/models/model_utils.py:57-104,141-192
//   temporal memory (SimpleConvRNN)              lib/models/temporal.py:51-139
//   skeleton encoder                             lib/models/skeleton_encoder.py:28-53
//   pose regressor + decoders + Procrustes       lib/models/regressor.py:76-186,
//                                                lib/models/model_utils.py:17-54
//   world-space recovery / right-hand un-mirror  lib/models/umetrack_model.py:77-97
// All feature maps are dense float32 [batch, C, 36] (NCHW with 6x6 flattened), which is the
// layout the reference's FTL reshape (C channels -> 3 x C/3*36 points) is defined on.
//
// The convolutions use one thread per (sample, output channel) holding all 36 output pixels in
// registers: per input channel a thread reads its weights once and the input map with 16-byte
// shared-memory broadcasts, so the inner loop is FMA bound.
#include <cmath>
#include <string>

#include "model.cuh"

namespace umt {

constexpr int kPix = 36;
constexpr int kSamplesPerBlock = 4;

// ------------------------------------------------------------------------------------------
// small matrix helpers (double precision, per thread)
// ------------------------------------------------------------------------------------------
__device__ void mat4_mul(const double* a, const double* b, double* c) {
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      double s = 0;
      for (int k = 0; k < 4; ++k) s += a[i * 4 + k] * b[k * 4 + j];
      c[i * 4 + j] = s;
    }
}

// general 4x4 inverse by Gauss-Jordan with partial pivoting; returns false if singular
__device__ bool mat4_inv(const double* a, double* inv) {
  double m[4][8];
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      m[i][j] = a[i * 4 + j];
      m[i][4 + j] = (i == j) ? 1.0 : 0.0;
    }
  for (int c = 0; c < 4; ++c) {
    int piv = c;
    double best = fabs(m[c][c]);
    for (int r = c + 1; r < 4; ++r)
      if (fabs(m[r][c]) > best) {
        best = fabs(m[r][c]);
        piv = r;
      }
    if (best < 1e-300) return false;
    if (piv != c)
      for (int j = 0; j < 8; ++j) {
        double t = m[c][j];
        m[c][j] = m[piv][j];
        m[piv][j] = t;
      }
    const double d = 1.0 / m[c][c];
    for (int j = 0; j < 8; ++j) m[c][j] *= d;
    for (int r = 0; r < 4; ++r) {
      if (r == c) continue;
      const double f = m[r][c];
      if (f != 0.0)
        for (int j = 0; j < 8; ++j) m[r][j] -= f * m[c][j];
    }
  }
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) inv[i * 4 + j] = m[i][4 + j];
  return true;
}

__device__ void load_mat4(const float* p, double* m) {
  for (int i = 0; i < 16; ++i) m[i] = (double)p[i];
}

// FTL on a [3*n] point cloud stored as three consecutive blocks X|Y|Z of n floats
// (lib/models/model_utils.py:83-95): out = R * p + t, with xf the top 3x4 of a 4x4 in float32.
__device__ __forceinline__ void ftl_point(const float* xf, float x, float y, float z, float& ox,
                                          float& oy, float& oz) {
  ox = xf[0] * x + xf[1] * y + xf[2] * z + xf[3];
  oy = xf[4] * x + xf[5] * y + xf[6] * z + xf[7];
  oz = xf[8] * x + xf[9] * y + xf[10] * z + xf[11];
}

// ------------------------------------------------------------------------------------------
// 1. views -> canonical space, channel-concatenated for the fusion convs
//    (feature_extractor.py:61-87,121-133).  One CTA per sample.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
fuse_prepare_kernel(const float* __restrict__ feats, const float* __restrict__ intrinsics,
                    const float* __restrict__ extrinsics, const int64_t* __restrict__ sample_range,
                    float* __restrict__ canon /* [B,144,36] */) {
  __shared__ float sxf[2][12];
  const int b = blockIdx.x;
  const int r0 = (int)sample_range[2 * b], r1 = (int)sample_range[2 * b + 1];
  float* dst = canon + (size_t)b * 144 * kPix;
  if (r1 - r0 != 2) {  // single-view samples skip the fusion convs (feature_extractor.py:89-94)
    for (int i = threadIdx.x; i < 144 * kPix; i += blockDim.x) dst[i] = 0.f;
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
    float* d = dst + (size_t)v * 72 * kPix;
    d[j] = ox;
    d[864 + j] = oy;
    d[1728 + j] = oz;
  }
}

// ------------------------------------------------------------------------------------------
// 2. convolutions on 6x6 maps.  Thread = (sample, output channel), 36 accumulators.
// ------------------------------------------------------------------------------------------
template <int KS>
__global__ void __launch_bounds__(kSamplesPerBlock * 112)
conv6_kernel(const float* __restrict__ in, int in_stride /* floats per sample */, int cin,
             const float* __restrict__ w /* [cout][cin][KS*KS] */, const float* __restrict__ bias,
             const float* __restrict__ res, int res_stride, float* __restrict__ out, int out_stride,
             int cout, int relu, int n_samples) {
  // KS==1: sIn[sample][ci][36]; KS==3: sIn[sample][ci][64] zero-padded 8x8 tiles
  extern __shared__ __align__(16) float sIn[];
  constexpr int kTile = (KS == 1) ? 36 : 64;
  const int s0 = blockIdx.x * kSamplesPerBlock;
  const int ns = min(kSamplesPerBlock, n_samples - s0);
  const int per = cin * kTile;
  for (int i = threadIdx.x; i < kSamplesPerBlock * per; i += blockDim.x) {
    const int s = i / per, r = i % per, ci = r / kTile, p = r % kTile;
    float v = 0.f;
    if (s < ns) {
      if (KS == 1) {
        v = in[(size_t)(s0 + s) * in_stride + ci * kPix + p];
      } else {
        const int y = (p >> 3) - 1, x = (p & 7) - 1;
        if (y >= 0 && y < 6 && x >= 0 && x < 6) v = in[(size_t)(s0 + s) * in_stride + ci * kPix + y * 6 + x];
      }
    }
    sIn[i] = v;
  }
  __syncthreads();
  const int s = threadIdx.x / 112, co = threadIdx.x % 112;
  if (s >= ns || co >= cout) return;
  float acc[36];
  const float b0 = bias ? bias[co] : 0.f;
#pragma unroll
  for (int p = 0; p < 36; ++p) acc[p] = b0;
  const float* xs = sIn + (size_t)s * per;
  const float* wr = w + (size_t)co * cin * KS * KS;
  for (int ci = 0; ci < cin; ++ci) {
    if (KS == 1) {
      const float ww = __ldg(wr + ci);
      const float4* x4 = reinterpret_cast<const float4*>(xs + ci * 36);
#pragma unroll
      for (int q = 0; q < 9; ++q) {
        const float4 x = x4[q];
        acc[4 * q + 0] = fmaf(x.x, ww, acc[4 * q + 0]);
        acc[4 * q + 1] = fmaf(x.y, ww, acc[4 * q + 1]);
        acc[4 * q + 2] = fmaf(x.z, ww, acc[4 * q + 2]);
        acc[4 * q + 3] = fmaf(x.w, ww, acc[4 * q + 3]);
      }
    } else {
      float wk[9];
#pragma unroll
      for (int k = 0; k < 9; ++k) wk[k] = __ldg(wr + ci * 9 + k);
      float t[64];
      const float4* x4 = reinterpret_cast<const float4*>(xs + ci * 64);
#pragma unroll
      for (int q = 0; q < 16; ++q) {
        const float4 x = x4[q];
        t[4 * q] = x.x; t[4 * q + 1] = x.y; t[4 * q + 2] = x.z; t[4 * q + 3] = x.w;
      }
#pragma unroll
      for (int y = 0; y < 6; ++y)
#pragma unroll
        for (int x = 0; x < 6; ++x) {
          float a = acc[y * 6 + x];
#pragma unroll
          for (int ky = 0; ky < 3; ++ky)
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) a = fmaf(t[(y + ky) * 8 + x + kx], wk[ky * 3 + kx], a);
          acc[y * 6 + x] = a;
        }
    }
  }
  float* o = out + (size_t)(s0 + s) * out_stride + (size_t)co * kPix;
  const float* r = res ? res + (size_t)(s0 + s) * res_stride + (size_t)co * kPix : nullptr;
#pragma unroll
  for (int p = 0; p < 36; ++p) {
    float v = acc[p];
    if (r) v += r[p];
    if (relu) v = fmaxf(v, 0.f);
    o[p] = v;
  }
}

template <int KS>
static int launch_conv6(const float* in, int in_stride, int cin, const float* w, const float* b,
                        const float* res, int res_stride, float* out, int out_stride, int cout,
                        int relu, int n, cudaStream_t st) {
  if (cout > 112) return fail(UMT_ERR_UNSUPPORTED, "conv6: cout %d > 112", cout);
  const size_t smem = (size_t)kSamplesPerBlock * cin * ((KS == 1) ? 36 : 64) * sizeof(float);
  UMT_CUDA_TRY(cudaFuncSetAttribute(conv6_kernel<KS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int blocks = (n + kSamplesPerBlock - 1) / kSamplesPerBlock;
  conv6_kernel<KS><<<blocks, kSamplesPerBlock * 112, smem, st>>>(in, in_stride, cin, w, b, res, res_stride,
                                                                 out, out_stride, cout, relu, n);
  return check_launch("conv6_kernel");
}

// ------------------------------------------------------------------------------------------
// 3. canonical -> cam0 (feature_extractor.py:135-139) or single-view FTL (:89-94), memory warp
//    and bookkeeping of SimpleConvRNN.transform_memory_features (temporal.py:51-78), and the
//    channel concat [memory(18) | image features(72)] (temporal.py:83-85).  One CTA per sample.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
temporal_prepare_kernel(const float* __restrict__ fused_raw /* [B,72,36] stride 144*36 */,
                        const float* __restrict__ feats, const float* __restrict__ intrinsics,
                        const float* __restrict__ extrinsics, const int64_t* __restrict__ sample_range,
                        const int64_t* __restrict__ memory_idx, const uint8_t* __restrict__ use_memory,
                        float* __restrict__ mem_state, float* __restrict__ prev_extr, int n_slots,
                        float* __restrict__ tin /* [B,90,36] stride 144*36 */) {
  __shared__ float sS[12];
  __shared__ float sM[12];
  __shared__ int sUse;
  const int b = blockIdx.x;
  const int r0 = (int)sample_range[2 * b], r1 = (int)sample_range[2 * b + 1];
  const int slot = (int)memory_idx[b];
  if (slot < 0 || slot >= n_slots) return;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 12; ++i) sS[i] = 0.f;
    sS[0] = sS[5] = 1.f;
    sS[10] = intrinsics[(size_t)r0 * 9] / 200.0f;
    int use = use_memory[b] ? 1 : 0;
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
  __syncthreads();
  float* out = tin + (size_t)b * 144 * kPix;
  float* mem = mem_state + (size_t)slot * 18 * kPix;
  // memory part: 18 channels = 216 points
  for (int j = threadIdx.x; j < 216; j += blockDim.x) {
    float ox = 0.f, oy = 0.f, oz = 0.f;
    if (sUse) ftl_point(sM, mem[j], mem[216 + j], mem[432 + j], ox, oy, oz);
    out[j] = ox;
    out[216 + j] = oy;
    out[432 + j] = oz;
  }
  // image-feature part: 72 channels = 864 points, transformed by S of view 0
  const bool multi = (r1 - r0) == 2;
  const float* src = multi ? fused_raw + (size_t)b * 144 * kPix : feats + (size_t)r0 * 72 * kPix;
  float* o2 = out + 18 * kPix;
  for (int j = threadIdx.x; j < 864; j += blockDim.x) {
    float ox, oy, oz;
    ftl_point(sS, src[j], src[864 + j], src[1728 + j], ox, oy, oz);
    o2[j] = ox;
    o2[864 + j] = oy;
    o2[1728 + j] = oz;
  }
  // prev_extrinsics[memory_idx] = cur_extrinsics (temporal.py:77)
  if (threadIdx.x < 16) prev_extr[(size_t)slot * 16 + threadIdx.x] = extrinsics[(size_t)r0 * 16 + threadIdx.x];
}

// ------------------------------------------------------------------------------------------
// 4. skeleton encoder: Linear(132->144) -> view(4,6,6) -> BN(4) -> ReLU, BN folded on the host.
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
// 5. write back the memory channels and assemble the regressor input
//    (temporal.py:87-91,134; umetrack_model.py:202-210)
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
temporal_finish_kernel(const float* __restrict__ tout /* [B,90,36] stride 144*36 */,
                       const int64_t* __restrict__ memory_idx, float* __restrict__ mem_state,
                       int n_slots, const float* __restrict__ skel, int n_skel,
                       float* __restrict__ rin /* [B,76|72,36] stride 144*36 */) {
  const int b = blockIdx.x;
  const int slot = (int)memory_idx[b];
  const float* t = tout + (size_t)b * 144 * kPix;
  if (slot >= 0 && slot < n_slots) {
    float* mem = mem_state + (size_t)slot * 18 * kPix;
    for (int i = threadIdx.x; i < 18 * kPix; i += blockDim.x) mem[i] = t[i];
  }
  float* r = rin + (size_t)b * 144 * kPix;
  for (int i = threadIdx.x; i < 72 * kPix; i += blockDim.x) r[i] = t[18 * kPix + i];
  if (skel) {
    const float* sk = skel + (size_t)(n_skel == 1 ? 0 : b) * 4 * kPix;
    for (int i = threadIdx.x; i < 4 * kPix; i += blockDim.x) r[72 * kPix + i] = sk[i];
  }
}

// ------------------------------------------------------------------------------------------
// 6. head: conv1x1(C -> n_out) + global average pool, decoders, Procrustes, world recovery
// ------------------------------------------------------------------------------------------
__device__ void jacobi_svd3(const double* A, double* U, double* S, double* V) {
  // one-sided Jacobi: rotate columns of W = A*V until orthogonal; then W = U*diag(S)
  double W[9];
  for (int i = 0; i < 9; ++i) {
    W[i] = A[i];
    V[i] = (i % 4 == 0) ? 1.0 : 0.0;
  }
  for (int sweep = 0; sweep < 30; ++sweep) {
    double off = 0.0;
    for (int p = 0; p < 2; ++p)
      for (int q = p + 1; q < 3; ++q) {
        double a = 0, b = 0, c = 0;
        for (int i = 0; i < 3; ++i) {
          a += W[i * 3 + p] * W[i * 3 + p];
          b += W[i * 3 + q] * W[i * 3 + q];
          c += W[i * 3 + p] * W[i * 3 + q];
        }
        off = fmax(off, fabs(c) / (sqrt(a * b) + 1e-300));
        if (fabs(c) < 1e-300) continue;
        const double zeta = (b - a) / (2.0 * c);
        const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
        const double cs = 1.0 / sqrt(1.0 + t * t), sn = cs * t;
        for (int i = 0; i < 3; ++i) {
          const double wp = W[i * 3 + p], wq = W[i * 3 + q];
          W[i * 3 + p] = cs * wp - sn * wq;
          W[i * 3 + q] = sn * wp + cs * wq;
          const double vp = V[i * 3 + p], vq = V[i * 3 + q];
          V[i * 3 + p] = cs * vp - sn * vq;
          V[i * 3 + q] = sn * vp + cs * vq;
        }
      }
    if (off < 1e-15) break;
  }
  for (int j = 0; j < 3; ++j) {
    S[j] = sqrt(W[j] * W[j] + W[3 + j] * W[3 + j] + W[6 + j] * W[6 + j]);
  }
  // sort singular values descending (torch.svd order) so the reflection fix hits the smallest
  int order[3] = {0, 1, 2};
  for (int i = 0; i < 2; ++i)
    for (int j = i + 1; j < 3; ++j)
      if (S[order[j]] > S[order[i]]) {
        int t = order[i];
        order[i] = order[j];
        order[j] = t;
      }
  double Ws[9], Vs[9], Ss[3];
  for (int j = 0; j < 3; ++j) {
    Ss[j] = S[order[j]];
    for (int i = 0; i < 3; ++i) {
      Ws[i * 3 + j] = W[i * 3 + order[j]];
      Vs[i * 3 + j] = V[i * 3 + order[j]];
    }
  }
  for (int j = 0; j < 3; ++j) S[j] = Ss[j];
  for (int i = 0; i < 9; ++i) V[i] = Vs[i];
  // U columns; a vanishing third singular value gets u2 = u0 x u1
  for (int j = 0; j < 3; ++j) {
    if (Ss[j] > 1e-12 * (Ss[0] + 1e-300)) {
      for (int i = 0; i < 3; ++i) U[i * 3 + j] = Ws[i * 3 + j] / Ss[j];
    } else if (j == 2) {
      U[2] = U[3] * U[7] - U[6] * U[4];
      U[5] = U[6] * U[1] - U[0] * U[7];
      U[8] = U[0] * U[4] - U[3] * U[1];
    } else {
      for (int i = 0; i < 3; ++i) U[i * 3 + j] = (i == j) ? 1.0 : 0.0;
    }
  }
}

__device__ double det3(const double* m) {
  return m[0] * (m[4] * m[8] - m[5] * m[7]) - m[1] * (m[3] * m[8] - m[5] * m[6]) +
         m[2] * (m[3] * m[7] - m[4] * m[6]);
}

__global__ void __launch_bounds__(128)
head_kernel(const float* __restrict__ x /* [B,C,36] stride 144*36 */, int C,
            const float* __restrict__ w /* [n_out][C] */, const float* __restrict__ bias, int n_out,
            int with_scale, const float* __restrict__ extrinsics,
            const int64_t* __restrict__ sample_range, const int64_t* __restrict__ hand_idx,
            float* __restrict__ out_angles, float* __restrict__ out_wrist,
            float* __restrict__ out_scale, float* __restrict__ out_sigma) {
  __shared__ float sMean[128];
  __shared__ float sRaw[64];
  const int b = blockIdx.x;
  const float* xb = x + (size_t)b * 144 * kPix;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float s = 0.f;
    for (int p = 0; p < kPix; ++p) s += xb[c * kPix + p];
    sMean[c] = s * (1.0f / 36.0f);
  }
  __syncthreads();
  // conv1x1 and the average pool commute (both linear); pooling first saves 36x the work
  for (int o = threadIdx.x; o < n_out; o += blockDim.x) {
    float a = bias[o];
    for (int c = 0; c < C; ++c) a = fmaf(sMean[c], w[(size_t)o * C + c], a);
    sRaw[o] = a;
  }
  __syncthreads();
  const int t = threadIdx.x;
  if (t < 22) out_angles[(size_t)b * 22 + t] = (t < 20) ? sRaw[t] : 0.f;  // regressor.py:76-86
  const int sig0 = with_scale ? 42 : 41;
  if (t < 21) {  // clamp(softplus(x), 1e-5)  (regressor.py:117-121); torch threshold = 20
    const float v = sRaw[sig0 + t];
    const float sp = (v > 20.f) ? v : log1pf(expf(v));
    out_sigma[(size_t)b * 21 + t] = fmaxf(sp, 1e-5f);
  }
  if (t == 32 && with_scale && out_scale) out_scale[b] = expf(sRaw[41]);  // regressor.py:106-114
  if (t == 64) {
    // Procrustes (model_utils.py:17-54) of the 7 template points (regressor.py:19-47) onto the
    // predicted ones, then inv(E0) * X and the right-hand flip (umetrack_model.py:77-90)
    const double k = 0.1, d = 0.1 / sqrt(2.0);
    const double P[21] = {0, 0, 0, k, 0, 0, 0, k, 0, 0, 0, k, -d, -d, 0, -d, 0, -d, 0, -d, -d};
    float Pf[21], Q[21];
    for (int i = 0; i < 21; ++i) {
      Pf[i] = (float)P[i];  // the reference keeps the template in float32
      Q[i] = sRaw[20 + i];
    }
    double mp[3] = {0, 0, 0}, mq[3] = {0, 0, 0};
    for (int i = 0; i < 7; ++i)
      for (int c = 0; c < 3; ++c) {
        mp[c] += Pf[i * 3 + c];
        mq[c] += Q[i * 3 + c];
      }
    for (int c = 0; c < 3; ++c) {
      mp[c] /= 7.0;
      mq[c] /= 7.0;
    }
    double Hm[9] = {0};
    for (int i = 0; i < 7; ++i)
      for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) Hm[r * 3 + c] += (Pf[i * 3 + r] - mp[r]) * (Q[i * 3 + c] - mq[c]);
    double U[9], S[3], V[9];
    jacobi_svd3(Hm, U, S, V);
    double VUt[9];
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) {
        double s = 0;
        for (int j = 0; j < 3; ++j) s += V[r * 3 + j] * U[c * 3 + j];
        VUt[r * 3 + c] = s;
      }
    const double dsign = det3(VUt);
    double R[9];
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) {
        double s = 0;
        for (int j = 0; j < 3; ++j) s += V[r * 3 + j] * (j == 2 ? dsign : 1.0) * U[c * 3 + j];
        R[r * 3 + c] = s;
      }
    double X[16] = {0};
    for (int r = 0; r < 3; ++r) {
      for (int c = 0; c < 3; ++c) X[r * 4 + c] = R[r * 3 + c];
      X[r * 4 + 3] = mq[r] - (R[r * 3] * mp[0] + R[r * 3 + 1] * mp[1] + R[r * 3 + 2] * mp[2]);
    }
    X[15] = 1.0;
    double E0[16], Einv[16], Wd[16];
    const int r0 = (int)sample_range[2 * b];
    load_mat4(extrinsics + (size_t)r0 * 16, E0);
    const bool ok = mat4_inv(E0, Einv);
    mat4_mul(Einv, X, Wd);
    const double flip = (hand_idx[b] == 1) ? -1.0 : 1.0;
    for (int r = 0; r < 4; ++r)
      for (int c = 0; c < 4; ++c) {
        double v = Wd[r * 4 + c] * (c == 0 ? flip : 1.0);
        out_wrist[(size_t)b * 16 + r * 4 + c] = ok ? (float)v : nanf("");
      }
  }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
static const float* tp(const float* const* t, const std::string& name) {
  const int i = tensor_index(name.c_str());
  return i < 0 ? nullptr : t[i];
}

// fold BN (or nothing) into a [cout][cin*ks*ks] weight and bias
static void fold(const float* const* t, const std::string& conv, const std::string& bn, int cout,
                 int per, bool has_bias, std::vector<float>& w, std::vector<float>& b) {
  const float* cw = tp(t, conv + ".weight");
  const float* cb = has_bias ? tp(t, conv + ".bias") : nullptr;
  w.resize((size_t)cout * per);
  b.resize(cout);
  for (int co = 0; co < cout; ++co) {
    float scale = 1.f, shift = 0.f;
    if (!bn.empty()) {
      scale = tp(t, bn + ".weight")[co] / std::sqrt(tp(t, bn + ".running_var")[co] + 1e-5f);
      shift = tp(t, bn + ".bias")[co] - tp(t, bn + ".running_mean")[co] * scale;
    }
    for (int i = 0; i < per; ++i) w[(size_t)co * per + i] = cw[(size_t)co * per + i] * scale;
    b[co] = (cb ? cb[co] * scale : 0.f) + shift;
  }
}

int tail_build(umt_model* m, const float* const* t) {
  std::vector<float> w, b;
  const std::string fu = "_feature_extractor._multi_view_fusion.";
  const int fcin[3] = {144, 108, 72}, fcout[3] = {108, 72, 72};
  const char* fconv[3] = {"0", "3", "6"};
  const char* fbn[3] = {"1", "4", ""};
  for (int i = 0; i < 3; ++i) {
    fold(t, fu + fconv[i], fbn[i][0] ? fu + fbn[i] : std::string(), fcout[i], fcin[i], true, w, b);
    m->tw.fus_w[i] = upload(m, w);
    m->tw.fus_b[i] = upload(m, b);
  }
  const std::string te = "_temporal._temporal_module.";
  const char* tconv[3] = {"0", "2", "4"};
  for (int i = 0; i < 3; ++i) {
    fold(t, te + tconv[i], "", 90, 90, true, w, b);
    m->tw.tmp_w[i] = upload(m, w);
    m->tw.tmp_b[i] = upload(m, b);
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
  const char* heads[2] = {"_regressor_k", "_regressor_u"};
  const int hc[2] = {76, 72}, hout[2] = {62, 63};
  for (int h = 0; h < 2; ++h) {
    const std::string p = std::string(heads[h]) + "._pose_regression_layers.";
    for (int blk = 0; blk < 2; ++blk)
      for (int c = 0; c < 2; ++c) {
        const std::string pre = p + std::to_string(blk);
        fold(t, pre + (c ? ".conv2" : ".conv1"), pre + (c ? ".bn2" : ".bn1"), hc[h], hc[h] * 9, false, w, b);
        m->tw.reg_w[h][blk * 2 + c] = upload(m, w);
        m->tw.reg_b[h][blk * 2 + c] = upload(m, b);
      }
    fold(t, p + "2", "", hout[h], hc[h], true, w, b);
    m->tw.head_w[h] = upload(m, w);
    m->tw.head_b[h] = upload(m, b);
  }
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
  const int stride = 144 * kPix;
  float *A = m->tail_a, *Bf = m->tail_b, *Cf = m->tail_c;
  const TailWeights& tw = m->tw;
  fuse_prepare_kernel<<<B, 128, 0, st>>>(feats, intrinsics, extrinsics, sample_range, A);
  UMT_TRY(check_launch("fuse_prepare_kernel"));
  UMT_TRY(launch_conv6<1>(A, stride, 144, tw.fus_w[0], tw.fus_b[0], nullptr, 0, Bf, stride, 108, 1, B, st));
  UMT_TRY(launch_conv6<1>(Bf, stride, 108, tw.fus_w[1], tw.fus_b[1], nullptr, 0, Cf, stride, 72, 1, B, st));
  UMT_TRY(launch_conv6<1>(Cf, stride, 72, tw.fus_w[2], tw.fus_b[2], nullptr, 0, A, stride, 72, 0, B, st));
  temporal_prepare_kernel<<<B, 128, 0, st>>>(A, feats, intrinsics, extrinsics, sample_range, memory_idx,
                                             use_memory, mem_state, prev_extr, n_slots, Bf);
  UMT_TRY(check_launch("temporal_prepare_kernel"));
  UMT_TRY(launch_conv6<1>(Bf, stride, 90, tw.tmp_w[0], tw.tmp_b[0], nullptr, 0, Cf, stride, 90, 1, B, st));
  UMT_TRY(launch_conv6<1>(Cf, stride, 90, tw.tmp_w[1], tw.tmp_b[1], nullptr, 0, A, stride, 90, 1, B, st));
  UMT_TRY(launch_conv6<1>(A, stride, 90, tw.tmp_w[2], tw.tmp_b[2], nullptr, 0, Bf, stride, 90, 0, B, st));
  m->launches += 8;
  const float* skel = nullptr;
  int C = 72;
  if (head == UMT_HEAD_KNOWN_SKELETON) {
    skeleton_kernel<<<n_skel, 144, 0, st>>>(skel_axes, skel_rest, tw.skel_w, tw.skel_b, m->tail_skel);
    UMT_TRY(check_launch("skeleton_kernel"));
    ++m->launches;
    skel = m->tail_skel;
    C = 76;
  }
  temporal_finish_kernel<<<B, 128, 0, st>>>(Bf, memory_idx, mem_state, n_slots, skel, n_skel, Cf);
  UMT_TRY(check_launch("temporal_finish_kernel"));
  // two BasicBlocks (backbone_resnet.py:56-72): x=Cf -> A -> Bf(+Cf) -> A -> Cf(+Bf)
  const int h = head;
  UMT_TRY(launch_conv6<3>(Cf, stride, C, tw.reg_w[h][0], tw.reg_b[h][0], nullptr, 0, A, stride, C, 1, B, st));
  UMT_TRY(launch_conv6<3>(A, stride, C, tw.reg_w[h][1], tw.reg_b[h][1], Cf, stride, Bf, stride, C, 1, B, st));
  UMT_TRY(launch_conv6<3>(Bf, stride, C, tw.reg_w[h][2], tw.reg_b[h][2], nullptr, 0, A, stride, C, 1, B, st));
  UMT_TRY(launch_conv6<3>(A, stride, C, tw.reg_w[h][3], tw.reg_b[h][3], Bf, stride, Cf, stride, C, 1, B, st));
  head_kernel<<<B, 128, 0, st>>>(Cf, C, tw.head_w[h], tw.head_b[h], head == UMT_HEAD_KNOWN_SKELETON ? 62 : 63,
                                 head == UMT_HEAD_UNKNOWN_SKELETON ? 1 : 0, extrinsics, sample_range,
                                 hand_idx, out_joint_angles, out_wrist, out_scale, out_sigma);
  UMT_TRY(check_launch("head_kernel"));
  m->launches += 6;
  return UMT_OK;
}

}  // namespace umt
