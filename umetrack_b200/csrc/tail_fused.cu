// The whole tail of ONE hand-frame in ONE CTA, for small batches: the per-frame tracker and the time
// scan of the batched driver run the tail on 1-32 samples per step, where the fifteen launches of
// tail.cu (ten of them tensor-core convolutions on a few hundred positions) are launch latency
// (0.17 ms per step, 34 ms of host time per 256-step scan).  Here a sample's activations never leave
// shared memory and the convolutions are plain float32 FMAs on the CUDA cores: two launches per step
// (skeleton encoder + this kernel).  Measured on B200 (UMT_TAIL_DBG=1): 278 k cycles per sample, of which
// the four 3x3 regressor convolutions are 202 k -- 7.5 M FMAs on one SM's FP32 pipes (three-register FFMA
// issues every other cycle per scheduler: 64 FMA/clk), i.e. the kernel is FP32-issue bound at 0.16 ms,
// on par with the 15-launch path in device time and 7x cheaper on the host.  Splitting a sample over a
// cluster of CTAs is the next step if the scan has to get faster.
//   feature-transform layer + two-view fusion   lib/models/feature_extractor.py:61-141,
//                                                lib/models/model_utils.py:57-104,141-192
//   temporal memory (SimpleConvRNN)              lib/models/temporal.py:51-139
//   pose regressor + decoders + Procrustes       lib/models/regressor.py:76-186, model_utils.py:17-54
//   world-space recovery / right-hand un-mirror  lib/models/umetrack_model.py:77-97
// Arithmetic: float32 products and sums like the reference (no operand splitting needed on this path);
// results agree with the tensor-core tail to ~1e-6 relative.
#include <cstdio>
#include <string>

#include "model.cuh"
#include "tail_math.cuh"

namespace umt {

constexpr int kFusedThreads = 256;
constexpr int kPad = 64;                       // a 6x6 map with a zero border: 8 x 8
constexpr int kPoolFloats = 3 * 76 * kPad;     // three padded 76-channel maps; the 1x1 stages alias it
constexpr int kFusedSmem = (kPoolFloats + 256) * 4;

// out[co][p] = act(bias[co] + sum_ci w[ci][co] * in[ci][p]) for the 36 pixels of a sample.
// Work item = (output channel, half of the pixels); consecutive lanes = consecutive channels, so the
// weight reads coalesce and the activation reads are shared-memory broadcasts.
__device__ __forceinline__ void conv1x1_cta(const float* __restrict__ in, int cin, float* __restrict__ out, int cout,
                                            const float* __restrict__ w, const float* __restrict__ bias, bool relu) {
  for (int item = threadIdx.x; item < 2 * cout; item += kFusedThreads) {
    const int g = item / cout, co = item - g * cout;
    float acc[18];
    const float b = __ldg(bias + co);
#pragma unroll
    for (int i = 0; i < 18; ++i) acc[i] = b;
    const float* ip = in + 18 * g;
    // every layer's input width is a multiple of 18: fetch 18 weights at once (independent loads in flight;
    // a single weight load per iteration made the loop pure L2 latency)
    float wn[18];
#pragma unroll
    for (int k = 0; k < 18; ++k) wn[k] = __ldg(w + (size_t)k * cout + co);
    for (int c0 = 0; c0 < cin; c0 += 18) {
      float wv[18];
#pragma unroll
      for (int k = 0; k < 18; ++k) wv[k] = wn[k];
      if (c0 + 18 < cin) {
#pragma unroll
        for (int k = 0; k < 18; ++k) wn[k] = __ldg(w + (size_t)(c0 + 18 + k) * cout + co);
      }
#pragma unroll
      for (int k = 0; k < 18; ++k) {
        const float2* row = reinterpret_cast<const float2*>(ip + (c0 + k) * kPix);
#pragma unroll
        for (int i = 0; i < 9; ++i) {
          const float2 v = row[i];
          acc[2 * i] = fmaf(wv[k], v.x, acc[2 * i]);
          acc[2 * i + 1] = fmaf(wv[k], v.y, acc[2 * i + 1]);
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 18; ++i) out[co * kPix + 18 * g + i] = relu ? fmaxf(acc[i], 0.f) : acc[i];
  }
}

// 3x3 convolution (padding 1) on zero-bordered 8x8 maps: out = relu(bias + conv(in) [+ res]).
// Work item = (output channel, two image rows); per input channel the 4x8 input window is read once into
// registers and feeds 108 FMAs.
__device__ __forceinline__ void conv3x3_cta(const float* __restrict__ in, float* __restrict__ out, int c,
                                            const float* __restrict__ w, const float* __restrict__ bias,
                                            const float* __restrict__ res) {
  for (int item = threadIdx.x; item < 3 * c; item += kFusedThreads) {
    const int g = item / c, co = item - g * c;
    float acc[2][6];
    const float b = __ldg(bias + co);
#pragma unroll
    for (int y = 0; y < 2; ++y)
#pragma unroll
      for (int x = 0; x < 6; ++x) acc[y][x] = b;
    // weights are requested three input channels ahead (an L2 round trip is ~3 iterations of this loop)
    float wq[3][9];
#pragma unroll
    for (int d = 0; d < 3; ++d)
#pragma unroll
      for (int k = 0; k < 9; ++k) wq[d][k] = (d < c) ? __ldg(w + ((size_t)d * 9 + k) * c + co) : 0.f;
    for (int ci = 0; ci < c; ++ci) {
      float wc[9];
#pragma unroll
      for (int k = 0; k < 9; ++k) { wc[k] = wq[0][k]; wq[0][k] = wq[1][k]; wq[1][k] = wq[2][k]; }
      if (ci + 3 < c) {
        const float* wp = w + (size_t)(ci + 3) * 9 * c + co;
#pragma unroll
        for (int k = 0; k < 9; ++k) wq[2][k] = __ldg(wp + k * c);
      }
      float win[4][8];
      const float4* src = reinterpret_cast<const float4*>(in + ci * kPad + 2 * g * 8);
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const float4 lo = src[2 * r], hi = src[2 * r + 1];
        win[r][0] = lo.x; win[r][1] = lo.y; win[r][2] = lo.z; win[r][3] = lo.w;
        win[r][4] = hi.x; win[r][5] = hi.y; win[r][6] = hi.z; win[r][7] = hi.w;
      }
#pragma unroll
      for (int ky = 0; ky < 3; ++ky)
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
          const float wv = wc[ky * 3 + kx];
#pragma unroll
          for (int y = 0; y < 2; ++y)
#pragma unroll
            for (int x = 0; x < 6; ++x) acc[y][x] = fmaf(wv, win[y + ky][x + kx], acc[y][x]);
        }
    }
#pragma unroll
    for (int y = 0; y < 2; ++y)
#pragma unroll
      for (int x = 0; x < 6; ++x) {
        const int o = co * kPad + (2 * g + y + 1) * 8 + x + 1;
        float v = acc[y][x];
        if (res) v += res[o];
        out[o] = fmaxf(v, 0.f);
      }
  }
}

struct FusedParams {
  const float *feats, *intrinsics, *extrinsics;
  const int64_t *sample_range, *memory_idx, *hand_idx;
  const uint8_t* use_memory;
  const float* skel;   // [n_skel,4,36] skeleton feature maps (known-skeleton head) or null
  int n_skel, n_slots, head, c, n_out;
  float *mem_state, *prev_extr;
  const float *fus_w[3], *fus_b[3], *tmp_w[3], *tmp_b[3], *reg_w[4], *reg_b[4], *head_w, *head_b;
  float *out_angles, *out_wrist, *out_scale, *out_sigma;
  unsigned* status;
  int dbg;   // UMT_TAIL_DBG: block 0 prints the cycles of its phases
};

__global__ void __launch_bounds__(kFusedThreads)
tail_fused_kernel(const FusedParams P) {
  extern __shared__ __align__(16) float pool[];
  float* A = pool;                    // up to 144 x 36
  float* B = pool + 144 * kPix;       // up to 108 x 36
  float* R0 = pool;                   // padded 76 x 64 maps of the regressor
  float* R1 = pool + 76 * kPad;
  float* R2 = pool + 2 * 76 * kPad;
  float* sxf = pool + kPoolFloats;    // [2][12] view -> canonical
  float* sS = sxf + 24;               // canonical -> cam0 (or single view -> cam0)
  float* sM = sS + 12;                // previous cam0 -> current cam0
  float* sMean = sM + 12;             // [80]
  float* sRaw = sMean + 80;           // [64]
  __shared__ int sUse;
  const int b = blockIdx.x, tid = threadIdx.x;
  const int r0 = (int)P.sample_range[2 * b], r1 = (int)P.sample_range[2 * b + 1];
  const bool multi = (r1 - r0) == 2;
  const int slot = (int)P.memory_idx[b];
  const bool slot_ok = slot >= 0 && slot < P.n_slots;
  long long tk[8];
  int nk = 0;
  auto tick = [&]() { if (P.dbg && nk < 8) tk[nk++] = clock64(); };
  tick();

  // ---- transforms (feature_extractor.py:61-87, model_utils.py:166-192, temporal.py:59-69) ----
  if (tid < 2 && multi) {
    const int v = tid;
    double E0[16], Ev[16], Einv[16], S[16] = {0}, S0inv[16] = {0}, t1[16], t2[16], t3[16];
    load_mat4(P.extrinsics + (size_t)r0 * 16, E0);
    load_mat4(P.extrinsics + (size_t)(r0 + v) * 16, Ev);
    const float f0 = P.intrinsics[(size_t)r0 * 9] / 200.0f;
    const float fv = P.intrinsics[(size_t)(r0 + v) * 9] / 200.0f;
    S[0] = S[5] = S[15] = 1.0; S[10] = (double)fv;
    S0inv[0] = S0inv[5] = S0inv[15] = 1.0; S0inv[10] = (double)(1.0f / f0);
    const bool ok = mat4_inv(Ev, Einv);
    mat4_mul(Einv, S, t1);
    mat4_mul(E0, t1, t2);
    mat4_mul(S0inv, t2, t3);
    for (int i = 0; i < 12; ++i) sxf[v * 12 + i] = ok ? (float)t3[i] : nanf("");
  }
  if (tid == 32) {
    if (!slot_ok) atomicOr(P.status, (unsigned)UMT_STATUS_MEMORY_INDEX);
    for (int i = 0; i < 12; ++i) sS[i] = 0.f;
    sS[0] = sS[5] = 1.f;
    sS[10] = P.intrinsics[(size_t)r0 * 9] / 200.0f;
    int use = (slot_ok && P.use_memory[b]) ? 1 : 0;
    if (use) {
      double Pm[16], Pinv[16], C[16], X[16];
      load_mat4(P.prev_extr + (size_t)slot * 16, Pm);
      load_mat4(P.extrinsics + (size_t)r0 * 16, C);
      if (mat4_inv(Pm, Pinv)) {
        mat4_mul(C, Pinv, X);
        for (int i = 0; i < 12; ++i) sM[i] = (float)X[i];
      } else {
        use = 0;
      }
    }
    sUse = use;
  }
  __syncthreads();
  tick();

  // ---- two-view fusion (feature_extractor.py:121-133) or the single view's features ----
  if (multi) {
    for (int i = tid; i < 2 * 864; i += kFusedThreads) {
      const int v = i / 864, j = i - v * 864;
      const float* src = P.feats + (size_t)(r0 + v) * 72 * kPix;
      float ox, oy, oz;
      ftl_point(sxf + v * 12, src[j], src[864 + j], src[1728 + j], ox, oy, oz);
      float* d = A + v * 72 * kPix;
      d[j] = ox; d[864 + j] = oy; d[1728 + j] = oz;
    }
    __syncthreads();
    conv1x1_cta(A, 144, B, 108, P.fus_w[0], P.fus_b[0], true);
    __syncthreads();
    conv1x1_cta(B, 108, A, 72, P.fus_w[1], P.fus_b[1], true);
    __syncthreads();
    conv1x1_cta(A, 72, B, 72, P.fus_w[2], P.fus_b[2], false);
  } else {
    const float* src = P.feats + (size_t)r0 * 72 * kPix;
    for (int i = tid; i < 72 * kPix; i += kFusedThreads) B[i] = src[i];
  }
  __syncthreads();
  tick();

  // ---- temporal input: [warped memory (18) | features in cam0 (72)] (temporal.py:65-85) ----
  {
    const float* mem = slot_ok ? P.mem_state + (size_t)slot * 18 * kPix : nullptr;
    for (int j = tid; j < 216; j += kFusedThreads) {
      float ox = 0.f, oy = 0.f, oz = 0.f;
      if (sUse) ftl_point(sM, mem[j], mem[216 + j], mem[432 + j], ox, oy, oz);
      A[j] = ox; A[216 + j] = oy; A[432 + j] = oz;
    }
    float* o2 = A + 18 * kPix;
    for (int j = tid; j < 864; j += kFusedThreads) {
      float ox, oy, oz;
      ftl_point(sS, B[j], B[864 + j], B[1728 + j], ox, oy, oz);
      o2[j] = ox; o2[864 + j] = oy; o2[1728 + j] = oz;
    }
    if (slot_ok && tid < 16) P.prev_extr[(size_t)slot * 16 + tid] = P.extrinsics[(size_t)r0 * 16 + tid];   // temporal.py:77
  }
  __syncthreads();
  conv1x1_cta(A, 90, B, 90, P.tmp_w[0], P.tmp_b[0], true);
  __syncthreads();
  conv1x1_cta(B, 90, A, 90, P.tmp_w[1], P.tmp_b[1], true);
  __syncthreads();
  conv1x1_cta(A, 90, B, 90, P.tmp_w[2], P.tmp_b[2], false);
  __syncthreads();
  tick();

  // ---- memory write-back, regressor input [72 features | 4 skeleton] as a zero-bordered map ----
  if (slot_ok) {
    float* mem = P.mem_state + (size_t)slot * 18 * kPix;
    for (int i = tid; i < 18 * kPix; i += kFusedThreads) mem[i] = B[i];       // temporal.py:134
  }
  for (int i = tid; i < 76 * kPad; i += kFusedThreads) { R2[i] = 0.f; R0[i] = 0.f; }
  __syncthreads();
  {
    const float* sk = P.skel ? P.skel + (size_t)(P.n_skel == 1 ? 0 : b) * 4 * kPix : nullptr;
    for (int i = tid; i < P.c * kPix; i += kFusedThreads) {
      const int ch = i / kPix, p = i - ch * kPix;
      const float v = ch < 72 ? B[18 * kPix + i] : (sk ? sk[(ch - 72) * kPix + p] : 0.f);
      R2[ch * kPad + (p / 6 + 1) * 8 + (p % 6) + 1] = v;
    }
  }
  __syncthreads();
  for (int i = tid; i < 76 * kPad; i += kFusedThreads) R1[i] = 0.f;       // (B is dead now)
  __syncthreads();

  tick();
  // ---- two BasicBlocks (backbone_resnet.py:56-72) ----
  conv3x3_cta(R2, R0, P.c, P.reg_w[0], P.reg_b[0], nullptr);
  __syncthreads();
  conv3x3_cta(R0, R1, P.c, P.reg_w[1], P.reg_b[1], R2);
  __syncthreads();
  conv3x3_cta(R1, R0, P.c, P.reg_w[2], P.reg_b[2], nullptr);
  __syncthreads();
  conv3x3_cta(R0, R2, P.c, P.reg_w[3], P.reg_b[3], R1);
  __syncthreads();
  tick();

  // ---- average pool, 1x1 head (they commute), decoders ----
  if (tid < 80) {
    float s = 0.f;
    if (tid < P.c)
      for (int p = 0; p < kPix; ++p) s += R2[tid * kPad + (p / 6 + 1) * 8 + (p % 6) + 1];
    sMean[tid] = s * (1.0f / 36.0f);
  }
  __syncthreads();
  for (int o = tid; o < P.n_out; o += kFusedThreads) {
    float a = P.head_b[o];
    for (int c = 0; c < 80; ++c) a = fmaf(sMean[c], P.head_w[(size_t)o * 80 + c], a);
    sRaw[o] = a;
  }
  __syncthreads();
  head_decode(tid, sRaw, b, P.head == UMT_HEAD_UNKNOWN_SKELETON ? 1 : 0, P.extrinsics, P.sample_range, P.hand_idx,
              P.out_angles, P.out_wrist, P.out_scale, P.out_sigma);
  if (P.dbg) {
    __syncthreads();
    tick();
    if (b == 0 && tid == 0)
      printf("tail_fused cycles: transforms %lld, fusion %lld, temporal %lld, pack %lld, regressor %lld, head %lld\n",
             tk[1] - tk[0], tk[2] - tk[1], tk[3] - tk[2], tk[4] - tk[3], tk[5] - tk[4], tk[6] - tk[5]);
  }
}

// Largest batch that takes this path (UMT_TAIL_FUSED_MAX overrides; 0 switches it off).  Beyond one wave
// of CTAs the tensor-core tail of tail.cu is faster (256 samples: 0.18 ms against 0.25 ms).
int tail_fused_max_samples() {
  const char* e = getenv("UMT_TAIL_FUSED_MAX");
  return e ? atoi(e) : 128;   // up to one CTA per SM: a single wave
}

int tail_fused_forward(umt_model* m, int head, const float* feats, const float* intrinsics,
                       const float* extrinsics, const int64_t* sample_range, const int64_t* memory_idx,
                       const uint8_t* use_memory, const int64_t* hand_idx, int n_samples, const float* skel,
                       int n_skel, float* mem_state, float* prev_extr, int n_slots, float* out_joint_angles,
                       float* out_wrist, float* out_scale, float* out_sigma, cudaStream_t st) {
  FusedParams p{};
  p.feats = feats; p.intrinsics = intrinsics; p.extrinsics = extrinsics;
  p.sample_range = sample_range; p.memory_idx = memory_idx; p.hand_idx = hand_idx; p.use_memory = use_memory;
  p.skel = skel; p.n_skel = n_skel; p.n_slots = n_slots; p.head = head;
  p.c = head == UMT_HEAD_KNOWN_SKELETON ? 76 : 72;
  p.n_out = head == UMT_HEAD_KNOWN_SKELETON ? 62 : 63;
  p.mem_state = mem_state; p.prev_extr = prev_extr;
  const TailWeights& tw = m->tw;
  for (int i = 0; i < 3; ++i) { p.fus_w[i] = tw.fus_w[i]; p.fus_b[i] = tw.fus_b[i]; p.tmp_w[i] = tw.tmp_w[i]; p.tmp_b[i] = tw.tmp_b[i]; }
  for (int i = 0; i < 4; ++i) { p.reg_w[i] = tw.reg_w[head][i]; p.reg_b[i] = tw.reg_b[head][i]; }
  p.head_w = tw.head_w[head]; p.head_b = tw.head_b[head];
  p.out_angles = out_joint_angles; p.out_wrist = out_wrist; p.out_scale = out_scale; p.out_sigma = out_sigma;
  p.status = m->d_status;
  p.dbg = getenv("UMT_TAIL_DBG") ? 1 : 0;
  {  // opt-in shared-memory size: once per device
    static bool done[64] = {};
    int dev = 0;
    UMT_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !done[dev]) {
      UMT_CUDA_TRY(cudaFuncSetAttribute(tail_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kFusedSmem));
      if (dev >= 0 && dev < 64) done[dev] = true;
    }
  }
  tail_fused_kernel<<<n_samples, kFusedThreads, kFusedSmem, st>>>(p);
  UMT_TRY(check_launch("tail_fused_kernel"));
  ++m->launches;
  return UMT_OK;
}

}  // namespace umt
