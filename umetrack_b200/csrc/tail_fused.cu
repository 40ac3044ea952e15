// The whole tail of ONE hand-frame in ONE launch, for small batches: the per-frame tracker and the time
// scan of the batched driver run the tail on 1-128 samples per step, where the fifteen launches of
// tail.cu (ten of them tensor-core convolutions on a few hundred positions) are launch latency
// (0.17 ms per step, 34 ms of host time per 256-step scan).  Here a sample's activations never leave
// shared memory and the convolutions are plain float32 FMAs on the CUDA cores: two launches per step
// (skeleton encoder + this kernel).
//  * `tail_fused_kernel<1>`: one CTA per sample.  Measured on B200 (UMT_TAIL_DBG=1): 239 k cycles per
//    sample, of which the four 3x3 regressor convolutions (7.5 M FMAs) are 163 k with their weights staged
//    through shared memory one chunk of 8 input channels ahead (read through L2 as they were needed they
//    cost one round trip per input channel: 200 k) and the six 1x1 convolutions 57 k: 0.144 ms per step
//    whatever the batch, up to one CTA per SM -- faster than the 15-launch path (0.17 ms) and 7x cheaper
//    on the host.
//  * `tail_fused_kernel<8>`, for up to 24 samples (a frame of the per-frame tracker, a frame index of a
//    small time scan): a sample is spread over a CLUSTER of 8 CTAs.  Every CTA keeps the whole
//    activation set in its own shared memory, computes one eighth of each layer's output channels from a
//    weight slice that cp.async brought into shared memory during the previous layer, and after one
//    cluster barrier per layer fetches the other seven eighths from its peers' shared memory
//    (ld.shared::cluster, 16 bytes).  98 k cycles per sample: 0.07 ms per step up to 12 samples (one wave of
//    clusters), 0.12 ms up to 24.
// Each output element is produced by one thread with the same FMA order in both variants, so they give
// bit-identical results (tests/test_gpu_model.py).
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
constexpr int kChunkCi = 8;                    // input channels per staged weight chunk of the one-CTA 3x3 convolution
constexpr int kChunkFloats = kChunkCi * 9 * 76;
constexpr int kFusedSmem = (kPoolFloats + 256 + 2 * kChunkFloats) * 4;   // pool, small arrays, two weight chunks
constexpr int kClusterBase = (kPoolFloats + 256) * 4;                      // (the cluster variant adds two weight slices instead)

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
    // (unrolled by the depth of the queue, so that its rotation is register renaming and not moves that
    // wait for the newest load)
#pragma unroll 3
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

// The same convolution with the weights staged through shared memory: chunks of kChunkCi input channels
// ([ci][tap][co], contiguous in global memory) arrive by cp.async one chunk ahead of the FMAs, so the loop
// waits for no L2 round trip (conv3x3_cta above does, once per input channel: 50 k cycles per layer against
// 33 k here).  Same work items, same FMA order per output element.  Called by all threads of the CTA.
__device__ __forceinline__ void stage_chunk3x3(float* dst, const float* __restrict__ w, int c, int ci0) {
  const int nci = (c - ci0 < kChunkCi) ? c - ci0 : kChunkCi;
  const int n = nci * 9 * c;                         // 9 * c is a multiple of 4 for c = 72, 76
  const float* src = w + (size_t)ci0 * 9 * c;
  for (int i = 4 * (int)threadIdx.x; i < n; i += 4 * kFusedThreads) {
    const uint32_t sa = (uint32_t)__cvta_generic_to_shared(dst + i);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(src + i) : "memory");
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
}
__device__ __forceinline__ void conv3x3_cta_staged(const float* __restrict__ in, float* __restrict__ out, int c,
                                                   const float* __restrict__ w, const float* __restrict__ bias,
                                                   const float* __restrict__ res, float* wbuf) {
  const int item = threadIdx.x;
  const bool active = item < 3 * c;
  const int g = active ? item / c : 0, co = active ? item - g * c : 0;
  float acc[2][6];
  const float b = __ldg(bias + co);
#pragma unroll
  for (int y = 0; y < 2; ++y)
#pragma unroll
    for (int x = 0; x < 6; ++x) acc[y][x] = b;
  const int n_chunks = (c + kChunkCi - 1) / kChunkCi;
  stage_chunk3x3(wbuf, w, c, 0);
  for (int ch = 0; ch < n_chunks; ++ch) {
    if (ch + 1 < n_chunks) {
      stage_chunk3x3(wbuf + ((ch + 1) & 1) * kChunkFloats, w, c, (ch + 1) * kChunkCi);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();                                   // chunk ch is in shared memory for every thread
    if (active) {
      const float* wb = wbuf + (ch & 1) * kChunkFloats + co;
      const int ci0 = ch * kChunkCi, nci = (c - ci0 < kChunkCi) ? c - ci0 : kChunkCi;
#pragma unroll 2
      for (int cl = 0; cl < nci; ++cl) {
        float wc[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) wc[k] = wb[(cl * 9 + k) * c];
        float win[4][8];
        const float4* src = reinterpret_cast<const float4*>(in + (ci0 + cl) * kPad + 2 * g * 8);
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
    }
    __syncthreads();                                   // chunk ch's buffer may be refilled
  }
  if (active) {
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

// ---- cluster variant: helpers (kClusterCtas, cluster_slice_floats: model.cuh) ----

__device__ __forceinline__ unsigned cluster_rank() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
template <int CS>
__device__ __forceinline__ void sync_all() {
  if constexpr (CS == 1) __syncthreads(); else cluster_sync_all();
}
// After a layer: fetch the other CTAs' output channels of `buf` ([cout][stride] floats, stride 36 or 64)
// from their shared memory into the same place of this CTA's.  Up to 5 x 16 bytes per thread, loads first.
template <int CS>
__device__ __forceinline__ void pull_slices(float* buf, int cout, int stride, int rank) {
  const int per = (cout + CS - 1) / CS, n4 = cout * stride / 4;
  const uint32_t base = (uint32_t)__cvta_generic_to_shared(buf);
  float4 v[5];
  bool take[5];
#pragma unroll
  for (int it = 0; it < 5; ++it) {
    const int i = (int)threadIdx.x + it * kFusedThreads;
    const int owner = i < n4 ? (i * 4 / stride) / per : rank;
    take[it] = owner != rank;
    if (take[it]) {
      uint32_t ra;
      asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(base + 16u * (uint32_t)i), "r"(owner));
      asm volatile("ld.shared::cluster.v4.f32 {%0,%1,%2,%3}, [%4];"
                   : "=f"(v[it].x), "=f"(v[it].y), "=f"(v[it].z), "=f"(v[it].w) : "r"(ra) : "memory");
    }
  }
#pragma unroll
  for (int it = 0; it < 5; ++it)
    if (take[it]) reinterpret_cast<float4*>(buf)[(int)threadIdx.x + it * kFusedThreads] = v[it];
}

// Cluster variant: CTA `rank` owns the output channels [rank * per, rank * per + per) of every layer
// (per = ceil(cout / 8) <= 14; the last CTA may own fewer).  tail.cu packs each layer's weights per CTA:
// slice r = [rows][16] weights (rows = input channels, x 9 taps for 3x3) followed by [16] biases.  A slice
// (<= 44 KB) is copied to shared memory with cp.async while the previous layer computes, so the inner loops
// see no global-memory latency (the one-CTA variant is bound by exactly that: one L2 round trip per input
// channel).
constexpr int kWBufFloats = cluster_slice_floats(76 * 9);      // largest slice: a 3x3 layer of the k-head
constexpr int kClusterSmem = kClusterBase + 2 * kWBufFloats * 4;
struct Slice { int co0, nco, per; };
template <int CS>
__device__ __forceinline__ Slice slice_of(int cout, int rank) {
  const int per = (cout + CS - 1) / CS, co0 = rank * per;
  const int nco = cout - co0 < per ? cout - co0 : per;
  return {co0, nco < 0 ? 0 : nco, per};
}
__device__ __forceinline__ void stage_slice(float* dst, const float* __restrict__ packed, int rows, int rank) {
  const int n = cluster_slice_floats(rows);
  const float* src = packed + (size_t)rank * n;
  for (int i = 4 * (int)threadIdx.x; i < n; i += 4 * kFusedThreads) {
    const uint32_t sa = (uint32_t)__cvta_generic_to_shared(dst + i);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(src + i) : "memory");
  }
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// 1x1 convolution, cluster variant: work item = (own output channel, 6 consecutive pixels), results into
// this CTA's copy of `out`.  `wb` = this CTA's slice in shared memory: [cin][16] weights, [16] biases.
__device__ __forceinline__ void conv1x1_cl(const float* __restrict__ in, int cin, float* __restrict__ out, Slice s,
                                           const float* __restrict__ wb, bool relu) {
  const int t = threadIdx.x;
  if (t >= 6 * s.nco) return;
  const int g = t / s.nco, j = t - g * s.nco, co = s.co0 + j;
  float acc[6];
  const float b = wb[cin * kSliceStride + j];
#pragma unroll
  for (int i = 0; i < 6; ++i) acc[i] = b;
  const float* ip = in + 6 * g;
  const float* wp = wb + j;
#pragma unroll 6
  for (int ci = 0; ci < cin; ++ci) {
    const float wv = wp[ci * kSliceStride];
    const float2* row = reinterpret_cast<const float2*>(ip + ci * kPix);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const float2 v = row[i];
      acc[2 * i] = fmaf(wv, v.x, acc[2 * i]);
      acc[2 * i + 1] = fmaf(wv, v.y, acc[2 * i + 1]);
    }
  }
  float2* o = reinterpret_cast<float2*>(out + co * kPix + 6 * g);
#pragma unroll
  for (int i = 0; i < 3; ++i)
    o[i] = relu ? make_float2(fmaxf(acc[2 * i], 0.f), fmaxf(acc[2 * i + 1], 0.f)) : make_float2(acc[2 * i], acc[2 * i + 1]);
}

// 3x3 convolution, cluster variant: work item = (own output channel, image row, half of the row); the two
// halves are dealt to different warps (threads 0-63 / 64-127) so that the window columns are compile-time
// register indices.  `wb` = [c][9][16] weights, [16] biases.
template <int HALF>
__device__ __forceinline__ void conv3x3_half(const float* __restrict__ in, float* __restrict__ out, int c, Slice s,
                                             const float* __restrict__ wb, const float* __restrict__ res, int item) {
  const int y = item / s.nco, j = item - y * s.nco, co = s.co0 + j;
  float acc[3];
  const float b = wb[c * 9 * kSliceStride + j];
#pragma unroll
  for (int x = 0; x < 3; ++x) acc[x] = b;
  const float* wp = wb + j;
  const float* ip = in + y * 8;
#pragma unroll 4
  for (int ci = 0; ci < c; ++ci) {
    float wc[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) wc[k] = wp[(ci * 9 + k) * kSliceStride];
    float win[3][5];      // padded columns 3*HALF .. 3*HALF+4 of the rows y .. y+2
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const float* row = ip + ci * kPad + r * 8;
      if (HALF == 0) {
        const float4 v = *reinterpret_cast<const float4*>(row);
        win[r][0] = v.x; win[r][1] = v.y; win[r][2] = v.z; win[r][3] = v.w; win[r][4] = row[4];
      } else {
        const float4 v = *reinterpret_cast<const float4*>(row + 4);
        win[r][0] = row[3]; win[r][1] = v.x; win[r][2] = v.y; win[r][3] = v.z; win[r][4] = v.w;
      }
    }
#pragma unroll
    for (int ky = 0; ky < 3; ++ky)
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const float wv = wc[ky * 3 + kx];
#pragma unroll
        for (int x = 0; x < 3; ++x) acc[x] = fmaf(wv, win[ky][x + kx], acc[x]);
      }
  }
#pragma unroll
  for (int x = 0; x < 3; ++x) {
    const int o = co * kPad + (y + 1) * 8 + 3 * HALF + x + 1;
    float v = acc[x];
    if (res) v += res[o];
    out[o] = fmaxf(v, 0.f);
  }
}
__device__ __forceinline__ void conv3x3_cl(const float* __restrict__ in, float* __restrict__ out, int c, Slice s,
                                           const float* __restrict__ wb, const float* __restrict__ res) {
  const int t = threadIdx.x, item = t & 63;
  if (t >= 128 || item >= 6 * s.nco) return;      // per <= 10 output channels: at most 60 items per half
  if (t < 64) conv3x3_half<0>(in, out, c, s, wb, res, item);
  else conv3x3_half<1>(in, out, c, s, wb, res, item);
}

struct FusedParams {
  const float *feats, *intrinsics, *extrinsics;
  const int64_t *sample_range, *memory_idx, *hand_idx;
  const uint8_t* use_memory;
  const float* skel;   // [n_skel,4,36] skeleton feature maps (known-skeleton head) or null
  int n_skel, n_slots, head, c, n_out;
  float *mem_state, *prev_extr;
  const float *fus_w[3], *fus_b[3], *tmp_w[3], *tmp_b[3], *reg_w[4], *reg_b[4], *head_w, *head_b;
  const float *fus_cl[3], *tmp_cl[3], *reg_cl[4];   // the same layers packed per CTA for the cluster variant
  float *out_angles, *out_wrist, *out_scale, *out_sigma;
  unsigned* status;
  int dbg;   // UMT_TAIL_DBG: block 0 prints the cycles of its phases
};

// One layer's weights as the cluster variant stages them.
struct LayerW { const float* w; int rows, cout; };

template <int CS>
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
  const int b = blockIdx.x / CS, tid = threadIdx.x;
  const int rank = CS == 1 ? 0 : (int)cluster_rank();      // cluster = CS consecutive CTAs = one sample
  const bool lead = rank == 0;                              // the CTA that writes to global memory
  const int r0 = (int)P.sample_range[2 * b], r1 = (int)P.sample_range[2 * b + 1];
  const bool multi = (r1 - r0) == 2;
  const int slot = (int)P.memory_idx[b];
  const bool slot_ok = slot >= 0 && slot < P.n_slots;
  // cluster variant: the layers this sample runs, and the staging of their weight slices (two buffers)
  float* wbuf = pool + kPoolFloats + 256;
  LayerW Ls[10];
  int nl = 0, li = 0;
  auto stage = [&](int i) { stage_slice(wbuf + (i & 1) * kWBufFloats, Ls[i].w, Ls[i].rows, rank); };
  if constexpr (CS > 1) {
    if (multi) {
      Ls[nl++] = {P.fus_cl[0], 144, 108}; Ls[nl++] = {P.fus_cl[1], 108, 72}; Ls[nl++] = {P.fus_cl[2], 72, 72};
    }
    for (int i = 0; i < 3; ++i) Ls[nl++] = {P.tmp_cl[i], 90, 90};
    for (int i = 0; i < 4; ++i) Ls[nl++] = {P.reg_cl[i], P.c * 9, P.c};
    stage(0);
    cp_async_commit();
  }
  // In front of a convolution layer.  Cluster variant: this layer's weight slice has landed, and the next
  // layer's starts moving into the buffer the previous layer has released.  `exchange` is set where other
  // CTAs may still be fetching the buffer this layer writes (see the call sites).
  long long dbg_wait = 0, dbg_sync = 0, dbg_stage = 0, dbg_c1 = 0, dbg_c3 = 0, dbg_pull = 0;
  auto layer_begin = [&](bool exchange) -> const float* {
    if constexpr (CS == 1) {
      __syncthreads();
      return nullptr;
    } else {
      const long long t0 = P.dbg ? clock64() : 0;
      cp_async_wait_all();
      if (exchange) cluster_sync_all(); else __syncthreads();
      const long long t1 = P.dbg ? clock64() : 0;
      const float* wb = wbuf + (li & 1) * kWBufFloats;
      if (++li < nl) stage(li);
      cp_async_commit();
      if (P.dbg) { const long long t2 = clock64(); dbg_wait += t1 - t0; dbg_stage += t2 - t1; }
      return wb;
    }
  };
  // Behind a layer of the cluster variant: one cluster barrier (every CTA's slice is written), then the
  // other CTAs' slices are fetched.  The caller's next barrier makes them visible to the whole CTA.
  auto layer_end = [&](float* out, int cout, int stride) {
    if constexpr (CS > 1) {
      const long long t0 = P.dbg ? clock64() : 0;
      cluster_sync_all();
      const long long t1 = P.dbg ? clock64() : 0;
      pull_slices<CS>(out, cout, stride, rank);
      if (P.dbg) { const long long t2 = clock64(); dbg_sync += t1 - t0; dbg_pull += t2 - t1; }
    }
  };
  auto c1 = [&](const float* in, int cin, float* out, int cout, const float* w, const float* bias, bool relu,
                bool exchange = false) {
    const float* wb = layer_begin(exchange);
    const long long t0 = P.dbg ? clock64() : 0;
    if constexpr (CS == 1) conv1x1_cta(in, cin, out, cout, w, bias, relu);
    else conv1x1_cl(in, cin, out, slice_of<CS>(cout, rank), wb, relu);
    if (P.dbg) dbg_c1 += clock64() - t0;
    layer_end(out, cout, kPix);
  };
  auto c3 = [&](const float* in, float* out, const float* w, const float* bias, const float* res) {
    const float* wb = layer_begin(false);
    const long long t0 = P.dbg ? clock64() : 0;
    if constexpr (CS == 1) conv3x3_cta_staged(in, out, P.c, w, bias, res, wbuf);
    else conv3x3_cl(in, out, P.c, slice_of<CS>(P.c, rank), wb, res);
    if (P.dbg) dbg_c3 += clock64() - t0;
    layer_end(out, P.c, kPad);
  };
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
    if (!slot_ok && lead) atomicOr(P.status, (unsigned)UMT_STATUS_MEMORY_INDEX);
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
  sync_all<CS>();     // (cluster: every CTA is running and has read prev_extr before the leader rewrites it)
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
    c1(A, 144, B, 108, P.fus_w[0], P.fus_b[0], true);
    c1(B, 108, A, 72, P.fus_w[1], P.fus_b[1], true);
    c1(A, 72, B, 72, P.fus_w[2], P.fus_b[2], false);
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
    if (slot_ok && lead && tid < 16) P.prev_extr[(size_t)slot * 16 + tid] = P.extrinsics[(size_t)r0 * 16 + tid];   // temporal.py:77
  }
  // (the barrier in front of the layer also keeps other CTAs from storing into B while it is still read above)
  c1(A, 90, B, 90, P.tmp_w[0], P.tmp_b[0], true, /*exchange=*/true);   // other CTAs may still be fetching B (fusion output)
  c1(B, 90, A, 90, P.tmp_w[1], P.tmp_b[1], true);
  c1(A, 90, B, 90, P.tmp_w[2], P.tmp_b[2], false);
  __syncthreads();
  tick();

  // ---- memory write-back, regressor input [72 features | 4 skeleton] as a zero-bordered map ----
  if (slot_ok && lead) {
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
  sync_all<CS>();     // (cluster: nobody is still fetching B, which R1 overlaps)
  for (int i = tid; i < 76 * kPad; i += kFusedThreads) R1[i] = 0.f;       // (B is dead now)
  tick();
  // ---- two BasicBlocks (backbone_resnet.py:56-72); the barrier in front of the first one also makes sure
  // every CTA has cleared its maps before results are stored into them ----
  c3(R2, R0, P.reg_w[0], P.reg_b[0], nullptr);
  c3(R0, R1, P.reg_w[1], P.reg_b[1], R2);
  c3(R1, R0, P.reg_w[2], P.reg_b[2], nullptr);
  c3(R0, R2, P.reg_w[3], P.reg_b[3], R1);
  sync_all<CS>();     // last access to other CTAs' shared memory
  tick();
  if (!lead) return;

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
      printf("tail_fused<%d> cycles: transforms %lld, fusion %lld, temporal %lld, pack %lld, regressor %lld, head %lld\n",
             CS, tk[1] - tk[0], tk[2] - tk[1], tk[3] - tk[2], tk[4] - tk[3], tk[5] - tk[4], tk[6] - tk[5]);
    if (b == 0 && tid == 0)
      printf("  layers: weight wait %lld, staging %lld, cluster barrier %lld, fetch %lld; thread 0 in 1x1 layers %lld, 3x3 %lld\n",
             dbg_wait, dbg_stage, dbg_sync, dbg_pull, dbg_c1, dbg_c3);
  }
}

// Largest batch that takes this path (UMT_TAIL_FUSED_MAX overrides; 0 switches it off).  Beyond one wave
// of CTAs the tensor-core tail of tail.cu is faster (256 samples: 0.18 ms against 0.25 ms).
int tail_fused_max_samples() {
  const char* e = getenv("UMT_TAIL_FUSED_MAX");
  return e ? atoi(e) : 128;   // up to one CTA per SM: a single wave
}

// Up to this many samples a sample is spread over a cluster of 8 CTAs (UMT_TAIL_CLUSTER_MAX overrides; 0
// switches the cluster variant off).  12-14 clusters of 8 CTAs with 147 KB of shared memory are resident at a
// time: 0.07 ms per wave against 0.16 ms for one CTA per sample, so two waves still win.
int tail_cluster_max_samples() {
  const char* e = getenv("UMT_TAIL_CLUSTER_MAX");
  return e ? atoi(e) : 24;
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
  for (int i = 0; i < 4; ++i) { p.reg_w[i] = tw.reg_w[head][i]; p.reg_b[i] = tw.reg_b[head][i]; p.reg_cl[i] = tw.reg_cl[head][i]; }
  for (int i = 0; i < 3; ++i) { p.fus_cl[i] = tw.fus_cl[i]; p.tmp_cl[i] = tw.tmp_cl[i]; }
  p.head_w = tw.head_w[head]; p.head_b = tw.head_b[head];
  p.out_angles = out_joint_angles; p.out_wrist = out_wrist; p.out_scale = out_scale; p.out_sigma = out_sigma;
  p.status = m->d_status;
  p.dbg = getenv("UMT_TAIL_DBG") ? 1 : 0;
  {  // opt-in shared-memory size: once per device
    static bool done[64] = {};
    int dev = 0;
    UMT_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !done[dev]) {
      UMT_CUDA_TRY(cudaFuncSetAttribute(tail_fused_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFusedSmem));
      UMT_CUDA_TRY(cudaFuncSetAttribute(tail_fused_kernel<kClusterCtas>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        kClusterSmem));
      if (dev >= 0 && dev < 64) done[dev] = true;
    }
  }
  if (n_samples <= tail_cluster_max_samples()) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)(n_samples * kClusterCtas));
    cfg.blockDim = dim3(kFusedThreads);
    cfg.dynamicSmemBytes = kClusterSmem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = kClusterCtas;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    UMT_CUDA_TRY(cudaLaunchKernelEx(&cfg, tail_fused_kernel<kClusterCtas>, p));
  } else {
    tail_fused_kernel<1><<<n_samples, kFusedThreads, kFusedSmem, st>>>(p);
  }
  UMT_TRY(check_launch("tail_fused_kernel"));
  ++m->launches;
  return UMT_OK;
}

}  // namespace umt
