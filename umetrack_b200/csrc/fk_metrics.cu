// Stage (c): forward kinematics + linear-blend skinning of the 21 landmarks, and the
// MPJPE / keypoint-acceleration reductions (see include/umetrack_b200.h).
//
// Warp-level kernels: one warp owns one hand pose.  Lanes 0..19 build the 20 finger-joint local
// transforms (clamped-angle Rodrigues, lib/common/hand_skinning.py:35-53), lanes 0..4 walk the
// five 4-joint finger chains (hand_skinning.py:17-32,100-127), lanes 0..20 blend one landmark
// each (hand_skinning.py:56-97,130-186).  The metric kernels also use one warp per pose and
// finish with a shuffle reduction and one float64 atomic per CTA.
#include "common.cuh"

namespace {

constexpr int kWarpsPerBlock = 4;

struct Xf {  // 3x4 rigid/affine transform, row major
  float m[12];
};

__device__ __forceinline__ Xf compose(const Xf& a, const Xf& b) {
  Xf c;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float s = a.m[i * 4 + 0] * b.m[0 * 4 + j];
      s = fmaf(a.m[i * 4 + 1], b.m[1 * 4 + j], s);
      s = fmaf(a.m[i * 4 + 2], b.m[2 * 4 + j], s);
      if (j == 3) s += a.m[i * 4 + 3];
      c.m[i * 4 + j] = s;
    }
  }
  return c;
}

__global__ void __launch_bounds__(kWarpsPerBlock * 32)
fk_kernel(const float* __restrict__ axes, const float* __restrict__ rest,
          const float* __restrict__ lm_rest, const float* __restrict__ lm_idx,
          const float* __restrict__ lm_w, const float* __restrict__ joint_angles,
          const float* __restrict__ wrist, const uint8_t* __restrict__ mirror, int n,
          float* __restrict__ out) {
  __shared__ Xf sLocal[kWarpsPerBlock][20];
  __shared__ Xf sFrame[kWarpsPerBlock][17];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int pose = blockIdx.x * kWarpsPerBlock + warp;
  if (pose >= n) return;

  // wrist transform (root and wrist frames share it); right hands flip column 0
  Xf W;
  {
    const float* w = wrist + (size_t)pose * 16;
    const float sgn = (mirror && mirror[pose]) ? -1.0f : 1.0f;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      W.m[i * 4 + 0] = w[i * 4 + 0] * sgn;
      W.m[i * 4 + 1] = w[i * 4 + 1];
      W.m[i * 4 + 2] = w[i * 4 + 2];
      W.m[i * 4 + 3] = w[i * 4 + 3];
    }
  }
  if (lane < 20) {
    const float a = joint_angles[(size_t)pose * UMT_N_JOINTS + lane];
    const float vx = axes[lane * 3 + 0] * a, vy = axes[lane * 3 + 1] * a, vz = axes[lane * 3 + 2] * a;
    const float nrm2 = fmaxf(vx * vx + vy * vy + vz * vz, 1e-4f);
    const float th = sqrtf(nrm2);
    const float f1 = sinf(th) / th;
    const float f2 = (1.0f - cosf(th)) / (th * th);
    // K = hat(v), K2 = K*K
    const float k2xx = -(vy * vy + vz * vz), k2yy = -(vx * vx + vz * vz), k2zz = -(vx * vx + vy * vy);
    const float k2xy = vx * vy, k2xz = vx * vz, k2yz = vy * vz;
    float R[9];
    R[0] = 1.0f + f2 * k2xx;
    R[1] = f1 * (-vz) + f2 * k2xy;
    R[2] = f1 * (vy) + f2 * k2xz;
    R[3] = f1 * (vz) + f2 * k2xy;
    R[4] = 1.0f + f2 * k2yy;
    R[5] = f1 * (-vx) + f2 * k2yz;
    R[6] = f1 * (-vy) + f2 * k2xz;
    R[7] = f1 * (vx) + f2 * k2yz;
    R[8] = 1.0f + f2 * k2zz;
    const float rx = rest[lane * 3 + 0], ry = rest[lane * 3 + 1], rz = rest[lane * 3 + 2];
    Xf L;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      L.m[i * 4 + 0] = R[i * 3 + 0];
      L.m[i * 4 + 1] = R[i * 3 + 1];
      L.m[i * 4 + 2] = R[i * 3 + 2];
      const float rr = (i == 0) ? rx : (i == 1 ? ry : rz);
      L.m[i * 4 + 3] = rr - (R[i * 3 + 0] * rx + R[i * 3 + 1] * ry + R[i * 3 + 2] * rz);
    }
    sLocal[warp][lane] = L;
  }
  __syncwarp();
  if (lane < 5) {
    Xf cur = compose(W, sLocal[warp][4 * lane + 0]);
    cur = compose(cur, sLocal[warp][4 * lane + 1]);
    sFrame[warp][2 + 3 * lane + 0] = cur;
    cur = compose(cur, sLocal[warp][4 * lane + 2]);
    sFrame[warp][2 + 3 * lane + 1] = cur;
    cur = compose(cur, sLocal[warp][4 * lane + 3]);
    sFrame[warp][2 + 3 * lane + 2] = cur;
  } else if (lane == 5) {
    sFrame[warp][0] = W;
    sFrame[warp][1] = W;
  }
  __syncwarp();
  if (lane < UMT_N_LANDMARKS) {
    const float px = lm_rest[lane * 3 + 0], py = lm_rest[lane * 3 + 1], pz = lm_rest[lane * 3 + 2];
    int bi[3];
    float bw[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      bi[k] = (int)lm_idx[lane * 3 + k];
      bw[k] = lm_w[lane * 3 + k];
    }
    // the reference scatters non-zero weights into a [21,17] matrix: a later entry with the same
    // bone index overwrites an earlier one (hand_skinning.py:87-94)
    if (bw[0] != 0.f && ((bw[1] != 0.f && bi[1] == bi[0]) || (bw[2] != 0.f && bi[2] == bi[0]))) bw[0] = 0.f;
    if (bw[1] != 0.f && (bw[2] != 0.f && bi[2] == bi[1])) bw[1] = 0.f;
    float ox = 0.f, oy = 0.f, oz = 0.f;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      if (bw[k] == 0.f || bi[k] < 0 || bi[k] >= 17) continue;
      const Xf& T = sFrame[warp][bi[k]];
      const float wx = px * bw[k], wy = py * bw[k], wz = pz * bw[k], ww = bw[k];
      ox += T.m[0] * wx + T.m[1] * wy + T.m[2] * wz + T.m[3] * ww;
      oy += T.m[4] * wx + T.m[5] * wy + T.m[6] * wz + T.m[7] * ww;
      oz += T.m[8] * wx + T.m[9] * wy + T.m[10] * wz + T.m[11] * ww;
    }
    float* o = out + ((size_t)pose * UMT_N_LANDMARKS + lane) * 3;
    o[0] = ox;
    o[1] = oy;
    o[2] = oz;
  }
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// one warp per (pose); CTA partial sums -> one atomic pair per CTA
__global__ void __launch_bounds__(256)
mpjpe_kernel(const float* __restrict__ pred, const float* __restrict__ gt,
             const uint8_t* __restrict__ valid, int n, double* __restrict__ sums) {
  __shared__ double sErr[8], sCnt[8];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double err = 0.0, cnt = 0.0;
  for (int pose = blockIdx.x * 8 + warp; pose < n; pose += gridDim.x * 8) {
    if (valid && !valid[pose]) continue;
    double d = 0.0;
    if (lane < UMT_N_LANDMARKS) {
      const size_t o = ((size_t)pose * UMT_N_LANDMARKS + lane) * 3;
      const double dx = (double)gt[o] - (double)pred[o], dy = (double)gt[o + 1] - (double)pred[o + 1],
                   dz = (double)gt[o + 2] - (double)pred[o + 2];
      d = sqrt(dx * dx + dy * dy + dz * dz);
    }
    d = warp_sum(d);
    err += d / (double)UMT_N_LANDMARKS;
    cnt += 1.0;
  }
  if (lane == 0) {
    sErr[warp] = err;
    sCnt[warp] = cnt;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double e = 0.0, c = 0.0;
    for (int i = 0; i < 8; ++i) {
      e += sErr[i];
      c += sCnt[i];
    }
    if (c > 0.0) {
      atomicAdd(&sums[0], e);
      atomicAdd(&sums[1], c);
    }
  }
}

__global__ void __launch_bounds__(256)
accel_kernel(const float* __restrict__ pts, const uint8_t* __restrict__ valid, int n_seq, int T,
             double* __restrict__ sums) {
  __shared__ double sAcc[8], sCnt[8];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int per = T - 2;
  const long long total = (long long)n_seq * (per > 0 ? per : 0);
  double acc = 0.0, cnt = 0.0;
  for (long long w = (long long)blockIdx.x * 8 + warp; w < total; w += (long long)gridDim.x * 8) {
    const int s = (int)(w / per), t = (int)(w % per);
    const uint8_t* v = valid + (size_t)s * T + t;
    if (!(v[0] && v[1] && v[2])) continue;
    double d = 0.0;
    if (lane < UMT_N_LANDMARKS) {
      const size_t stride = (size_t)UMT_N_LANDMARKS * 3;
      const float* p0 = pts + ((size_t)s * T + t) * stride + lane * 3;
      const float* p1 = p0 + stride;
      const float* p2 = p1 + stride;
      double a[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) a[k] = (double)p0[k] + (double)p2[k] - 2.0 * (double)p1[k];
      d = sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]);
    }
    d = warp_sum(d);
    acc += d / (double)UMT_N_LANDMARKS;
    cnt += 1.0;
  }
  if (lane == 0) {
    sAcc[warp] = acc;
    sCnt[warp] = cnt;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double e = 0.0, c = 0.0;
    for (int i = 0; i < 8; ++i) {
      e += sAcc[i];
      c += sCnt[i];
    }
    if (c > 0.0) {
      atomicAdd(&sums[0], e);
      atomicAdd(&sums[1], c);
    }
  }
}

}  // namespace

extern "C" int umt_fk_landmarks(const float* axes, const float* rest, const float* lm_rest,
                                const float* lm_idx, const float* lm_w, const float* joint_angles,
                                const float* wrist, const uint8_t* mirror, int n, float* out,
                                void* stream) {
  if (n < 0) return umt::fail(UMT_ERR_INVALID, "umt_fk_landmarks: n=%d", n);
  if (n == 0) return UMT_OK;
  if (!axes || !rest || !lm_rest || !lm_idx || !lm_w || !joint_angles || !wrist || !out)
    return umt::fail(UMT_ERR_INVALID, "umt_fk_landmarks: null pointer");
  const int blocks = (n + kWarpsPerBlock - 1) / kWarpsPerBlock;
  fk_kernel<<<blocks, kWarpsPerBlock * 32, 0, (cudaStream_t)stream>>>(
      axes, rest, lm_rest, lm_idx, lm_w, joint_angles, wrist, mirror, n, out);
  return umt::check_launch("fk_kernel");
}

// err[i] = mean over 21 landmarks of |gt - pred|, float64, in numpy's evaluation order:
// norm = sqrt((dx*dx + dy*dy) + dz*dz); mean = pairwise_sum(21 values) / 21 where pairwise_sum
// keeps eight partial sums r[j] = a[j] + a[8+j], combines them as ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7))
// and adds a[16..20] one by one.  No FMA contraction anywhere.  One warp per hand-frame.
__global__ void __launch_bounds__(256)
keypoint_error_kernel(const float* __restrict__ pred, const float* __restrict__ gt, int n,
                      double* __restrict__ err) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int pose = blockIdx.x * 8 + warp; pose < n; pose += gridDim.x * 8) {
    double d = 0.0;
    if (lane < UMT_N_LANDMARKS) {
      const size_t o = ((size_t)pose * UMT_N_LANDMARKS + lane) * 3;
      const double dx = __dsub_rn((double)gt[o], (double)pred[o]);
      const double dy = __dsub_rn((double)gt[o + 1], (double)pred[o + 1]);
      const double dz = __dsub_rn((double)gt[o + 2], (double)pred[o + 2]);
      d = __dsqrt_rn(__dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz)));
    }
    // r[j] for j < 8 lives in lane j
    double r = __dadd_rn(d, __shfl_down_sync(0xffffffffu, d, 8));
    const double s1 = __dadd_rn(r, __shfl_down_sync(0xffffffffu, r, 1));      // lanes 0,2,4,6: r[j]+r[j+1]
    const double s2 = __dadd_rn(s1, __shfl_down_sync(0xffffffffu, s1, 2));    // lanes 0,4
    double res = __dadd_rn(s2, __shfl_down_sync(0xffffffffu, s2, 4));         // lane 0
    for (int i = 16; i < UMT_N_LANDMARKS; ++i) res = __dadd_rn(res, __shfl_sync(0xffffffffu, d, i));
    if (lane == 0) err[pose] = __ddiv_rn(res, (double)UMT_N_LANDMARKS);
  }
}

// counts[k] += #{i valid : err[i] <= thr[k]}.  Each block histograms "first threshold that
// admits the error" in shared memory (binary search), then adds the running totals.
__global__ void __launch_bounds__(256)
pck_kernel(const double* __restrict__ err, const uint8_t* __restrict__ valid, long long n,
           const double* __restrict__ thr, int n_thr, double* __restrict__ counts) {
  __shared__ unsigned int hist[1025];
  __shared__ double sthr[1024];
  for (int i = threadIdx.x; i <= n_thr; i += blockDim.x) hist[i] = 0;
  for (int i = threadIdx.x; i < n_thr; i += blockDim.x) sthr[i] = thr[i];
  __syncthreads();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    if (valid && !valid[i]) continue;
    const double e = err[i];
    int lo = 0, hi = n_thr;   // first k in [0, n_thr] with e <= thr[k] (n_thr = none; NaN -> none)
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (e <= sthr[mid]) hi = mid; else lo = mid + 1;
    }
    atomicAdd(&hist[lo], 1u);
  }
  __syncthreads();
  for (int k = threadIdx.x; k < n_thr; k += blockDim.x) {
    unsigned long long c = 0;
    for (int j = 0; j <= k; ++j) c += hist[j];
    if (c) atomicAdd(&counts[k], (double)c);
  }
}

extern "C" int umt_mpjpe_accumulate(const float* pred, const float* gt, const uint8_t* valid,
                                    int n, double* sums, void* stream) {
  if (n < 0) return umt::fail(UMT_ERR_INVALID, "umt_mpjpe_accumulate: n=%d", n);
  if (n == 0) return UMT_OK;
  if (!pred || !gt || !sums) return umt::fail(UMT_ERR_INVALID, "umt_mpjpe_accumulate: null pointer");
  int blocks = (n + 7) / 8;
  if (blocks > 148 * 8) blocks = 148 * 8;
  mpjpe_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(pred, gt, valid, n, sums);
  return umt::check_launch("mpjpe_kernel");
}

extern "C" int umt_accel_accumulate(const float* pts, const uint8_t* valid, int n_seq, int T,
                                    double* sums, void* stream) {
  if (n_seq < 0 || T < 0) return umt::fail(UMT_ERR_INVALID, "umt_accel_accumulate: bad sizes");
  if (n_seq == 0 || T < 3) return UMT_OK;
  if (!pts || !valid || !sums) return umt::fail(UMT_ERR_INVALID, "umt_accel_accumulate: null pointer");
  long long total = (long long)n_seq * (T - 2);
  long long blocks = (total + 7) / 8;
  if (blocks > 148 * 8) blocks = 148 * 8;
  accel_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(pts, valid, n_seq, T, sums);
  return umt::check_launch("accel_kernel");
}

extern "C" int umt_keypoint_errors(const float* pred, const float* gt, int n, double* err, void* stream) {
  if (n < 0) return umt::fail(UMT_ERR_INVALID, "umt_keypoint_errors: n=%d", n);
  if (n == 0) return UMT_OK;
  if (!pred || !gt || !err) return umt::fail(UMT_ERR_INVALID, "umt_keypoint_errors: null pointer");
  int blocks = (n + 7) / 8;
  if (blocks > 148 * 8) blocks = 148 * 8;
  keypoint_error_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(pred, gt, n, err);
  return umt::check_launch("keypoint_error_kernel");
}

extern "C" int umt_pck_accumulate(const double* err, const uint8_t* valid, long long n,
                                  const double* thresholds, int n_thresholds, double* counts, void* stream) {
  if (n < 0 || n_thresholds < 1 || n_thresholds > 1024)
    return umt::fail(UMT_ERR_INVALID, "umt_pck_accumulate: n=%lld, n_thresholds=%d (1..1024)", n, n_thresholds);
  if (n == 0) return UMT_OK;
  if (!err || !thresholds || !counts) return umt::fail(UMT_ERR_INVALID, "umt_pck_accumulate: null pointer");
  long long blocks = (n + 255) / 256;
  if (blocks > 148 * 4) blocks = 148 * 4;
  pck_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(err, valid, n, thresholds, n_thresholds, counts);
  return umt::check_launch("pck_kernel");
}
