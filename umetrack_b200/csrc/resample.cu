// Stage (a): fisheye -> pinhole crop resampling (see include/umetrack_b200.h).
//
// One CTA produces one 32x32 tile of one 96x96 crop; each of its 256 threads owns four
// horizontally adjacent output pixels, so the float32 crop is written with 16-byte stores that
// a warp coalesces into four full 128-byte rows.  Per pixel the thread evaluates the
// reference's float64 camera chain (lib/tracker/tracker.py:61-75, lib/common/camera.py) with
// the same operation order and no fused multiply-adds where numpy has none, casts to float32
// and quantises to 1/32 pixel exactly like cv2.remap's 8-bit bilinear path.  The CTA then
// reduces the bounding box of all taps it needs, stages that source window in shared memory
// with coalesced byte loads (zero filled outside the image = BORDER_CONSTANT 0) and gathers the
// four taps from there; windows that do not fit fall back to bounds-checked global gathers.
#include "common.cuh"

namespace {

constexpr int kCrop = UMT_CROP;
constexpr int kTile = 32;
constexpr int kThreads = 256;
constexpr int kStageBytes = 40 * 1024;  // shared window capacity

struct PixelCoord {
  int ix, iy, fx, fy;
};

__device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }

// Float64 map for one destination pixel; returns float32 window coordinates in the source.
__device__ __forceinline__ void source_coord(const double* __restrict__ P, int px, int py,
                                             float& mx, float& my) {
  const double sfx = P[0], sfy = P[1], scx = P[2], scy = P[3];
  const double k1 = P[4], k2 = P[5], k3 = P[6], k4 = P[7], p1 = P[8], p2 = P[9], k5 = P[10],
               k6 = P[11];
  const double* S = P + 12;  // source camera_to_world, row major
  const double cfx = P[28], cfy = P[29], ccx = P[30], ccy = P[31];
  const double* C = P + 32;  // crop camera_to_world

  // window_to_eye: (w - c) / f, append 1, normalise (camera.py:275-281,68-75; affine.py:22-24)
  double qx = __ddiv_rn(add((double)px, -ccx), cfx);
  double qy = __ddiv_rn(add((double)py, -ccy), cfy);
  double nn = add(add(mul(qx, qx), mul(qy, qy)), 1.0);
  double d = fmax(5.43e-20, __dsqrt_rn(nn));
  double v0 = __ddiv_rn(qx, d), v1 = __ddiv_rn(qy, d), v2 = __ddiv_rn(1.0, d);
  // eye_to_world: v @ R^T + t (camera.py:263-267, affine.py:10-19)
  double w0 = add(__fma_rn(v2, C[2], __fma_rn(v1, C[1], mul(v0, C[0]))), C[3]);
  double w1 = add(__fma_rn(v2, C[6], __fma_rn(v1, C[5], mul(v0, C[4]))), C[7]);
  double w2 = add(__fma_rn(v2, C[10], __fma_rn(v1, C[9], mul(v0, C[8]))), C[11]);
  // world_to_eye: (w - t) @ R (camera.py:257-261)
  double u0 = add(w0, -S[3]), u1 = add(w1, -S[7]), u2 = add(w2, -S[11]);
  double ex = __fma_rn(u2, S[8], __fma_rn(u1, S[4], mul(u0, S[0])));
  double ey = __fma_rn(u2, S[9], __fma_rn(u1, S[5], mul(u0, S[1])));
  double ez = __fma_rn(u2, S[10], __fma_rn(u1, S[6], mul(u0, S[2])));
  // arctan projection (camera.py:79-85)
  double r = __dsqrt_rn(add(mul(ex, ex), mul(ey, ey)));
  double s = __ddiv_rn(atan2(r, ez), fmax(r, 2.938735877055719e-39 /* 2^-128 */));
  double ax = mul(ex, s), ay = mul(ey, s);
  // Fisheye62 distortion (camera.py:120-141)
  const double pi2 = 9.869604401089358;
  double r2 = add(mul(ax, ax), mul(ay, ay));
  r2 = fmin(fmax(r2, -pi2), pi2);
  double r4 = mul(r2, r2), r6 = mul(r2, r4), r8 = mul(r4, r4), r10 = mul(r4, r6),
         r12 = mul(r6, r6);
  double radial = add(add(add(add(add(add(1.0, mul(k1, r2)), mul(k2, r4)), mul(k3, r6)),
                              mul(k4, r8)), mul(k5, r10)), mul(k6, r12));
  double x = mul(ax, radial), y = mul(ay, radial);
  double x2 = mul(x, x), y2 = mul(y, y), xy = mul(x, y), rr = add(x2, y2);
  double xo = add(x, add(mul(mul(2.0, p2), xy), mul(p1, add(rr, mul(2.0, x2)))));
  double yo = add(y, add(mul(mul(2.0, p1), xy), mul(p2, add(rr, mul(2.0, y2)))));
  double wx = add(mul(xo, sfx), scx);
  double wy = add(mul(yo, sfy), scy);
  if (ez < 0.0) {  // tracker.py:71-73
    wx = -1.0;
    wy = -1.0;
  }
  mx = __double2float_rn(wx);
  my = __double2float_rn(wy);
}

__device__ __forceinline__ int sat16(int v) { return max(-32768, min(32767, v)); }

// cv2.remap fixed point: cvRound(m * 32) with round-half-even, split into tap origin + fraction
__device__ __forceinline__ PixelCoord quantise(float mx, float my) {
  PixelCoord c;
  float ax = mx * 32.0f, ay = my * 32.0f;
  // non-finite or out-of-int-range maps behave like x86 cvtss2si ("integer indefinite")
  int sx = (fabsf(ax) < 2147483648.0f) ? __float2int_rn(ax) : INT_MIN;
  int sy = (fabsf(ay) < 2147483648.0f) ? __float2int_rn(ay) : INT_MIN;
  c.ix = sat16(sx >> 5);
  c.iy = sat16(sy >> 5);
  c.fx = sx & 31;
  c.fy = sy & 31;
  return c;
}

// four CTAs per SM (64 registers): the float64 chain is latency bound, so occupancy -- not
// per-thread ILP -- sets the rate (1 CTA/SM at 139 registers ran 2.3x slower)
__global__ void __launch_bounds__(kThreads, 4)
resample_kernel(const uint8_t* __restrict__ src_base, const int64_t* __restrict__ src_offset,
                int src_pitch, int src_w, int src_h, const double* __restrict__ params,
                uint8_t* __restrict__ out_u8, float* __restrict__ out_f32,
                int32_t* __restrict__ out_idx) {
  __shared__ double sP[UMT_RESAMPLE_PARAMS];
  __shared__ int sBox[4];
  extern __shared__ uint8_t sWin[];

  const int crop = blockIdx.x / 9;
  const int tile = blockIdx.x % 9;
  const int tx0 = (tile % 3) * kTile, ty0 = (tile / 3) * kTile;
  const int tid = threadIdx.x;
  if (tid < UMT_RESAMPLE_PARAMS) sP[tid] = params[(size_t)crop * UMT_RESAMPLE_PARAMS + tid];
  if (tid < 4) sBox[tid] = (tid < 2) ? INT_MAX : INT_MIN;
  __syncthreads();

  const int px0 = tx0 + (tid & 7) * 4;
  const int py = ty0 + (tid >> 3);
  PixelCoord pc[4];
  int bx0 = INT_MAX, by0 = INT_MAX, bx1 = INT_MIN, by1 = INT_MIN;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    float mx, my;
    source_coord(sP, px0 + k, py, mx, my);
    pc[k] = quantise(mx, my);
    // taps (ix..ix+1, iy..iy+1); only pixels with at least one tap inside the image matter
    bool touches = pc[k].ix + 1 >= 0 && pc[k].ix < src_w && pc[k].iy + 1 >= 0 && pc[k].iy < src_h;
    if (touches) {
      bx0 = min(bx0, pc[k].ix);
      bx1 = max(bx1, pc[k].ix + 1);
      by0 = min(by0, pc[k].iy);
      by1 = max(by1, pc[k].iy + 1);
    }
  }
  // block-wide bounding box of the taps
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    bx0 = min(bx0, __shfl_xor_sync(0xffffffffu, bx0, o));
    by0 = min(by0, __shfl_xor_sync(0xffffffffu, by0, o));
    bx1 = max(bx1, __shfl_xor_sync(0xffffffffu, bx1, o));
    by1 = max(by1, __shfl_xor_sync(0xffffffffu, by1, o));
  }
  if ((tid & 31) == 0) {
    atomicMin(&sBox[0], bx0);
    atomicMin(&sBox[1], by0);
    atomicMax(&sBox[2], bx1);
    atomicMax(&sBox[3], by1);
  }
  __syncthreads();
  bx0 = sBox[0];
  by0 = sBox[1];
  bx1 = sBox[2];
  by1 = sBox[3];
  const uint8_t* __restrict__ img = src_base + src_offset[crop];
  const bool any = bx1 >= bx0;
  const int bw = any ? (bx1 - bx0 + 1) : 0, bh = any ? (by1 - by0 + 1) : 0;
  const int pitch = (bw + 3) & ~3;
  const bool staged = any && (size_t)pitch * bh <= (size_t)kStageBytes;
  if (staged) {
    for (int i = tid; i < pitch * bh; i += kThreads) {
      int yy = by0 + i / pitch, xx = bx0 + i % pitch;
      uint8_t v = 0;
      if (xx >= 0 && xx < src_w && yy >= 0 && yy < src_h) v = __ldg(img + (size_t)yy * src_pitch + xx);
      sWin[i] = v;
    }
  }
  __syncthreads();

  uint8_t res[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int ix = pc[k].ix, iy = pc[k].iy, fx = pc[k].fx, fy = pc[k].fy;
    int acc = 0;
#pragma unroll
    for (int dy = 0; dy < 2; ++dy) {
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        const int xx = ix + dx, yy = iy + dy;
        int p = 0;
        if (xx >= 0 && xx < src_w && yy >= 0 && yy < src_h) {
          p = staged ? sWin[(yy - by0) * pitch + (xx - bx0)]
                     : __ldg(img + (size_t)yy * src_pitch + xx);
        }
        acc += p * (dx ? fx : 32 - fx) * (dy ? fy : 32 - fy);
      }
    }
    res[k] = (uint8_t)((acc + 512) >> 10);
  }
  const size_t o = (size_t)crop * kCrop * kCrop + (size_t)py * kCrop + px0;
  if (out_u8) *reinterpret_cast<uchar4*>(out_u8 + o) = make_uchar4(res[0], res[1], res[2], res[3]);
  if (out_f32) {
    // `crop.astype(float32) / 255.0` (tracker.py:211): IEEE division, not a reciprocal multiply
    float4 f = make_float4(__fdiv_rn((float)res[0], 255.0f), __fdiv_rn((float)res[1], 255.0f),
                           __fdiv_rn((float)res[2], 255.0f), __fdiv_rn((float)res[3], 255.0f));
    *reinterpret_cast<float4*>(out_f32 + o) = f;
  }
  if (out_idx) {
#pragma unroll
    for (int k = 0; k < 4; ++k)
      *reinterpret_cast<int4*>(out_idx + (o + k) * 4) = make_int4(pc[k].ix, pc[k].iy, pc[k].fx, pc[k].fy);
  }
}

// ------------------------------------------------------------------------------------------
// torch_data variant (lib/batched_dataset/data_transform.py:78-144): pinhole -> pinhole
// homography resampling of float32 images.  One thread per destination pixel, float4 stores over
// four adjacent pixels.  Coordinates are float64 like numpy's float32 x int32 promotion.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
homography_kernel(const float* __restrict__ src, int src_w, int src_h, const float* __restrict__ xfs,
                  int n, int out_w, int out_h, int divide, float* __restrict__ out) {
  const int groups_per_row = out_w >> 2;
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = (long long)n * out_h * groups_per_row;
  if (gid >= total) return;
  const int gx = (int)(gid % groups_per_row);
  const int v = (int)((gid / groups_per_row) % out_h);
  const int img = (int)(gid / ((long long)groups_per_row * out_h));
  const float* M = xfs + (size_t)img * 16;
  const float* im = src + (size_t)img * src_w * src_h;
  float res[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const double u = (double)(gx * 4 + k), vv = (double)v;
    // r[i,0]*u + r[i,1]*v + r[i,2]*1, then + t[i]
    const double X = add(add(add(mul((double)M[0], u), mul((double)M[1], vv)), (double)M[2]), (double)M[3]);
    const double Y = add(add(add(mul((double)M[4], u), mul((double)M[5], vv)), (double)M[6]), (double)M[7]);
    const double Z = add(add(add(mul((double)M[8], u), mul((double)M[9], vv)), (double)M[10]), (double)M[11]);
    const double x = __ddiv_rn(X, Z), y = __ddiv_rn(Y, Z);
    float r = 0.f;
    if (x >= 0.0 && x < (double)(src_w - 1) && y >= 0.0 && y < (double)(src_h - 1)) {
      const int x0 = (int)x, y0 = (int)y;   // truncation, like astype(int32)
      const double dx1 = (double)(x0 + 1) - x, dx0 = x - (double)x0;
      const double dy1 = (double)(y0 + 1) - y, dy0 = y - (double)y0;
      const double f00 = (double)__ldg(im + (size_t)y0 * src_w + x0), f10 = (double)__ldg(im + (size_t)y0 * src_w + x0 + 1);
      const double f01 = (double)__ldg(im + (size_t)(y0 + 1) * src_w + x0), f11 = (double)__ldg(im + (size_t)(y0 + 1) * src_w + x0 + 1);
      // fx_00*(x1-x)*(y1-y) + fx_10*(x-x0)*(y1-y) + fx_01*(x1-x)*(y-y0) + fx_11*(x-x0)*(y-y0)
      const double s = add(add(add(mul(mul(f00, dx1), dy1), mul(mul(f10, dx0), dy1)), mul(mul(f01, dx1), dy0)),
                           mul(mul(f11, dx0), dy0));
      r = __double2float_rn(s);
    }
    res[k] = divide ? __fdiv_rn(r, 255.0f) : r;
  }
  *reinterpret_cast<float4*>(out + ((size_t)img * out_h + v) * out_w + gx * 4) = make_float4(res[0], res[1], res[2], res[3]);
}

}  // namespace

extern "C" int umt_resample_crops(const uint8_t* src_base, const int64_t* src_offset,
                                  int src_pitch, int src_w, int src_h, const double* params,
                                  int n_crops, uint8_t* out_u8, float* out_f32, int32_t* out_idx,
                                  void* stream) {
  if (n_crops < 0 || src_w <= 0 || src_h <= 0 || src_pitch < src_w)
    return umt::fail(UMT_ERR_INVALID, "umt_resample_crops: bad geometry (n=%d w=%d h=%d pitch=%d)",
                     n_crops, src_w, src_h, src_pitch);
  if (n_crops == 0) return UMT_OK;
  if (!src_base || !src_offset || !params)
    return umt::fail(UMT_ERR_INVALID, "umt_resample_crops: null input pointer");
  {  // once per device: the attribute belongs to the device
    static bool done[64] = {};
    int dev = 0;
    UMT_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !done[dev]) {
      UMT_CUDA_TRY(cudaFuncSetAttribute(resample_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kStageBytes));
      if (dev >= 0 && dev < 64) done[dev] = true;
    }
  }
  resample_kernel<<<n_crops * 9, kThreads, kStageBytes, (cudaStream_t)stream>>>(
      src_base, src_offset, src_pitch, src_w, src_h, params, out_u8, out_f32, out_idx);
  return umt::check_launch("resample_kernel");
}

extern "C" int umt_resample_homography(const float* src, int src_w, int src_h, const float* xfs, int n,
                                       int out_w, int out_h, int divide_by_255, float* out, void* stream) {
  if (n < 0 || src_w < 2 || src_h < 2 || out_w <= 0 || out_h <= 0 || (out_w & 3))
    return umt::fail(UMT_ERR_INVALID, "umt_resample_homography: bad geometry (n=%d src %dx%d out %dx%d, out_w %% 4 must be 0)",
                     n, src_w, src_h, out_w, out_h);
  if (n == 0) return UMT_OK;
  if (!src || !xfs || !out) return umt::fail(UMT_ERR_INVALID, "umt_resample_homography: null pointer");
  const long long total = (long long)n * out_h * (out_w / 4);
  homography_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      src, src_w, src_h, xfs, n, out_w, out_h, divide_by_255, out);
  return umt::check_launch("homography_kernel");
}
