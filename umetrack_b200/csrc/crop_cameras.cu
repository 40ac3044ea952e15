// Batched crop-camera generation on the device (float64), the host loop of the reference's
// HandTracker.gen_crop_cameras made data parallel:
//   rank_hand_visibility_in_cameras      lib/tracker/perspective_crop.py:54-86
//   gen_crop_cameras_from_pose           lib/tracker/perspective_crop.py:118-162 (sort_camera_index=True,
//                                        max_view_num=2 as HandTracker uses it, tracker.py:116-129)
//   gen_crop_parameters_from_points      lib/common/crop.py:15-82
//   make_look_at_matrix / from_two_vectors lib/common/affine.py:34-66
// One thread per (frame, hand) item.  Inputs are the 63 float32 crop points of the item (FK of the
// given, neutral and open pose; the first 21 are the landmarks used for visibility) and the
// recording's cameras.  Outputs are everything the next stages need: the 48-double parameter rows
// of umt_resample_crops, and the model's float32 intrinsics / extrinsics (metres).
#include "common.cuh"

namespace {

struct M4 {
  double m[16];
};

__device__ bool inv4(const double* a, double* out) {
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
      if (fabs(m[r][c]) > best) { best = fabs(m[r][c]); piv = r; }
    if (best < 1e-300) return false;
    if (piv != c)
      for (int j = 0; j < 8; ++j) { double t = m[c][j]; m[c][j] = m[piv][j]; m[piv][j] = t; }
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
    for (int j = 0; j < 4; ++j) out[i * 4 + j] = m[i][4 + j];
  return true;
}

__device__ void mul3(const double* a, const double* b, double* c) {  // 3x3 row major
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) c[i * 3 + j] = a[i * 3] * b[j] + a[i * 3 + 1] * b[3 + j] + a[i * 3 + 2] * b[6 + j];
}

// FishEye62 window coordinates of an eye-space point (lib/common/camera.py:79-85,120-141,269-273)
__device__ void fisheye_window(const double* K /* fx fy cx cy k1 k2 k3 k4 p1 p2 k5 k6 */, double ex,
                               double ey, double ez, double& wx, double& wy) {
  const double r = sqrt(ex * ex + ey * ey);
  const double s = atan2(r, ez) / fmax(r, 2.938735877055719e-39);
  const double ax = ex * s, ay = ey * s;
  double r2 = ax * ax + ay * ay;
  const double pi2 = 9.869604401089358;
  r2 = fmin(fmax(r2, -pi2), pi2);
  const double r4 = r2 * r2, r6 = r2 * r4;
  const double radial = 1 + K[4] * r2 + K[5] * r4 + K[6] * r6 + K[7] * (r4 * r4) + K[10] * (r4 * r6) + K[11] * (r6 * r6);
  const double x = ax * radial, y = ay * radial;
  const double xy = x * y, x2 = x * x, y2 = y * y, rr = x2 + y2;
  const double xo = x + (2 * K[9] * xy + K[8] * (rr + 2 * x2));
  const double yo = y + (2 * K[8] * xy + K[9] * (rr + 2 * y2));
  wx = xo * K[0] + K[2];
  wy = yo * K[1] + K[3];
}

__global__ void __launch_bounds__(64)
crop_camera_kernel(const float* __restrict__ crop_points /* [n,63,3] */, const int32_t* __restrict__ frame_idx,
                   const uint8_t* __restrict__ hand_is_right, const uint8_t* __restrict__ valid,
                   const double* __restrict__ cam_intr /* [n_cam,12] */, const double* __restrict__ cam_to_world /* [n_frames,n_cam,16] */,
                   const double* __restrict__ cam_angles_deg, int n_cam, int cam_w, int cam_h, int n,
                   int n_points, int min_vis, double focal_multiplier, int crop_size,
                   int32_t* __restrict__ sel_count, int32_t* __restrict__ sel_cam /* [n,2] */,
                   double* __restrict__ resample_params /* [n,2,48] */, float* __restrict__ intr_out /* [n,2,9] */,
                   float* __restrict__ extr_out /* [n,2,16] */, int32_t* __restrict__ error_flag) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  sel_count[i] = 0;
  sel_cam[2 * i] = sel_cam[2 * i + 1] = -1;
  if (valid && !valid[i]) return;
  const float* pts = crop_points + (size_t)i * n_points * 3;
  const int f = frame_idx[i];
  // crop centre from float32 min/max, like numpy on the float32 landmark array (crop.py:59)
  float mn[3] = {pts[0], pts[1], pts[2]}, mx[3] = {pts[0], pts[1], pts[2]};
  for (int p = 1; p < n_points; ++p)
    for (int k = 0; k < 3; ++k) {
      mn[k] = fminf(mn[k], pts[p * 3 + k]);
      mx[k] = fmaxf(mx[k], pts[p * 3 + k]);
    }
  const float centre_f[3] = {(mn[0] + mx[0]) / 2.0f, (mn[1] + mx[1]) / 2.0f, (mn[2] + mx[2]) / 2.0f};
  int n_sel = 0;
  for (int c = 0; c < n_cam && n_sel < 2; ++c) {
    const double* K = cam_intr + (size_t)c * 12;
    const double* C2W = cam_to_world + ((size_t)f * n_cam + c) * 16;
    // visibility of the 21 landmarks of the given pose (perspective_crop.py:61-80)
    int vis = 0;
    for (int p = 0; p < 21; ++p) {
      const double dx = (double)pts[p * 3] - C2W[3], dy = (double)pts[p * 3 + 1] - C2W[7], dz = (double)pts[p * 3 + 2] - C2W[11];
      const double ex = dx * C2W[0] + dy * C2W[4] + dz * C2W[8];
      const double ey = dx * C2W[1] + dy * C2W[5] + dz * C2W[9];
      const double ez = dx * C2W[2] + dy * C2W[6] + dz * C2W[10];
      double wx, wy;
      fisheye_window(K, ex, ey, ez, wx, wy);
      if (wx >= 0 && wx <= cam_w - 1 && wy >= 0 && wy <= cam_h - 1 && ez > 0) ++vis;
    }
    if (vis < min_vis) continue;
    // look-at camera at the same position (affine.py:47-66)
    double w2e[16], e2w[16];
    if (!inv4(C2W, w2e) || !inv4(w2e, e2w)) { atomicExch(error_flag, 2); return; }
    const double cx = centre_f[0], cy = centre_f[1], cz = centre_f[2];
    double lx = w2e[0] * cx + w2e[1] * cy + w2e[2] * cz + w2e[3];
    double ly = w2e[4] * cx + w2e[5] * cy + w2e[6] * cz + w2e[7];
    double lz = w2e[8] * cx + w2e[9] * cy + w2e[10] * cz + w2e[11];
    const double ln = sqrt(lx * lx + ly * ly + lz * lz);
    lx /= ln; ly /= ln; lz /= ln;
    const double bn = fmax(5.43e-20, sqrt(lx * lx + ly * ly + lz * lz));
    const double bx = lx / bn, by = ly / bn, bz = lz / bn;
    const double vx = -by, vy = bx;              // cross((0,0,1), b)
    const double s = sqrt(vx * vx + vy * vy), cth = bz;
    const double Vm[9] = {0, 0, vy, 0, 0, -vx, -vy, vx, 0};   // skew(v), v = (vx, vy, 0)
    double V2[9];
    mul3(Vm, Vm, V2);
    const double kf = (1 - cth) / fmax(s * s, 1e-15);
    double dR[9];
    for (int k = 0; k < 9; ++k) dR[k] = ((k % 4 == 0) ? 1.0 : 0.0) + Vm[k] + V2[k] * kf;
    double R0[9] = {e2w[0], e2w[1], e2w[2], e2w[4], e2w[5], e2w[6], e2w[8], e2w[9], e2w[10]}, R1[9], R2[9];
    mul3(R0, dR, R1);
    const double ang = cam_angles_deg[c] * 0.017453292519943295;
    const double Rz[9] = {cos(ang), -sin(ang), 0, sin(ang), cos(ang), 0, 0, 0, 1};
    mul3(R1, Rz, R2);
    double ne2w[16] = {R2[0], R2[1], R2[2], e2w[3], R2[3], R2[4], R2[5], e2w[7], R2[6], R2[7], R2[8], e2w[11], 0, 0, 0, 1};
    double nw2e[16];
    if (!inv4(ne2w, nw2e)) { atomicExch(error_flag, 2); return; }
    if (hand_is_right[i])
      for (int k = 0; k < 4; ++k) nw2e[k] = -nw2e[k];      // mirrorx @ new_world_to_eye (crop.py:63-66)
    // focal length from the bounding points (crop.py:15-28)
    double max_ndc = 0.0;
    bool bad = false;
    for (int p = 0; p < n_points; ++p) {
      const double px = pts[p * 3], py = pts[p * 3 + 1], pz = pts[p * 3 + 2];
      const double ex = nw2e[0] * px + nw2e[1] * py + nw2e[2] * pz + nw2e[3];
      const double ey = nw2e[4] * px + nw2e[5] * py + nw2e[6] * pz + nw2e[7];
      const double ez = nw2e[8] * px + nw2e[9] * py + nw2e[10] * pz + nw2e[11];
      if (ez < 0.0001) bad = true;
      max_ndc = fmax(max_ndc, fmax(fabs(ex / ez), fabs(ey / ez)));
    }
    const double cxy = (crop_size - 1) / 2.0;
    const double fxy = cxy / max_ndc;
    if (bad || fxy < 5.0) { atomicExch(error_flag, 1); return; }   // ValueError("Unable to create crop camera")
    const double fo = focal_multiplier * fxy;
    double crop_c2w[16];
    if (!inv4(nw2e, crop_c2w)) { atomicExch(error_flag, 2); return; }
    // outputs
    double* P = resample_params + ((size_t)i * 2 + n_sel) * 48;
    for (int k = 0; k < 12; ++k) P[k] = K[k];
    for (int k = 0; k < 16; ++k) P[12 + k] = C2W[k];
    P[28] = fo; P[29] = fo; P[30] = cxy; P[31] = cxy;
    for (int k = 0; k < 16; ++k) P[32 + k] = crop_c2w[k];
    float* I = intr_out + ((size_t)i * 2 + n_sel) * 9;
    I[0] = (float)fo; I[1] = 0; I[2] = (float)cxy; I[3] = 0; I[4] = (float)fo; I[5] = (float)cxy; I[6] = 0; I[7] = 0; I[8] = 1;
    // world -> crop eye with the translation in metres (tracker.py:214-216): inv(crop_c2w)
    double cw2e[16];
    if (!inv4(crop_c2w, cw2e)) { atomicExch(error_flag, 2); return; }
    float* E = extr_out + ((size_t)i * 2 + n_sel) * 16;
    for (int k = 0; k < 16; ++k) E[k] = (float)((k % 4 == 3 && k < 12) ? cw2e[k] * 0.001 : cw2e[k]);
    sel_cam[2 * i + n_sel] = c;
    ++n_sel;
  }
  sel_count[i] = n_sel;
}

}  // namespace

extern "C" int umt_crop_cameras(const float* crop_points, const int32_t* frame_idx, const uint8_t* hand_is_right,
                                const uint8_t* valid, const double* cam_intr, const double* cam_to_world,
                                const double* cam_angles_deg, int n_cam, int cam_w, int cam_h, int n, int n_points,
                                int min_vis, double focal_multiplier, int32_t* sel_count, int32_t* sel_cam,
                                double* resample_params, float* intr_out, float* extr_out, int32_t* error_flag,
                                void* stream) {
  if (n < 0 || n_cam <= 0 || n_points < 21 || n_points > 63 || cam_w <= 0 || cam_h <= 0)
    return umt::fail(UMT_ERR_INVALID, "umt_crop_cameras: bad sizes (n=%d n_cam=%d n_points=%d)", n, n_cam, n_points);
  if (n == 0) return UMT_OK;
  if (!crop_points || !frame_idx || !hand_is_right || !cam_intr || !cam_to_world || !cam_angles_deg || !sel_count ||
      !sel_cam || !resample_params || !intr_out || !extr_out || !error_flag)
    return umt::fail(UMT_ERR_INVALID, "umt_crop_cameras: null pointer");
  crop_camera_kernel<<<(n + 63) / 64, 64, 0, (cudaStream_t)stream>>>(
      crop_points, frame_idx, hand_is_right, valid, cam_intr, cam_to_world, cam_angles_deg, n_cam, cam_w, cam_h, n,
      n_points, min_vis, focal_multiplier, UMT_CROP, sel_count, sel_cam, resample_params, intr_out, extr_out,
      error_flag);
  return umt::check_launch("crop_camera_kernel");
}
