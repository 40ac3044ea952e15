// Device math shared by the tail kernels (tail.cu: per-stage kernels around the tensor-core convolutions;
// tail_fused.cu: the whole tail of one sample in one CTA): 4x4 transforms in double precision, the
// feature-transform layer, a 3x3 Jacobi SVD for the Procrustes step, the head decoders.
#pragma once
#include <cmath>

#include "model.cuh"

namespace umt {

constexpr int kPix = 36;
constexpr int kPos = 49;   // positions per sample in the 6x6 planar flat raster (7 x 7)

// ------------------------------------------------------------------------------------------
// small matrix helpers (double precision, per thread)
// ------------------------------------------------------------------------------------------
__device__ inline void mat4_mul(const double* a, const double* b, double* c) {
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      double s = 0;
      for (int k = 0; k < 4; ++k) s += a[i * 4 + k] * b[k * 4 + j];
      c[i * 4 + j] = s;
    }
}

// general 4x4 inverse by Gauss-Jordan with partial pivoting; returns false if singular
__device__ inline bool mat4_inv(const double* a, double* inv) {
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

__device__ inline void load_mat4(const float* p, double* m) {
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

__device__ inline void jacobi_svd3(const double* A, double* U, double* S, double* V) {
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

__device__ inline double det3(const double* m) {
  return m[0] * (m[4] * m[8] - m[5] * m[7]) - m[1] * (m[3] * m[8] - m[5] * m[6]) +
         m[2] * (m[3] * m[7] - m[4] * m[6]);
}


// Decoders of the regression head for sample b (regressor.py:76-121), Procrustes alignment of the 7 rigid
// points (model_utils.py:17-54), world recovery and right-hand un-mirroring (umetrack_model.py:77-90).
// Called by every thread of the CTA with t = its index (needs t = 0..64); sRaw = the n_out raw outputs.
__device__ inline void head_decode(int t, const float* sRaw, int b, int with_scale, const float* __restrict__ extrinsics,
                                   const int64_t* __restrict__ sample_range, const int64_t* __restrict__ hand_idx,
                                   float* __restrict__ out_angles, float* __restrict__ out_wrist,
                                   float* __restrict__ out_scale, float* __restrict__ out_sigma) {
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

}  // namespace umt
