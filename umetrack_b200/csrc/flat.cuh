// Device helpers for the planar flat activation layout (see model.cuh): 16-byte rows of 8 bf16
// channels, values kept as hi (+ lo) bf16 pairs.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "model.cuh"

namespace umt {

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float bf16_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t u) { return __uint_as_float(u & 0xFFFF0000u); }

__device__ __forceinline__ uint32_t pack_f16x2(float a, float b) {
  __half2 t = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float2 unpack_f16x2(uint32_t u) {
  return __half22float2(*reinterpret_cast<__half2*>(&u));
}

// Numeric modes (UMT_MODE_*): 0 = split bf16 (hi + lo planes), 1 = bf16 (hi plane only),
// 2 = fp16 (one plane).  Store 8 floats as one 16-byte row (two rows in mode 0).
template <int MODE>
__device__ __forceinline__ void store_row8(const FlatTensor& t, int plane, long long row,
                                           const float* v) {
  uint8_t* p = t.base + (long long)plane * t.plane_stride + (t.guard + row) * 16;
  uint4 h;
  if (MODE == 2) {
    h.x = pack_f16x2(v[0], v[1]);
    h.y = pack_f16x2(v[2], v[3]);
    h.z = pack_f16x2(v[4], v[5]);
    h.w = pack_f16x2(v[6], v[7]);
    *reinterpret_cast<uint4*>(p) = h;
    return;
  }
  h.x = pack_bf16x2(v[0], v[1]);
  h.y = pack_bf16x2(v[2], v[3]);
  h.z = pack_bf16x2(v[4], v[5]);
  h.w = pack_bf16x2(v[6], v[7]);
  *reinterpret_cast<uint4*>(p) = h;
  if (MODE == 0) {
    uint4 l;
    l.x = pack_bf16x2(v[0] - bf16_lo(h.x), v[1] - bf16_hi(h.x));
    l.y = pack_bf16x2(v[2] - bf16_lo(h.y), v[3] - bf16_hi(h.y));
    l.z = pack_bf16x2(v[4] - bf16_lo(h.z), v[5] - bf16_hi(h.z));
    l.w = pack_bf16x2(v[6] - bf16_lo(h.w), v[7] - bf16_hi(h.w));
    *reinterpret_cast<uint4*>(p + t.lo_offset) = l;
  }
}

template <int MODE>
__device__ __forceinline__ void load_row8(const FlatTensor& t, int plane, long long row, float* v) {
  const uint8_t* p = t.base + (long long)plane * t.plane_stride + (t.guard + row) * 16;
  uint4 h = *reinterpret_cast<const uint4*>(p);
  if (MODE == 2) {
    float2 a = unpack_f16x2(h.x), b = unpack_f16x2(h.y), c = unpack_f16x2(h.z), d = unpack_f16x2(h.w);
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
    return;
  }
  v[0] = bf16_lo(h.x); v[1] = bf16_hi(h.x); v[2] = bf16_lo(h.y); v[3] = bf16_hi(h.y);
  v[4] = bf16_lo(h.z); v[5] = bf16_hi(h.z); v[6] = bf16_lo(h.w); v[7] = bf16_hi(h.w);
  if (MODE == 0) {
    uint4 l = *reinterpret_cast<const uint4*>(p + t.lo_offset);
    v[0] += bf16_lo(l.x); v[1] += bf16_hi(l.x); v[2] += bf16_lo(l.y); v[3] += bf16_hi(l.y);
    v[4] += bf16_lo(l.z); v[5] += bf16_hi(l.z); v[6] += bf16_lo(l.w); v[7] += bf16_hi(l.w);
  }
}

// decode one already-loaded row (hi [+ lo]) into 8 floats
template <int MODE>
__device__ __forceinline__ void decode_row8(const uint4& h, const uint4& l, float* v) {
  if (MODE == 2) {
    float2 a = unpack_f16x2(h.x), b = unpack_f16x2(h.y), c = unpack_f16x2(h.z), d = unpack_f16x2(h.w);
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
    return;
  }
  v[0] = bf16_lo(h.x); v[1] = bf16_hi(h.x); v[2] = bf16_lo(h.y); v[3] = bf16_hi(h.y);
  v[4] = bf16_lo(h.z); v[5] = bf16_hi(h.z); v[6] = bf16_lo(h.w); v[7] = bf16_hi(h.w);
  if (MODE == 0) {
    v[0] += bf16_lo(l.x); v[1] += bf16_hi(l.x); v[2] += bf16_lo(l.y); v[3] += bf16_hi(l.y);
    v[4] += bf16_lo(l.z); v[5] += bf16_hi(l.z); v[6] += bf16_lo(l.w); v[7] += bf16_hi(l.w);
  }
}

// one element of a stored row (test hooks)
__device__ __forceinline__ float load_elem(const FlatTensor& t, int mode, int c, long long row) {
  const uint8_t* p = t.base + (long long)(c >> 3) * t.plane_stride + (t.guard + row) * 16;
  if (mode == 2) return __half2float(reinterpret_cast<const __half*>(p)[c & 7]);
  float v = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[c & 7]);
  if (mode == 0) v += __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p + t.lo_offset)[c & 7]);
  return v;
}

}  // namespace umt
