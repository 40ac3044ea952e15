// Stage (b), part 1: the per-view CNN encoder (24 3x3 convolutions with 3 fused 1x1 downsample
// branches, 99 % of the FLOPs) as
// implicit-GEMM convolutions on the 5th-generation tensor cores (tcgen05.mma, accumulators in
// TMEM), plus the stem (conv 1->32 + BN + ReLU + 2x2 max-pool) and the 256->72 projection on CUDA
// cores.  Reference: lib/models/backbone_resnet.py:14-158, lib/models/model_utils.py:107-138.
//
// How a 3x3 convolution becomes a GEMM without materialising im2col
// -----------------------------------------------------------------
// Activations use the "planar flat" layout described in model.cuh: in that layout tap (dy,dx) of
// output position q is input position q + dy*(W+1) + dx for EVERY q, so the A operand of tap t is
// the same shared-memory patch viewed through a descriptor whose start address is moved by
// shift_t rows.  A CTA therefore loads each 16-channel slice of its input patch ONCE (one
// contiguous cp.async.bulk per 8-channel plane, no tensor map needed) and issues nine
// M=128 x N=Cout x K=16 MMAs per 128-row sub-tile against it.  The UMMA shared-memory layout is
// the no-swizzle K-major canonical layout: 8 rows x 16 bytes core matrices, SBO = 128 bytes
// between 8-row groups, LBO = one plane between the two 8-channel halves of a K=16 step; a row
// shift is then a plain +16-byte-per-row change of the descriptor start address.
// Stride-2 convolutions read four parity-phase copies of their input (written by the producing
// layer's epilogue), for which every tap is again a constant shift (0 or -1 rows/columns); the
// 1x1 stride-2 downsample branch of a block is folded into the same accumulator as extra K.
//
// Numerics
// --------
// UMT_MODE_FP32_EQUIV keeps every activation and weight as a bf16 pair (hi, lo = x - hi) and
// issues three products per K step, hi*hi + hi*lo + lo*hi, accumulating in fp32 in TMEM
// (relative error ~2^-17 per product).  UMT_MODE_BF16 issues only hi*hi.
//
// Warp roles (256 threads): warp 0 lane 0 = A-patch producer, warp 1 lane 0 = weight-slab
// producer, warp 2 = TMEM allocator + single-thread MMA issuer, warps 4..7 = epilogue
// (tcgen05.ld -> bias/residual/ReLU -> bf16 split -> coalesced 16-byte stores).
#include <cuda_bf16.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <string>

#include "model.cuh"
#include "flat.cuh"
#include "umma.cuh"

namespace umt {

static const int kStageHW[4] = {48, 24, 12, 6};
static const int kStageC[4] = {32, 64, 128, 256};
static const int kStageBlocks[4] = {2, 3, 5, 2};

// ------------------------------------------------------------------------------------------
// The implicit-GEMM convolution kernel
// ------------------------------------------------------------------------------------------
constexpr int kConvThreads = 256;   // 4 control warps + 4 epilogue warps


// Issue the MMAs of `n` consecutive taps (one weight stage) for all S sub-tiles.  Descriptors are
// advanced by adding to their low word: +shift rows for a tap, +128 (2048 B) per sub-tile, +slab
// bytes per tap on the weight side; everything else is loop invariant.
// Split mode, weight slab = [plane 0: hi rows, lo rows][plane 1: hi rows, lo rows], so the hi rows alone
// (N = COUT) and hi|lo together (N = 2 COUT) are the same descriptor with a different N.
//   MERGE (COUT <= 64): two MMAs per tap and sub-tile instead of three:
//       [D1 | D2] += a_hi * [w_hi | w_lo]   (N = 2 COUT)        D1 += a_lo * w_hi   (N = COUT)
//   and the epilogue adds D1 + D2.  The narrow layers are bound by the shared-memory operand path
//   (every MMA re-reads its 4 KB A tile), so one A read fewer per tap is 27 % fewer operand bytes.
template <int S, int COUT, bool SPLIT, bool MERGE>
__device__ __forceinline__ void issue_taps(const ConvParams& P, int stage, int t0, int n, uint32_t a_lo32,
                                           uint32_t a_hi32, uint32_t a_plane_bytes, uint32_t b_lo32,
                                           uint32_t b_hi32, uint32_t d_base, uint32_t idesc, uint32_t any,
                                           bool skip_hilo) {
  constexpr uint32_t kBHalf16 = (2u * COUT * 16u) >> 4;
  constexpr uint32_t kBSlab16 = SPLIT ? 2u * kBHalf16 : kBHalf16;
  constexpr uint32_t kLoRows16 = COUT;                           // lo rows follow the hi rows inside a plane (16-byte units)
  constexpr uint32_t kDCols = MERGE ? 2u * COUT : (uint32_t)COUT;   // accumulator columns per sub-tile
  constexpr uint32_t kIdescWide = (uint32_t)((2 * COUT) >> 3) << 17;
  constexpr uint32_t kIdescNMask = 0x3Fu << 17;
  const uint32_t a_lo_plane16 = (2u * a_plane_bytes) >> 4;
  const int tap0 = P.st_tap0[stage] + t0;
#pragma unroll 1
  for (int j = 0; j < n; ++j) {
    // rows == 16-byte units; merged stages: the tap's patch sits tap_src patches into the plane
    const uint32_t shift16 = (uint32_t)((int)P.tap_shift[tap0 + j] - P.lo_shift + (int)P.tap_src[tap0 + j] * P.patch_rows);
    const uint32_t a0 = a_lo32 + shift16;
    const uint64_t b_hi = desc64(b_lo32 + (uint32_t)j * kBSlab16, b_hi32);
    const uint64_t b_lo = desc64(b_lo32 + (uint32_t)j * kBSlab16 + kLoRows16, b_hi32);
#pragma unroll
    for (int s = 0; s < S; ++s) {
      const uint64_t a_hi = desc64(a0 + (uint32_t)s * 128u, a_hi32);
      const uint32_t d = d_base + (uint32_t)s * kDCols;
      if (MERGE) {
        const uint64_t a_lo = desc64(a0 + (uint32_t)s * 128u + a_lo_plane16, a_hi32);
        tc_mma_bf16(d, a_hi, b_hi, skip_hilo ? idesc : ((idesc & ~kIdescNMask) | kIdescWide), any);
        tc_mma_bf16(d, a_lo, b_hi, idesc, 1u);
      } else {
        tc_mma_bf16(d, a_hi, b_hi, idesc, any);
        if (SPLIT) {
          const uint64_t a_lo = desc64(a0 + (uint32_t)s * 128u + a_lo_plane16, a_hi32);
          if (!skip_hilo) tc_mma_bf16(d, a_hi, b_lo, idesc, 1u);
          tc_mma_bf16(d, a_lo, b_hi, idesc, 1u);
        }
      }
    }
    any = 1u;
  }
}

// Persistent, warp-specialised kernel.  Each CTA (one per SM) walks tiles t = blockIdx.x,
// blockIdx.x + gridDim.x, ...; a tile is S sub-tiles of 128 positions x all Cout channels.
// Pipelines: A patches (n_a slots) and weight slabs (n_b slots) keep streaming across tile
// boundaries; the accumulator lives in n_buf TMEM buffers so the epilogue of tile i overlaps the
// MMAs of tile i+1.  With a cluster of `cluster` CTAs the weight slabs -- identical for every
// tile -- are fetched from L2 once per cluster: each CTA loads 1/cluster of a slab and multicasts
// it into the shared memory of all CTAs of the cluster.
// EPI selects the epilogue (kept out of the loop body at compile time: the epilogue warps run one
// warp per scheduler, so every instruction they fetch is on the path that frees the accumulator):
// 0 = bias/ReLU -> planar rows, 1 = + identity residual, 2 = float32 NCHW output (projection).
template <int COUT, int MODE, int EPI>
__global__ void __launch_bounds__(kConvThreads, 2)
conv_umma_kernel(const __grid_constant__ ConvParams P) {
  constexpr bool SPLIT = (MODE == UMT_MODE_FP32_EQUIV);
  constexpr bool MERGE = SPLIT && COUT <= 64;               // see issue_taps (backbone widths 32 and 64)
  constexpr int kDCols = MERGE ? 2 * COUT : COUT;           // accumulator columns per sub-tile
  extern __shared__ __align__(128) uint8_t smem[];
  // [0,1024): barriers + tmem pointer; [1024, 1024+COUT*4): bias; then A stages, then B stages
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + 1000);
  float* s_bias = reinterpret_cast<float*>(smem + 1024);
  constexpr int kBiasBytes = ((COUT * 4 + 127) / 128) * 128;
  const int n_a = P.n_a, n_b = P.n_b, n_buf = P.n_buf;
  const uint32_t a_patch_bytes = (uint32_t)P.patch_rows * 16u;          // one source's patch in one plane
  const uint32_t a_plane_bytes = a_patch_bytes * (uint32_t)P.a_nsrc;    // merged stages: a_nsrc patches per plane
  const uint32_t a_stage_bytes = a_plane_bytes * (SPLIT ? 4u : 2u);
  const uint32_t a_stage_stride = (a_stage_bytes + 127u) & ~127u;
  constexpr uint32_t kBHalf = 2u * COUT * 16u;            // one of {w_hi, w_lo}: 2 planes
  constexpr uint32_t kBSlab = SPLIT ? 2u * kBHalf : kBHalf;   // bytes of one (stage, tap) slab in this mode
  const int G = P.bgroup;                                    // slabs per B pipeline stage
  const uint32_t b_stage_bytes = (uint32_t)G * kBSlab;
  uint8_t* a_base = smem + 1024 + kBiasBytes;
  uint8_t* b_base = a_base + (size_t)n_a * a_stage_stride;

  const uint32_t bar0 = smem_u32(bars);
  auto a_full = [&](int i) { return bar0 + 8u * (uint32_t)i; };
  auto a_empty = [&](int i) { return bar0 + 8u * (uint32_t)(n_a + i); };
  auto b_full = [&](int i) { return bar0 + 8u * (uint32_t)(2 * n_a + i); };
  auto b_empty = [&](int i) { return bar0 + 8u * (uint32_t)(2 * n_a + n_b + i); };
  auto acc_full = [&](int i) { return bar0 + 8u * (uint32_t)(2 * n_a + 2 * n_b + i); };
  auto acc_empty = [&](int i) { return bar0 + 8u * (uint32_t)(2 * n_a + 2 * n_b + 2 + i); };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cl = (uint32_t)P.cluster;
  uint32_t cta_rank = 0;
  if (cl > 1) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(cta_rank));
  // output-channel slices (ConvParams::n_slices): this CTA's slice, and its place among the CTAs that walk tiles
  const int n_sl = P.n_slices > 1 ? P.n_slices : 1;
  const int slice = n_sl > 1 ? (int)blockIdx.x % n_sl : 0;
  const int vblock = n_sl > 1 ? (int)blockIdx.x / n_sl : (int)blockIdx.x;
  const int vgrid = n_sl > 1 ? (int)gridDim.x / n_sl : (int)gridDim.x;
  const uint8_t* const weights = n_sl > 1 ? P.slice_weights[slice] : P.weights;
  const float* const bias = n_sl > 1 ? P.slice_bias[slice] : P.bias;
  const int pl0 = slice * (COUT / 8);      // first output / residual plane of the slice
  const uint16_t cta_mask = (uint16_t)((1u << cl) - 1u);
  const int mc = P.S * 128;
  const uint32_t buf_cols = (uint32_t)(P.S * kDCols);
  const uint32_t want_cols = buf_cols * (uint32_t)n_buf;
  const uint32_t tmem_cols = want_cols <= 32u ? 32u : want_cols <= 64u ? 64u : want_cols <= 128u ? 128u
                             : want_cols <= 256u ? 256u : 512u;

  if (threadIdx.x == 0) {
    for (int i = 0; i < n_a; ++i) { mbar_init(a_full(i), 1); mbar_init(a_empty(i), 1); }
    for (int i = 0; i < n_b; ++i) { mbar_init(b_full(i), 1); mbar_init(b_empty(i), cl); }
    for (int i = 0; i < 2; ++i) { mbar_init(acc_full(i), 1); mbar_init(acc_empty(i), 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(tmem_slot)),
                 "r"(tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (warp == 3) {
    for (int c = lane; c < COUT; c += 32) s_bias[c] = bias[c];
  }
  tc_fence_before();
  __syncthreads();
  if (cl > 1) cluster_sync_all();   // peers' barriers must be initialised before anything is multicast
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // The next layer may now set itself up and stream its weights while this one runs.  (Only after this
  // CTA owns its TMEM columns: a dependent CTA that took them first would wait for this grid while this
  // CTA waits for the columns.)
  grid_launch_dependents();

  if (warp == 0 && lane == 0) {
    // ===== A-patch producer: one 16-channel slice (2 planes hi [+2 planes lo]) per stage =====
    grid_dependency_wait();     // the activations are the previous launch's output
    int it = 0;
    long long w_aempty = 0, t_begin = clock64();
    for (int tile = vblock; tile < P.n_tiles; tile += vgrid) {
      const long long q0 = (long long)tile * mc;
      for (int i = 0; i < P.n_stages; ++i, ++it) {
        const int slot = it % n_a;
        const long long c0 = P.dbg ? clock64() : 0;
        mbar_wait(a_empty(slot), ((uint32_t)(it / n_a) & 1u) ^ 1u);
        if (P.dbg) w_aempty += clock64() - c0;
        mbar_expect_tx(a_full(slot), a_stage_bytes);
        for (int js = 0; js < P.a_nsrc; ++js) {
          const FlatTensor& T = P.src[P.st_src[i] + js];
          const uint8_t* g = T.base + (long long)(2 * P.st_chunk[i]) * T.plane_stride +
                             ((long long)T.guard + q0 + P.lo_shift) * 16;
          const uint32_t dst = smem_u32(a_base + (size_t)slot * a_stage_stride) + (uint32_t)js * a_patch_bytes;
          bulk_g2s(dst, g, a_patch_bytes, a_full(slot));
          bulk_g2s(dst + a_plane_bytes, g + T.plane_stride, a_patch_bytes, a_full(slot));
          if (SPLIT) {
            bulk_g2s(dst + 2 * a_plane_bytes, g + T.lo_offset, a_patch_bytes, a_full(slot));
            bulk_g2s(dst + 3 * a_plane_bytes, g + T.lo_offset + T.plane_stride, a_patch_bytes,
                     a_full(slot));
          }
        }
      }
    }
    if (P.dbg) { P.dbg[blockIdx.x * 16 + 0] = w_aempty; P.dbg[blockIdx.x * 16 + 1] = clock64() - t_begin; }
  } else if (warp == 1 && lane == 0) {
    // ===== weight-slab producer: one (stage, tap) slab per B stage, multicast over the cluster =====
    // (up to G consecutive taps of one stage travel as one bulk copy; the packed blob of this
    // numeric mode is contiguous in program order)
    int kb = 0;
    long long w_bempty = 0;
    for (int tile = vblock; tile < P.n_tiles; tile += vgrid) {
      if (P.b_resident && tile != vblock) break;   // resident weights: loaded with the first tile only
      int k = 0;
      for (int i = 0; i < P.n_stages; ++i) {
        const int nt = P.st_ntaps[i];
        for (int t0 = 0; t0 < nt; t0 += G, ++kb) {
          const int n = min(G, nt - t0);
          const int slot = kb % n_b;
          const long long c0 = P.dbg ? clock64() : 0;
          mbar_wait(b_empty(slot), ((uint32_t)(kb / n_b) & 1u) ^ 1u);
          if (P.dbg) w_bempty += clock64() - c0;
          const uint32_t bytes = (uint32_t)n * kBSlab;
          mbar_expect_tx(b_full(slot), bytes);
          const uint32_t dst = smem_u32(b_base + (size_t)slot * b_stage_bytes);
          const uint8_t* src = weights + (size_t)k * kBSlab;
          if (cl == 1) {
            bulk_g2s(dst, src, bytes, b_full(slot));
          } else {
            const uint32_t part = bytes / cl;
            bulk_g2s_multicast(dst + cta_rank * part, src + cta_rank * part, part, b_full(slot), cta_mask);
          }
          k += n;
        }
      }
    }
    if (P.dbg) P.dbg[blockIdx.x * 16 + 2] = w_bempty;
  } else if (warp == 2) {
    // ===== MMA issuer: the whole warp walks the program (so every value stays warp-uniform and
    // lives in uniform registers); one elected lane issues tcgen05.mma / tcgen05.commit =====
    constexpr uint32_t idesc = (1u << 4) /* D = f32 */ | ((MODE == UMT_MODE_FP16 ? 0u : 1u) << 7) /* A */ |
                               ((MODE == UMT_MODE_FP16 ? 0u : 1u) << 10) /* B: 0 = f16, 1 = bf16 */ | ((uint32_t)(COUT >> 3) << 17) |
                               ((128u >> 4) << 24);
    const bool leader = elect_one();
    const uint64_t desc_hi_bits = ((uint64_t)((128u >> 4) & 0x3FFF) << 32) | (1ull << 46);
    const uint64_t a_desc_base = desc_hi_bits | ((uint64_t)((a_plane_bytes >> 4) & 0x3FFF) << 16);
    // plane distance of a weight slab: COUT rows (16-bit modes) or hi + lo rows (split mode)
    const uint64_t b_desc_base = desc_hi_bits | ((uint64_t)((((SPLIT ? 2u : 1u) * COUT * 16u) >> 4) & 0x3FFF) << 16);
    int it = 0, kb = 0, tc = 0;
    long long w_acc = 0, w_a = 0, w_b = 0, t_begin = clock64();
    unsigned long long g_begin = 0;
    if (P.dbg) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g_begin));
    for (int tile = vblock; tile < P.n_tiles; tile += vgrid, ++tc) {
      const int buf = tc % n_buf;
      long long c0 = P.dbg ? clock64() : 0;
      mbar_wait(acc_empty(buf), ((uint32_t)(tc / n_buf) & 1u) ^ 1u);   // epilogue drained this buffer
      if (P.dbg) w_acc += clock64() - c0;
      tc_fence_after();
      const uint32_t d_base = tmem_base + (uint32_t)buf * buf_cols;
      uint32_t any = 0;
      for (int i = 0; i < P.n_stages; ++i, ++it) {
        const int aslot = it % n_a;
        c0 = P.dbg ? clock64() : 0;
        mbar_wait(a_full(aslot), (uint32_t)(it / n_a) & 1u);
        if (P.dbg) w_a += clock64() - c0;
        const uint32_t a_s = smem_u32(a_base + (size_t)aslot * a_stage_stride);
        const bool skip_hilo = SPLIT && (P.st_flags[i] & 1);   // identity (residual) slabs have w_lo == 0
        const int nt = P.st_ntaps[i];
        // low 32 bits of the A descriptors of this stage (start address field, 16-byte units);
        // tap shift and sub-tile only add to it
        const uint32_t a_lo32 = (uint32_t)(a_desc_base & 0xFFFFFFFFu) | ((a_s >> 4) & 0x3FFFu);
        const uint32_t a_hi32 = (uint32_t)(a_desc_base >> 32);
        for (int t0 = 0; t0 < nt; t0 += G, ++kb) {
          const int n = min(G, nt - t0);
          const int bslot = kb % n_b;
          c0 = P.dbg ? clock64() : 0;
          if (!P.b_resident || tile == vblock) mbar_wait(b_full(bslot), (uint32_t)(kb / n_b) & 1u);
          if (P.dbg) w_b += clock64() - c0;
          tc_fence_after();
          const uint32_t b_s0 = smem_u32(b_base + (size_t)bslot * b_stage_bytes);
          const uint32_t b_lo32 = (uint32_t)(b_desc_base & 0xFFFFFFFFu) | ((b_s0 >> 4) & 0x3FFFu);
          const uint32_t b_hi32 = (uint32_t)(b_desc_base >> 32);
          if (leader) {
            switch (P.S) {
              case 1: issue_taps<1, COUT, SPLIT, MERGE>(P, i, t0, n, a_lo32, a_hi32, a_plane_bytes, b_lo32, b_hi32, d_base, idesc, any, skip_hilo); break;
              case 2: issue_taps<2, COUT, SPLIT, MERGE>(P, i, t0, n, a_lo32, a_hi32, a_plane_bytes, b_lo32, b_hi32, d_base, idesc, any, skip_hilo); break;
              default: issue_taps<4, COUT, SPLIT, MERGE>(P, i, t0, n, a_lo32, a_hi32, a_plane_bytes, b_lo32, b_hi32, d_base, idesc, any, skip_hilo); break;
            }
            if (P.b_resident) {}                                  // resident weights are never released
            else if (cl == 1) tc_commit(b_empty(bslot));
            else tc_commit_multicast(b_empty(bslot), cta_mask);   // frees the slot in every CTA of the cluster
          }
          any = 1u;
          __syncwarp();
        }
        if (leader) tc_commit(a_empty(aslot));
        __syncwarp();
      }
      if (leader) tc_commit(acc_full(buf));
      __syncwarp();
    }
    if (P.dbg && leader) {
      P.dbg[blockIdx.x * 16 + 3] = w_acc; P.dbg[blockIdx.x * 16 + 4] = w_a; P.dbg[blockIdx.x * 16 + 5] = w_b;
      P.dbg[blockIdx.x * 16 + 6] = clock64() - t_begin; P.dbg[blockIdx.x * 16 + 7] = tc;
      unsigned long long g_end;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g_end));
      unsigned smid;
      asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
      P.dbg[blockIdx.x * 16 + 10] = (long long)g_begin; P.dbg[blockIdx.x * 16 + 11] = (long long)g_end;
      P.dbg[blockIdx.x * 16 + 12] = smid;
    }
  } else if (warp >= 4) {
    // ===== epilogue: TMEM -> registers -> bias / ReLU -> split bf16 -> global =====
    grid_dependency_wait();     // residual rows and the output tensor belong to earlier launches until they are done
    const int lgrp = warp & 3;  // TMEM lane group this warp may read
    int tc = 0;
    bool fp16_over = false;
    long long w_full = 0, t_begin = clock64();
    for (int tile = vblock; tile < P.n_tiles; tile += vgrid, ++tc) {
      const int buf = tc % n_buf;
      const long long q0 = (long long)tile * mc;
      constexpr int kStep = (COUT % 32 == 0 && !MERGE) ? 32 : 16;   // columns per TMEM round trip
      constexpr int kRows = kStep / 8;
      // identity residual: the rows of the NEXT chunk are requested before the current chunk is
      // processed (and the first chunk's before the accumulator is awaited)
      uint4 rn_h[kRows], rn_l[kRows];
      auto res_fetch = [&](long long q, bool valid, int c0) {
        const uint8_t* g = P.res.base + (long long)(pl0 + (c0 >> 3)) * P.res.plane_stride + (P.res.guard + q) * 16;
#pragma unroll
        for (int h = 0; h < kRows; ++h) {
          if (valid) {
            rn_h[h] = __ldg(reinterpret_cast<const uint4*>(g + (long long)h * P.res.plane_stride));
            if (SPLIT) rn_l[h] = __ldg(reinterpret_cast<const uint4*>(g + (long long)h * P.res.plane_stride + P.res.lo_offset));
          }
        }
      };
      OutPos o_next = map_output(q0 + lgrp * 32 + lane, P.H, P.W, P.n_pos, P.phase_out);
      if (EPI == kEpiResidual) {
        res_fetch(q0 + lgrp * 32 + lane, o_next.valid, 0);
        // the accumulator is still being computed: pull the rest of this tile's residual rows
        // into L2 now (they were written two layers ago and have usually been evicted)
        if (q0 + (long long)P.S * 128 <= P.n_pos) {
          for (int s = 0; s < P.S; ++s) {
            const uint8_t* g = P.res.base + (long long)pl0 * P.res.plane_stride + (P.res.guard + q0 + s * 128 + lgrp * 32 + lane) * 16;
#pragma unroll 4
            for (int pl = 0; pl < COUT / 8; ++pl) {
              asm volatile("prefetch.global.L2 [%0];" ::"l"(g + (long long)pl * P.res.plane_stride));
              if (SPLIT) asm volatile("prefetch.global.L2 [%0];" ::"l"(g + (long long)pl * P.res.plane_stride + P.res.lo_offset));
            }
          }
        }
      }
      const long long c0 = P.dbg ? clock64() : 0;
      mbar_wait(acc_full(buf), (uint32_t)(tc / n_buf) & 1u);
      if (P.dbg) w_full += clock64() - c0;
      tc_fence_after();
      for (int s = 0; s < P.S; ++s) {
        const long long q = q0 + s * 128 + lgrp * 32 + lane;
        const OutPos o = o_next;
        if (s + 1 < P.S) o_next = map_output(q + 128, P.H, P.W, P.n_pos, P.phase_out);
        const FlatTensor& T = P.out[o.tensor];
        const float lo_clamp = P.relu ? 0.0f : -INFINITY;
#pragma unroll 1
        for (int c0 = 0; c0 < COUT; c0 += kStep) {
          uint32_t v[kStep];
          uint32_t v2[MERGE ? kStep : 1];   // second accumulator (a_hi * w_lo) of the merged form
          const uint32_t taddr = tmem_base + ((uint32_t)(lgrp * 32) << 16) +
                                 (uint32_t)buf * buf_cols + (uint32_t)(s * kDCols + c0);
          tc_ld16(taddr, v);
          if (kStep == 32) tc_ld16(taddr + 16, v + 16);
          if (MERGE) tc_ld16(taddr + COUT, v2);
          uint4 rc_h[kRows], rc_l[kRows];
          if (EPI == kEpiResidual) {
#pragma unroll
            for (int h = 0; h < kRows; ++h) { rc_h[h] = rn_h[h]; if (SPLIT) rc_l[h] = rn_l[h]; }
            if (c0 + kStep < COUT) res_fetch(q, o.valid, c0 + kStep);
            else if (s + 1 < P.S) res_fetch(q + 128, o_next.valid, 0);
          }
          tc_ld_wait();
          if (o.valid) {
#pragma unroll
            for (int h = 0; h < kStep / 8; ++h) {
              float f[8];
#pragma unroll
              for (int e = 0; e < 8; ++e) {
                f[e] = __uint_as_float(v[h * 8 + e]);
                if (MERGE) f[e] += __uint_as_float(v2[h * 8 + e]);
                f[e] += s_bias[c0 + h * 8 + e];
              }
              if (EPI == kEpiResidual) {
                float r[8];
                decode_row8<MODE>(rc_h[h], rc_l[h], r);
#pragma unroll
                for (int e = 0; e < 8; ++e) f[e] += r[e];
              }
#pragma unroll
              for (int e = 0; e < 8; ++e) f[e] = fmaxf(f[e], lo_clamp);
              if (MODE == UMT_MODE_FP16 && EPI != kEpiDense) {   // a value beyond the half range would be stored as inf
                float mx = 0.f;
#pragma unroll
                for (int e = 0; e < 8; ++e) mx = fmaxf(mx, fabsf(f[e]));
                fp16_over |= (mx > 65504.f);
              }
              if (EPI == kEpiDense) {
                // float32 NCHW output [n, dense_c, H, W]: lanes hold consecutive pixels, so for each
                // channel the warp writes (nearly) consecutive floats
                const long long hw = (long long)P.H * P.W, img = o.dense / hw, pix = o.dense - img * hw;
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                  const int c = c0 + h * 8 + e;
                  if (c < P.dense_c) P.dense_out[(img * P.dense_c + c) * hw + pix] = f[e];
                }
              } else {
                store_row8<MODE>(T, pl0 + (c0 >> 3) + h, o.row, f);
              }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_empty(buf));   // 4 arrivals (one per epilogue warp) free the buffer
    }
    if (MODE == UMT_MODE_FP16 && P.status && __any_sync(0xFFFFFFFFu, fp16_over) && lane == 0)
      atomicOr(P.status, (unsigned)UMT_STATUS_FP16_RANGE);
    if (P.dbg && warp == 4 && lane == 0) { P.dbg[blockIdx.x * 16 + 8] = w_full; P.dbg[blockIdx.x * 16 + 9] = clock64() - t_begin; }
  }
  tc_fence_before();
  __syncthreads();
  if (cl > 1) cluster_sync_all();   // nobody leaves while a peer may still multicast into its smem
  if (warp == 2) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                 "r"(tmem_cols)
                 : "memory");
  }
}

static size_t conv_smem_bytes(const ConvParams& p, int cout, bool split) {
  size_t bias = ((size_t)cout * 4 + 127) / 128 * 128;
  size_t a = (((size_t)p.patch_rows * p.a_nsrc * 16 * (split ? 4 : 2)) + 127) & ~(size_t)127;
  size_t b = (size_t)2 * cout * 16 * (split ? 2 : 1) * p.bgroup;
  return 1024 + bias + a * p.n_a + b * p.n_b;
}

constexpr int kMaxDevices = 64;
static int env_int(const char* name, int dflt);
static long long* g_conv_dbg = nullptr;   // set by umt_debug_set_conv_timeline (tools only)
static int g_conv_dbg_launch = 0;         // UMT_CONV_DBG_STRIDE=n: launch i writes its counters at buffer + i * n
void set_conv_debug_buffer(long long* p) { g_conv_dbg = p; g_conv_dbg_launch = 0; }
// counters of the next launch (nullptr in production); conv_pair.cu asks here
long long* next_conv_debug_slot() {
  if (!g_conv_dbg) return nullptr;
  const int stride = env_int("UMT_CONV_DBG_STRIDE", 0);
  return stride > 0 ? g_conv_dbg + (size_t)(g_conv_dbg_launch++) * stride : g_conv_dbg;
}

static int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
  }
  return n;
}

template <int COUT, int MODE>
static int launch_conv_t(const ConvParams& p, cudaStream_t st) {
  const size_t smem = conv_smem_bytes(p, COUT, MODE == UMT_MODE_FP32_EQUIV);
  void (*kernel)(ConvParams) = conv_umma_kernel<COUT, MODE, kEpiPlain>;
  if (p.dense_out) {
    if constexpr (COUT == 80) kernel = conv_umma_kernel<COUT, MODE, kEpiDense>;
    else return fail(UMT_ERR_UNSUPPORTED, "float32 output epilogue only for the projection width");
    if (p.res_epi) return fail(UMT_ERR_INVALID, "float32 output epilogue has no residual");
  } else if (p.res_epi) {
    kernel = conv_umma_kernel<COUT, MODE, kEpiResidual>;
  }
  ConvParams pd = p;
  pd.dbg = next_conv_debug_slot();
  {  // the opt-in shared-memory size is a per-device attribute of each kernel: set it once (and when it grows)
    static size_t have[3][kMaxDevices] = {};
    const int epi = p.dense_out ? kEpiDense : p.res_epi ? kEpiResidual : kEpiPlain;
    int dev = 0;
    UMT_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices || have[epi][dev] < smem) {
      UMT_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      if (dev >= 0 && dev < kMaxDevices) have[epi][dev] = smem;
    }
  }
  int grid = std::min(p.n_tiles, sm_count() * std::max(1, p.occ));
  grid = std::max(p.cluster, grid / p.cluster * p.cluster);
  if (p.n_slices > 1) grid = p.n_tiles * p.n_slices;      // few tiles: one CTA per (tile, output-channel slice)
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(kConvThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)p.cluster;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = p.overlap_launch ? 2 : 1;
  UMT_CUDA_TRY(cudaLaunchKernelEx(&cfg, kernel, pd));
  return check_launch("conv_umma_kernel");
}

template <int COUT>
static int launch_conv_c(const ConvParams& p, int mode, cudaStream_t st) {
  switch (mode) {
    case UMT_MODE_FP32_EQUIV: return launch_conv_t<COUT, UMT_MODE_FP32_EQUIV>(p, st);
    case UMT_MODE_BF16: return launch_conv_t<COUT, UMT_MODE_BF16>(p, st);
    case UMT_MODE_FP16: return launch_conv_t<COUT, UMT_MODE_FP16>(p, st);
  }
  return fail(UMT_ERR_INVALID, "conv: bad mode %d", mode);
}

int launch_conv(const ConvParams& p, int cout, int mode, cudaStream_t st) {
  if (p.pair_weights) return launch_conv_pair(p, cout, mode, st);   // 128 / 256 channels: CTA pairs (conv_pair.cu)
  switch (cout) {
    case 32: return launch_conv_c<32>(p, mode, st);
    case 64: return launch_conv_c<64>(p, mode, st);
    case 128: return launch_conv_c<128>(p, mode, st);
    case 256: return launch_conv_c<256>(p, mode, st);
    // widths of the fusion / temporal / regressor layers (always fp32-equivalent)
    case 80: return launch_conv_c<80>(p, mode, st);   // also the 256 -> 72 projection (all modes)
    case 96: if (mode == UMT_MODE_FP32_EQUIV) return launch_conv_t<96, UMT_MODE_FP32_EQUIV>(p, st); break;
    case 112: if (mode == UMT_MODE_FP32_EQUIV) return launch_conv_t<112, UMT_MODE_FP32_EQUIV>(p, st); break;
  }
  return fail(UMT_ERR_UNSUPPORTED, "conv: unsupported Cout %d (mode %d)", cout, mode);
}

// ------------------------------------------------------------------------------------------
// CUDA-core check kernel (test hook): same program, same packed weights, plain fp32 FMAs,
// dense [n, C, H, W] float output.  One thread per (position, output channel).
// ------------------------------------------------------------------------------------------
__global__ void conv_simt_check_kernel(const ConvParams P, int cout, int mode,
                                       float* __restrict__ out /* [n,cout,H,W] */) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int Wp = P.W + 1, S = (P.H + 1) * Wp;
  const long long total = (long long)P.n_pos * cout;
  if (idx >= total) return;
  const int co = (int)(idx % cout);
  const long long q = idx / cout;
  const int img = (int)(q / S), rem = (int)(q % S), y = rem / Wp, x = rem % Wp;
  if (x >= P.W || y >= P.H) return;
  float acc = 0.f;
  int k = 0;
  const int split = (mode == 0);
  const size_t half = (size_t)2 * cout * 16, slab = split ? 2 * half : half;
  for (int i = 0; i < P.n_stages; ++i) {
    for (int t = 0; t < P.st_ntaps[i]; ++t, ++k) {
      const FlatTensor& T = P.src[P.st_src[i] + P.tap_src[P.st_tap0[i] + t]];
      const long long row = q + P.tap_shift[P.st_tap0[i] + t];
      const uint8_t* w = P.weights + (size_t)k * slab;
      for (int pl = 0; pl < 2; ++pl) {
        // split mode: [plane][hi rows | lo rows]; 16-bit modes: [plane][rows]
        const uint8_t* wb = w + (size_t)pl * (split ? 2 : 1) * cout * 16 + (size_t)co * 16;
        for (int e = 0; e < 8; ++e) {
          const float a = load_elem(T, mode, (2 * P.st_chunk[i] + pl) * 8 + e, row);
          float b;
          if (mode == 2) {
            b = __half2float(reinterpret_cast<const __half*>(wb)[e]);
          } else {
            b = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(wb)[e]);
            if (split) b += __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(wb + (size_t)cout * 16)[e]);
          }
          acc = fmaf(a, b, acc);
        }
      }
    }
  }
  acc += P.bias[co];
  if (P.res_epi) acc += load_elem(P.res, mode, co, q);
  if (P.relu) acc = fmaxf(acc, 0.f);
  out[(((size_t)img * cout + co) * P.H + y) * P.W + x] = acc;
}

// planar flat (or 4 parity-phase tensors) -> dense [n, C, H, W] float (test hook)
__global__ void densify_kernel(const ConvParams P, int C, int mode, float* __restrict__ out) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = (long long)P.n_pos * C;
  if (idx >= total) return;
  const int c = (int)(idx % C);
  const long long q = idx / C;
  const OutPos o = map_output(q, P.H, P.W, P.n_pos, P.phase_out);
  if (!o.valid) return;
  const FlatTensor& T = P.out[o.tensor];
  const float v = load_elem(T, mode, c, o.row);
  const int Wp = P.W + 1, S = (P.H + 1) * Wp;
  const int img = (int)(q / S), rem = (int)(q % S), y = rem / Wp, x = rem % Wp;
  out[(((size_t)img * C + c) * P.H + y) * P.W + x] = v;
}

// ------------------------------------------------------------------------------------------
// Stem: conv3x3(1->32, bias) + BN + ReLU + MaxPool2 (lib/models/model_utils.py:119-124), BN folded.
// One thread per pooled output pixel; writes the 32 channels as four 16-byte rows (hi [+ lo]).
// ------------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(256)
stem_kernel(const float* __restrict__ images, int n_img, const float* __restrict__ w /* [32][9] */,
            const float* __restrict__ b, FlatTensor out) {
  __shared__ float sw[32 * 9];
  __shared__ float sb[32];
  for (int i = threadIdx.x; i < 288; i += blockDim.x) sw[i] = w[i];
  if (threadIdx.x < 32) sb[threadIdx.x] = b[threadIdx.x];
  __syncthreads();
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = (long long)n_img * 48 * 48;
  if (idx >= total) return;
  const int img = (int)(idx / 2304), rem = (int)(idx % 2304), oy = rem / 48, ox = rem % 48;
  const float* im = images + (size_t)img * 9216;
  float patch[4][4];
#pragma unroll
  for (int dy = 0; dy < 4; ++dy) {
#pragma unroll
    for (int dx = 0; dx < 4; ++dx) {
      const int yy = 2 * oy + dy - 1, xx = 2 * ox + dx - 1;
      patch[dy][dx] = (yy >= 0 && yy < 96 && xx >= 0 && xx < 96) ? __ldg(im + yy * 96 + xx) : 0.f;
    }
  }
  const long long row = (long long)img * 49 * 49 + oy * 49 + ox;
#pragma unroll 1
  for (int g = 0; g < 4; ++g) {
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int c = g * 8 + e;
      float best = 0.f;  // ReLU then max == max(0, .)
#pragma unroll
      for (int py = 0; py < 2; ++py) {
#pragma unroll
        for (int px = 0; px < 2; ++px) {
          float a = sb[c];
#pragma unroll
          for (int ky = 0; ky < 3; ++ky)
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) a = fmaf(patch[py + ky][px + kx], sw[c * 9 + ky * 3 + kx], a);
          best = fmaxf(best, a);
        }
      }
      v[e] = best;
    }
    store_row8<MODE>(out, g, row, v);
  }
}

// ------------------------------------------------------------------------------------------
// Stem on the tensor cores.  A pooled output pixel (oy,ox) sees the 4x4 input patch at rows
// 2oy-1..2oy+2, columns 2ox-1..2ox+2; the four conv outputs under its 2x2 pooling window are
// four different 3x3 sub-windows of that patch.  So with K = the 16 patch pixels and
// N = 4 pool members x 32 channels = 128, ONE M=128,N=128,K=16 MMA evaluates the convolution
// for 128 pooled pixels (three MMAs in the hi/lo split mode), and the epilogue takes the max
// over the four 32-column groups, adds the bias, applies ReLU and writes the planar rows.
// Warp roles: warp 0 issues the MMAs, warps 1-4 build the im2col operand in shared memory
// (one patch row each), warps 5-8 drain the accumulator; two operand and two accumulator
// buffers keep the three roles overlapped.
// ------------------------------------------------------------------------------------------
constexpr int kStemThreads = 288;
constexpr int kStemDepth = 6;                                  // staging tiles in flight per CTA
constexpr int kStemStageOff = 1024 + 8192 + 2 * 8192;
constexpr int kStemSmem = kStemStageOff + kStemDepth * 8192;   // dynamic shared memory

template <int MODE>
__global__ void __launch_bounds__(kStemThreads, 2)
stem_umma_kernel(const float* __restrict__ images, int n_img, const uint8_t* __restrict__ wpack,
                 const float* __restrict__ bias, FlatTensor out, unsigned* __restrict__ status) {
  constexpr bool SPLIT = (MODE == UMT_MODE_FP32_EQUIV);
  extern __shared__ __align__(128) uint8_t smem[];
  float* s_bias = reinterpret_cast<float*>(smem + 256);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + 128);
  uint8_t* b_smem = smem + 1024;               // [hi|lo][2 planes][128][8] 16-bit
  uint8_t* a_smem = smem + 1024 + 8192;        // 2 buffers x [hi|lo][2 planes][128 rows][8]
  const uint32_t bar0 = smem_u32(smem);
  auto a_full = [&](int i) { return bar0 + 8u * (uint32_t)i; };
  auto a_empty = [&](int i) { return bar0 + 8u * (uint32_t)(2 + i); };
  auto acc_full = [&](int i) { return bar0 + 8u * (uint32_t)(4 + i); };
  auto acc_empty = [&](int i) { return bar0 + 8u * (uint32_t)(6 + i); };
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long P = (long long)n_img * 2304;
  const int n_tiles = (int)((P + 127) / 128);

  if (threadIdx.x == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(a_full(i), 4); mbar_init(a_empty(i), 1); mbar_init(acc_full(i), 1); mbar_init(acc_empty(i), 4);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(256u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  {
    constexpr int kWBytes = SPLIT ? 8192 : 4096;
    for (int i = threadIdx.x; i < kWBytes / 16; i += blockDim.x)
      reinterpret_cast<uint4*>(b_smem)[i] = __ldg(reinterpret_cast<const uint4*>(wpack) + i);
    if (threadIdx.x < 32) s_bias[threadIdx.x] = bias[threadIdx.x];
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy stores -> tensor-core reads
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===== MMA issuer =====
    constexpr uint32_t idesc = (1u << 4) | ((MODE == UMT_MODE_FP16 ? 0u : 1u) << 7) |
                               ((MODE == UMT_MODE_FP16 ? 0u : 1u) << 10) | ((128u >> 3) << 17) | ((128u >> 4) << 24);
    const bool leader = elect_one();
    const uint32_t b_s = smem_u32(b_smem);
    int it = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
      const int buf = it & 1;
      const uint32_t ph = (uint32_t)(it >> 1) & 1u;
      mbar_wait(acc_empty(buf), ph ^ 1u);
      mbar_wait(a_full(buf), ph);
      tc_fence_after();
      if (leader) {
        const uint32_t a_s = smem_u32(a_smem + buf * 8192);
        const uint32_t d = tmem_base + (uint32_t)buf * 128u;
        tc_mma_bf16(d, umma_desc(a_s, 2048, 128), umma_desc(b_s, 2048, 128), idesc, 0u);
        if (SPLIT) {
          tc_mma_bf16(d, umma_desc(a_s, 2048, 128), umma_desc(b_s + 4096, 2048, 128), idesc, 1u);
          tc_mma_bf16(d, umma_desc(a_s + 4096, 2048, 128), umma_desc(b_s, 2048, 128), idesc, 1u);
        }
        tc_commit(a_empty(buf));
        tc_commit(acc_full(buf));
      }
      __syncwarp();
    }
  } else if (warp <= 4) {
    // ===== im2col: one pooled pixel (= one operand row of 16 patch pixels) per thread =====
    const int r = (warp - 1) * 32 + lane;
    // The patch pixels travel global -> shared with 4-byte cp.async into a ring of kStemDepth
    // staging tiles ([16 pixels][128 rows] float, zero-filled outside the image), so the loads of
    // several tiles are in flight without holding registers; each thread only ever reads back
    // the row it requested itself, so cp.async.wait_group is all the synchronisation needed.
    float* stage = reinterpret_cast<float*>(smem + kStemStageOff);
    auto request = [&](int tile, int slot) {
      const long long p = (long long)tile * 128 + r;
      const bool ok = tile < n_tiles && p < P;
      const int img = ok ? (int)(p / 2304) : 0;
      const int rem = ok ? (int)(p - (long long)img * 2304) : 0, oy = rem / 48, ox = rem - oy * 48;
      const float* im = images + (size_t)img * 9216;
      const uint32_t dst0 = smem_u32(stage + (size_t)slot * 2048 + r);
#pragma unroll
      for (int dy = 0; dy < 4; ++dy) {
        const int yy = 2 * oy + dy - 1;
#pragma unroll
        for (int dx = 0; dx < 4; ++dx) {
          const int xx = 2 * ox + dx - 1;
          const bool in = ok && yy >= 0 && yy < 96 && xx >= 0 && xx < 96;
          const float* src = in ? im + yy * 96 + xx : images;
          asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst0 + (uint32_t)(dy * 4 + dx) * 512u),
                       "l"(src), "r"(in ? 4u : 0u) : "memory");
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
    for (int d = 0; d < kStemDepth; ++d) request(blockIdx.x + d * (int)gridDim.x, d);
    int it = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
      const int buf = it & 1, slot = it % kStemDepth;
      asm volatile("cp.async.wait_group %0;" ::"n"(kStemDepth - 1) : "memory");
      float px[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) px[i] = stage[(size_t)slot * 2048 + i * 128 + r];
      uint4 h[2], l[2];
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const float* v = px + 8 * k;
        if (MODE == UMT_MODE_FP16) {
          h[k] = make_uint4(pack_f16x2(v[0], v[1]), pack_f16x2(v[2], v[3]), pack_f16x2(v[4], v[5]), pack_f16x2(v[6], v[7]));
        } else {
          h[k] = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
          if (SPLIT)
            l[k] = make_uint4(pack_bf16x2(v[0] - bf16_lo(h[k].x), v[1] - bf16_hi(h[k].x)),
                              pack_bf16x2(v[2] - bf16_lo(h[k].y), v[3] - bf16_hi(h[k].y)),
                              pack_bf16x2(v[4] - bf16_lo(h[k].z), v[5] - bf16_hi(h[k].z)),
                              pack_bf16x2(v[6] - bf16_lo(h[k].w), v[7] - bf16_hi(h[k].w)));
        }
      }
      request(tile + kStemDepth * (int)gridDim.x, slot);   // refill the staging slot just consumed
      mbar_wait(a_empty(buf), ((uint32_t)(it >> 1) & 1u) ^ 1u);
      uint8_t* a = a_smem + buf * 8192 + r * 16;
      *reinterpret_cast<uint4*>(a) = h[0];
      *reinterpret_cast<uint4*>(a + 2048) = h[1];
      if (SPLIT) {
        *reinterpret_cast<uint4*>(a + 4096) = l[0];
        *reinterpret_cast<uint4*>(a + 6144) = l[1];
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(a_full(buf));
    }
  } else {
    // ===== epilogue: max over the four pool members, + bias, ReLU, planar rows =====
    const int lgrp = warp & 3;
    int it = 0;
    bool fp16_over = false;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
      const int buf = it & 1;
      const long long p = (long long)tile * 128 + lgrp * 32 + lane;
      const bool valid = p < P;
      const int img = (int)(p / 2304), rem = (int)(p - (long long)img * 2304), oy = rem / 48, ox = rem - oy * 48;
      const long long row = (long long)img * 2401 + oy * 49 + ox;
      mbar_wait(acc_full(buf), (uint32_t)(it >> 1) & 1u);
      tc_fence_after();
      const uint32_t t0 = tmem_base + ((uint32_t)(lgrp * 32) << 16) + (uint32_t)buf * 128u;
#pragma unroll 1
      for (int c0 = 0; c0 < 32; c0 += 16) {
        uint32_t v0[16], v1[16], v2[16], v3[16];
        tc_ld16(t0 + c0, v0);
        tc_ld16(t0 + 32 + c0, v1);
        tc_ld16(t0 + 64 + c0, v2);
        tc_ld16(t0 + 96 + c0, v3);
        tc_ld_wait();
        if (valid) {
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            float f[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const int i = h * 8 + e;
              const float m = fmaxf(fmaxf(__uint_as_float(v0[i]), __uint_as_float(v1[i])),
                                    fmaxf(__uint_as_float(v2[i]), __uint_as_float(v3[i])));
              f[e] = fmaxf(m + s_bias[c0 + i], 0.f);
              if (MODE == UMT_MODE_FP16) fp16_over |= (f[e] > 65504.f);
            }
            store_row8<MODE>(out, (c0 >> 3) + h, row, f);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_empty(buf));
    }
    if (MODE == UMT_MODE_FP16 && status && __any_sync(0xFFFFFFFFu, fp16_over) && lane == 0)
      atomicOr(status, (unsigned)UMT_STATUS_FP16_RANGE);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(256u) : "memory");
  }
}

// ------------------------------------------------------------------------------------------
// Projection conv1x1(256->72, bias) (lib/models/model_utils.py:134): planar flat (6x6x256) ->
// dense float32 [n,72,6,6].  One CTA per image: activations in shared memory, each thread
// produces one output channel for six pixels (weights read once per six FMAs).
// ------------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(432)
proj_kernel(FlatTensor in, int n_img, const float* __restrict__ w /* [256][72], transposed */,
            const float* __restrict__ b, float* __restrict__ out) {
  __shared__ float sx[36][257];
  const int img = blockIdx.x;
  for (int i = threadIdx.x; i < 36 * 32; i += blockDim.x) {
    const int pix = i / 32, plane = i % 32;
    const long long row = (long long)img * 49 + (pix / 6) * 7 + (pix % 6);
    float v[8];
    load_row8<MODE>(in, plane, row, v);
#pragma unroll
    for (int e = 0; e < 8; ++e) sx[pix][plane * 8 + e] = v[e];
  }
  __syncthreads();
  const int co = threadIdx.x % 72, pg = threadIdx.x / 72;  // pixel group of 6
  float acc[6];
#pragma unroll
  for (int j = 0; j < 6; ++j) acc[j] = b[co];
  for (int ci = 0; ci < 256; ++ci) {
    const float ww = __ldg(w + ci * 72 + co);   // consecutive lanes read consecutive channels
#pragma unroll
    for (int j = 0; j < 6; ++j) acc[j] = fmaf(sx[pg * 6 + j][ci], ww, acc[j]);
  }
#pragma unroll
  for (int j = 0; j < 6; ++j) out[((size_t)img * 72 + co) * 36 + pg * 6 + j] = acc[j];
}

// ------------------------------------------------------------------------------------------
// Host side: BN folding, operand packing, programs, workspace layout, layer schedule
// ------------------------------------------------------------------------------------------
static inline uint16_t f2bf(float f) {  // round to nearest even
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7F800000u) == 0x7F800000u) return (uint16_t)(u >> 16);
  u += 0x7FFFu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
static inline uint16_t f2h(float f) {  // float -> IEEE half, round to nearest even, saturating
  uint32_t x;
  memcpy(&x, &f, 4);
  const uint32_t sign = (x >> 16) & 0x8000u;
  x &= 0x7FFFFFFFu;
  if (x >= 0x47800000u) return (uint16_t)(sign | (x > 0x7F800000u ? 0x7E00u : 0x7BFFu));   // NaN / clamp to 65504
  if (x < 0x38800000u) {  // subnormal half or zero
    if (x < 0x33000000u) return (uint16_t)sign;
    const int shift = 126 - (int)(x >> 23);          // 14..24
    uint32_t mant = (x & 0x7FFFFFu) | 0x800000u;
    const uint32_t rounded = (mant + (1u << (shift - 1)) - 1u + ((mant >> shift) & 1u)) >> shift;
    return (uint16_t)(sign | rounded);
  }
  const uint32_t r = x + 0xFFFu + ((x >> 13) & 1u);   // round mantissa to 10 bits
  return (uint16_t)(sign | ((r - 0x38000000u) >> 13));
}
static inline float bf2f(uint16_t h) {
  uint32_t u = (uint32_t)h << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}


static Folded fold_conv_bn(const float* w, const float* conv_bias, const float* gamma,
                           const float* beta, const float* mean, const float* var, int cout,
                           int cin, int ks) {
  Folded f;
  f.cout = cout; f.cin = cin; f.ks = ks;
  f.w.resize((size_t)cout * cin * ks * ks);
  f.b.resize(cout);
  for (int co = 0; co < cout; ++co) {
    float scale = 1.f, shift = 0.f;
    if (gamma) {
      // same arithmetic as eval-mode batch_norm: (x - mean) / sqrt(var + eps) * gamma + beta
      scale = gamma[co] / std::sqrt(var[co] + 1e-5f);
      shift = beta[co] - mean[co] * scale;
    }
    for (size_t i = 0; i < (size_t)cin * ks * ks; ++i)
      f.w[(size_t)co * cin * ks * ks + i] = w[(size_t)co * cin * ks * ks + i] * scale;
    f.b[co] = (conv_bias ? conv_bias[co] * scale : 0.f) + shift;
  }
  return f;
}

static const float* tensor_ptr(const float* const* t, const std::string& name) {
  int i = tensor_index(name.c_str());
  return i < 0 ? nullptr : t[i];
}

static Folded folded_from(const float* const* t, const std::string& conv, const std::string& bn,
                          int cout, int cin, int ks, bool has_bias) {
  return fold_conv_bn(tensor_ptr(t, conv + ".weight"), has_bias ? tensor_ptr(t, conv + ".bias") : nullptr,
                      bn.empty() ? nullptr : tensor_ptr(t, bn + ".weight"),
                      bn.empty() ? nullptr : tensor_ptr(t, bn + ".bias"),
                      bn.empty() ? nullptr : tensor_ptr(t, bn + ".running_mean"),
                      bn.empty() ? nullptr : tensor_ptr(t, bn + ".running_var"), cout, cin, ks);
}

// append one operand slab: [plane 0: hi [cout][8], lo [cout][8]] [plane 1: same]
static inline size_t slab_index(int cout, int pl, int lo, int co, int e) {
  return (((size_t)pl * 2 + lo) * cout + co) * 8 + e;
}
static void pack_slab(std::vector<uint16_t>& blob, const Folded& f, int chunk, int ky, int kx) {
  const size_t base = blob.size();
  blob.resize(base + (size_t)4 * f.cout * 8);
  for (int pl = 0; pl < 2; ++pl)
    for (int co = 0; co < f.cout; ++co)
      for (int e = 0; e < 8; ++e) {
        const int ci = chunk * 16 + pl * 8 + e;
        const float v = f.w[(((size_t)co * f.cin + ci) * f.ks + ky) * f.ks + kx];
        const uint16_t hi = f2bf(v);
        const uint16_t lo = f2bf(v - bf2f(hi));
        blob[base + slab_index(f.cout, pl, 0, co, e)] = hi;
        blob[base + slab_index(f.cout, pl, 1, co, e)] = lo;
      }
}

static int env_int(const char* name, int dflt) {
  const char* s = getenv(name);
  return s ? atoi(s) : dflt;
}

// identity slab: adds source channel 16*chunk + k to output channel of the same index
static void pack_identity_slab(std::vector<uint16_t>& blob, int cout, int chunk) {
  const size_t base = blob.size();
  blob.resize(base + (size_t)4 * cout * 8, 0);
  for (int pl = 0; pl < 2; ++pl)
    for (int e = 0; e < 8; ++e) {
      const int co = chunk * 16 + pl * 8 + e;
      blob[base + slab_index(cout, pl, 0, co, e)] = 0x3F80;  // bf16(1.0)
    }
}

// Stage table of a stride-2 3x3 convolution: tap (ky,kx) reads parity phase ((ky+1)&1, (kx+1)&1) of
// the input, shifted by -1 row/col iff k == 0.  The four phases contribute 1, 2, 2 and 4 taps; stages
// that small are dominated by their pipeline hand-shake, so `merge` (1, 2 or 4) consecutive phases share
// one A stage: their patches sit back to back in each plane and a tap's descriptor shift also selects
// the patch.  merge = 2 gives stages of 3 and 6 taps, merge = 4 one 9-tap stage per 16 channels (four
// patches per stage: too much shared memory for the hi/lo split mode).
static void s2_stage_table(ConvParams& p, int cin, int Wp, int merge, int& ns, int& nt, int& lo, int& hi) {
  merge = (merge == 4 || merge == 2) ? merge : 1;
  p.a_nsrc = merge;
  ns = 0; nt = 0;
  int tap0[4], cnt[4];
  for (int ph = 0; ph < 4; ++ph) {
    tap0[ph] = nt;
    cnt[ph] = 0;
    const int py = ph >> 1, px = ph & 1;
    for (int ky = 0; ky < 3; ++ky) {
      if (((ky + 1) & 1) != py) continue;
      for (int kx = 0; kx < 3; ++kx) {
        if (((kx + 1) & 1) != px) continue;
        const int shift = (ky == 0 ? -1 : 0) * Wp + (kx == 0 ? -1 : 0);
        p.tap_src[nt] = (uint8_t)(ph % merge);
        p.tap_shift[nt++] = (short)shift;
        lo = std::min(lo, shift);
        hi = std::max(hi, shift);
        ++cnt[ph];
      }
    }
  }
  for (int c = 0; c < cin / 16; ++c)
    for (int g0 = 0; g0 < 4; g0 += merge) {
      int n = 0;
      for (int ph = g0; ph < g0 + merge; ++ph) n += cnt[ph];
      p.st_src[ns] = (uint8_t)g0; p.st_chunk[ns] = (uint8_t)c; p.st_ntaps[ns] = (uint8_t)n;
      p.st_tap0[ns] = (uint8_t)tap0[g0];
      ++ns;
    }
}

// Build program + packed weights of one backbone convolution.
int build_conv(umt_model* m, ConvLayer& L, const Folded& main, const Folded* ds) {
  ConvParams& p = L.prog;
  memset(&p, 0, sizeof(p));
  const int H = kStageHW[L.geom], W = H, Wp = W + 1;
  p.H = H; p.W = W;
  p.relu = L.relu;
  std::vector<uint16_t> blob;
  int ns = 0, nt = 0, lo = 0, hi = 0;
  auto add_tap = [&](int shift) {
    p.tap_shift[nt++] = (short)shift;
    lo = std::min(lo, shift);
    hi = std::max(hi, shift);
  };
  if (main.ks == 1) {   // pointwise convolution: one tap, no halo
    add_tap(0);
    for (int c = 0; c < main.cin / 16; ++c) {
      p.st_src[ns] = 0; p.st_chunk[ns] = (uint8_t)c; p.st_ntaps[ns] = 1; p.st_tap0[ns] = 0;
      ++ns;
      pack_slab(blob, main, c, 0, 0);
    }
  } else if (L.stride == 1) {
    for (int ky = 0; ky < 3; ++ky)
      for (int kx = 0; kx < 3; ++kx) add_tap((ky - 1) * Wp + (kx - 1));
    for (int c = 0; c < main.cin / 16; ++c) {
      p.st_src[ns] = 0; p.st_chunk[ns] = (uint8_t)c; p.st_ntaps[ns] = 9; p.st_tap0[ns] = 0;
      ++ns;
      for (int ky = 0; ky < 3; ++ky)
        for (int kx = 0; kx < 3; ++kx) pack_slab(blob, main, c, ky, kx);
    }
    // identity residual (backbone_resnet.py:66-69).  Two ways, chosen per launch in
    // finalize_conv_params: (a) the epilogue re-reads x (prefetched one chunk ahead) and adds it in
    // float32, or (b) x enters the accumulator as extra K through identity weight slabs (source 2,
    // centre tap), exactly as hi + lo.  The identity stages are the LAST stages of the program so
    // that (a) simply stops before them.
    if (L.res_buf >= 0) {
      p.res_stages = L.cout / 16;
      for (int c = 0; c < L.cout / 16; ++c) {
        p.st_src[ns] = 2; p.st_chunk[ns] = (uint8_t)c; p.st_ntaps[ns] = 1; p.st_tap0[ns] = 4;
        p.st_flags[ns] = 1;
        ++ns;
        pack_identity_slab(blob, L.cout, c);
      }
    }
    if (ds && L.res_buf >= 0) return fail(UMT_ERR_INVALID, "conv: identity residual and downsample branch together");
    if (ds) {  // fused 1x1 stride-2 downsample: source 1 = parity-phase (0,0) of the block input
      for (int c = 0; c < ds->cin / 16; ++c) {
        p.st_src[ns] = 1; p.st_chunk[ns] = (uint8_t)c; p.st_ntaps[ns] = 1; p.st_tap0[ns] = 4;
        ++ns;
        pack_slab(blob, *ds, c, 0, 0);
      }
    }
  } else {
    // stride 2 (see s2_stage_table): weights are packed in phase order, which every grouping shares
    s2_stage_table(p, main.cin, Wp, env_int("UMT_CONV_S2_MERGE", 2), ns, nt, lo, hi);
    for (int c = 0; c < main.cin / 16; ++c)
      for (int ph = 0; ph < 4; ++ph) {
        const int py = ph >> 1, px = ph & 1;
        for (int ky = 0; ky < 3; ++ky) {
          if (((ky + 1) & 1) != py) continue;
          for (int kx = 0; kx < 3; ++kx) {
            if (((kx + 1) & 1) != px) continue;
            pack_slab(blob, main, c, ky, kx);
          }
        }
      }
  }
  if (p.a_nsrc < 1) p.a_nsrc = 1;
  if (ns > kMaxStages || nt > kMaxTaps) return fail(UMT_ERR_INVALID, "conv program too large");
  p.n_stages = ns;
  p.lo_shift = lo;
  // tile shape: S sub-tiles of 128 positions.  Two accumulator buffers of S*Cout TMEM columns
  // (epilogue/MMA overlap) except for Cout = 256, where one 512-column buffer halves the weight
  // traffic per position instead.
  // Tuned per output width; override with UMT_CONV_CFG<cout>="S,occ,G,cluster,nbuf" (tools/sweep_conv.py)
  // (B200 sweep, round 1: two co-resident CTAs per SM beat one larger CTA for every width; as many
  // taps per weight stage as shared memory allows)
  // (taps per weight stage: 9 where they fit; 5 / 3 for the 128 / 256-channel layers, whose stages would
  // otherwise be cut unevenly by the shared-memory budget -- +2 % / +1 %, tools/ab_conv.py)
  int S = (L.cout == 32) ? 4 : (L.cout == 256 || L.cout % 32 != 0 || L.cout == 96) ? 1 : 2, occ = 2,
      G = (L.cout == 128) ? 5 : (L.cout == 256) ? 3 : 9, clu = (L.cout == 64 || L.cout == 128) ? 1 : 2, nbuf = 2;
  // (weight multicast over a CTA pair: neutral for 32 / 256 channels, -1.7 % for 64 / 128 -- the L2 is
  // not the limit and the cluster hand-shakes cost more than the halved fetch saves)
  {
    const std::string name = "UMT_CONV_CFG" + std::to_string(L.cout);
    if (const char* e = getenv(name.c_str())) sscanf(e, "%d,%d,%d,%d,%d", &S, &occ, &G, &clu, &nbuf);
  }
  occ = std::max(1, std::min(2, occ));
  while (S > 1 && S * L.cout * occ > 512) S >>= 1;
  p.S = S;
  p.n_buf = (2 * S * L.cout * occ <= 512) ? std::max(1, std::min(2, nbuf)) : 1;
  p.occ = occ;
  p.bgroup = std::max(1, std::min(9, G));
  p.cluster = std::max(1, std::min(2, clu));
  p.halo_rows = hi - lo;
  p.patch_rows = S * 128 + p.halo_rows;
  p.phase_out = L.phase_out;
  L.n_slabs = (int)(blob.size() / ((size_t)4 * L.cout * 8));
  std::vector<float> bias = main.b;
  if (ds) for (int i = 0; i < L.cout; ++i) bias[i] += ds->b[i];
  UMT_CUDA_TRY(cudaMalloc(&L.d_weights, blob.size() * 2));
  m->owned.push_back(L.d_weights);
  UMT_CUDA_TRY(cudaMemcpy(L.d_weights, blob.data(), blob.size() * 2, cudaMemcpyHostToDevice));
  {  // hi halves only, contiguous, for the single-product mode
    const size_t half = (size_t)2 * L.cout * 8;
    std::vector<uint16_t> hi((size_t)L.n_slabs * half);
    for (int k = 0; k < L.n_slabs; ++k)
      for (int pl = 0; pl < 2; ++pl)
        memcpy(&hi[(size_t)k * half + (size_t)pl * L.cout * 8], &blob[(size_t)k * 2 * half + slab_index(L.cout, pl, 0, 0, 0)],
               (size_t)L.cout * 8 * 2);
    UMT_CUDA_TRY(cudaMalloc(&L.d_weights_hi, hi.size() * 2));
    m->owned.push_back(L.d_weights_hi);
    UMT_CUDA_TRY(cudaMemcpy(L.d_weights_hi, hi.data(), hi.size() * 2, cudaMemcpyHostToDevice));
  }
  {  // fp16 weights (UMT_MODE_FP16).  Values beyond the half range saturate.
    const size_t half = (size_t)2 * L.cout * 8;
    std::vector<uint16_t> h1((size_t)L.n_slabs * half);
    for (int k = 0; k < L.n_slabs; ++k)
      for (int pl = 0; pl < 2; ++pl)
        for (int co = 0; co < L.cout; ++co)
          for (int e = 0; e < 8; ++e) {
            const size_t src = (size_t)k * 2 * half;
            const float v = bf2f(blob[src + slab_index(L.cout, pl, 0, co, e)]) +
                            bf2f(blob[src + slab_index(L.cout, pl, 1, co, e)]);   // hi + lo == the folded fp32 weight (to 2^-17)
            const uint16_t hv = f2h(v);
            if (std::fabs(v) > 65504.f) m->fp16_weights_ok = false;
            h1[(size_t)k * half + ((size_t)pl * L.cout + co) * 8 + e] = hv;
          }
    UMT_CUDA_TRY(cudaMalloc(&L.d_weights_h16, h1.size() * 2));
    m->owned.push_back(L.d_weights_h16);
    UMT_CUDA_TRY(cudaMemcpy(L.d_weights_h16, h1.data(), h1.size() * 2, cudaMemcpyHostToDevice));
  }
  if (!L.no_pair && main.ks == 3 && L.stride == 1 && p.a_nsrc == 1 && (L.cout == 32 || L.cout == 64 || L.cout == 128 || L.cout == 256)) {
    // weights repacked per CTA half for the pair kernel (conv_pair.cu); the trailing identity stages are
    // not part of it (these widths always add the residual in the epilogue)
    std::vector<int> ntaps;
    for (int i = 0; i < ns - p.res_stages; ++i) ntaps.push_back(p.st_ntaps[i]);
    for (int kind = 0; kind < 3; ++kind) {
      L.pair_G[kind] = (kind == 0 || L.cout == 256) ? 3 : 9;
      const std::string gname = std::string(kind == 0 ? "UMT_PAIR_G_SPLIT" : "UMT_PAIR_G16") + std::to_string(L.cout);
      const char* e = getenv(gname.c_str());
      if (!e) e = getenv(kind == 0 ? "UMT_PAIR_G_SPLIT" : "UMT_PAIR_G16");
      if (e) { const int g = atoi(e); if (g == 1 || g == 3 || g == 9) L.pair_G[kind] = g; }
      std::vector<uint16_t> pb;
      pack_pair_blob(blob, L.cout, ntaps, L.pair_G[kind], kind, f2h, bf2f, pb);
      UMT_CUDA_TRY(cudaMalloc(&L.d_pair[kind], pb.size() * 2));
      m->owned.push_back(L.d_pair[kind]);
      UMT_CUDA_TRY(cudaMemcpy(L.d_pair[kind], pb.data(), pb.size() * 2, cudaMemcpyHostToDevice));
    }
  }
  if (main.ks == 3 && L.stride == 2 && env_int("UMT_CONV_S2_MERGE", 2) == 2) {
    // the 16-bit modes have the shared memory for all four phase patches in one stage
    L.prog16 = p;
    int ns16 = 0, nt16 = 0, lo16 = 0, hi16 = 0;
    s2_stage_table(L.prog16, main.cin, Wp, 4, ns16, nt16, lo16, hi16);
    L.prog16.n_stages = ns16;
    L.has_prog16 = 1;
  }
  L.d_bias = upload(m, bias);
  if (!L.d_bias) return UMT_ERR_CUDA;
  return UMT_OK;
}

// 128 / 256 output channels: the same layer once more as 2 / 4 layers of 64 output channels (own weight
// blobs in every numeric mode, own biases), for launches on few positions -- a frame of the per-frame
// tracker has one or two 128-position tiles in these layers, and a tile of the whole layer is 20-30 us of
// tensor-core time on one SM; as (tile, slice) CTAs the layer spreads over 8-12 SMs (ConvParams::n_slices).
static int build_slices(umt_model* m, ConvLayer& L, const Folded& main, const Folded* ds) {
  constexpr int kSlice = 64;
  L.slices.clear();
  if (L.cout < 128 || L.cout % kSlice != 0 || L.cout / kSlice > 4 || !env_int("UMT_CONV_SLICES", 1)) return UMT_OK;
  auto cut = [&](const Folded& f, int j) {
    Folded g;
    g.cout = kSlice; g.cin = f.cin; g.ks = f.ks;
    const size_t row = (size_t)f.cin * f.ks * f.ks;
    g.w.assign(f.w.begin() + (size_t)j * kSlice * row, f.w.begin() + (size_t)(j + 1) * kSlice * row);
    g.b.assign(f.b.begin() + (size_t)j * kSlice, f.b.begin() + (size_t)(j + 1) * kSlice);
    return g;
  };
  for (int j = 0; j < L.cout / kSlice; ++j) {
    ConvLayer S;
    S.cout = kSlice; S.cin = L.cin; S.stride = L.stride; S.geom = L.geom; S.ds_cin = L.ds_cin;
    S.src_kind = L.src_kind; S.src_buf = L.src_buf; S.out_buf = L.out_buf; S.res_buf = L.res_buf;
    S.phase_out = L.phase_out; S.relu = L.relu; S.no_pair = 1;
    const Folded fj = cut(main, j);
    if (ds) {
      const Folded dj = cut(*ds, j);
      UMT_TRY(build_conv(m, S, fj, &dj));
    } else {
      UMT_TRY(build_conv(m, S, fj, nullptr));
    }
    L.slices.push_back(S);
  }
  return UMT_OK;
}

int encoder_build(umt_model* m, const float* const* t) {
  const std::string bb = "_feature_extractor._image_backbone.0._layers.";
  m->convs.clear();
  int cin = 32;
  for (int g = 0; g < 4; ++g) {
    const int cout = kStageC[g];
    int cur = 0;  // index of the ping/pong buffer holding the block input
    for (int b = 0; b < kStageBlocks[g]; ++b) {
      const std::string pre = bb + std::to_string(g + 1) + "." + std::to_string(b);
      const bool first_ds = (b == 0 && g > 0);
      const bool last = (b == kStageBlocks[g] - 1);
      ConvLayer c1, c2;
      c1.cout = cout; c1.cin = first_ds ? cin : cout; c1.stride = first_ds ? 2 : 1; c1.geom = g;
      c1.src_kind = first_ds ? 1 : 0; c1.src_buf = cur; c1.out_buf = 2; c1.res_buf = -1;
      Folded f1 = folded_from(t, pre + ".conv1", pre + ".bn1", cout, c1.cin, 3, false);
      UMT_TRY(build_conv(m, c1, f1, nullptr));
      UMT_TRY(build_slices(m, c1, f1, nullptr));
      m->convs.push_back(c1);
      c2.cout = cout; c2.cin = cout; c2.stride = 1; c2.geom = g; c2.src_kind = 0; c2.src_buf = 2;
      c2.phase_out = (last && g < 3) ? 1 : 0;
      Folded f2 = folded_from(t, pre + ".conv2", pre + ".bn2", cout, cout, 3, false);
      if (first_ds) {
        c2.ds_cin = cin;
        c2.res_buf = -1;
        c2.out_buf = 0;
        Folded fd = folded_from(t, pre + ".downsample.0", pre + ".downsample.1", cout, cin, 1, false);
        UMT_TRY(build_conv(m, c2, f2, &fd));
        UMT_TRY(build_slices(m, c2, f2, &fd));
        cur = 0;
      } else {
        c2.res_buf = cur;
        c2.out_buf = 1 - cur;
        UMT_TRY(build_conv(m, c2, f2, nullptr));
        UMT_TRY(build_slices(m, c2, f2, nullptr));
        cur = 1 - cur;
      }
      m->convs.push_back(c2);
    }
    cin = cout;
  }
  // stem + projection (CUDA cores, float32)
  Folded st = folded_from(t, bb + "0.0", bb + "0.1", 32, 1, 3, true);
  m->tw.stem_w = upload(m, st.w);
  m->tw.stem_b = upload(m, st.b);
  {  // tensor-core stem: B[n = pool member (py,px) * 32 + c][k = patch pixel (dy,dx)] = w[c][dy-py][dx-px]
    std::vector<uint16_t> bf((size_t)2 * 2 * 128 * 8, 0), h16((size_t)2 * 128 * 8, 0);
    for (int pp = 0; pp < 4; ++pp)
      for (int c = 0; c < 32; ++c)
        for (int k = 0; k < 16; ++k) {
          const int py = pp >> 1, px = pp & 1, ky = (k >> 2) - py, kx = (k & 3) - px;
          const float w = (ky >= 0 && ky < 3 && kx >= 0 && kx < 3) ? st.w[(size_t)c * 9 + ky * 3 + kx] : 0.f;
          const int n = pp * 32 + c;
          const size_t o = ((size_t)(k >> 3) * 128 + n) * 8 + (k & 7);
          const uint16_t hi = f2bf(w);
          bf[o] = hi;
          bf[(size_t)2 * 128 * 8 + o] = f2bf(w - bf2f(hi));
          h16[o] = f2h(w);
          if (std::fabs(w) > 65504.f) m->fp16_weights_ok = false;
        }
    UMT_CUDA_TRY(cudaMalloc(&m->tw.stem_pack_bf, bf.size() * 2));
    m->owned.push_back(m->tw.stem_pack_bf);
    UMT_CUDA_TRY(cudaMemcpy(m->tw.stem_pack_bf, bf.data(), bf.size() * 2, cudaMemcpyHostToDevice));
    UMT_CUDA_TRY(cudaMalloc(&m->tw.stem_pack_h16, h16.size() * 2));
    m->owned.push_back(m->tw.stem_pack_h16);
    UMT_CUDA_TRY(cudaMemcpy(m->tw.stem_pack_h16, h16.data(), h16.size() * 2, cudaMemcpyHostToDevice));
  }
  Folded pj = folded_from(t, "_feature_extractor._image_backbone.1", "", 72, 256, 1, true);
  {  // the projection also as a tensor-core layer: 1x1, 256 -> 72 (padded to 80), no ReLU, float32 output
    Folded pp;
    pp.cout = 80; pp.cin = 256; pp.ks = 1;
    pp.w.assign((size_t)80 * 256, 0.f);
    pp.b.assign(80, 0.f);
    for (int co = 0; co < 72; ++co) {
      for (int ci = 0; ci < 256; ++ci) pp.w[(size_t)co * 256 + ci] = pj.w[(size_t)co * 256 + ci];
      pp.b[co] = pj.b[co];
    }
    ConvLayer L;
    L.cout = 80; L.cin = 256; L.stride = 1; L.geom = 3; L.relu = 0; L.res_buf = -1;
    UMT_TRY(build_conv(m, L, pp, nullptr));
    m->proj_conv = L;
  }
  std::vector<float> pjt((size_t)256 * 72);
  for (int co = 0; co < 72; ++co)
    for (int ci = 0; ci < 256; ++ci) pjt[(size_t)ci * 72 + co] = pj.w[(size_t)co * 256 + ci];
  m->tw.proj_w = upload(m, pjt);
  m->tw.proj_b = upload(m, pj.b);
  if (!m->tw.stem_w || !m->tw.stem_b || !m->tw.proj_w || !m->tw.proj_b) return UMT_ERR_CUDA;
  return UMT_OK;
}

// ---- workspace layout -------------------------------------------------------------------
static long long plane_rows(int H, int n_img, int* guard_out) {
  const int Wp = H + 1;
  const int guard = ((Wp + 2 + 7) / 8) * 8;
  const long long pos = (long long)n_img * Wp * Wp;
  const long long padded = (pos + 1023) / 1024 * 1024;   // whole tiles for every cluster (512 x 2)
  *guard_out = guard;
  return guard + padded + guard + 8;
}

size_t flat_tensor_bytes(int H, int C, int n_img) {
  int g;
  return (size_t)plane_rows(H, n_img, &g) * 16 * (C / 8) * 2;
}
FlatTensor make_flat_tensor(uint8_t* base, int H, int C, int n_img);

static size_t flat_bytes(int H, int C, int n_img) {
  int g;
  return (size_t)plane_rows(H, n_img, &g) * 16 * (C / 8) * 2;
}

FlatTensor make_flat_tensor(uint8_t* base, int H, int C, int n_img) {
  FlatTensor t;
  int g;
  const long long rows = plane_rows(H, n_img, &g);
  t.base = base;
  t.plane_stride = rows * 16;
  t.lo_offset = t.plane_stride * (C / 8);
  t.guard = g;
  t.pad_ = 0;
  return t;
}

static FlatTensor make_flat(uint8_t* base, int H, int C, int n_img) {
  FlatTensor t;
  int g;
  const long long rows = plane_rows(H, n_img, &g);
  t.base = base;
  t.plane_stride = rows * 16;
  t.lo_offset = t.plane_stride * (C / 8);
  t.guard = g;
  t.pad_ = 0;
  return t;
}

size_t encoder_workspace_bytes(int chunk_images) {
  size_t total = 0;
  const int n = chunk_images;
  for (int g = 0; g < 4; ++g) {
    total += 3 * flat_bytes(kStageHW[g], kStageC[g], n);
    if (g > 0) total += 4 * flat_bytes(kStageHW[g], kStageC[g - 1], n);
  }
  return (total + 255) & ~(size_t)255;
}

void encoder_bind(umt_model* m, uint8_t* base) {
  const int n = m->chunk_images;
  uint8_t* p = base;
  for (int g = 0; g < 4; ++g) {
    for (int b = 0; b < 3; ++b) {
      m->buf[g][b] = make_flat(p, kStageHW[g], kStageC[g], n);
      p += flat_bytes(kStageHW[g], kStageC[g], n);
    }
    if (g > 0)
      for (int ph = 0; ph < 4; ++ph) {
        m->phase[g][ph] = make_flat(p, kStageHW[g], kStageC[g - 1], n);
        p += flat_bytes(kStageHW[g], kStageC[g - 1], n);
      }
  }
}

// pipeline depths etc. once tensors, weights and n_pos are known
void finalize_conv_params(ConvParams& p, const ConvLayer& L, int mode);

static void fill_tensors(const umt_model* m, const ConvLayer& L, int n_img, int mode,
                         ConvParams& p, int force_S = 0) {
  p = (L.has_prog16 && mode != UMT_MODE_FP32_EQUIV) ? L.prog16 : L.prog;
  p.force_S = force_S;
  const int g = L.geom;
  if (L.src_kind == 1) {
    for (int ph = 0; ph < 4; ++ph) p.src[ph] = m->phase[g][ph];
  } else {
    p.src[0] = m->buf[g][L.src_buf];
    if (L.ds_cin) p.src[1] = m->phase[g][0];
  }
  if (L.phase_out) {
    for (int ph = 0; ph < 4; ++ph) p.out[ph] = m->phase[g + 1][ph];
  } else {
    p.out[0] = m->buf[g][L.out_buf];
  }
  if (L.res_buf >= 0) { p.src[2] = m->buf[g][L.res_buf]; p.res = p.src[2]; }
  p.bias = L.d_bias;
  p.status = m->d_status;
  p.n_pos = n_img * (p.H + 1) * (p.W + 1);
  finalize_conv_params(p, L, mode);
}

void finalize_conv_params(ConvParams& p, const ConvLayer& L, int mode) {
  const bool split = (mode == UMT_MODE_FP32_EQUIV);
  p.weights = split ? L.d_weights : (mode == UMT_MODE_FP16 ? L.d_weights_h16 : L.d_weights_hi);
  // identity residual: through the epilogue or through identity slabs?  (B200, round 1: the
  // one-MMA identity stages cost a full pipeline hand-shake each, so the epilogue wins wherever
  // it is not itself the bottleneck -- everywhere in the 3-product mode, and for Cout >= 64 in
  // the 16-bit modes.)  UMT_RES_MMA=0/1 forces one way.
  if (p.res_stages > 0) {
    int force = env_int("UMT_RES_MMA", -1);
    force = env_int(("UMT_RES_MMA" + std::to_string(L.cout)).c_str(), force);   // per width: UMT_RES_MMA32=1 ...
    const bool epi = force >= 0 ? force == 0 : (split || L.cout >= 64);
    if (epi) { p.n_stages -= p.res_stages; p.res_epi = 1; }
  }
  // 128 / 256 channels, stride 1: the CTA-pair kernel (UMT_CONV_PAIR=0 falls back to one CTA per tile)
  p.pair_weights = nullptr;
  if (L.d_pair[0] && pair_kernel_enabled(L.cout, mode) && !p.dense_out && (p.res_stages == 0 || p.res_epi)) {
    const int kind = split ? 0 : (mode == UMT_MODE_FP16 ? 2 : 1);
    p.pair_weights = L.d_pair[kind];
    p.pair_G = L.pair_G[kind];
  }
  // tile shape per numeric mode.  The merged split form (issue_taps) keeps two accumulators per sub-tile;
  // for the 32-channel layers two buffers of S = 2 sub-tiles (epilogue overlapped inside the CTA) beat one
  // buffer of S = 4 by 28 % (tools/ab_conv.py, B200).  UMT_CONV_SPLIT_CFG<cout>="S,nbuf" overrides.
  if (p.force_S > 0 && p.force_S < p.S) {
    p.S = p.force_S;
    p.patch_rows = p.S * 128 + p.halo_rows;
    p.n_buf = (2 * p.S * L.cout * p.occ <= 512) ? 2 : 1;
  }
  if (split && L.cout <= 64) {
    int S = (L.cout == 32) ? 2 : p.S, nbuf = 2;
    const std::string name = "UMT_CONV_SPLIT_CFG" + std::to_string(L.cout);
    if (const char* e = getenv(name.c_str())) sscanf(e, "%d,%d", &S, &nbuf);
    if (p.force_S > 0) S = std::min(S, p.force_S);
    S = std::max(1, std::min(4, S));
    while (S > 1 && S * 2 * L.cout * p.occ > 512) S >>= 1;
    p.S = S;
    p.patch_rows = S * 128 + p.halo_rows;
    p.n_buf = (2 * S * 2 * L.cout * p.occ <= 512) ? std::max(1, std::min(2, nbuf)) : 1;
  }
  // pipeline depths from the shared-memory budget
  const int mc = p.S * 128;
  p.n_tiles = ((p.n_pos + mc - 1) / mc + p.cluster - 1) / p.cluster * p.cluster;
  // 112 KB per CTA: two CTAs plus the per-CTA reserve fill the SM's 228 KB (+4 % on the 256-channel
  // layers over 108 KB, whose weight stages are the largest; tools/ab_conv.py)
  const size_t budget = (size_t)env_int("UMT_CONV_SMEM_KB", L.prog.occ == 2 ? 112 : 200) * 1024;
  const size_t a = (((size_t)p.patch_rows * p.a_nsrc * 16 * (split ? 4 : 2)) + 127) & ~(size_t)127;
  const size_t fixed = 1024 + ((size_t)L.cout * 4 + 127) / 128 * 128;
  const size_t slab = (size_t)2 * L.cout * 16 * (split ? 2 : 1);
  while (p.bgroup > 1 && fixed + a * 2 + slab * p.bgroup * 3 > budget) --p.bgroup;   // keep >= 3 B stages
  const size_t b = slab * p.bgroup;
  int n_a = env_int("UMT_CONV_NA", 3);
  while (n_a > 2 && fixed + a * n_a + b * 4 > budget) --n_a;
  long long nb = ((long long)budget - (long long)fixed - (long long)(a * n_a)) / (long long)b;
  p.n_a = n_a;
  p.n_b = (int)std::max(3LL, std::min((long long)env_int("UMT_CONV_NB_MAX", 24), nb));
  p.b_resident = 0;
  // Narrow layers: the whole weight set is a few tens of KB.  If it fits beside three A stages, every CTA
  // loads it once and keeps it -- no weight pipeline, no per-stage hand-shake, more room for A stages
  // (UMT_CONV_RESIDENT=0 switches this off).
  if (env_int("UMT_CONV_RESIDENT", 1) && L.cout <= 64) {
    int groups = 0;
    for (int i = 0; i < p.n_stages; ++i) groups += (p.st_ntaps[i] + 8) / 9;
    const size_t need = fixed + a * 3 + (size_t)groups * 9 * slab;
    if (need <= budget && 2 * 3 + 2 * groups + 4 <= 120) {
      p.b_resident = 1;
      p.bgroup = 9;
      p.n_b = groups;
      p.n_a = 3 + (int)std::min<size_t>(3, (budget - need) / a);
    }
  }
}

int encoder_run_chunk(umt_model* m, int mode, const float* images, int n_img, float* features,
                      int stop_after_layer, cudaStream_t st, cudaEvent_t* ev) {
  if (!m->ws) return fail(UMT_ERR_INVALID, "no workspace bound (umt_model_bind_workspace)");
  if (n_img <= 0 || n_img > m->chunk_images)
    return fail(UMT_ERR_INVALID, "encoder chunk of %d images (capacity %d)", n_img, m->chunk_images);
  const long long px = (long long)n_img * 2304;
  const int blocks = (int)((px + 255) / 256);
  if (ev) cudaEventRecord(ev[0], st);
  if (!env_int("UMT_STEM_SIMT", 0)) {
    const int tiles = (int)((px + 127) / 128);
    const int grid = std::max(1, std::min(tiles, 2 * sm_count()));
    {  // once per device and mode (the attribute belongs to the device)
      static bool done[3][kMaxDevices] = {};
      int dev = 0;
      UMT_CUDA_TRY(cudaGetDevice(&dev));
      if (dev < 0 || dev >= kMaxDevices || !done[mode][dev]) {
        if (mode == UMT_MODE_FP32_EQUIV)
          UMT_CUDA_TRY(cudaFuncSetAttribute(stem_umma_kernel<UMT_MODE_FP32_EQUIV>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStemSmem));
        else if (mode == UMT_MODE_BF16)
          UMT_CUDA_TRY(cudaFuncSetAttribute(stem_umma_kernel<UMT_MODE_BF16>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStemSmem));
        else
          UMT_CUDA_TRY(cudaFuncSetAttribute(stem_umma_kernel<UMT_MODE_FP16>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStemSmem));
        if (dev >= 0 && dev < kMaxDevices) done[mode][dev] = true;
      }
    }
    if (mode == UMT_MODE_FP32_EQUIV)
      stem_umma_kernel<UMT_MODE_FP32_EQUIV><<<grid, kStemThreads, kStemSmem, st>>>(images, n_img, m->tw.stem_pack_bf, m->tw.stem_b, m->buf[0][0], m->d_status);
    else if (mode == UMT_MODE_BF16)
      stem_umma_kernel<UMT_MODE_BF16><<<grid, kStemThreads, kStemSmem, st>>>(images, n_img, m->tw.stem_pack_bf, m->tw.stem_b, m->buf[0][0], m->d_status);
    else
      stem_umma_kernel<UMT_MODE_FP16><<<grid, kStemThreads, kStemSmem, st>>>(images, n_img, m->tw.stem_pack_h16, m->tw.stem_b, m->buf[0][0], m->d_status);
  } else if (mode == UMT_MODE_FP32_EQUIV)
    stem_kernel<UMT_MODE_FP32_EQUIV><<<blocks, 256, 0, st>>>(images, n_img, m->tw.stem_w, m->tw.stem_b, m->buf[0][0]);
  else if (mode == UMT_MODE_BF16)
    stem_kernel<UMT_MODE_BF16><<<blocks, 256, 0, st>>>(images, n_img, m->tw.stem_w, m->tw.stem_b, m->buf[0][0]);
  else
    stem_kernel<UMT_MODE_FP16><<<blocks, 256, 0, st>>>(images, n_img, m->tw.stem_w, m->tw.stem_b, m->buf[0][0]);
  UMT_TRY(check_launch("stem_kernel"));
  ++m->launches;
  const int overlap_max = env_int("UMT_CONV_OVERLAP_MAX", 64);
  const long long slice_rows = env_int("UMT_CONV_SLICE_ROWS", 2048);   // wide layers run sliced up to this many positions
  for (size_t i = 0; i < m->convs.size(); ++i) {
    if (stop_after_layer >= 0 && (int)i > stop_after_layer) return UMT_OK;
    ConvParams p;
    const ConvLayer& L = m->convs[i];
    const long long rows = (long long)n_img * (kStageHW[L.geom] + 1) * (kStageHW[L.geom] + 1);
    if (!L.slices.empty() && rows <= slice_rows) {
      // few positions: (tile, 64-channel slice) CTAs of the one-CTA kernel instead of whole-layer tiles
      fill_tensors(m, L.slices[0], n_img, mode, p, 1);
      p.pair_weights = nullptr;
      p.cluster = 1;
      p.n_slices = (int)L.slices.size();
      for (int j = 0; j < p.n_slices; ++j) {
        p.slice_weights[j] = mode == UMT_MODE_FP32_EQUIV ? L.slices[j].d_weights
                             : mode == UMT_MODE_FP16   ? L.slices[j].d_weights_h16 : L.slices[j].d_weights_hi;
        p.slice_bias[j] = L.slices[j].d_bias;
      }
      p.n_tiles = (p.n_pos + p.S * 128 - 1) / (p.S * 128);
    } else {
      fill_tensors(m, L, n_img, mode, p);
    }
    if (ev) cudaEventRecord(ev[1 + i], st);
    // Small batches (a frame of the per-frame tracker): a layer takes ~18 us of which most is set-up and
    // the weight stream of one or two CTAs; launched as programmatic dependents, the layers overlap that
    // part with their predecessors (UMT_CONV_OVERLAP_MAX = largest crop count that does, 0 = never).
    p.overlap_launch = (!ev && n_img <= overlap_max) ? 1 : 0;
    UMT_TRY(launch_conv(p, p.n_slices > 1 ? 64 : L.cout, mode, st));
    ++m->launches;
  }
  if (ev) cudaEventRecord(ev[1 + m->convs.size()], st);
  if (features && !env_int("UMT_PROJ_SIMT", 0)) {
    const ConvLayer& last = m->convs.back();
    ConvParams p = m->proj_conv.prog;
    p.src[0] = m->buf[3][last.out_buf];
      p.bias = m->proj_conv.d_bias;
    p.n_pos = n_img * 49;
    p.dense_out = features;
    p.dense_c = 72;
    p.overlap_launch = (!ev && n_img <= overlap_max) ? 1 : 0;
    finalize_conv_params(p, m->proj_conv, mode);
    UMT_TRY(launch_conv(p, 80, mode, st));
    ++m->launches;
  } else if (features) {
    const ConvLayer& last = m->convs.back();
    if (mode == UMT_MODE_FP32_EQUIV)
      proj_kernel<UMT_MODE_FP32_EQUIV><<<n_img, 432, 0, st>>>(m->buf[3][last.out_buf], n_img, m->tw.proj_w, m->tw.proj_b, features);
    else if (mode == UMT_MODE_BF16)
      proj_kernel<UMT_MODE_BF16><<<n_img, 432, 0, st>>>(m->buf[3][last.out_buf], n_img, m->tw.proj_w, m->tw.proj_b, features);
    else
      proj_kernel<UMT_MODE_FP16><<<n_img, 432, 0, st>>>(m->buf[3][last.out_buf], n_img, m->tw.proj_w, m->tw.proj_b, features);
    UMT_TRY(check_launch("proj_kernel"));
    ++m->launches;
  }
  if (ev) cudaEventRecord(ev[2 + m->convs.size()], st);
  return UMT_OK;
}

// Test hook: count the 16-byte rows that must always read as zero but do not -- guard rows before
// position 0 and after the padded end of every plane, the halo positions (x == W or y == H) of every
// image the workspace can hold, and the padding positions beyond its capacity.
__global__ void halo_check_kernel(FlatTensor T, int planes, int H, int n_cap, unsigned long long* count) {
  const long long rows = T.plane_stride / 16;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= rows * planes) return;
  const int pl = (int)(idx / rows);
  const long long r = idx - (long long)pl * rows, q = r - T.guard;
  const int Wp = H + 1, S = Wp * Wp;
  bool must_be_zero = q < 0 || q >= (long long)n_cap * S;
  if (!must_be_zero) {
    const int rem = (int)(q % S), y = rem / Wp, x = rem - y * Wp;
    must_be_zero = (x == H) || (y == H);
  }
  if (!must_be_zero) return;
  // planes [0, planes/2) are the hi planes, the rest the lo planes (lo_offset = plane_stride * C/8)
  const uint4 v = *reinterpret_cast<const uint4*>(T.base + (long long)pl * T.plane_stride + r * 16);
  if (v.x | v.y | v.z | v.w) atomicAdd(count, 1ull);
}

int encoder_halo_check(umt_model* m, long long* host_nonzero, cudaStream_t st) {
  if (!m->ws) return fail(UMT_ERR_INVALID, "no workspace bound");
  unsigned long long* d = nullptr;
  UMT_CUDA_TRY(cudaMalloc(&d, sizeof(*d)));          // test hook: may allocate
  cudaMemsetAsync(d, 0, sizeof(*d), st);
  auto run = [&](const FlatTensor& T, int C, int H) {
    const int planes = 2 * (C / 8);
    const long long total = T.plane_stride / 16 * planes;
    halo_check_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(T, planes, H, m->chunk_images, d);
  };
  for (int g = 0; g < 4; ++g) {
    for (int b = 0; b < 3; ++b) run(m->buf[g][b], kStageC[g], kStageHW[g]);
    if (g > 0)
      for (int ph = 0; ph < 4; ++ph) run(m->phase[g][ph], kStageC[g - 1], kStageHW[g]);
  }
  unsigned long long h = 0;
  cudaError_t e = cudaMemcpyAsync(&h, d, sizeof(h), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  cudaFree(d);
  if (e != cudaSuccess) return fail(UMT_ERR_CUDA, "halo check: %s", cudaGetErrorString(e));
  *host_nonzero = (long long)h;
  return UMT_OK;
}

int encoder_debug_layer(umt_model* m, int mode, int layer, const float* images, int n_img,
                        float* out_simt, float* out_umma, int* shape, cudaStream_t st) {
  if (layer < 0 || layer >= (int)m->convs.size()) return fail(UMT_ERR_INVALID, "layer %d out of range", layer);
  // inputs of `layer`: run everything before it
  UMT_TRY(encoder_run_chunk(m, mode, images, n_img, nullptr, layer - 1, st, nullptr));
  const ConvLayer& L = m->convs[layer];
  ConvParams p;
  fill_tensors(m, L, n_img, mode, p);
  if (shape) { shape[0] = L.cout; shape[1] = p.H; shape[2] = p.W; }
  const long long total = (long long)p.n_pos * L.cout;
  const int blocks = (int)((total + 255) / 256);
  const size_t dense = (size_t)n_img * L.cout * p.H * p.W * sizeof(float);
  if (out_simt) {
    UMT_CUDA_TRY(cudaMemsetAsync(out_simt, 0, dense, st));
    ConvParams pc = p;   // the check kernel reads the one-CTA weight packing
    pc.weights = mode == UMT_MODE_FP32_EQUIV ? L.d_weights : (mode == UMT_MODE_FP16 ? L.d_weights_h16 : L.d_weights_hi);
    conv_simt_check_kernel<<<blocks, 256, 0, st>>>(pc, L.cout, mode, out_simt);
    UMT_TRY(check_launch("conv_simt_check_kernel"));
  }
  if (out_umma) {
    UMT_TRY(launch_conv(p, L.cout, mode, st));
    UMT_CUDA_TRY(cudaMemsetAsync(out_umma, 0, dense, st));
    densify_kernel<<<blocks, 256, 0, st>>>(p, L.cout, mode, out_umma);
    UMT_TRY(check_launch("densify_kernel"));
  }
  return UMT_OK;
}

}  // namespace umt
