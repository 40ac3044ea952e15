// The 128- and 256-channel 3x3 convolutions of the backbone (58 % of its FLOPs;
// lib/models/backbone_resnet.py:56-72 at 12x12 and 6x6) on CTA PAIRS: tcgen05.mma.cta_group::2.
//
// Why pairs.  With one CTA per tile (encoder.cu) every M = 128 MMA reads its whole B operand
// (Cout x 16 weights) from that CTA's shared memory: 8 KB of operands per 64 clk of math at Cout = 128,
// which the shared-memory path cannot deliver (measured 94 clk per MMA), and a single 128-column
// accumulator per CTA whose epilogue the MMA warp has to wait for.  A CTA pair issues ONE
// M = 256 MMA for two 128-position tiles: each CTA keeps its own 128 rows of A and only HALF of
// the weight rows (the hardware reads the other half from the peer), so per SM the B read and the
// weight fetch from L2 are halved, the weight stages shrink to half (deeper pipelines in the same
// shared memory) and the 128-channel layers get two accumulator buffers in their 256 TMEM columns.
//
// Same data layout and the same arithmetic as conv_umma_kernel (planar flat activations, tap =
// constant row shift, K-major no-swizzle operands, hi/lo split products in the fp32-equivalent
// mode); what changes is who loads what and who signals whom:
//   * operands arrive by tensor-map TMA (cp.async.bulk.tensor ... cta_group::2), because only that
//     form lets the PEER's copy complete on the LEADER's mbarrier (address with the peer bit
//     cleared); plain bulk copies would need a relay thread (round 1: 20-35 % slower);
//   * A patch: 3-D map {8 channels, rows, planes} of an activation tensor, box {8, patch rows, 2};
//     weights: 2-D map {256 bytes, rows} over a blob repacked per CTA half, so one stage is one
//     contiguous wide box;
//   * the leader's MMA warp frees stages and publishes accumulators with
//     tcgen05.commit.cta_group::2 ... multicast::cluster (both CTAs' barriers); the peer's epilogue
//     warps return the accumulator with a remote mbarrier.arrive on the leader.
#include <cuda.h>

#include <algorithm>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "model.cuh"
#include "flat.cuh"
#include "umma.cuh"

namespace umt {

constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;   // shared::cluster address of the same offset in the pair's CTA 0
constexpr int kPairThreads = 256;
constexpr int kPairMaxStages = 24;

struct alignas(64) PairParams {
  CUtensorMap mapA[2];      // [0] main input, [1] input of the fused 1x1 stride-2 downsample branch
  CUtensorMap mapB[2];      // weight blob seen as rows of 256 bytes: box of G taps, box of one tap
  FlatTensor out[4];
  FlatTensor res;
  const float* bias;
  unsigned* status;
  int H, W, n_pos;
  int n_tiles;              // pair tiles of 2 * S * 128 positions
  int n_a, n_b, n_buf, G;
  int S;                    // 128-row sub-tiles per CTA (each MMA covers sub-tile s of both CTAs)
  int box_rows, n_box;      // an A plane arrives as n_box TMA boxes of box_rows rows (box limit: 256 rows)
  int lo_shift, patch_rows; // patch_rows = n_box * box_rows >= S*128 + halo: rows of one plane in shared memory
  int guard[2], planes[2];  // guard rows / hi planes (C/8) of the two sources
  int n_stages;
  int relu, res_epi, phase_out;
  int overlap_launch;       // host only: launch with programmatic stream serialization
  long long* dbg;           // optional [pairs][16] cycle counters, slots as in conv_umma_kernel (tools only)
  uint8_t st_src[kPairMaxStages], st_chunk[kPairMaxStages], st_ntaps[kPairMaxStages];
  short tap_shift[kMaxTaps];
  uint8_t st_tap0[kPairMaxStages];
};

// ---- PTX specific to the pair kernel ---------------------------------------------------------------
__device__ __forceinline__ void tma_load_3d_pair(uint32_t dst, const CUtensorMap* map, int c0, int c1, int c2,
                                                 uint32_t leader_bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(leader_bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* map, int c0, int c1,
                                                 uint32_t leader_bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(leader_bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tc_mma_pair(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the barrier at this offset in BOTH CTAs of the pair once all MMAs issued so far are done
__device__ __forceinline__ void tc_commit_pair(uint32_t bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
      ::"r"(bar), "h"((uint16_t)3)
      : "memory");
}
__device__ __forceinline__ void mbar_arrive_leader(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(bar & kPeerBitMask) : "memory");
}

// EPI: 0 = bias/ReLU -> planar rows, 1 = + identity residual (same meaning as in encoder.cu)
template <int COUT, int MODE, int EPI, int S>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kPairThreads, 2)
conv_pair_kernel(const __grid_constant__ PairParams P) {
  constexpr bool SPLIT = (MODE == UMT_MODE_FP32_EQUIV);
  // Narrow layers in the split mode use the merged form of encoder.cu's issue_taps: per tap
  //   [D1 | D2] += a_hi * [w_hi | w_lo]  (N = 2 COUT)      D1 += a_lo * w_hi  (N = COUT)
  // A pair splits the N rows of B between its CTAs, so for the first MMA CTA 0 holds all hi rows and CTA 1
  // all lo rows (region X, COUT rows), and for the second each holds its half of the hi rows (region Y).
  constexpr bool MERGE = SPLIT && COUT <= 64;
  constexpr int kDCols = MERGE ? 2 * COUT : COUT;                  // accumulator columns per buffer
  constexpr int kHalf = COUT / 2;                                  // weight rows of one CTA for an N = COUT MMA
  constexpr uint32_t kBPart = (uint32_t)kHalf * 16u;               // one of {hi, lo} of one plane (plain form)
  constexpr uint32_t kBPlane = MERGE ? 3u * kBPart : (SPLIT ? 2u : 1u) * kBPart;   // plane distance (descriptor LBO)
  constexpr uint32_t kBTap = 2u * kBPlane;                         // bytes of one tap in one CTA
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + 1000);
  float* s_bias = reinterpret_cast<float*>(smem + 1024);
  constexpr int kBiasBytes = ((COUT * 4 + 127) / 128) * 128;
  const int n_a = P.n_a, n_b = P.n_b, n_buf = P.n_buf, G = P.G;
  const uint32_t a_plane_bytes = (uint32_t)P.patch_rows * 16u;
  const uint32_t a_stage_bytes = a_plane_bytes * (SPLIT ? 4u : 2u);
  const uint32_t a_stage_stride = (a_stage_bytes + 127u) & ~127u;
  const uint32_t b_stage_bytes = (uint32_t)G * kBTap;
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
  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  const bool leader_cta = rank == 0;
  const int n_pairs = (int)(gridDim.x >> 1), pair_id = (int)(blockIdx.x >> 1);
  constexpr uint32_t buf_cols = (uint32_t)(S * kDCols);
  const uint32_t want_cols = buf_cols * (uint32_t)n_buf;
  const uint32_t tmem_cols = want_cols <= 32u ? 32u : want_cols <= 64u ? 64u : want_cols <= 128u ? 128u
                             : want_cols <= 256u ? 256u : 512u;

  if (threadIdx.x == 0) {
    // Every CTA initialises its own copy of every barrier; the FULL and ACC_EMPTY barriers are only
    // ever completed in the leader (the peer's copies and remote arrivals target the leader).
    for (int i = 0; i < n_a; ++i) { mbar_init(a_full(i), 1); mbar_init(a_empty(i), 1); }
    for (int i = 0; i < n_b; ++i) { mbar_init(b_full(i), 1); mbar_init(b_empty(i), 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(acc_full(i), 1); mbar_init(acc_empty(i), 8); }   // 4 epilogue warps x 2 CTAs
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(tmem_cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  if (warp == 3) {
    for (int c = lane; c < COUT; c += 32) s_bias[c] = P.bias[c];
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // the peer's barriers exist before anything is signalled across the pair
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  grid_launch_dependents();   // (after the TMEM allocation, see conv_umma_kernel)

  if (warp == 0 && lane == 0) {
    // ===== A-patch producer (both CTAs): own 128 positions + halo, one 16-channel slice per stage =====
    grid_dependency_wait();     // the activations are the previous launch's output
    int it = 0;
    for (int tile = pair_id; tile < P.n_tiles; tile += n_pairs) {
      const long long q0 = ((long long)tile * 2 + rank) * (S * 128);
      for (int i = 0; i < P.n_stages; ++i, ++it) {
        const int slot = it % n_a;
        mbar_wait(a_empty(slot), ((uint32_t)(it / n_a) & 1u) ^ 1u);
        if (leader_cta) mbar_expect_tx(a_full(slot), 2u * a_stage_bytes);   // both CTAs' bytes land on this barrier
        const int src = P.st_src[i];
        const int row = (int)(P.guard[src] + q0 + P.lo_shift);
        const uint32_t dst = smem_u32(a_base + (size_t)slot * a_stage_stride);
        const uint32_t bar = a_full(slot) & kPeerBitMask;
        const uint32_t box_bytes = (uint32_t)P.box_rows * 16u;
        // shared-memory planes: hi plane 0, hi plane 1 [, lo plane 0, lo plane 1]; each in n_box boxes
        if (P.n_box == 1) {   // the box covers both planes of the slice
          tma_load_3d_pair(dst, &P.mapA[src], 0, row, 2 * P.st_chunk[i], bar);
          if (SPLIT) tma_load_3d_pair(dst + 2 * a_plane_bytes, &P.mapA[src], 0, row, P.planes[src] + 2 * P.st_chunk[i], bar);
        } else {
          for (int pl = 0; pl < (SPLIT ? 4 : 2); ++pl) {
            const int gplane = (pl >> 1) * P.planes[src] + 2 * P.st_chunk[i] + (pl & 1);
            for (int bx = 0; bx < P.n_box; ++bx)
              tma_load_3d_pair(dst + (uint32_t)pl * a_plane_bytes + (uint32_t)bx * box_bytes, &P.mapA[src], 0,
                               row + bx * P.box_rows, gplane, bar);
          }
        }
      }
    }
  } else if (warp == 1 && lane == 0) {
    // ===== weight producer (both CTAs): this CTA's half of the rows of up to G taps per stage =====
    int kb = 0;
    for (int tile = pair_id; tile < P.n_tiles; tile += n_pairs) {
      int k = 0;   // taps before this group (all taps of all stages, program order)
      for (int i = 0; i < P.n_stages; ++i) {
        const int nt = P.st_ntaps[i];
        for (int t0 = 0; t0 < nt; t0 += G, ++kb) {
          const int n = min(G, nt - t0);   // G or 1 by construction (host checks)
          const int slot = kb % n_b;
          mbar_wait(b_empty(slot), ((uint32_t)(kb / n_b) & 1u) ^ 1u);
          if (leader_cta) mbar_expect_tx(b_full(slot), 2u * (uint32_t)n * kBTap);
          // blob: [group][CTA half][tap][plane][hi|lo][rows][16 B], viewed as rows of 256 bytes
          const uint32_t byte0 = (uint32_t)k * 2u * kBTap + rank * (uint32_t)n * kBTap;
          tma_load_2d_pair(smem_u32(b_base + (size_t)slot * b_stage_bytes), &P.mapB[n == 1 && G != 1 ? 1 : 0], 0,
                           (int)(byte0 >> 8), b_full(slot) & kPeerBitMask);
          k += n;
        }
      }
    }
  } else if (warp == 2 && leader_cta) {
    // ===== MMA issuer (leader CTA only): M = 256 over the pair =====
    constexpr uint32_t idesc = (1u << 4) /* D = f32 */ | ((MODE == UMT_MODE_FP16 ? 0u : 1u) << 7) |
                               ((MODE == UMT_MODE_FP16 ? 0u : 1u) << 10) | ((uint32_t)(COUT >> 3) << 17) |
                               ((256u >> 4) << 24);
    const bool leader = elect_one();
    const uint64_t desc_hi_bits = ((uint64_t)((128u >> 4) & 0x3FFF) << 32) | (1ull << 46);
    const uint64_t a_desc_base = desc_hi_bits | ((uint64_t)((a_plane_bytes >> 4) & 0x3FFF) << 16);
    const uint64_t b_desc_base = desc_hi_bits | ((uint64_t)((kBPlane >> 4) & 0x3FFF) << 16);
    const uint32_t a_hi32 = (uint32_t)(a_desc_base >> 32), b_hi32 = (uint32_t)(b_desc_base >> 32);
    const uint32_t a_lo_plane16 = (2u * a_plane_bytes) >> 4;
    int it = 0, kb = 0, tc = 0;
    long long w_acc = 0, w_a = 0, w_b = 0;
    const long long t_begin = clock64();
    unsigned long long g_begin = 0;
    if (P.dbg) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g_begin));
    for (int tile = pair_id; tile < P.n_tiles; tile += n_pairs, ++tc) {
      const int buf = tc % n_buf;
      const long long c_acc = P.dbg ? clock64() : 0;
      mbar_wait(acc_empty(buf), ((uint32_t)(tc / n_buf) & 1u) ^ 1u);   // both CTAs' epilogues drained this buffer
      if (P.dbg) w_acc += clock64() - c_acc;
      tc_fence_after();
      const uint32_t d0 = tmem_base + (uint32_t)buf * buf_cols;
      uint32_t any = 0;
      for (int i = 0; i < P.n_stages; ++i, ++it) {
        const int aslot = it % n_a;
        const long long c_a = P.dbg ? clock64() : 0;
        mbar_wait(a_full(aslot), (uint32_t)(it / n_a) & 1u);
        if (P.dbg) w_a += clock64() - c_a;
        const uint32_t a_s = smem_u32(a_base + (size_t)aslot * a_stage_stride);
        const uint32_t a_lo32 = (uint32_t)(a_desc_base & 0xFFFFFFFFu) | ((a_s >> 4) & 0x3FFFu);
        const int nt = P.st_ntaps[i], tap0 = P.st_tap0[i];
        for (int t0 = 0; t0 < nt; t0 += G, ++kb) {
          const int n = min(G, nt - t0);
          const int bslot = kb % n_b;
          const long long c_b = P.dbg ? clock64() : 0;
          mbar_wait(b_full(bslot), (uint32_t)(kb / n_b) & 1u);
          if (P.dbg) w_b += clock64() - c_b;
          tc_fence_after();
          const uint32_t b_s0 = smem_u32(b_base + (size_t)bslot * b_stage_bytes);
          const uint32_t b_lo32 = (uint32_t)(b_desc_base & 0xFFFFFFFFu) | ((b_s0 >> 4) & 0x3FFFu);
          if (leader) {
#pragma unroll 1
            for (int j = 0; j < n; ++j) {
              const uint32_t a00 = a_lo32 + (uint32_t)((int)P.tap_shift[tap0 + t0 + j] - P.lo_shift);
              const uint64_t b_hi = desc64(b_lo32 + (uint32_t)j * (kBTap >> 4), b_hi32);
#pragma unroll
              for (int sb = 0; sb < S; ++sb) {
              const uint32_t a0 = a00 + (uint32_t)sb * 128u;
              const uint32_t d = d0 + (uint32_t)sb * (uint32_t)kDCols;
              const uint64_t a_hi = desc64(a0, a_hi32);
              if (MERGE) {
                constexpr uint32_t idesc_wide = (idesc & ~(0x3Fu << 17)) | ((uint32_t)((2 * COUT) >> 3) << 17);
                const uint64_t a_lo = desc64(a0 + a_lo_plane16, a_hi32);
                const uint64_t b_y = desc64(b_lo32 + (uint32_t)j * (kBTap >> 4) + ((2u * kBPart) >> 4), b_hi32);
                tc_mma_pair(d, a_hi, b_hi, idesc_wide, any);   // region X: [w_hi | w_lo] over the pair
                tc_mma_pair(d, a_lo, b_y, idesc, 1u);          // region Y: w_hi halves
              } else {
                tc_mma_pair(d, a_hi, b_hi, idesc, any);
                if (SPLIT) {
                  const uint64_t a_lo = desc64(a0 + a_lo_plane16, a_hi32);
                  const uint64_t b_lo = desc64(b_lo32 + (uint32_t)j * (kBTap >> 4) + (kBPart >> 4), b_hi32);
                  tc_mma_pair(d, a_hi, b_lo, idesc, 1u);
                  tc_mma_pair(d, a_lo, b_hi, idesc, 1u);
                }
              }
              }
              any = 1u;
            }
            tc_commit_pair(b_empty(bslot));
          }
          any = 1u;
          __syncwarp();
        }
        if (leader) tc_commit_pair(a_empty(aslot));
        __syncwarp();
      }
      if (leader) tc_commit_pair(acc_full(buf));
      __syncwarp();
    }
    if (P.dbg && leader) {
      long long* o = P.dbg + pair_id * 16;
      o[3] = w_acc; o[4] = w_a; o[5] = w_b; o[6] = clock64() - t_begin; o[7] = tc;
      unsigned long long g_end;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g_end));
      unsigned smid;
      asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
      o[10] = (long long)g_begin; o[11] = (long long)g_end; o[12] = smid;
    }
  } else if (warp >= 4) {
    // ===== epilogue (both CTAs): own 128 rows of the accumulator -> bias / residual / ReLU -> global =====
    grid_dependency_wait();     // residual rows and the output tensor belong to earlier launches until they are done
    const int lgrp = warp & 3;
    int tc = 0;
    bool fp16_over = false;
    long long w_full = 0;
    const long long e_begin = clock64();
    constexpr int kStep = MERGE ? 16 : 32, kRows = kStep / 8;
    for (int tile = pair_id; tile < P.n_tiles; tile += n_pairs, ++tc) {
      const int buf = tc % n_buf;
      for (int sb = 0; sb < S; ++sb) {
      const long long q = ((long long)tile * 2 + rank) * (S * 128) + sb * 128 + lgrp * 32 + lane;
      const OutPos o = map_output(q, P.H, P.W, P.n_pos, P.phase_out);
      uint4 rn_h[kRows], rn_l[kRows];
      auto res_fetch = [&](int c0) {
        const uint8_t* g = P.res.base + (long long)(c0 >> 3) * P.res.plane_stride + (P.res.guard + q) * 16;
#pragma unroll
        for (int h = 0; h < kRows; ++h) {
          if (o.valid) {
            rn_h[h] = __ldg(reinterpret_cast<const uint4*>(g + (long long)h * P.res.plane_stride));
            if (SPLIT) rn_l[h] = __ldg(reinterpret_cast<const uint4*>(g + (long long)h * P.res.plane_stride + P.res.lo_offset));
          }
        }
      };
      if (EPI == kEpiResidual) {
        res_fetch(0);
        if (o.valid) {   // pull the rest of this row's residual into L2 while the accumulator is still being computed
          const uint8_t* g = P.res.base + (P.res.guard + q) * 16;
#pragma unroll 4
          for (int pl = kRows; pl < COUT / 8; ++pl) {
            asm volatile("prefetch.global.L2 [%0];" ::"l"(g + (long long)pl * P.res.plane_stride));
            if (SPLIT) asm volatile("prefetch.global.L2 [%0];" ::"l"(g + (long long)pl * P.res.plane_stride + P.res.lo_offset));
          }
        }
      }
      if (sb == 0) {
        const long long c_f = P.dbg ? clock64() : 0;
        mbar_wait(acc_full(buf), (uint32_t)(tc / n_buf) & 1u);
        if (P.dbg) w_full += clock64() - c_f;
        tc_fence_after();
      }
      const FlatTensor& T = P.out[o.tensor];
      const float lo_clamp = P.relu ? 0.0f : -INFINITY;
#pragma unroll 1
      for (int c0 = 0; c0 < COUT; c0 += kStep) {
        uint32_t v[kStep];
        uint32_t v2[MERGE ? kStep : 1];   // D2 = a_hi * w_lo of the merged form
        const uint32_t taddr = tmem_base + ((uint32_t)(lgrp * 32) << 16) + (uint32_t)buf * buf_cols +
                               (uint32_t)(sb * kDCols + c0);
        tc_ld16(taddr, v);
        if (kStep == 32) tc_ld16(taddr + 16, v + 16);
        if (MERGE) tc_ld16(taddr + COUT, v2);
        uint4 rc_h[kRows], rc_l[kRows];
        if (EPI == kEpiResidual) {
#pragma unroll
          for (int h = 0; h < kRows; ++h) { rc_h[h] = rn_h[h]; if (SPLIT) rc_l[h] = rn_l[h]; }
          if (c0 + kStep < COUT) res_fetch(c0 + kStep);
        }
        tc_ld_wait();
        if (o.valid) {
#pragma unroll
          for (int h = 0; h < kRows; ++h) {
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
            if (MODE == UMT_MODE_FP16) {   // a value beyond the half range would be stored as inf
              float mx = 0.f;
#pragma unroll
              for (int e = 0; e < 8; ++e) mx = fmaxf(mx, fabsf(f[e]));
              fp16_over |= (mx > 65504.f);
            }
            store_row8<MODE>(T, (c0 >> 3) + h, o.row, f);
          }
        }
      }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_leader(acc_empty(buf));   // 8 arrivals (4 warps x 2 CTAs) free the buffer
    }
    if (MODE == UMT_MODE_FP16 && P.status && __any_sync(0xFFFFFFFFu, fp16_over) && lane == 0)
      atomicOr(P.status, (unsigned)UMT_STATUS_FP16_RANGE);
    if (P.dbg && leader_cta && warp == 4 && lane == 0) {
      P.dbg[pair_id * 16 + 8] = w_full;
      P.dbg[pair_id * 16 + 9] = clock64() - e_begin;
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // nobody leaves (or frees TMEM) while the pair's MMAs or signals may still touch it
  if (warp == 2) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols) : "memory");
  }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

static int make_map(CUtensorMap* m, int rank, void* base, const cuuint64_t* dims, const cuuint64_t* strides,
                    const cuuint32_t* box) {
  EncodeTiledFn fn = encode_tiled();
  if (!fn) return fail(UMT_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_UINT16, (cuuint32_t)rank, base, dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(UMT_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return UMT_OK;
}

// 3-D view of an activation tensor: {8 channels of a row, rows of a plane, planes (hi then lo)}
static int make_map_a(CUtensorMap* m, const FlatTensor& T, int box_rows, int box_planes) {
  const cuuint64_t planes = (cuuint64_t)(2 * (T.lo_offset / T.plane_stride));
  const cuuint64_t dims[3] = {8, (cuuint64_t)(T.plane_stride / 16), planes};
  const cuuint64_t strides[2] = {16, (cuuint64_t)T.plane_stride};
  const cuuint32_t box[3] = {8, (cuuint32_t)box_rows, (cuuint32_t)box_planes};
  return make_map(m, 3, T.base, dims, strides, box);
}

static int make_map_b(CUtensorMap* m, const uint8_t* blob, size_t bytes, int box_rows) {
  const cuuint64_t dims[2] = {128, (cuuint64_t)(bytes / 256)};
  const cuuint64_t strides[1] = {256};
  const cuuint32_t box[2] = {128, (cuuint32_t)box_rows};
  return make_map(m, 2, const_cast<uint8_t*>(blob), dims, strides, box);
}

// bytes of one tap in one CTA (see the kernel's kBTap)
static size_t pair_tap_bytes(int cout, bool split) {
  const size_t part = (size_t)(cout / 2) * 16;
  return 2 * ((split && cout <= 64) ? 3 * part : (split ? 2 : 1) * part);
}

static size_t pair_smem_bytes(const PairParams& p, int cout, bool split) {
  const size_t bias = ((size_t)cout * 4 + 127) / 128 * 128;
  const size_t a = (((size_t)p.patch_rows * 16 * (split ? 4 : 2)) + 127) & ~(size_t)127;
  const size_t b = (size_t)p.G * pair_tap_bytes(cout, split);
  return 1024 + bias + a * p.n_a + b * p.n_b;
}

static int pair_sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
  }
  return n;
}

template <int COUT, int MODE, int S>
static int launch_pair_s(const PairParams& p, cudaStream_t st) {
  const size_t smem = pair_smem_bytes(p, COUT, MODE == UMT_MODE_FP32_EQUIV);
  void (*kernel)(PairParams) = p.res_epi ? conv_pair_kernel<COUT, MODE, kEpiResidual, S> : conv_pair_kernel<COUT, MODE, kEpiPlain, S>;
  {  // opt-in shared-memory size: once per device and kernel (and when it grows)
    static size_t have[2][64] = {};
    int dev = 0;
    UMT_CUDA_TRY(cudaGetDevice(&dev));
    const int epi = p.res_epi ? 1 : 0;
    if (dev < 0 || dev >= 64 || have[epi][dev] < smem) {
      UMT_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      if (dev >= 0 && dev < 64) have[epi][dev] = smem;
    }
  }
  const int pairs = std::max(1, std::min(p.n_tiles, pair_sm_count()));   // two CTAs per SM = one pair per SM on average
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)(2 * pairs));
  cfg.blockDim = dim3(kPairThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = p.overlap_launch ? 1 : 0;      // (the cluster shape is a compile-time attribute of the kernel)
  UMT_CUDA_TRY(cudaLaunchKernelEx(&cfg, kernel, p));
  return check_launch("conv_pair_kernel");
}

// UMT_CONV_PAIR: 0 = never, 1 = default selection, "32,64,..." = exactly these widths (all modes).
// Read per launch so that tools can alternate (tools/ab_conv.py).
// sub-tiles per CTA that exist as kernels: accumulators of S sub-tiles must fit 256 TMEM columns
template <int COUT, int MODE>
static int launch_pair_t(const PairParams& p, cudaStream_t st) {
  constexpr int kDCols = (MODE == UMT_MODE_FP32_EQUIV && COUT <= 64) ? 2 * COUT : COUT;
  if (p.S == 1) return launch_pair_s<COUT, MODE, 1>(p, st);
  if constexpr (2 * kDCols <= 256) {
    if (p.S == 2) return launch_pair_s<COUT, MODE, 2>(p, st);
  }
  if constexpr (4 * kDCols <= 256) {
    if (p.S == 4) return launch_pair_s<COUT, MODE, 4>(p, st);
  }
  return fail(UMT_ERR_INVALID, "pair conv: S = %d not available for Cout %d", p.S, COUT);
}

int pair_kernel_enabled(int cout, int mode) {
  const char* e = getenv("UMT_CONV_PAIR");
  // default (B200, tools/ab_conv.py): pairs win for 128 / 256 channels in every mode (+14 % / +8 % in the
  // 16-bit modes, +17 % / +4 % in the split mode) and for 64 channels in the split mode (+8 %); the
  // 32-channel layers (and 64 channels in the 16-bit modes) are faster with one CTA per tile, whose
  // four sub-tiles per CTA keep the halo of the 48x48 raster at 20 % of the A traffic
  if (!e || !strcmp(e, "1")) return cout >= 128 || (cout == 64 && mode == UMT_MODE_FP32_EQUIV);
  if (!strcmp(e, "0")) return 0;
  if (!strcmp(e, "all")) return 1;
  char key[16];
  snprintf(key, sizeof key, "%d", cout);
  for (const char* q = e; (q = strstr(q, key)); q += strlen(key))
    if ((q == e || q[-1] == ',') && (q[strlen(key)] == ',' || q[strlen(key)] == 0)) return 1;
  return 0;
}

// Launch one eligible convolution (see finalize_conv_params) through the pair kernel.
int launch_conv_pair(const ConvParams& c, int cout, int mode, cudaStream_t st) {
  const bool split = (mode == UMT_MODE_FP32_EQUIV);
  PairParams p;
  memset(&p, 0, sizeof(p));
  p.H = c.H; p.W = c.W; p.n_pos = c.n_pos;
  // sub-tiles per CTA: the halo of a 128-row tile is 78 % extra A traffic at 48x48 and 41 % at 24x24, so the
  // wide-raster layers take more rows per CTA (UMT_PAIR_S<cout> overrides)
  const int dcols = (split && cout <= 64) ? 2 * cout : cout;
  int S = (cout == 32) ? (split ? 2 : 4) : (cout == 64 && !split) ? 2 : 1;
  {
    const std::string name = "UMT_PAIR_S" + std::to_string(cout);
    if (const char* e = getenv(name.c_str())) S = atoi(e);
  }
  S = (S >= 4) ? 4 : (S >= 2) ? 2 : 1;
  while (S > 1 && S * dcols > 256) S >>= 1;
  p.S = S;
  p.n_tiles = (c.n_pos + 256 * S - 1) / (256 * S);
  p.G = c.pair_G;
  p.lo_shift = c.lo_shift;
  const int need_rows = S * 128 + c.halo_rows;
  p.n_box = (need_rows + 255) / 256;
  p.box_rows = ((need_rows + p.n_box - 1) / p.n_box + 7) / 8 * 8;   // multiple of 8 rows: 128-byte aligned box destinations
  if (p.n_box == 1) {
    // one two-plane box: only the box itself must be 128-byte aligned (rows % 4 == 0).  Keep the plane
    // distance at 64 mod 128 bytes: with a multiple of 128 the two K halves of the A operand start in the
    // same banks and the MMA's operand fetch slows down (measured -20 % on the 128-channel layers)
    p.box_rows = (need_rows + 3) / 4 * 4;
    if (p.box_rows % 8 == 0 && !getenv("UMT_PAIR_NO_SKEW")) p.box_rows += 4;
  }
  p.patch_rows = p.n_box * p.box_rows;
  p.relu = c.relu; p.res_epi = c.res_epi; p.phase_out = c.phase_out;
  p.overlap_launch = c.overlap_launch;
  p.dbg = next_conv_debug_slot();
  p.bias = c.bias;
  p.status = c.status;
  p.res = c.res;
  for (int i = 0; i < 4; ++i) p.out[i] = c.out[i];
  if (c.n_stages > kPairMaxStages) return fail(UMT_ERR_INVALID, "pair conv: %d stages", c.n_stages);
  p.n_stages = c.n_stages;
  int n_taps_total = 0;
  for (int i = 0; i < c.n_stages; ++i) {
    if (c.st_src[i] > 1 || c.tap_src[c.st_tap0[i]] != 0) return fail(UMT_ERR_INVALID, "pair conv: source %d", (int)c.st_src[i]);
    if (!(c.st_ntaps[i] == 1 || c.st_ntaps[i] % p.G == 0))
      return fail(UMT_ERR_INVALID, "pair conv: %d taps per stage with groups of %d", (int)c.st_ntaps[i], p.G);
    p.st_src[i] = c.st_src[i]; p.st_chunk[i] = c.st_chunk[i]; p.st_ntaps[i] = c.st_ntaps[i]; p.st_tap0[i] = c.st_tap0[i];
    n_taps_total += c.st_ntaps[i];
  }
  for (int t = 0; t < kMaxTaps; ++t) p.tap_shift[t] = c.tap_shift[t];
  if (p.box_rows > 256) return fail(UMT_ERR_INVALID, "pair conv: box of %d rows exceeds the TMA limit", p.box_rows);
  // tensor maps (host-side encode, a few hundred nanoseconds each)
  const size_t tap_bytes = pair_tap_bytes(cout, split);
  const size_t blob_bytes = (size_t)n_taps_total * 2 * tap_bytes;
  for (int s = 0; s < 2; ++s) {
    const FlatTensor& T = c.src[s];
    if (!T.base) continue;
    p.guard[s] = T.guard;
    p.planes[s] = (int)(T.lo_offset / T.plane_stride);
    UMT_TRY(make_map_a(&p.mapA[s], T, p.box_rows, p.n_box == 1 ? 2 : 1));
  }
  UMT_TRY(make_map_b(&p.mapB[0], c.pair_weights, blob_bytes, (int)(p.G * tap_bytes / 256)));
  UMT_TRY(make_map_b(&p.mapB[1], c.pair_weights, blob_bytes, (int)(tap_bytes / 256)));
  // pipeline depths from the shared-memory budget (two CTAs per SM)
  const size_t budget = 112 * 1024;
  const size_t a = (((size_t)p.patch_rows * 16 * (split ? 4 : 2)) + 127) & ~(size_t)127;
  const size_t fixed = 1024 + ((size_t)cout * 4 + 127) / 128 * 128;
  const size_t b = (size_t)p.G * tap_bytes;
  int n_a = split ? 3 : 4;   // split mode: deep weight pipelines matter more than deep A pipelines (tools/ab_conv.py: +4..12 % over 4-6 A stages)
  if (const char* e = getenv("UMT_PAIR_NA")) n_a = std::max(2, atoi(e));
  while (n_a > 2 && fixed + a * n_a + b * 2 > budget) --n_a;
  long long nb = ((long long)budget - (long long)fixed - (long long)(a * n_a)) / (long long)b;
  if (nb < 2) return fail(UMT_ERR_INVALID, "pair conv: weight stages do not fit");
  p.n_a = n_a;
  p.n_b = (int)std::min(12LL, nb);
  p.n_buf = (2 * S * dcols <= 256) ? 2 : 1;
  switch (cout * 4 + mode) {
    case 32 * 4 + UMT_MODE_FP32_EQUIV: return launch_pair_t<32, UMT_MODE_FP32_EQUIV>(p, st);
    case 32 * 4 + UMT_MODE_BF16: return launch_pair_t<32, UMT_MODE_BF16>(p, st);
    case 32 * 4 + UMT_MODE_FP16: return launch_pair_t<32, UMT_MODE_FP16>(p, st);
    case 64 * 4 + UMT_MODE_FP32_EQUIV: return launch_pair_t<64, UMT_MODE_FP32_EQUIV>(p, st);
    case 64 * 4 + UMT_MODE_BF16: return launch_pair_t<64, UMT_MODE_BF16>(p, st);
    case 64 * 4 + UMT_MODE_FP16: return launch_pair_t<64, UMT_MODE_FP16>(p, st);
    case 128 * 4 + UMT_MODE_FP32_EQUIV: return launch_pair_t<128, UMT_MODE_FP32_EQUIV>(p, st);
    case 128 * 4 + UMT_MODE_BF16: return launch_pair_t<128, UMT_MODE_BF16>(p, st);
    case 128 * 4 + UMT_MODE_FP16: return launch_pair_t<128, UMT_MODE_FP16>(p, st);
    case 256 * 4 + UMT_MODE_FP32_EQUIV: return launch_pair_t<256, UMT_MODE_FP32_EQUIV>(p, st);
    case 256 * 4 + UMT_MODE_BF16: return launch_pair_t<256, UMT_MODE_BF16>(p, st);
    case 256 * 4 + UMT_MODE_FP16: return launch_pair_t<256, UMT_MODE_FP16>(p, st);
  }
  return fail(UMT_ERR_UNSUPPORTED, "pair conv: Cout %d mode %d", cout, mode);
}

// Repack one layer's weight slabs (encoder.cu pack_slab order: [slab][plane][hi|lo][cout][8]) for the pair
// kernel: taps grouped as the kernel consumes them (groups of G, or single taps), inside a group first the
// rows of CTA 0 then those of CTA 1: [group][half][tap][plane][hi|lo][cout/2][8].  `kind`: 0 = split bf16
// (hi and lo), 1 = bf16 hi only, 2 = fp16 of hi + lo.
void pack_pair_blob(const std::vector<uint16_t>& blob, int cout, const std::vector<int>& stage_ntaps, int G, int kind,
                    uint16_t (*to_half)(float), float (*from_bf)(uint16_t), std::vector<uint16_t>& out) {
  const int parts = (kind == 0) ? 2 : 1, half = cout / 2;
  const bool merge = (kind == 0 && cout <= 64);   // [plane][X: all hi rows (CTA 0) / all lo rows (CTA 1)][Y: own half of the hi rows]
  const size_t slab = (size_t)4 * cout * 8;
  out.clear();
  int k = 0;
  for (int nt : stage_ntaps)
    for (int t0 = 0; t0 < nt; t0 += G) {
      const int n = std::min(G, nt - t0);
      for (int r = 0; r < 2; ++r)
        for (int j = 0; j < n; ++j) {
          const uint16_t* s = &blob[(size_t)(k + j) * slab];
          if (merge) {
            for (int pl = 0; pl < 2; ++pl) {
              for (int co = 0; co < cout; ++co)
                for (int e = 0; e < 8; ++e) out.push_back(s[(((size_t)pl * 2 + r) * cout + co) * 8 + e]);
              for (int co = r * half; co < (r + 1) * half; ++co)
                for (int e = 0; e < 8; ++e) out.push_back(s[(((size_t)pl * 2 + 0) * cout + co) * 8 + e]);
            }
            continue;
          }
          for (int pl = 0; pl < 2; ++pl)
            for (int part = 0; part < parts; ++part)
              for (int co = r * half; co < (r + 1) * half; ++co)
                for (int e = 0; e < 8; ++e) {
                  const uint16_t hi = s[(((size_t)pl * 2 + 0) * cout + co) * 8 + e];
                  const uint16_t lo = s[(((size_t)pl * 2 + 1) * cout + co) * 8 + e];
                  out.push_back(kind == 0 ? (part == 0 ? hi : lo) : kind == 1 ? hi : to_half(from_bf(hi) + from_bf(lo)));
                }
        }
      k += n;
    }
}

}  // namespace umt
