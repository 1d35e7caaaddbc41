// conv_flat.cu — "window" kernels for stride-1 convolutions on the flat grid (layout: gemm_tc.cu, "flat-grid
// convolution"; reference: caffe/src/caffe/layers/base_conv_layer.cpp:243-307 forward_cpu_gemm /
// backward_cpu_gemm + caffe/src/caffe/util/im2col.cu:9-58).
//
// Why a second kernel family.  A TF32 tcgen05.mma of 128 x N x 8 takes N/2 clocks, and its operands are
// 4-byte: a 128 x 256 tile fed chunk by chunk needs (128 + 256) x 128 B per 512 clocks = 96 B/clk per SM,
// while the L2 delivers ~42 B/clk per SM with all 148 SMs pulling (B300_MICROARCH.md "LTS throughput cap").
// The chunk-per-tap kernels of gemm_tc.cu are therefore L2-bound at < 45 % of the tensor peak.  On the flat
// grid a filter tap is a pure row shift of the packed activations, so here
//   * one TMA load per 32-channel block brings the rows [q0, q0 + MT*128 + span) of the grid ONCE (span =
//     (kh-1)*Wp + kw-1); every tap reads it through a shared-memory descriptor whose start address is
//     advanced by (ky*Wp + kx) rows — kh*kw fewer activation bytes through the L2;
//   * MT (1, 2 or 4) accumulators of 128 rows x n_mma columns live in TMEM side by side (MT * n_mma <= 512
//     columns), so every filter chunk that arrives is used by MT MMAs instead of one.
// Feed rate for AlexNet's conv3 (3x3, 256 -> 384): 34 B/clk per SM instead of 96.
//
//   warps 0-7  epilogue: TMEM -> registers -> NCHW (bias + ReLU fused), warp w owns TMEM lanes 32(w%4)..+31
//   warp 8     allocates TMEM, issues tcgen05.mma (one thread)
//   warp 9     issues the TMA loads (one thread)
#include <cuda.h>

#include <cstdio>
#include <cstdlib>

#include "gemm_ops.cuh"
#include "kernels.hpp"
#include "tc_prims.cuh"

namespace sn {
using namespace tc;

CUtensorMap tc_make_tmap(const float* base, uint64_t inner, uint64_t outer, uint64_t pitch_bytes, uint32_t box_outer, bool mn_major);

constexpr int kWEpiWarps = 8;
constexpr int kWThreads = (kWEpiWarps + 2) * 32;
constexpr int kWMaxStages = 8;
constexpr size_t kWSmemLimit = 227 * 1024;

struct WinArgs {
  int Ptot;                 // grid positions = rows of D
  int Wp, kw, T;            // tap t -> row shift (t / kw) * Wp + t % kw
  int row_off;              // TMA row coordinate of window row 0 of the tile that starts at grid position 0
  int MT, n_mma;            // accumulators per CTA, columns of each
  int ncb, red, slots;      // 32-channel blocks of the reduction, its true length, slot pitch of one tap in the filter
  int a_chan_group;         // channel coordinate of group g in the activation tensor = g * a_chan_group
  int b_row_group;          // row coordinate of group g in the filter tensor = g * b_row_group
  int box_rows, nbox;       // a window is nbox TMA boxes of box_rows grid rows
  int nstage;               // filter pipeline depth
};
// m = grid position, n = channel in group; NCHW destination with Hv x Wv valid positions per cell
struct WinEpi {
  float* dst; const float* bias; int Hv, Wv, Ctot, Cgrp, relu; FastDiv dGp, dWp;
};

__global__ void __launch_bounds__(kWThreads, 1)
conv_window_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const WinArgs a, const WinEpi ep) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t win_full[2], win_empty[2], b_full[kWMaxStages], b_empty[kWMaxStages], tmem_full;
  __shared__ uint32_t tmem_base_smem;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q0 = blockIdx.x * a.MT * 128, n0 = blockIdx.y * a.n_mma, grp = blockIdx.z;
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t win_bytes = (uint32_t)(a.nbox * a.box_rows) * 128u;      // multiple of 1024 (box_rows % 8 == 0)
  const uint32_t b_bytes = (uint32_t)a.n_mma * 128u;                      // multiple of 2048 (n_mma % 16 == 0)
  uint8_t* bst = smem + 2 * win_bytes;

  if (tid == 0) {
    for (int i = 0; i < 2; ++i) { mbar_init(&win_full[i], 1u); mbar_init(&win_empty[i], 1u); }
    for (int i = 0; i < a.nstage; ++i) { mbar_init(&b_full[i], 1u); mbar_init(&b_empty[i], 1u); }
    mbar_init(&tmem_full, 1u);
    fence_barrier_init();
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == kWEpiWarps) tmem_alloc<512>(&tmem_base_smem);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = tmem_base_smem;

  if (warp < kWEpiWarps) {
    // ---------------- epilogue ----------------
    mbar_wait(&tmem_full, 0);
    tc_fence_after_sync();
    const int quarter = warp & 3, half = warp >> 2, chan0 = grp * ep.Cgrp;
    const size_t plane = (size_t)ep.Hv * ep.Wv;
    for (int i = 0; i < a.MT; ++i) {
      const int m = q0 + i * 128 + 32 * quarter + lane;
      bool ok = m < a.Ptot;
      float* p = ep.dst;
      if (ok) {
        uint32_t img, rem, h, w;
        ep.dGp.divmod((uint32_t)m, img, rem);
        ep.dWp.divmod(rem, h, w);
        ok = (int)h < ep.Hv && (int)w < ep.Wv;
        p = ep.dst + (((size_t)img * ep.Ctot + chan0) * ep.Hv + h) * ep.Wv + w;
      }
      for (int c0 = 16 * half; c0 < a.n_mma; c0 += 32) {
        uint32_t v[16];
        tmem_ld_32x32b_x16(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(i * a.n_mma + c0), v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int n = n0 + c0 + j;
          if (ok && n < ep.Cgrp) {
            float acc = __uint_as_float(v[j]);
            if (ep.bias) acc += __ldg(ep.bias + chan0 + n);
            if (ep.relu) acc = fmaxf(acc, 0.f);
            p[(size_t)n * plane] = acc;
          }
        }
      }
    }
  } else if (warp == kWEpiWarps) {
    // ---------------- MMA issuer ----------------
    if (lane == 0) {
      const uint32_t idesc = make_idesc_tf32(128, a.n_mma, 0, 0);
      const uint32_t win_addr = smem_u32(smem), b_addr0 = smem_u32(bst);
      int it = 0;
      for (int cb = 0; cb < a.ncb; ++cb) {
        const int wb = cb & 1;
        mbar_wait(&win_full[wb], (cb >> 1) & 1);
        tc_fence_after_sync();
        const int left = a.red - cb * 32, ks = left >= 32 ? 4 : (left + 7) / 8;
        int ky = 0, kx = 0;
        for (int tap = 0; tap < a.T; ++tap, ++it) {
          const int s = it % a.nstage;
          mbar_wait(&b_full[s], (it / a.nstage) & 1);
          tc_fence_after_sync();
          const uint32_t b_addr = b_addr0 + (uint32_t)s * b_bytes;
          const uint32_t shift = (uint32_t)(ky * a.Wp + kx);
          // The 128-byte swizzle is a function of the absolute shared-memory address (bits 4-6 ^= bits 7-9), for the
          // TMA write and for the MMA read alike, so a descriptor may start at any row of the window: the
          // base-offset field stays 0 (verified on B200 with tools/debug_window.py).
          for (int i = 0; i < a.MT; ++i) {
            const uint32_t a_addr = win_addr + (uint32_t)wb * win_bytes + ((uint32_t)(i * 128) + shift) * 128u;
            for (int j = 0; j < ks; ++j) {
              const uint64_t adesc = make_smem_desc(a_addr + j * 32, 16, 1024, kLayoutSW128);
              const uint64_t bdesc = make_smem_desc(b_addr + j * 32, 16, 1024, kLayoutSW128);
              mma_tf32(tmem_base + (uint32_t)(i * a.n_mma), adesc, bdesc, idesc, (cb | tap | j) != 0 ? 1u : 0u);
            }
          }
          mma_commit(&b_empty[s]);
          if (++kx == a.kw) { kx = 0; ++ky; }
        }
        mma_commit(&win_empty[wb]);
      }
      mma_commit(&tmem_full);
    }
  } else {
    // ---------------- TMA issuer ----------------
    if (lane == 0) {
      int it = 0;
      const int a_chan0 = grp * a.a_chan_group, b_row0 = grp * a.b_row_group + n0;
      for (int cb = 0; cb < a.ncb; ++cb) {
        const int wb = cb & 1;
        mbar_wait(&win_empty[wb], ((cb >> 1) & 1) ^ 1);
        mbar_arrive_expect_tx(&win_full[wb], win_bytes);
        for (int b = 0; b < a.nbox; ++b)
          tma_load_2d(smem + (size_t)wb * win_bytes + (size_t)b * a.box_rows * 128, &tmA, &win_full[wb], a_chan0 + cb * 32,
                      q0 + a.row_off + b * a.box_rows);
        for (int tap = 0; tap < a.T; ++tap, ++it) {
          const int s = it % a.nstage;
          mbar_wait(&b_empty[s], ((it / a.nstage) & 1) ^ 1);
          mbar_arrive_expect_tx(&b_full[s], b_bytes);
          tma_load_2d(bst + (size_t)s * b_bytes, &tmB, &b_full[s], tap * a.slots + cb * 32, b_row0);
        }
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == kWEpiWarps) tmem_dealloc<512>(tmem_base);
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
static int g_force_mt = 0;
void tcw_set_debug(int what, int value) {
  if (what == 1) g_force_mt = value;
}

struct WinPlan { int MT, n_mma, nstage, box_rows, nbox; size_t smem; bool ok; };

static int win_n_mma(int N) {
  const int tiles = (N + 255) / 256;
  int w = (N + tiles - 1) / tiles;
  w = (w + 15) / 16 * 16;
  return w < 16 ? 16 : w;
}
// Pick the number of accumulators: fewest estimated clocks over all waves.  Per tile: MMA time MT*T*ksteps*n_mma/2,
// L2 time bytes / 40, plus the serial prologue / epilogue.
static WinPlan plan_window(int Ptot, int N, int groups, int T, int span, int red) {
  WinPlan best{};
  best.ok = false;
  const int n_mma = win_n_mma(N), n_tiles = (N + n_mma - 1) / n_mma, ncb = (red + 31) / 32;
  int ksteps = 0;
  for (int cb = 0; cb < ncb; ++cb) { const int left = red - cb * 32; ksteps += left >= 32 ? 4 : (left + 7) / 8; }
  double best_cost = 0;
  for (int MT = 4; MT >= 1; MT >>= 1) {
    if (MT * n_mma > 512) continue;
    if (g_force_mt && MT != g_force_mt) continue;
    const int win_rows = MT * 128 + span;
    const int nbox = (win_rows + 255) / 256;
    const int box_rows = ((win_rows + nbox - 1) / nbox + 7) / 8 * 8;
    const size_t win_bytes = (size_t)nbox * box_rows * 128, b_bytes = (size_t)n_mma * 128;
    const size_t fixed = 2 * win_bytes + 1024 + 512;             // + alignment slack + static shared (barriers)
    if (fixed + 3 * b_bytes > kWSmemLimit) continue;
    int nstage = (int)((kWSmemLimit - fixed) / b_bytes);
    if (nstage > kWMaxStages) nstage = kWMaxStages;
    const long long tiles = (long long)((Ptot + MT * 128 - 1) / (MT * 128)) * n_tiles * groups;
    const long long waves = (tiles + 147) / 148;
    const double t_mma = (double)MT * T * ksteps * n_mma / 2.0;
    const double t_l2 = ((double)ncb * win_bytes + (double)ncb * T * b_bytes) / 40.0;
    const double t_epi = (double)MT * 128 * n_mma * 4 / 48.0;
    const double cost = (double)waves * ((t_mma > t_l2 ? t_mma : t_l2) + t_epi + 4000.0);
    if (!best.ok || cost < best_cost) {
      best = WinPlan{MT, n_mma, nstage, box_rows, nbox, 2 * win_bytes + (size_t)nstage * b_bytes + 1024, true};
      best_cost = cost;
    }
  }
  return best;
}

static void launch_window(cudaStream_t s, const CUtensorMap& tmA, const CUtensorMap& tmB, const WinArgs& a, const WinEpi& ep,
                          const WinPlan& pl, int N, int groups) {
  static bool attr_done = false;
  if (!attr_done) {
    SN_CUDA(cudaFuncSetAttribute(conv_window_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWSmemLimit - 1024));
    attr_done = true;
  }
  dim3 grid((a.Ptot + a.MT * 128 - 1) / (a.MT * 128), (N + a.n_mma - 1) / a.n_mma, groups);
  conv_window_kernel<<<grid, kWThreads, pl.smem, s>>>(tmA, tmB, a, ep);
  count_launch();
  SN_CUDA(cudaGetLastError());
}

static int slots_of(int ch) { return (ch + 31) / 32 * 32; }

bool tcw_supported(const ConvGeom& g) {
  if (!tc_flat_supported(g)) return false;
  const FlatGeom f = tc_flat_geom(g);
  const int span = (g.kh - 1) * f.Wp + g.kw - 1;
  return plan_window(g.N * f.Gp, g.Og(), g.group, g.kh * g.kw, span, g.Cg()).ok &&
         plan_window(g.N * f.Gp, g.Cg(), g.group, g.kh * g.kw, span, g.Og()).ok;
}

// `xflat`: packed input (tc_flat_pack, input side); `wf`: tc_flat_pack_filter(dgrad = false)
void tcw_conv_fwd(cudaStream_t s, const ConvGeom& g, const float* xflat, const float* wf, const float* bias, float* y, bool relu) {
  const FlatGeom f = tc_flat_geom(g);
  const int Cg = g.Cg(), Og = g.Og(), T = g.kh * g.kw, Ptot = g.N * f.Gp, slots = slots_of(Cg), kp = T * slots;
  const int span = (g.kh - 1) * f.Wp + g.kw - 1;
  const WinPlan pl = plan_window(Ptot, Og, g.group, T, span, Cg);
  if (!pl.ok) fatal("window convolution: no tile plan for this layer (forward)");
  const CUtensorMap tmA = tc_make_tmap(xflat, (uint64_t)g.C, tc_flat_rows(g), (uint64_t)g.C * 4, (uint32_t)pl.box_rows, false);
  const CUtensorMap tmB = tc_make_tmap(wf, (uint64_t)kp, (uint64_t)g.Cout, (uint64_t)kp * 4, (uint32_t)pl.n_mma, false);
  WinArgs a{Ptot, f.Wp, g.kw, T, f.lead - (g.ph * f.Wp + g.pw), pl.MT, pl.n_mma, slots / 32, Cg, slots, Cg, Og,
            pl.box_rows, pl.nbox, pl.nstage};
  WinEpi ep{y, bias, g.Ho, g.Wo, g.Cout, Og, relu ? 1 : 0, FastDiv(f.Gp), FastDiv(f.Wp)};
  launch_window(s, tmA, tmB, a, ep, pl, Og, g.group);
}
// `dyflat`: packed top gradient (tc_flat_pack, output side); `wt`: tc_flat_pack_filter(dgrad = true)
void tcw_conv_dgrad(cudaStream_t s, const ConvGeom& g, const float* dyflat, const float* wt, float* dx) {
  const FlatGeom f = tc_flat_geom(g);
  const int Cg = g.Cg(), Og = g.Og(), T = g.kh * g.kw, Ptot = g.N * f.Gp, slots = slots_of(Og), kp = T * slots;
  const int span = (g.kh - 1) * f.Wp + g.kw - 1;
  const WinPlan pl = plan_window(Ptot, Cg, g.group, T, span, Og);
  if (!pl.ok) fatal("window convolution: no tile plan for this layer (backward data)");
  const CUtensorMap tmA = tc_make_tmap(dyflat, (uint64_t)g.Cout, tc_flat_rows(g), (uint64_t)g.Cout * 4, (uint32_t)pl.box_rows, false);
  const CUtensorMap tmB = tc_make_tmap(wt, (uint64_t)kp, (uint64_t)g.C, (uint64_t)kp * 4, (uint32_t)pl.n_mma, false);
  WinArgs a{Ptot, f.Wp, g.kw, T, f.lead - ((g.kh - 1 - g.ph) * f.Wp + (g.kw - 1 - g.pw)), pl.MT, pl.n_mma, slots / 32, Og, slots,
            Og, Cg, pl.box_rows, pl.nbox, pl.nstage};
  WinEpi ep{dx, nullptr, g.H, g.W, g.C, Cg, 0, FastDiv(f.Gp), FastDiv(f.Wp)};
  launch_window(s, tmA, tmB, a, ep, pl, Cg, g.group);
}

}  // namespace sn
