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
#include <array>
#include <map>
#include <mutex>
#include <vector>

#include "gemm_ops.cuh"
#include "kernels.hpp"
#include "tc_prims.cuh"

namespace sn {
using namespace tc;

CUtensorMap tc_make_tmap(const float* base, uint64_t inner, uint64_t outer, uint64_t pitch_bytes, uint32_t box_outer, bool mn_major);

constexpr int kWEpiWarps = 8;
// One thread issuing every tcgen05.mma is latency-bound on its own instruction stream (~150-250 dependent uniform-datapath
// instructions per tap at ~5 clocks each: measured 900-1250 clocks per tap where the tensor pipe needs 512-768).  Up to
// four warps issue instead, one accumulator (forward / backward-data) or every fourth tap (backward-filter) each.
constexpr int kWMmaWarps = 4;
constexpr int kWTmaWarp = kWEpiWarps + kWMmaWarps;
constexpr int kWThreads = (kWEpiWarps + kWMmaWarps + 1) * 32;
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
  long long* stamps;        // development aid: 8 clock64 stamps per CTA (null = off)
  // --- plain matrix products (inner-product layers run as 1-tap "convolutions" over the batch rows) ---
  int nwin;                 // window buffers (2 for convolutions: a window lives for T taps; up to 8 when T = 1)
  int splits, cb_per_split; // the reduction's channel blocks are cut into `splits` ranges (blockIdx.z = group * splits + split)
  int b_mn;                 // B operand stored [reduction][n] (n contiguous): MN-major tiles of 32 n x 32 rows per TMA box
  // --- persistent CTAs: cluster c works on tiles c, c + #clusters, ...; tile t = (m unit, n tile, z) with the m unit fastest ---
  int m_units, n_tiles, total_tiles;   // m unit = MT*128 rows per CTA, one per CTA of the cluster
  int tpg;                  // taps per pipeline stage: one barrier hand-shake serves tpg filter chunks (a stage = tpg * b_bytes)
  int acc_stages;           // 2 when two accumulator sets fit in TMEM (MT * n_mma <= 256): the epilogue of tile k overlaps the MMAs of tile k+1
  int b_resident;           // the launch has ONE filter tile (one group, one column tile, no split) and all its chunks fit in the
                            // stage buffers (nstage == ncb * ceil(T / tpg)): they are loaded for the cluster's first tile and stay;
                            // later tiles stream only their windows (AlexNet conv1: 83 KB of filter per tile left the L2 -> SM path)
};
// m = grid position, n = channel in group; NCHW destination with Hv x Wv valid positions per cell.
// ld > 0 selects the matrix epilogue instead: dst[split][m][ld] row-major (bias / ReLU only when splits == 1).
struct WinEpi {
  float* dst; const float* bias; int Hv, Wv, Ctot, Cgrp, relu; FastDiv dGp, dWp;
  int ld; size_t split_stride;
  const float* gate;        // backward-data only: same shape as dst; dst = (gate > 0) ? value : 0  (the backward of an in-place
                            // ReLU below this layer, relu_layer.cu:35-41, applied while the gradient is written)
  // forward only (nullable): the flat grid of the NEXT window convolution, which reads this layer's output.  The epilogue
  // writes every value a second time there, rounded to tf32, channels-last (64 contiguous bytes per thread):
  // flat[(flat_lead + img * flat_Gp + h * flat_Wp + w) * flat_C + channel] — the consumer's pack kernel has nothing left to do.
  float* flat; int flat_lead, flat_Gp, flat_Wp, flat_C;
};

// CG = 1: one CTA per tile.  CG = 2: a CTA pair (cluster of 2 along x) runs every MMA with M = 256: each CTA keeps its
// own grid rows (window + accumulators) and HALF of the filter rows of the chunk, so per SM the filter bytes through the
// L2 and the shared-memory bytes the tensor core reads per MMA are halved (the TF32 MMA of one CTA is otherwise bound by
// shared-memory bandwidth: (128 + N) x 32 B per N/2 clocks).
template <int CG>
__global__ void __launch_bounds__(kWThreads, 1)
conv_window_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const WinArgs a, const WinEpi ep) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t win_full[kWMaxStages], win_empty[kWMaxStages], b_full[kWMaxStages], b_empty[kWMaxStages], tmem_full[2], tmem_empty[2];
  __shared__ uint32_t tmem_base_smem;
  SN_GRID_DEP_LAUNCH();

  const long long t_enter = clock64();
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rank = CG == 2 ? (int)cluster_ctarank() : 0;
  const int cluster_id = blockIdx.x / CG, n_clusters = gridDim.x / CG;
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t win_bytes = (uint32_t)(a.nbox * a.box_rows) * 128u;      // multiple of 1024 (box_rows % 8 == 0)
  const int nb_local = a.n_mma / CG;                                      // filter rows this CTA stages per chunk
  const uint32_t b_bytes = (uint32_t)nb_local * 128u;                     // multiple of 1024 (n_mma % 16 == 0)
  uint8_t* bst = smem + a.nwin * win_bytes;

  if (tid == 0) {
    // every accumulator has its own issuing warp: MT commits free a stage / complete the tile
    for (int i = 0; i < a.nwin; ++i) { mbar_init(&win_full[i], 1u); mbar_init(&win_empty[i], (uint32_t)a.MT); }
    for (int i = 0; i < a.nstage; ++i) { mbar_init(&b_full[i], 1u); mbar_init(&b_empty[i], (uint32_t)a.MT); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], (uint32_t)a.MT);
      mbar_init(&tmem_empty[i], (uint32_t)(CG * kWEpiWarps));   // every epilogue warp of the cluster, once per tile (rank 0's barrier counts)
    }
    fence_barrier_init();
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == kWEpiWarps) {
    if (CG == 2) tmem_alloc_pair<512>(&tmem_base_smem);
    else tmem_alloc<512>(&tmem_base_smem);
  }
  tc_fence_before_sync();
  if (CG == 2) cluster_sync_all();
  else __syncthreads();
  tc_fence_after_sync();
  SN_GRID_DEP_WAIT();                        // barriers and TMEM are set up; everything below touches the operands / outputs
  const uint32_t tmem_base = tmem_base_smem;
  long long* stamp = a.stamps ? a.stamps + 8 * (size_t)blockIdx.x : nullptr;      // stamps describe the CTA's FIRST tile
  if (stamp && tid == 0) { stamp[0] = t_enter; stamp[1] = clock64(); }
  // tile t -> rows q0.., columns n0.., group, reduction range
  struct Tile { int q0, n0, grp, split, cb_begin, cb_end; };
  auto decode = [&](int t) {
    Tile tl;
    const int mu = t % a.m_units, r = t / a.m_units, nt = r % a.n_tiles, z = r / a.n_tiles;
    tl.q0 = (mu * CG + rank) * a.MT * 128;
    tl.n0 = nt * a.n_mma;
    tl.grp = z / a.splits;
    tl.split = z % a.splits;
    tl.cb_begin = tl.split * a.cb_per_split;
    tl.cb_end = min(a.ncb, tl.cb_begin + a.cb_per_split);       // never empty (host)
    return tl;
  };

  if (warp < kWEpiWarps) {
    // ---------------- epilogue ----------------
    const int quarter = warp & 3, half = warp >> 2;
    const size_t plane = (size_t)ep.Hv * ep.Wv;
    int ti = 0;
    for (int t = cluster_id; t < a.total_tiles; t += n_clusters, ++ti) {
    const Tile tl = decode(t);
    const int q0 = tl.q0, n0 = tl.n0, split = tl.split, chan0 = tl.grp * ep.Cgrp;
    const int buf = ti % a.acc_stages;
    const uint32_t tmem_acc = tmem_base + (uint32_t)buf * 256u;
    mbar_wait(&tmem_full[buf], (uint32_t)(ti / a.acc_stages) & 1u);
    tc_fence_after_sync();
    if (stamp && tid == 0 && t == cluster_id) stamp[5] = clock64();
    const int ncols = min(a.n_mma, ep.Cgrp - n0);                 // valid columns of this tile (multiple of 16)
    if (ep.ld > 0) {
      // matrix epilogue: row m of D goes to dst[split][m][n0 ..] (16 consecutive floats per TMEM load)
      const bool plain = a.splits == 1;
      for (int i = 0; i < a.MT; ++i) {
        const int m = q0 + i * 128 + 32 * quarter + lane;
        float* p = ep.dst + (size_t)split * ep.split_stride + (size_t)min(m, a.Ptot - 1) * ep.ld + n0;
        for (int c0 = 16 * half; c0 < ncols; c0 += 32) {
          uint32_t v[16];
          tmem_ld_32x32b_x16(tmem_acc + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(i * a.n_mma + c0), v);
          tmem_ld_wait();
          if (m < a.Ptot) {
#pragma unroll
            for (int j = 0; j < 16; j += 4) {
              if (c0 + j >= ncols) break;                       // ncols is a multiple of 4 (N % 4 == 0)
              float4 o = make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
              if (plain && ep.bias) {
                const float* bp = ep.bias + n0 + c0 + j;
                o.x += __ldg(bp); o.y += __ldg(bp + 1); o.z += __ldg(bp + 2); o.w += __ldg(bp + 3);
              }
              if (plain && ep.relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
              *(float4*)(p + c0 + j) = o;
            }
          }
        }
      }
    } else
    for (int c0 = 16 * half; c0 < ncols; c0 += 32) {
      float bias[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) bias[j] = ep.bias ? __ldg(ep.bias + chan0 + n0 + c0 + j) : 0.f;
      for (int i = 0; i < a.MT; ++i) {
        uint32_t v[16];
        tmem_ld_32x32b_x16(tmem_acc + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(i * a.n_mma + c0), v);
        const int m = q0 + i * 128 + 32 * quarter + lane;
        bool ok = m < a.Ptot;
        float* p = ep.dst;
        if (ok) {
          uint32_t img, rem, h, w;
          ep.dGp.divmod((uint32_t)m, img, rem);
          ep.dWp.divmod(rem, h, w);
          ok = (int)h < ep.Hv && (int)w < ep.Wv;
          p = ep.dst + ((((size_t)img * ep.Ctot + chan0 + n0 + c0) * ep.Hv + h) * ep.Wv + w);
        }
        tmem_ld_wait();
        if (ok && ep.gate) {
          const float* gp = ep.gate + (p - ep.dst);
          float gv[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) gv[j] = __ldg(gp + (size_t)j * plane);
#pragma unroll
          for (int j = 0; j < 16; ++j) p[(size_t)j * plane] = __uint_as_float(v[j]) * (gv[j] > 0.f ? 1.f : 0.f);   // relu_g() of kernels.cu
        } else if (ok) {
          float o[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            float acc = __uint_as_float(v[j]) + bias[j];
            if (ep.relu) acc = fmaxf(acc, 0.f);
            p[(size_t)j * plane] = acc;
            o[j] = acc;
          }
          if (ep.flat) {
            uint32_t img, rem, h, w;
            ep.dGp.divmod((uint32_t)m, img, rem);
            ep.dWp.divmod(rem, h, w);
            float4* f = reinterpret_cast<float4*>(ep.flat + ((size_t)ep.flat_lead + (size_t)img * ep.flat_Gp + (size_t)h * ep.flat_Wp + w) * ep.flat_C +
                                                  chan0 + n0 + c0);
#pragma unroll
            for (int j = 0; j < 16; j += 4)
              f[j >> 2] = make_float4(to_tf32_exact(o[j]), to_tf32_exact(o[j + 1]), to_tf32_exact(o[j + 2]), to_tf32_exact(o[j + 3]));
          }
        }
      }
    }
    // this warp has read its share of the accumulators: the next tile's MMAs may overwrite them
    tc_fence_before_sync();
    __syncwarp();
    if (lane == 0) {
      if (stamp && warp == 0 && t == cluster_id) stamp[6] = clock64();
      if (CG == 2) mbar_arrive_rank0(&tmem_empty[buf]);
      else mbar_arrive(&tmem_empty[buf]);
    }
    }
  } else if (warp < kWTmaWarp) {
    // ---------------- MMA issuers (rank 0 of a pair): warp mi owns accumulator mi; the whole warp walks the pipeline,
    // one elected lane issues ----------------
    const int mi = warp - kWEpiWarps;
    if (rank == 0 && mi < a.MT) {
      const uint32_t idesc = make_idesc_tf32(128 * CG, a.n_mma, 0, a.b_mn);
      // descriptor = {lo: start >> 4 | LBO << 16, hi: SBO | version 1 | layout}; only lo moves.
      // K-major (SWIZZLE_128B): LBO 16 B (unused), SBO 1024 B, +32 B per k-step of 8.
      // MN-major B (SWIZZLE_128B_BASE32B): 32-column blocks 4096 B apart (LBO), 4-row atoms 512 B apart (SBO), +1024 B per k-step.
      const uint32_t desc_hi = (uint32_t)(make_smem_desc(0, 16, 1024, kLayoutSW128) >> 32);
      const uint32_t b_hi = a.b_mn ? (uint32_t)(make_smem_desc(0, 4096, 512, kLayoutSW128Base32) >> 32) : desc_hi;
      const uint32_t win_lo = (smem_u32(smem) >> 4) | (1u << 16), b_lo0 = (smem_u32(bst) >> 4) | (a.b_mn ? (4096u >> 4) << 16 : 1u << 16);
      const uint32_t win_step = win_bytes >> 4, b_step = b_bytes >> 4, b_kstep = a.b_mn ? 64u : 2u;
      int s = 0, wb = 0;
      uint32_t sphase = 0, wphase = 0;
      long long wait_win = 0, wait_b = 0;
      int ti = 0;
      for (int t = cluster_id; t < a.total_tiles; t += n_clusters, ++ti) {
      const Tile tl = decode(t);
      const int cb_begin = tl.cb_begin, cb_end = tl.cb_end;
      int it = 0;
      const int buf = ti % a.acc_stages;
      // the epilogue has drained what this accumulator set held before (passes at once the first time round)
      mbar_wait(&tmem_empty[buf], ((uint32_t)(ti / a.acc_stages) & 1u) ^ 1u);
      tc_fence_after_sync();
      for (int cb = cb_begin; cb < cb_end; ++cb) {
        long long tw = (stamp && mi == 0) ? clock64() : 0;
        mbar_wait(&win_full[wb], wphase);
        if (stamp && mi == 0) wait_win += clock64() - tw;
        const int left = a.red - cb * 32, ks = left >= 32 ? 4 : (left + 7) / 8;
        uint32_t shift_row = 0;                                     // (ky * Wp) in rows
        int kx = 0;
        for (int tap0 = 0; tap0 < a.T; tap0 += a.tpg) {
          const int ng = min(a.tpg, a.T - tap0);                   // filter chunks in this stage
          tw = (stamp && mi == 0) ? clock64() : 0;
          if (!a.b_resident || ti == 0) {                           // resident filter: the chunks of the first tile stay valid
            mbar_wait(&b_full[s], sphase);
            tc_fence_after_sync();
          }
          if (stamp && mi == 0) wait_b += clock64() - tw;
          // The 128-byte swizzle is a function of the absolute shared-memory address (bits 4-6 ^= bits 7-9), for the
          // TMA write and for the MMA read alike, so a descriptor may start at any row of the window: the
          // base-offset field stays 0 (verified on B200 with tools/debug_window.py).
          if (elect_one_sync()) {
            const uint32_t d = tmem_base + (uint32_t)buf * 256u + (uint32_t)(mi * a.n_mma);
            uint32_t b_lo = b_lo0 + (uint32_t)s * b_step * (uint32_t)a.tpg, sr = shift_row;
            int kxl = kx, itl = it;
            for (int g = 0; g < ng; ++g, ++itl, b_lo += b_step) {
              const uint32_t a_lo = win_lo + (uint32_t)wb * win_step + (sr + (uint32_t)kxl) * 8u + (uint32_t)(mi * 1024);   // 128 B per row >> 4
#pragma unroll 4
              for (int j = 0; j < ks; ++j) {
                if (CG == 2) mma_tf32_lohi2_pair(d, a_lo + (uint32_t)(2 * j), desc_hi, b_lo + b_kstep * (uint32_t)j, b_hi, idesc, (itl | j) != 0 ? 1u : 0u);
                else mma_tf32_lohi2(d, a_lo + (uint32_t)(2 * j), desc_hi, b_lo + b_kstep * (uint32_t)j, b_hi, idesc, (itl | j) != 0 ? 1u : 0u);
              }
              if (++kxl == a.kw) { kxl = 0; sr += (uint32_t)a.Wp; }
            }
            const bool last_tap = tap0 + ng == a.T, last = last_tap && cb == cb_end - 1;
            if (CG == 2) {
              if (!a.b_resident) mma_commit_pair(&b_empty[s]);
              if (last_tap) mma_commit_pair(&win_empty[wb]);
              if (last) mma_commit_pair(&tmem_full[buf]);
            } else {
              if (!a.b_resident) mma_commit(&b_empty[s]);
              if (last_tap) mma_commit(&win_empty[wb]);
              if (last) mma_commit(&tmem_full[buf]);
            }
          }
          __syncwarp();
          for (int g = 0; g < ng; ++g) {                           // every lane tracks the tap position
            ++it;
            if (++kx == a.kw) { kx = 0; shift_row += (uint32_t)a.Wp; }
          }
          if (++s == a.nstage) { s = 0; sphase ^= 1u; }
        }
        if (++wb == a.nwin) { wb = 0; wphase ^= 1u; }
      }
      if (stamp && lane == 0 && mi == 0 && t == cluster_id) { stamp[2] = wait_win; stamp[3] = wait_b; stamp[4] = clock64(); }
      }
    }
  } else {
    // ---------------- TMA issuer (both CTAs of a pair; completion bytes land on rank 0's barriers) ----------------
    if (lane == 0) {
      int s = 0, wb = 0;
      uint32_t sphase = 0, wphase = 0;
      for (int t = cluster_id; t < a.total_tiles; t += n_clusters) {
      const Tile tl = decode(t);
      const int q0 = tl.q0, cb_begin = tl.cb_begin, cb_end = tl.cb_end;
      const int a_chan0 = tl.grp * a.a_chan_group, b_row0 = tl.grp * a.b_row_group + tl.n0 + rank * nb_local;
      for (int cb = cb_begin; cb < cb_end; ++cb) {
        mbar_wait(&win_empty[wb], wphase ^ 1u);
        if (rank == 0) mbar_arrive_expect_tx(&win_full[wb], win_bytes * CG);
        for (int b = 0; b < a.nbox; ++b) {
          uint8_t* dst = smem + (size_t)wb * win_bytes + (size_t)b * a.box_rows * 128;
          if (CG == 2) tma_load_2d_pair(dst, &tmA, &win_full[wb], a_chan0 + cb * 32, q0 + a.row_off + b * a.box_rows);
          else tma_load_2d(dst, &tmA, &win_full[wb], a_chan0 + cb * 32, q0 + a.row_off + b * a.box_rows);
        }
        for (int tap0 = 0; tap0 < a.T; tap0 += a.tpg) {
          const int ng = min(a.tpg, a.T - tap0);
          if (a.b_resident && t != cluster_id) {                    // already in shared memory since this cluster's first tile
            if (++s == a.nstage) { s = 0; sphase ^= 1u; }
            continue;
          }
          mbar_wait(&b_empty[s], sphase ^ 1u);
          if (rank == 0) mbar_arrive_expect_tx(&b_full[s], b_bytes * CG * ng);
          for (int g = 0; g < ng; ++g) {
            const int tap = tap0 + g;
            uint8_t* dst = bst + ((size_t)s * a.tpg + g) * b_bytes;
            if (a.b_mn) {
              // B stored [reduction row][n]: one box of 32 n x 32 rows per 32 columns of this CTA's share
              for (int blk = 0; blk < nb_local / 32; ++blk) {
                if (CG == 2) tma_load_2d_pair(dst + blk * 4096, &tmB, &b_full[s], b_row0 + 32 * blk, cb * 32);
                else tma_load_2d(dst + blk * 4096, &tmB, &b_full[s], b_row0 + 32 * blk, cb * 32);
              }
            } else if (CG == 2) tma_load_2d_pair(dst, &tmB, &b_full[s], tap * a.slots + cb * 32, b_row0);
            else tma_load_2d(dst, &tmB, &b_full[s], tap * a.slots + cb * 32, b_row0);
          }
          if (++s == a.nstage) { s = 0; sphase ^= 1u; }
        }
        if (++wb == a.nwin) { wb = 0; wphase ^= 1u; }
      }
      }
    }
  }
  tc_fence_before_sync();
  if (CG == 2) cluster_sync_all();
  else __syncthreads();
  if (stamp && tid == 0) stamp[7] = clock64();
  if (warp == kWEpiWarps) {
    if (CG == 2) tmem_dealloc_pair<512>(tmem_base);
    else tmem_dealloc<512>(tmem_base);
  }
}

// Development aid: clocks per tcgen05.mma (TF32, operands in shared memory) when one thread issues `iters` of them
// back to back on resident tiles — the ceiling the pipelines above are compared with.  mode 0: A K-major, B K-major.
template <int CG>
__global__ void __launch_bounds__(128, 1) mma_rate_kernel(int n_mma, int iters, int shift_rows, int nacc, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t done, scratch_bar;
  __shared__ uint32_t tmem_base_smem;
  const int tid = threadIdx.x, warp = tid >> 5;
  const int rank = CG == 2 ? (int)cluster_ctarank() : 0;
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  for (int i = tid; i < 216 * 1024 / 4; i += blockDim.x) ((float*)smem)[i] = 0.f;
  if (tid == 0) { mbar_init(&done, 1u); mbar_init(&scratch_bar, 1u << 20); fence_barrier_init(); }
  if (warp == 0) { if (CG == 2) tmem_alloc_pair<512>(&tmem_base_smem); else tmem_alloc<512>(&tmem_base_smem); }
  fence_proxy_async_smem();
  tc_fence_before_sync();
  if (CG == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = tmem_base_smem;
  if (warp == 1 && rank == 0) {
    const uint32_t idesc = make_idesc_tf32(128 * CG, n_mma, 0, 0);
    const uint32_t desc_hi = (uint32_t)(make_smem_desc(0, 16, 1024, kLayoutSW128) >> 32);
    const uint32_t a_lo = ((smem_u32(smem) >> 4) | (1u << 16)) + (uint32_t)shift_rows * 8u;
    const uint32_t b_lo = ((smem_u32(smem) + 88 * 1024) >> 4) | (1u << 16);
    long long t0 = 0, t1 = 0;
    if (elect_one_sync()) {
      t0 = clock64();
      // shift_rows >= 1000 selects the pipeline-like pattern: per "tap" nacc accumulators x 4 k-steps, a commit after every
      // tap (to a barrier nobody waits on), the A start row and the B stage changing from tap to tap
      const bool piped = shift_rows >= 1000;
      for (int it = 0; it < iters; ++it) {
        if (!piped) {
          const uint32_t d = tmem_base + (uint32_t)((it % nacc) * n_mma);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (CG == 2) mma_tf32_lohi_pair(d, a_lo + 2 * j, b_lo + 2 * j, desc_hi, idesc, 1u);
            else mma_tf32_lohi(d, a_lo + 2 * j, b_lo + 2 * j, desc_hi, idesc, 1u);
          }
        } else {
          const uint32_t ash = (uint32_t)((it % 9) * 3) * 8u, bsh = (uint32_t)(it % 4) * (uint32_t)(n_mma * 128 / CG / 16);
          for (int i = 0; i < nacc; ++i) {
            const uint32_t d = tmem_base + (uint32_t)(i * n_mma);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              if (CG == 2) mma_tf32_lohi_pair(d, a_lo + ash + (uint32_t)(i * 1024 + 2 * j), b_lo + bsh + 2 * j, desc_hi, idesc, 1u);
              else mma_tf32_lohi(d, a_lo + ash + (uint32_t)(i * 1024 + 2 * j), b_lo + bsh + 2 * j, desc_hi, idesc, 1u);
            }
          }
          if (shift_rows == 1001) { if (CG == 2) mma_commit_pair(&scratch_bar); else mma_commit(&scratch_bar); }
        }
      }
      t1 = clock64();
      if (CG == 2) mma_commit_pair(&done); else mma_commit(&done);
    }
    __syncwarp();
    mbar_wait(&done, 0);
    const long long t2 = clock64();
    if (elect_one_sync()) { out[blockIdx.x * 2] = t1 - t0; out[blockIdx.x * 2 + 1] = t2 - t0; }
  } else if (warp == 1) {
    mbar_wait(&done, 0);
  }
  tc_fence_before_sync();
  if (CG == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 0) { if (CG == 2) tmem_dealloc_pair<512>(tmem_base); else tmem_dealloc<512>(tmem_base); }
}
// returns clocks per MMA (issue .. all complete) averaged over the CTAs; out2 (nullable) gets the issue-only figure
float tcw_mma_rate(int cg, int n_mma, int iters, int shift_rows, int nacc, int ctas, float* issue_only) {
  long long* d = nullptr;
  SN_CUDA(cudaMalloc(&d, sizeof(long long) * 2 * ctas));
  SN_CUDA(cudaMemset(d, 0, sizeof(long long) * 2 * ctas));
  const size_t smem = 217 * 1024 + 1024;
  if (cg == 2) {
    SN_CUDA(cudaFuncSetAttribute(mma_rate_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(ctas); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem; cfg.stream = 0;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    SN_CUDA(cudaLaunchKernelEx(&cfg, mma_rate_kernel<2>, n_mma, iters, shift_rows, nacc, d));
  } else {
    SN_CUDA(cudaFuncSetAttribute(mma_rate_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    mma_rate_kernel<1><<<ctas, 128, smem>>>(n_mma, iters, shift_rows, nacc, d);
  }
  SN_CUDA(cudaDeviceSynchronize());
  std::vector<long long> h(2 * ctas);
  SN_CUDA(cudaMemcpy(h.data(), d, sizeof(long long) * 2 * ctas, cudaMemcpyDeviceToHost));
  cudaFree(d);
  double a = 0, b = 0; int n = 0;
  for (int i = 0; i < ctas; i += cg) { a += (double)h[2 * i]; b += (double)h[2 * i + 1]; ++n; }
  const double per = shift_rows >= 1000 ? iters * 4.0 * nacc : iters * 4.0;
  if (issue_only) *issue_only = (float)(a / n / per);
  return (float)(b / n / per);
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
static int g_force_mt = 0, g_force_cg = 0, g_force_n = 0;
static long long* g_stamps = nullptr;
static int g_stamp_ctas = 0;
constexpr int kMaxStampCtas = 4096;
void tcw_set_debug(int what, int value) {
  if (what == 1) g_force_mt = value;
  if (what == 4) g_force_n = value;                   // force the column tile width of the window kernels (0 = planner's choice)
  if (what == 3) g_force_cg = value;                  // 1 = single-CTA MMAs only, 2 = CTA pairs whenever possible, 0 = auto
  if (what == 2) {                                      // per-CTA clock stamps of the next window launches on / off
    if (value && !g_stamps) SN_CUDA(cudaMalloc(&g_stamps, sizeof(long long) * 8 * kMaxStampCtas));
    if (!value && g_stamps) { cudaFree(g_stamps); g_stamps = nullptr; }
  }
}
// copies the stamps of the most recent window launch (8 per CTA) to `out`; returns the number of CTAs
int tcw_read_stamps(long long* out, int max_ctas) {
  if (!g_stamps) return 0;
  SN_CUDA(cudaDeviceSynchronize());
  const int n = g_stamp_ctas < max_ctas ? g_stamp_ctas : max_ctas;
  SN_CUDA(cudaMemcpy(out, g_stamps, sizeof(long long) * 8 * n, cudaMemcpyDeviceToHost));
  return n;
}

struct WinPlan { int MT, n_mma, nstage, box_rows, nbox, cg; size_t smem; bool ok; };

static int win_n_mma(int N) {
  const int tiles = (N + 255) / 256;
  int w = (N + tiles - 1) / tiles;
  w = (w + 15) / 16 * 16;
  return w < 16 ? 16 : w;
}
// Pick the column tile width and the number of accumulators per CTA: fewest estimated clocks for the busiest cluster.
// The per-tile model below is calibrated on the sweep of tools/sweep_window.py (within ~10 % for AlexNet's layers); the
// epilogue costs output bytes / 19 per clock, and tiles whose accumulators fit twice in TMEM (MT * n_mma <= 256) hide it
// behind the next tile's MMAs.
static WinPlan plan_window_impl(int Ptot, int N, int groups, int T, int span, int red, bool forced) {
  WinPlan best{};
  best.ok = false;
  const int ncb = (red + 31) / 32;
  double best_cost = 0;
  const int cg = g_force_cg == 1 ? 1 : 2;                       // CTA pairs: half the filter bytes per SM
  for (int n_tiles = (N + 255) / 256; n_tiles <= (N + 31) / 32; ++n_tiles) {
    const int n_mma = ((N + n_tiles - 1) / n_tiles + 15) / 16 * 16;
    if (n_mma > 256 || (n_tiles > 1 && n_mma * (n_tiles - 1) >= N)) continue;
    if (forced && g_force_n && n_mma != g_force_n) continue;
    for (int MT = 4; MT >= 1; MT >>= 1) {
      if (MT * n_mma > 512) continue;
      if (forced && g_force_mt && MT != g_force_mt) continue;
      const int win_rows = MT * 128 + span;
      const int nbox = (win_rows + 255) / 256;
      const int box_rows = ((win_rows + nbox - 1) / nbox + 7) / 8 * 8;
      const size_t win_bytes = (size_t)nbox * box_rows * 128, b_bytes = (size_t)n_mma * 128 / cg;
      const size_t fixed = 2 * win_bytes + 1024 + 512;             // + alignment slack + static shared (barriers)
      if (fixed + 3 * b_bytes > kWSmemLimit) continue;
      int nstage = (int)((kWSmemLimit - fixed) / b_bytes);
      if (nstage > kWMaxStages) nstage = kWMaxStages;
      const long long m_ctas = (Ptot + MT * 128 - 1) / (MT * 128);
      const long long tiles = (m_ctas + cg - 1) / cg * n_tiles * groups;
      const long long per_cluster = (tiles + 148 / cg - 1) / (148 / cg);
      // per tap (measured with tools/sweep_window.py): ~300 clocks of pipeline bookkeeping at least (several taps share one stage), else 1.2 x the MMA
      // clocks of the MT accumulators (one MMA = max(n_mma / 2, 42) clocks), else the L2 bytes of the tap at 25 B/clk
      const double mma1 = n_mma / 2.0 > 42.0 ? n_mma / 2.0 : 42.0;
      const double l2_tap = ((double)win_bytes / T + (double)b_bytes) / 25.0;
      double t_main = 0;
      for (int cb = 0; cb < ncb; ++cb) {
        const int left = red - cb * 32, ks = left >= 32 ? 4 : (left + 7) / 8;
        double tap = 1.2 * MT * ks * mma1;
        if (tap < 300.0) tap = 300.0;
        if (tap < l2_tap) tap = l2_tap;
        t_main += T * tap;
      }
      const double t_epi = (double)MT * 128 * n_mma * 4 / 19.0;
      const bool dbl = MT * n_mma <= 256;
      const double cost = dbl ? per_cluster * (t_main > t_epi ? t_main : t_epi) + t_epi + 6000.0
                              : per_cluster * (t_main + t_epi) + 6000.0;
      if (!best.ok || cost < best_cost) {
        best = WinPlan{MT, n_mma, nstage, box_rows, nbox, cg, 2 * win_bytes + (size_t)nstage * b_bytes + 1024, true};
        best_cost = cost;
      }
    }
  }
  return best;
}

// the development switches (forced tile shape) apply where they can be met, else the planner decides
// Plans are memoised: the launchers run every iteration, the enumeration is host time a small net would notice.
static WinPlan plan_window(int Ptot, int N, int groups, int T, int span, int red) {
  static std::map<std::array<int, 9>, WinPlan> memo;
  static std::mutex mu;
  std::lock_guard<std::mutex> lock(mu);
  const std::array<int, 9> key{Ptot, N, groups, T, span, red, g_force_n, g_force_mt, g_force_cg};
  auto it = memo.find(key);
  if (it != memo.end()) return it->second;
  WinPlan p = plan_window_impl(Ptot, N, groups, T, span, red, true);
  if (!p.ok && (g_force_n || g_force_mt)) p = plan_window_impl(Ptot, N, groups, T, span, red, false);
  memo[key] = p;
  return p;
}

static void launch_window(cudaStream_t s, const CUtensorMap& tmA, const CUtensorMap& tmB, const WinArgs& a, const WinEpi& ep,
                          const WinPlan& pl, int N, int groups) {
  static bool attr_done = false;
  if (!attr_done) {
    SN_CUDA(cudaFuncSetAttribute(conv_window_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWSmemLimit - 1024));
    SN_CUDA(cudaFuncSetAttribute(conv_window_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWSmemLimit - 1024));
    attr_done = true;
  }
  // persistent CTAs: one cluster per SM (pair), each walking the tile list with stride #clusters
  WinArgs aa = a;
  const int m_ctas = (a.Ptot + a.MT * 128 - 1) / (a.MT * 128);
  aa.m_units = (m_ctas + pl.cg - 1) / pl.cg;              // whole pairs: an odd last CTA works on rows past the end (zero-filled)
  aa.n_tiles = (N + a.n_mma - 1) / a.n_mma;
  aa.total_tiles = aa.m_units * aa.n_tiles * groups * a.splits;
  aa.acc_stages = a.MT * a.n_mma <= 256 ? 2 : 1;
  size_t smem_bytes = pl.smem;
  {
    // taps per stage: a divisor of T up to 5 when at least three stages still fit, else the largest group that does
    const size_t b_bytes = (size_t)a.n_mma * 128 / pl.cg, win_total = (size_t)a.nwin * a.nbox * a.box_rows * 128;
    const size_t avail = kWSmemLimit - 2048 - win_total;
    int best_g = 1;
    for (int g = 5; g >= 2; --g) {
      if (g > a.T || avail / (g * b_bytes) < 3) continue;
      if (a.T % g == 0) { best_g = g; break; }
      if (best_g == 1) best_g = g;
    }
    static const int forced_tpg = getenv("SPARKNET_TPG") ? atoi(getenv("SPARKNET_TPG")) : 0;
    if (forced_tpg > 0) best_g = forced_tpg < a.T ? forced_tpg : a.T;
    if (avail / (best_g * b_bytes) < 2) best_g = 1;
    aa.tpg = best_g;
    int nst = (int)(avail / (best_g * b_bytes));
    aa.nstage = nst > kWMaxStages ? kWMaxStages : nst;
    smem_bytes = win_total + (size_t)aa.nstage * best_g * b_bytes + 1024;
  }
  aa.b_resident = 0;
  {
    static const int allow = getenv("SPARKNET_B_RESIDENT") ? atoi(getenv("SPARKNET_B_RESIDENT")) : 1;
    const size_t b_bytes = (size_t)a.n_mma * 128 / pl.cg, win_total = (size_t)a.nwin * a.nbox * a.box_rows * 128;
    if (allow && aa.n_tiles * groups * a.splits == 1 && !a.b_mn && aa.total_tiles > 148 / pl.cg) {
      // taps per stage g: all ncb * ceil(T / g) stages must exist at once (<= kWMaxStages barriers) and fit beside the windows
      int best = 0;
      size_t best_bytes = 0;
      for (int g = 1; g <= a.T; ++g) {
        const int nst = a.ncb * ((a.T + g - 1) / g);
        const size_t bytes = (size_t)nst * g * b_bytes;
        if (nst > kWMaxStages || win_total + bytes + 2048 > kWSmemLimit) continue;
        if (!best || bytes <= best_bytes) { best = g; best_bytes = bytes; }
      }
      if (best) {
        aa.b_resident = 1;
        aa.tpg = best;
        aa.nstage = a.ncb * ((a.T + best - 1) / best);
        smem_bytes = win_total + best_bytes + 1024;
      }
    }
  }
  static const bool plan_debug = getenv("SPARKNET_PLAN_DEBUG") != nullptr;
  if (plan_debug && aa.b_resident) fprintf(stderr, "window plan: filter resident in shared memory (%d stages of %d taps)\n", aa.nstage, aa.tpg);
  if (plan_debug)
    fprintf(stderr, "window plan: rows %d N %d T %d ncb %d -> MT %d n_mma %d cg %d stages %d tiles %d acc_stages %d splits %d\n", a.Ptot, N, a.T,
            a.ncb, a.MT, a.n_mma, pl.cg, aa.nstage, aa.total_tiles, aa.acc_stages, a.splits);
  if (plan_debug) fprintf(stderr, "   taps per stage %d, smem %zu\n", aa.tpg, smem_bytes);
  const int max_clusters = 148 / pl.cg;
  const int clusters = aa.total_tiles < max_clusters ? aa.total_tiles : max_clusters;
  dim3 grid(clusters * pl.cg, 1, 1);
  const int ctas = (int)grid.x;
  aa.stamps = (g_stamps && ctas <= kMaxStampCtas) ? g_stamps : nullptr;
  if (aa.stamps) g_stamp_ctas = ctas;
  if (pl.cg == 2) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(kWThreads);
    cfg.dynamicSmemBytes = smem_bytes;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    SN_CUDA(cudaLaunchKernelEx(&cfg, conv_window_kernel<2>, tmA, tmB, aa, ep));
  } else {
    conv_window_kernel<1><<<grid, kWThreads, smem_bytes, s>>>(tmA, tmB, aa, ep);
  }
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
void tcw_conv_fwd(cudaStream_t s, const ConvGeom& g, const float* xflat, const float* wf, const float* bias, float* y, bool relu,
                  const FlatOut* next) {
  const FlatGeom f = tc_flat_geom(g);
  const int Cg = g.Cg(), Og = g.Og(), T = g.kh * g.kw, Ptot = g.N * f.Gp, slots = slots_of(Cg), kp = T * slots;
  const int span = (g.kh - 1) * f.Wp + g.kw - 1;
  const WinPlan pl = plan_window(Ptot, Og, g.group, T, span, Cg);
  if (!pl.ok) fatal("window convolution: no tile plan for this layer (forward)");
  const CUtensorMap tmA = tc_make_tmap(xflat, (uint64_t)g.C, tc_flat_rows(g), (uint64_t)g.C * 4, (uint32_t)pl.box_rows, false);
  const CUtensorMap tmB = tc_make_tmap(wf, (uint64_t)kp, (uint64_t)g.Cout, (uint64_t)kp * 4, (uint32_t)(pl.n_mma / pl.cg), false);
  WinArgs a{Ptot, f.Wp, g.kw, T, f.lead - (g.ph * f.Wp + g.pw), pl.MT, pl.n_mma, slots / 32, Cg, slots, Cg, Og,
            pl.box_rows, pl.nbox, pl.nstage, nullptr, 2, 1, slots / 32, 0};
  WinEpi ep{y, bias, g.Ho, g.Wo, g.Cout, Og, relu ? 1 : 0, FastDiv(f.Gp), FastDiv(f.Wp), 0, 0, nullptr, nullptr, 0, 0, 0, 0};
  if (next && next->flat) {
    if (next->C != g.Cout || (next->C % 4) != 0) fatal("window convolution: the consumer's flat grid has %d channels, this layer writes %d", next->C, g.Cout);
    ep.flat = next->flat; ep.flat_lead = next->lead; ep.flat_Gp = next->Gp; ep.flat_Wp = next->Wp; ep.flat_C = next->C;
  }
  launch_window(s, tmA, tmB, a, ep, pl, Og, g.group);
}
// `dyflat`: packed top gradient (tc_flat_pack, output side); `wt`: tc_flat_pack_filter(dgrad = true)
void tcw_conv_dgrad(cudaStream_t s, const ConvGeom& g, const float* dyflat, const float* wt, float* dx, const float* relu_gate) {
  const FlatGeom f = tc_flat_geom(g);
  const int Cg = g.Cg(), Og = g.Og(), T = g.kh * g.kw, Ptot = g.N * f.Gp, slots = slots_of(Og), kp = T * slots;
  const int span = (g.kh - 1) * f.Wp + g.kw - 1;
  const WinPlan pl = plan_window(Ptot, Cg, g.group, T, span, Og);
  if (!pl.ok) fatal("window convolution: no tile plan for this layer (backward data)");
  const CUtensorMap tmA = tc_make_tmap(dyflat, (uint64_t)g.Cout, tc_flat_rows(g), (uint64_t)g.Cout * 4, (uint32_t)pl.box_rows, false);
  const CUtensorMap tmB = tc_make_tmap(wt, (uint64_t)kp, (uint64_t)g.C, (uint64_t)kp * 4, (uint32_t)(pl.n_mma / pl.cg), false);
  WinArgs a{Ptot, f.Wp, g.kw, T, f.lead - ((g.kh - 1 - g.ph) * f.Wp + (g.kw - 1 - g.pw)), pl.MT, pl.n_mma, slots / 32, Og, slots,
            Og, Cg, pl.box_rows, pl.nbox, pl.nstage, nullptr, 2, 1, slots / 32, 0};
  WinEpi ep{dx, nullptr, g.H, g.W, g.C, Cg, 0, FastDiv(f.Gp), FastDiv(f.Wp), 0, 0, relu_gate, nullptr, 0, 0, 0, 0};
  launch_window(s, tmA, tmB, a, ep, pl, Cg, g.group);
}


// ------------------------------------------------------------------------------------------
// Plain matrix products for the inner-product layers (inner_product_layer.cpp:78-117) on the same kernel: a one-tap
// "convolution" whose grid positions are the batch rows.  out[m][n] = sum_k A[m][k] * B(n, k) (+ bias[n], ReLU);
//   A  [M][K] row-major, already rounded to tf32;   B  [N][K] row-major (b_mn = false) or [K][N] row-major (b_mn = true).
// The weight matrix streams from HBM exactly once; CTA pairs split its rows, the reduction is cut into ranges so that
// ~148 CTAs are busy, and the slices are summed in range order (deterministic) with bias / ReLU applied there.
// ------------------------------------------------------------------------------------------
__global__ void mat_reduce_kernel(const float* __restrict__ ws, int splits, size_t total, int N, const float* __restrict__ bias, int relu,
                                  float* __restrict__ out) {
  SN_GRID_DEP_LAUNCH_ALWAYS();
  SN_GRID_DEP_WAIT();
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    float acc = ws[i];
    for (int z = 1; z < splits; ++z) acc += ws[(size_t)z * total + i];
    if (bias) acc += __ldg(bias + (int)(i % N));
    if (relu) acc = fmaxf(acc, 0.f);
    out[i] = acc;
  }
}
bool tcw_matmul_supported(int M, int N, int K, const float* A, const float* B, bool b_mn, const float* out) {
  if (M < 1 || N < 16 || K < 8 || N % 4 != 0 || K % 4 != 0) return false;
  if (((uintptr_t)A | (uintptr_t)B | (uintptr_t)out) & 15) return false;
  (void)b_mn;
  return true;
}
void tcw_matmul(cudaStream_t s, int M, int N, int K, const float* A, const float* B, bool b_mn, const float* bias, bool relu, float* out) {
  WinPlan pl{};
  pl.cg = (M > 128 && g_force_cg != 1) ? 2 : 1;
  pl.MT = M > 128 * pl.cg ? 2 : 1;                     // 2 accumulators of 256 columns fill the 512 TMEM columns
  const int gran = b_mn ? 32 * pl.cg : 16;
  int n_mma = N >= 256 ? 256 : (N + gran - 1) / gran * gran;
  if (pl.MT * n_mma > 512) n_mma = 256;
  pl.n_mma = n_mma;
  const int win_rows = pl.MT * 128;
  pl.nbox = 1; pl.box_rows = win_rows;
  const size_t win_bytes = (size_t)win_rows * 128, b_bytes = (size_t)n_mma * 128 / pl.cg;
  int depth = (int)((kWSmemLimit - 2048) / (win_bytes + b_bytes));
  if (depth > kWMaxStages) depth = kWMaxStages;
  pl.nstage = depth;
  pl.smem = (size_t)depth * (win_bytes + b_bytes) + 1024;
  const int ncb = (K + 31) / 32, m_ctas = (M + win_rows - 1) / win_rows, n_tiles = (N + n_mma - 1) / n_mma;
  const int tiles = ((pl.cg == 2 ? (m_ctas + 1) / 2 * 2 : m_ctas)) * n_tiles;
  int splits = 148 / tiles;
  if (splits < 1) splits = 1;
  if (splits > ncb / 4) splits = ncb / 4 > 0 ? ncb / 4 : 1;
  const int cps = (ncb + splits - 1) / splits;
  splits = (ncb + cps - 1) / cps;
  const CUtensorMap tmA = tc_make_tmap(A, (uint64_t)K, (uint64_t)M, (uint64_t)K * 4, (uint32_t)pl.box_rows, false);
  const CUtensorMap tmB = b_mn ? tc_make_tmap(B, (uint64_t)N, (uint64_t)K, (uint64_t)N * 4, 32, true)
                               : tc_make_tmap(B, (uint64_t)K, (uint64_t)N, (uint64_t)K * 4, (uint32_t)(n_mma / pl.cg), false);
  WinArgs a{M, 1, 1, 1, 0, pl.MT, n_mma, ncb, K, 0, 0, 0, pl.box_rows, pl.nbox, pl.nstage, nullptr, depth, splits, cps, b_mn ? 1 : 0};
  float* dst = out;
  size_t stride = 0;
  if (splits > 1) { stride = (size_t)M * N; dst = get_workspace(stride * splits); }
  WinEpi ep{dst, bias, 1, 1, N, N, relu ? 1 : 0, FastDiv(1), FastDiv(1), N, stride, nullptr, nullptr, 0, 0, 0, 0};
  launch_window(s, tmA, tmB, a, ep, pl, N, 1);
  if (splits > 1) {
    const size_t total = (size_t)M * N;
    int blocks = (int)((total + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    mat_reduce_kernel<<<blocks, 256, 0, s>>>(dst, splits, total, N, bias, relu ? 1 : 0, out);
    count_launch();
  }
  SN_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------------------------------
// Backward-filter on the flat grid:  dW[co][c][tap] = sum_q dyflat[q][co] * xflat[q + shift(tap) - pad][c].
// The reduction runs over grid positions q, 32 per chunk.  A = top-gradient tile (32 q x 128 co, MN-major), B = the
// rows [q0 + shift(t0) - pad, + 32 + span) of the packed input for n_mma channels (MN-major, one TMA box per 32
// channels): the taps t0 .. t0 + nt - 1 of one CTA read it through descriptors whose start advances by
// (shift(t) - shift(t0)) rows, each tap into its own accumulator (nt * n_mma <= 512 TMEM columns).  One load of the
// top-gradient tile therefore serves nt taps (the chunk-per-tap kernel re-read it for every 256 / slots taps and
// re-read the input rows for every tap).  Slices of the reduction go to the workspace as [slice][co][tap][c] and
// wgrad_reduce_kernel sums them in slice order into dW's (co, c, tap) layout: deterministic, no atomics.
// ------------------------------------------------------------------------------------------
struct WgArgs {
  int Ptot_chunks;          // number of R*32-row chunks of the reduction
  int chunks_per_split;
  int tmem_cols;            // TMEM columns to allocate: nt * n_mma rounded up to a power of two (>= 32)
  int R;                    // 32-row top-gradient tiles per pipeline stage: they share ONE input window of R*32 + span rows, so the
                            // span rows (116 for AlexNet conv1's folded 3x3) cross the L2 -> SM path once per R tiles instead of per tile
  int Wp, kw, T;
  int nt, n_groups_t;       // taps per CTA, number of tap groups
  int m_tiles, n_tiles, n_mma, nblk;      // 128-row tiles of Og, channel tiles of Cg (n_mma wide, nblk = ceil(n_mma / 32) boxes)
  int Og, Cg;
  int lead, pad_off;        // leading zero rows of the packed tensors; ph * Wp + pw
  int KR;                   // rows of one input window (multiple of 8)
  int nstage;
  float* ws;                // [split][Cout][T][Cg]
  size_t split_stride;      // Cout * T * Cg
};

__global__ void __launch_bounds__(kWThreads, 1)
wgrad_window_kernel(const __grid_constant__ CUtensorMap tmDY, const __grid_constant__ CUtensorMap tmX, const WgArgs a) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full[kWMaxStages], empty[kWMaxStages], tmem_full;
  __shared__ uint32_t tmem_base_smem;
  SN_GRID_DEP_LAUNCH();

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // work item: (group, m tile, n tile, tap group)
  int w = blockIdx.x;
  const int tg = w % a.n_groups_t; w /= a.n_groups_t;
  const int nti = w % a.n_tiles; w /= a.n_tiles;
  const int mti = w % a.m_tiles; w /= a.m_tiles;
  const int grp = w;
  const int t0 = tg * a.nt, nt = min(a.nt, a.T - t0);
  const int m0 = mti * 128, n0 = nti * a.n_mma;
  const int kc_begin = blockIdx.y * a.chunks_per_split, kc_end = min(a.Ptot_chunks, kc_begin + a.chunks_per_split);
  const int nchunks = max(kc_end - kc_begin, 0);

  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t a_bytes = (uint32_t)a.R * 4u * 4096u, blk_bytes = (uint32_t)a.KR * 128u, b_bytes = (uint32_t)a.nblk * blk_bytes;
  const uint32_t stage_bytes = a_bytes + b_bytes;                          // multiple of 1024 (KR % 8 == 0)

  if (tid == 0) {
    const uint32_t issuers = (uint32_t)min(kWMmaWarps, nt);            // MMA warps that commit on every stage
    for (int i = 0; i < a.nstage; ++i) { mbar_init(&full[i], 1u); mbar_init(&empty[i], issuers); }
    mbar_init(&tmem_full, issuers);
    fence_barrier_init();
    tma_prefetch_desc(&tmDY);
    tma_prefetch_desc(&tmX);
  }
  if (warp == kWEpiWarps) tmem_alloc_cols(&tmem_base_smem, (uint32_t)a.tmem_cols);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  SN_GRID_DEP_WAIT();                        // prologue done; operands and the workspace slice are touched from here on
  const uint32_t tmem_base = tmem_base_smem;
  const int sh0 = (t0 / a.kw) * a.Wp + t0 % a.kw;                          // row shift of the group's first tap

  if (warp < kWEpiWarps) {
    // ---------------- epilogue: TMEM -> workspace slice [co][tap][c] ----------------
    if (nchunks > 0) {
      mbar_wait(&tmem_full, 0);
      tc_fence_after_sync();
    }
    const int quarter = warp & 3, half = warp >> 2;
    const int row = m0 + 32 * quarter + lane;                              // output channel within the group
    const int ncols = min(a.n_mma, a.Cg - n0);
    const bool live = row < a.Og;                                          // the TMEM loads are warp-collective: every lane runs them
    float* dst_row = a.ws + (size_t)blockIdx.y * a.split_stride + (size_t)(grp * a.Og + (live ? row : 0)) * a.T * a.Cg + n0;
    for (int ts = 0; ts < nt; ++ts) {
      float* dst = dst_row + (size_t)(t0 + ts) * a.Cg;
      for (int c0 = 16 * half; c0 < ncols; c0 += 32) {
        uint32_t v[16];
        if (nchunks > 0) {
          tmem_ld_32x32b_x16(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(ts * a.n_mma + c0), v);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = 0u;
        }
        if (live) {
#pragma unroll
          for (int j = 0; j < 16; j += 4)
            *(float4*)(dst + c0 + j) = make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
        }
      }
    }
  } else if (warp < kWTmaWarp) {
    // ---------------- MMA issuers: warp mi takes the taps t0 + mi, t0 + mi + issuers, ... (one accumulator each) ----------------
    const int mi = warp - kWEpiWarps, issuers = min(kWMmaWarps, nt);
    if (mi < issuers) {
      const uint32_t idesc = make_idesc_tf32(128, a.n_mma, 1, 1);
      // MN-major descriptors: A blocks (32 co) are 4096 B apart, B blocks (32 c) KR * 128 B apart; 4-row atoms 512 B apart
      const uint32_t a_hi = (uint32_t)(make_smem_desc(0, 4096, 512, kLayoutSW128Base32) >> 32);
      const uint32_t a_lo_flags = (4096u >> 4) << 16, b_lo_flags = (blk_bytes >> 4) << 16;
      constexpr int kMaxMine = 16;                                          // taps per issuing warp (nt <= 64)
      uint32_t shift8[kMaxMine];                                            // (row shift of my k-th tap) * 128 B >> 4
      int mine = 0;
      for (int ts = mi; ts < nt && mine < kMaxMine; ts += issuers, ++mine) {
        const int t = t0 + ts;
        shift8[mine] = (uint32_t)((t / a.kw) * a.Wp + t % a.kw - sh0) * 8u;
      }
      int s = 0;
      uint32_t phase = 0;
      for (int kc = 0; kc < nchunks; ++kc) {
        mbar_wait(&full[s], phase);
        tc_fence_after_sync();
        if (elect_one_sync()) {
          const uint32_t a_addr = smem_u32(smem) + (uint32_t)s * stage_bytes, b_addr = a_addr + a_bytes;
          const uint32_t a_lo = (a_addr >> 4) | a_lo_flags, b_lo0 = (b_addr >> 4) | b_lo_flags;
#pragma unroll 1
          for (int r = 0; r < a.R; ++r) {                          // tile r: top-gradient rows 32r.., window rows 32r + shift..
            const uint32_t a_lo_r = a_lo + (uint32_t)r * (16384u >> 4), b_lo_r = b_lo0 + (uint32_t)r * 32u * 8u;
#pragma unroll 1
            for (int k = 0; k < mine; ++k) {
              const uint32_t b_lo = b_lo_r + shift8[k];
              const uint32_t d = tmem_base + (uint32_t)((mi + k * issuers) * a.n_mma);
#pragma unroll
              for (int j = 0; j < 4; ++j) mma_tf32_lohi(d, a_lo_r + 64u * j, b_lo + 64u * j, a_hi, idesc, (kc | r | j) != 0 ? 1u : 0u);
            }
          }
          mma_commit(&empty[s]);
          if (kc == nchunks - 1) mma_commit(&tmem_full);
        }
        __syncwarp();
        if (++s == a.nstage) { s = 0; phase ^= 1u; }
      }
    }
  } else {
    // ---------------- TMA issuer ----------------
    if (lane == 0) {
      int s = 0;
      uint32_t phase = 0;
      const int a_chan = grp * a.Og + m0, b_chan = grp * a.Cg + n0;
      for (int kc = 0; kc < nchunks; ++kc) {
        const int q0 = (kc_begin + kc) * 32 * a.R;
        mbar_wait(&empty[s], phase ^ 1u);
        uint8_t* a_tile = smem + (size_t)s * stage_bytes;
        uint8_t* b_tile = a_tile + a_bytes;
        mbar_arrive_expect_tx(&full[s], stage_bytes);
        for (int r = 0; r < a.R; ++r)
          for (int i = 0; i < 4; ++i) tma_load_2d(a_tile + r * 16384 + i * 4096, &tmDY, &full[s], a_chan + 32 * i, a.lead + q0 + 32 * r);
        for (int j = 0; j < a.nblk; ++j)
          tma_load_2d(b_tile + (size_t)j * blk_bytes, &tmX, &full[s], b_chan + 32 * j, a.lead + q0 + sh0 - a.pad_off);
        if (++s == a.nstage) { s = 0; phase ^= 1u; }
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == kWEpiWarps) tmem_dealloc_cols(tmem_base, (uint32_t)a.tmem_cols);
}

// dw[(co, c, tap)] (+)= sum over slices of ws[slice][(co, tap, c)], slices in ascending order
__global__ void wgrad_reduce_kernel(const float* __restrict__ ws, int splits, size_t split_stride, int T, int Cg, float* __restrict__ dw,
                                    int accumulate) {
  SN_GRID_DEP_LAUNCH_ALWAYS();
  SN_GRID_DEP_WAIT();
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < split_stride; i += (size_t)gridDim.x * blockDim.x) {
    // i walks the workspace layout (co, tap, c): coalesced reads
    const size_t co = i / ((size_t)T * Cg);
    const int rem = (int)(i - co * (size_t)T * Cg), tap = rem / Cg, c = rem - tap * Cg;
    float acc = ws[i];
    for (int z = 1; z < splits; ++z) acc += ws[(size_t)z * split_stride + i];
    float* p = dw + (co * Cg + c) * T + tap;
    *p = accumulate ? *p + acc : acc;
  }
}

// Many slices over a small filter (conv1: 147 slices x 41 k elements): one thread per element would walk `splits`
// dependent round trips.  Eight lanes share an element: lane j sums slices j, j + 8, ... in ascending order, then the eight
// partial sums meet in a fixed butterfly (same order on every run).
__global__ void wgrad_reduce_wide_kernel(const float* __restrict__ ws, int splits, size_t split_stride, int T, int Cg, float* __restrict__ dw,
                                         int accumulate) {
  SN_GRID_DEP_LAUNCH_ALWAYS();
  SN_GRID_DEP_WAIT();
  const size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const size_t i = t >> 3;
  const int j = (int)(t & 7);
  float acc = 0.f;
  if (i < split_stride) {
    float a0 = 0.f, a1 = 0.f;
    int z = j;
    for (; z + 8 < splits; z += 16) { a0 += ws[(size_t)z * split_stride + i]; a1 += ws[(size_t)(z + 8) * split_stride + i]; }
    if (z < splits) a0 += ws[(size_t)z * split_stride + i];
    acc = a0 + a1;
  }
  acc += __shfl_xor_sync(0xffffffffu, acc, 1);
  acc += __shfl_xor_sync(0xffffffffu, acc, 2);
  acc += __shfl_xor_sync(0xffffffffu, acc, 4);
  if (i < split_stride && j == 0) {
    const size_t co = i / ((size_t)T * Cg);
    const int rem = (int)(i - co * (size_t)T * Cg), tap = rem / Cg, c = rem - tap * Cg;
    float* p = dw + (co * Cg + c) * T + tap;
    *p = accumulate ? *p + acc : acc;
  }
}

struct WgPlan { int n_mma, nt, KR, nstage, splits, chunks_per_split; size_t smem; bool ok; int R; };
static int span_of_taps(int t0, int nt, int kw, int Wp) {
  const int t1 = t0 + nt - 1;
  return (t1 / kw) * Wp + t1 % kw - ((t0 / kw) * Wp + t0 % kw);
}
static WgPlan plan_wgrad_impl(int Ptot, int Og, int Cg, int groups, int T, int kw, int Wp);
static WgPlan plan_wgrad(int Ptot, int Og, int Cg, int groups, int T, int kw, int Wp) {
  static std::map<std::array<int, 7>, WgPlan> memo;
  static std::mutex mu;
  std::lock_guard<std::mutex> lock(mu);
  const std::array<int, 7> key{Ptot, Og, Cg, groups, T, kw, Wp};
  auto it = memo.find(key);
  if (it != memo.end()) return it->second;
  return memo[key] = plan_wgrad_impl(Ptot, Og, Cg, groups, T, kw, Wp);
}
static WgPlan plan_wgrad_impl(int Ptot, int Og, int Cg, int groups, int T, int kw, int Wp) {
  WgPlan best{};
  best.ok = false;
  double best_cost = 0;
  const int chunks = (Ptot + 31) / 32, m_tiles = (Og + 127) / 128;
  for (int n_tiles = 1; n_tiles <= 512; ++n_tiles) {
    int n_mma = ((Cg + n_tiles - 1) / n_tiles + 15) / 16 * 16;
    if (n_mma < 16) break;
    if (n_mma > 256) continue;
    if (n_tiles > 1 && n_mma * (n_tiles - 1) >= Cg) continue;          // an empty last tile
    const int nt_max = 512 / n_mma < T ? 512 / n_mma : T;
    const int n_groups_t = (T + nt_max - 1) / nt_max, nt = (T + n_groups_t - 1) / n_groups_t;
    int span = 0;
    for (int t0 = 0; t0 < T; t0 += nt) { const int sp = span_of_taps(t0, nt < T - t0 ? nt : T - t0, kw, Wp); if (sp > span) span = sp; }
    static const int forced_R = getenv("SPARKNET_WGRAD_R") ? atoi(getenv("SPARKNET_WGRAD_R")) : 0;
    for (int R = 1; R <= 4; R *= 2) {                                  // top-gradient tiles per stage (they share one input window)
      if (forced_R && R != forced_R) continue;
      if (R > 1 && span == 0) break;                                   // nothing to share (one tap, inner products)
      const int KR = (32 * R + span + 7) / 8 * 8;
      if (KR > 256) continue;
      const int nblk = (n_mma + 31) / 32;
      const size_t stage = (size_t)R * 16384 + (size_t)nblk * KR * 128;
      if ((R == 1 ? 3 : 2) * stage + 2048 > kWSmemLimit) continue;
      int nstage = (int)((kWSmemLimit - 2048) / stage);
      if (nstage > kWMaxStages) nstage = kWMaxStages;
      const int chunksR = (chunks + R - 1) / R;                        // chunks of R * 32 rows
      const long long tiles = (long long)m_tiles * n_tiles * n_groups_t * groups;
      int splits = (int)(148 / tiles);
      if (splits < 1) splits = 1;
      const int min_cps = 8 / R > 1 ? 8 / R : 1;
      if (splits > chunksR / min_cps) splits = chunksR / min_cps > 0 ? chunksR / min_cps : 1;
      const int cps = (chunksR + splits - 1) / splits;
      splits = (chunksR + cps - 1) / cps;
      const long long waves = (tiles * splits + 147) / 148;
      const double mma_clk = R * nt * 4.0 * (n_mma / 2.0 > 44.0 ? n_mma / 2.0 : 44.0);
      const double l2_clk = (double)stage / 25.0;
      double per_chunk = mma_clk > l2_clk ? mma_clk : l2_clk;
      if (nstage < 3) per_chunk *= 1.1;                                // two buffers hide less of the load latency
      if (R > 1 && span >= 32) per_chunk *= 0.95;                      // measured (AlexNet conv2, span 56-66): the shared window pays
                                                                       // 20 us even where this model calls the chunk MMA-bound
      const double cost = (double)waves * (cps * per_chunk + 128.0 * nt * n_mma * 4 / 30.0 + 5000.0);
      if (!best.ok || cost < best_cost) {
        best = WgPlan{n_mma, nt, KR, nstage, splits, cps, (size_t)nstage * stage + 1024, true, R};
        best_cost = cost;
      }
    }
  }
  return best;
}

bool tcw_wgrad_supported(const ConvGeom& g) {
  if (!tc_flat_supported(g)) return false;
  const FlatGeom f = tc_flat_geom(g);
  return plan_wgrad(g.N * f.Gp, g.Og(), g.Cg(), g.group, g.kh * g.kw, g.kw, f.Wp).ok;
}

// x: [rows][x_chan] and dy: [rows][dy_chan] packed tensors (tf32-rounded, channels contiguous); Ptot reduction rows
static void wgrad_core(cudaStream_t s, const float* xflat, int x_chan, const float* dyflat, int dy_chan, size_t rows_total, int Ptot, int Og, int Cg,
                       int groups, int T, int kw, int Wp, int lead, int pad_off, float* dw, bool accumulate) {
  const WgPlan pl = plan_wgrad(Ptot, Og, Cg, groups, T, kw, Wp);
  if (!pl.ok) fatal("window convolution: no tile plan for this layer (backward filter)");
  static bool attr_done = false;
  if (!attr_done) {
    SN_CUDA(cudaFuncSetAttribute(wgrad_window_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWSmemLimit - 1024));
    attr_done = true;
  }
  WgArgs a;
  a.R = pl.R;
  a.Ptot_chunks = ((Ptot + 31) / 32 + pl.R - 1) / pl.R;
  a.chunks_per_split = pl.chunks_per_split;
  a.Wp = Wp; a.kw = kw; a.T = T;
  a.nt = pl.nt; a.n_groups_t = (T + pl.nt - 1) / pl.nt;
  a.m_tiles = (Og + 127) / 128;
  a.n_mma = pl.n_mma; a.n_tiles = (Cg + pl.n_mma - 1) / pl.n_mma; a.nblk = (pl.n_mma + 31) / 32;
  a.Og = Og; a.Cg = Cg;
  a.lead = lead; a.pad_off = pad_off;
  a.KR = pl.KR; a.nstage = pl.nstage;
  a.tmem_cols = 32;
  while (a.tmem_cols < pl.nt * pl.n_mma) a.tmem_cols *= 2;
  size_t smem_bytes = pl.smem;
  {
    // Short reductions (inner products: 8 chunks per CTA) are dominated by each CTA's prologue and epilogue, not by its
    // main loop: two stages and half of the TMEM let TWO CTAs share an SM, one loading while the other drains.
    static const int allow2 = getenv("SPARKNET_WGRAD_2CTA") ? atoi(getenv("SPARKNET_WGRAD_2CTA")) : 1;
    const size_t stage = (pl.smem - 1024) / (size_t)pl.nstage;
    if (allow2 && pl.chunks_per_split <= 16 && a.tmem_cols <= 256 && 2 * stage + 1024 <= 110 * 1024) {
      a.nstage = (int)((110 * 1024 - 1024) / stage);
      if (a.nstage > pl.nstage) a.nstage = pl.nstage;
      smem_bytes = (size_t)a.nstage * stage + 1024;
    }
  }
  a.split_stride = (size_t)Og * groups * T * Cg;
  // one tap, one slice, plain store: the slice layout [co][tap][c] IS dW, so the kernel writes it directly
  const bool direct = T == 1 && pl.splits == 1 && !accumulate;
  a.ws = direct ? dw : get_workspace((size_t)pl.splits * a.split_stride);
  const CUtensorMap tmDY = tc_make_tmap(dyflat, (uint64_t)dy_chan, rows_total, (uint64_t)dy_chan * 4, 32, true);
  const CUtensorMap tmX = tc_make_tmap(xflat, (uint64_t)x_chan, rows_total, (uint64_t)x_chan * 4, (uint32_t)pl.KR, true);
  dim3 grid(a.m_tiles * a.n_tiles * a.n_groups_t * groups, pl.splits, 1);
  static const bool plan_debug = getenv("SPARKNET_PLAN_DEBUG") != nullptr;
  if (plan_debug)
    fprintf(stderr, "wgrad plan: rows %d Og %d Cg %d T %d -> n_mma %d taps/CTA %d tiles/stage %d window rows %d stages %d splits %d grid %u x %u\n", Ptot, Og,
            Cg, T, pl.n_mma, pl.nt, pl.R, pl.KR, pl.nstage, pl.splits, grid.x, grid.y);
  wgrad_window_kernel<<<grid, kWThreads, smem_bytes, s>>>(tmDY, tmX, a);
  count_launch();
  if (!direct) {
    int blocks = (int)((a.split_stride + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (pl.splits >= 32 && a.split_stride * 8 <= (size_t)148 * 8 * 256)
      wgrad_reduce_wide_kernel<<<(int)((a.split_stride * 8 + 255) / 256), 256, 0, s>>>(a.ws, pl.splits, a.split_stride, T, Cg, dw, accumulate ? 1 : 0);
    else
      wgrad_reduce_kernel<<<blocks, 256, 0, s>>>(a.ws, pl.splits, a.split_stride, T, Cg, dw, accumulate ? 1 : 0);
    count_launch();
  }
  SN_CUDA(cudaGetLastError());
}
void tcw_conv_wgrad(cudaStream_t s, const ConvGeom& g, const float* xflat, const float* dyflat, float* dw, bool accumulate) {
  const FlatGeom f = tc_flat_geom(g);
  wgrad_core(s, xflat, g.C, dyflat, g.Cout, tc_flat_rows(g), g.N * f.Gp, g.Og(), g.Cg(), g.group, g.kh * g.kw, g.kw, f.Wp, f.lead,
             g.ph * f.Wp + g.pw, dw, accumulate);
}
// inner-product backward-filter (inner_product_layer.cpp:96-102): dW[o][k] (+)= sum_b dY[b][o] * X[b][k]
bool tcw_matmul_wgrad_supported(int batch, int out, int K, const float* dy, const float* x, const float* dw) {
  if (batch < 1 || out % 4 != 0 || K % 16 != 0 || out < 16) return false;
  if (((uintptr_t)dy | (uintptr_t)x | (uintptr_t)dw) & 15) return false;
  return plan_wgrad(batch, out, K, 1, 1, 1, 1).ok;
}
void tcw_matmul_wgrad(cudaStream_t s, int batch, int out, int K, const float* dy_r, const float* x_r, float* dw, bool accumulate) {
  wgrad_core(s, x_r, K, dy_r, out, (size_t)batch, batch, out, K, 1, 1, 1, 1, 0, 0, dw, accumulate);
}

// ------------------------------------------------------------------------------------------
// Strided convolutions (AlexNet conv1: 11x11, stride 4) through the same window kernels: with s x s pixel blocks
// folded into channels ("space to depth"), a stride-s convolution with a k x k filter IS a stride-1 convolution
// with a ceil(k/s) x ceil(k/s) filter over C*s*s channels on the H/s x W/s grid:
//   y[n][co][ho][wo] = sum_{ky',kx'} sum_{c,dy,dx} xs[n][ho+ky'][wo+kx'][(c,dy,dx)] * w[co][c][s*ky'+dy][s*kx'+dx]
// (filter taps beyond k are zero).  conv1: 3 -> 48 channels, 11x11 -> 3x3, 227 -> 57, reduction 363 -> 432.
// ------------------------------------------------------------------------------------------
bool tcw_s2d_supported(const ConvGeom& g) {
  if (g.sh != g.sw || g.sh < 2 || g.sh > 8 || g.ph != 0 || g.pw != 0 || g.group != 1) return false;
  if ((g.C * g.sh * g.sw) % 16 != 0 || g.C * g.sh * g.sw > 256 || g.Cout % 16 != 0) return false;
  const ConvGeom e = tcw_s2d_geom(g);
  return tcw_supported(e) && tcw_wgrad_supported(e);
}
ConvGeom tcw_s2d_geom(const ConvGeom& g) {
  ConvGeom e = g;
  const int s = g.sh;
  e.C = g.C * s * s;
  e.H = (g.H + s - 1) / s; e.W = (g.W + s - 1) / s;
  e.kh = (g.kh + s - 1) / s; e.kw = (g.kw + s - 1) / s;
  e.sh = e.sw = 1; e.ph = e.pw = 0;
  // every output the strided layer has must exist on the folded grid: Ho <= H' - kh' + 1
  e.Ho = g.Ho; e.Wo = g.Wo;
  if (e.H - e.kh + 1 < g.Ho) e.H = g.Ho + e.kh - 1;
  if (e.W - e.kw + 1 < g.Wo) e.W = g.Wo + e.kw - 1;
  return e;
}
// one block per (image, folded row): s*C image rows in, W' positions x C*s*s channels out (contiguous).
// BYTES: the source is the byte-pixel staging buffer of a JavaData layer (runtime.cpp JavaDataLayer, byte input): the
// pixels are widened here, value = ((float)b - mean[(i / mean_div) % mean_period]) * scale with i the NCHW element index,
// the same two fp32 operations as kernels.cu widen_bytes — the fp32 copy of the minibatch is then never written or read.
struct ByteSrc { const uint8_t* b; int is_signed; const float* mean; FastDiv ddiv, dper; float scale; };
template <bool BYTES>
__global__ void __launch_bounds__(256) s2d_pack_kernel(const float* __restrict__ x, float* __restrict__ dst, int C, int H, int W, int s,
                                                        int Hs, int Ws, int Wp, int Gp, int lead, FastDiv dCs, FastDiv dS, ByteSrc bs) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  extern __shared__ float rows[];                      // [C*s][Ws*s]
  const int hs = blockIdx.x, n = blockIdx.y, Wpad = Ws * s, nrows = C * s;
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
  for (int r = ty; r < nrows; r += 4) {
    const int c = r / s, dy = r - c * s, h = hs * s + dy;
    const size_t row0 = (((size_t)n * C + c) * H + h) * W;
    for (int w = tx; w < Wpad; w += 64) {
      float v = 0.f;
      if (h < H && w < W) {
        if (BYTES) {
          const uint32_t i = (uint32_t)(row0 + w), raw = bs.b[i];
          v = bs.is_signed ? (float)(int)(signed char)raw : (float)raw;
          if (bs.mean) {
            const uint32_t q = bs.ddiv.div(i);
            v = __fsub_rn(v, __ldg(bs.mean + (q - bs.dper.div(q) * bs.dper.d)));
          }
          if (bs.scale != 1.f) v = __fmul_rn(v, bs.scale);
        } else {
          v = x[row0 + w];
        }
      }
      rows[r * Wpad + w] = v;
    }
  }
  __syncthreads();
  const int Cs = C * s * s;
  float* out = dst + ((size_t)lead + (size_t)n * Gp + (size_t)hs * Wp) * Cs;
  if (s == 4 && (((uintptr_t)out) & 15) == 0) {
    // the four dx of one (position, c, dy) are contiguous in the staged row AND in the output: float4 in, float4 out,
    // one division per four elements (the scalar loop below is bound by its index arithmetic, not by memory)
    const int R = Cs / 4;                              // (c, dy) pairs
    float4* out4 = reinterpret_cast<float4*>(out);
    for (int i4 = threadIdx.x; i4 < Ws * R; i4 += blockDim.x) {
      uint32_t ws, r;
      dCs.divmod((uint32_t)i4 * 4u, ws, r);            // i = 4 * i4 = ws * Cs + 4 * r
      r >>= 2;
      const float4 v = *reinterpret_cast<const float4*>(rows + r * Wpad + ws * 4);
      out4[i4] = make_float4(to_tf32_exact(v.x), to_tf32_exact(v.y), to_tf32_exact(v.z), to_tf32_exact(v.w));
    }
    return;
  }
  for (int i = threadIdx.x; i < Ws * Cs; i += blockDim.x) {
    uint32_t ws, ch, r, dx;
    dCs.divmod((uint32_t)i, ws, ch);
    dS.divmod(ch, r, dx);                              // r = c*s + dy
    out[i] = to_tf32_exact(rows[r * Wpad + ws * s + dx]);
  }
}
static void s2d_pack_launch(cudaStream_t st, const ConvGeom& g, const float* x, const ByteSrc* bs, float* xflat) {
  const ConvGeom e = tcw_s2d_geom(g);
  const FlatGeom f = tc_flat_geom(e);
  const int s = g.sh, Hs = (g.H + s - 1) / s, Ws = (g.W + s - 1) / s;
  const size_t smem = (size_t)g.C * s * Ws * s * sizeof(float);
  static size_t attr[2] = {48 * 1024, 48 * 1024};
  if (smem > attr[bs ? 1 : 0]) {
    if (bs) SN_CUDA(cudaFuncSetAttribute(s2d_pack_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    else SN_CUDA(cudaFuncSetAttribute(s2d_pack_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr[bs ? 1 : 0] = smem;
  }
  const FastDiv dCs((uint32_t)(g.C * s * s)), dS((uint32_t)s);
  if (bs) s2d_pack_kernel<true><<<dim3(Hs, g.N), 256, smem, st>>>(nullptr, xflat, g.C, g.H, g.W, s, Hs, Ws, f.Wp, f.Gp, f.lead, dCs, dS, *bs);
  else s2d_pack_kernel<false><<<dim3(Hs, g.N), 256, smem, st>>>(x, xflat, g.C, g.H, g.W, s, Hs, Ws, f.Wp, f.Gp, f.lead, dCs, dS, ByteSrc{});
  count_launch();
  SN_CUDA(cudaGetLastError());
}
void tcw_s2d_pack_input(cudaStream_t st, const ConvGeom& g, const float* x, float* xflat) { s2d_pack_launch(st, g, x, nullptr, xflat); }
void tcw_s2d_pack_input_bytes(cudaStream_t st, const ConvGeom& g, const uint8_t* bytes, bool is_signed, const float* mean, int mean_div,
                              int mean_period, float scale, float* xflat) {
  if ((size_t)g.N * g.C * g.H * g.W >= (1ull << 31)) fatal("s2d pack from bytes: tensor exceeds the 32-bit index range of the mean lookup");
  const ByteSrc bs{bytes, is_signed ? 1 : 0, mean, FastDiv((uint32_t)(mean_div > 0 ? mean_div : 1)), FastDiv((uint32_t)(mean_period > 0 ? mean_period : 1)), scale};
  s2d_pack_launch(st, g, nullptr, &bs, xflat);
}
// wf[co][tap'][slot]: slot = (c*s + dy)*s + dx < C*s*s holds w[co][c][s*ky'+dy][s*kx'+dx] (0 beyond the filter), tf32-rounded
__global__ void s2d_pack_filter_kernel(const float* __restrict__ w, float* __restrict__ wf, int Cout, int C, int kh, int kw, int s, int khs,
                                       int kws, int slots) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  const size_t total = (size_t)Cout * khs * kws * slots;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int slot = (int)(i % slots);
    size_t r = i / slots;
    const int tap = (int)(r % (khs * kws)), co = (int)(r / (khs * kws));
    float v = 0.f;
    if (slot < C * s * s) {
      const int dx = slot % s, dy = (slot / s) % s, c = slot / (s * s), ky = s * (tap / kws) + dy, kx = s * (tap % kws) + dx;
      if (ky < kh && kx < kw) v = w[(((size_t)co * C + c) * kh + ky) * kw + kx];
    }
    wf[i] = to_tf32_exact(v);
  }
}
void tcw_s2d_pack_filter(cudaStream_t st, const ConvGeom& g, const float* w, float* wf) {
  const ConvGeom e = tcw_s2d_geom(g);
  const int slots = slots_of(e.C);
  const size_t total = (size_t)g.Cout * e.kh * e.kw * slots;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  s2d_pack_filter_kernel<<<blocks, 256, 0, st>>>(w, wf, g.Cout, g.C, g.kh, g.kw, g.sh, e.kh, e.kw, slots);
  count_launch();
  SN_CUDA(cudaGetLastError());
}
// All stale filter copies of a net in ONE launch (Solver::step, start of an iteration): the weights of every layer
// change with every SGD update, and eight 5-microsecond pack kernels per CaffeNet iteration are pure launch latency.
// mode 0 / 1: tc_flat_pack_filter(dgrad = mode) layout; mode 2: the folded filter of s2d_pack_filter_kernel.
__global__ void filter_pack_batch_kernel(const FilterPackBatch b) {
  SN_GRID_DEP_LAUNCH_ALWAYS();
  SN_GRID_DEP_WAIT();
  const FilterPackJob& j = b.job[blockIdx.y];
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < j.total; i += (size_t)gridDim.x * blockDim.x) {
    float v = 0.f;
    if (j.mode == 2) {
      const int Cout = j.p[0], C = j.p[1], kh = j.p[2], kw = j.p[3], s = j.p[4], khs = j.p[5], kws = j.p[6], slots = j.p[7];
      (void)Cout;
      const int slot = (int)(i % slots);
      const size_t r = i / slots;
      const int tap = (int)(r % (khs * kws)), co = (int)(r / (khs * kws));
      if (slot < C * s * s) {
        const int dx = slot % s, dy = (slot / s) % s, c = slot / (s * s), ky = s * (tap / kws) + dy, kx = s * (tap % kws) + dx;
        if (ky < kh && kx < kw) v = j.w[(((size_t)co * C + c) * kh + ky) * kw + kx];
      }
    } else {
      const int Cg = j.p[0], Og = j.p[1], T = j.p[2], dgrad = j.mode;
      const int red = dgrad ? Og : Cg;
      const int slots = (red + 31) / 32 * 32, kp = T * slots;
      const int row = (int)(i / kp), k = (int)(i - (size_t)row * kp), tap = k / slots, ch = k - tap * slots;
      if (ch < red) {
        const int Csrc = j.p[3] > 0 ? j.p[3] : Cg;      // channel-padded first layers: the stored filter has Csrc < Cg channels
        if (!dgrad) v = ch < Csrc ? j.w[((size_t)row * Csrc + ch) * T + tap] : 0.f;
        else { const int grp = row / Cg, c = row - grp * Cg; v = j.w[((size_t)(grp * Og + ch) * Cg + c) * T + (T - 1 - tap)]; }
      }
    }
    j.dst[i] = to_tf32_exact(v);
  }
}
FilterPackJob tcw_filter_pack_job(const ConvGeom& g, int mode, const float* w, float* dst, int src_cg) {
  FilterPackJob j{};
  j.p[3] = mode == 0 ? src_cg : 0;
  j.w = w; j.dst = dst; j.mode = mode;
  if (mode == 2) {
    const ConvGeom e = tcw_s2d_geom(g);
    const int slots = slots_of(e.C);
    j.total = (size_t)g.Cout * e.kh * e.kw * slots;
    const int p[8] = {g.Cout, g.C, g.kh, g.kw, g.sh, e.kh, e.kw, slots};
    for (int i = 0; i < 8; ++i) j.p[i] = p[i];
  } else {
    j.total = tc_flat_filter_floats(g, mode == 1);
    j.p[0] = g.Cg(); j.p[1] = g.Og(); j.p[2] = g.kh * g.kw;
  }
  return j;
}
void tcw_filter_pack_batch(cudaStream_t st, const FilterPackBatch& b) {
  if (b.n <= 0) return;
  size_t most = 0;
  for (int i = 0; i < b.n; ++i) most = std::max(most, b.job[i].total);
  int blocks = (int)((most + 255) / 256);
  const int cap = std::max(148 * 16 / b.n, 32);       // ~2400 blocks in all: the largest copy (conv3 backward-data, 0.9 M floats) gets ~4 elements per thread
  if (blocks > cap) blocks = cap;
  filter_pack_batch_kernel<<<dim3(blocks, b.n), 256, 0, st>>>(b);
  count_launch();
  SN_CUDA(cudaGetLastError());
}
// Channel-padded first layers (C < 16 padded to 16 on the flat grid): dw[co][c][tap] (+)= dws[co][c][tap] for the real channels
__global__ void cpad_unpack_wgrad_kernel(const float* __restrict__ dws, float* __restrict__ dw, int Cout, int C, int Cpad, int T, int accumulate) {
  SN_GRID_DEP_LAUNCH_ALWAYS();
  SN_GRID_DEP_WAIT();
  const int total = Cout * C * T;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int r = i % (C * T), co = i / (C * T);
    const float v = dws[(size_t)co * Cpad * T + r];
    dw[i] = accumulate ? dw[i] + v : v;
  }
}
void tcw_cpad_unpack_wgrad(cudaStream_t st, const ConvGeom& g, int Cpad, const float* dws, float* dw, bool accumulate) {
  const int total = g.Cout * g.C * g.kh * g.kw;
  cpad_unpack_wgrad_kernel<<<(total + 255) / 256, 256, 0, st>>>(dws, dw, g.Cout, g.C, Cpad, g.kh * g.kw, accumulate ? 1 : 0);
  count_launch();
  SN_CUDA(cudaGetLastError());
}
// dw[co][c][ky][kx] (+)= dws[co][(c,dy,dx)][tap']  (the folded filter gradient, layout of tcw_conv_wgrad's output)
__global__ void s2d_unpack_wgrad_kernel(const float* __restrict__ dws, float* __restrict__ dw, int Cout, int C, int kh, int kw, int s, int khs,
                                        int kws, int accumulate) {
  SN_GRID_DEP_LAUNCH_ALWAYS();
  SN_GRID_DEP_WAIT();
  const size_t total = (size_t)Cout * C * kh * kw;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int kx = (int)(i % kw), ky = (int)((i / kw) % kh), c = (int)((i / ((size_t)kw * kh)) % C), co = (int)(i / ((size_t)kw * kh * C));
    const int ch = (c * s + ky % s) * s + kx % s, tap = (ky / s) * kws + kx / s;
    const float v = dws[((size_t)co * C * s * s + ch) * (khs * kws) + tap];
    dw[i] = accumulate ? dw[i] + v : v;
  }
}
void tcw_s2d_unpack_wgrad(cudaStream_t st, const ConvGeom& g, const float* dws, float* dw, bool accumulate) {
  const ConvGeom e = tcw_s2d_geom(g);
  const size_t total = (size_t)g.Cout * g.C * g.kh * g.kw;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  s2d_unpack_wgrad_kernel<<<blocks, 256, 0, st>>>(dws, dw, g.Cout, g.C, g.kh, g.kw, g.sh, e.kh, e.kw, accumulate ? 1 : 0);
  count_launch();
  SN_CUDA(cudaGetLastError());
}

}  // namespace sn
