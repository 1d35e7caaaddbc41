// gemm_tc.cu — TF32 tensor-core contractions for sm_100a: tcgen05.mma with the accumulator in TMEM,
// operands staged in 128-byte-swizzled shared memory by TMA (weights) and by gather warps (the
// implicit-GEMM views of NCHW activations), mbarrier producer/consumer pipeline, epilogue straight
// from TMEM (bias + ReLU fused).  Replaces the reference's im2col + cublasSgemm loops
// (caffe/src/caffe/layers/base_conv_layer.cpp:243-376, caffe/src/caffe/util/im2col.cu:9-58,
// caffe/src/caffe/layers/inner_product_layer.cu:10-56).
//
// Every contraction is  D[m][n] = sum_r A(m,r) * B(n,r)  with a 128 x BN accumulator tile per CTA:
//   warps 0-15 gather/convert (four groups of four warps, one group per pipeline stage) the operands that cannot come through TMA (fp32 -> tf32 with
//              round-to-nearest; TMA-fed weights come from a pre-rounded shadow copy), then run the
//              epilogue (tcgen05.ld: warp w owns TMEM lanes 32(w%4)..+31 = tile rows);
//   warp 16    allocates TMEM and issues tcgen05.mma (one thread), 4 x (128 x n_mma x 8) per 32-deep chunk;
//   warp 17    issues the TMA loads; warp 18 builds the per-chunk reduction-index tables.
// Reductions that leave too few tiles for 148 SMs are split along r; the slices are summed in a
// fixed order by splitk_reduce (deterministic, no atomics).
#include <cuda.h>

#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <tuple>
#include <type_traits>

#include "gemm_ops.cuh"
#include "kernels.hpp"
#include "tc_prims.cuh"

namespace sn {
using namespace tc;

constexpr int TBM = 128, TBK = 32, TSTAGES = 4;
constexpr int kATileBytes = TBM * TBK * 4;            // 16 KB

// ------------------------------------------------------------------------------------------
// Operand loaders.  kTMA: issued by the TMA warp.  Otherwise the operand is gathered by ONE gather group
// of GT = 128 threads per chunk; the NT = 512 gather threads form NG = 4 groups and group g fills pipeline
// stage g (chunks g, g+4, ...), so four chunks are in flight.  A gather is split in two phases:
//   fetch()   issues ALL global loads of the chunk into registers — it does not need the smem stage, so it
//             runs BEFORE the group waits for the stage to be free (load latency overlaps the MMA that still
//             reads the stage);
//   commit()  rounds to tf32 (one integer add) and stores into the swizzled tile.
// Loaders whose tile has more than 128 K-major rows run fetch/commit a second time (`pass` = 1).
// kMN: MN-major tile (element (mn, k) at off_mnmajor) else K-major (off_kmajor).  `rows` = tile rows
// actually needed (128 for A, n_mma for B); rows beyond the operand's extent are zero-filled.
// ------------------------------------------------------------------------------------------
constexpr int NT = 512;
constexpr int NW = NT / 32;                   // gather warps (all of them run the epilogue)
constexpr int GT = 128, GW = GT / 32;         // threads / warps per gather group
constexpr int NG = NT / GT;                   // gather groups == pipeline stages
struct NoState {};
struct NoRegs {};

__device__ __forceinline__ uint32_t tf32_bits(float x) { return __float_as_uint(x) + 0x1000u; }
__device__ __forceinline__ float4 tf32x4(float4 v) {
  v.x = to_tf32(v.x); v.y = to_tf32(v.y); v.z = to_tf32(v.z); v.w = to_tf32(v.w);
  return v;
}
// Keep a per-thread base pointer as an opaque 64-bit value: without this the compiler re-derives it from
// the kernel parameter and a 64-bit element offset at every use (4 instructions per gathered element).
__device__ __forceinline__ const float* opaque(const float* p) {
  asm volatile("" : "+l"(p));
  return p;
}

// weights through TMA, K-major: tensor {K (contiguous), rows}; box {32, rows_box}
struct LoadTmaK {
  static constexpr bool kTMA = true, kMN = false, kKTable = false;
  typedef NoState State;
  typedef NoRegs Regs;
  int rows_box;
  __device__ uint32_t bytes() const { return (uint32_t)rows_box * 128u; }
  __device__ void issue(const CUtensorMap* tm, uint8_t* tile, uint64_t* bar, int row0, int r0) const {
    tma_load_2d(tile, tm, bar, r0, row0);
  }
};
// weights through TMA, MN-major: tensor {MN (contiguous), R rows}; box {32, 32} with the 32-byte-atom
// 128-byte swizzle (CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B); one box per 32 rows
struct LoadTmaMN {
  static constexpr bool kTMA = true, kMN = true, kKTable = false;
  typedef NoState State;
  typedef NoRegs Regs;
  int rows;                                       // multiple of 32
  __device__ uint32_t bytes() const { return (uint32_t)(rows / 32) * 4096u; }
  __device__ void issue(const CUtensorMap* tm, uint8_t* tile, uint64_t* bar, int row0, int r0) const {
    for (int i = 0; i < rows / 32; ++i) tma_load_2d(tile + i * 4096, tm, bar, row0 + 32 * i, r0);
  }
};

// dense [row][r] (r contiguous), K-major tile.  128-bit path when the pitch allows it: thread t owns the
// 16-byte chunk t%8 of rows t/8 + 16*i.  One pass = 64 rows.
struct LoadDenseK {
  static constexpr bool kTMA = false, kMN = false, kKTable = false;
  typedef NoState State;
  struct Regs { float4 v[4]; };
  const float* p; int ld, nrows, nred, rows; int vec;
  __device__ int passes() const { return (rows + 63) / 64; }
  __device__ void init(State&, int, int) const {}
  __device__ void fetch(State&, Regs& rg, int pass, int row0, int r0, int r_end, int t, const void*) const {
    const int rb = pass * 64;
    if (vec) {
      const int c = t & 7, r = r0 + 4 * c;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int row = rb + (t >> 3) + 16 * i, gr = row0 + row;
        rg.v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row < rows && gr < nrows && r < r_end) {
          const float* q = p + (size_t)gr * ld + r;
          if (r + 3 < r_end) rg.v[i] = __ldg((const float4*)q);
          else { rg.v[i].x = __ldg(q); if (r + 1 < r_end) rg.v[i].y = __ldg(q + 1); if (r + 2 < r_end) rg.v[i].z = __ldg(q + 2); }
        }
      }
    } else {                                      // scalar: lane = k, warp w owns rows rb + w + 4*j
      const int k = t & 31, r = r0 + k, w = t >> 5;
      float* f = (float*)rg.v;
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int row = rb + w + GW * j, gr = row0 + row;
        f[j] = (row < rows && gr < nrows && r < r_end) ? __ldg(p + (size_t)gr * ld + r) : 0.f;
      }
    }
  }
  __device__ void commit(State&, const Regs& rg, int pass, uint8_t* tile, int t) const {
    const int rb = pass * 64;
    if (vec) {
      const int c = t & 7;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int row = rb + (t >> 3) + 16 * i;
        if (row < rows) *(float4*)(tile + off_kmajor(row, 4 * c)) = tf32x4(rg.v[i]);
      }
    } else {
      const int k = t & 31, w = t >> 5;
      const float* f = (const float*)rg.v;
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int row = rb + w + GW * j;
        if (row < rows) *(float*)(tile + off_kmajor(row, k)) = to_tf32(f[j]);
      }
    }
  }
};
// dense [r][mn] (mn contiguous), MN-major tile.  Warp w of the group owns k = 8w..8w+7; lanes run along
// mn in 16-byte chunks.  Pass p covers mn block p/2 (128 wide) and k half p%2 (4 of the warp's 8 k).
struct LoadDenseMN {
  static constexpr bool kTMA = false, kMN = true, kKTable = false;
  typedef NoState State;
  struct Regs { float4 v[4]; };
  const float* p; int ld, nrows, nred, rows; int vec;
  __device__ int passes() const { return 2 * ((rows + 127) / 128); }
  __device__ void init(State&, int, int) const {}
  __device__ void fetch(State&, Regs& rg, int pass, int row0, int r0, int r_end, int t, const void*) const {
    const int w = t >> 5, lane = t & 31, mb = pass >> 1, kh = (pass & 1) * 4;
    if (vec) {
      const int ch = rows >> 2, c = 32 * mb + lane, mn = row0 + 4 * c;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int r = r0 + 8 * w + kh + i;
        rg.v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (c < ch && r < r_end && mn < nrows) {
          const float* q = p + (size_t)r * ld + mn;
          if (mn + 3 < nrows) rg.v[i] = __ldg((const float4*)q);
          else { rg.v[i].x = __ldg(q); if (mn + 1 < nrows) rg.v[i].y = __ldg(q + 1); if (mn + 2 < nrows) rg.v[i].z = __ldg(q + 2); }
        }
      }
    } else {                                      // scalar: slot j = (kk, q): k = 8w + kh + kk, mn = 128*mb + lane + 32*q
      float* f = (float*)rg.v;
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int k = 8 * w + kh + (j >> 2), m = 128 * mb + lane + 32 * (j & 3), mn = row0 + m, r = r0 + k;
        f[j] = (m < rows && r < r_end && mn < nrows) ? __ldg(p + (size_t)r * ld + mn) : 0.f;
      }
    }
  }
  __device__ void commit(State&, const Regs& rg, int pass, uint8_t* tile, int t) const {
    const int w = t >> 5, lane = t & 31, mb = pass >> 1, kh = (pass & 1) * 4;
    if (vec) {
      const int ch = rows >> 2, c = 32 * mb + lane;
#pragma unroll
      for (int i = 0; i < 4; ++i)
        if (c < ch) *(float4*)(tile + off_mnmajor(4 * c, 8 * w + kh + i)) = tf32x4(rg.v[i]);
    } else {
      const float* f = (const float*)rg.v;
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int k = 8 * w + kh + (j >> 2), m = 128 * mb + lane + 32 * (j & 3);
        if (m < rows) *(float*)(tile + off_mnmajor(m, k)) = to_tf32(f[j]);
      }
    }
  }
};

// Any accessor of gemm_ops.cuh as an MN-major tile (thread = row): generic fallback (strided backward-data).
template <class Acc>
struct LoadAccMN {
  static constexpr bool kTMA = false, kMN = true, kKTable = false;
  typedef NoState State;
  struct Regs { float v[16]; };
  Acc acc; int rows;
  __device__ int passes() const { return 2; }
  __device__ void init(State&, int, int) const {}
  __device__ void fetch(State&, Regs& rg, int pass, int row0, int r0, int r_end, int t, const void*) const {
#pragma unroll 4
    for (int k = 0; k < 16; ++k) {
      const int r = r0 + 16 * pass + k;
      rg.v[k] = (t < rows && r < r_end) ? acc(row0 + t, r) : 0.f;
    }
  }
  __device__ void commit(State&, const Regs& rg, int pass, uint8_t* tile, int t) const {
    if (t >= rows) return;
#pragma unroll
    for (int k = 0; k < 16; ++k) *(float*)(tile + off_mnmajor(t, 16 * pass + k)) = to_tf32(rg.v[k]);
  }
};

// Reduction-index table for the two pixel-row gathers below, built once per CTA in shared memory for the
// CTA's whole reduction range: entry r = {input offset delta, tap mask}.  Filters with kh*kw <= 32 taps
// (SMALL) test ONE per-thread 2-D validity mask (bit kh_i*kw + kw_i set when that tap reads inside the
// image) against 1 << tap; larger filters (11x11) carry a second mask word (row mask in .y, column mask in
// .z against per-thread hmask / wmask).  An all-zero mask marks an index beyond the reduction.  Loads are
// unconditional plain LDGs: a padded tap reads a neighbouring row of the tensor or its guard band
// (runtime.hpp kGuardFloats) and the value is replaced by zero at commit time.
template <bool SMALL> struct KEntry { typedef int2 type; };
template <> struct KEntry<false> { typedef int4 type; };
__device__ __forceinline__ uint32_t mask2d(uint32_t hmask, uint32_t wmask, int kh, int kw) {
  uint32_t m = 0;
  for (int k = 0; k < kh; ++k) if ((hmask >> k) & 1u) m |= wmask << (k * kw);
  return m;
}
template <bool SMALL> __device__ __forceinline__ typename KEntry<SMALL>::type make_kentry(int delta, uint32_t tap, uint32_t ky, uint32_t kx);
template <> __device__ __forceinline__ int2 make_kentry<true>(int delta, uint32_t tap, uint32_t, uint32_t) { return make_int2(delta, 1 << tap); }
template <> __device__ __forceinline__ int4 make_kentry<false>(int delta, uint32_t, uint32_t ky, uint32_t kx) { return make_int4(delta, 1 << ky, 1 << kx, 0); }
__device__ __forceinline__ int2 zero_kentry(int2) { return make_int2(0, 0); }
__device__ __forceinline__ int4 zero_kentry(int4) { return make_int4(0, 0, 0, 0); }

// Convolution forward A operand: im2col rows (output pixels) x (c,kh,kw), MN-major.  Thread t of the
// group owns pixel row0 + t and all 32 reduction indices of the chunk.  Its (n,ho,wo), input base and
// padding masks are computed once per tile.  Lanes = consecutive pixels, so global reads are coalesced
// along W and the swizzled 128-byte rows are written conflict-free.
template <bool SMALL>
struct LoadIm2colRowsMN {
  static constexpr bool kTMA = false, kMN = true, kKTable = true;
  __device__ int passes() const { return 2; }
  typedef typename KEntry<SMALL>::type Entry;
  const float* x; ConvGeom g; int grp;
  FastDiv dHoWo, dWo, dKhKw, dKw;
  struct State { const float* base; uint32_t ma, mb; };
  struct Regs { float v[16]; uint32_t ok; };
  __device__ Entry table_entry(int r, int r_end) const {
    if (r >= r_end || r >= g.Kdim()) return zero_kentry(Entry());
    uint32_t c, tt, ky, kx;
    dKhKw.divmod((uint32_t)r, c, tt);
    dKw.divmod(tt, ky, kx);
    return make_kentry<SMALL>((int)c * g.H * g.W + (int)ky * g.W + (int)kx, tt, ky, kx);
  }
  __device__ void init(State& s, int row0, int t) const {
    const int row = row0 + t;
    const bool live = row < g.N * g.Ho * g.Wo;
    uint32_t n = 0, pix = 0, ho = 0, wo = 0;
    if (live) { dHoWo.divmod(row, n, pix); dWo.divmod(pix, ho, wo); }
    const int hi0 = (int)ho * g.sh - g.ph, wi0 = (int)wo * g.sw - g.pw;
    uint32_t hmask = 0, wmask = 0;
    for (int k = 0; k < g.kh; ++k) if (hi0 + k >= 0 && hi0 + k < g.H) hmask |= 1u << k;
    for (int k = 0; k < g.kw; ++k) if (wi0 + k >= 0 && wi0 + k < g.W) wmask |= 1u << k;
    if (!live || !wmask) hmask = 0;
    s.base = opaque(hmask ? x + ((ptrdiff_t)((int)n * g.C + grp * g.Cg()) * g.H * g.W + hi0 * g.W + wi0) : x);
    if (SMALL) { s.ma = mask2d(hmask, wmask, g.kh, g.kw); s.mb = 0; } else { s.ma = hmask; s.mb = wmask; }
  }
  __device__ void fetch(State& s, Regs& rg, int pass, int, int, int, int, const void* ktab_chunk) const {
    const Entry* kt = (const Entry*)ktab_chunk + 16 * pass;
    rg.ok = 0;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      const Entry e = kt[k];
      bool p;
      if constexpr (SMALL) p = (s.ma & (uint32_t)e.y) != 0;
      else p = ((s.ma & (uint32_t)e.y) != 0) & ((s.mb & (uint32_t)e.z) != 0);
      rg.v[k] = __ldg(s.base + e.x);
      rg.ok |= (p ? 1u : 0u) << k;
    }
  }
  __device__ void commit(State&, const Regs& rg, int pass, uint8_t* tile, int t) const {
#pragma unroll
    for (int k = 0; k < 16; ++k)
      *(uint32_t*)(tile + off_mnmajor(t, 16 * pass + k)) = ((rg.ok >> k) & 1u) ? tf32_bits(rg.v[k]) : 0u;
  }
};

// Convolution backward-data A operand (stride 1): rows = input pixels, r = (co,kh,kw), MN-major.
// value = dY[n][co][h+ph-kh][w+pw-kw] when that output position exists.
template <bool SMALL>
struct LoadDgradRowsMN {
  static constexpr bool kTMA = false, kMN = true, kKTable = true;
  __device__ int passes() const { return 2; }
  typedef typename KEntry<SMALL>::type Entry;
  const float* dy; ConvGeom g; int grp;
  FastDiv dHW, dW, dKhKw, dKw;
  struct State { const float* base; uint32_t ma, mb; };
  struct Regs { float v[16]; uint32_t ok; };
  __device__ Entry table_entry(int r, int r_end) const {
    if (r >= r_end || r >= g.Og() * g.kh * g.kw) return zero_kentry(Entry());
    uint32_t co, tt, ky, kx;
    dKhKw.divmod((uint32_t)r, co, tt);
    dKw.divmod(tt, ky, kx);
    return make_kentry<SMALL>((int)co * g.Ho * g.Wo - (int)ky * g.Wo - (int)kx, tt, ky, kx);
  }
  __device__ void init(State& s, int row0, int t) const {
    const int row = row0 + t;
    const bool live = row < g.N * g.H * g.W;
    uint32_t n = 0, pix = 0, h = 0, w = 0;
    if (live) { dHW.divmod(row, n, pix); dW.divmod(pix, h, w); }
    const int hb = (int)h + g.ph, wb = (int)w + g.pw;            // output row = hb - ky, col = wb - kx
    uint32_t hmask = 0, wmask = 0;
    for (int k = 0; k < g.kh; ++k) if (hb - k >= 0 && hb - k < g.Ho) hmask |= 1u << k;
    for (int k = 0; k < g.kw; ++k) if (wb - k >= 0 && wb - k < g.Wo) wmask |= 1u << k;
    if (!live || !wmask) hmask = 0;
    s.base = opaque(hmask ? dy + ((ptrdiff_t)((int)n * g.Cout + grp * g.Og()) * g.Ho * g.Wo + hb * g.Wo + wb) : dy);
    if (SMALL) { s.ma = mask2d(hmask, wmask, g.kh, g.kw); s.mb = 0; } else { s.ma = hmask; s.mb = wmask; }
  }
  __device__ void fetch(State& s, Regs& rg, int pass, int, int, int, int, const void* ktab_chunk) const {
    const Entry* kt = (const Entry*)ktab_chunk + 16 * pass;
    rg.ok = 0;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      const Entry e = kt[k];
      bool p;
      if constexpr (SMALL) p = (s.ma & (uint32_t)e.y) != 0;
      else p = ((s.ma & (uint32_t)e.y) != 0) & ((s.mb & (uint32_t)e.z) != 0);
      rg.v[k] = __ldg(s.base + e.x);
      rg.ok |= (p ? 1u : 0u) << k;
    }
  }
  __device__ void commit(State&, const Regs& rg, int pass, uint8_t* tile, int t) const {
#pragma unroll
    for (int k = 0; k < 16; ++k)
      *(uint32_t*)(tile + off_mnmajor(t, 16 * pass + k)) = ((rg.ok >> k) & 1u) ? tf32_bits(rg.v[k]) : 0u;
  }
};
// Backward-data B operand: filters viewed as [c][(co,kh,kw)], K-major.  lane = r: its (co, tap) is
// decomposed once per chunk; rows (input channels) are khkw apart.  Warp w of the group owns rows
// w + 4*j; one pass covers 128 rows.
struct LoadFilterDgradK {
  static constexpr bool kTMA = false, kMN = false, kKTable = false;
  typedef NoState State;
  struct Regs { float v[16]; int lim; };
  const float* w; ConvGeom g; int grp; int rows;
  __device__ int passes() const { return (rows + 63) / 64; }
  FastDiv dKhKw;
  __device__ void init(State&, int, int) const {}
  __device__ void fetch(State&, Regs& rg, int pass, int row0, int r0, int r_end, int t, const void*) const {
    const int khkw = g.kh * g.kw, Cg = g.Cg();
    const int k = t & 31, r = r0 + k, wp = t >> 5, rb = 64 * pass;
    const bool ok = r < r_end && r < g.Og() * khkw;
    uint32_t co = 0, tap = 0;
    if (ok) dKhKw.divmod(r, co, tap);
    const float* q = opaque(w + ((size_t)(grp * g.Og() + co) * Cg + row0) * khkw + tap);
    rg.lim = ok ? min(rows, Cg - row0) : 0;                       // valid tile rows
#pragma unroll
    for (int j = 0; j < 16; ++j) rg.v[j] = 0.f;
    if (rg.lim > rb) {
#pragma unroll
      for (int j = 0; j < 16; ++j) rg.v[j] = __ldg(q + (ptrdiff_t)min(rb + wp + GW * j, rg.lim - 1) * khkw);
    }
  }
  __device__ void commit(State&, const Regs& rg, int pass, uint8_t* tile, int t) const {
    const int k = t & 31, wp = t >> 5, rb = 64 * pass;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int row = rb + wp + GW * j;
      if (row < rows) *(uint32_t*)(tile + off_kmajor(row, k)) = row < rg.lim ? tf32_bits(rg.v[j]) : 0u;
    }
  }
};
// Backward-filter A operand: dY viewed as [co][(n,ho,wo)], K-major.  lane = r = (n, pix).
struct LoadDyRowsK {
  static constexpr bool kTMA = false, kMN = false, kKTable = false;
  typedef NoState State;
  struct Regs { float v[16]; int lim; };
  const float* dy; ConvGeom g; int grp;
  FastDiv dHoWo;
  __device__ int passes() const { return 2; }
  __device__ void init(State&, int, int) const {}
  __device__ void fetch(State&, Regs& rg, int pass, int row0, int r0, int r_end, int t, const void*) const {
    const int HoWo = g.Ho * g.Wo;
    const int k = t & 31, r = r0 + k, wp = t >> 5, rb = 64 * pass;
    const bool ok = r < r_end && r < g.N * HoWo;
    uint32_t n = 0, pix = 0;
    if (ok) dHoWo.divmod(r, n, pix);
    const float* q = opaque(dy + ((size_t)n * g.Cout + grp * g.Og() + row0) * HoWo + pix);
    rg.lim = ok ? min(TBM, g.Og() - row0) : 0;
#pragma unroll
    for (int j = 0; j < 16; ++j) rg.v[j] = 0.f;
    if (rg.lim > rb) {
#pragma unroll
      for (int j = 0; j < 16; ++j) rg.v[j] = __ldg(q + (ptrdiff_t)min(rb + wp + GW * j, rg.lim - 1) * HoWo);
    }
  }
  __device__ void commit(State&, const Regs& rg, int pass, uint8_t* tile, int t) const {
    const int k = t & 31, wp = t >> 5, rb = 64 * pass;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int row = rb + wp + GW * j;
      *(uint32_t*)(tile + off_kmajor(row, k)) = row < rg.lim ? tf32_bits(rg.v[j]) : 0u;
    }
  }
};
// Backward-filter B operand: im2col viewed as [(c,kh,kw)][(n,ho,wo)], K-major.  lane = r: its
// (n,ho,wo), input base and padding masks are computed once per chunk; the per-row (c,kh,kw) comes
// from a table built once per CTA in shared memory (same entry format as the reduction-index table;
// zero masks mark rows beyond the filter).  One pass covers 128 rows.
template <bool SMALL>
struct LoadIm2colColsK {
  static constexpr bool kTMA = false, kMN = false, kKTable = false;
  typedef typename KEntry<SMALL>::type Entry;
  const float* x; ConvGeom g; int grp; int rows;
  FastDiv dHoWo, dWo, dKhKw, dKw;
  struct State { const Entry* tab; };
  struct Regs { float v[16]; uint32_t ok; };
  __device__ int passes() const { return (rows + 63) / 64; }
  static constexpr int kTableRows = 256;
  __device__ void build_table(void* tab_, int row0, int t) const {
    Entry* tab = (Entry*)tab_;
    for (int row = t; row < kTableRows; row += kBuildThreads) {
      const int rr = row0 + row;
      Entry e = zero_kentry(Entry());
      if (row < rows && rr < g.Kdim()) {
        uint32_t c, tt, ky, kx;
        dKhKw.divmod(rr, c, tt);
        dKw.divmod(tt, ky, kx);
        e = make_kentry<SMALL>((int)c * g.H * g.W + (int)ky * g.W + (int)kx, tt, ky, kx);
      }
      tab[row] = e;
    }
  }
  static constexpr int kBuildThreads = NT;
  __device__ void init(State&, int, int) const {}
  __device__ void fetch(State& s, Regs& rg, int pass, int, int r0, int r_end, int t, const void*) const {
    const int k = t & 31, r = r0 + k, wp = t >> 5, rb = 64 * pass;
    const bool ok = r < r_end && r < g.N * g.Ho * g.Wo;
    uint32_t n = 0, pix = 0, ho = 0, wo = 0;
    if (ok) { dHoWo.divmod(r, n, pix); dWo.divmod(pix, ho, wo); }
    const int hi0 = (int)ho * g.sh - g.ph, wi0 = (int)wo * g.sw - g.pw;
    uint32_t hmask = 0, wmask = 0;
    for (int q = 0; q < g.kh; ++q) if (hi0 + q >= 0 && hi0 + q < g.H) hmask |= 1u << q;
    for (int q = 0; q < g.kw; ++q) if (wi0 + q >= 0 && wi0 + q < g.W) wmask |= 1u << q;
    if (!ok || !wmask) hmask = 0;
    uint32_t ma, mb;
    if (SMALL) { ma = mask2d(hmask, wmask, g.kh, g.kw); mb = 0; } else { ma = hmask; mb = wmask; }
    const float* base = opaque(hmask ? x + ((ptrdiff_t)((int)n * g.C + grp * g.Cg()) * g.H * g.W + hi0 * g.W + wi0) : x);
    rg.ok = 0;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const Entry e = s.tab[rb + wp + GW * j];                // rows beyond the filter: zero mask, offset 0
      bool p;
      if constexpr (SMALL) p = (ma & (uint32_t)e.y) != 0;
      else p = ((ma & (uint32_t)e.y) != 0) & ((mb & (uint32_t)e.z) != 0);
      rg.v[j] = __ldg(base + e.x);
      rg.ok |= (p ? 1u : 0u) << j;
    }
  }
  __device__ void commit(State&, const Regs& rg, int pass, uint8_t* tile, int t) const {
    const int k = t & 31, wp = t >> 5, rb = 64 * pass;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int row = rb + wp + GW * j;
      if (row < rows) *(uint32_t*)(tile + off_kmajor(row, k)) = ((rg.ok >> j) & 1u) ? tf32_bits(rg.v[j]) : 0u;
    }
  }
};

// ------------------------------------------------------------------------------------------
// Epilogues: row(m) prepares the per-row context once per thread, store(ctx, n, acc) writes D[m][n].
// ------------------------------------------------------------------------------------------
struct TEpiConvFwd {       // m = output pixel, n = co in group; y NCHW; bias + optional ReLU
  float* y; const float* bias; int P, HoWo, Cout, Og, grp, relu; FastDiv dHoWo;
  struct Ctx { float* p; bool ok; };
  __device__ Ctx row(int m) const {
    Ctx c; c.ok = m < P;
    uint32_t img = 0, pix = 0;
    if (c.ok) dHoWo.divmod(m, img, pix);
    c.p = y + ((size_t)img * Cout + grp * Og) * HoWo + pix;
    return c;
  }
  __device__ void store(const Ctx& c, int n, float v) const {
    if (!c.ok || n >= Og) return;
    if (bias) v += __ldg(bias + grp * Og + n);
    if (relu) v = fmaxf(v, 0.f);
    c.p[(size_t)n * HoWo] = v;
  }
};
struct TEpiConvDgrad {     // m = input pixel, n = c in group
  float* dx; int P, HW, C, Cg, grp; FastDiv dHW;
  struct Ctx { float* p; bool ok; };
  __device__ Ctx row(int m) const {
    Ctx c; c.ok = m < P;
    uint32_t img = 0, pix = 0;
    if (c.ok) dHW.divmod(m, img, pix);
    c.p = dx + ((size_t)img * C + grp * Cg) * HW + pix;
    return c;
  }
  __device__ void store(const Ctx& c, int n, float v) const {
    if (c.ok && n < Cg) c.p[(size_t)n * HW] = v;
  }
};
struct TEpiRowMajor {      // p[m*ld + n] (+)= v
  float* p; int ld, M, N, accumulate;
  struct Ctx { float* q; bool ok; };
  __device__ Ctx row(int m) const { Ctx c; c.ok = m < M; c.q = p + (size_t)m * ld; return c; }
  __device__ void store(const Ctx& c, int n, float v) const {
    if (c.ok && n < N) c.q[n] = accumulate ? c.q[n] + v : v;
  }
};
struct TEpiTransposed {    // p[n*ld + m] = v + bias[m] (+ReLU)
  float* p; int ld, M, N; const float* bias; int relu;
  struct Ctx { float* q; float b; bool ok; };
  __device__ Ctx row(int m) const {
    Ctx c; c.ok = m < M; c.q = p + m; c.b = (c.ok && bias) ? __ldg(bias + m) : 0.f;
    return c;
  }
  __device__ void store(const Ctx& c, int n, float v) const {
    if (!c.ok || n >= N) return;
    v += c.b;
    if (relu) v = fmaxf(v, 0.f);
    c.q[(size_t)n * ld] = v;
  }
};

// ------------------------------------------------------------------------------------------
// The kernel
// ------------------------------------------------------------------------------------------
template <class L> struct HasRowTable : std::false_type {};
template <bool S> struct HasRowTable<LoadIm2colColsK<S>> : std::true_type {};
template <class L, bool K = L::kKTable> struct KTabEntryBytes { static constexpr int value = 0; };
template <class L> struct KTabEntryBytes<L, true> { static constexpr int value = (int)sizeof(typename L::Entry); };

template <int BN>
struct TcCfg {
  static constexpr int kBTileBytes = BN * TBK * 4;
  static constexpr int kStageBytes = kATileBytes + kBTileBytes;
  static constexpr int kSmemBytes = TSTAGES * kStageBytes + 1024;     // + slack to align the tiles to 1024 B
  static constexpr uint32_t kTmemCols = BN <= 32 ? 32 : BN <= 64 ? 64 : BN <= 128 ? 128 : 256;
};

constexpr int kTcThreads = NT + 64;          // 16 gather/epilogue warps + MMA warp + TMA warp
constexpr int kMmaWarp = NW, kTmaWarp = NW + 1;

template <class AL, class BL, class EP, int BN, bool SPLIT>
__global__ void __launch_bounds__(kTcThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, AL al, BL bl, EP ep,
               int M, int N, int R, int r_split_len, int n_mma, float* __restrict__ ws) {
  typedef TcCfg<BN> Cfg;
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full_bar[TSTAGES], empty_bar[TSTAGES], tmem_full_bar;
  __shared__ uint32_t tmem_base_smem;
  __shared__ int4 row_table[256];                              // backward-filter: per-CTA row table
  SN_GRID_DEP_LAUNCH();

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int m0 = blockIdx.x * TBM, n0 = blockIdx.y * n_mma;
  const int r_begin = blockIdx.z * r_split_len;
  const int r_end = min(R, r_begin + r_split_len);
  const int nchunks = r_end > r_begin ? (r_end - r_begin + TBK - 1) / TBK : 0;

  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* ktab = smem + TSTAGES * Cfg::kStageBytes;           // reduction-index table (pixel-row gathers)
  constexpr bool kAnySimt = !AL::kTMA || !BL::kTMA, kAnyTma = AL::kTMA || BL::kTMA;
  static_assert(!BL::kKTable, "the reduction-index table is wired for the A operand only");

  if (tid == 0) {
    for (int s = 0; s < TSTAGES; ++s) {
      mbar_init(&full_bar[s], (kAnySimt ? (uint32_t)GT : 0u) + (kAnyTma ? 1u : 0u));
      mbar_init(&empty_bar[s], 1u);
    }
    mbar_init(&tmem_full_bar, 1u);
    fence_barrier_init();
    if (AL::kTMA) tma_prefetch_desc(&tmA);
    if (BL::kTMA) tma_prefetch_desc(&tmB);
  }
  if (warp == kMmaWarp) tmem_alloc<Cfg::kTmemCols>(&tmem_base_smem);
  if constexpr (HasRowTable<BL>::value) {
    if (warp < NW) bl.build_table(row_table, n0, tid);
  }
  if constexpr (AL::kKTable) {
    typedef typename AL::Entry Entry;
    for (int i = tid; i < nchunks * TBK; i += kTcThreads) ((Entry*)ktab)[i] = al.table_entry(r_begin + i, r_end);
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  SN_GRID_DEP_WAIT();                        // barriers, TMEM and the index tables (pure geometry) are set up; operands are read below
  const uint32_t tmem_base = tmem_base_smem;

  if (warp < NW) {
    // ---------------- gather producers ----------------
    if (kAnySimt && nchunks > 0) {
      static_assert(NG == TSTAGES, "one gather group per pipeline stage");
      const int grp = warp / GW, tg = tid & (GT - 1);          // group g fills stage g: chunks g, g+NG, ...
      typename AL::State sa;
      typename BL::State sb;
      typename AL::Regs ra;
      typename BL::Regs rb;
      if constexpr (!AL::kTMA) al.init(sa, m0, tg);
      if constexpr (!BL::kTMA) {
        bl.init(sb, n0, tg);
        if constexpr (HasRowTable<BL>::value) sb.tab = (decltype(sb.tab))row_table;
      }
      int a_passes = 0, b_passes = 0;
      if constexpr (!AL::kTMA) a_passes = al.passes();
      if constexpr (!BL::kTMA) b_passes = bl.passes();
      const int max_passes = max(a_passes, b_passes);
      const int s = grp;
      uint8_t* a_tile = smem + s * Cfg::kStageBytes;
      uint8_t* b_tile = a_tile + kATileBytes;
      for (int kc = grp; kc < nchunks; kc += NG) {
        const int r0 = r_begin + kc * TBK;
        const void* kt = ktab + (size_t)kc * TBK * KTabEntryBytes<AL>::value;
        // pass 0 global loads first: they do not need the stage, so their latency overlaps the MMA still reading it
        if constexpr (!AL::kTMA) al.fetch(sa, ra, 0, m0, r0, r_end, tg, kt);
        if constexpr (!BL::kTMA) bl.fetch(sb, rb, 0, n0, r0, r_end, tg, nullptr);
        mbar_wait(&empty_bar[s], ((kc / TSTAGES) & 1) ^ 1);
        if constexpr (!AL::kTMA) al.commit(sa, ra, 0, a_tile, tg);
        if constexpr (!BL::kTMA) bl.commit(sb, rb, 0, b_tile, tg);
        for (int pass = 1; pass < max_passes; ++pass) {
          if constexpr (!AL::kTMA) { if (pass < a_passes) al.fetch(sa, ra, pass, m0, r0, r_end, tg, kt); }
          if constexpr (!BL::kTMA) { if (pass < b_passes) bl.fetch(sb, rb, pass, n0, r0, r_end, tg, nullptr); }
          if constexpr (!AL::kTMA) { if (pass < a_passes) al.commit(sa, ra, pass, a_tile, tg); }
          if constexpr (!BL::kTMA) { if (pass < b_passes) bl.commit(sb, rb, pass, b_tile, tg); }
        }
        fence_proxy_async_smem();
        mbar_arrive(&full_bar[s]);
      }
    }
    // ---------------- epilogue ----------------
    if (nchunks > 0) {
      mbar_wait(&tmem_full_bar, 0);
      tc_fence_after_sync();
    }
    // warp w reads TMEM lanes 32*(w%4)..+31 (= tile rows); the four warps sharing a lane quarter take
    // every fourth 16-column group
    const int m = m0 + 32 * (warp & 3) + lane;
    typename EP::Ctx ctx;
    if constexpr (!SPLIT) ctx = ep.row(m);
    for (int c0 = 16 * (warp >> 2); c0 < n_mma; c0 += 16 * (NW / 4)) {
      uint32_t v[16];
      if (nchunks > 0) {
        tmem_ld_32x32b_x16(tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)c0, v);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = 0u;
      }
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int n = n0 + c0 + j;
        const float acc = __uint_as_float(v[j]);
        if constexpr (SPLIT) {
          if (m < M && n < N) ws[((size_t)blockIdx.z * M + m) * N + n] = acc;
        } else {
          ep.store(ctx, n, acc);
        }
      }
    }
  } else if (warp == kMmaWarp) {
    // ---------------- MMA issuer ----------------
    if (lane == 0 && nchunks > 0) {
      const uint32_t idesc = make_idesc_tf32(TBM, n_mma, AL::kMN ? 1 : 0, BL::kMN ? 1 : 0);
      for (int kc = 0; kc < nchunks; ++kc) {
        const int s = kc % TSTAGES;
        mbar_wait(&full_bar[s], (kc / TSTAGES) & 1);
        tc_fence_after_sync();
        const uint32_t a_addr = smem_u32(smem + s * Cfg::kStageBytes);
        const uint32_t b_addr = a_addr + kATileBytes;
#pragma unroll
        for (int j = 0; j < TBK / 8; ++j) {
          const uint64_t adesc = AL::kMN ? make_smem_desc(a_addr + j * 1024, 4096, 512, kLayoutSW128Base32)
                                         : make_smem_desc(a_addr + j * 32, 16, 1024, kLayoutSW128);
          const uint64_t bdesc = BL::kMN ? make_smem_desc(b_addr + j * 1024, 4096, 512, kLayoutSW128Base32)
                                         : make_smem_desc(b_addr + j * 32, 16, 1024, kLayoutSW128);
          mma_tf32(tmem_base, adesc, bdesc, idesc, (kc | j) != 0 ? 1u : 0u);
        }
        mma_commit(&empty_bar[s]);          // smem stage free once these MMAs have read it
      }
      mma_commit(&tmem_full_bar);           // accumulator complete
    }
  } else {
    // ---------------- TMA issuer ----------------
    if (kAnyTma && lane == 0 && nchunks > 0) {
      uint32_t bytes = 0;
      if constexpr (AL::kTMA) bytes += al.bytes();
      if constexpr (BL::kTMA) bytes += bl.bytes();
      for (int kc = 0; kc < nchunks; ++kc) {
        const int s = kc % TSTAGES;
        mbar_wait(&empty_bar[s], ((kc / TSTAGES) & 1) ^ 1);
        uint8_t* a_tile = smem + s * Cfg::kStageBytes;
        uint8_t* b_tile = a_tile + kATileBytes;
        const int r0 = r_begin + kc * TBK;
        mbar_arrive_expect_tx(&full_bar[s], bytes);
        if constexpr (AL::kTMA) al.issue(&tmA, a_tile, &full_bar[s], m0, r0);
        if constexpr (BL::kTMA) bl.issue(&tmB, b_tile, &full_bar[s], n0, r0);
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == kMmaWarp) tmem_dealloc<Cfg::kTmemCols>(tmem_base);
}

// deterministic slice reduction (ascending z) + the per-element epilogue of gemm_ops.cuh
template <class E>
__global__ void tc_splitk_reduce_kernel(const float* __restrict__ ws, int splits, int M, int N, E epi) {
  SN_GRID_DEP_LAUNCH_ALWAYS();
  SN_GRID_DEP_WAIT();
  const size_t total = (size_t)M * N;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    float acc = ws[i];
    for (int z = 1; z < splits; ++z) acc += ws[(size_t)z * total + i];
    epi((int)(i / N), (int)(i % N), acc);
  }
}
// Many slices over a small output (first-layer backward-filter: 200 slices x 2400 elements): one thread per element
// would walk `splits` dependent loads.  Here 8 lanes share an element: lane j sums slices j, j+8, ... in ascending
// order, then the eight partial sums are combined in a fixed butterfly — still the same order on every run.
template <class E>
__global__ void tc_splitk_reduce_wide_kernel(const float* __restrict__ ws, int splits, int M, int N, E epi) {
  SN_GRID_DEP_LAUNCH_ALWAYS();
  SN_GRID_DEP_WAIT();
  const size_t total = (size_t)M * N;
  const size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const size_t i = t >> 3;
  const int j = (int)(t & 7);
  float acc = 0.f;
  if (i < total) {
    float a0 = 0.f, a1 = 0.f;
    int z = j;
    for (; z + 8 < splits; z += 16) { a0 += ws[(size_t)z * total + i]; a1 += ws[(size_t)(z + 8) * total + i]; }
    if (z < splits) a0 += ws[(size_t)z * total + i];
    acc = a0 + a1;
  }
  acc += __shfl_xor_sync(0xffffffffu, acc, 1);
  acc += __shfl_xor_sync(0xffffffffu, acc, 2);
  acc += __shfl_xor_sync(0xffffffffu, acc, 4);
  if (i < total && j == 0) epi((int)(i / N), (int)(i % N), acc);
}
template <class E>
static void launch_splitk_reduce(cudaStream_t s, const float* ws, int splits, int M, int N, E epi) {
  const size_t total = (size_t)M * N;
  if (splits >= 16 && total * 8 <= (size_t)148 * 8 * 256) {
    tc_splitk_reduce_wide_kernel<<<(int)((total * 8 + 255) / 256), 256, 0, s>>>(ws, splits, M, N, epi);
    return;
  }
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  tc_splitk_reduce_kernel<<<blocks, 256, 0, s>>>(ws, splits, M, N, epi);
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    SN_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres));
    if (!p || qres != cudaDriverEntryPointSuccess) fatal("cuTensorMapEncodeTiled is not available from this driver");
    fn = (EncodeTiledFn)p;
  });
  return fn;
}
// 2-D fp32 tensor {inner (contiguous), outer} with pitch (bytes, multiple of 16); box {box_inner = 32, box_outer}
static CUtensorMap make_tmap(const float* base, uint64_t inner, uint64_t outer, uint64_t pitch_bytes, uint32_t box_outer,
                             bool mn_major = false) {
  CUtensorMap m;
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {pitch_bytes};
  cuuint32_t box[2] = {32, box_outer};
  cuuint32_t estr[2] = {1, 1};
  if (((uintptr_t)base & 15) || (pitch_bytes & 15)) fatal("TMA tensor map needs 16-byte aligned base and pitch");
  CUresult r = encode_fn()(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)base, dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE,
                           mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) fatal("cuTensorMapEncodeTiled failed with %d (inner %llu outer %llu pitch %llu box %u)", (int)r,
                               (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)pitch_bytes, box_outer);
  return m;
}
// Encoded descriptors are cached: buffers and shapes repeat every iteration, cuTensorMapEncodeTiled is ~1 us of host time.
CUtensorMap tc_make_tmap(const float* base, uint64_t inner, uint64_t outer, uint64_t pitch_bytes, uint32_t box_outer, bool mn_major) {
  struct Key {
    const float* base; uint64_t inner, outer, pitch; uint32_t box; bool mn;
    bool operator<(const Key& o) const { return std::tie(base, inner, outer, pitch, box, mn) < std::tie(o.base, o.inner, o.outer, o.pitch, o.box, o.mn); }
  };
  static std::map<Key, CUtensorMap> cache;
  static std::mutex mu;
  std::lock_guard<std::mutex> lock(mu);
  const Key k{base, inner, outer, pitch_bytes, box_outer, mn_major};
  auto it = cache.find(k);
  if (it != cache.end()) return it->second;
  if (cache.size() > 4096) cache.clear();
  return cache[k] = make_tmap(base, inner, outer, pitch_bytes, box_outer, mn_major);
}
static CUtensorMap dummy_tmap() {
  CUtensorMap m;
  memset(&m, 0, sizeof(m));
  return m;
}

static int pick_bn(int N) { return N <= 64 ? 64 : (N <= 128 ? 128 : 256); }
// N tile width: split N evenly into ceil(N/256) tiles, rounded up to the MMA granularity of 16
static int pick_n_mma(int N) {
  const int tiles = (N + 255) / 256;
  int w = (N + tiles - 1) / tiles;
  w = (w + 15) / 16 * 16;
  return w < 16 ? 16 : w;
}

template <class AL, class BL, class EP, class RE, int BN>
static void launch_bn(cudaStream_t s, const CUtensorMap& tmA, const CUtensorMap& tmB, AL al, BL bl, EP ep, RE reduce_epi,
                      int M, int N, int R, int n_mma) {
  typedef TcCfg<BN> Cfg;
  dim3 grid((M + TBM - 1) / TBM, (N + n_mma - 1) / n_mma, 1);
  // one CTA of this kernel per SM: 148 slices are one wave (cifar conv1 backward-filter, 102400 reduction rows: 200 slices ran
  // as 148 + 52, 72.7 -> 58.4 us with 148)
  static const int split_waves = getenv("SPARKNET_TC_SPLIT_WAVES") ? atoi(getenv("SPARKNET_TC_SPLIT_WAVES")) : 1;
  const int splits = splitk_slices((int)(grid.x * grid.y), R, split_waves);
  // largest dynamic shared-memory size opted in so far, per device (cudaFuncSetAttribute is per device) and under a lock:
  // sparknet_tc_gemm_static_scratch may be called on any device from any thread and shares these instantiations
  static size_t attr_split_dev[32] = {0}, attr_dev[32] = {0};
  static std::mutex attr_mu;
  int dev = 0;
  SN_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 32) fatal("tensor-core contraction: device index %d out of range", dev);
  constexpr size_t kMaxSmem = 227 * 1024;
  if (splits == 1) {
    const int len = (R + TBK - 1) / TBK * TBK;
    const size_t smem = Cfg::kSmemBytes + (size_t)len * KTabEntryBytes<AL>::value;
    if (smem > kMaxSmem) fatal("tensor-core contraction: reduction of %d needs %zu bytes of shared memory (> %zu)", R, smem, kMaxSmem);
    auto kfn = gemm_tc_kernel<AL, BL, EP, BN, false>;
    {
      std::lock_guard<std::mutex> lock(attr_mu);
      if (smem > attr_dev[dev]) { SN_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); attr_dev[dev] = smem; }
    }
    kfn<<<grid, kTcThreads, smem, s>>>(tmA, tmB, al, bl, ep, M, N, R, len, n_mma, nullptr);
    count_launch();
  } else {
    int len = (R + splits - 1) / splits;
    len = (len + TBK - 1) / TBK * TBK;
    grid.z = (R + len - 1) / len;
    float* ws = get_workspace((size_t)grid.z * M * N);
    const size_t smem = Cfg::kSmemBytes + (size_t)len * KTabEntryBytes<AL>::value;
    if (smem > kMaxSmem) fatal("tensor-core contraction: reduction slice of %d needs %zu bytes of shared memory (> %zu)", len, smem, kMaxSmem);
    auto kfn = gemm_tc_kernel<AL, BL, EP, BN, true>;
    {
      std::lock_guard<std::mutex> lock(attr_mu);
      if (smem > attr_split_dev[dev]) { SN_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); attr_split_dev[dev] = smem; }
    }
    kfn<<<grid, kTcThreads, smem, s>>>(tmA, tmB, al, bl, ep, M, N, R, len, n_mma, ws);
    launch_splitk_reduce(s, ws, (int)grid.z, M, N, reduce_epi);
    count_launch(2);
  }
  SN_CUDA(cudaGetLastError());
}
// rows_setter(n_mma) lets the caller size loaders / TMA boxes for the chosen tile width
template <class AL, class BL, class EP, class RE>
static void launch_tc(cudaStream_t s, const CUtensorMap& tmA, const CUtensorMap& tmB, AL al, BL bl, EP ep, RE re, int M, int N,
                      int R, int n_mma) {
  switch (pick_bn(n_mma)) {
    case 64: launch_bn<AL, BL, EP, RE, 64>(s, tmA, tmB, al, bl, ep, re, M, N, R, n_mma); break;
    case 128: launch_bn<AL, BL, EP, RE, 128>(s, tmA, tmB, al, bl, ep, re, M, N, R, n_mma); break;
    default: launch_bn<AL, BL, EP, RE, 256>(s, tmA, tmB, al, bl, ep, re, M, N, R, n_mma); break;
  }
}

bool tc_available() { return true; }

// Rounded (tf32, round-to-nearest) shadow copies of weight blobs for the TMA-fed operands.  Keyed by
// the fp32 pointer; refreshed by tc_refresh_shadow whenever the runtime knows the weights changed.
__global__ void round_tf32_kernel(size_t n, const float* __restrict__ src, float* __restrict__ dst) {
  SN_GRID_DEP_LAUNCH_ALWAYS();
  SN_GRID_DEP_WAIT();
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = to_tf32_exact(src[i]);
}
void tc_round_to_tf32(cudaStream_t s, size_t n, const float* src, float* dst) {
  int blocks = (int)((n + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  round_tf32_kernel<<<blocks, 256, 0, s>>>(n, src, dst);
  count_launch();
}

// ---- convolution -------------------------------------------------------------------------------
void tc_conv_fwd(cudaStream_t s, const ConvGeom& g, const float* x, const float* w_tf32, const float* bias, float* y, bool relu) {
  const int P = g.N * g.Ho * g.Wo, K = g.Kdim(), Og = g.Og();
  const int n_mma = pick_n_mma(Og);
  for (int grp = 0; grp < g.group; ++grp) {
    TEpiConvFwd ep{y, bias, P, g.Ho * g.Wo, g.Cout, Og, grp, relu ? 1 : 0, FastDiv(g.Ho * g.Wo)};
    EpiConvFwd re{y, bias, g, grp, relu ? 1 : 0, FastDiv(g.Ho * g.Wo)};
    const float* wg = w_tf32 + (size_t)grp * Og * K;
    auto go = [&](auto al) {
      if (K % 4 == 0 && ((uintptr_t)wg & 15) == 0) {
        CUtensorMap tmB = make_tmap(wg, (uint64_t)K, (uint64_t)Og, (uint64_t)K * 4, (uint32_t)n_mma);
        launch_tc(s, dummy_tmap(), tmB, al, LoadTmaK{n_mma}, ep, re, P, Og, K, n_mma);
      } else {
        LoadDenseK bl{wg, K, Og, K, n_mma, 0};
        launch_tc(s, dummy_tmap(), dummy_tmap(), al, bl, ep, re, P, Og, K, n_mma);
      }
    };
    if (g.kh * g.kw <= 32) go(LoadIm2colRowsMN<true>{x, g, grp, FastDiv(g.Ho * g.Wo), FastDiv(g.Wo), FastDiv(g.kh * g.kw), FastDiv(g.kw)});
    else go(LoadIm2colRowsMN<false>{x, g, grp, FastDiv(g.Ho * g.Wo), FastDiv(g.Wo), FastDiv(g.kh * g.kw), FastDiv(g.kw)});
  }
}
void tc_conv_dgrad(cudaStream_t s, const ConvGeom& g, const float* dy, const float* w, float* dx) {
  const int P = g.N * g.H * g.W, R = g.Og() * g.kh * g.kw, Cg = g.Cg();
  const int n_mma = pick_n_mma(Cg);
  for (int grp = 0; grp < g.group; ++grp) {
    TEpiConvDgrad ep{dx, P, g.H * g.W, g.C, Cg, grp, FastDiv(g.H * g.W)};
    EpiConvDgrad re{dx, g, grp, FastDiv(g.H * g.W)};
    LoadFilterDgradK bl{w, g, grp, n_mma, FastDiv(g.kh * g.kw)};
    if (g.sh == 1 && g.sw == 1 && g.kh * g.kw <= 32) {
      LoadDgradRowsMN<true> al{dy, g, grp, FastDiv(g.H * g.W), FastDiv(g.W), FastDiv(g.kh * g.kw), FastDiv(g.kw)};
      launch_tc(s, dummy_tmap(), dummy_tmap(), al, bl, ep, re, P, Cg, R, n_mma);
    } else if (g.sh == 1 && g.sw == 1) {
      LoadDgradRowsMN<false> al{dy, g, grp, FastDiv(g.H * g.W), FastDiv(g.W), FastDiv(g.kh * g.kw), FastDiv(g.kw)};
      launch_tc(s, dummy_tmap(), dummy_tmap(), al, bl, ep, re, P, Cg, R, n_mma);
    } else {
      LoadAccMN<DgradRows> al{make_dgrad_rows(dy, g, grp), TBM};
      launch_tc(s, dummy_tmap(), dummy_tmap(), al, bl, ep, re, P, Cg, R, n_mma);
    }
  }
}
void tc_conv_wgrad(cudaStream_t s, const ConvGeom& g, const float* x, const float* dy, float* dw, bool accumulate) {
  const int R = g.N * g.Ho * g.Wo, K = g.Kdim(), Og = g.Og();
  const int n_mma = pick_n_mma(K);
  for (int grp = 0; grp < g.group; ++grp) {
    float* dwg = dw + (size_t)grp * Og * K;
    LoadDyRowsK al{dy, g, grp, FastDiv(g.Ho * g.Wo)};
    TEpiRowMajor ep{dwg, K, Og, K, accumulate ? 1 : 0};
    EpiRowMajor re{dwg, K, Og, K, accumulate ? 1 : 0};
    if (g.kh * g.kw <= 32) {
      LoadIm2colColsK<true> bl{x, g, grp, n_mma, FastDiv(g.Ho * g.Wo), FastDiv(g.Wo), FastDiv(g.kh * g.kw), FastDiv(g.kw)};
      launch_tc(s, dummy_tmap(), dummy_tmap(), al, bl, ep, re, Og, K, R, n_mma);
    } else {
      LoadIm2colColsK<false> bl{x, g, grp, n_mma, FastDiv(g.Ho * g.Wo), FastDiv(g.Wo), FastDiv(g.kh * g.kw), FastDiv(g.kw)};
      launch_tc(s, dummy_tmap(), dummy_tmap(), al, bl, ep, re, Og, K, R, n_mma);
    }
  }
}

// ---- flat-grid convolution (stride 1): every operand through TMA ------------------------------------
// The gathers above spend ~8 instructions per staged element.  For stride-1 filters the im2col view can
// instead be expressed as plain TMA boxes if activations are first repacked channels-last on a "flat grid":
//   flat[lead + n*Gp + h*Wp + w][c] = tf32(t[n][c][h][w]),  Hp = H + 2*ph = Ho + kh - 1,  Wp likewise, Gp = Hp*Wp,
// images top-left aligned in their Hp x Wp cell, everything else zero (zeroed once at allocation; the pack
// kernel only writes image positions).  With output positions on the SAME grid (q = n*Gp + ho*Wp + wo),
//   y(q)  = sum_{c,ky,kx} xflat[q + ky*Wp + kx - (ph*Wp + pw)][c] * W[co][c][ky][kx]
//   dx(q) = sum_{co,ky',kx'} dyflat[q + ky'*Wp + kx' - (pad'*Wp + pad')][co] * W[co][c][kh-1-ky'][kw-1-kx'],  pad' = k-1-p
//   dW[co][c][ky][kx] = sum_q dyflat[q][co] * xflat[q + ky*Wp + kx - (ph*Wp + pw)][c]
// so a filter tap is a pure shift of the grid row: the zero bands to the right of / below each image double
// as the left / top padding of the next row / image.  The shift is on the OUTER TMA coordinate (the inner
// one must stay 16-byte aligned), which is why the packed layout is channels-last.  Grid positions that are
// not real outputs (w >= Wo or h >= Ho) are computed and dropped by the epilogue: (Hp*Wp)/(Ho*Wo) = 1.3x
// extra MMA work for AlexNet's 3x3 / 5x5 layers, in exchange for no gather.  The reduction of fwd / dgrad
// runs over (tap, channel slot) with the group's channels padded to a multiple of 32 slots; the filters are
// repacked to match with zeros in the extra slots (which read the neighbouring group's channels or TMA
// zero fill).

static int flat_slots(int ch) { return (ch + 31) / 32 * 32; }

bool tc_flat_supported(const ConvGeom& g) {
  const int Cg = g.Cg(), Og = g.Og();
  return g.sh == 1 && g.sw == 1 && Cg % 16 == 0 && Og % 16 == 0 && Cg <= 256 && g.kh * g.kw <= 64 && g.ph < g.kh &&
         g.pw < g.kw && (long long)g.N * (g.H + g.kh) * (g.W + g.kw) < (1ll << 30);
}
// Cell geometry.  The zero band to the right of / below each image is shared: it is the right / bottom
// padding of this image row and the left / top padding of the next one, so it only has to be as wide as the
// larger of the padding the forward taps need (p) and the padding the backward-data taps need (k - 1 - p).
FlatGeom tc_flat_geom(const ConvGeom& g) {
  FlatGeom f;
  f.bw = g.pw > g.kw - 1 - g.pw ? g.pw : g.kw - 1 - g.pw;
  f.bh = g.ph > g.kh - 1 - g.ph ? g.ph : g.kh - 1 - g.ph;
  f.Wp = g.W + f.bw;
  f.Hp = g.H + f.bh;
  f.Gp = f.Wp * f.Hp;
  f.lead = (f.bh * f.Wp + f.bw + 31) / 32 * 32;        // zero rows in front of cell 0: no TMA coordinate is negative
  return f;
}
static int flat_lead(const ConvGeom& g) { return tc_flat_geom(g).lead; }
size_t tc_flat_rows(const ConvGeom& g) {
  const FlatGeom f = tc_flat_geom(g);
  const size_t cells = (size_t)g.N * f.Gp;
  return (cells + 127) / 128 * 128 + 128 + (size_t)f.lead;
}
size_t tc_flat_filter_floats(const ConvGeom& g, bool dgrad) {
  return dgrad ? (size_t)g.C * flat_slots(g.Og()) * g.kh * g.kw : (size_t)g.Cout * flat_slots(g.Cg()) * g.kh * g.kw;
}

// NCHW -> grid rows x channels (32 x 32 transposes through shared memory).  One block owns 32 channels of one image
// and walks its pixels; with BIAS it also sums every value it reads per channel (the bias gradient of this image,
// conv_layer.cpp:52-55: gemv against the all-ones multiplier) into part[c][n], which bias_grad_final_kernel adds up
// in image order.  The top gradient is therefore read from HBM once for both purposes.
template <bool BIAS>
__global__ void __launch_bounds__(256) flat_pack_kernel(const float* __restrict__ src, float* __restrict__ dst, float* __restrict__ part,
                                                         int N, int C, int H, int W, int Wp, int Gp, int lead, FastDiv dW, int span, int Csrc) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  // Csrc <= C: the source tensor has Csrc channels, the packed rows C (the rest are written as zeros: channel padding of
  // first layers with fewer than 16 input channels)
  __shared__ float tile[32][129];
  const int HW = H * W, c0 = blockIdx.x * 32, n = blockIdx.y;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const float* sp = src + ((size_t)n * Csrc + c0) * HW;
  float* dp = dst + ((size_t)lead + (size_t)n * Gp) * C + c0 + tx;
  float bsum[4] = {0.f, 0.f, 0.f, 0.f};
  const int p_begin = blockIdx.z * span, p_end = min(HW, p_begin + span);     // this block's share of the image's pixels
  for (int p0 = p_begin; p0 < p_end; p0 += 128) {
    float v[4][4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int j = ty + 8 * k, p = p0 + tx + 32 * q;
        v[k][q] = (c0 + j < Csrc && p < p_end) ? sp[(size_t)j * HW + p] : 0.f;
      }
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        tile[ty + 8 * k][tx + 32 * q] = v[k][q];
        if (BIAS) bsum[k] += v[k][q];
      }
    __syncthreads();
    if (c0 + tx < C) {
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        const int pl = ty + 8 * k, q = p0 + pl;
        if (q < p_end) {
          uint32_t h, w;
          dW.divmod((uint32_t)q, h, w);
          dp[(size_t)(h * Wp + w) * C] = to_tf32_exact(tile[tx][pl]);
        }
      }
    }
    __syncthreads();
  }
  if (BIAS) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float v = bsum[k];
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (tx == 0 && c0 + ty + 8 * k < C) part[(size_t)(c0 + ty + 8 * k) * (N * gridDim.z) + n * gridDim.z + blockIdx.z] = v;
    }
  }
}
// one warp per channel: lane l adds partials l, l + 32, ... in ascending order, then a fixed butterfly: deterministic
__global__ void bias_grad_final_kernel(int C, int S, const float* __restrict__ part, float* __restrict__ db, int accumulate) {
  SN_GRID_DEP_LAUNCH_ALWAYS();
  SN_GRID_DEP_WAIT();
  const int c = (int)((blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (c >= C) return;
  float t = 0.f;
  for (int s = lane; s < S; s += 32) t += part[(size_t)c * S + s];
  for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  if (lane == 0) db[c] = accumulate ? db[c] + t : t;
}
// db (nullable): bias gradient, summed here when the packed tensor is the top gradient
void tc_flat_pack(cudaStream_t s, const ConvGeom& g, bool output_side, const float* src, float* dst, float* db, bool accumulate, int src_channels) {
  const int C = output_side ? g.Cout : g.C, H = output_side ? g.Ho : g.H, W = output_side ? g.Wo : g.W;
  const int Csrc = src_channels > 0 ? src_channels : C;
  const FlatGeom f = tc_flat_geom(g);
  const int HW = H * W;
  int Z = (HW + 1023) / 1024;                          // large planes: several blocks per (image, channel block)
  if (Z > 4) Z = 4;
  const int span = ((HW + Z - 1) / Z + 127) / 128 * 128;
  Z = (HW + span - 1) / span;
  dim3 grid((C + 31) / 32, g.N, Z);
  if (db) {
    float* part = get_workspace((size_t)C * g.N * Z);
    flat_pack_kernel<true><<<grid, 256, 0, s>>>(src, dst, part, g.N, C, H, W, f.Wp, f.Gp, f.lead, FastDiv(W), span, Csrc);
    bias_grad_final_kernel<<<(C * 32 + 127) / 128, 128, 0, s>>>(C, g.N * Z, part, db, accumulate ? 1 : 0);
    count_launch(2);
  } else {
    flat_pack_kernel<false><<<grid, 256, 0, s>>>(src, dst, nullptr, g.N, C, H, W, f.Wp, f.Gp, f.lead, FastDiv(W), span, Csrc);
    count_launch();
  }
  SN_CUDA(cudaGetLastError());
}
// dst[row][tap*slots + ch]: fwd   (dgrad = 0): row = co (all groups), ch = c in group,  value W[co][c][tap]
//                           dgrad (dgrad = 1): row = grp*Cg + c,      ch = co in group, value W[grp*Og + co][c][T-1-tap]
__global__ void flat_pack_filter_kernel(const float* __restrict__ w, float* __restrict__ dst, int Cg, int Og, int T, int groups, int dgrad, int Csrc) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  const int red = dgrad ? Og : Cg, rows = dgrad ? Cg * groups : Og * groups;
  const int slots = (red + 31) / 32 * 32, kp = T * slots;
  const size_t total = (size_t)rows * kp;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int row = (int)(i / kp), k = (int)(i - (size_t)row * kp), tap = k / slots, ch = k - tap * slots;
    float v = 0.f;
    if (ch < red) {
      if (!dgrad) v = ch < Csrc ? w[((size_t)row * Csrc + ch) * T + tap] : 0.f;   // Csrc < Cg: zero slots for padded channels
      else { const int grp = row / Cg, c = row - grp * Cg; v = w[((size_t)(grp * Og + ch) * Cg + c) * T + (T - 1 - tap)]; }
    }
    dst[i] = to_tf32_exact(v);
  }
}
void tc_flat_pack_filter(cudaStream_t s, const ConvGeom& g, bool dgrad, const float* w, float* dst, int src_cg) {
  const size_t total = tc_flat_filter_floats(g, dgrad);
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  flat_pack_filter_kernel<<<blocks, 256, 0, s>>>(w, dst, g.Cg(), g.Og(), g.kh * g.kw, g.group, dgrad ? 1 : 0, (src_cg > 0 && !dgrad) ? src_cg : g.Cg());
  count_launch();
  SN_CUDA(cudaGetLastError());
}

// A operand of flat fwd / dgrad: K-major, tile row = grid position, chunk = one tap x 32 channel slots.
// tensor {channels (contiguous), grid rows}; box {32, 128}
struct LoadTmaFlatRowsK {
  static constexpr bool kTMA = true, kMN = false, kKTable = false;
  typedef NoState State;
  typedef NoRegs Regs;
  int Wp, kw, off, nblk, chan0;          // off = row offset of tap (0,0), lead folded in
  __device__ uint32_t bytes() const { return (uint32_t)kATileBytes; }
  __device__ void issue(const CUtensorMap* tm, uint8_t* tile, uint64_t* bar, int row0, int r0) const {
    const int kc = r0 >> 5, tap = kc / nblk, cb = kc - tap * nblk, ky = tap / kw, kx = tap - ky * kw;
    tma_load_2d(tile, tm, bar, chan0 + cb * 32, row0 + ky * Wp + kx + off);
  }
};
// A operand of flat wgrad: MN-major, tile row = output channel, reduction = grid position.  box {32, 32}
struct LoadTmaFlatDyMN {
  static constexpr bool kTMA = true, kMN = true, kKTable = false;
  typedef NoState State;
  typedef NoRegs Regs;
  int chan0, lead;
  __device__ uint32_t bytes() const { return (uint32_t)kATileBytes; }
  __device__ void issue(const CUtensorMap* tm, uint8_t* tile, uint64_t* bar, int row0, int r0) const {
    for (int i = 0; i < TBM / 32; ++i) tma_load_2d(tile + i * 4096, tm, bar, chan0 + row0 + 32 * i, lead + r0);
  }
};
// B operand of flat wgrad: MN-major, tile row = (tap, channel slot), reduction = grid position.  box {32, 32}
struct LoadTmaFlatTapsMN {
  static constexpr bool kTMA = true, kMN = true, kKTable = false;
  typedef NoState State;
  typedef NoRegs Regs;
  int Wp, kw, T, off, slots, nt, chan0;
  __device__ uint32_t bytes() const { return (uint32_t)(nt * slots) * 128u; }
  __device__ void issue(const CUtensorMap* tm, uint8_t* tile, uint64_t* bar, int row0, int r0) const {
    const int t0 = row0 / slots, nblk = slots >> 5;
    for (int j = 0; j < nt; ++j) {
      const int tap = min(t0 + j, T - 1), ky = tap / kw, kx = tap - ky * kw;
      for (int b = 0; b < nblk; ++b)
        tma_load_2d(tile + (j * nblk + b) * 4096, tm, bar, chan0 + 32 * b, r0 + ky * Wp + kx + off);
    }
  }
};
// m = grid position, n = channel in group; NCHW destination with Hv x Wv valid positions per cell
struct TEpiConvFlat {
  float* dst; const float* bias; int Ptot, Hv, Wv, Ctot, Cgrp, chan0, relu; FastDiv dGp, dWp;
  struct Ctx { float* p; bool ok; };
  __device__ Ctx row(int m) const {
    Ctx c; c.ok = m < Ptot; c.p = dst;
    if (c.ok) {
      uint32_t n, rem, h, w;
      dGp.divmod((uint32_t)m, n, rem);
      dWp.divmod(rem, h, w);
      c.ok = (int)h < Hv && (int)w < Wv;
      c.p = dst + (((size_t)n * Ctot + chan0) * Hv + h) * Wv + w;
    }
    return c;
  }
  __device__ void store(const Ctx& c, int n, float v) const {
    if (!c.ok || n >= Cgrp) return;
    if (bias) v += __ldg(bias + chan0 + n);
    if (relu) v = fmaxf(v, 0.f);
    c.p[(size_t)n * Hv * Wv] = v;
  }
  __device__ void operator()(int m, int n, float v) const { store(row(m), n, v); }
};
// m = co in group, n = tap*slots + c;  dw[co][c][tap] (+)= v
struct TEpiWgradFlat {
  float* dw; int Og, Cg, T, accumulate; FastDiv dSlots;
  struct Ctx { float* q; bool ok; };
  __device__ Ctx row(int m) const { Ctx c; c.ok = m < Og; c.q = dw + (size_t)m * Cg * T; return c; }
  __device__ void store(const Ctx& c, int n, float v) const {
    uint32_t tap, ch;
    dSlots.divmod((uint32_t)n, tap, ch);
    if (!c.ok || (int)tap >= T || (int)ch >= Cg) return;
    float* p = c.q + ch * T + tap;
    *p = accumulate ? *p + v : v;
  }
  __device__ void operator()(int m, int n, float v) const { store(row(m), n, v); }
};

// `xflat`: packed input (tc_flat_pack, input side); `wf`: tc_flat_pack_filter(dgrad = false)
void tc_conv_fwd_flat(cudaStream_t s, const ConvGeom& g, const float* xflat, const float* wf, const float* bias, float* y, bool relu) {
  const int Cg = g.Cg(), Og = g.Og(), T = g.kh * g.kw, Wp = tc_flat_geom(g).Wp, Gp = tc_flat_geom(g).Gp, Ptot = g.N * Gp;
  const int slots = flat_slots(Cg), kp = T * slots, n_mma = pick_n_mma(Og), lead = flat_lead(g);
  CUtensorMap tmA = make_tmap(xflat, (uint64_t)g.C, tc_flat_rows(g), (uint64_t)g.C * 4, TBM);
  for (int grp = 0; grp < g.group; ++grp) {
    CUtensorMap tmB = make_tmap(wf + (size_t)grp * Og * kp, (uint64_t)kp, (uint64_t)Og, (uint64_t)kp * 4, (uint32_t)n_mma);
    LoadTmaFlatRowsK al{Wp, g.kw, lead - (g.ph * Wp + g.pw), slots / 32, grp * Cg};
    TEpiConvFlat ep{y, bias, Ptot, g.Ho, g.Wo, g.Cout, Og, grp * Og, relu ? 1 : 0, FastDiv(Gp), FastDiv(Wp)};
    launch_tc(s, tmA, tmB, al, LoadTmaK{n_mma}, ep, ep, Ptot, Og, kp, n_mma);
  }
}
// `dyflat`: packed top gradient (tc_flat_pack, output side); `wt`: tc_flat_pack_filter(dgrad = true)
void tc_conv_dgrad_flat(cudaStream_t s, const ConvGeom& g, const float* dyflat, const float* wt, float* dx) {
  const int Cg = g.Cg(), Og = g.Og(), T = g.kh * g.kw, Wp = tc_flat_geom(g).Wp, Gp = tc_flat_geom(g).Gp, Ptot = g.N * Gp;
  const int slots = flat_slots(Og), kp = T * slots, n_mma = pick_n_mma(Cg), lead = flat_lead(g);
  CUtensorMap tmA = make_tmap(dyflat, (uint64_t)g.Cout, tc_flat_rows(g), (uint64_t)g.Cout * 4, TBM);
  for (int grp = 0; grp < g.group; ++grp) {
    CUtensorMap tmB = make_tmap(wt + (size_t)grp * Cg * kp, (uint64_t)kp, (uint64_t)Cg, (uint64_t)kp * 4, (uint32_t)n_mma);
    LoadTmaFlatRowsK al{Wp, g.kw, lead - ((g.kh - 1 - g.ph) * Wp + (g.kw - 1 - g.pw)), slots / 32, grp * Og};
    TEpiConvFlat ep{dx, nullptr, Ptot, g.H, g.W, g.C, Cg, grp * Cg, 0, FastDiv(Gp), FastDiv(Wp)};
    launch_tc(s, tmA, tmB, al, LoadTmaK{n_mma}, ep, ep, Ptot, Cg, kp, n_mma);
  }
}
void tc_conv_wgrad_flat(cudaStream_t s, const ConvGeom& g, const float* xflat, const float* dyflat, float* dw, bool accumulate) {
  const int Cg = g.Cg(), Og = g.Og(), T = g.kh * g.kw, Wp = tc_flat_geom(g).Wp, Gp = tc_flat_geom(g).Gp, Ptot = g.N * Gp;
  const int slots = flat_slots(Cg), nt = 256 / slots < T ? 256 / slots : T, n_mma = nt * slots, lead = flat_lead(g);
  CUtensorMap tmA = make_tmap(dyflat, (uint64_t)g.Cout, tc_flat_rows(g), (uint64_t)g.Cout * 4, 32, true);
  CUtensorMap tmB = make_tmap(xflat, (uint64_t)g.C, tc_flat_rows(g), (uint64_t)g.C * 4, 32, true);
  for (int grp = 0; grp < g.group; ++grp) {
    LoadTmaFlatDyMN al{grp * Og, lead};
    LoadTmaFlatTapsMN bl{Wp, g.kw, T, lead - (g.ph * Wp + g.pw), slots, nt, grp * Cg};
    TEpiWgradFlat ep{dw + (size_t)grp * Og * Cg * T, Og, Cg, T, accumulate ? 1 : 0, FastDiv(slots)};
    launch_tc(s, tmA, tmB, al, bl, ep, ep, Og, (T + nt - 1) / nt * n_mma, Ptot, n_mma);
  }
}

// ---- inner product -------------------------------------------------------------------------------
// D[m = out][n = batch] = sum_k W[m][k] X[n][k];  Y[n][m] = D + bias[m]
void tc_ip_fwd(cudaStream_t s, int batch, int out, int K, const float* x, const float* w_tf32, const float* bias, float* y, bool relu) {
  const int n_mma = pick_n_mma(batch);
  LoadDenseK bl{x, K, batch, K, n_mma, (K % 4 == 0 && ((uintptr_t)x & 15) == 0) ? 1 : 0};
  TEpiTransposed ep{y, out, out, batch, bias, relu ? 1 : 0};
  EpiTransposed re{y, out, out, batch, bias, relu ? 1 : 0};
  if (K % 4 == 0 && ((uintptr_t)w_tf32 & 15) == 0) {
    CUtensorMap tmA = make_tmap(w_tf32, (uint64_t)K, (uint64_t)out, (uint64_t)K * 4, TBM);
    launch_tc(s, tmA, dummy_tmap(), LoadTmaK{TBM}, bl, ep, re, out, batch, K, n_mma);
  } else {
    LoadDenseK al{w_tf32, K, out, K, TBM, 0};
    launch_tc(s, dummy_tmap(), dummy_tmap(), al, bl, ep, re, out, batch, K, n_mma);
  }
}
// D[m = k_in][n = batch] = sum_o W[o][m] dY[n][o];  dX[n][m] = D
void tc_ip_dgrad(cudaStream_t s, int batch, int out, int K, const float* dy, const float* w_tf32, float* dx) {
  const int n_mma = pick_n_mma(batch);
  LoadDenseK bl{dy, out, batch, out, n_mma, (out % 4 == 0 && ((uintptr_t)dy & 15) == 0) ? 1 : 0};
  TEpiTransposed ep{dx, K, K, batch, nullptr, 0};
  EpiTransposed re{dx, K, K, batch, nullptr, 0};
  if (K % 4 == 0 && ((uintptr_t)w_tf32 & 15) == 0) {
    CUtensorMap tmA = make_tmap(w_tf32, (uint64_t)K, (uint64_t)out, (uint64_t)K * 4, 32, true);
    launch_tc(s, tmA, dummy_tmap(), LoadTmaMN{TBM}, bl, ep, re, K, batch, out, n_mma);
  } else {
    LoadDenseMN al{w_tf32, K, K, out, TBM, 0};
    launch_tc(s, dummy_tmap(), dummy_tmap(), al, bl, ep, re, K, batch, out, n_mma);
  }
}
// D[m = out][n = k_in] = sum_b dY[b][m] X[b][n];  dW[m][n] (+)= D
void tc_ip_wgrad(cudaStream_t s, int batch, int out, int K, const float* x, const float* dy, float* dw, bool accumulate) {
  const int n_mma = pick_n_mma(K);
  LoadDenseMN al{dy, out, out, batch, TBM, (out % 4 == 0 && ((uintptr_t)dy & 15) == 0) ? 1 : 0};
  LoadDenseMN bl{x, K, K, batch, n_mma, (K % 4 == 0 && ((uintptr_t)x & 15) == 0) ? 1 : 0};
  TEpiRowMajor ep{dw, K, out, K, accumulate ? 1 : 0};
  EpiRowMajor re{dw, K, out, K, accumulate ? 1 : 0};
  launch_tc(s, dummy_tmap(), dummy_tmap(), al, bl, ep, re, out, K, batch, n_mma);
}

// Test hook (tools/debug_tc_gemm.py, tests): D[M][N] = A x B^T through the tensor-core kernel with
// plain dense operands.  a_mn / b_mn select the operand storage: 0 = [rows][R] (K-major), 1 = [R][rows]
// (MN-major).  tma_a: feed A through TMA instead of the gather warps.  All pointers are device pointers.
void tc_test_gemm(cudaStream_t s, int a_mn, int b_mn, int tma_a, int M, int N, int R, const float* A, const float* B, float* D) {
  const int n_mma = pick_n_mma(N);
  TEpiRowMajor ep{D, N, M, N, 0};
  EpiRowMajor re{D, N, M, N, 0};
  auto with_b = [&](auto al, const CUtensorMap& tmA) {
    if (b_mn) {
      LoadDenseMN bl{B, N, N, R, n_mma, (N % 4 == 0) ? 1 : 0};
      launch_tc(s, tmA, dummy_tmap(), al, bl, ep, re, M, N, R, n_mma);
    } else {
      LoadDenseK bl{B, R, N, R, n_mma, (R % 4 == 0) ? 1 : 0};
      launch_tc(s, tmA, dummy_tmap(), al, bl, ep, re, M, N, R, n_mma);
    }
  };
  if (tma_a) {
    if (a_mn) with_b(LoadTmaMN{TBM}, make_tmap(A, (uint64_t)M, (uint64_t)R, (uint64_t)M * 4, 32, true));
    else with_b(LoadTmaK{TBM}, make_tmap(A, (uint64_t)R, (uint64_t)M, (uint64_t)R * 4, TBM));
  } else {
    if (a_mn) with_b(LoadDenseMN{A, M, M, R, TBM, (M % 4 == 0) ? 1 : 0}, dummy_tmap());
    else with_b(LoadDenseK{A, R, M, R, TBM, (R % 4 == 0) ? 1 : 0}, dummy_tmap());
  }
}

// ------------------------------------------------------------------------------------------
// Split-reduction GEMM whose scratch lives in a static device array (no allocation at call time).
//
// Who needs this: callers that must not allocate device memory while they run — code between
// cudaStreamBeginCapture/EndCapture, real-time or latency-critical threads where cudaMalloc's implicit
// device synchronisation is unacceptable, and hosts that hand the library a fixed memory budget.  The
// regular path takes its split-K slices from a workspace that grows on demand (cudaMalloc/cudaFree);
// this variant keeps them in `g_static_scratch` below instead.
//
// Cost and limits: the array is part of the module image, so every GPU that loads libccaffe.so holds
// kStaticScratchFloats * 4 bytes (8 MiB) until the process exits, whether or not this entry point is
// ever called.  Problems whose slices need more than that are refused, never truncated.
// All calls share the one array and it keeps whatever the previous call left in it.  Nothing has to
// start at zero: each slice element the reduce kernel reads is written by the GEMM kernel of the same
// call first.  Two calls whose kernels can overlap on the GPU (different streams, no dependency) would
// write the same slices concurrently — that is a data race; serialise such calls.
// ------------------------------------------------------------------------------------------
constexpr size_t kStaticScratchFloats = 2u * 1024u * 1024u;       // 8 MiB
static __device__ __align__(256) float g_static_scratch[kStaticScratchFloats];

// Device address of g_static_scratch for the current device.  In host code the array's name is only a
// host-side stand-in for the device symbol, so it must never be passed to a kernel directly; the address
// is looked up once per device with that device current.
static float* static_scratch_ptr(int* err) {
  constexpr int kMaxDevices = 64;
  static float* ptrs[kMaxDevices] = {};
  static std::mutex mu;
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    fprintf(stderr, "libccaffe(b200): static scratch: cannot identify the current device (%s)\n", cudaGetErrorString(e));
    *err = -10;
    return nullptr;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || !p) {
      fprintf(stderr, "libccaffe(b200): cudaGetSymbolAddress(g_static_scratch) failed on device %d: %s\n", dev, cudaGetErrorString(e));
      *err = -11;
      return nullptr;
    }
    ptrs[dev] = (float*)p;
  }
  return ptrs[dev];
}

template <class AL, class BL, int BN>
static int launch_static_bn(cudaStream_t s, AL al, BL bl, int M, int N, int R, int n_mma, float* D, float* scratch) {
  typedef TcCfg<BN> Cfg;
  TEpiRowMajor ep{D, N, M, N, 0};
  EpiRowMajor re{D, N, M, N, 0};
  dim3 grid((M + TBM - 1) / TBM, (N + n_mma - 1) / n_mma, 1);
  const int splits = splitk_slices((int)(grid.x * grid.y), R, 1);   // as launch_bn: one CTA per SM, one wave
  cudaError_t e;
  if (splits == 1) {                                    // no scratch needed at all
    auto kfn = gemm_tc_kernel<AL, BL, TEpiRowMajor, BN, false>;
    e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
    if (e != cudaSuccess) return -20;
    kfn<<<grid, kTcThreads, Cfg::kSmemBytes, s>>>(dummy_tmap(), dummy_tmap(), al, bl, ep, M, N, R, (R + TBK - 1) / TBK * TBK, n_mma, nullptr);
  } else {
    int len = (R + splits - 1) / splits;
    len = (len + TBK - 1) / TBK * TBK;
    grid.z = (R + len - 1) / len;
    const size_t need = (size_t)grid.z * M * N;
    if (need > kStaticScratchFloats) {
      fprintf(stderr, "libccaffe(b200): sparknet_tc_gemm_static_scratch: %d x %d x %d needs %zu scratch floats in %u slices, "
                      "the static array holds %zu; refused\n", M, N, R, need, grid.z, kStaticScratchFloats);
      return -2;
    }
    auto kfn = gemm_tc_kernel<AL, BL, TEpiRowMajor, BN, true>;
    e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
    if (e != cudaSuccess) return -20;
    kfn<<<grid, kTcThreads, Cfg::kSmemBytes, s>>>(dummy_tmap(), dummy_tmap(), al, bl, ep, M, N, R, len, n_mma, scratch);
    launch_splitk_reduce(s, scratch, (int)grid.z, M, N, re);
  }
  e = cudaGetLastError();
  if (e != cudaSuccess) {
    fprintf(stderr, "libccaffe(b200): sparknet_tc_gemm_static_scratch: launch failed: %s\n", cudaGetErrorString(e));
    return -21;
  }
  return 0;
}
template <class AL, class BL>
static int launch_static(cudaStream_t s, AL al, BL bl, int M, int N, int R, int n_mma, float* D, float* scratch) {
  switch (pick_bn(n_mma)) {
    case 64: return launch_static_bn<AL, BL, 64>(s, al, bl, M, N, R, n_mma, D, scratch);
    case 128: return launch_static_bn<AL, BL, 128>(s, al, bl, M, N, R, n_mma, D, scratch);
    default: return launch_static_bn<AL, BL, 256>(s, al, bl, M, N, R, n_mma, D, scratch);
  }
}

}  // namespace sn

// D[M][N] = A x B^T in TF32 on the tensor cores, split along the reduction when that is needed to fill
// the GPU, with NO allocation of any kind: the slice scratch is the static array above.
//   a_mn / b_mn   storage of the operands: 0 = [rows][R] (reduction index contiguous), 1 = [R][rows]
//   A, B, D       DEVICE pointers supplied by the caller (D is written, A and B only read)
//   stream        cudaStream_t to launch on (NULL = the default stream); the call does not synchronise
// Returns 0 on success; -1 bad arguments; -2 the problem needs more scratch than the static array holds
// (message on stderr says how much); -10/-11 the scratch address could not be resolved; -20/-21 launch errors.
// The first call on each device resolves the device address of the scratch array
// (cudaGetSymbolAddress) and sets the kernels' shared-memory attribute; make that first call before any
// stream capture begins.  Calls that may overlap on the GPU race on the shared scratch (see above).
extern "C" int sparknet_tc_gemm_static_scratch(int a_mn, int b_mn, int M, int N, int R, const float* A, const float* B,
                                               float* D, void* stream) {
  using namespace sn;
  if (M <= 0 || N <= 0 || R <= 0 || !A || !B || !D) {
    fprintf(stderr, "libccaffe(b200): sparknet_tc_gemm_static_scratch: bad arguments (M %d N %d R %d)\n", M, N, R);
    return -1;
  }
  int err = 0;
  float* scratch = static_scratch_ptr(&err);
  if (!scratch) return err;
  cudaStream_t s = (cudaStream_t)stream;
  const int n_mma = pick_n_mma(N);
  const int avec = ((uintptr_t)A & 15) == 0, bvec = ((uintptr_t)B & 15) == 0;
  if (a_mn) {
    LoadDenseMN al{A, M, M, R, TBM, (avec && M % 4 == 0) ? 1 : 0};
    if (b_mn) return launch_static(s, al, LoadDenseMN{B, N, N, R, n_mma, (bvec && N % 4 == 0) ? 1 : 0}, M, N, R, n_mma, D, scratch);
    return launch_static(s, al, LoadDenseK{B, R, N, R, n_mma, (bvec && R % 4 == 0) ? 1 : 0}, M, N, R, n_mma, D, scratch);
  }
  LoadDenseK al{A, R, M, R, TBM, (avec && R % 4 == 0) ? 1 : 0};
  if (b_mn) return launch_static(s, al, LoadDenseMN{B, N, N, R, n_mma, (bvec && N % 4 == 0) ? 1 : 0}, M, N, R, n_mma, D, scratch);
  return launch_static(s, al, LoadDenseK{B, R, N, R, n_mma, (bvec && R % 4 == 0) ? 1 : 0}, M, N, R, n_mma, D, scratch);
}
