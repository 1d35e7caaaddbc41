// kernels.cu — the bandwidth-bound kernels of the hot path (sm_100a).
//
// Each kernel replaces a group of reference kernels / cuBLAS level-1 calls
// and is bounded by HBM bandwidth; DESIGN.md lists the algorithmic bytes per element used for the
// roofline.  Common rules: 128-bit accesses where the layout allows, grid = multiple of the SM
// count with grid-stride loops, warp-shuffle reductions, no atomics on floating point (fixed
// reduction order => run-to-run deterministic), IEEE arithmetic (no -use_fast_math).
#include <algorithm>
#include <cfloat>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>

#include "kernels.hpp"

namespace sn {

static constexpr int kSMs = 148;

static inline int grid_for(size_t work_items, int block, int max_waves = 8) {
  size_t b = (work_items + block - 1) / block;
  size_t cap = (size_t)kSMs * max_waves;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

__device__ __forceinline__ float round_tf32_rna(float x) {      // cvt.rna.tf32.f32: what the packed tensor-core operands hold
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ------------------------------------------------------------------------------------------
// fill / axpy / split_add
// ------------------------------------------------------------------------------------------
__global__ void fill_kernel(size_t n4, size_t n, float4* __restrict__ p4, float* __restrict__ p, float v) {
  SN_GRID_DEP_LAUNCH_ALWAYS();
  SN_GRID_DEP_WAIT();
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const float4 v4 = make_float4(v, v, v, v);
  for (size_t j = i; j < n4; j += stride) p4[j] = v4;
  for (size_t j = n4 * 4 + i; j < n; j += stride) p[j] = v;
}
void fill(cudaStream_t s, size_t n, float* p, float v) {
  if (!n) return;
  const size_t n4 = (((uintptr_t)p & 15) == 0) ? n / 4 : 0;
  fill_kernel<<<grid_for(n4 ? n4 : n, 256), 256, 0, s>>>(n4, n, (float4*)p, p, v);
  count_launch();
}
__global__ void split_add_kernel(size_t n, const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ o) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    o[i] = a[i] + b[i];
}
void split_add(cudaStream_t s, size_t n, const float* a, const float* b, float* out) {
  split_add_kernel<<<grid_for(n, 256), 256, 0, s>>>(n, a, b, out);
  count_launch();
}
__global__ void axpy_kernel(size_t n, float a, const float* __restrict__ x, float* __restrict__ y) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    y[i] = __fadd_rn(y[i], __fmul_rn(a, x[i]));
}
void axpy(cudaStream_t s, size_t n, float a, const float* x, float* y) {
  axpy_kernel<<<grid_for(n, 256), 256, 0, s>>>(n, a, x, y);
  count_launch();
}

// ------------------------------------------------------------------------------------------
// Byte pixels to fp32 — the device side of the byte-input path (runtime.cpp JavaDataLayer).  The reference widens on
// the host inside its data callback (ByteImage.java:86-104: float = (signed) byte; ImageNetApp.scala:158-175 then
// subtracts the mean image in fp32); the same two fp32 operations here, so the blob is bit-identical to the fp32
// callback path.  5 B per element: 1 read + 4 written.
// ------------------------------------------------------------------------------------------
template <bool SIGNED>
__global__ void __launch_bounds__(256) widen_bytes_kernel(size_t n, const uint4* __restrict__ src16, const uint8_t* __restrict__ src,
                                                          const float* __restrict__ mean, FastDiv ddiv, FastDiv dper, float scale,
                                                          float4* __restrict__ dst4, float* __restrict__ dst, int vec) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  const size_t n16 = vec ? n >> 4 : 0;                       // unaligned buffers: every element through the scalar tail loop
  auto one = [&](uint32_t byte, size_t i) -> float {
    float v = SIGNED ? (float)(int)(signed char)byte : (float)byte;
    if (mean) {
      const uint32_t q = ddiv.div((uint32_t)i);
      v = __fsub_rn(v, __ldg(mean + (q - dper.div(q) * dper.d)));
    }
    return scale == 1.f ? v : __fmul_rn(v, scale);
  };
  for (size_t g = blockIdx.x * (size_t)blockDim.x + threadIdx.x; g < n16; g += (size_t)gridDim.x * blockDim.x) {
    const uint4 b = __ldg(src16 + g);
    const uint32_t w[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const size_t i = (g << 4) + (size_t)(k << 2);
      dst4[(g << 2) + k] = make_float4(one(w[k] & 0xffu, i), one((w[k] >> 8) & 0xffu, i + 1), one((w[k] >> 16) & 0xffu, i + 2), one(w[k] >> 24, i + 3));
    }
  }
  for (size_t i = (n16 << 4) + blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = one(src[i], i);
}
void widen_bytes(cudaStream_t s, size_t n, const uint8_t* src, bool is_signed, const float* mean, int mean_div, int mean_period,
                 float scale, float* dst) {
  if (!n) return;
  if (n >= (1ull << 31)) fatal("widen_bytes: %zu elements exceed the 32-bit index range of the mean lookup", n);
  const int vec = (((uintptr_t)src | (uintptr_t)dst) & 15) == 0 ? 1 : 0;
  const FastDiv ddiv((uint32_t)std::max(mean_div, 1)), dper((uint32_t)std::max(mean_period, 1));
  const int grid = grid_for(vec ? n / 16 + 1 : n, 256, 16);
  if (is_signed) widen_bytes_kernel<true><<<grid, 256, 0, s>>>(n, (const uint4*)src, src, mean, ddiv, dper, scale, (float4*)dst, dst, vec);
  else widen_bytes_kernel<false><<<grid, 256, 0, s>>>(n, (const uint4*)src, src, mean, ddiv, dper, scale, (float4*)dst, dst, vec);
  count_launch();
}

// ------------------------------------------------------------------------------------------
// ReLU — reference ReLUForward/ReLUBackward caffe/src/caffe/layers/relu_layer.cu:9-14,35-41
// (CPU form relu_layer.cpp:16-17,32-33).  fwd 8 B/elt, bwd 12 B/elt.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float relu1(float x, float slope) { return fmaxf(x, 0.f) + slope * fminf(x, 0.f); }
__global__ void relu_fwd_kernel(size_t n4, size_t n, const float4* __restrict__ x4, float4* __restrict__ y4,
                                const float* __restrict__ x, float* __restrict__ y, float slope) {
  SN_GRID_DEP_LAUNCH_ALWAYS();
  SN_GRID_DEP_WAIT();
  const size_t stride = (size_t)gridDim.x * blockDim.x, i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  for (size_t j = i; j < n4; j += stride) {
    float4 v = x4[j];
    v.x = relu1(v.x, slope); v.y = relu1(v.y, slope); v.z = relu1(v.z, slope); v.w = relu1(v.w, slope);
    y4[j] = v;
  }
  for (size_t j = n4 * 4 + i; j < n; j += stride) y[j] = relu1(x[j], slope);
}
void relu_fwd(cudaStream_t s, size_t n, const float* x, float* y, float slope) {
  const size_t n4 = n / 4;
  relu_fwd_kernel<<<grid_for(n4 + 1, 256), 256, 0, s>>>(n4, n, (const float4*)x, (float4*)y, x, y, slope);
  count_launch();
}
__device__ __forceinline__ float relu_g(float dy, float x, float slope) {
  return dy * ((x > 0.f ? 1.f : 0.f) + slope * (x <= 0.f ? 1.f : 0.f));
}
__global__ void relu_bwd_kernel(size_t n4, size_t n, const float4* __restrict__ dy4, const float4* __restrict__ x4,
                                float4* __restrict__ dx4, const float* __restrict__ dy, const float* __restrict__ x,
                                float* __restrict__ dx, float slope) {
  SN_GRID_DEP_LAUNCH_ALWAYS();
  SN_GRID_DEP_WAIT();
  const size_t stride = (size_t)gridDim.x * blockDim.x, i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  for (size_t j = i; j < n4; j += stride) {
    const float4 g = dy4[j], v = x4[j];
    dx4[j] = make_float4(relu_g(g.x, v.x, slope), relu_g(g.y, v.y, slope), relu_g(g.z, v.z, slope), relu_g(g.w, v.w, slope));
  }
  for (size_t j = n4 * 4 + i; j < n; j += stride) dx[j] = relu_g(dy[j], x[j], slope);
}
void relu_bwd(cudaStream_t s, size_t n, const float* dy, const float* x, float* dx, float slope) {
  const size_t n4 = n / 4;
  relu_bwd_kernel<<<grid_for(n4 + 1, 256), 256, 0, s>>>(n4, n, (const float4*)dy, (const float4*)x, (float4*)dx, dy, x, dx, slope);
  count_launch();
}

// ------------------------------------------------------------------------------------------
// Dropout — reference DropoutForward/Backward caffe/src/caffe/layers/dropout_layer.cu:10-16,39-45
// with curandGenerate replaced by in-kernel Philox4x32-10 (counter = element/4 + offset, key = seed).
// mask = (u32 > ratio*UINT_MAX) as in dropout_layer.cu:14,27; y = x * mask * 1/(1-ratio).
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(M0, c.x), lo0 = M0 * c.x;
    const uint32_t hi1 = __umulhi(M1, c.z), lo1 = M1 * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += W0;
    k.y += W1;
  }
  return c;
}
__global__ void dropout_fwd_kernel(size_t n, const float* __restrict__ x, float* __restrict__ y, uint8_t* __restrict__ mask,
                                   uint32_t thres, float scale, bool draw, uint64_t seed, uint64_t offset,
                                   const unsigned long long* __restrict__ offset_dev) {
  SN_GRID_DEP_LAUNCH_ALWAYS();
  SN_GRID_DEP_WAIT();
  const size_t n4 = (n + 3) / 4;
  if (offset_dev) offset = *offset_dev;          // per-iteration value published on the device (replayed graphs)
  for (size_t q = blockIdx.x * (size_t)blockDim.x + threadIdx.x; q < n4; q += (size_t)gridDim.x * blockDim.x) {
    uint32_t r[4];
    if (draw) {
      const uint64_t ctr = q + offset;
      const uint4 v = philox4x32_10(make_uint4((uint32_t)ctr, (uint32_t)(ctr >> 32), 0u, 0u),
                                    make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
      r[0] = v.x; r[1] = v.y; r[2] = v.z; r[3] = v.w;
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const size_t i = q * 4 + e;
      if (i < n) {
        uint8_t m;
        if (draw) { m = r[e] > thres ? 1 : 0; mask[i] = m; } else { m = mask[i]; }
        y[i] = x[i] * (float)m * scale;
      }
    }
  }
}
void dropout_fwd(cudaStream_t s, size_t n, const float* x, float* y, uint8_t* mask, float ratio, float scale, bool draw,
                 uint64_t seed, uint64_t offset, const unsigned long long* offset_dev) {
  const uint32_t thres = (uint32_t)((double)0xFFFFFFFFu * ratio);   // dropout_layer.cpp:18 uint_thres_
  dropout_fwd_kernel<<<grid_for((n + 3) / 4, 256), 256, 0, s>>>(n, x, y, mask, thres, scale, draw, seed, offset, offset_dev);
  count_launch();
}
__global__ void dropout_bwd_kernel(size_t n, const float* __restrict__ dy, const uint8_t* __restrict__ mask,
                                   float* __restrict__ dx, float scale) {
  SN_GRID_DEP_LAUNCH_ALWAYS();
  SN_GRID_DEP_WAIT();
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    dx[i] = dy[i] * (float)mask[i] * scale;
}
void dropout_bwd(cudaStream_t s, size_t n, const float* dy, const uint8_t* mask, float* dx, float scale) {
  dropout_bwd_kernel<<<grid_for(n, 256), 256, 0, s>>>(n, dy, mask, dx, scale);
  count_launch();
}

// ------------------------------------------------------------------------------------------
// Pooling — reference MaxPoolForward/MaxPoolBackward/AvePoolForward/AvePoolBackward
// caffe/src/caffe/layers/pooling_layer.cu:11-47,50-80,217-260,263-296 (CPU: pooling_layer.cpp:136-292).
// Tie rule: strict '>' scanning h then w (first maximum wins), start value -FLT_MAX, mask = h*W+w.
// AVE divisor = window area clipped to H+pad / W+pad (counts the padding), true division.
// ------------------------------------------------------------------------------------------
// ---- tile kernels: every window shape the fixed-shape kernels below do not cover, and both directions of AVE ----
// A work unit is a CONTIGUOUS span of the input tensor: either `ppb` whole (n, c) planes (small planes: several per unit,
// so a CTA always has a few thousand elements to chew on) or, for planes too large for shared memory, the input rows that
// one band of output rows needs.  The span is staged in shared memory with 16-byte loads (ragged head / tail by scalar
// loads; the staging keeps the span's 16-byte phase so both sides of the copy are aligned), every window is then scanned
// there in the reference's order (h outer, w inner, strict '>' for MAX, plain fp32 sum for AVE), and the results of a
// unit — also contiguous in memory — leave as 16-byte stores where the destination phase allows.  Global traffic is the
// algorithmic 4*in + 4*out (+ 4*out mask) bytes; strided window reads never leave the SM.
struct PoolTile {
  int ppb;             // planes per unit (> 1 only when bands == 1)
  int band;            // rows per unit: output rows (forward) / input rows (backward)
  int bands;           // units per plane group
  unsigned units;
  int vec;             // 1: source and destination pointers are 16-byte aligned
};
// sm[(goff & 3) + i] = src[goff + i], i < len
__device__ __forceinline__ void pool_stage_span(float* __restrict__ sm, const float* __restrict__ src, size_t goff, int len, bool vec) {
  const int lead = (int)(goff & 3);
  const float* base = src + (goff - lead);                  // 16-byte aligned when src is
  const int total = lead + len;
  if (vec) {
    const int n4 = total >> 2;                              // whole float4s inside [base, base + total): never past the span's end
    const float4* b4 = reinterpret_cast<const float4*>(base);
    float4* s4 = reinterpret_cast<float4*>(sm);
    for (int k = threadIdx.x; k < n4; k += blockDim.x) s4[k] = __ldg(b4 + k);
    for (int i = (n4 << 2) + threadIdx.x; i < total; i += blockDim.x) sm[i] = __ldg(base + i);
  } else {
    for (int i = lead + threadIdx.x; i < total; i += blockDim.x) sm[i] = __ldg(base + i);
  }
}
// Results of a unit: element j of `n` goes to dst[off + j].  Threads own groups of four elements aligned to the
// DESTINATION's 16-byte phase; `f(j)` computes one element.
template <class T4, class T, class F>
__device__ __forceinline__ void pool_store_span(T* __restrict__ dst, size_t off, int n, bool vec, F f) {
  const int lead = (int)(off & 3);
  T* base = dst + (off - lead);
  const int groups = (lead + n + 3) >> 2;
  for (int k = threadIdx.x; k < groups; k += blockDim.x) {
    const int j0 = (k << 2) - lead;
    if (vec && j0 >= 0 && j0 + 3 < n) {
      T4 v;
      v.x = f(j0); v.y = f(j0 + 1); v.z = f(j0 + 2); v.w = f(j0 + 3);
      reinterpret_cast<T4*>(base)[k] = v;
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e)
        if (j0 + e >= 0 && j0 + e < n) base[(k << 2) + e] = f(j0 + e);
    }
  }
}
// forward: MAX writes (value, argmax), AVE the window mean with the padded-area divisor (pooling_layer.cpp:179-212)
template <bool AVE>
__global__ void __launch_bounds__(256) pool_fwd_tile_kernel(PoolGeom g, PoolTile t, FastDiv dPW, FastDiv dPH, const float* __restrict__ x,
                                                            float* __restrict__ y, int* __restrict__ mask) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  extern __shared__ __align__(16) float sm_pool[];
  const int HW = g.H * g.W, P = g.PH * g.PW;
  const unsigned planes = (unsigned)(g.N * g.C);
  for (unsigned u = blockIdx.x; u < t.units; u += gridDim.x) {
    const unsigned pg = u / (unsigned)t.bands, b = u - pg * (unsigned)t.bands;
    const unsigned p0 = pg * (unsigned)t.ppb;
    const int np = (int)min((unsigned)t.ppb, planes - p0);
    const int r0 = (int)b * t.band, r1 = min(r0 + t.band, g.PH);
    int hs0 = 0, rows_in = g.H;
    if (t.bands > 1) {                                       // one plane, the input rows of output rows [r0, r1)
      hs0 = max(r0 * g.sh - g.ph, 0);
      rows_in = max(min((r1 - 1) * g.sh - g.ph + g.kh, g.H) - hs0, 0);
    }
    const size_t goff = (size_t)p0 * HW + (size_t)hs0 * g.W;
    const int len = t.bands > 1 ? rows_in * g.W : np * HW;
    __syncthreads();                                         // the previous unit's readers are done with the tile
    pool_stage_span(sm_pool, x, goff, len, t.vec != 0);
    __syncthreads();
    const float* tile = sm_pool + (goff & 3) - hs0 * g.W;   // tile[pl * HW + h * W + w]
    const int nout = np * (r1 - r0) * g.PW;
    const size_t yoff = (size_t)p0 * P + (size_t)r0 * g.PW;
    auto window = [&](int j, int& arg) -> float {
      uint32_t q, pw, pl, r;
      dPW.divmod((uint32_t)j, q, pw);
      if (t.bands > 1) { pl = 0; r = q; } else dPH.divmod(q, pl, r);
      const int ph = r0 + (int)r;
      int hs = ph * g.sh - g.ph, ws = (int)pw * g.sw - g.pw;
      const float* tp = tile + (int)pl * HW;
      if (AVE) {
        int he = min(hs + g.kh, g.H + g.ph), we = min(ws + g.kw, g.W + g.pw);
        const int pool_size = (he - hs) * (we - ws);
        hs = max(hs, 0); ws = max(ws, 0); he = min(he, g.H); we = min(we, g.W);
        float acc = 0.f;
        for (int h = hs; h < he; ++h)
          for (int w = ws; w < we; ++w) acc += tp[h * g.W + w];
        return __fdiv_rn(acc, (float)pool_size);
      }
      const int he = min(hs + g.kh, g.H), we = min(ws + g.kw, g.W);
      hs = max(hs, 0); ws = max(ws, 0);
      float best = -FLT_MAX;
      int bi = -1;
      for (int h = hs; h < he; ++h)
        for (int w = ws; w < we; ++w) {
          const float v = tp[h * g.W + w];
          if (v > best) { best = v; bi = h * g.W + w; }
        }
      arg = bi;
      return best;
    };
    if (AVE) {
      pool_store_span<float4>(y, yoff, nout, t.vec != 0, [&](int j) { int a; return window(j, a); });
    } else {
      // value and argmax index come from one scan; both leave as 16-byte stores (same group walk as pool_store_span)
      const int lead = (int)(yoff & 3), groups = (lead + nout + 3) >> 2;
      for (int k0 = 0; k0 < groups; k0 += blockDim.x) {
        const int k = k0 + threadIdx.x;
        if (k < groups) {
          const int j0 = (k << 2) - lead;
          float v[4]; int a[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) { a[e] = -1; v[e] = (j0 + e >= 0 && j0 + e < nout) ? window(j0 + e, a[e]) : 0.f; }
          float* yb = y + (yoff - lead);
          int* mb = mask + (yoff - lead);
          if (t.vec && j0 >= 0 && j0 + 3 < nout) {
            reinterpret_cast<float4*>(yb)[k] = make_float4(v[0], v[1], v[2], v[3]);
            reinterpret_cast<int4*>(mb)[k] = make_int4(a[0], a[1], a[2], a[3]);
          } else {
#pragma unroll
            for (int e = 0; e < 4; ++e)
              if (j0 + e >= 0 && j0 + e < nout) { yb[(k << 2) + e] = v[e]; mb[(k << 2) + e] = a[e]; }
          }
        }
      }
    }
  }
}
static constexpr int kPoolTileFloats = 6144;                 // staged span per unit: 24 KB, eight CTAs per SM
static constexpr int kPoolTileMaxFloats = 50 * 1024;         // opt-in ceiling (200 KB of the 227 KB an SM has)
// rows_for(band) = rows of the staged tensor a band needs; row_floats = floats per staged row
template <class RowsFor>
static PoolTile plan_pool_tile(size_t planes, int plane_floats, int row_floats, int out_rows, RowsFor rows_for, size_t* smem_floats,
                               int work_per_plane) {
  PoolTile t{};
  if (plane_floats <= kPoolTileFloats) {
    size_t ppb = std::max<size_t>(1, (size_t)kPoolTileFloats / (size_t)plane_floats);
    ppb = std::min(ppb, std::max<size_t>(1, planes / (size_t)(kSMs * 4)));       // small layers: keep every SM busy
    ppb = std::min(ppb, std::max<size_t>(1, (size_t)32768 / (size_t)std::max(work_per_plane, 1)));
    t.ppb = (int)ppb; t.band = out_rows; t.bands = 1;
    *smem_floats = (size_t)t.ppb * plane_floats + 4;
  } else {
    int band = out_rows;
    while (band > 1 && (size_t)rows_for(band) * row_floats > (size_t)kPoolTileFloats) band = (band + 1) / 2;
    if ((size_t)rows_for(band) * row_floats > (size_t)kPoolTileMaxFloats)
      fatal("pooling: a window band of %d rows x %d floats does not fit in shared memory", rows_for(band), row_floats);
    t.ppb = 1; t.band = band; t.bands = (out_rows + band - 1) / band;
    *smem_floats = (size_t)rows_for(band) * row_floats + 4;
  }
  t.units = (unsigned)(((planes + t.ppb - 1) / t.ppb) * (size_t)t.bands);
  return t;
}
template <class K>
static void pool_tile_smem_attr(K kernel, size_t bytes) {
  if (bytes > 48 * 1024) SN_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
}
template <bool AVE>
static void pool_fwd_tile(cudaStream_t s, const PoolGeom& g, const float* x, float* y, int* mask) {
  size_t smf = 0;
  const size_t planes = (size_t)g.N * g.C;
  PoolTile t = plan_pool_tile(planes, g.H * g.W, g.W, g.PH, [&](int band) { return std::min((band - 1) * g.sh + g.kh, g.H); }, &smf,
                              g.PH * g.PW * g.kh * g.kw / 4);
  t.vec = (((uintptr_t)x | (uintptr_t)y | (uintptr_t)mask) & 15) == 0 ? 1 : 0;
  const size_t bytes = (smf * sizeof(float) + 15) / 16 * 16;
  pool_tile_smem_attr(pool_fwd_tile_kernel<AVE>, bytes);
  const int blocks = (int)std::min<size_t>(t.units, (size_t)kSMs * 8);
  pool_fwd_tile_kernel<AVE><<<blocks, 256, bytes, s>>>(g, t, FastDiv((uint32_t)g.PW), FastDiv((uint32_t)g.PH), x, y, mask);
}

// Thread-per-output form for the window shapes of the four nets (compile-time kernel size and stride, no padding):
// 32-bit index arithmetic with multiply-shift division, all window loads issued before the compare chain.  Window
// elements are read straight from global memory: neighbouring outputs share them through the L1, HBM sees them once.
template <int KH, int KW, int SH, int SW>
__global__ void __launch_bounds__(256) pool_max_fwd_fixed_kernel(PoolGeom g, uint32_t total, FastDiv dPW, FastDiv dPH, const float* __restrict__ x,
                                                                 float* __restrict__ y, int* __restrict__ mask) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  const int HW = g.H * g.W;
  for (uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    uint32_t t, pw, plane, ph;
    dPW.divmod(idx, t, pw);
    dPH.divmod(t, plane, ph);
    const int hs = (int)ph * SH, ws = (int)pw * SW;
    const float* xp = x + (size_t)plane * HW + hs * g.W + ws;
    float v[KH][KW];
#pragma unroll
    for (int a = 0; a < KH; ++a)
#pragma unroll
      for (int b = 0; b < KW; ++b) v[a][b] = (hs + a < g.H && ws + b < g.W) ? __ldg(xp + a * g.W + b) : -FLT_MAX;
    float best = -FLT_MAX;
    int bi = -1;
#pragma unroll
    for (int a = 0; a < KH; ++a)
#pragma unroll
      for (int b = 0; b < KW; ++b)
        if (v[a][b] > best) { best = v[a][b]; bi = (hs + a) * g.W + ws + b; }
    y[idx] = best;
    mask[idx] = bi;
  }
}
void pool_max_fwd(cudaStream_t s, const PoolGeom& g, const float* x, float* y, int* mask) {
  const size_t total = (size_t)g.N * g.C * g.PH * g.PW;
  static const bool force_tile = getenv("SPARKNET_POOL_TILE") && atoi(getenv("SPARKNET_POOL_TILE")) == 1;   // experiments: every shape on the tile kernels
  const bool fixed = !force_tile && g.ph == 0 && g.pw == 0 && total < (1ull << 31) && (size_t)g.N * g.C * g.H * g.W < (1ull << 31);
  if (fixed && g.kh == 3 && g.kw == 3 && g.sh == 2 && g.sw == 2) {
    pool_max_fwd_fixed_kernel<3, 3, 2, 2><<<grid_for(total, 256, 32), 256, 0, s>>>(g, (uint32_t)total, FastDiv((uint32_t)g.PW), FastDiv((uint32_t)g.PH), x, y, mask);
  } else if (fixed && g.kh == 2 && g.kw == 2 && g.sh == 2 && g.sw == 2) {
    pool_max_fwd_fixed_kernel<2, 2, 2, 2><<<grid_for(total, 256, 32), 256, 0, s>>>(g, (uint32_t)total, FastDiv((uint32_t)g.PW), FastDiv((uint32_t)g.PH), x, y, mask);
  } else {
    pool_fwd_tile<false>(s, g, x, y, mask);
  }
  count_launch();
}
// backward, gather form (pooling_layer.cpp:253-270 MAX, :272-292 AVE scatter the top diffs window by window in (ph, pw)
// order; here every input element adds the windows that cover it in that same order, so the fp32 sums are identical).
// A unit stages the pooled rows (dy, and the argmax indices for MAX) that its input rows can receive from; the window
// ranges of every input row / column and the AVE divisors' row / column extents are tabulated once per CTA.
// `gate` (nullable, MAX) = the pooled OUTPUT values: when the pooling input is the output of an in-place ReLU whose
// backward would run next (relu_layer.cpp:26-36: dx = dy * (y > 0)), that test is applied while staging — the element
// that won a window is > 0 exactly when the pooled value is — and the separate 12 B/element ReLU pass disappears.
template <bool AVE>
__global__ void __launch_bounds__(256) pool_bwd_tile_kernel(PoolGeom g, PoolTile t, FastDiv dW, FastDiv dH, const float* __restrict__ dy,
                                                            const int* __restrict__ mask, const float* __restrict__ gate,
                                                            float* __restrict__ dx, int span_floats) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  extern __shared__ __align__(16) float sm_pool[];
  const int P = g.PH * g.PW, HW = g.H * g.W;
  float* sdy = sm_pool;
  int* smask = reinterpret_cast<int*>(sm_pool + span_floats);
  int* rowr = smask + (AVE ? 0 : span_floats);              // packed (first window | one-past-last window << 16) per input row
  int* colr = rowr + g.H;
  int* rext = colr + g.W;                                    // AVE: clipped-to-padding window extent per pooled row / column
  int* cext = rext + g.PH;
  for (int h = threadIdx.x; h < g.H; h += blockDim.x) {
    const int phs = (h + g.ph < g.kh) ? 0 : (h + g.ph - g.kh) / g.sh + 1, phe = min((h + g.ph) / g.sh + 1, g.PH);
    rowr[h] = phs | (phe << 16);
  }
  for (int w = threadIdx.x; w < g.W; w += blockDim.x) {
    const int pws = (w + g.pw < g.kw) ? 0 : (w + g.pw - g.kw) / g.sw + 1, pwe = min((w + g.pw) / g.sw + 1, g.PW);
    colr[w] = pws | (pwe << 16);
  }
  if (AVE) {
    for (int p = threadIdx.x; p < g.PH; p += blockDim.x) { const int hs = p * g.sh - g.ph; rext[p] = min(hs + g.kh, g.H + g.ph) - hs; }
    for (int p = threadIdx.x; p < g.PW; p += blockDim.x) { const int ws = p * g.sw - g.pw; cext[p] = min(ws + g.kw, g.W + g.pw) - ws; }
  }
  const unsigned planes = (unsigned)(g.N * g.C);
  for (unsigned u = blockIdx.x; u < t.units; u += gridDim.x) {
    const unsigned pg = u / (unsigned)t.bands, b = u - pg * (unsigned)t.bands;
    const unsigned p0 = pg * (unsigned)t.ppb;
    const int np = (int)min((unsigned)t.ppb, planes - p0);
    const int h0 = (int)b * t.band, h1 = min(h0 + t.band, g.H);
    __syncthreads();                                         // tables written / previous unit's readers done
    int ps0 = 0, prows = g.PH;
    if (t.bands > 1) {
      ps0 = rowr[h0] & 0xffff;
      prows = max((rowr[h1 - 1] >> 16) - ps0, 0);
    }
    const size_t goff = (size_t)p0 * P + (size_t)ps0 * g.PW;
    const int len = t.bands > 1 ? prows * g.PW : np * P;
    pool_stage_span(sdy, dy, goff, len, t.vec != 0);
    if (!AVE) pool_stage_span(reinterpret_cast<float*>(smask), reinterpret_cast<const float*>(mask), goff, len, t.vec != 0);
    if (!AVE && gate) {
      __syncthreads();
      const int lead = (int)(goff & 3);
      for (int i = threadIdx.x; i < len; i += blockDim.x)
        if (!(__ldg(gate + goff + i) > 0.f)) sdy[lead + i] = 0.f;
    }
    __syncthreads();
    const float* tdy = sdy + (goff & 3) - ps0 * g.PW;       // tdy[pl * P + ph * PW + pw]
    const int* tmask = smask + (goff & 3) - ps0 * g.PW;
    const int nin = np * (h1 - h0) * g.W;
    const size_t xoff = (size_t)p0 * HW + (size_t)h0 * g.W;
    pool_store_span<float4>(dx, xoff, nin, t.vec != 0, [&](int j) -> float {
      uint32_t q, w, pl, r;
      dW.divmod((uint32_t)j, q, w);
      if (t.bands > 1) { pl = 0; r = q; } else dH.divmod(q, pl, r);
      const int h = h0 + (int)r;
      const int rr = rowr[h], cr = colr[w];
      const int phs = rr & 0xffff, phe = rr >> 16, pws = cr & 0xffff, pwe = cr >> 16;
      const int me = h * g.W + (int)w;
      const float* dp = tdy + (int)pl * P;
      const int* mp = tmask + (int)pl * P;
      float acc = 0.f;
      for (int ph = phs; ph < phe; ++ph)
        for (int pw = pws; pw < pwe; ++pw) {
          if (AVE) acc += __fdiv_rn(dp[ph * g.PW + pw], (float)(rext[ph] * cext[pw]));
          else if (mp[ph * g.PW + pw] == me) acc += dp[ph * g.PW + pw];
        }
      return acc;
    });
  }
}
template <bool AVE>
static void pool_bwd_tile(cudaStream_t s, const PoolGeom& g, const float* dy, const int* mask, const float* gate, float* dx) {
  size_t span = 0;
  const size_t planes = (size_t)g.N * g.C;
  // pooled rows that a band of `band` input rows can receive from
  auto rows_for = [&](int band) { return std::min((band - 1) / g.sh + (g.kh + g.sh - 1) / g.sh + 1, g.PH); };
  PoolTile t = plan_pool_tile(planes, g.PH * g.PW, g.PW, g.H, rows_for, &span, g.H * g.W);
  t.vec = (((uintptr_t)dy | (uintptr_t)dx | (AVE ? 0 : (uintptr_t)mask)) & 15) == 0 ? 1 : 0;
  span = (span + 3) / 4 * 4;
  const size_t bytes = (span * (AVE ? 1 : 2) + (size_t)(g.H + g.W + g.PH + g.PW)) * sizeof(float);
  pool_tile_smem_attr(pool_bwd_tile_kernel<AVE>, bytes);
  const int blocks = (int)std::min<size_t>(t.units, (size_t)kSMs * 8);
  pool_bwd_tile_kernel<AVE><<<blocks, 256, bytes, s>>>(g, t, FastDiv((uint32_t)g.W), FastDiv((uint32_t)g.H), dy, mask, gate, dx, (int)span);
}
// Thread-per-input-element gather for the same fixed window shapes: at most ceil(K/S)^2 = 4 candidate windows, tested
// in the reference's (ph, pw) order; 32-bit index arithmetic, divisions by compile-time strides.
template <int KH, int KW, int SH, int SW>
__global__ void __launch_bounds__(256) pool_max_bwd_fixed_kernel(PoolGeom g, uint32_t total, FastDiv dW, FastDiv dH, const float* __restrict__ dy,
                                                                 const int* __restrict__ mask, const float* __restrict__ gate,
                                                                 float* __restrict__ dx) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  constexpr int NH = (KH + SH - 1) / SH, NW = (KW + SW - 1) / SW;
  const int P = g.PH * g.PW;
  for (uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    uint32_t t, w, plane, h;
    dW.divmod(idx, t, w);
    dH.divmod(t, plane, h);
    const int phe = min((int)h / SH + 1, g.PH), pwe = min((int)w / SW + 1, g.PW);       // one past the last covering window
    const int phs = (int)h < KH ? 0 : ((int)h - KH) / SH + 1, pws = (int)w < KW ? 0 : ((int)w - KW) / SW + 1;
    const int me = (int)h * g.W + (int)w;
    const int* mp = mask + (size_t)plane * P;
    const float* gp = dy + (size_t)plane * P;
    int m[NH][NW];
#pragma unroll
    for (int a = 0; a < NH; ++a)
#pragma unroll
      for (int b = 0; b < NW; ++b) m[a][b] = (phs + a < phe && pws + b < pwe) ? __ldg(mp + (phs + a) * g.PW + pws + b) : -2;
    float acc = 0.f;
#pragma unroll
    for (int a = 0; a < NH; ++a)
#pragma unroll
      for (int b = 0; b < NW; ++b)
        if (m[a][b] == me) {
          const int j = (phs + a) * g.PW + pws + b;
          float v = __ldg(gp + j);
          if (gate && !(__ldg(gate + (size_t)plane * P + j) > 0.f)) v = 0.f;
          acc += v;
        }
    dx[idx] = acc;
  }
}
// 3x3 / stride 2 / no padding (every max pooling of the four nets but LeNet's): one thread owns a 2x2 block of
// input elements (rows 2a, 2a+1; columns 2b, 2b+1).  Only the four windows (a-1|a, b-1|b) can cover them, so four
// (mask, dy, gate) loads serve four outputs, with no branches: the even/even element tests all four windows, the
// even/odd and odd/even elements two, the odd/odd element one, always in the reference's (ph, pw) order.
__global__ void __launch_bounds__(256) pool_max_bwd_3x3s2_kernel(PoolGeom g, uint32_t total, FastDiv dB, FastDiv dA, const float* __restrict__ dy,
                                                                 const int* __restrict__ mask, const float* __restrict__ gate,
                                                                 float* __restrict__ dx) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  const int P = g.PH * g.PW, HW = g.H * g.W;
  for (uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    uint32_t t, b, plane, a;
    dB.divmod(idx, t, b);
    dA.divmod(t, plane, a);
    const int h0 = 2 * (int)a, w0 = 2 * (int)b;
    const int* mp = mask + (size_t)plane * P;
    const float* gp = dy + (size_t)plane * P;
    const float* tp = gate ? gate + (size_t)plane * P : nullptr;
    int m[2][2];
    float v[2][2];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int ph = (int)a - 1 + i, pw = (int)b - 1 + j;
        const bool ok = ph >= 0 && ph < g.PH && pw >= 0 && pw < g.PW;
        const int o = ph * g.PW + pw;
        m[i][j] = ok ? __ldg(mp + o) : -2;
        float d = ok ? __ldg(gp + o) : 0.f;
        if (tp && ok && !(__ldg(tp + o) > 0.f)) d = 0.f;
        v[i][j] = d;
      }
    const int me00 = h0 * g.W + w0;
    float* dp = dx + (size_t)plane * HW + me00;
    float acc;
    // (2a, 2b): windows (a-1,b-1), (a-1,b), (a,b-1), (a,b)
    acc = (m[0][0] == me00 ? v[0][0] : 0.f);
    acc += (m[0][1] == me00 ? v[0][1] : 0.f);
    acc += (m[1][0] == me00 ? v[1][0] : 0.f);
    acc += (m[1][1] == me00 ? v[1][1] : 0.f);
    dp[0] = acc;
    if (w0 + 1 < g.W) {                               // (2a, 2b+1): windows (a-1,b), (a,b)
      acc = (m[0][1] == me00 + 1 ? v[0][1] : 0.f);
      acc += (m[1][1] == me00 + 1 ? v[1][1] : 0.f);
      dp[1] = acc;
    }
    if (h0 + 1 < g.H) {
      acc = (m[1][0] == me00 + g.W ? v[1][0] : 0.f);  // (2a+1, 2b): windows (a,b-1), (a,b)
      acc += (m[1][1] == me00 + g.W ? v[1][1] : 0.f);
      dp[g.W] = acc;
      if (w0 + 1 < g.W) dp[g.W + 1] = (m[1][1] == me00 + g.W + 1 ? v[1][1] : 0.f);   // (2a+1, 2b+1): window (a,b)
    }
  }
}
// Strip form of the same 3x3 / stride 2 backward: one thread owns the two input columns (2b, 2b+1) of a plane and walks
// down the rows two at a time, keeping the pooled row above (a-1) in registers.  Per 2x2 block that is two new (mask, dy,
// gate) triples instead of four and no integer division (one per thread, not one per block): the block form above is
// bound by instruction issue, not by memory.  Same windows, same (ph, pw) order of the additions: identical results.
__global__ void __launch_bounds__(256) pool_max_bwd_3x3s2_strip_kernel(PoolGeom g, uint32_t total, FastDiv dB, int A, const float* __restrict__ dy,
                                                                       const int* __restrict__ mask, const float* __restrict__ gate,
                                                                       float* __restrict__ dx) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  const int P = g.PH * g.PW, HW = g.H * g.W;
  for (uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    uint32_t plane, b;
    dB.divmod(idx, plane, b);
    const int w0 = 2 * (int)b;
    const int* mp = mask + (size_t)plane * P;
    const float* gp = dy + (size_t)plane * P;
    const float* tp = gate ? gate + (size_t)plane * P : nullptr;
    const bool okl = (int)b - 1 >= 0, okr = (int)b < g.PW;      // window columns b-1 and b exist
    const bool has_w1 = w0 + 1 < g.W;
    float* dp = dx + (size_t)plane * HW + w0;
    int m0l = -2, m0r = -2;                                      // pooled row a-1 (none above the first)
    float v0l = 0.f, v0r = 0.f;
    // pooled row a, loaded one step ahead
    auto load_row = [&](int a, int& ml, int& mr, float& vl, float& vr) {
      ml = mr = -2; vl = vr = 0.f;
      if (a < g.PH) {
        const int o = a * g.PW + (int)b;
        if (okl) { ml = __ldg(mp + o - 1); vl = __ldg(gp + o - 1); if (tp && !(__ldg(tp + o - 1) > 0.f)) vl = 0.f; }
        if (okr) { mr = __ldg(mp + o); vr = __ldg(gp + o); if (tp && !(__ldg(tp + o) > 0.f)) vr = 0.f; }
      }
    };
    int m1l, m1r; float v1l, v1r;
    load_row(0, m1l, m1r, v1l, v1r);
    for (int a = 0; a < A; ++a) {
      int m2l, m2r; float v2l, v2r;
      load_row(a + 1, m2l, m2r, v2l, v2r);                       // next step's loads are in flight during this step's stores
      const int h0 = 2 * a, me00 = h0 * g.W + w0;
      float acc;
      acc = (m0l == me00 ? v0l : 0.f);                            // (2a, 2b): windows (a-1,b-1), (a-1,b), (a,b-1), (a,b)
      acc += (m0r == me00 ? v0r : 0.f);
      acc += (m1l == me00 ? v1l : 0.f);
      acc += (m1r == me00 ? v1r : 0.f);
      dp[me00 - w0] = acc;
      if (has_w1) {                                               // (2a, 2b+1): windows (a-1,b), (a,b)
        acc = (m0r == me00 + 1 ? v0r : 0.f);
        acc += (m1r == me00 + 1 ? v1r : 0.f);
        dp[me00 - w0 + 1] = acc;
      }
      if (h0 + 1 < g.H) {
        acc = (m1l == me00 + g.W ? v1l : 0.f);                    // (2a+1, 2b): windows (a,b-1), (a,b)
        acc += (m1r == me00 + g.W ? v1r : 0.f);
        dp[me00 - w0 + g.W] = acc;
        if (has_w1) dp[me00 - w0 + g.W + 1] = (m1r == me00 + g.W + 1 ? v1r : 0.f);   // (2a+1, 2b+1): window (a,b)
      }
      m0l = m1l; m0r = m1r; v0l = v1l; v0r = v1r;
      m1l = m2l; m1r = m2r; v1l = v2l; v1r = v2r;
    }
  }
}
// Row form of the same backward: a group of G = 8 / 16 / 32 lanes owns one plane (G >= ceil(W / 2)), lane b the input
// columns (2b, 2b+1), and the group walks down the plane two input rows per step.  Per step a lane loads ONE pooled
// element (mask, dy, gate of window (a, b)) — the left neighbour's comes by shuffle, the row above stays in registers —
// and the two finished input rows leave through shuffles so that every store instruction writes a contiguous run of the
// row (the block and strip forms write every other float per instruction).  Same windows, same order of additions.
template <int G, int DEPTH>
__global__ void __launch_bounds__(256) pool_max_bwd_3x3s2_row_kernel(PoolGeom g, uint32_t planes, int A, const float* __restrict__ dy,
                                                                     const int* __restrict__ mask, const float* __restrict__ gate,
                                                                     float* __restrict__ dx) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  const int P = g.PH * g.PW, HW = g.H * g.W;
  const int lane = threadIdx.x & 31, b = lane & (G - 1);
  const uint32_t groups_per_block = 256 / G;
  const uint32_t first = blockIdx.x * groups_per_block + threadIdx.x / G, stride = gridDim.x * groups_per_block;
  const int w0 = 2 * b;
  const bool okw = b < g.PW;                            // window column b exists
  for (uint32_t plane0 = first; plane0 - (plane0 % (32 / G)) < planes; plane0 += stride) {   // whole warps stay in the loop (shuffles)
    const bool live = plane0 < planes;
    const uint32_t plane = live ? plane0 : planes - 1;
    const int* mp = mask + (size_t)plane * P;
    const float* gp = dy + (size_t)plane * P;
    const float* tp = gate ? gate + (size_t)plane * P : nullptr;
    float* dp = dx + (size_t)plane * HW;
    int m0l = -2, m0r = -2;
    float v0l = 0.f, v0r = 0.f;
    auto load_row = [&](int a, int& mr, float& vr) {
      mr = -2; vr = 0.f;
      if (a < g.PH && okw) {
        const int o = a * g.PW + b;
        mr = __ldg(mp + o); vr = __ldg(gp + o);
        if (tp && !(__ldg(tp + o) > 0.f)) vr = 0.f;
      }
    };
    // pooled rows are fetched kDepth steps ahead through a rotating set of registers: slot k holds row a0 + k and is
    // refilled with row a0 + k + kDepth as soon as step a0 + k has read it.  Measured on B200 (CaffeNet pool1 / pool2 /
    // pool5 backward): depth 1 132 / 88 / 25 us, depth 2 154 / 91 / 27 us, depth 3 168 / 101 / 28 us — the kernel is not
    // waiting on memory latency, the extra registers only cost occupancy; depth 1 is the default (SPARKNET_POOL_DEPTH).
    constexpr int kDepth = DEPTH;
    int mq[kDepth]; float vq[kDepth];
#pragma unroll
    for (int k = 0; k < kDepth; ++k) load_row(k, mq[k], vq[k]);
    for (int a0 = 0; a0 < A; a0 += kDepth) {
#pragma unroll
      for (int k = 0; k < kDepth; ++k) {
        const int a = a0 + k;
        if (a >= A) break;                                       // uniform across the warp
        const int m1r = mq[k];
        const float v1r = vq[k];
        load_row(a + kDepth, mq[k], vq[k]);
        // window (a, b-1) lives in the lane to the left
        int m1l = __shfl_up_sync(0xffffffffu, m1r, 1, G);
        float v1l = __shfl_up_sync(0xffffffffu, v1r, 1, G);
        if (b == 0) { m1l = -2; v1l = 0.f; }
        const int h0 = 2 * a, me00 = h0 * g.W + w0;
        float x00, x01, x10, x11;
        x00 = (m0l == me00 ? v0l : 0.f);                       // (2a, 2b): windows (a-1,b-1), (a-1,b), (a,b-1), (a,b)
        x00 += (m0r == me00 ? v0r : 0.f);
        x00 += (m1l == me00 ? v1l : 0.f);
        x00 += (m1r == me00 ? v1r : 0.f);
        x01 = (m0r == me00 + 1 ? v0r : 0.f);                   // (2a, 2b+1): windows (a-1,b), (a,b)
        x01 += (m1r == me00 + 1 ? v1r : 0.f);
        x10 = (m1l == me00 + g.W ? v1l : 0.f);                 // (2a+1, 2b): windows (a,b-1), (a,b)
        x10 += (m1r == me00 + g.W ? v1r : 0.f);
        x11 = (m1r == me00 + g.W + 1 ? v1r : 0.f);             // (2a+1, 2b+1): window (a,b)
        // column c of a row sits in lane c >> 1, component c & 1: lane j stores columns j and G + j
        const int src0 = b >> 1, src1 = (G >> 1) + (b >> 1);
        const bool odd = b & 1;
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          const float e = r ? x10 : x00, o = r ? x11 : x01;
          const float lo_e = __shfl_sync(0xffffffffu, e, src0, G), lo_o = __shfl_sync(0xffffffffu, o, src0, G);
          const float hi_e = __shfl_sync(0xffffffffu, e, src1, G), hi_o = __shfl_sync(0xffffffffu, o, src1, G);
          const int h = h0 + r;
          if (live && h < g.H) {
            float* row = dp + h * g.W;
            if (b < g.W) row[b] = odd ? lo_o : lo_e;
            if (G + b < g.W) row[G + b] = odd ? hi_o : hi_e;
          }
        }
        m0l = m1l; m0r = m1r; v0l = v1l; v0r = v1r;
      }
    }
  }
}
void pool_max_bwd(cudaStream_t s, const PoolGeom& g, const float* dy, const int* mask, float* dx, const float* gate) {
  const size_t total_in = (size_t)g.N * g.C * g.H * g.W;
  static const bool force_tile = getenv("SPARKNET_POOL_TILE") && atoi(getenv("SPARKNET_POOL_TILE")) == 1;
  const bool fixed = !force_tile && g.ph == 0 && g.pw == 0 && total_in < (1ull << 31);
  if (fixed && g.kh == 3 && g.kw == 3 && g.sh == 2 && g.sw == 2) {
    const int A = (g.H + 1) / 2, B = (g.W + 1) / 2;
    const size_t total = (size_t)g.N * g.C * A * B, strips = (size_t)g.N * g.C * B;
    const char* e = getenv("SPARKNET_POOL_STRIP");      // 0: block form only, 2: strip form always, 3: row form always (tests), default: by size
    const int strip_mode = e ? atoi(e) : 1;
    const uint32_t planes = (uint32_t)(g.N * g.C);
    if ((strip_mode == 3 || (strip_mode == 1 && strips >= (size_t)148 * 2048)) && B <= 32) {
      // large layers: the row form (coalesced stores, one load per window)
      static const int depth = getenv("SPARKNET_POOL_DEPTH") ? atoi(getenv("SPARKNET_POOL_DEPTH")) : 1;
#define SN_POOL_ROW(GV, DV) pool_max_bwd_3x3s2_row_kernel<GV, DV><<<grid_for((size_t)planes * GV, 256, 32), 256, 0, s>>>(g, planes, A, dy, mask, gate, dx)
#define SN_POOL_ROW_D(GV) do { if (depth == 2) SN_POOL_ROW(GV, 2); else if (depth == 3) SN_POOL_ROW(GV, 3); else SN_POOL_ROW(GV, 1); } while (0)
      if (B > 16) SN_POOL_ROW_D(32);
      else if (B > 8) SN_POOL_ROW_D(16);
      else SN_POOL_ROW_D(8);
#undef SN_POOL_ROW_D
#undef SN_POOL_ROW
    } else if (strip_mode == 2 || (strip_mode == 1 && strips >= (size_t)148 * 2048))   // the strip form: half the instructions per element
      pool_max_bwd_3x3s2_strip_kernel<<<grid_for(strips, 256, 32), 256, 0, s>>>(g, (uint32_t)strips, FastDiv((uint32_t)B), A, dy, mask, gate, dx);
    else
      pool_max_bwd_3x3s2_kernel<<<grid_for(total, 256, 32), 256, 0, s>>>(g, (uint32_t)total, FastDiv((uint32_t)B), FastDiv((uint32_t)A), dy, mask, gate, dx);
  } else if (fixed && g.kh == 2 && g.kw == 2 && g.sh == 2 && g.sw == 2) {
    pool_max_bwd_fixed_kernel<2, 2, 2, 2><<<grid_for(total_in, 256, 32), 256, 0, s>>>(g, (uint32_t)total_in, FastDiv((uint32_t)g.W), FastDiv((uint32_t)g.H), dy, mask, gate, dx);
  } else {
    pool_bwd_tile<false>(s, g, dy, mask, gate, dx);
  }
  count_launch();
}
void pool_ave_fwd(cudaStream_t s, const PoolGeom& g, const float* x, float* y) {
  pool_fwd_tile<true>(s, g, x, y, nullptr);
  count_launch();
}
void pool_ave_bwd(cudaStream_t s, const PoolGeom& g, const float* dy, float* dx) {
  pool_bwd_tile<true>(s, g, dy, nullptr, nullptr, dx);
  count_launch();
}

// ------------------------------------------------------------------------------------------
// LRN across channels — reference LRNFillScale / LRNComputeOutput / LRNComputeDiff
// caffe/src/caffe/layers/lrn_layer.cu:9-52,72-77,122-177 (CPU: lrn_layer.cpp:108-152,180-232).
// Consecutive threads are consecutive pixels, so every channel step is a coalesced row access.
// Algorithmic bytes: fwd 12 B/elt (x, scale, y), bwd 20 B/elt; moved here: 8 / 12 (no scale tensor).
// ------------------------------------------------------------------------------------------
// Chunked form (odd local_size up to 11): one thread owns CH consecutive channels of one pixel, loads them plus the halo in
// one batch (independent loads in flight instead of a serial walk over all C channels), and sums each window directly:
//   scale_c = k + alpha/n * sum_{|j-c| <= pre} x_j^2 ,   y_c = x_c * scale_c^-beta .
// Nothing but y is written: the backward kernel recomputes scale from x (cheaper than 8 B/element of scale traffic),
// so the forward pass moves 8 B per element.
__device__ __forceinline__ float sqrt_approx(float x) {
  float r;
  asm("sqrt.approx.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
// sc^-beta and 1/sc from two special-function instructions when beta = 0.75 (AlexNet / CaffeNet): q = sc^-1/2,
// sc^-3/4 = q * sqrt(q), 1/sc = q * q; relative error ~3e-7, far inside the 1e-5 of the fp32 validation mode
__device__ __forceinline__ float lrn_pow_neg(float sc, float beta, bool is075, float* inv_sc = nullptr) {
  if (is075) {
    const float q = rsqrtf(sc);
    if (inv_sc) *inv_sc = q * q;
    return q * sqrt_approx(q);
  }
  if (inv_sc) *inv_sc = __fdividef(1.f, sc);
  return __powf(sc, -beta);
}
// channels per thread: 16 by default (SPARKNET_LRN_CH = 8 / 16 / 32 selects another instantiation)
static int lrn_chunk() {
  static const int ch = getenv("SPARKNET_LRN_CH") ? atoi(getenv("SPARKNET_LRN_CH")) : 16;
  return ch == 8 || ch == 32 ? ch : 16;
}
template <int SIZE, int CH>
__global__ void __launch_bounds__(128) lrn_across_fwd_chunk_kernel(LrnGeom g, int chunks, const float* __restrict__ x, float* __restrict__ y,
                                                                   FlatOut fo, FastDiv dW) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  constexpr int PRE = (SIZE - 1) / 2;
  const int HW = g.H * g.W;
  const float aos = g.alpha / g.size;
  const bool is075 = g.beta == 0.75f;
  const size_t total = (size_t)g.N * chunks * HW;
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int pix = (int)(idx % HW);
    const int ck = (int)((idx / HW) % chunks), n = (int)(idx / ((size_t)HW * chunks));
    const int c0 = ck * CH;
    const float* xp = x + ((size_t)n * g.C) * HW + pix;
    float v[CH + 2 * PRE];
#pragma unroll
    for (int j = 0; j < CH + 2 * PRE; ++j) {
      const int c = c0 - PRE + j;
      v[j] = (c >= 0 && c < g.C) ? xp[(size_t)c * HW] : 0.f;
    }
    float* yp = y + ((size_t)n * g.C) * HW + pix;
    float o[CH];
#pragma unroll
    for (int j = 0; j < CH; ++j) {
      o[j] = 0.f;
      if (c0 + j < g.C) {
        float acc = 0.f;
#pragma unroll
        for (int t = 0; t < SIZE; ++t) acc += v[j + t] * v[j + t];
        const float sc = g.k + acc * aos;
        o[j] = v[j + PRE] * lrn_pow_neg(sc, g.beta, is075);
        yp[(size_t)(c0 + j) * HW] = o[j];
      }
    }
    if (fo.flat) {
      // second copy for the window convolution that reads this blob: its flat grid, channels-last, rounded to tf32 — this
      // thread's CH channels are contiguous there (the consumer's pack kernel then has nothing to do)
      uint32_t h, w;
      dW.divmod((uint32_t)pix, h, w);
      float4* f = reinterpret_cast<float4*>(fo.flat + ((size_t)fo.lead + (size_t)n * fo.Gp + (size_t)h * fo.Wp + w) * fo.C + c0);
#pragma unroll
      for (int j = 0; j < CH; j += 4)
        if (c0 + j + 3 < g.C) f[j >> 2] = make_float4(round_tf32_rna(o[j]), round_tf32_rna(o[j + 1]), round_tf32_rna(o[j + 2]), round_tf32_rna(o[j + 3]));
    }
  }
}
static void lrn_size_check(const LrnGeom& g) {
  if (g.size != 3 && g.size != 5 && g.size != 7 && g.size != 9 && g.size != 11)
    fatal("LRN ACROSS_CHANNELS: local_size %d is not on the supported path (odd sizes 3..11)", g.size);
}
void lrn_across_fwd(cudaStream_t s, const LrnGeom& g, const float* x, float* y, const FlatOut* next) {
  lrn_size_check(g);
  const int CHn = g.size > 5 ? 16 : lrn_chunk(), chunks = (g.C + CHn - 1) / CHn;
  const size_t total = (size_t)g.N * chunks * g.H * g.W;
  FlatOut fo{nullptr, 0, 0, 0, 0};
  if (next && next->flat) {
    if (next->C != g.C || g.C % 4 != 0) fatal("lrn: the consumer's flat grid has %d channels, this layer writes %d", next->C, g.C);
    fo = *next;
  }
  const FastDiv dW((uint32_t)g.W);
#define SN_LRN_F(SZ, CHV) lrn_across_fwd_chunk_kernel<SZ, CHV><<<grid_for(total, 128, 32), 128, 0, s>>>(g, chunks, x, y, fo, dW)
  if (g.size == 5) { if (CHn == 8) SN_LRN_F(5, 8); else if (CHn == 32) SN_LRN_F(5, 32); else SN_LRN_F(5, 16); }
  else if (g.size == 3) { if (CHn == 8) SN_LRN_F(3, 8); else if (CHn == 32) SN_LRN_F(3, 32); else SN_LRN_F(3, 16); }
  else if (g.size == 7) SN_LRN_F(7, 16);
  else if (g.size == 9) SN_LRN_F(9, 16);
  else SN_LRN_F(11, 16);
#undef SN_LRN_F
  count_launch();
}
// Chunked backward, reference LRNComputeDiff (lrn_layer.cu:122-177) / CrossChannelBackward_cpu
// (lrn_layer.cpp:180-232):   dx_c = dy_c * scale_c^-beta - (2 alpha beta / n) * x_c * sum_{|j-c| <= pre} dy_j * y_j / scale_j
// with scale and y recomputed from x (halo 2*pre).  Reads x and dy, writes dx: 12 B per element.
template <int SIZE, int CH>
__global__ void __launch_bounds__(128) lrn_across_bwd_chunk_kernel(LrnGeom g, int chunks, const float* __restrict__ x,
                                                                   const float* __restrict__ dy, float* __restrict__ dx) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  constexpr int PRE = (SIZE - 1) / 2;
  const int HW = g.H * g.W;
  const float aos = g.alpha / g.size;
  const float cache_ratio = 2.f * g.alpha * g.beta / g.size;
  const bool is075 = g.beta == 0.75f;
  const size_t total = (size_t)g.N * chunks * HW;
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int pix = (int)(idx % HW);
    const int ck = (int)((idx / HW) % chunks), n = (int)(idx / ((size_t)HW * chunks));
    const int c0 = ck * CH;
    const size_t base = ((size_t)n * g.C) * HW + pix;
    float v[CH + 4 * PRE], r[CH + 2 * PRE], p[CH + 2 * PRE];
#pragma unroll
    for (int j = 0; j < CH + 4 * PRE; ++j) {
      const int c = c0 - 2 * PRE + j;
      v[j] = (c >= 0 && c < g.C) ? x[base + (size_t)c * HW] : 0.f;
    }
#pragma unroll
    for (int j = 0; j < CH + 2 * PRE; ++j) {
      const int c = c0 - PRE + j;
      r[j] = (c >= 0 && c < g.C) ? dy[base + (size_t)c * HW] : 0.f;          // r holds dy for now
    }
#pragma unroll
    for (int j = 0; j < CH + 2 * PRE; ++j) {                             // channel c0 - PRE + j
      float acc = 0.f;
#pragma unroll
      for (int t = 0; t < SIZE; ++t) acc += v[j + t] * v[j + t];
      const float sc = g.k + acc * aos;
      float inv;
      const float pw = lrn_pow_neg(sc, g.beta, is075, &inv);
      p[j] = pw;
      r[j] = r[j] * (v[j + PRE] * pw) * inv;                                  // dy * y / scale; the raw dy is re-read below
    }
#pragma unroll
    for (int j = 0; j < CH; ++j) {
      const int c = c0 + j;
      if (c < g.C) {
        float acc = 0.f;
#pragma unroll
        for (int t = 0; t < SIZE; ++t) acc += r[j + t];
        const float g_c = dy[base + (size_t)c * HW];                           // L1 hit: loaded a few instructions ago
        dx[base + (size_t)c * HW] = g_c * p[j + PRE] - cache_ratio * v[j + 2 * PRE] * acc;
      }
    }
  }
}
void lrn_across_bwd(cudaStream_t s, const LrnGeom& g, const float* x, const float* dy, float* dx) {
  lrn_size_check(g);
  const int CHn = g.size > 5 ? 16 : lrn_chunk(), chunks = (g.C + CHn - 1) / CHn;
  const size_t total = (size_t)g.N * chunks * g.H * g.W;
#define SN_LRN_B(SZ, CHV) lrn_across_bwd_chunk_kernel<SZ, CHV><<<grid_for(total, 128, 32), 128, 0, s>>>(g, chunks, x, dy, dx)
  if (g.size == 5) { if (CHn == 8) SN_LRN_B(5, 8); else if (CHn == 32) SN_LRN_B(5, 32); else SN_LRN_B(5, 16); }
  else if (g.size == 3) { if (CHn == 8) SN_LRN_B(3, 8); else if (CHn == 32) SN_LRN_B(3, 32); else SN_LRN_B(3, 16); }
  else if (g.size == 7) SN_LRN_B(7, 16);
  else if (g.size == 9) SN_LRN_B(9, 16);
  else SN_LRN_B(11, 16);
#undef SN_LRN_B
  count_launch();
}

// ------------------------------------------------------------------------------------------
// LRN within channel — the reference composes Split + Power(x^2) + AVE pool (k=size, pad=(size-1)/2,
// stride 1) + Power((1+alpha*x)^-beta) + Eltwise PROD (caffe/src/caffe/layers/lrn_layer.cpp:17-65,
// 155-162,235-247; power_layer.cpp:20-93; eltwise_layer.cpp:53-58,111-123; split_layer.cpp:26-49):
// ~12 launches and 4 scratch blobs.  Here: one fused stencil kernel each way, one CTA per (n,c)
// plane staged in shared memory.  fwd 8 B/elt, bwd 12 B/elt, no intermediates kept.
// ------------------------------------------------------------------------------------------
constexpr int kLrnWithinMaxPlane = 1024;
__global__ void lrn_within_fwd_kernel(LrnGeom g, const float* __restrict__ x, float* __restrict__ y) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  __shared__ float xs[kLrnWithinMaxPlane];
  const int HW = g.H * g.W, pre = (g.size - 1) / 2;
  const size_t off = (size_t)blockIdx.x * HW;
  for (int i = threadIdx.x; i < HW; i += blockDim.x) xs[i] = x[off + i];
  __syncthreads();
  const float area = (float)(g.size * g.size);    // clipped-to-(H+pad) area is size^2 everywhere for stride 1
  for (int i = threadIdx.x; i < HW; i += blockDim.x) {
    const int h = i / g.W, w = i % g.W;
    const int hs = max(h - pre, 0), he = min(h - pre + g.size, g.H), ws = max(w - pre, 0), we = min(w - pre + g.size, g.W);
    float acc = 0.f;
    for (int a = hs; a < he; ++a)
      for (int b = ws; b < we; ++b) { const float v = xs[a * g.W + b]; acc += powf(v, 2.f); }
    const float avg = __fdiv_rn(acc, area);
    y[off + i] = xs[i] * powf(__fadd_rn(__fmul_rn(avg, g.alpha), 1.f), -g.beta);
  }
}
void lrn_within_fwd(cudaStream_t s, const LrnGeom& g, const float* x, float* y) {
  if (g.H * g.W > kLrnWithinMaxPlane) fatal("lrn_within: plane %dx%d exceeds the %d-element shared-memory tile", g.H, g.W, kLrnWithinMaxPlane);
  const int threads = g.H * g.W >= 256 ? 256 : (g.H * g.W >= 128 ? 128 : 64);
  lrn_within_fwd_kernel<<<g.N * g.C, threads, 0, s>>>(g, x, y);
  count_launch();
}
__global__ void lrn_within_bwd_kernel(LrnGeom g, const float* __restrict__ x, const float* __restrict__ dy,
                                      float* __restrict__ dx) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  __shared__ float xs[kLrnWithinMaxPlane], ps[kLrnWithinMaxPlane], qs[kLrnWithinMaxPlane];
  const int HW = g.H * g.W, pre = (g.size - 1) / 2;
  const size_t off = (size_t)blockIdx.x * HW;
  for (int i = threadIdx.x; i < HW; i += blockDim.x) xs[i] = x[off + i];
  __syncthreads();
  const float area = (float)(g.size * g.size);
  const float diff_scale = (-g.beta) * g.alpha;
  for (int i = threadIdx.x; i < HW; i += blockDim.x) {
    const int h = i / g.W, w = i % g.W;
    const int hs = max(h - pre, 0), he = min(h - pre + g.size, g.H), ws = max(w - pre, 0), we = min(w - pre + g.size, g.W);
    float acc = 0.f;
    for (int a = hs; a < he; ++a)
      for (int b = ws; b < we; ++b) { const float v = xs[a * g.W + b]; acc += powf(v, 2.f); }
    const float base = __fadd_rn(__fmul_rn(__fdiv_rn(acc, area), g.alpha), 1.f);
    const float p = powf(base, -g.beta);
    const float gy = dy[off + i];
    ps[i] = p * gy;                                            // d(product input) = power_out * dy
    qs[i] = (xs[i] * gy) * (__fdiv_rn(p, base) * diff_scale);  // d(pool out) = (x*dy) * diff_scale * y_pow / base
  }
  __syncthreads();
  for (int i = threadIdx.x; i < HW; i += blockDim.x) {
    const int h = i / g.W, w = i % g.W;
    // windows j=(a,b) that contain pixel i: a in [h+pre-size+1, h+pre] intersect [0,H)
    const int as = max(h + pre - g.size + 1, 0), ae = min(h + pre + 1, g.H);
    const int bs = max(w + pre - g.size + 1, 0), be = min(w + pre + 1, g.W);
    float acc = 0.f;
    for (int a = as; a < ae; ++a)
      for (int b = bs; b < be; ++b) acc += __fdiv_rn(qs[a * g.W + b], area);
    dx[off + i] = __fadd_rn(ps[i], __fmul_rn(acc, __fmul_rn(2.f, xs[i])));
  }
}
void lrn_within_bwd(cudaStream_t s, const LrnGeom& g, const float* x, const float* dy, float* dx) {
  if (g.H * g.W > kLrnWithinMaxPlane) fatal("lrn_within: plane %dx%d exceeds the %d-element shared-memory tile", g.H, g.W, kLrnWithinMaxPlane);
  const int threads = g.H * g.W >= 256 ? 256 : (g.H * g.W >= 128 ? 128 : 64);
  lrn_within_bwd_kernel<<<g.N * g.C, threads, 0, s>>>(g, x, dy, dx);
  count_launch();
}

// ------------------------------------------------------------------------------------------
// SoftmaxWithLoss — reference: kernel_channel_max/subtract/exp/sum/div + SoftmaxLossForwardGPU +
// 2x cublasSasum (+ cublasSdot in Layer::Forward) = ~7 launches and 3 host syncs per iteration
// (caffe/src/caffe/layers/softmax_layer.cu:13-120, softmax_loss_layer.cu:11-120, layer.hpp:468-480).
// Here: one warp per row with shuffle reductions, the loss reduced on the device in a fixed order,
// nothing read back unless the host asks for the loss blob.
// ------------------------------------------------------------------------------------------
__global__ void softmax_rows_kernel(int rows, int classes, const float* __restrict__ x, const float* __restrict__ label,
                                    float* __restrict__ prob, float* __restrict__ row_loss) {
  SN_GRID_DEP_LAUNCH_ALWAYS();
  SN_GRID_DEP_WAIT();
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const float* xr = x + (size_t)warp * classes;
  float* pr = prob + (size_t)warp * classes;
  float m = -FLT_MAX;
  for (int c = lane; c < classes; c += 32) m = fmaxf(m, xr[c]);
  m = warp_max(m);
  float sum = 0.f;
  for (int c = lane; c < classes; c += 32) {
    const float e = expf(xr[c] - m);
    pr[c] = e;
    sum += e;
  }
  sum = warp_sum(sum);
  for (int c = lane; c < classes; c += 32) pr[c] = __fdiv_rn(pr[c], sum);
  __syncwarp();
  if (lane == 0) {
    const int lv = (int)label[warp];
    const float p = (lv >= 0 && lv < classes) ? pr[lv] : 1.f;
    row_loss[warp] = -logf(fmaxf(p, FLT_MIN));
  }
}
// out[0] = (sum_i v[i]) / denom with a fixed-shape tree (deterministic)
__global__ void reduce_rows_kernel(int n, const float* __restrict__ v, float denom, float* __restrict__ out,
                                   float* __restrict__ out2) {
  SN_GRID_DEP_LAUNCH_ALWAYS();
  SN_GRID_DEP_WAIT();
  __shared__ float sh[32];
  float acc = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) acc += v[i];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : 0.f;
    t = warp_sum(t);
    if (threadIdx.x == 0) {
      const float r = __fdiv_rn(t, denom);
      out[0] = r;
      if (out2) out2[0] = r;
    }
  }
}
void softmax_loss_fwd(cudaStream_t s, int rows, int classes, const float* x, const float* label, float* prob,
                      float* loss_out, float* row_scratch, float* loss_log) {
  const int wpb = 4;
  softmax_rows_kernel<<<(rows + wpb - 1) / wpb, wpb * 32, 0, s>>>(rows, classes, x, label, prob, row_scratch);
  reduce_rows_kernel<<<1, 256, 0, s>>>(rows, row_scratch, (float)rows, loss_out, loss_log);
  count_launch(2);
}
__global__ void softmax_loss_bwd_kernel(int rows, int classes, const float* __restrict__ prob, const float* __restrict__ label,
                                        const float* __restrict__ lw, float* __restrict__ dx) {
  SN_GRID_DEP_LAUNCH_ALWAYS();
  SN_GRID_DEP_WAIT();
  const float scale = __fdiv_rn(lw[0], (float)rows);     // loss_weight / count   (softmax_loss_layer.cpp:110-116)
  const size_t total = (size_t)rows * classes;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / classes), c = (int)(i % classes);
    float v = prob[i];
    if (c == (int)label[r]) v -= 1.f;
    dx[i] = v * scale;
  }
}
void softmax_loss_bwd(cudaStream_t s, int rows, int classes, const float* prob, const float* label,
                      const float* loss_weight_dev, float* dx) {
  softmax_loss_bwd_kernel<<<grid_for((size_t)rows * classes, 256), 256, 0, s>>>(rows, classes, prob, label, loss_weight_dev, dx);
  count_launch();
}
// Accuracy top-1 (caffe/src/caffe/layers/accuracy_layer.cpp:48-103): the reference sorts
// (value, index) pairs with std::greater, i.e. among equal values the larger index ranks first.
__global__ void accuracy_rows_kernel(int rows, int classes, const float* __restrict__ x, const float* __restrict__ label,
                                     float* __restrict__ hit) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const float* xr = x + (size_t)warp * classes;
  float bv = -FLT_MAX;
  int bi = -1;
  for (int c = lane; c < classes; c += 32) {
    const float v = xr[c];
    if (v > bv || (v == bv && c > bi)) { bv = v; bi = c; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > bv || (ov == bv && oi > bi)) { bv = ov; bi = oi; }
  }
  if (lane == 0) hit[warp] = (bi == (int)label[warp]) ? 1.f : 0.f;
}
void accuracy_top1(cudaStream_t s, int rows, int classes, const float* x, const float* label, float* acc_out,
                   float* row_scratch) {
  const int wpb = 4;
  accuracy_rows_kernel<<<(rows + wpb - 1) / wpb, wpb * 32, 0, s>>>(rows, classes, x, label, row_scratch);
  reduce_rows_kernel<<<1, 256, 0, s>>>(rows, row_scratch, (float)rows, acc_out, nullptr);
  count_launch(2);
}

// ------------------------------------------------------------------------------------------
// bias gradients
// ------------------------------------------------------------------------------------------
// db[c] (+)= sum_n sum_pix dy[n][c][pix].  Two deterministic stages so that small channel counts still
// fill the GPU: block (c, s) sums the images n = s, s+S, ...; a second tiny kernel adds the S partials of
// each channel in ascending order.
__global__ void conv_bias_grad_partial_kernel(int N, int C, int HW, int S, const float* __restrict__ dy, float* __restrict__ part) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  __shared__ float sh[32];
  const int c = blockIdx.x, s = blockIdx.y;
  float acc = 0.f;
  for (int n = s; n < N; n += S) {
    const float* p = dy + ((size_t)n * C + c) * HW;
    for (int i = threadIdx.x; i < HW; i += blockDim.x) acc += p[i];
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : 0.f;
    t = warp_sum(t);
    if (threadIdx.x == 0) part[(size_t)c * S + s] = t;
  }
}
__global__ void conv_bias_grad_final_kernel(int C, int S, const float* __restrict__ part, float* __restrict__ db, int accumulate) {
  SN_GRID_DEP_LAUNCH_ALWAYS();
  SN_GRID_DEP_WAIT();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float t = 0.f;
  for (int s = 0; s < S; ++s) t += part[(size_t)c * S + s];
  db[c] = accumulate ? db[c] + t : t;
}
void conv_bias_grad(cudaStream_t s, int N, int C, int HW, const float* dy, float* db, bool accumulate) {
  int S = (4 * kSMs + C - 1) / C;
  if (S > N) S = N;
  if (S < 1) S = 1;
  float* part = get_workspace((size_t)C * S);
  const int threads = HW >= 256 ? 256 : (HW >= 128 ? 128 : 64);
  conv_bias_grad_partial_kernel<<<dim3(C, S), threads, 0, s>>>(N, C, HW, S, dy, part);
  conv_bias_grad_final_kernel<<<(C + 127) / 128, 128, 0, s>>>(C, S, part, db, accumulate ? 1 : 0);
  count_launch(2);
}
__global__ void ip_bias_grad_kernel(int rows, int cols, const float* __restrict__ dy, float* __restrict__ db, int accumulate) {
  SN_GRID_DEP_LAUNCH_ALWAYS();
  SN_GRID_DEP_WAIT();
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= cols) return;
  float acc = 0.f;
  for (int i = 0; i < rows; ++i) acc += dy[(size_t)i * cols + j];
  db[j] = accumulate ? db[j] + acc : acc;
}
// InnerProduct backward on the TMA path needs the top gradient twice: rounded to tf32 as the tensor-core operand and
// summed over the batch rows as the bias gradient (inner_product_layer.cpp:100-104 caffe_cpu_gemv).  One pass does both:
// block = 32 float4 columns x 8 row groups, each thread walks rows rg, rg + 8, ..., the eight partial sums are added
// in row-group order (fixed order, no atomics).  cols % 4 == 0.
__global__ void __launch_bounds__(256) ip_round_bias_kernel(int rows, int cols4, const float4* __restrict__ dy, float4* __restrict__ dy_r,
                                                            float* __restrict__ db, int accumulate) {
  SN_GRID_DEP_LAUNCH_ALWAYS();
  SN_GRID_DEP_WAIT();
  __shared__ float4 part[8][32];
  const int cx = threadIdx.x & 31, rg = threadIdx.x >> 5, c4 = blockIdx.x * 32 + cx;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (c4 < cols4) {
    int r = rg;
    for (; r + 24 < rows; r += 32) {                  // four rows in flight per thread: the walk is a chain of memory round trips
      float4 v[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) v[k] = dy[(size_t)(r + 8 * k) * cols4 + c4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        acc.x += v[k].x; acc.y += v[k].y; acc.z += v[k].z; acc.w += v[k].w;
        dy_r[(size_t)(r + 8 * k) * cols4 + c4] = make_float4(round_tf32_rna(v[k].x), round_tf32_rna(v[k].y), round_tf32_rna(v[k].z), round_tf32_rna(v[k].w));
      }
    }
    for (; r < rows; r += 8) {
      const float4 v = dy[(size_t)r * cols4 + c4];
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      dy_r[(size_t)r * cols4 + c4] = make_float4(round_tf32_rna(v.x), round_tf32_rna(v.y), round_tf32_rna(v.z), round_tf32_rna(v.w));
    }
  }
  part[rg][cx] = acc;
  __syncthreads();
  if (rg == 0 && c4 < cols4 && db) {
    float4 t = part[0][cx];
#pragma unroll
    for (int k = 1; k < 8; ++k) { const float4 u = part[k][cx]; t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w; }
    float4* out = reinterpret_cast<float4*>(db) + c4;
    if (accumulate) { const float4 o = *out; t.x += o.x; t.y += o.y; t.z += o.z; t.w += o.w; }
    *out = t;
  }
}
void ip_round_bias(cudaStream_t s, int rows, int cols, const float* dy, float* dy_r, float* db, bool accumulate) {
  const int cols4 = cols / 4;
  ip_round_bias_kernel<<<(cols4 + 31) / 32, 256, 0, s>>>(rows, cols4, (const float4*)dy, (float4*)dy_r, db, accumulate ? 1 : 0);
  count_launch();
}
void ip_bias_grad(cudaStream_t s, int rows, int cols, const float* dy, float* db, bool accumulate) {
  ip_bias_grad_kernel<<<(cols + 127) / 128, 128, 0, s>>>(rows, cols, dy, db, accumulate ? 1 : 0);
  count_launch();
}

// ------------------------------------------------------------------------------------------
// Fused SGD update — reference: per learnable blob cublasSaxpy (L2) + cublasSscal + cublasSaxpy
// (axpby) + cudaMemcpy (history -> diff) + cublasSaxpy (w -= diff): 6 passes, ~96 launches/iter for
// CaffeNet (caffe/src/caffe/solvers/sgd_solver.cpp:145-239, caffe/src/caffe/blob.cpp:154-177,
// caffe/src/caffe/util/math_functions.cu:94-98).  Here: one launch over the whole parameter arena,
// 3 reads + 3 writes per parameter (24 B), rounding order of the CPU path kept:
//   g <- g + decay*w ; h <- (mu*h) + (rate*g) ; g <- h ; w <- w - h     (mul and add rounded separately)
// Blobs start on 1024-element boundaries in the arena, so chunk_table[i/1024] gives the blob's
// (lr_mult, local_decay); local_rate = rate * lr_mult is formed in fp32 as in the reference.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float round_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
__global__ void sgd_update_kernel(size_t n4, float4* __restrict__ w, float4* __restrict__ g, float4* __restrict__ h, float4* __restrict__ shadow,
                                  const SgdChunk* __restrict__ table, float rate, float mu, const float* __restrict__ rate_dev) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  if (rate_dev) rate = *rate_dev;                // this iteration's learning rate, published on the device (replayed graphs)
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    const SgdChunk ck = table[i >> 8];           // 1024 floats = 256 float4
    const float local_rate = __fmul_rn(rate, ck.lr_mult);   // sgd_solver.cpp:210-211
    float4 wv = w[i], gv = g[i], hv = h[i];
#define SN_SGD1(F)                                                        \
    {                                                                     \
      float gg = gv.F;                                                    \
      if (ck.decay != 0.f) gg = __fadd_rn(gg, __fmul_rn(ck.decay, wv.F)); \
      const float hh = __fadd_rn(__fmul_rn(mu, hv.F), __fmul_rn(local_rate, gg)); \
      hv.F = hh; gv.F = hh; wv.F = __fsub_rn(wv.F, hh);                   \
    }
    SN_SGD1(x) SN_SGD1(y) SN_SGD1(z) SN_SGD1(w)
#undef SN_SGD1
    w[i] = wv; g[i] = gv; h[i] = hv;
    // tf32-rounded copy for the TMA-fed tensor-core operands (saves a separate 8 B/parameter rounding pass per step)
    if (shadow) shadow[i] = make_float4(round_tf32(wv.x), round_tf32(wv.y), round_tf32(wv.z), round_tf32(wv.w));
  }
}
void sgd_update(cudaStream_t s, size_t n, float* w, float* g, float* h, float* shadow, const SgdChunk* chunk_table, float rate,
                float momentum, const float* rate_dev) {
  if (n % 1024) fatal("sgd_update: arena length %zu is not a multiple of 1024", n);
  sgd_update_kernel<<<grid_for(n / 4, 256), 256, 0, s>>>(n / 4, (float4*)w, (float4*)g, (float4*)h, (float4*)shadow, chunk_table, rate, momentum, rate_dev);
  count_launch();
}

// Per-iteration scalars that a replayed CUDA graph cannot take as kernel arguments (learning rate, dropout counters):
// one tiny launch OUTSIDE the graph writes them into device memory, the kernels inside read them from there.
__global__ void publish_step_scalars_kernel(StepScalars* dst, StepScalars v) {
  SN_GRID_DEP_LAUNCH_ALWAYS();
  SN_GRID_DEP_WAIT(); *dst = v; }
void publish_step_scalars(cudaStream_t s, StepScalars* dst, const StepScalars& v) {
  publish_step_scalars_kernel<<<1, 1, 0, s>>>(dst, v);
  count_launch();
}

// ------------------------------------------------------------------------------------------
// Weight averaging over N replicas resident on this device (used for the emulated multi-rank test
// and as the reduce stage of the multi-GPU exchange): sequential fp32 sum in replica order, then
// one IEEE division by (float)N.
// ------------------------------------------------------------------------------------------
__global__ void average_replicas_kernel(const float* const* __restrict__ reps, int nrep, size_t n, float* __restrict__ out) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  const float denom = (float)nrep;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    float acc = reps[0][i];
    for (int r = 1; r < nrep; ++r) acc = __fadd_rn(acc, reps[r][i]);
    out[i] = __fdiv_rn(acc, denom);
  }
}
void average_replicas(cudaStream_t s, const float* const* replicas_dev, int nrep, size_t n, float* out) {
  average_replicas_kernel<<<grid_for(n, 256), 256, 0, s>>>(replicas_dev, nrep, n, out);
  count_launch();
}

// ------------------------------------------------------------------------------------------
// Weight averaging over peer memory: the whole exchange of one round in ONE kernel per GPU
// (replaces collect / reduce(WeightCollection.add) / scalarDivide / broadcast / setWeights,
// CifarApp.scala:92-99,133-134, Net.scala:14-47,132-172).  Every GPU owns one slice of the weight
// arena.  It reads that slice from every replica through NVLink loads IN RANK ORDER, folds it
// sequentially in fp32 (((w0 + w1) + w2) + ...), divides once by (float)N and stores the result into
// the same slice of every replica's arena — reduce-scatter and all-gather in one pass, 2(N-1)/N·P bytes
// over NVLink per GPU instead of the (N-1)·P of an all-gather followed by a local fold.
//
// Ordering across GPUs is by flags in each GPU's own memory that the peers write:
//   ready[p] = epoch   "replica p finished its local iterations; its arena may be read and written"
//   done[p]  = epoch   "replica p has stored its slice into every arena"
// A block starts reading once all ready flags show this epoch (every block polls its OWN GPU's
// memory); the last block to finish publishes `done` to all peers and the kernel ends only after all
// peers' done flags arrived, so the next kernel on the stream sees the complete averaged arena.
// Slices are disjoint: rank r alone reads and writes slice r of any arena during the exchange.
// No grid-wide co-residency is needed.  A wait that exceeds `timeout_ns` sets *err and gives up
// (the result is then undefined, the host reports it) instead of hanging the GPU.
// Plain ld/st only: no TMA, no multicast touches another GPU's memory here.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float4 ld_peer_f4(const float4* p) {
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ float ld_peer_f(const float* p) {
  float v;
  asm volatile("ld.volatile.global.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ unsigned ld_flag(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_flag(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// flags layout per GPU (32 words): [0..7] ready, [8..15] done, [16] finished-block counter, [17] error,
// [20..21] entry time of the current round, [22..23] time all replicas were ready, then three u64 sums over the
// rounds so far: [24..25] ns spent waiting for the slowest replica, [26..27] ns from "all ready" to "all done"
// (the exchange proper), [28..29] rounds
__device__ bool wait_flags(const unsigned* mine, int base, int n, unsigned epoch, unsigned long long timeout_ns) {
  const unsigned long long t0 = global_ns();
  for (int p = 0; p < n; ++p) {
    while ((int)(ld_flag(mine + base + p) - epoch) < 0) {
      if (global_ns() - t0 > timeout_ns) return false;
      __nanosleep(200);
    }
  }
  return true;
}
template <int N>
__global__ void __launch_bounds__(512) average_peers_kernel(PeerTable t, int rank, size_t count, unsigned epoch,
                                                            unsigned long long timeout_ns) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  unsigned* mine = t.flags[rank];
  __shared__ int ok;
  if (threadIdx.x == 0) {
    unsigned long long* stamps = reinterpret_cast<unsigned long long*>(mine + 20);
    if (blockIdx.x == 0) {
      stamps[0] = global_ns();
      __threadfence_system();
      for (int p = 0; p < N; ++p) st_flag(t.flags[p] + rank, epoch);
    }
    ok = wait_flags(mine, 0, N, epoch, timeout_ns) ? 1 : 0;
    if (!ok) mine[17] = 1u;
    if (blockIdx.x == 0) stamps[1] = global_ns();
  }
  __syncthreads();
  if (ok) {
    // slice of this rank, in float4 units; the arena's tail (count % 4) is folded by rank N-1
    const size_t n4 = count / 4;
    const size_t per = (n4 + N - 1) / N;
    const size_t lo = per * (size_t)rank, hi = lo + per < n4 ? lo + per : n4;
    const float denom = (float)N;
    for (size_t i = lo + blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < hi; i += (size_t)gridDim.x * blockDim.x) {
      float4 v[N];
#pragma unroll
      for (int r = 0; r < N; ++r) v[r] = ld_peer_f4(reinterpret_cast<const float4*>(t.w[r]) + i);
      float4 a = v[0];
#pragma unroll
      for (int r = 1; r < N; ++r) {
        a.x = __fadd_rn(a.x, v[r].x); a.y = __fadd_rn(a.y, v[r].y);
        a.z = __fadd_rn(a.z, v[r].z); a.w = __fadd_rn(a.w, v[r].w);
      }
      a.x = __fdiv_rn(a.x, denom); a.y = __fdiv_rn(a.y, denom); a.z = __fdiv_rn(a.z, denom); a.w = __fdiv_rn(a.w, denom);
#pragma unroll
      for (int r = 0; r < N; ++r) reinterpret_cast<float4*>(t.w[r])[i] = a;
    }
    if (rank == N - 1 && blockIdx.x == 0) {
      for (size_t i = n4 * 4 + threadIdx.x; i < count; i += blockDim.x) {
        float a = ld_peer_f(t.w[0] + i);
        for (int r = 1; r < N; ++r) a = __fadd_rn(a, ld_peer_f(t.w[r] + i));
        a = __fdiv_rn(a, denom);
        for (int r = 0; r < N; ++r) t.w[r][i] = a;
      }
    }
  }
  // last block out tells the peers that this rank's slice is in place everywhere, then waits for theirs
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned prev = atomicAdd(mine + 16, 1u);
    if (prev == gridDim.x - 1) {
      mine[16] = 0u;                               // next call starts from zero
      __threadfence_system();
      for (int p = 0; p < N; ++p) st_flag(t.flags[p] + 8 + rank, epoch);
      if (!wait_flags(mine, 8, N, epoch, timeout_ns)) mine[17] = 1u;
      unsigned long long* stamps = reinterpret_cast<unsigned long long*>(mine + 20);
      const unsigned long long t2 = global_ns();
      stamps[2] += stamps[1] - stamps[0];
      stamps[3] += t2 - stamps[1];
      stamps[4] += 1ull;
    }
  }
}
void average_peers(cudaStream_t s, const PeerTable& t, int nranks, int rank, size_t count, unsigned epoch,
                   unsigned long long timeout_ns, int max_grid) {
  const size_t per4 = (count / 4 + nranks - 1) / nranks;
  int grid = (int)std::min<size_t>((per4 + 511) / 512 + 1, (size_t)148 * 4);
  if (max_grid > 0 && grid > max_grid) grid = max_grid;
  if (grid < 1) grid = 1;
  switch (nranks) {
#define SN_PEERS(NR) case NR: average_peers_kernel<NR><<<grid, 512, 0, s>>>(t, rank, count, epoch, timeout_ns); break;
    SN_PEERS(1) SN_PEERS(2) SN_PEERS(3) SN_PEERS(4) SN_PEERS(5) SN_PEERS(6) SN_PEERS(7) SN_PEERS(8)
#undef SN_PEERS
    default: fatal("average_peers: 1..8 replicas per box, got %d", nranks);
  }
  count_launch();
}


}  // namespace sn
