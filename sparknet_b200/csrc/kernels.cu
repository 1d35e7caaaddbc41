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
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    o[i] = a[i] + b[i];
}
void split_add(cudaStream_t s, size_t n, const float* a, const float* b, float* out) {
  split_add_kernel<<<grid_for(n, 256), 256, 0, s>>>(n, a, b, out);
  count_launch();
}
__global__ void axpy_kernel(size_t n, float a, const float* __restrict__ x, float* __restrict__ y) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    y[i] = __fadd_rn(y[i], __fmul_rn(a, x[i]));
}
void axpy(cudaStream_t s, size_t n, float a, const float* x, float* y) {
  axpy_kernel<<<grid_for(n, 256), 256, 0, s>>>(n, a, x, y);
  count_launch();
}

// ------------------------------------------------------------------------------------------
// ReLU — reference ReLUForward/ReLUBackward caffe/src/caffe/layers/relu_layer.cu:9-14,35-41
// (CPU form relu_layer.cpp:16-17,32-33).  fwd 8 B/elt, bwd 12 B/elt.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float relu1(float x, float slope) { return fmaxf(x, 0.f) + slope * fminf(x, 0.f); }
__global__ void relu_fwd_kernel(size_t n4, size_t n, const float4* __restrict__ x4, float4* __restrict__ y4,
                                const float* __restrict__ x, float* __restrict__ y, float slope) {
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
__global__ void pool_max_fwd_kernel(PoolGeom g, size_t total, const float* __restrict__ x, float* __restrict__ y,
                                    int* __restrict__ mask) {
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int pw = (int)(idx % g.PW), ph = (int)((idx / g.PW) % g.PH);
    const size_t plane = idx / ((size_t)g.PW * g.PH);
    int hs = ph * g.sh - g.ph, ws = pw * g.sw - g.pw;
    const int he = min(hs + g.kh, g.H), we = min(ws + g.kw, g.W);
    hs = max(hs, 0); ws = max(ws, 0);
    const float* xp = x + plane * (size_t)(g.H * g.W);
    float best = -FLT_MAX;
    int bi = -1;
    for (int h = hs; h < he; ++h)
      for (int w = ws; w < we; ++w) {
        const float v = __ldg(xp + h * g.W + w);
        if (v > best) { best = v; bi = h * g.W + w; }
      }
    y[idx] = best;
    mask[idx] = bi;
  }
}
// Plane-per-CTA form: the input plane is staged in shared memory with linear, coalesced loads; every pooled
// element then scans its window there (same scan order and strict '>' as above).  Global traffic is the algorithmic
// 4*in + 8*out bytes; the strided window reads never leave the SM.
__global__ void __launch_bounds__(256) pool_max_fwd_plane_kernel(PoolGeom g, FastDiv dPW, const float* __restrict__ x, float* __restrict__ y,
                                                                 int* __restrict__ mask) {
  extern __shared__ float sm_plane[];
  const int HW = g.H * g.W, P = g.PH * g.PW;
  for (size_t plane = blockIdx.x; plane < (size_t)g.N * g.C; plane += gridDim.x) {
    const float* xp = x + plane * (size_t)HW;
    for (int i = threadIdx.x; i < HW; i += blockDim.x) sm_plane[i] = xp[i];
    __syncthreads();
    for (int j = threadIdx.x; j < P; j += blockDim.x) {
      uint32_t ph, pw;
      dPW.divmod((uint32_t)j, ph, pw);
      int hs = (int)ph * g.sh - g.ph, ws = (int)pw * g.sw - g.pw;
      const int he = min(hs + g.kh, g.H), we = min(ws + g.kw, g.W);
      hs = max(hs, 0); ws = max(ws, 0);
      float best = -FLT_MAX;
      int bi = -1;
      for (int h = hs; h < he; ++h)
        for (int w = ws; w < we; ++w) {
          const float v = sm_plane[h * g.W + w];
          if (v > best) { best = v; bi = h * g.W + w; }
        }
      y[plane * P + j] = best;
      mask[plane * P + j] = bi;
    }
    __syncthreads();
  }
}
// Thread-per-output form for the window shapes of the four nets (compile-time kernel size and stride, no padding):
// 32-bit index arithmetic with multiply-shift division, all window loads issued before the compare chain.  Window
// elements are read straight from global memory: neighbouring outputs share them through the L1, HBM sees them once.
template <int KH, int KW, int SH, int SW>
__global__ void __launch_bounds__(256) pool_max_fwd_fixed_kernel(PoolGeom g, uint32_t total, FastDiv dPW, FastDiv dPH, const float* __restrict__ x,
                                                                 float* __restrict__ y, int* __restrict__ mask) {
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
  const size_t total = (size_t)g.N * g.C * g.PH * g.PW, smem = (size_t)g.H * g.W * sizeof(float);
  const bool fixed = g.ph == 0 && g.pw == 0 && total < (1ull << 31) && (size_t)g.N * g.C * g.H * g.W < (1ull << 31);
  if (fixed && g.kh == 3 && g.kw == 3 && g.sh == 2 && g.sw == 2) {
    pool_max_fwd_fixed_kernel<3, 3, 2, 2><<<grid_for(total, 256, 32), 256, 0, s>>>(g, (uint32_t)total, FastDiv((uint32_t)g.PW), FastDiv((uint32_t)g.PH), x, y, mask);
  } else if (fixed && g.kh == 2 && g.kw == 2 && g.sh == 2 && g.sw == 2) {
    pool_max_fwd_fixed_kernel<2, 2, 2, 2><<<grid_for(total, 256, 32), 256, 0, s>>>(g, (uint32_t)total, FastDiv((uint32_t)g.PW), FastDiv((uint32_t)g.PH), x, y, mask);
  } else if (smem <= 40 * 1024 && g.H * g.W >= 64) {
    const size_t planes = (size_t)g.N * g.C;
    const int blocks = (int)std::min<size_t>(planes, (size_t)kSMs * 16);
    pool_max_fwd_plane_kernel<<<blocks, 256, smem, s>>>(g, FastDiv((uint32_t)g.PW), x, y, mask);
  } else {
    pool_max_fwd_kernel<<<grid_for(total, 256, 16), 256, 0, s>>>(g, total, x, y, mask);
  }
  count_launch();
}
// gather form: every input element sums the top diffs of the windows whose argmax it is
__global__ void pool_max_bwd_kernel(PoolGeom g, size_t total, const float* __restrict__ dy, const int* __restrict__ mask,
                                    float* __restrict__ dx) {
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int w = (int)(idx % g.W), h = (int)((idx / g.W) % g.H);
    const size_t plane = idx / ((size_t)g.W * g.H);
    const int phs = (h + g.ph < g.kh) ? 0 : (h + g.ph - g.kh) / g.sh + 1;
    const int phe = min((h + g.ph) / g.sh + 1, g.PH);
    const int pws = (w + g.pw < g.kw) ? 0 : (w + g.pw - g.kw) / g.sw + 1;
    const int pwe = min((w + g.pw) / g.sw + 1, g.PW);
    const size_t off = plane * (size_t)(g.PH * g.PW);
    const int me = h * g.W + w;
    float acc = 0.f;
    for (int ph = phs; ph < phe; ++ph)
      for (int pw = pws; pw < pwe; ++pw)
        if (__ldg(mask + off + ph * g.PW + pw) == me) acc += __ldg(dy + off + ph * g.PW + pw);
    dx[idx] = acc;
  }
}
// Plane-per-CTA form: the pooled plane's (dy, mask) pair is staged in shared memory with coalesced loads, the
// window range of every input row / column is tabulated once per CTA (no integer division per element), and the
// plane is then written linearly: every input element checks the <= ceil(k/s)^2 windows that cover it, in the
// reference's (ph, pw) order.  Global traffic is exactly the algorithmic 8*out + 4*in bytes.
// `gate` (nullable) = the pooled OUTPUT values: when the pooling input is the output of an in-place ReLU whose
// backward would run next (relu_layer.cpp:26-36: dx = dy * (y > 0)), that test is applied here — the element that
// won a window is > 0 exactly when the pooled value is — and the separate 12 B/element ReLU pass disappears.
__global__ void __launch_bounds__(256) pool_max_bwd_plane_kernel(PoolGeom g, FastDiv dW, const float* __restrict__ dy,
                                                                 const int* __restrict__ mask, const float* __restrict__ gate,
                                                                 float* __restrict__ dx) {
  extern __shared__ float sm_pool[];
  const int P = g.PH * g.PW, HW = g.H * g.W;
  float* sdy = sm_pool;
  int* smask = (int*)(sm_pool + P);
  int* rowr = smask + P;                           // packed (first window | one-past-last window << 16) per input row
  int* colr = rowr + g.H;
  for (int h = threadIdx.x; h < g.H; h += blockDim.x) {
    const int phs = (h + g.ph < g.kh) ? 0 : (h + g.ph - g.kh) / g.sh + 1, phe = min((h + g.ph) / g.sh + 1, g.PH);
    rowr[h] = phs | (phe << 16);
  }
  for (int w = threadIdx.x; w < g.W; w += blockDim.x) {
    const int pws = (w + g.pw < g.kw) ? 0 : (w + g.pw - g.kw) / g.sw + 1, pwe = min((w + g.pw) / g.sw + 1, g.PW);
    colr[w] = pws | (pwe << 16);
  }
  for (size_t plane = blockIdx.x; plane < (size_t)g.N * g.C; plane += gridDim.x) {
    __syncthreads();
    for (int i = threadIdx.x; i < P; i += blockDim.x) {
      float v = dy[plane * P + i];
      if (gate && !(gate[plane * P + i] > 0.f)) v = 0.f;
      sdy[i] = v;
      smask[i] = mask[plane * P + i];
    }
    __syncthreads();
    float* dxp = dx + plane * (size_t)HW;
    for (int me = threadIdx.x; me < HW; me += blockDim.x) {
      uint32_t h, w;
      dW.divmod((uint32_t)me, h, w);
      const int rr = rowr[h], cr = colr[w];
      const int phs = rr & 0xffff, phe = rr >> 16, pws = cr & 0xffff, pwe = cr >> 16;
      float acc = 0.f;
      for (int ph = phs; ph < phe; ++ph)
        for (int pw = pws; pw < pwe; ++pw)
          if (smask[ph * g.PW + pw] == me) acc += sdy[ph * g.PW + pw];
      dxp[me] = acc;
    }
  }
}
// Thread-per-input-element gather for the same fixed window shapes: at most ceil(K/S)^2 = 4 candidate windows, tested
// in the reference's (ph, pw) order; 32-bit index arithmetic, divisions by compile-time strides.
template <int KH, int KW, int SH, int SW>
__global__ void __launch_bounds__(256) pool_max_bwd_fixed_kernel(PoolGeom g, uint32_t total, FastDiv dW, FastDiv dH, const float* __restrict__ dy,
                                                                 const int* __restrict__ mask, const float* __restrict__ gate,
                                                                 float* __restrict__ dx) {
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
__global__ void pool_gate_kernel(size_t n, const float* __restrict__ dy, const float* __restrict__ gate, float* __restrict__ out) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    out[i] = gate[i] > 0.f ? dy[i] : 0.f;
}
void pool_max_bwd(cudaStream_t s, const PoolGeom& g, const float* dy, const int* mask, float* dx, const float* gate) {
  const size_t smem = (size_t)g.PH * g.PW * 8 + (size_t)(g.H + g.W) * 4;
  const size_t total_in = (size_t)g.N * g.C * g.H * g.W;
  const bool fixed = g.ph == 0 && g.pw == 0 && total_in < (1ull << 31);
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
  } else if (smem <= 40 * 1024) {
    const size_t planes = (size_t)g.N * g.C;
    const int blocks = (int)std::min<size_t>(planes, (size_t)kSMs * 16);
    pool_max_bwd_plane_kernel<<<blocks, 256, smem, s>>>(g, FastDiv((uint32_t)g.W), dy, mask, gate, dx);
  } else {
    const size_t total = (size_t)g.N * g.C * g.H * g.W;
    if (gate) {                                    // large planes: gate into the workspace first
      const size_t n = (size_t)g.N * g.C * g.PH * g.PW;
      float* tmp = get_workspace(n);
      pool_gate_kernel<<<grid_for(n, 256), 256, 0, s>>>(n, dy, gate, tmp);
      count_launch();
      dy = tmp;
    }
    pool_max_bwd_kernel<<<grid_for(total, 256, 16), 256, 0, s>>>(g, total, dy, mask, dx);
  }
  count_launch();
}
__global__ void pool_ave_fwd_kernel(PoolGeom g, size_t total, const float* __restrict__ x, float* __restrict__ y) {
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int pw = (int)(idx % g.PW), ph = (int)((idx / g.PW) % g.PH);
    const size_t plane = idx / ((size_t)g.PW * g.PH);
    int hs = ph * g.sh - g.ph, ws = pw * g.sw - g.pw;
    int he = min(hs + g.kh, g.H + g.ph), we = min(ws + g.kw, g.W + g.pw);
    const int pool_size = (he - hs) * (we - ws);
    hs = max(hs, 0); ws = max(ws, 0); he = min(he, g.H); we = min(we, g.W);
    const float* xp = x + plane * (size_t)(g.H * g.W);
    float acc = 0.f;
    for (int h = hs; h < he; ++h)
      for (int w = ws; w < we; ++w) acc += __ldg(xp + h * g.W + w);
    y[idx] = __fdiv_rn(acc, (float)pool_size);
  }
}
void pool_ave_fwd(cudaStream_t s, const PoolGeom& g, const float* x, float* y) {
  const size_t total = (size_t)g.N * g.C * g.PH * g.PW;
  pool_ave_fwd_kernel<<<grid_for(total, 256, 16), 256, 0, s>>>(g, total, x, y);
  count_launch();
}
__global__ void pool_ave_bwd_kernel(PoolGeom g, size_t total, const float* __restrict__ dy, float* __restrict__ dx) {
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int w = (int)(idx % g.W) + g.pw, h = (int)((idx / g.W) % g.H) + g.ph;
    const size_t plane = idx / ((size_t)g.W * g.H);
    const int phs = (h < g.kh) ? 0 : (h - g.kh) / g.sh + 1, phe = min(h / g.sh + 1, g.PH);
    const int pws = (w < g.kw) ? 0 : (w - g.kw) / g.sw + 1, pwe = min(w / g.sw + 1, g.PW);
    const size_t off = plane * (size_t)(g.PH * g.PW);
    float acc = 0.f;
    for (int ph = phs; ph < phe; ++ph)
      for (int pw = pws; pw < pwe; ++pw) {
        const int hs = ph * g.sh - g.ph, ws = pw * g.sw - g.pw;
        const int he = min(hs + g.kh, g.H + g.ph), we = min(ws + g.kw, g.W + g.pw);
        acc += __fdiv_rn(__ldg(dy + off + ph * g.PW + pw), (float)((he - hs) * (we - ws)));
      }
    dx[idx] = acc;
  }
}
void pool_ave_bwd(cudaStream_t s, const PoolGeom& g, const float* dy, float* dx) {
  const size_t total = (size_t)g.N * g.C * g.H * g.W;
  pool_ave_bwd_kernel<<<grid_for(total, 256, 16), 256, 0, s>>>(g, total, dy, dx);
  count_launch();
}

// ------------------------------------------------------------------------------------------
// LRN across channels — reference LRNFillScale / LRNComputeOutput / LRNComputeDiff
// caffe/src/caffe/layers/lrn_layer.cu:9-52,72-77,122-177 (CPU: lrn_layer.cpp:108-152,180-232).
// One thread per (n,h,w) walks the channels with a running window; consecutive threads are
// consecutive w, so every step is a coalesced row access.  fwd 12 B/elt (x, scale, y), bwd 20 B/elt.
// ------------------------------------------------------------------------------------------
__global__ void lrn_across_fwd_kernel(LrnGeom g, int nthreads, const float* __restrict__ x, float* __restrict__ y,
                                      float* __restrict__ scale) {
  const int HW = g.H * g.W;
  const int pre = (g.size - 1) / 2, post = g.size - pre - 1;
  const float aos = g.alpha / g.size;
  const float nbeta = -g.beta;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < nthreads; idx += gridDim.x * blockDim.x) {
    const int pix = idx % HW, n = idx / HW;
    const size_t off = (size_t)n * g.C * HW + pix;
    const float* xi = x + off;
    float* yi = y + off;
    float* si = scale + off;
    int head = 0;
    float accum = 0.f;
    while (head < post && head < g.C) {
      const float v = xi[(size_t)head * HW];
      accum += v * v;
      ++head;
    }
    while (head < g.C) {
      const float v = xi[(size_t)head * HW];
      accum += v * v;
      if (head - g.size >= 0) {
        const float t = xi[(size_t)(head - g.size) * HW];
        accum -= t * t;
      }
      const int c = head - post;
      const float sc = g.k + accum * aos;
      si[(size_t)c * HW] = sc;
      yi[(size_t)c * HW] = xi[(size_t)c * HW] * powf(sc, nbeta);
      ++head;
    }
    while (head < g.C + post) {
      if (head - g.size >= 0) {
        const float t = xi[(size_t)(head - g.size) * HW];
        accum -= t * t;
      }
      const int c = head - post;
      if (c >= 0 && c < g.C) {
        const float sc = g.k + accum * aos;
        si[(size_t)c * HW] = sc;
        yi[(size_t)c * HW] = xi[(size_t)c * HW] * powf(sc, nbeta);
      }
      ++head;
    }
  }
}
// Chunked form (local_size 3 or 5): one thread owns CH consecutive channels of one pixel, loads them plus the halo in
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
__global__ void __launch_bounds__(128) lrn_across_fwd_chunk_kernel(LrnGeom g, int chunks, const float* __restrict__ x, float* __restrict__ y) {
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
#pragma unroll
    for (int j = 0; j < CH; ++j) {
      if (c0 + j < g.C) {
        float acc = 0.f;
#pragma unroll
        for (int t = 0; t < SIZE; ++t) acc += v[j + t] * v[j + t];
        const float sc = g.k + acc * aos;
        yp[(size_t)(c0 + j) * HW] = v[j + PRE] * lrn_pow_neg(sc, g.beta, is075);
      }
    }
  }
}
void lrn_across_fwd(cudaStream_t s, const LrnGeom& g, const float* x, float* y, float* scale) {
  if (g.size == 5 || g.size == 3) {
    const int CHn = lrn_chunk(), chunks = (g.C + CHn - 1) / CHn;
    const size_t total = (size_t)g.N * chunks * g.H * g.W;
#define SN_LRN_F(SZ, CHV) lrn_across_fwd_chunk_kernel<SZ, CHV><<<grid_for(total, 128, 32), 128, 0, s>>>(g, chunks, x, y)
    if (g.size == 5) { if (CHn == 8) SN_LRN_F(5, 8); else if (CHn == 32) SN_LRN_F(5, 32); else SN_LRN_F(5, 16); }
    else { if (CHn == 8) SN_LRN_F(3, 8); else if (CHn == 32) SN_LRN_F(3, 32); else SN_LRN_F(3, 16); }
#undef SN_LRN_F
    count_launch();
    return;
  }
  const int nthreads = g.N * g.H * g.W;
  lrn_across_fwd_kernel<<<grid_for(nthreads, 128, 16), 128, 0, s>>>(g, nthreads, x, y, scale);
  count_launch();
}
__global__ void lrn_across_bwd_kernel(LrnGeom g, int nthreads, const float* __restrict__ x, const float* __restrict__ y,
                                      const float* __restrict__ scale, const float* __restrict__ dy,
                                      float* __restrict__ dx) {
  const int HW = g.H * g.W;
  const int pre = g.size - (g.size + 1) / 2, post = g.size - pre - 1;
  const float cache_ratio = 2.f * g.alpha * g.beta / g.size;
  const float nbeta = -g.beta;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < nthreads; idx += gridDim.x * blockDim.x) {
    const int pix = idx % HW, n = idx / HW;
    const size_t off = (size_t)n * g.C * HW + pix;
    const float *xi = x + off, *yi = y + off, *si = scale + off, *gi = dy + off;
    float* di = dx + off;
    int head = 0;
    float accum = 0.f;
    while (head < post && head < g.C) {
      accum += gi[(size_t)head * HW] * yi[(size_t)head * HW] / si[(size_t)head * HW];
      ++head;
    }
    while (head < g.C + post) {
      if (head < g.C) accum += gi[(size_t)head * HW] * yi[(size_t)head * HW] / si[(size_t)head * HW];
      if (head - g.size >= 0)
        accum -= gi[(size_t)(head - g.size) * HW] * yi[(size_t)(head - g.size) * HW] / si[(size_t)(head - g.size) * HW];
      const int c = head - post;
      if (c >= 0 && c < g.C)
        di[(size_t)c * HW] = gi[(size_t)c * HW] * powf(si[(size_t)c * HW], nbeta) - cache_ratio * xi[(size_t)c * HW] * accum;
      ++head;
    }
  }
}
// Chunked backward (local_size 3 or 5), reference LRNComputeDiff (lrn_layer.cu:122-177) / CrossChannelBackward_cpu
// (lrn_layer.cpp:180-232):   dx_c = dy_c * scale_c^-beta - (2 alpha beta / n) * x_c * sum_{|j-c| <= pre} dy_j * y_j / scale_j
// with scale and y recomputed from x (halo 2*pre).  Reads x and dy, writes dx: 12 B per element.
template <int SIZE, int CH>
__global__ void __launch_bounds__(128) lrn_across_bwd_chunk_kernel(LrnGeom g, int chunks, const float* __restrict__ x,
                                                                   const float* __restrict__ dy, float* __restrict__ dx) {
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
void lrn_across_bwd(cudaStream_t s, const LrnGeom& g, const float* x, const float* y, const float* scale,
                    const float* dy, float* dx) {
  if (g.size == 5 || g.size == 3) {
    const int CHn = lrn_chunk(), chunks = (g.C + CHn - 1) / CHn;
    const size_t total = (size_t)g.N * chunks * g.H * g.W;
#define SN_LRN_B(SZ, CHV) lrn_across_bwd_chunk_kernel<SZ, CHV><<<grid_for(total, 128, 32), 128, 0, s>>>(g, chunks, x, dy, dx)
    if (g.size == 5) { if (CHn == 8) SN_LRN_B(5, 8); else if (CHn == 32) SN_LRN_B(5, 32); else SN_LRN_B(5, 16); }
    else { if (CHn == 8) SN_LRN_B(3, 8); else if (CHn == 32) SN_LRN_B(3, 32); else SN_LRN_B(3, 16); }
#undef SN_LRN_B
    count_launch();
    return;
  }
  const int nthreads = g.N * g.H * g.W;
  lrn_across_bwd_kernel<<<grid_for(nthreads, 128, 16), 128, 0, s>>>(g, nthreads, x, y, scale, dy, dx);
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
__device__ __forceinline__ float round_tf32_rna(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
__global__ void __launch_bounds__(256) ip_round_bias_kernel(int rows, int cols4, const float4* __restrict__ dy, float4* __restrict__ dy_r,
                                                            float* __restrict__ db, int accumulate) {
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
__global__ void publish_step_scalars_kernel(StepScalars* dst, StepScalars v) { *dst = v; }
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
