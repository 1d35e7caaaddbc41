// gemm_ops.cuh — operand accessors and epilogues shared by the two contraction back ends
// (gemm_ffma.cu: fp32 FFMA validation mode; gemm_tc.cu: TF32 tcgen05/TMEM mode).
//
// Every dense contraction on the hot path is written as   D[m][n] = sum_r  A(m, r) * B(n, r)
// where A and B are *accessors* over the Caffe-layout tensors (NCHW activations, (Cout,Cin/g,kh,kw)
// filters, (out,in) inner-product weights) — no im2col buffer is ever materialised.  This replaces
// the reference's per-image im2col + cublasSgemm loops (caffe/src/caffe/layers/base_conv_layer.cpp:243-376,
// caffe/src/caffe/util/im2col.cu:9-58, caffe/src/caffe/layers/inner_product_layer.cu:10-56).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace sn {

struct ConvGeom {
  int N, C, H, W;          // input
  int Cout, kh, kw, sh, sw, ph, pw, Ho, Wo, group;
  __host__ __device__ int Cg() const { return C / group; }
  __host__ __device__ int Og() const { return Cout / group; }
  __host__ __device__ int Kdim() const { return Cg() * kh * kw; }   // reduction length of fwd
};

// Programmatic dependent launch: inside a recorded iteration (runtime.cpp Solver::step) consecutive kernels are joined by
// programmatic edges, so a kernel's CTAs are scheduled — and run their prologue — while the previous kernel's last CTAs
// drain.  Every kernel executes this before its first access to global memory: it returns once the previous kernel of
// the stream has completed and its writes are visible (a no-op for a kernel launched the ordinary way).
#define SN_GRID_DEP_WAIT() asm volatile("griddepcontrol.wait;" ::: "memory")
// ... and this lets the NEXT kernel of a recorded iteration be scheduled as soon as every CTA of this one has started
// (instead of when they have all exited): its CTAs take free slots and sit in their own wait, so the launch latency
// between two kernels is off the critical path.  Only for launches that leave SMs free (fewer CTAs than the GPU has SMs —
// the small nets: cifar10_quick / full -4 % per iteration); measured on CaffeNet, whose kernels fill the GPU, early
// release of every kernel cost 1.8 % (3.73 -> 3.80 ms), so full-size launches release their successor when they exit.
#define SN_GRID_DEP_LAUNCH()                                                                              \
  do {                                                                                                    \
    if (gridDim.x * gridDim.y * gridDim.z < 148u) asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); \
  } while (0)

// Short helper kernels (slice reductions, rounding passes, bias-gradient partial sums ...: a few microseconds each) release
// their successor at entry whatever their grid size.
#define SN_GRID_DEP_LAUNCH_ALWAYS() asm volatile("griddepcontrol.launch_dependents;" ::: "memory")

// Division by a runtime constant via float reciprocal-free 32-bit magic (exact for the ranges used
// here: numerators < 2^31, divisors < 2^16).  Falls back to '/' on the host.
struct FastDiv {
  uint32_t d, mul, shr;
  __host__ FastDiv() : d(1), mul(0), shr(0) {}
  __host__ explicit FastDiv(uint32_t div) : d(div) {
    if (div == 1) { mul = 0; shr = 0; return; }
    uint32_t l = 0;
    while ((1u << l) < div) ++l;          // ceil(log2 d)
    uint64_t m = ((uint64_t(1) << (31 + l)) + div - 1) / div;   // works for n < 2^31
    mul = (uint32_t)m;
    shr = l - 1;
  }
  __device__ __forceinline__ uint32_t div(uint32_t n) const {
    return d == 1 ? n : (uint32_t)(__umulhi(n, mul) >> shr);
  }
  __device__ __forceinline__ void divmod(uint32_t n, uint32_t& q, uint32_t& r) const {
    q = div(n);
    r = n - q * d;
  }
};

// ---------------------------------------------------------------------------------------------
// Accessors: value(row, r).  `rows()` / `red()` give the logical extents (out-of-range => 0).
// ---------------------------------------------------------------------------------------------

// Dense row-major [row][r] ("K-major")
struct DenseK {
  const float* p; int ld, nrows, nred;
  __device__ __forceinline__ float operator()(int row, int r) const {
    return (row < nrows && r < nred) ? __ldg(p + (size_t)row * ld + r) : 0.f;
  }
};
// Dense [r][row] ("MN-major": the row index is the contiguous one)
struct DenseMN {
  const float* p; int ld, nrows, nred;
  __device__ __forceinline__ float operator()(int row, int r) const {
    return (row < nrows && r < nred) ? __ldg(p + (size_t)r * ld + row) : 0.f;
  }
};

// im2col view for the forward pass, rows = output pixels (n,ho,wo), r = (c,kh,kw) within group g.
// Reference semantics: im2col_cpu caffe/src/caffe/util/im2col.cpp:9-31 (zero padding).
struct Im2colRows {
  const float* x; ConvGeom g; int grp;
  FastDiv dHoWo, dWo, dKhKw, dKw;
  __device__ __forceinline__ float operator()(int row, int r) const {
    if (row >= g.N * g.Ho * g.Wo || r >= g.Kdim()) return 0.f;
    uint32_t n, pix, ho, wo, c, t, ky, kx;
    dHoWo.divmod(row, n, pix);
    dWo.divmod(pix, ho, wo);
    dKhKw.divmod(r, c, t);
    dKw.divmod(t, ky, kx);
    int hi = (int)ho * g.sh - g.ph + (int)ky, wi = (int)wo * g.sw - g.pw + (int)kx;
    if (hi < 0 || hi >= g.H || wi < 0 || wi >= g.W) return 0.f;
    return __ldg(x + (((size_t)n * g.C + grp * g.Cg() + c) * g.H + hi) * g.W + wi);
  }
};

// Backward-data view: rows = input pixels (n,h,w), r = (co,kh,kw) within group g; value = dY at the
// output position that used this input pixel with tap (kh,kw), or 0 (col2im semantics,
// caffe/src/caffe/util/im2col.cpp col2im_cpu).
struct DgradRows {
  const float* dy; ConvGeom g; int grp;
  FastDiv dHW, dW, dKhKw, dKw, dSh, dSw;
  __device__ __forceinline__ float operator()(int row, int r) const {
    if (row >= g.N * g.H * g.W || r >= g.Og() * g.kh * g.kw) return 0.f;
    uint32_t n, pix, h, w, co, t, ky, kx;
    dHW.divmod(row, n, pix);
    dW.divmod(pix, h, w);
    dKhKw.divmod(r, co, t);
    dKw.divmod(t, ky, kx);
    int hn = (int)h + g.ph - (int)ky, wn = (int)w + g.pw - (int)kx;
    if (hn < 0 || wn < 0) return 0.f;
    uint32_t ho, hr, wo, wr;
    dSh.divmod((uint32_t)hn, ho, hr);
    dSw.divmod((uint32_t)wn, wo, wr);
    if (hr != 0 || wr != 0 || (int)ho >= g.Ho || (int)wo >= g.Wo) return 0.f;
    return __ldg(dy + (((size_t)n * g.Cout + grp * g.Og() + co) * g.Ho + ho) * g.Wo + wo);
  }
};

// Filter view for backward-data: rows = input channel c within group, r = (co,kh,kw)
struct FilterDgrad {
  const float* w; ConvGeom g; int grp;
  FastDiv dKhKw;
  __device__ __forceinline__ float operator()(int row, int r) const {
    const int khkw = g.kh * g.kw;
    if (row >= g.Cg() || r >= g.Og() * khkw) return 0.f;
    uint32_t co, t;
    dKhKw.divmod(r, co, t);
    return __ldg(w + ((size_t)(grp * g.Og() + co) * g.Cg() + row) * khkw + t);
  }
};

// Backward-filter views: reduction index r = (n,ho,wo) flattened.
struct DyRows {           // rows = output channel co within group
  const float* dy; ConvGeom g; int grp;
  FastDiv dHoWo;
  __device__ __forceinline__ float operator()(int row, int r) const {
    if (row >= g.Og() || r >= g.N * g.Ho * g.Wo) return 0.f;
    uint32_t n, pix;
    dHoWo.divmod(r, n, pix);
    return __ldg(dy + ((size_t)n * g.Cout + grp * g.Og() + row) * (g.Ho * g.Wo) + pix);
  }
};
struct Im2colCols {       // rows = (c,kh,kw) within group, r = (n,ho,wo)
  const float* x; ConvGeom g; int grp;
  FastDiv dHoWo, dWo, dKhKw, dKw;
  __device__ __forceinline__ float operator()(int row, int r) const {
    Im2colRows v{x, g, grp, dHoWo, dWo, dKhKw, dKw};
    return v(r, row);
  }
};

// ---------------------------------------------------------------------------------------------
// Epilogues: store(m, n, acc)
// ---------------------------------------------------------------------------------------------
// conv forward: m = output pixel, n = co within group.  Y[n_img][grp*Og+co][pix] = acc + bias (+ReLU)
// (bias: base_conv_layer.cpp:263-266; fused ReLU only when the next layer is an in-place ReLU).
struct EpiConvFwd {
  float* y; const float* bias; ConvGeom g; int grp; int relu;
  FastDiv dHoWo;
  __device__ __forceinline__ void operator()(int m, int n, float acc) const {
    if (m >= g.N * g.Ho * g.Wo || n >= g.Og()) return;
    uint32_t img, pix;
    dHoWo.divmod(m, img, pix);
    const int co = grp * g.Og() + n;
    float v = bias ? acc + __ldg(bias + co) : acc;
    if (relu) v = fmaxf(v, 0.f);
    y[((size_t)img * g.Cout + co) * (g.Ho * g.Wo) + pix] = v;
  }
};
// conv backward-data: m = input pixel, n = c within group.  dX[n_img][grp*Cg+c][pix] = acc
struct EpiConvDgrad {
  float* dx; ConvGeom g; int grp;
  FastDiv dHW;
  __device__ __forceinline__ void operator()(int m, int n, float acc) const {
    if (m >= g.N * g.H * g.W || n >= g.Cg()) return;
    uint32_t img, pix;
    dHW.divmod(m, img, pix);
    dx[((size_t)img * g.C + grp * g.Cg() + n) * (g.H * g.W) + pix] = acc;
  }
};
// row-major store with optional accumulate (weight gradients: beta = 1, base_conv_layer.cpp:295-298,
// inner_product_layer.cpp:100-103)
struct EpiRowMajor {
  float* p; int ld, M, N; int accumulate;
  __device__ __forceinline__ void operator()(int m, int n, float acc) const {
    if (m >= M || n >= N) return;
    float* q = p + (size_t)m * ld + n;
    *q = accumulate ? *q + acc : acc;
  }
};
// transposed store: D[m][n] -> p[n*ld + m] = acc + bias[m] (+ReLU).  Inner-product forward
// (Y[batch][out], m = out) and backward-data (dX[batch][in], m = in).
struct EpiTransposed {
  float* p; int ld, M, N; const float* bias; int relu;
  __device__ __forceinline__ void operator()(int m, int n, float acc) const {
    if (m >= M || n >= N) return;
    float v = bias ? acc + __ldg(bias + m) : acc;
    if (relu) v = fmaxf(v, 0.f);
    p[(size_t)n * ld + m] = v;
  }
};

// host-side constructors
inline Im2colRows make_im2col_rows(const float* x, const ConvGeom& g, int grp) {
  return Im2colRows{x, g, grp, FastDiv(g.Ho * g.Wo), FastDiv(g.Wo), FastDiv(g.kh * g.kw), FastDiv(g.kw)};
}
inline Im2colCols make_im2col_cols(const float* x, const ConvGeom& g, int grp) {
  return Im2colCols{x, g, grp, FastDiv(g.Ho * g.Wo), FastDiv(g.Wo), FastDiv(g.kh * g.kw), FastDiv(g.kw)};
}
inline DgradRows make_dgrad_rows(const float* dy, const ConvGeom& g, int grp) {
  return DgradRows{dy, g, grp, FastDiv(g.H * g.W), FastDiv(g.W), FastDiv(g.kh * g.kw), FastDiv(g.kw),
                   FastDiv(g.sh), FastDiv(g.sw)};
}
inline FilterDgrad make_filter_dgrad(const float* w, const ConvGeom& g, int grp) {
  return FilterDgrad{w, g, grp, FastDiv(g.kh * g.kw)};
}
inline DyRows make_dy_rows(const float* dy, const ConvGeom& g, int grp) {
  return DyRows{dy, g, grp, FastDiv(g.Ho * g.Wo)};
}

}  // namespace sn
