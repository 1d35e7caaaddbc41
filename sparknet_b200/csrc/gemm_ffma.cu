// gemm_ffma.cu — fp32 FFMA "validation mode" contraction (north_star: "an FFMA fp32 validation
// mode must match within 1e-5").  SIMT, shared-memory tiled, register-blocked 8x8, no tensor
// cores.  The operands are the accessors of gemm_ops.cuh, so convolution / inner-product forward,
// backward-data and backward-filter all run as implicit GEMMs on the Caffe-layout tensors.
//
// It is the arbiter for the TF32 tcgen05 kernels (gemm_tc.cu) and the bring-up path; it is not the
// performance path.
#include "gemm_ops.cuh"
#include "kernels.hpp"

namespace sn {

constexpr int BM = 128, BN = 128, BK = 8, TM = 8, TN = 8, NT = 256;

// SPLIT = false: the epilogue is applied directly.  SPLIT = true (split along the reduction): each
// z-slice writes its partial tile to `ws[z][M][N]`; splitk_reduce_kernel then sums the slices in
// ascending z order (deterministic) and applies the epilogue.
template <bool SPLIT, class A, class B, class E>
__global__ void __launch_bounds__(NT) gemm_ffma_kernel(A a, B b, E epi, int M, int N, int R, int r_split_len,
                                                       float* __restrict__ ws) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  __shared__ float As[2][BK][BM + 4];
  __shared__ float Bs[2][BK][BN + 4];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  const int r_begin = blockIdx.z * r_split_len;
  const int r_end = min(R, r_begin + r_split_len);
  // loader mapping: thread owns row (tid % 128) for k-slots (tid / 128) + {0,2,4,6}
  const int lrow = tid & 127, lk = tid >> 7;
  // compute mapping: 16 x 16 threads, each 8x8 (strided by 16 to keep smem reads conflict-free)
  const int tx = tid & 15, ty = tid >> 4;
  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  float ra[4], rb[4];
  auto gload = [&](int r0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int k = lk + 2 * i;
      const int r = r0 + k;
      ra[i] = (r < r_end) ? a(m0 + lrow, r) : 0.f;
      rb[i] = (r < r_end) ? b(n0 + lrow, r) : 0.f;
    }
  };
  auto sstore = [&](int buf) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      As[buf][lk + 2 * i][lrow] = ra[i];
      Bs[buf][lk + 2 * i][lrow] = rb[i];
    }
  };
  int buf = 0;
  if (r_begin < r_end) {
    gload(r_begin);
    sstore(0);
  }
  __syncthreads();
  for (int r0 = r_begin; r0 < r_end; r0 += BK) {
    const bool more = r0 + BK < r_end;
    if (more) gload(r0 + BK);
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float av[TM], bv[TN];
#pragma unroll
      for (int i = 0; i < TM; ++i) av[i] = As[buf][k][ty + 16 * i];
#pragma unroll
      for (int j = 0; j < TN; ++j) bv[j] = Bs[buf][k][tx + 16 * j];
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    if (more) sstore(buf ^ 1);
    __syncthreads();
    buf ^= 1;
  }
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      const int m = m0 + ty + 16 * i, n = n0 + tx + 16 * j;
      if (SPLIT) {
        if (m < M && n < N) ws[((size_t)blockIdx.z * M + m) * N + n] = acc[i][j];
      } else {
        epi(m, n, acc[i][j]);
      }
    }
}

template <class E>
__global__ void splitk_reduce_kernel(const float* __restrict__ ws, int splits, int M, int N, E epi) {
  SN_GRID_DEP_LAUNCH();
  SN_GRID_DEP_WAIT();
  const size_t total = (size_t)M * N;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    float acc = ws[i];
    for (int z = 1; z < splits; ++z) acc += ws[(size_t)z * total + i];
    epi((int)(i / N), (int)(i % N), acc);
  }
}

// Number of reduction slices: as many CTAs as fit in two waves of 148 SMs (never a partial third wave),
// each slice >= 512 long.
int splitk_slices(int tiles, int R, int ctas_per_sm) {
  if (tiles >= 148 || R < 2048) return 1;
  int want = (ctas_per_sm * 148) / tiles;
  int maxs = R / 512;
  int s = want < maxs ? want : maxs;
  return s < 1 ? 1 : s;
}

template <class A, class B, class E>
static void launch(cudaStream_t s, A a, B b, E e, int M, int N, int R) {
  dim3 grid((M + BM - 1) / BM, (N + BN - 1) / BN, 1);
  const int splits = splitk_slices((int)(grid.x * grid.y), R);
  if (splits == 1) {
    gemm_ffma_kernel<false><<<grid, NT, 0, s>>>(a, b, e, M, N, R, R, nullptr);
    count_launch();
    return;
  }
  int len = (R + splits - 1) / splits;
  len = (len + BK - 1) / BK * BK;
  grid.z = (R + len - 1) / len;
  float* ws = get_workspace((size_t)grid.z * M * N);
  gemm_ffma_kernel<true><<<grid, NT, 0, s>>>(a, b, e, M, N, R, len, ws);
  const size_t total = (size_t)M * N;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  splitk_reduce_kernel<<<blocks, 256, 0, s>>>(ws, (int)grid.z, M, N, e);
  count_launch(2);
}

// ------------------------------------------------------------------------------------------
// public entry points (kernels.hpp)
// ------------------------------------------------------------------------------------------
void ffma_conv_fwd(cudaStream_t s, const ConvGeom& g, const float* x, const float* w, const float* bias, float* y,
                   bool relu) {
  const int P = g.N * g.Ho * g.Wo, K = g.Kdim();
  for (int grp = 0; grp < g.group; ++grp) {
    DenseK B{w + (size_t)grp * g.Og() * K, K, g.Og(), K};
    EpiConvFwd E{y, bias, g, grp, relu ? 1 : 0, FastDiv(g.Ho * g.Wo)};
    launch(s, make_im2col_rows(x, g, grp), B, E, P, g.Og(), K);
  }
}

void ffma_conv_dgrad(cudaStream_t s, const ConvGeom& g, const float* dy, const float* w, float* dx) {
  const int P = g.N * g.H * g.W, R = g.Og() * g.kh * g.kw;
  for (int grp = 0; grp < g.group; ++grp) {
    EpiConvDgrad E{dx, g, grp, FastDiv(g.H * g.W)};
    launch(s, make_dgrad_rows(dy, g, grp), make_filter_dgrad(w, g, grp), E, P, g.Cg(), R);
  }
}

void ffma_conv_wgrad(cudaStream_t s, const ConvGeom& g, const float* x, const float* dy, float* dw, bool accumulate) {
  const int R = g.N * g.Ho * g.Wo, K = g.Kdim();
  for (int grp = 0; grp < g.group; ++grp) {
    EpiRowMajor E{dw + (size_t)grp * g.Og() * K, K, g.Og(), K, accumulate ? 1 : 0};
    launch(s, make_dy_rows(dy, g, grp), make_im2col_cols(x, g, grp), E, g.Og(), K, R);
  }
}

// Y[batch][out] = X[batch][K] * W[out][K]^T + bias   (inner_product_layer.cpp:78-91)
void ffma_ip_fwd(cudaStream_t s, int batch, int out, int K, const float* x, const float* w, const float* bias, float* y,
                 bool relu) {
  DenseK A{w, K, out, K}, B{x, K, batch, K};
  EpiTransposed E{y, out, out, batch, bias, relu ? 1 : 0};
  launch(s, A, B, E, out, batch, K);
}
// dX[batch][K] = dY[batch][out] * W[out][K]      (:110-116)
void ffma_ip_dgrad(cudaStream_t s, int batch, int out, int K, const float* dy, const float* w, float* dx) {
  DenseMN A{w, K, K, out};
  DenseK B{dy, out, batch, out};
  EpiTransposed E{dx, K, K, batch, nullptr, 0};
  launch(s, A, B, E, K, batch, out);
}
// dW[out][K] (+)= dY[batch][out]^T * X[batch][K]  (:96-103)
void ffma_ip_wgrad(cudaStream_t s, int batch, int out, int K, const float* x, const float* dy, float* dw,
                   bool accumulate) {
  DenseMN A{dy, out, out, batch}, B{x, K, K, batch};
  EpiRowMajor E{dw, K, out, K, accumulate ? 1 : 0};
  launch(s, A, B, E, out, K, batch);
}

}  // namespace sn
