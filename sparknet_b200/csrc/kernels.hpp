// kernels.hpp — host-callable launchers of every device kernel on the hot path.
// All launch on the given stream, never synchronise, never allocate (scratch comes from
// get_workspace, owned by the runtime).  Reference anchors are given per kernel in kernels.cu.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>
#include "gemm_ops.cuh"

namespace sn {

// ---- runtime services implemented in runtime.cpp -------------------------------------------------
void fatal(const char* fmt, ...) __attribute__((noreturn, format(printf, 1, 2)));
float* get_workspace(size_t nfloats);           // per-device scratch, grown on demand (never shrinks)
void count_launch(int n = 1);                   // bookkeeping for bench.py's gpu_launches
#define SN_CUDA(expr)                                                                        \
  do {                                                                                       \
    cudaError_t e_ = (expr);                                                                 \
    if (e_ != cudaSuccess) sn::fatal("CUDA error %s at %s:%d: %s", cudaGetErrorName(e_), __FILE__, __LINE__, \
                                     cudaGetErrorString(e_));                                \
  } while (0)

// ---- FFMA validation contractions (gemm_ffma.cu) ---------------------------------------------
void ffma_conv_fwd(cudaStream_t s, const ConvGeom& g, const float* x, const float* w, const float* bias, float* y, bool relu);
void ffma_conv_dgrad(cudaStream_t s, const ConvGeom& g, const float* dy, const float* w, float* dx);
void ffma_conv_wgrad(cudaStream_t s, const ConvGeom& g, const float* x, const float* dy, float* dw, bool accumulate);
void ffma_ip_fwd(cudaStream_t s, int batch, int out, int K, const float* x, const float* w, const float* bias, float* y, bool relu);
void ffma_ip_dgrad(cudaStream_t s, int batch, int out, int K, const float* dy, const float* w, float* dx);
void ffma_ip_wgrad(cudaStream_t s, int batch, int out, int K, const float* x, const float* dy, float* dw, bool accumulate);
int splitk_slices(int tiles, int R, int ctas_per_sm = 2);   // slices of the reduction so that tiles * slices fills the GPU

// ---- TF32 tcgen05 contractions (gemm_tc.cu) --------------------------------------------------
bool tc_available();
// dst[i] = tf32(src[i]) (round to nearest): the pre-rounded weight copy the TMA-fed operands read
void tc_round_to_tf32(cudaStream_t s, size_t n, const float* src, float* dst);
// `w_tf32`: weights already rounded to tf32 (tc_round_to_tf32); all other operands are rounded on the fly
void tc_conv_fwd(cudaStream_t s, const ConvGeom& g, const float* x, const float* w_tf32, const float* bias, float* y, bool relu);
void tc_test_gemm(cudaStream_t s, int a_mn, int b_mn, int tma_a, int M, int N, int R, const float* A, const float* B, float* D);
void tc_conv_dgrad(cudaStream_t s, const ConvGeom& g, const float* dy, const float* w, float* dx);
void tc_conv_wgrad(cudaStream_t s, const ConvGeom& g, const float* x, const float* dy, float* dw, bool accumulate);
// flat-grid path for stride-1 convolutions (all operands through TMA); see gemm_tc.cu
bool tc_flat_supported(const ConvGeom& g);
struct FlatGeom { int Wp, Hp, Gp, lead, bw, bh; };                    // cell width / height / size, leading zero rows, band widths
FlatGeom tc_flat_geom(const ConvGeom& g);
size_t tc_flat_rows(const ConvGeom& g);                                // grid rows of a packed tensor (x channels = floats)
size_t tc_flat_filter_floats(const ConvGeom& g, bool dgrad);
// dst zeroed once by the owner; db (nullable) = bias gradient summed in the same pass over the top gradient
void tc_flat_pack(cudaStream_t s, const ConvGeom& g, bool output_side, const float* src, float* dst, float* db = nullptr,
                  bool accumulate = false, int src_channels = 0);   // src_channels < g.C: the remaining packed channels are zeros
void tc_flat_pack_filter(cudaStream_t s, const ConvGeom& g, bool dgrad, const float* w, float* dst, int src_cg = 0);   // src_cg < g.Cg() (forward layout): zero slots
void tc_conv_fwd_flat(cudaStream_t s, const ConvGeom& g, const float* xflat, const float* wf, const float* bias, float* y, bool relu);
void tc_conv_dgrad_flat(cudaStream_t s, const ConvGeom& g, const float* dyflat, const float* wt, float* dx);
void tc_conv_wgrad_flat(cudaStream_t s, const ConvGeom& g, const float* xflat, const float* dyflat, float* dw, bool accumulate);
// window kernels (conv_flat.cu): same packed operands, one TMA load of the grid rows per channel block serves every tap
// A producer that also writes its output into the flat grid of the window convolution that reads it (tf32-rounded,
// channels-last; bands untouched = zero): flat[(lead + img*Gp + h*Wp + w) * C + c].  Saves the consumer's pack launch.
struct FlatOut { float* flat; int lead, Gp, Wp, C; };
void tcw_conv_fwd(cudaStream_t s, const ConvGeom& g, const float* xflat, const float* wf, const float* bias, float* y, bool relu,
                  const FlatOut* next = nullptr);
void tcw_conv_dgrad(cudaStream_t s, const ConvGeom& g, const float* dyflat, const float* wt, float* dx, const float* relu_gate = nullptr);
bool tcw_supported(const ConvGeom& g);
bool tcw_wgrad_supported(const ConvGeom& g);
void tcw_conv_wgrad(cudaStream_t s, const ConvGeom& g, const float* xflat, const float* dyflat, float* dw, bool accumulate);
// strided layers folded to stride 1 ("space to depth", conv_flat.cu): e = tcw_s2d_geom(g) is the equivalent stride-1 layer
bool tcw_s2d_supported(const ConvGeom& g);
ConvGeom tcw_s2d_geom(const ConvGeom& g);
void tcw_s2d_pack_input(cudaStream_t s, const ConvGeom& g, const float* x, float* xflat);
// the same from a byte-pixel staging buffer, widened on the fly: ((float)b - mean[(i / mean_div) % mean_period]) * scale
void tcw_s2d_pack_input_bytes(cudaStream_t s, const ConvGeom& g, const uint8_t* bytes, bool is_signed, const float* mean, int mean_div,
                              int mean_period, float scale, float* xflat);
void tcw_s2d_pack_filter(cudaStream_t s, const ConvGeom& g, const float* w, float* wf);
// Several filter copies in one launch (conv_flat.cu filter_pack_batch_kernel): mode 0 / 1 = tc_flat_pack_filter(dgrad = mode),
// mode 2 = tcw_s2d_pack_filter.
struct FilterPackJob { const float* w; float* dst; size_t total; int mode; int p[8]; };
struct FilterPackBatch { static constexpr int kMax = 16; FilterPackJob job[kMax]; int n; };
FilterPackJob tcw_filter_pack_job(const ConvGeom& g, int mode, const float* w, float* dst, int src_cg = 0);
void tcw_cpad_unpack_wgrad(cudaStream_t s, const ConvGeom& g, int Cpad, const float* dws, float* dw, bool accumulate);
void tcw_filter_pack_batch(cudaStream_t s, const FilterPackBatch& b);
void tcw_s2d_unpack_wgrad(cudaStream_t s, const ConvGeom& g, const float* dws, float* dw, bool accumulate);
// out[m][n] = sum_k A[m][k] * B(n,k) (+ bias[n], ReLU) on the window kernel; A [M][K] tf32-rounded, B [N][K] or (b_mn) [K][N]
bool tcw_matmul_supported(int M, int N, int K, const float* A, const float* B, bool b_mn, const float* out);
void tcw_matmul(cudaStream_t s, int M, int N, int K, const float* A, const float* B, bool b_mn, const float* bias, bool relu, float* out);
// dW[o][k] (+)= sum_b dY[b][o] * X[b][k], both operands tf32-rounded row-major matrices
bool tcw_matmul_wgrad_supported(int batch, int out, int K, const float* dy, const float* x, const float* dw);
void tcw_matmul_wgrad(cudaStream_t s, int batch, int out, int K, const float* dy_r, const float* x_r, float* dw, bool accumulate);
void tcw_set_debug(int what, int value);
int tcw_read_stamps(long long* out, int max_ctas);
float tcw_mma_rate(int cg, int n_mma, int iters, int shift_rows, int nacc, int ctas, float* issue_only);
void tc_ip_fwd(cudaStream_t s, int batch, int out, int K, const float* x, const float* w_tf32, const float* bias, float* y, bool relu);
void tc_ip_dgrad(cudaStream_t s, int batch, int out, int K, const float* dy, const float* w_tf32, float* dx);
void tc_ip_wgrad(cudaStream_t s, int batch, int out, int K, const float* x, const float* dy, float* dw, bool accumulate);

// ---- bandwidth-bound kernels (kernels.cu) ----------------------------------------------------
struct PoolGeom { int N, C, H, W, kh, kw, sh, sw, ph, pw, PH, PW; };
void pool_max_fwd(cudaStream_t s, const PoolGeom& g, const float* x, float* y, int* mask);
// gate (nullable): pooled output values; windows whose value is <= 0 pass no gradient (fused in-place ReLU backward)
void pool_max_bwd(cudaStream_t s, const PoolGeom& g, const float* dy, const int* mask, float* dx, const float* gate = nullptr);
void pool_ave_fwd(cudaStream_t s, const PoolGeom& g, const float* x, float* y);
void pool_ave_bwd(cudaStream_t s, const PoolGeom& g, const float* dy, float* dx);

void relu_fwd(cudaStream_t s, size_t n, const float* x, float* y, float slope);
void relu_bwd(cudaStream_t s, size_t n, const float* dy, const float* x_or_y, float* dx, float slope);

// mask: one byte per element (1 = keep).  If `draw` the kernel draws it (Philox4x32-10 keyed by
// seed/offset); otherwise it uses the mask as given (injected by tests).
void dropout_fwd(cudaStream_t s, size_t n, const float* x, float* y, uint8_t* mask, float ratio, float scale,
                 bool draw, uint64_t seed, uint64_t offset, const unsigned long long* offset_dev = nullptr);
void dropout_bwd(cudaStream_t s, size_t n, const float* dy, const uint8_t* mask, float* dx, float scale);

struct LrnGeom { int N, C, H, W, size; float alpha, beta, k; };
// across channels, odd local_size 3..11; nothing but y is kept: the backward pass recomputes the scale from x
void lrn_across_fwd(cudaStream_t s, const LrnGeom& g, const float* x, float* y, const FlatOut* next = nullptr);
void lrn_across_bwd(cudaStream_t s, const LrnGeom& g, const float* x, const float* dy, float* dx);
void lrn_within_fwd(cudaStream_t s, const LrnGeom& g, const float* x, float* y);
void lrn_within_bwd(cudaStream_t s, const LrnGeom& g, const float* x, const float* dy, float* dx);

// prob[rows][classes]; loss_out[0] = -(1/rows) sum log max(p[label], FLT_MIN)   (normalize = true)
// loss_log (nullable): second destination for the scalar (per-iteration loss history kept on the device)
void softmax_loss_fwd(cudaStream_t s, int rows, int classes, const float* x, const float* label, float* prob,
                      float* loss_out, float* row_scratch, float* loss_log);
void softmax_loss_bwd(cudaStream_t s, int rows, int classes, const float* prob, const float* label,
                      const float* loss_weight_dev, float* dx);
void accuracy_top1(cudaStream_t s, int rows, int classes, const float* x, const float* label, float* acc_out,
                   float* row_scratch);

// bias gradients: db[c] (+)= sum over (n, pix) dY[n][c][pix]   (base_conv_layer.cpp:303-307)
void conv_bias_grad(cudaStream_t s, int N, int C, int HW, const float* dy, float* db, bool accumulate);
// db[j] (+)= sum_i dY[i][j]                                   (inner_product_layer.cpp:104-109)
void ip_bias_grad(cudaStream_t s, int rows, int cols, const float* dy, float* db, bool accumulate);
// dy_r = tf32(dy) and (db nullable) db (+)= column sums of dy, one pass; cols % 4 == 0, 16-byte aligned pointers
void ip_round_bias(cudaStream_t s, int rows, int cols, const float* dy, float* dy_r, float* db, bool accumulate);

// fused SGD step over the whole parameter arena (sgd_solver.cpp:145-239 + blob.cpp:154-162):
//   g += decay*w ; h = momentum*h + rate*g ; g = h ; w -= h       per 1024-element chunk table
struct SgdChunk { float lr_mult, decay; };     // per-chunk lr_mult and local_decay (= weight_decay * decay_mult)
// shadow (nullable): receives the tf32-rounded new weights, same layout as w
void sgd_update(cudaStream_t s, size_t n, float* w, float* g, float* h, float* shadow, const SgdChunk* chunk_table, float rate,
                float momentum, const float* rate_dev = nullptr);
// Scalars that change every iteration, kept in device memory so that a captured iteration can be replayed
// (Solver::step): the learning rate and the Philox counters of up to kMaxDropout dropout layers.
struct StepScalars {
  static constexpr int kMaxDropout = 6;
  float rate;
  unsigned int n_dropout;
  unsigned long long dropout_offset[kMaxDropout];
};
void publish_step_scalars(cudaStream_t s, StepScalars* dst, const StepScalars& v);

// byte pixels -> fp32 (JavaData byte input): dst[i] = ((float)src[i] - mean[(i / mean_div) % mean_period]) * scale, src read as
// signed (Java byte) or unsigned; mean nullable; 16 bytes in, four 16-byte stores out per thread and step
void widen_bytes(cudaStream_t s, size_t n, const uint8_t* src, bool is_signed, const float* mean, int mean_div, int mean_period,
                 float scale, float* dst);
void fill(cudaStream_t s, size_t n, float* p, float v);
void split_add(cudaStream_t s, size_t n, const float* a, const float* b, float* out);
void axpy(cudaStream_t s, size_t n, float a, const float* x, float* y);

// averaging: out[i] = (((r0[i] + r1[i]) + r2[i]) + ...) / (float)nrep, IEEE fp32, rank order
// (src/main/java/libs/JavaNDArray.java:84-117, src/main/scala/apps/CifarApp.scala:133-134)
void average_replicas(cudaStream_t s, const float* const* replicas_dev, int nrep, size_t n, float* out);
// Peer-memory exchange (kernels.cu "Weight averaging over peer memory"): w[r] = replica r's weight arena and
// flags[r] = replica r's 32-word flag block, both addressable from this GPU (own memory or an IPC mapping).
struct PeerTable {
  float* w[8];
  unsigned* flags[8];
};
void average_peers(cudaStream_t s, const PeerTable& t, int nranks, int rank, size_t count, unsigned epoch,
                   unsigned long long timeout_ns, int max_grid = 0);

}  // namespace sn
