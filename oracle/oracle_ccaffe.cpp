// oracle_ccaffe.cpp — CPU ORACLE.  TEST INFRASTRUCTURE ONLY.
//
// A plain C++17 restatement of the reference's CPU_ONLY Caffe path behind the same C ABI as
// libccaffe (libccaffe/ccaffe.h), so that one ctypes harness can drive oracle and product.
// Nothing in the product (sparknet_b200/, libccaffe.so) includes, links or calls this file;
// only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do.
//
// Why a restatement: the reference cannot be compiled in this image (it needs protoc,
// libprotobuf, boost, glog, gflags, OpenBLAS dev files, hdf5, lmdb, leveldb — none present,
// see DESIGN.md).  Parity pin: the reference's own known-answer tests, restated in
// tests/test_oracle_golden.py (test_util_blas.cpp:20-131, test_pooling_layer.cpp:45-569,
// test_convolution_layer.cpp:17-489, test_lrn_layer.cpp:50-229, test_softmax_*,
// test_gradient_based_solver.cpp:225-399, NDArraySpec.scala:34-56).  What stays unpinned:
// (1) the BLAS summation order (OpenBLAS, version unpinned in the reference) — this file uses
// a fixed k-ascending order; (2) the RNG streams (boost::mt19937) of fillers and dropout —
// tests copy weights and inject masks instead; (3) the N>=3 cross-worker summation order
// (Spark task-completion order) — fixed here to rank order as BASELINE.json's north_star says.
//
// Every function cites the reference file:line it follows (paths relative to
// /root/reference/caffe/ unless stated).
#include <algorithm>
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <map>
#include <memory>
#include <random>
#include <set>
#include <sstream>
#include <string>
#include <thread>
#include <unistd.h>
#include <utility>
#include <vector>
#include <dlfcn.h>

#include "../sparknet_b200/csrc/caffe_proto.hpp"

namespace cp = caffe_proto;
using std::string;
using std::vector;

#define OCHECK(cond, msg)                                                                  \
  do {                                                                                     \
    if (!(cond)) {                                                                         \
      std::ostringstream os_;                                                              \
      os_ << "oracle: check failed: " #cond " — " << msg << " (" << __FILE__ << ":" << __LINE__ << ")"; \
      fprintf(stderr, "%s\n", os_.str().c_str());                                          \
      abort();                                                                             \
    }                                                                                      \
  } while (0)

// ------------------------------------------------------------------------------------------
// BLAS restatement (src/caffe/util/math_functions.cpp:13-127, include/caffe/util/mkl_alternate.hpp:34-88)
// Fixed k-ascending summation order.  An optional real sgemm (dlopen'ed OpenBLAS from the
// scipy wheel) can be switched on for the *timed* CPU baseline only (oracle_use_blas).
// ------------------------------------------------------------------------------------------
enum Trans { NoTrans = 0, DoTrans = 1 };
typedef void (*cblas_sgemm_t)(int, int, int, int, int, int, float, const float*, int, const float*, int, float,
                              float*, int);
static cblas_sgemm_t g_sgemm = nullptr;

// Row-parallel helper (std::thread; each output row is still reduced by one thread in a fixed
// order, so results do not depend on the thread count).  g_threads = 1 keeps it serial.
static int g_threads = 1;
template <class F>
static void parallel_rows(int rows, long work, F fn) {
  int nt = std::min(g_threads, rows);
  if (nt <= 1 || work < 400000) {
    for (int i = 0; i < rows; ++i) fn(i);
    return;
  }
  vector<std::thread> th;
  for (int t = 0; t < nt; ++t)
    th.emplace_back([=]() {
      for (int i = (int)((long)rows * t / nt); i < (int)((long)rows * (t + 1) / nt); ++i) fn(i);
    });
  for (auto& x : th) x.join();
}

// C[M,N] = alpha * op(A)[M,K] * op(B)[K,N] + beta * C   (row major)
static void cpu_gemm(Trans ta, Trans tb, int M, int N, int K, float alpha, const float* A, const float* B, float beta,
                     float* C) {
  if (g_sgemm) {  // CblasRowMajor=101, NoTrans=111, Trans=112
    int lda = (ta == NoTrans) ? K : M, ldb = (tb == NoTrans) ? N : K;
    g_sgemm(101, ta == NoTrans ? 111 : 112, tb == NoTrans ? 111 : 112, M, N, K, alpha, A, lda, B, ldb, beta, C, N);
    return;
  }
  parallel_rows(M, (long)M * N * K, [&](int i) {
    float* c = C + (size_t)i * N;
    if (beta == 0.f) {
      for (int j = 0; j < N; ++j) c[j] = 0.f;
    } else if (beta != 1.f) {
      for (int j = 0; j < N; ++j) c[j] *= beta;
    }
    if (tb == NoTrans) {
      // i-k-j order: each c[j] accumulates k = 0..K-1 in ascending order
      for (int k = 0; k < K; ++k) {
        const float a = alpha * ((ta == NoTrans) ? A[(size_t)i * K + k] : A[(size_t)k * M + i]);
        const float* b = B + (size_t)k * N;
        for (int j = 0; j < N; ++j) c[j] += a * b[j];
      }
    } else {
      for (int j = 0; j < N; ++j) {
        const float* b = B + (size_t)j * K;
        float acc = 0.f;
        if (ta == NoTrans) {
          const float* a = A + (size_t)i * K;
          for (int k = 0; k < K; ++k) acc += a[k] * b[k];
        } else {
          for (int k = 0; k < K; ++k) acc += A[(size_t)k * M + i] * b[k];
        }
        c[j] += alpha * acc;
      }
    }
  });
}

// y = alpha * op(A) x + beta * y, A is M x N row major (math_functions.cpp:37-50)
static void cpu_gemv(Trans ta, int M, int N, float alpha, const float* A, const float* x, float beta, float* y) {
  if (ta == NoTrans) {
    for (int i = 0; i < M; ++i) {
      float acc = 0.f;
      for (int j = 0; j < N; ++j) acc += A[(size_t)i * N + j] * x[j];
      y[i] = alpha * acc + (beta == 0.f ? 0.f : beta * y[i]);
    }
  } else {
    for (int j = 0; j < N; ++j) y[j] = (beta == 0.f ? 0.f : beta * y[j]);
    for (int i = 0; i < M; ++i) {
      const float ax = alpha * x[i];
      for (int j = 0; j < N; ++j) y[j] += ax * A[(size_t)i * N + j];
    }
  }
}
static void cpu_axpy(size_t n, float a, const float* x, float* y) {
  for (size_t i = 0; i < n; ++i) y[i] += a * x[i];
}
static void cpu_scal(size_t n, float a, float* x) {
  for (size_t i = 0; i < n; ++i) x[i] *= a;
}
// axpby = scal then axpy (mkl_alternate.hpp:83-88)
static void cpu_axpby(size_t n, float a, const float* x, float b, float* y) {
  cpu_scal(n, b, y);
  cpu_axpy(n, a, x, y);
}
static float cpu_dot(size_t n, const float* x, const float* y) {
  float acc = 0.f;
  for (size_t i = 0; i < n; ++i) acc += x[i] * y[i];
  return acc;
}

// ------------------------------------------------------------------------------------------
// Blob (src/caffe/blob.cpp:23-177).  Host only: data + diff.
// ------------------------------------------------------------------------------------------
struct Blob {
  vector<int> shape;
  vector<float> data, diff;
  vector<int> idata;  // int payload (pooling mask / dropout mask), same count
  size_t count() const {
    size_t c = 1;
    for (int d : shape) c *= (size_t)d;
    return c;
  }
  size_t count(int from, int to) const {
    size_t c = 1;
    for (int i = from; i < to; ++i) c *= (size_t)shape[i];
    return c;
  }
  size_t count_from(int from) const { return count(from, (int)shape.size()); }
  void reshape(const vector<int>& s) {
    shape = s;
    size_t c = count();
    if (data.size() != c) {
      data.assign(c, 0.f);
      diff.assign(c, 0.f);
    }
  }
  int dim(int i) const { return shape[(size_t)(i < 0 ? i + (int)shape.size() : i)]; }
  int num() const { return shape.size() > 0 ? shape[0] : 1; }
  int channels() const { return shape.size() > 1 ? shape[1] : 1; }
  int height() const { return shape.size() > 2 ? shape[2] : 1; }
  int width() const { return shape.size() > 3 ? shape[3] : 1; }
};
typedef std::shared_ptr<Blob> BlobP;

extern "C" {
typedef void (*java_callback_t)(void* data, int batch_size, int num_dims, int* shape);
typedef void (*error_callback_t)(const char* msg);
}

// ------------------------------------------------------------------------------------------
// RNG (src/caffe/common.cpp:23-63, include/caffe/util/rng.hpp).  std::mt19937 stands in for
// boost::mt19937 (same engine); the distributions differ from boost's, hence "unpinned".
// ------------------------------------------------------------------------------------------
static std::mt19937& rng() {
  static thread_local std::mt19937 g(std::random_device{}());
  return g;
}
static void rng_seed(uint32_t s) { rng().seed(s); }

// Fillers (include/caffe/filler.hpp:31-46 constant, 65-95 gaussian, 144-166 xavier, 48-63 uniform)
static void fill_blob(const cp::FillerParameter& f, Blob* b) {
  size_t n = b->count();
  OCHECK(n > 0, "empty blob in filler");
  if (f.type == "constant") {
    for (size_t i = 0; i < n; ++i) b->data[i] = f.value;
  } else if (f.type == "gaussian") {
    std::normal_distribution<float> d(f.mean, f.std);
    for (size_t i = 0; i < n; ++i) b->data[i] = d(rng());
    OCHECK(f.sparse == -1, "sparse gaussian filler not on the path");
  } else if (f.type == "xavier") {
    int fan_in = (int)(n / (size_t)b->num());
    int fan_out = (int)(n / (size_t)b->channels());
    float nn = (float)fan_in;
    if (f.variance_norm == 2) nn = (fan_in + fan_out) / 2.f;
    else if (f.variance_norm == 1) nn = (float)fan_out;
    float scale = std::sqrt(3.f / nn);
    std::uniform_real_distribution<float> d(-scale, scale);
    for (size_t i = 0; i < n; ++i) b->data[i] = d(rng());
  } else if (f.type == "uniform") {
    std::uniform_real_distribution<float> d(f.min, f.max);
    for (size_t i = 0; i < n; ++i) b->data[i] = d(rng());
  } else {
    OCHECK(false, "filler type '" << f.type << "' is not on the supported path");
  }
}

// ------------------------------------------------------------------------------------------
// Layers
// ------------------------------------------------------------------------------------------
struct Layer {
  cp::LayerParameter lp;
  int phase = cp::TRAIN;
  vector<BlobP> blobs;                 // learnable parameters
  vector<bool> param_propagate_down;
  vector<float> loss;                  // per-top loss weight
  explicit Layer(const cp::LayerParameter& p) : lp(p) {}
  virtual ~Layer() {}
  virtual const char* type() const = 0;
  virtual void setup(const vector<Blob*>& bottom, const vector<Blob*>& top) = 0;  // LayerSetUp + Reshape
  virtual void forward(const vector<Blob*>& bottom, const vector<Blob*>& top) = 0;
  virtual void backward(const vector<Blob*>& top, const vector<bool>& propagate_down, const vector<Blob*>& bottom) = 0;
  virtual bool is_loss() const { return false; }
  // Layer::SetLossWeights (include/caffe/layer.hpp:405-427): loss weights live in top diff
  void set_loss_weights(const vector<Blob*>& top) {
    loss.assign(top.size(), 0.f);
    for (size_t i = 0; i < lp.loss_weight.size() && i < top.size(); ++i) {
      float w = lp.loss_weight[i];
      if (w == 0.f) continue;
      loss[i] = w;
      std::fill(top[i]->diff.begin(), top[i]->diff.end(), w);
    }
  }
  // Layer::Forward wrapper (include/caffe/layer.hpp:451-487): loss = sum dot(top, loss weights)
  float forward_with_loss(const vector<Blob*>& bottom, const vector<Blob*>& top) {
    forward(bottom, top);
    float l = 0.f;
    for (size_t i = 0; i < top.size(); ++i) {
      if (i >= loss.size() || loss[i] == 0.f) continue;
      l += cpu_dot(top[i]->count(), top[i]->data.data(), top[i]->diff.data());
    }
    return l;
  }
};

// JavaDataLayer (src/caffe/layers/java_data_layer.cpp:10-44, include/caffe/data_layers.hpp:324-354)
struct JavaDataLayer : Layer {
  java_callback_t cb = nullptr;
  vector<int> shape, java_shape;
  vector<float> buffer;
  using Layer::Layer;
  const char* type() const override { return "JavaData"; }
  void setup(const vector<Blob*>&, const vector<Blob*>& top) override {
    OCHECK(lp.has_java_data_param, "JavaData layer " << lp.name << " needs java_data_param");
    for (int64_t d : lp.java_data_shape.dim) shape.push_back((int)d);
    OCHECK(shape.size() >= 1, "JavaData shape needs a batch dimension");
    java_shape.assign(shape.begin() + 1, shape.end());
    top[0]->reshape(shape);
    OCHECK(top[0]->count() > 0, "batch_size, channels, height, and width must be specified and positive");
    buffer.assign(top[0]->count(), 0.f);
  }
  void forward(const vector<Blob*>&, const vector<Blob*>& top) override {
    OCHECK(cb, "JavaDataLayer needs to be initalized by calling SetCallback");
    cb((void*)buffer.data(), shape[0], (int)shape.size() - 1, java_shape.data());
    top[0]->reshape(shape);
    memcpy(top[0]->data.data(), buffer.data(), buffer.size() * sizeof(float));  // set_cpu_data(buffer_)
  }
  void backward(const vector<Blob*>&, const vector<bool>&, const vector<Blob*>&) override {}
};

// im2col / col2im (src/caffe/util/im2col.cpp:9-31 and its col2im twin)
static void im2col(const float* im, int C, int H, int W, int kh, int kw, int ph, int pw, int sh, int sw, float* col) {
  const int Ho = (H + 2 * ph - kh) / sh + 1, Wo = (W + 2 * pw - kw) / sw + 1, CC = C * kh * kw;
  for (int cc = 0; cc < CC; ++cc) {
    int wo = cc % kw, ho = (cc / kw) % kh, c = cc / kh / kw;
    for (int h = 0; h < Ho; ++h)
      for (int w = 0; w < Wo; ++w) {
        int hi = h * sh - ph + ho, wi = w * sw - pw + wo;
        col[((size_t)cc * Ho + h) * Wo + w] =
            (hi >= 0 && wi >= 0 && hi < H && wi < W) ? im[((size_t)c * H + hi) * W + wi] : 0.f;
      }
  }
}
static void col2im(const float* col, int C, int H, int W, int kh, int kw, int ph, int pw, int sh, int sw, float* im) {
  std::fill(im, im + (size_t)C * H * W, 0.f);
  const int Ho = (H + 2 * ph - kh) / sh + 1, Wo = (W + 2 * pw - kw) / sw + 1, CC = C * kh * kw;
  for (int cc = 0; cc < CC; ++cc) {
    int wo = cc % kw, ho = (cc / kw) % kh, c = cc / kh / kw;
    for (int h = 0; h < Ho; ++h)
      for (int w = 0; w < Wo; ++w) {
        int hi = h * sh - ph + ho, wi = w * sw - pw + wo;
        if (hi >= 0 && hi < H && wi >= 0 && wi < W)
          im[((size_t)c * H + hi) * W + wi] += col[((size_t)cc * Ho + h) * Wo + w];
      }
  }
}

// ConvolutionLayer (src/caffe/layers/conv_layer.cpp:8-71, base_conv_layer.cpp:12-240 setup, 243-307 gemm helpers)
struct ConvolutionLayer : Layer {
  int kh, kw, sh, sw, ph, pw, group, num_output, channels, H, W, Ho, Wo, num;
  bool bias_term;
  vector<float> col, ones;
  using Layer::Layer;
  const char* type() const override { return "Convolution"; }
  void setup(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    const cp::ConvolutionParameter& c = lp.convolution_param;
    OCHECK(bottom[0]->shape.size() == 4, "Convolution input must be 4-D (N,C,H,W)");
    OCHECK(c.axis == 1, "only channel axis 1 supported");
    if (c.has_kernel_h || c.has_kernel_w) {
      OCHECK(c.kernel_size.empty(), "Either kernel_size or kernel_h/w should be specified; not both.");
      kh = (int)c.kernel_h; kw = (int)c.kernel_w;
    } else {
      OCHECK(!c.kernel_size.empty(), "kernel_size must be specified");
      kh = (int)c.kernel_size[0];
      kw = (int)(c.kernel_size.size() > 1 ? c.kernel_size[1] : c.kernel_size[0]);
    }
    OCHECK(kh > 0 && kw > 0, "Filter dimensions must be nonzero.");
    if (c.has_stride_h || c.has_stride_w) { sh = (int)c.stride_h; sw = (int)c.stride_w; }
    else if (c.stride.empty()) { sh = sw = 1; }
    else { sh = (int)c.stride[0]; sw = (int)(c.stride.size() > 1 ? c.stride[1] : c.stride[0]); }
    if (c.has_pad_h || c.has_pad_w) { ph = (int)c.pad_h; pw = (int)c.pad_w; }
    else if (c.pad.empty()) { ph = pw = 0; }
    else { ph = (int)c.pad[0]; pw = (int)(c.pad.size() > 1 ? c.pad[1] : c.pad[0]); }
    channels = bottom[0]->channels();
    num_output = (int)c.num_output;
    OCHECK(num_output > 0, "num_output must be positive");
    group = (int)c.group;
    OCHECK(channels % group == 0 && num_output % group == 0, "channels / num_output must be multiples of group");
    bias_term = c.bias_term;
    if (blobs.empty()) {
      blobs.push_back(std::make_shared<Blob>());
      blobs[0]->reshape({num_output, channels / group, kh, kw});
      fill_blob(c.weight_filler, blobs[0].get());
      if (bias_term) {
        blobs.push_back(std::make_shared<Blob>());
        blobs[1]->reshape({num_output});
        fill_blob(c.bias_filler, blobs[1].get());
      }
    }
    param_propagate_down.assign(blobs.size(), true);
    num = bottom[0]->num(); H = bottom[0]->height(); W = bottom[0]->width();
    Ho = (H + 2 * ph - kh) / sh + 1;   // conv_layer.cpp:9-21 compute_output_shape (floor)
    Wo = (W + 2 * pw - kw) / sw + 1;
    top[0]->reshape({num, num_output, Ho, Wo});
    col.assign((size_t)channels * kh * kw * Ho * Wo, 0.f);
    ones.assign((size_t)Ho * Wo, 1.f);
  }
  void forward(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    const int kdim = channels / group * kh * kw, osp = Ho * Wo, og = num_output / group;
    const float* w = blobs[0]->data.data();
    for (int n = 0; n < num; ++n) {
      const float* in = bottom[0]->data.data() + (size_t)n * channels * H * W;
      float* out = top[0]->data.data() + (size_t)n * num_output * osp;
      im2col(in, channels, H, W, kh, kw, ph, pw, sh, sw, col.data());
      for (int g = 0; g < group; ++g)   // base_conv_layer.cpp:253-258
        cpu_gemm(NoTrans, NoTrans, og, osp, kdim, 1.f, w + (size_t)g * og * kdim, col.data() + (size_t)g * kdim * osp,
                 0.f, out + (size_t)g * og * osp);
      if (bias_term)                    // :263-266 rank-1 update with the all-ones multiplier
        cpu_gemm(NoTrans, NoTrans, num_output, osp, 1, 1.f, blobs[1]->data.data(), ones.data(), 1.f, out);
    }
  }
  void backward(const vector<Blob*>& top, const vector<bool>& pd, const vector<Blob*>& bottom) override {
    const int kdim = channels / group * kh * kw, osp = Ho * Wo, og = num_output / group;
    const float* w = blobs[0]->data.data();
    float* wd = blobs[0]->diff.data();
    const float* td = top[0]->diff.data();
    if (bias_term && param_propagate_down[1]) {   // conv_layer.cpp:44-50 + base_conv_layer.cpp:303-307
      for (int n = 0; n < num; ++n)
        cpu_gemv(NoTrans, num_output, osp, 1.f, td + (size_t)n * num_output * osp, ones.data(), 1.f,
                 blobs[1]->diff.data());
    }
    if (param_propagate_down[0] || pd[0]) {       // conv_layer.cpp:51-66
      for (int n = 0; n < num; ++n) {
        const float* in = bottom[0]->data.data() + (size_t)n * channels * H * W;
        const float* od = td + (size_t)n * num_output * osp;
        if (param_propagate_down[0]) {            // weight_cpu_gemm :285-300 (beta = 1: accumulates)
          im2col(in, channels, H, W, kh, kw, ph, pw, sh, sw, col.data());
          for (int g = 0; g < group; ++g)
            cpu_gemm(NoTrans, DoTrans, og, kdim, osp, 1.f, od + (size_t)g * og * osp,
                     col.data() + (size_t)g * kdim * osp, 1.f, wd + (size_t)g * og * kdim);
        }
        if (pd[0]) {                              // backward_cpu_gemm :268-283
          for (int g = 0; g < group; ++g)
            cpu_gemm(DoTrans, NoTrans, kdim, osp, og, 1.f, w + (size_t)g * og * kdim, od + (size_t)g * og * osp, 0.f,
                     col.data() + (size_t)g * kdim * osp);
          col2im(col.data(), channels, H, W, kh, kw, ph, pw, sh, sw,
                 bottom[0]->diff.data() + (size_t)n * channels * H * W);
        }
      }
    }
  }
};

// InnerProductLayer (src/caffe/layers/inner_product_layer.cpp:10-117)
struct InnerProductLayer : Layer {
  int M, N, K;
  bool bias_term;
  vector<float> ones;
  using Layer::Layer;
  const char* type() const override { return "InnerProduct"; }
  void setup(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    const cp::InnerProductParameter& p = lp.inner_product_param;
    N = (int)p.num_output;
    bias_term = p.bias_term;
    int axis = p.axis < 0 ? p.axis + (int)bottom[0]->shape.size() : p.axis;
    K = (int)bottom[0]->count_from(axis);
    M = (int)bottom[0]->count(0, axis);
    if (blobs.empty()) {
      blobs.push_back(std::make_shared<Blob>());
      blobs[0]->reshape({N, K});
      fill_blob(p.weight_filler, blobs[0].get());
      if (bias_term) {
        blobs.push_back(std::make_shared<Blob>());
        blobs[1]->reshape({N});
        fill_blob(p.bias_filler, blobs[1].get());
      }
    }
    param_propagate_down.assign(blobs.size(), true);
    vector<int> ts(bottom[0]->shape.begin(), bottom[0]->shape.begin() + axis);
    ts.push_back(N);
    top[0]->reshape(ts);
    ones.assign((size_t)M, 1.f);
  }
  void forward(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    cpu_gemm(NoTrans, DoTrans, M, N, K, 1.f, bottom[0]->data.data(), blobs[0]->data.data(), 0.f, top[0]->data.data());
    if (bias_term)
      cpu_gemm(NoTrans, NoTrans, M, N, 1, 1.f, ones.data(), blobs[1]->data.data(), 1.f, top[0]->data.data());
  }
  void backward(const vector<Blob*>& top, const vector<bool>& pd, const vector<Blob*>& bottom) override {
    if (param_propagate_down[0])
      cpu_gemm(DoTrans, NoTrans, N, K, M, 1.f, top[0]->diff.data(), bottom[0]->data.data(), 1.f,
               blobs[0]->diff.data());
    if (bias_term && param_propagate_down[1])
      cpu_gemv(DoTrans, M, N, 1.f, top[0]->diff.data(), ones.data(), 1.f, blobs[1]->diff.data());
    if (pd[0])
      cpu_gemm(NoTrans, NoTrans, M, K, N, 1.f, top[0]->diff.data(), blobs[0]->data.data(), 0.f,
               bottom[0]->diff.data());
  }
};

// PoolingLayer (src/caffe/layers/pooling_layer.cpp:17-307)
struct PoolingLayer : Layer {
  int kh, kw, sh, sw, ph, pw, C, H, W, PH, PW, method;
  bool global_pooling;
  vector<int> max_idx;
  using Layer::Layer;
  const char* type() const override { return "Pooling"; }
  void setup(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    const cp::PoolingParameter& p = lp.pooling_param;
    method = p.pool;
    OCHECK(method == 0 || method == 1, "only MAX and AVE pooling are on the path");
    global_pooling = p.global_pooling;
    if (global_pooling) {
      kh = bottom[0]->height(); kw = bottom[0]->width();
    } else if (p.has_kernel_size) {
      kh = kw = (int)p.kernel_size;
    } else {
      OCHECK(p.has_kernel_h && p.has_kernel_w, "For non-square filters both kernel_h and kernel_w are required.");
      kh = (int)p.kernel_h; kw = (int)p.kernel_w;
    }
    OCHECK(kh > 0 && kw > 0, "Filter dimensions cannot be zero.");
    if (p.has_pad_h || p.has_pad_w) { ph = (int)p.pad_h; pw = (int)p.pad_w; } else { ph = pw = (int)p.pad; }
    if (p.has_stride_h || p.has_stride_w) { sh = (int)p.stride_h; sw = (int)p.stride_w; } else { sh = sw = (int)p.stride; }
    if (ph != 0 || pw != 0) {
      OCHECK(ph < kh && pw < kw, "pad must be smaller than kernel");
    }
    OCHECK(bottom[0]->shape.size() == 4, "Pooling input must have 4 axes");
    C = bottom[0]->channels(); H = bottom[0]->height(); W = bottom[0]->width();
    // :90-105 — ceil, then clip the last window if it would start in the padding
    PH = (int)std::ceil((float)(H + 2 * ph - kh) / sh) + 1;
    PW = (int)std::ceil((float)(W + 2 * pw - kw) / sw) + 1;
    if (ph || pw) {
      if ((PH - 1) * sh >= H + ph) --PH;
      if ((PW - 1) * sw >= W + pw) --PW;
      OCHECK((PH - 1) * sh < H + ph && (PW - 1) * sw < W + pw, "bad pooled size");
    }
    top[0]->reshape({bottom[0]->num(), C, PH, PW});
    if (method == 0) max_idx.assign(top[0]->count(), -1);
  }
  void forward(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    const float* bd = bottom[0]->data.data();
    float* td = top[0]->data.data();
    const size_t tc = top[0]->count();
    const int planes = bottom[0]->num() * C;
    if (method == 0) {  // MAX :136-178
      int* mask = max_idx.data();
      std::fill(mask, mask + tc, -1);
      std::fill(td, td + tc, -FLT_MAX);
      for (int pl = 0; pl < planes; ++pl) {
        for (int y = 0; y < PH; ++y)
          for (int x = 0; x < PW; ++x) {
            int hs = y * sh - ph, ws = x * sw - pw;
            int he = std::min(hs + kh, H), we = std::min(ws + kw, W);
            hs = std::max(hs, 0); ws = std::max(ws, 0);
            const int pi = y * PW + x;
            for (int h = hs; h < he; ++h)
              for (int w = ws; w < we; ++w) {
                const int idx = h * W + w;
                if (bd[idx] > td[pi]) { td[pi] = bd[idx]; mask[pi] = idx; }
              }
          }
        bd += H * W; td += PH * PW; mask += PH * PW;
      }
    } else {            // AVE :179-212
      std::fill(td, td + tc, 0.f);
      for (int pl = 0; pl < planes; ++pl) {
        for (int y = 0; y < PH; ++y)
          for (int x = 0; x < PW; ++x) {
            int hs = y * sh - ph, ws = x * sw - pw;
            int he = std::min(hs + kh, H + ph), we = std::min(ws + kw, W + pw);
            int pool_size = (he - hs) * (we - ws);
            hs = std::max(hs, 0); ws = std::max(ws, 0);
            he = std::min(he, H); we = std::min(we, W);
            for (int h = hs; h < he; ++h)
              for (int w = ws; w < we; ++w) td[y * PW + x] += bd[h * W + w];
            td[y * PW + x] /= pool_size;
          }
        bd += H * W; td += PH * PW;
      }
    }
  }
  void backward(const vector<Blob*>& top, const vector<bool>& pd, const vector<Blob*>& bottom) override {
    if (!pd[0]) return;
    const float* tdiff = top[0]->diff.data();
    float* bdiff = bottom[0]->diff.data();
    std::fill(bdiff, bdiff + bottom[0]->count(), 0.f);
    const int planes = top[0]->num() * C;
    if (method == 0) {  // :235-262
      const int* mask = max_idx.data();
      for (int pl = 0; pl < planes; ++pl) {
        for (int i = 0; i < PH * PW; ++i) bdiff[mask[i]] += tdiff[i];
        bdiff += H * W; tdiff += PH * PW; mask += PH * PW;
      }
    } else {            // :263-292
      for (int pl = 0; pl < planes; ++pl) {
        for (int y = 0; y < PH; ++y)
          for (int x = 0; x < PW; ++x) {
            int hs = y * sh - ph, ws = x * sw - pw;
            int he = std::min(hs + kh, H + ph), we = std::min(ws + kw, W + pw);
            int pool_size = (he - hs) * (we - ws);
            hs = std::max(hs, 0); ws = std::max(ws, 0);
            he = std::min(he, H); we = std::min(we, W);
            for (int h = hs; h < he; ++h)
              for (int w = ws; w < we; ++w) bdiff[h * W + w] += tdiff[y * PW + x] / pool_size;
          }
        bdiff += H * W; tdiff += PH * PW;
      }
    }
  }
};

// ReLULayer (src/caffe/layers/relu_layer.cpp:9-36)
struct ReLULayer : Layer {
  using Layer::Layer;
  const char* type() const override { return "ReLU"; }
  void setup(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    if (top[0] != bottom[0]) top[0]->reshape(bottom[0]->shape);
  }
  void forward(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    const float s = lp.relu_negative_slope;
    const size_t n = bottom[0]->count();
    const float* b = bottom[0]->data.data();
    float* t = top[0]->data.data();
    for (size_t i = 0; i < n; ++i) t[i] = std::max(b[i], 0.f) + s * std::min(b[i], 0.f);
  }
  void backward(const vector<Blob*>& top, const vector<bool>& pd, const vector<Blob*>& bottom) override {
    if (!pd[0]) return;
    const float s = lp.relu_negative_slope;
    const size_t n = bottom[0]->count();
    const float* b = bottom[0]->data.data();
    const float* td = top[0]->diff.data();
    float* bd = bottom[0]->diff.data();
    for (size_t i = 0; i < n; ++i) bd[i] = td[i] * ((b[i] > 0) + s * (b[i] <= 0));
  }
};

// DropoutLayer (src/caffe/layers/dropout_layer.cpp:11-65).  Mask = bernoulli(1-p) ints unless injected.
struct DropoutLayer : Layer {
  float threshold, scale;
  vector<unsigned> mask;
  bool injected = false;
  using Layer::Layer;
  const char* type() const override { return "Dropout"; }
  void setup(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    threshold = lp.dropout_ratio;
    scale = (float)(1. / (1. - threshold));
    if (top[0] != bottom[0]) top[0]->reshape(bottom[0]->shape);
    mask.assign(bottom[0]->count(), 0u);
  }
  void forward(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    const size_t n = bottom[0]->count();
    const float* b = bottom[0]->data.data();
    float* t = top[0]->data.data();
    if (phase == cp::TRAIN) {
      if (!injected) {
        std::bernoulli_distribution d(1. - threshold);
        for (size_t i = 0; i < n; ++i) mask[i] = d(rng()) ? 1u : 0u;
      }
      for (size_t i = 0; i < n; ++i) t[i] = b[i] * mask[i] * scale;
    } else if (t != b) {
      memcpy(t, b, n * sizeof(float));
    }
  }
  void backward(const vector<Blob*>& top, const vector<bool>& pd, const vector<Blob*>& bottom) override {
    if (!pd[0]) return;
    const size_t n = bottom[0]->count();
    const float* td = top[0]->diff.data();
    float* bd = bottom[0]->diff.data();
    if (phase == cp::TRAIN) {
      for (size_t i = 0; i < n; ++i) bd[i] = td[i] * mask[i] * scale;
    } else if (bd != td) {
      memcpy(bd, td, n * sizeof(float));
    }
  }
};

// SplitLayer (src/caffe/layers/split_layer.cpp:10-49): tops share the bottom's data; diffs are summed.
struct SplitLayer : Layer {
  using Layer::Layer;
  const char* type() const override { return "Split"; }
  void setup(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    for (Blob* t : top) t->reshape(bottom[0]->shape);
  }
  void forward(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    for (Blob* t : top) t->data = bottom[0]->data;  // ShareData
  }
  void backward(const vector<Blob*>& top, const vector<bool>& pd, const vector<Blob*>& bottom) override {
    if (!pd[0]) return;
    const size_t n = bottom[0]->count();
    float* bd = bottom[0]->diff.data();
    if (top.size() == 1) { memcpy(bd, top[0]->diff.data(), n * sizeof(float)); return; }
    for (size_t i = 0; i < n; ++i) bd[i] = top[0]->diff[i] + top[1]->diff[i];   // caffe_add :41
    for (size_t k = 2; k < top.size(); ++k) cpu_axpy(n, 1.f, top[k]->diff.data(), bd);
  }
};

// LRNLayer (src/caffe/layers/lrn_layer.cpp:9-247).  ACROSS_CHANNELS restates :108-152 / :180-232.
// WITHIN_CHANNEL restates the five-sub-layer composition :17-65,155-162,235-247 with
// PowerLayer (power_layer.cpp:20-93), AVE PoolingLayer, EltwiseLayer PROD with stable gradient
// (eltwise_layer.cpp:53-58,111-123) and SplitLayer (split_layer.cpp:26-49).
struct LRNLayer : Layer {
  int size, pre_pad, num, C, H, W, region;
  float alpha, beta, k;
  Blob scale;                                     // ACROSS
  // WITHIN: blobs of the sub-net
  Blob square_out, pool_out, power_out;           // x^2 ; avepool ; (1+alpha*pool)^-beta
  vector<float> product_input_diff, square_input_diff;
  using Layer::Layer;
  const char* type() const override { return "LRN"; }
  void setup(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    const cp::LRNParameter& p = lp.lrn_param;
    size = (int)p.local_size;
    OCHECK(size % 2 == 1, "LRN only supports odd values for local_size");
    pre_pad = (size - 1) / 2;
    alpha = p.alpha; beta = p.beta; k = p.k; region = p.norm_region;
    OCHECK(bottom[0]->shape.size() == 4, "LRN input must have 4 axes");
    num = bottom[0]->num(); C = bottom[0]->channels(); H = bottom[0]->height(); W = bottom[0]->width();
    top[0]->reshape(bottom[0]->shape);
    if (region == 0) {
      scale.reshape(bottom[0]->shape);
    } else {
      square_out.reshape(bottom[0]->shape);
      pool_out.reshape(bottom[0]->shape);   // k=size, pad=pre_pad, stride 1 keeps H x W
      power_out.reshape(bottom[0]->shape);
      product_input_diff.assign(bottom[0]->count(), 0.f);
      square_input_diff.assign(bottom[0]->count(), 0.f);
    }
  }
  // AVE pool, kernel=size, pad=pre_pad, stride=1 (pooling_layer.cpp:179-212)
  void avepool_fwd(const float* in, float* out) const {
    for (int pl = 0; pl < num * C; ++pl) {
      for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
          int hs = y - pre_pad, ws = x - pre_pad;
          int he = std::min(hs + size, H + pre_pad), we = std::min(ws + size, W + pre_pad);
          int pool_size = (he - hs) * (we - ws);
          hs = std::max(hs, 0); ws = std::max(ws, 0); he = std::min(he, H); we = std::min(we, W);
          float acc = 0.f;
          for (int h = hs; h < he; ++h)
            for (int w = ws; w < we; ++w) acc += in[h * W + w];
          out[y * W + x] = acc / pool_size;
        }
      in += H * W; out += H * W;
    }
  }
  void avepool_bwd(const float* tdiff, float* bdiff) const {
    std::fill(bdiff, bdiff + (size_t)num * C * H * W, 0.f);
    for (int pl = 0; pl < num * C; ++pl) {
      for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
          int hs = y - pre_pad, ws = x - pre_pad;
          int he = std::min(hs + size, H + pre_pad), we = std::min(ws + size, W + pre_pad);
          int pool_size = (he - hs) * (we - ws);
          hs = std::max(hs, 0); ws = std::max(ws, 0); he = std::min(he, H); we = std::min(we, W);
          for (int h = hs; h < he; ++h)
            for (int w = ws; w < we; ++w) bdiff[h * W + w] += tdiff[y * W + x] / pool_size;
        }
      tdiff += H * W; bdiff += H * W;
    }
  }
  void forward(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    const float* bd = bottom[0]->data.data();
    float* td = top[0]->data.data();
    const size_t n = bottom[0]->count();
    if (region == 0) {  // CrossChannelForward_cpu :108-152 (running window: add head, subtract tail)
      float* sd = scale.data.data();
      for (size_t i = 0; i < n; ++i) sd[i] = k;
      const int HW = H * W;
      vector<float> padded((size_t)(C + size - 1) * HW, 0.f);
      const float aos = alpha / size;
      for (int nn = 0; nn < num; ++nn) {
        const float* b = bd + (size_t)nn * C * HW;
        float* s = sd + (size_t)nn * C * HW;
        for (size_t i = 0; i < (size_t)C * HW; ++i) padded[(size_t)pre_pad * HW + i] = b[i] * b[i];
        for (int c = 0; c < size; ++c) cpu_axpy(HW, aos, padded.data() + (size_t)c * HW, s);
        for (int c = 1; c < C; ++c) {
          memcpy(s + (size_t)c * HW, s + (size_t)(c - 1) * HW, HW * sizeof(float));
          cpu_axpy(HW, aos, padded.data() + (size_t)(c + size - 1) * HW, s + (size_t)c * HW);
          cpu_axpy(HW, -aos, padded.data() + (size_t)(c - 1) * HW, s + (size_t)c * HW);
        }
      }
      for (size_t i = 0; i < n; ++i) td[i] = std::pow(sd[i], -beta);
      for (size_t i = 0; i < n; ++i) td[i] = td[i] * bd[i];
    } else {            // WithinChannelForward :155-162
      float* sq = square_out.data.data();
      // square layer: Power(power=2, scale=1, shift=0) -> copy then powx(x, 2) (power_layer.cpp:30-40)
      for (size_t i = 0; i < n; ++i) sq[i] = std::pow(bd[i], 2.f);
      avepool_fwd(sq, pool_out.data.data());
      // power layer: (shift=1 + scale=alpha * x)^(-beta): copy, scal, add_scalar, powx
      float* po = power_out.data.data();
      const float* pl = pool_out.data.data();
      for (size_t i = 0; i < n; ++i) {
        float v = pl[i];
        v *= alpha;
        v += 1.f;
        po[i] = std::pow(v, -beta);
      }
      for (size_t i = 0; i < n; ++i) td[i] = bd[i] * po[i];   // Eltwise PROD
    }
  }
  void backward(const vector<Blob*>& top, const vector<bool>& pd, const vector<Blob*>& bottom) override {
    if (!pd[0]) return;
    const float* tdiff = top[0]->diff.data();
    const float* tdata = top[0]->data.data();
    const float* bdata = bottom[0]->data.data();
    float* bdiff = bottom[0]->diff.data();
    const size_t n = bottom[0]->count();
    if (region == 0) {  // CrossChannelBackward_cpu :180-232
      const float* sd = scale.data.data();
      const int HW = H * W;
      vector<float> padded_ratio((size_t)(C + size - 1) * HW, 0.f), accum(HW), accum_tb(HW);
      const float cache_ratio = (float)(2. * alpha * beta / size);
      for (size_t i = 0; i < n; ++i) bdiff[i] = std::pow(sd[i], -beta);
      for (size_t i = 0; i < n; ++i) bdiff[i] = tdiff[i] * bdiff[i];
      const int inverse_pre_pad = size - (size + 1) / 2;
      for (int nn = 0; nn < num; ++nn) {
        const size_t off = (size_t)nn * C * HW;
        float* pr = padded_ratio.data() + (size_t)inverse_pre_pad * HW;
        for (size_t i = 0; i < (size_t)C * HW; ++i) pr[i] = tdiff[off + i] * tdata[off + i];
        for (size_t i = 0; i < (size_t)C * HW; ++i) pr[i] = pr[i] / sd[off + i];
        std::fill(accum.begin(), accum.end(), 0.f);
        for (int c = 0; c < size - 1; ++c) cpu_axpy(HW, 1.f, padded_ratio.data() + (size_t)c * HW, accum.data());
        for (int c = 0; c < C; ++c) {
          cpu_axpy(HW, 1.f, padded_ratio.data() + (size_t)(c + size - 1) * HW, accum.data());
          for (int i = 0; i < HW; ++i) accum_tb[i] = bdata[off + (size_t)c * HW + i] * accum[i];
          cpu_axpy(HW, -cache_ratio, accum_tb.data(), bdiff + off + (size_t)c * HW);
          cpu_axpy(HW, -1.f, padded_ratio.data() + (size_t)c * HW, accum.data());
        }
      }
    } else {            // WithinChannelBackward :235-247
      // product (stable): d(product_input) = power_out * dy ; d(power_out) = x * dy
      float* pid = product_input_diff.data();
      float* pod = power_out.diff.data();
      const float* po = power_out.data.data();
      for (size_t i = 0; i < n; ++i) { pid[i] = po[i]; pid[i] = pid[i] * tdiff[i]; }
      for (size_t i = 0; i < n; ++i) { pod[i] = bdata[i]; pod[i] = pod[i] * tdiff[i]; }
      // power layer backward, general branch (power_layer.cpp:73-86): diff_scale = -beta*alpha
      //   bdiff = copy(x); scal(alpha); add_scalar(1); div(y, bdiff); scal(diff_scale); mul(top_diff)
      const float diff_scale = (-beta) * alpha;
      float* pld = pool_out.diff.data();
      const float* pl = pool_out.data.data();
      for (size_t i = 0; i < n; ++i) {
        float v = pl[i];
        v *= alpha;
        v += 1.f;
        v = po[i] / v;
        if (diff_scale != 1.f) v *= diff_scale;
        pld[i] = pod[i] * v;
      }
      avepool_bwd(pld, square_out.diff.data());
      // square layer backward (power==2 special case :57-65): axpby(diff_scale*scale = 2, x, 0) then * top_diff
      float* sid = square_input_diff.data();
      const float* sqd = square_out.diff.data();
      for (size_t i = 0; i < n; ++i) {
        float v = 2.f * bdata[i];
        sid[i] = sqd[i] * v;
      }
      // split backward: sum of the two branch diffs (split_layer.cpp:41)
      for (size_t i = 0; i < n; ++i) bdiff[i] = pid[i] + sid[i];
    }
  }
};

// SoftmaxLayer forward (src/caffe/layers/softmax_layer.cpp:27-60), axis=1 with inner_num
static void softmax_forward(const float* bd, float* td, int outer, int channels, int inner) {
  const int dim = channels * inner;
  vector<float> sc(inner);
  memcpy(td, bd, (size_t)outer * dim * sizeof(float));
  for (int i = 0; i < outer; ++i) {
    float* t = td + (size_t)i * dim;
    memcpy(sc.data(), bd + (size_t)i * dim, inner * sizeof(float));
    for (int j = 0; j < channels; ++j)
      for (int k = 0; k < inner; ++k) sc[k] = std::max(sc[k], bd[(size_t)i * dim + j * inner + k]);
    for (int j = 0; j < channels; ++j)
      for (int k = 0; k < inner; ++k) t[j * inner + k] += -1.f * sc[k];        // gemm(-1, ones, scale, +1)
    for (int j = 0; j < dim; ++j) t[j] = std::exp(t[j]);
    for (int k = 0; k < inner; ++k) {                                          // gemv(Trans): sum over channels
      float acc = 0.f;
      for (int j = 0; j < channels; ++j) acc += t[j * inner + k];
      sc[k] = acc;
    }
    for (int j = 0; j < channels; ++j)
      for (int k = 0; k < inner; ++k) t[j * inner + k] = t[j * inner + k] / sc[k];
  }
}

// SoftmaxWithLossLayer (src/caffe/layers/softmax_loss_layer.cpp:11-119, loss_layer.cpp:8-23)
struct SoftmaxWithLossLayer : Layer {
  Blob prob;
  int outer, inner, channels;
  bool has_ignore, normalize;
  int ignore_label;
  using Layer::Layer;
  const char* type() const override { return "SoftmaxWithLoss"; }
  bool is_loss() const override { return true; }
  void setup(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    if (lp.loss_weight.empty()) lp.loss_weight.push_back(1.f);   // LossLayer::LayerSetUp
    OCHECK(bottom.size() == 2, "SoftmaxWithLoss needs (prediction, label)");
    OCHECK(bottom[0]->num() == bottom[1]->num(), "The data and label should have the same number.");
    int axis = lp.softmax_axis < 0 ? lp.softmax_axis + (int)bottom[0]->shape.size() : lp.softmax_axis;
    outer = (int)bottom[0]->count(0, axis);
    channels = bottom[0]->shape[axis];
    inner = (int)bottom[0]->count_from(axis + 1);
    OCHECK((size_t)outer * inner == bottom[1]->count(), "Number of labels must match number of predictions");
    prob.reshape(bottom[0]->shape);
    top[0]->reshape({});   // 0-axis scalar (loss_layer.cpp:19-22)
    has_ignore = lp.loss_param.has_ignore_label;
    ignore_label = lp.loss_param.ignore_label;
    normalize = lp.loss_param.normalize;
  }
  void forward(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    softmax_forward(bottom[0]->data.data(), prob.data.data(), outer, channels, inner);
    const float* p = prob.data.data();
    const float* label = bottom[1]->data.data();
    const int dim = channels * inner;
    int count = 0;
    float loss = 0.f;
    for (int i = 0; i < outer; ++i)
      for (int j = 0; j < inner; ++j) {
        const int lv = (int)label[i * inner + j];
        if (has_ignore && lv == ignore_label) continue;
        OCHECK(lv >= 0 && lv < channels, "label " << lv << " out of range");
        loss -= std::log(std::max(p[(size_t)i * dim + lv * inner + j], FLT_MIN));
        ++count;
      }
    top[0]->data[0] = normalize ? loss / count : loss / outer;
  }
  void backward(const vector<Blob*>& top, const vector<bool>& pd, const vector<Blob*>& bottom) override {
    OCHECK(pd.size() < 2 || !pd[1], "SoftmaxWithLoss Layer cannot backpropagate to label inputs.");
    if (!pd[0]) return;
    float* bd = bottom[0]->diff.data();
    const size_t n = prob.count();
    memcpy(bd, prob.data.data(), n * sizeof(float));
    const float* label = bottom[1]->data.data();
    const int dim = channels * inner;
    int count = 0;
    for (int i = 0; i < outer; ++i)
      for (int j = 0; j < inner; ++j) {
        const int lv = (int)label[i * inner + j];
        if (has_ignore && lv == ignore_label) {
          for (int c = 0; c < channels; ++c) bd[(size_t)i * dim + c * inner + j] = 0.f;
        } else {
          bd[(size_t)i * dim + lv * inner + j] -= 1.f;
          ++count;
        }
      }
    const float lw = top[0]->diff[0];
    cpu_scal(n, normalize ? lw / count : lw / outer, bd);
  }
};

// AccuracyLayer (src/caffe/layers/accuracy_layer.cpp:13-103), top-k with (value,index) ordering
struct AccuracyLayer : Layer {
  int outer, inner, labels, top_k;
  using Layer::Layer;
  const char* type() const override { return "Accuracy"; }
  void setup(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    top_k = (int)lp.accuracy_param.top_k;
    int axis = lp.accuracy_param.axis < 0 ? lp.accuracy_param.axis + (int)bottom[0]->shape.size() : lp.accuracy_param.axis;
    outer = (int)bottom[0]->count(0, axis);
    labels = bottom[0]->shape[axis];
    inner = (int)bottom[0]->count_from(axis + 1);
    OCHECK(top_k <= labels, "top_k must be less than or equal to the number of classes.");
    OCHECK((size_t)outer * inner == bottom[1]->count(), "Number of labels must match number of predictions");
    top[0]->reshape({});
  }
  void forward(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    float acc = 0.f;
    const float* bd = bottom[0]->data.data();
    const float* bl = bottom[1]->data.data();
    const int dim = labels * inner;
    int count = 0;
    for (int i = 0; i < outer; ++i)
      for (int j = 0; j < inner; ++j) {
        const int lv = (int)bl[i * inner + j];
        if (lp.accuracy_param.has_ignore_label && lv == lp.accuracy_param.ignore_label) continue;
        vector<std::pair<float, int>> v;
        for (int kk = 0; kk < labels; ++kk) v.push_back({bd[(size_t)i * dim + kk * inner + j], kk});
        std::partial_sort(v.begin(), v.begin() + top_k, v.end(), std::greater<std::pair<float, int>>());
        for (int kk = 0; kk < top_k; ++kk)
          if (v[kk].second == lv) { ++acc; break; }
        ++count;
      }
    top[0]->data[0] = acc / count;
  }
  void backward(const vector<Blob*>&, const vector<bool>& pd, const vector<Blob*>&) override {
    for (bool b : pd) OCHECK(!b, "Accuracy layer cannot backpropagate");
  }
};

static std::shared_ptr<Layer> create_layer(const cp::LayerParameter& lp) {
  const string& t = lp.type;
  if (t == "JavaData") return std::make_shared<JavaDataLayer>(lp);
  if (t == "Convolution") return std::make_shared<ConvolutionLayer>(lp);
  if (t == "InnerProduct") return std::make_shared<InnerProductLayer>(lp);
  if (t == "Pooling") return std::make_shared<PoolingLayer>(lp);
  if (t == "ReLU") return std::make_shared<ReLULayer>(lp);
  if (t == "Dropout") return std::make_shared<DropoutLayer>(lp);
  if (t == "LRN") return std::make_shared<LRNLayer>(lp);
  if (t == "SoftmaxWithLoss") return std::make_shared<SoftmaxWithLossLayer>(lp);
  if (t == "Accuracy") return std::make_shared<AccuracyLayer>(lp);
  if (t == "Split") return std::make_shared<SplitLayer>(lp);
  OCHECK(false, "layer type '" << t << "' is outside the supported hot path (layer " << lp.name << ")");
  return nullptr;
}

// ------------------------------------------------------------------------------------------
// InsertSplits (src/caffe/util/insert_splits.cpp:12-137)
// ------------------------------------------------------------------------------------------
static string split_layer_name(const string& layer, const string& blob, int idx) {
  std::ostringstream s;
  s << blob << "_" << layer << "_" << idx << "_split";
  return s.str();
}
static string split_blob_name(const string& layer, const string& blob, int idx, int split) {
  std::ostringstream s;
  s << blob << "_" << layer << "_" << idx << "_split_" << split;
  return s.str();
}
static vector<cp::LayerParameter> insert_splits(const vector<cp::LayerParameter>& layers) {
  typedef std::pair<int, int> PI;
  std::map<string, PI> last_top;
  std::map<PI, PI> bottom_src;
  std::map<PI, int> top_count, split_idx;
  std::map<PI, float> top_lw;
  for (int i = 0; i < (int)layers.size(); ++i) {
    const cp::LayerParameter& lp = layers[i];
    for (int j = 0; j < (int)lp.bottom.size(); ++j) {
      auto it = last_top.find(lp.bottom[j]);
      OCHECK(it != last_top.end(), "Unknown bottom blob '" << lp.bottom[j] << "' (layer '" << lp.name << "')");
      bottom_src[{i, j}] = it->second;
      ++top_count[it->second];
    }
    for (int j = 0; j < (int)lp.top.size(); ++j) last_top[lp.top[j]] = {i, j};
    const int last_loss = std::min((int)lp.loss_weight.size(), (int)lp.top.size());
    for (int j = 0; j < last_loss; ++j) {
      PI t = last_top[lp.top[j]];
      top_lw[t] = lp.loss_weight[j];
      if (top_lw[t]) ++top_count[t];
    }
  }
  vector<cp::LayerParameter> out;
  for (int i = 0; i < (int)layers.size(); ++i) {
    cp::LayerParameter lp = layers[i];
    for (int j = 0; j < (int)lp.bottom.size(); ++j) {
      PI t = bottom_src[{i, j}];
      if (top_count[t] > 1) lp.bottom[j] = split_blob_name(layers[t.first].name, lp.bottom[j], t.second, split_idx[t]++);
    }
    vector<cp::LayerParameter> splits;
    for (int j = 0; j < (int)lp.top.size(); ++j) {
      PI t{i, j};
      const int sc = top_count[t];
      if (sc > 1) {
        cp::LayerParameter sp;
        sp.bottom.push_back(lp.top[j]);
        sp.name = split_layer_name(lp.name, lp.top[j], j);
        sp.type = "Split";
        const float lw = top_lw[t];
        for (int k = 0; k < sc; ++k) {
          sp.top.push_back(split_blob_name(lp.name, lp.top[j], j, k));
          if (lw) sp.loss_weight.push_back(k == 0 ? lw : 0.f);
        }
        if (lw) { lp.loss_weight.clear(); split_idx[t]++; }
        splits.push_back(sp);
      }
    }
    out.push_back(lp);
    for (auto& s : splits) out.push_back(s);
  }
  return out;
}

// ------------------------------------------------------------------------------------------
// Net (src/caffe/net.cpp:40-284 Init, 385-562 Append*, 565-581 ForwardFromTo, 635-645 BackwardFromTo,
// 737-766 ShareTrainedLayersWith, 983-1008 Update / ClearParamDiffs)
// ------------------------------------------------------------------------------------------
struct Net {
  string name;
  int phase;
  vector<std::shared_ptr<Layer>> layers;
  vector<string> layer_names;
  vector<BlobP> blobs;
  vector<string> blob_names;
  vector<bool> blob_need_backward;
  vector<vector<Blob*>> bottom_vecs, top_vecs;
  vector<vector<int>> bottom_ids, top_ids;
  vector<vector<bool>> bottom_need_backward;
  vector<bool> layer_need_backward;
  vector<Blob*> learnable_params;
  vector<float> params_lr, params_decay;
  vector<Blob*> output_blobs;
  vector<int> output_blob_indices;

  Net(const cp::NetParameter& in, int phase_) : phase(phase_) {
    cp::NetState st = in.state;
    st.phase = phase_;
    vector<cp::LayerParameter> param = insert_splits(cp::filter_net(in, st));
    name = in.name;
    OCHECK(in.input.empty(), "net-level `input` blobs are not on the supported path");
    std::map<string, int> name_to_idx;
    std::set<string> available;
    const int L = (int)param.size();
    bottom_vecs.resize(L); top_vecs.resize(L); bottom_ids.resize(L); top_ids.resize(L);
    bottom_need_backward.resize(L);
    for (int li = 0; li < L; ++li) {
      cp::LayerParameter lp = param[li];
      if (!lp.has_phase) { lp.phase = phase; lp.has_phase = true; }
      std::shared_ptr<Layer> layer = create_layer(lp);
      layer->phase = lp.phase;
      layers.push_back(layer);
      layer_names.push_back(lp.name);
      bool need_backward = false;
      for (int b = 0; b < (int)lp.bottom.size(); ++b) {     // AppendBottom :437-467
        const string& bn = lp.bottom[b];
        OCHECK(available.count(bn), "Unknown bottom blob '" << bn << "' (layer '" << lp.name << "')");
        const int id = name_to_idx[bn];
        bottom_vecs[li].push_back(blobs[id].get());
        bottom_ids[li].push_back(id);
        available.erase(bn);
        bool pd = true;
        if (!lp.propagate_down.empty()) pd = lp.propagate_down[b];
        const bool nb = blob_need_backward[id] && pd;
        bottom_need_backward[li].push_back(nb);
        need_backward |= blob_need_backward[id];
      }
      for (int t = 0; t < (int)lp.top.size(); ++t) {        // AppendTop :385-434
        const string& tn = lp.top[t];
        if (t < (int)lp.bottom.size() && lp.bottom[t] == tn) {   // in-place
          const int id = name_to_idx[tn];
          top_vecs[li].push_back(blobs[id].get());
          top_ids[li].push_back(id);
        } else {
          OCHECK(!name_to_idx.count(tn), "Top blob '" << tn << "' produced by multiple sources.");
          blobs.push_back(std::make_shared<Blob>());
          const int id = (int)blobs.size() - 1;
          blob_names.push_back(tn);
          blob_need_backward.push_back(false);
          name_to_idx[tn] = id;
          top_vecs[li].push_back(blobs[id].get());
          top_ids[li].push_back(id);
        }
        available.insert(tn);
      }
      layer->setup(bottom_vecs[li], top_vecs[li]);
      layer->set_loss_weights(top_vecs[li]);
      // AppendParam :470-562 (no sharing by name on this path)
      const int nparam = (int)layer->blobs.size();
      OCHECK((int)lp.param.size() <= nparam, "Too many params specified for layer " << lp.name);
      for (int p = 0; p < nparam; ++p) {
        cp::ParamSpec spec;
        if (p < (int)lp.param.size()) spec = lp.param[p];
        OCHECK(spec.name.empty(), "named (shared) params are not on the supported path");
        const bool pnb = spec.lr_mult != 0.f;
        need_backward |= pnb;
        layer->param_propagate_down[p] = pnb;
        learnable_params.push_back(layer->blobs[p].get());
        params_lr.push_back(spec.lr_mult);
        params_decay.push_back(spec.decay_mult);
      }
      layer_need_backward.push_back(need_backward);
      if (need_backward)
        for (int id : top_ids[li]) blob_need_backward[id] = true;
    }
    // backward pruning :191-247
    std::set<string> under_loss, skip_bp;
    for (int li = L - 1; li >= 0; --li) {
      bool contributes = false, skip_pd = true;
      for (int t = 0; t < (int)top_vecs[li].size(); ++t) {
        const string& bn = blob_names[top_ids[li][t]];
        if ((t < (int)layers[li]->loss.size() && layers[li]->loss[t]) || under_loss.count(bn)) contributes = true;
        if (!skip_bp.count(bn)) skip_pd = false;
        if (contributes && !skip_pd) break;
      }
      if (layer_need_backward[li] && skip_pd) {
        layer_need_backward[li] = false;
        for (size_t b = 0; b < bottom_need_backward[li].size(); ++b) bottom_need_backward[li][b] = false;
      }
      if (!contributes) layer_need_backward[li] = false;
      for (int b = 0; b < (int)bottom_vecs[li].size(); ++b) {
        const string& bn = blob_names[bottom_ids[li][b]];
        if (contributes) under_loss.insert(bn); else bottom_need_backward[li][b] = false;
        if (!bottom_need_backward[li][b]) skip_bp.insert(bn);
      }
    }
    if (in.force_backward) {   // :248-265; AllowForceBackward is false only for a loss layer's label bottom
      for (int li = 0; li < L; ++li) {
        layer_need_backward[li] = true;
        for (size_t b = 0; b < bottom_need_backward[li].size(); ++b) {
          const bool allow = !(layers[li]->is_loss() && b == 1) && string(layers[li]->type()) != "Accuracy";
          bottom_need_backward[li][b] = bottom_need_backward[li][b] || allow;
          blob_need_backward[bottom_ids[li][b]] = blob_need_backward[bottom_ids[li][b]] || bottom_need_backward[li][b];
        }
        for (size_t p = 0; p < layers[li]->param_propagate_down.size(); ++p) layers[li]->param_propagate_down[p] = true;
      }
    }
    for (const string& bn : available) {   // std::set order => sorted by name (:266-274)
      output_blobs.push_back(blobs[name_to_idx[bn]].get());
      output_blob_indices.push_back(name_to_idx[bn]);
    }
  }
  float forward_from_to(int start, int end) {
    float loss = 0.f;
    for (int i = start; i <= end; ++i) loss += layers[i]->forward_with_loss(bottom_vecs[i], top_vecs[i]);
    return loss;
  }
  void backward_from_to(int start, int end) {
    for (int i = start; i >= end; --i)
      if (layer_need_backward[i]) layers[i]->backward(top_vecs[i], bottom_need_backward[i], bottom_vecs[i]);
  }
  void clear_param_diffs() {
    for (Blob* b : learnable_params) std::fill(b->diff.begin(), b->diff.end(), 0.f);
  }
  void update() {   // Blob::Update (blob.cpp:154-162): data += -1 * diff
    for (Blob* b : learnable_params) cpu_axpy(b->count(), -1.f, b->diff.data(), b->data.data());
  }
  // net.cpp:737-766: target layers with the same name share the source's blobs
  void share_trained_layers_with(const Net& other) {
    for (size_t i = 0; i < other.layers.size(); ++i) {
      const string& src = other.layer_names[i];
      size_t t = 0;
      while (t != layer_names.size() && layer_names[t] != src) ++t;
      if (t == layer_names.size()) continue;
      OCHECK(layers[t]->blobs.size() == other.layers[i]->blobs.size(), "Incompatible number of blobs for layer " << src);
      for (size_t j = 0; j < layers[t]->blobs.size(); ++j) {
        OCHECK(layers[t]->blobs[j]->shape == other.layers[i]->blobs[j]->shape, "Cannot share param " << j << " of layer " << src);
        layers[t]->blobs[j] = other.layers[i]->blobs[j];
      }
    }
  }
};

// ------------------------------------------------------------------------------------------
// SGDSolver (src/caffe/solver.cpp:44-190 Init, 193-282 Step, 413-444 TestAndStoreResult;
// src/caffe/solvers/sgd_solver.cpp:27-63, 66-78, 102-116, 145-172, 207-222)
// ------------------------------------------------------------------------------------------
struct Solver {
  cp::SolverParameter param;
  std::unique_ptr<Net> net, test_net;
  int iter = 0, current_step = 0;
  vector<vector<float>> history;
  float last_loss = 0.f, smoothed_loss = 0.f;
  vector<float> losses;
  explicit Solver(const cp::SolverParameter& p) : param(p) {
    OCHECK(param.type == "SGD", "only the SGD solver is on the path (libccaffe/ccaffe.cpp:131)");
    if (param.random_seed >= 0) rng_seed((uint32_t)param.random_seed);   // solver.cpp:52-54
    OCHECK(param.has_net_param || param.has_train_net_param, "SolverParameter must embed net_param (ProtoLoader.scala:40)");
    const cp::NetParameter& np = param.has_train_net_param ? param.train_net_param : param.net_param;
    net.reset(new Net(np, cp::TRAIN));
    if (param.has_net_param) {   // InitTestNets :111-190 (generic net => one test net)
      test_net.reset(new Net(param.net_param, cp::TEST));
    } else if (!param.test_net_param.empty()) {
      test_net.reset(new Net(param.test_net_param[0], cp::TEST));
    }
    for (Blob* b : net->learnable_params) history.emplace_back(b->count(), 0.f);   // PreSolve
  }
  float get_learning_rate() {
    const string& pol = param.lr_policy;
    float rate;
    if (pol == "fixed") {
      rate = param.base_lr;
    } else if (pol == "step") {
      current_step = iter / param.stepsize;
      rate = (float)(param.base_lr * std::pow(param.gamma, current_step));
    } else if (pol == "exp") {
      rate = (float)(param.base_lr * std::pow(param.gamma, iter));
    } else if (pol == "inv") {
      rate = param.base_lr * std::pow(1.f + param.gamma * iter, -param.power);
    } else if (pol == "multistep") {
      if (current_step < (int)param.stepvalue.size() && iter >= param.stepvalue[current_step]) current_step++;
      rate = (float)(param.base_lr * std::pow(param.gamma, current_step));
    } else if (pol == "poly") {
      rate = param.base_lr * std::pow(1.f - ((float)iter / (float)param.max_iter), param.power);
    } else if (pol == "sigmoid") {
      rate = param.base_lr * (1.f / (1.f + std::exp(-param.gamma * ((float)iter - (float)param.stepsize))));
    } else {
      OCHECK(false, "Unknown learning rate policy: " << pol);
      rate = 0.f;
    }
    return rate;
  }
  void apply_update() {
    const float rate = get_learning_rate();
    OCHECK(param.clip_gradients < 0, "clip_gradients is not on the supported path");
    OCHECK(param.iter_size == 1, "iter_size != 1 is not on the supported path");
    OCHECK(param.regularization_type == "L2", "only L2 regularization is on the path");
    for (size_t i = 0; i < net->learnable_params.size(); ++i) {
      Blob* b = net->learnable_params[i];
      const size_t n = b->count();
      const float local_decay = param.weight_decay * net->params_decay[i];
      if (local_decay) cpu_axpy(n, local_decay, b->data.data(), b->diff.data());        // Regularize
      const float local_rate = rate * net->params_lr[i];
      cpu_axpby(n, local_rate, b->diff.data(), param.momentum, history[i].data());      // ComputeUpdateValue
      memcpy(b->diff.data(), history[i].data(), n * sizeof(float));
    }
    net->update();
  }
  void step(int iters) {
    const int start_iter = iter, stop_iter = iter + iters;
    const int average_loss = param.average_loss;
    losses.clear();
    smoothed_loss = 0.f;
    while (iter < stop_iter) {
      net->clear_param_diffs();
      float loss = net->forward_from_to(0, (int)net->layers.size() - 1);
      net->backward_from_to((int)net->layers.size() - 1, 0);
      if ((int)losses.size() < average_loss) {
        losses.push_back(loss);
        const int size = (int)losses.size();
        smoothed_loss = (smoothed_loss * (size - 1) + loss) / size;
      } else {
        const int idx = (iter - start_iter) % average_loss;
        smoothed_loss += (loss - losses[idx]) / average_loss;
        losses[idx] = loss;
      }
      last_loss = loss;
      apply_update();
      ++iter;
    }
  }
  void test_and_store(int num_steps, vector<float>* score) {
    OCHECK(test_net, "no test net");
    test_net->share_trained_layers_with(*net);
    for (int i = 0; i < num_steps; ++i) {
      test_net->forward_from_to(0, (int)test_net->layers.size() - 1);
      int idx = 0;
      for (Blob* r : test_net->output_blobs)
        for (size_t kk = 0; kk < r->count(); ++kk) {
          if (i == 0) score->push_back(r->data[kk]); else (*score)[idx++] += r->data[kk];
        }
    }
  }
};

// ------------------------------------------------------------------------------------------
// C ABI (libccaffe/ccaffe.h:5-77, libccaffe/ccaffe.cpp:22-304)
// ------------------------------------------------------------------------------------------
struct caffenet_state {
  Net* net = nullptr;
  Solver* solver = nullptr;
  Net* testnet = nullptr;
  vector<float> test_score;
  vector<char> buffer;
};

extern "C" {
error_callback_t global_caffe_error_callback;

void init_logging(const char*, int) {}
void set_basepath(const char* path) { if (chdir(path) != 0) perror("set_basepath"); }
int get_int_size() { return (int)sizeof(int); }
int get_dtype_size() { return (int)sizeof(float); }

caffenet_state* create_state() { return new caffenet_state(); }
void destroy_state(caffenet_state* s) {
  if (s->solver) delete s->solver; else if (s->net) delete s->net;
  delete s;
}
void load_solver_from_protobuf(caffenet_state* s, const char* p, int n) {
  cp::SolverParameter sp;
  sp.parse(cp::Reader(p, (size_t)n));
  s->solver = new Solver(sp);
  s->net = s->solver->net.get();
  s->testnet = s->solver->test_net.get();
}
void load_net_from_protobuf(caffenet_state* s, const char* p, int n) {
  cp::NetParameter np;
  np.parse(cp::Reader(p, (size_t)n));
  s->net = new Net(np, np.has_state ? np.state.phase : cp::TEST);   // NetState.phase default TEST
}
static int set_data_callback(Net* net, int idx, java_callback_t cb) {
  if (idx < 0 || idx >= (int)net->layers.size()) return -1;
  JavaDataLayer* l = dynamic_cast<JavaDataLayer*>(net->layers[idx].get());
  if (!l) return -2;
  l->cb = cb;
  return 0;
}
int set_train_data_callback(caffenet_state* s, int idx, java_callback_t cb) { return set_data_callback(s->net, idx, cb); }
int set_test_data_callback(caffenet_state* s, int idx, java_callback_t cb) { return set_data_callback(s->testnet, idx, cb); }
void forward(caffenet_state* s) { s->net->forward_from_to(0, (int)s->net->layers.size() - 1); }
void backward(caffenet_state* s) { s->net->backward_from_to((int)s->net->layers.size() - 1, 0); }
int solver_step(caffenet_state* s, int step) { s->solver->step(step); return 0; }
void solver_test(caffenet_state* s, int num_steps) {
  s->test_score.clear();
  s->solver->test_and_store(num_steps, &s->test_score);
}
int num_test_scores(caffenet_state* s) { return (int)s->test_score.size(); }
float get_test_score(caffenet_state* s, int i) { return s->test_score[(size_t)i]; }
int num_layers(caffenet_state* s) { return (int)s->net->layers.size(); }
const char* layer_name(caffenet_state* s, int i) { return s->net->layers[(size_t)i]->lp.name.c_str(); }
int num_layer_weights(caffenet_state* s, int i) { return (int)s->net->layers[(size_t)i]->blobs.size(); }
int num_data_blobs(caffenet_state* s) { return (int)s->net->blobs.size(); }
void* get_data_blob(caffenet_state* s, int i) { return s->net->blobs[(size_t)i].get(); }
const char* get_data_blob_name(caffenet_state* s, int i) { return s->net->blob_names[(size_t)i].c_str(); }
void* get_weight_blob(caffenet_state* s, int l, int j) { return s->net->layers[(size_t)l]->blobs[(size_t)j].get(); }
float* get_data(void* b) { return ((Blob*)b)->data.data(); }
float* get_diff(void* b) { return ((Blob*)b)->diff.data(); }
int get_num_axes(void* b) { return (int)((Blob*)b)->shape.size(); }
int get_axis_shape(void* b, int a) { return ((Blob*)b)->dim(a); }
void set_global_error_callback(error_callback_t c) { global_caffe_error_callback = c; }
void set_mode_cpu() {}
void set_mode_gpu() {}   // the oracle is the CPU path; accepted so one harness drives both libraries
void set_device(int) {}

// ---- oracle/test extensions (same names as the product's sparknet_* superset where they overlap) ----
// Net::ForwardFromTo / BackwardFromTo (src/caffe/net.cpp:565-581, 635-645)
int sparknet_forward_range(caffenet_state* s, int start, int end) {
  if (start < 0 || end >= (int)s->net->layers.size() || start > end) return -1;
  s->net->forward_from_to(start, end);
  return 0;
}
int sparknet_backward_range(caffenet_state* s, int start, int end) {
  if (end < 0 || start >= (int)s->net->layers.size() || start < end) return -1;
  s->net->backward_from_to(start, end);
  return 0;
}
int sparknet_get_iter(caffenet_state* s) { return s->solver ? s->solver->iter : -1; }
float sparknet_last_loss(caffenet_state* s) { return s->solver ? s->solver->last_loss : 0.f; }
float sparknet_smoothed_loss(caffenet_state* s) { return s->solver ? s->solver->smoothed_loss : 0.f; }
// Use `mask` (count unsigned 0/1 values) for this Dropout layer instead of drawing one; NULL restores drawing.
int sparknet_inject_dropout_mask(caffenet_state* s, int layer_idx, const unsigned* mask, int count) {
  if (layer_idx < 0 || layer_idx >= (int)s->net->layers.size()) return -1;
  DropoutLayer* d = dynamic_cast<DropoutLayer*>(s->net->layers[(size_t)layer_idx].get());
  if (!d) return -2;
  if (!mask) { d->injected = false; return 0; }
  if (count != (int)d->mask.size()) return -3;
  d->mask.assign(mask, mask + count);
  d->injected = true;
  return 0;
}
// int32 argmax mask of a MAX pooling layer (PoolingLayer::max_idx_), for known-answer tests
const int* oracle_pool_mask(caffenet_state* s, int layer_idx) {
  PoolingLayer* p = dynamic_cast<PoolingLayer*>(s->net->layers[(size_t)layer_idx].get());
  return (p && p->method == 0) ? p->max_idx.data() : nullptr;
}
// Momentum history of learnable param i (SGDSolver::history_)
float* oracle_history(caffenet_state* s, int param_idx) { return s->solver->history[(size_t)param_idx].data(); }
int oracle_num_learnable(caffenet_state* s) { return (int)s->net->learnable_params.size(); }

// Weight averaging restated from the JVM side: WeightCollection.add = NDArray.plus (copy w0, then
// in-place fp32 add; src/main/java/libs/JavaNDArray.java:84-94, src/main/scala/libs/Net.scala:27-46)
// folded left over workers 0..n-1 in rank order, then scalarDivide by (float)n
// (JavaNDArray.java:108-117, src/main/scala/apps/CifarApp.scala:133-134).
void oracle_average(const float* const* replicas, int n, long count, float* out) {
  for (long i = 0; i < count; ++i) {
    float acc = replicas[0][i];
    for (int r = 1; r < n; ++r) acc = acc + replicas[r][i];
    out[i] = acc / (float)n;
  }
}
// Raw BLAS restatement, for the reference's literal GEMM/GEMV vectors (test_util_blas.cpp:20-131)
void oracle_gemm(int ta, int tb, int M, int N, int K, float alpha, const float* A, const float* B, float beta, float* C) {
  cpu_gemm(ta ? DoTrans : NoTrans, tb ? DoTrans : NoTrans, M, N, K, alpha, A, B, beta, C);
}
void oracle_gemv(int ta, int M, int N, float alpha, const float* A, const float* x, float beta, float* y) {
  cpu_gemv(ta ? DoTrans : NoTrans, M, N, alpha, A, x, beta, y);
}
void oracle_seed(unsigned s) { rng_seed(s); }
// Threads used by the row-parallel sgemm (timed baseline; results are thread-count independent).
void oracle_set_threads(int n) { g_threads = n < 1 ? 1 : n; }
// Timed-baseline only: route sgemm through a real BLAS (e.g. scipy's bundled OpenBLAS,
// symbol scipy_cblas_sgemm).  Returns 0 on success.  Parity tests never call this.
int oracle_use_blas(const char* so_path, const char* sgemm_symbol) {
  if (!so_path) { g_sgemm = nullptr; return 0; }
  void* h = dlopen(so_path, RTLD_NOW | RTLD_LOCAL);
  if (!h) return -1;
  void* f = dlsym(h, sgemm_symbol);
  if (!f) return -2;
  g_sgemm = (cblas_sgemm_t)f;
  return 0;
}
}  // extern "C"
