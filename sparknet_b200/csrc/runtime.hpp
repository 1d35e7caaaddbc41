// runtime.hpp — host runtime of the B200 libccaffe: blobs with Caffe's host/device mirror
// semantics, layers, the net graph builder and the SGD solver.  C++ only; every compute step is a
// launch from kernels.hpp.  There is no CPU compute path: set_mode_cpu() is accepted and ignored.
//
// Reference counterparts (what this replaces): caffe/src/caffe/{blob,syncedmem,net,solver}.cpp,
// caffe/src/caffe/solvers/sgd_solver.cpp, caffe/src/caffe/layers/*.{cpp,cu} (on-path subset),
// caffe/src/caffe/util/insert_splits.cpp.
#pragma once
#include <cuda_runtime.h>

#include <functional>
#include <map>
#include <memory>
#include <set>
#include <string>
#include <vector>

#include "caffe_proto.hpp"
#include "kernels.hpp"

extern "C" {
typedef void (*java_callback_t)(void* data, int batch_size, int num_dims, int* shape);
typedef void (*error_callback_t)(const char* msg);
}

namespace sn {
namespace cp = caffe_proto;

// Per-process device context (one GPU per process is the deployment model; set_device() picks it).
struct Context {
  int device = 0;
  bool initialised = false;
  cudaStream_t compute = nullptr;   // all kernels
  cudaStream_t copy = nullptr;      // JavaData H2D staging
  cudaStream_t side = nullptr;      // backward-filter kernels inside solver_step (see runtime.cpp side_begin)
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  bool side_used = false;           // work has been queued on `side` since the last join
  // scratch: [0] for the compute stream, [1] for the side stream (they run concurrently)
  float* workspace[2] = {nullptr, nullptr};
  size_t workspace_floats[2] = {0, 0};
  long launches = 0;
  long param_d2h_copies = 0, param_h2d_copies = 0;   // bulk transfers between a parameter arena and its pinned mirror
  int numa_node = -1;               // NUMA node of the GPU (sysfs), -1 unknown
  bool numa_pinned = false;         // the calling thread was restricted to that node's CPUs at init
  void init();
};
Context& ctx();       // initialises the device context on first use
Context& raw_ctx();   // the same object, without initialising it
float* alloc_guarded(size_t count, float** raw);   // cudaMalloc with zeroed guard bands on both sides

// ------------------------------------------------------------------------------------------
// Storage: one buffer with a device copy and a lazily created pinned host mirror.  State machine
// of caffe/src/caffe/syncedmem.cpp:25-153 (UNINITIALIZED / HEAD_AT_CPU / HEAD_AT_GPU / SYNCED).
// ------------------------------------------------------------------------------------------
struct ParamArena;
struct Storage {
  enum Head { UNINIT, AT_HOST, AT_DEVICE, SYNCED };
  size_t count = 0;
  float* dev = nullptr;
  float* dev_raw = nullptr;      // owned allocation: dev = dev_raw + kGuardFloats
  float* host = nullptr;
  bool owns_dev = true;
  bool owns_host = true;
  ParamArena* arena = nullptr;   // set for weight blobs: dev / host point into the net's arena and its pinned mirror
  // Every activation buffer is surrounded by guard bands: the implicit-GEMM gathers read padded filter
  // taps unconditionally (the value is then discarded), which may touch a few rows before/after a tensor.
  static constexpr size_t kGuardFloats = 4096;
  Head head = UNINIT;
  uint64_t gen = 0;              // bumped whenever the contents may have changed (host or device write access)
  explicit Storage(size_t n) : count(n) {}
  Storage(size_t n, float* arena_ptr) : count(n), dev(arena_ptr), owns_dev(false) {}
  ~Storage();
  const float* gpu();            // device copy current (uploads if the host copy is newer)
  float* mutable_gpu();          // + device copy becomes authoritative
  float* mutable_cpu();          // host copy current and authoritative (ccaffe.cpp:181-187 get_data/get_diff)
  void ensure_dev();
  void ensure_host();
};

// The weights of a net live in ONE device arena with ONE pinned host mirror of the same layout.  The JVM
// driver reads / writes weights blob by blob through get_data (Net.scala:132-172); moving the arena as a whole makes that
// one bulk D2H when the first blob is asked for after device-side updates (every other blob is then SYNCED and costs
// nothing), and one bulk H2D before the next device use of host-written weights.
struct ParamArena {
  float* dev = nullptr;
  float* host = nullptr;          // pinned, arena-sized
  size_t floats = 0;
  std::vector<Storage*> members;  // in offset order
  std::vector<size_t> offsets;
  void download();                // every AT_DEVICE member -> SYNCED, contiguous runs in one copy each; synchronises
  void upload();                  // every AT_HOST member -> SYNCED, contiguous runs in one copy each; asynchronous (compute stream)
};

struct Blob {
  std::vector<int> shape;
  std::shared_ptr<Storage> data, diff;
  size_t count() const {
    size_t c = 1;
    for (int d : shape) c *= (size_t)d;
    return c;
  }
  size_t count(int from, int to) const {
    size_t c = 1;
    for (int i = from; i < to; ++i) c *= (size_t)shape[(size_t)i];
    return c;
  }
  int dim(int i) const { return shape[(size_t)(i < 0 ? i + (int)shape.size() : i)]; }
  int num() const { return shape.size() > 0 ? shape[0] : 1; }
  int channels() const { return shape.size() > 1 ? shape[1] : 1; }
  int height() const { return shape.size() > 2 ? shape[2] : 1; }
  int width() const { return shape.size() > 3 ? shape[3] : 1; }
  void reshape(const std::vector<int>& s);
  void share_data(const Blob& o) { data = o.data; }
};
typedef std::shared_ptr<Blob> BlobP;

enum Precision { PREC_FFMA = 0, PREC_TF32 = 1 };

struct Net;
struct Layer {
  cp::LayerParameter lp;
  int phase = cp::TRAIN;
  Net* net = nullptr;
  std::vector<BlobP> blobs;
  std::vector<bool> param_propagate_down;
  std::vector<float> loss;
  explicit Layer(const cp::LayerParameter& p) : lp(p) {}
  virtual ~Layer();
  virtual const char* type() const = 0;
  virtual void setup(const std::vector<Blob*>& bottom, const std::vector<Blob*>& top) = 0;
  virtual void forward(const std::vector<Blob*>& bottom, const std::vector<Blob*>& top) = 0;
  virtual void backward(const std::vector<Blob*>& top, const std::vector<bool>& pd, const std::vector<Blob*>& bottom) = 0;
  virtual bool is_loss() const { return false; }
  void set_loss_weights(const std::vector<Blob*>& top);
  // tf32-rounded copy of blobs[0] for the TMA-fed tensor-core operands, re-rounded when the weights changed
  float* w_shadow = nullptr;
  bool w_shadow_in_arena = false;       // points into the net's shadow arena, which the fused SGD update keeps current
  uint64_t w_shadow_gen = ~0ull;
  const float* weights_tf32();
};

struct Net {
  std::string name;
  int phase;
  int precision = PREC_FFMA;
  bool overwrite_param_diffs = false;   // true inside solver_step: weight gradients overwrite (ClearParamDiffs folded away)
  // set by Solver::step: called by a layer right after it queued its backward-filter on the side stream, when everything
  // that reads the layer's weights in this iteration has been queued before the fork (Solver::update_layer_early)
  std::function<void(Layer*)> on_param_grads_ready;
  float* loss_log = nullptr;            // device slot for this iteration's loss (solver-owned), or null
  uint64_t rng_seed = 0;
  std::vector<std::shared_ptr<Layer>> layers;
  std::vector<std::string> layer_names;
  std::vector<BlobP> blobs;
  std::vector<std::string> blob_names;
  std::vector<bool> blob_need_backward;
  std::vector<std::vector<Blob*>> bottom_vecs, top_vecs;
  std::vector<std::vector<int>> bottom_ids, top_ids;
  std::vector<std::vector<bool>> bottom_need_backward;
  std::vector<bool> layer_need_backward;
  std::vector<Blob*> learnable_params;
  std::vector<float> params_lr, params_decay;
  std::vector<Blob*> output_blobs;
  std::vector<int> output_blob_indices;
  // parameter arena (weights | gradients), blobs start on 1024-float boundaries
  float* arena_w = nullptr;
  float* arena_g = nullptr;
  float* arena_s = nullptr;             // tf32-rounded shadow of arena_w (same layout), written by sgd_update
  size_t arena_floats = 0;
  std::vector<size_t> param_offsets;
  bool owns_arena = false;
  ParamArena param_arena;               // pinned host mirror of arena_w behind every get_weight_blob handle
  const Net* shares_with = nullptr;     // TEST net of a solver: layers of the same name use the TRAIN net's blobs (no own arena)

  // share_from (nullable): net.cpp:737-766 ShareTrainedLayersWith applied while building — same-named layers adopt the
  // other net's parameter blobs instead of allocating and filling their own
  Net(const cp::NetParameter& param, int phase, uint64_t seed, const Net* share_from = nullptr);
  ~Net();
  void params_to_device();              // host-written weights -> device in as few copies as possible
  void mark_shadows_current();
  void prepack_filters();               // all stale tf32 filter copies of the convolution layers in one launch (Solver::step)
  void forward_from_to(int start, int end);
  void backward_from_to(int start, int end);
  void share_trained_layers_with(const Net& other);
  float total_loss();                   // sum over loss tops of dot(data, loss weights) — reads back (syncs)
};

struct Solver {
  cp::SolverParameter param;
  std::unique_ptr<Net> net, test_net;
  int iter = 0, current_step = 0;
  float* arena_h = nullptr;             // momentum history, same layout as the net's arena
  SgdChunk* chunk_table = nullptr;      // device, one entry per 1024 floats
  float* loss_hist = nullptr;           // device ring of per-iteration losses
  float* loss_hist_host = nullptr;      // pinned mirror: every iteration copies its loss here asynchronously (4 bytes D2H per step)
  static constexpr int kLossHist = 4096;
  float smoothed_loss = 0.f;
  // Steady-state iterations are captured once into a CUDA graph and replayed.  The first iteration of
  // every solver_step call runs eagerly (it picks up whatever the caller changed through the ABI in between: weights
  // written through get_data pointers, host-side heads, stale shadows), later ones replay the graph recorded for the
  // current input buffers.  Per-iteration scalars live in `scalars_dev` (see kernels.hpp StepScalars).
  struct GraphEntry {
    std::vector<const void*> key;       // device pointers of the data layers' tops the recording reads
    cudaGraphExec_t exec = nullptr;
    long launches = 0;                  // kernel launches inside (bookkeeping for sparknet_kernel_launches)
  };
  std::vector<GraphEntry> graphs;
  StepScalars* scalars_dev = nullptr;
  StepScalars scalars_last{};
  bool scalars_published = false;
  int graph_mode = -1;                  // -1: SPARKNET_GRAPH environment variable (default on), 0 off, 1 on
  int eager_iters = 0;                  // eager iterations since the graphs were last dropped (capture needs >= 2)
  uint64_t graphs_epoch = 0;            // value of the configuration epoch the recordings belong to
  long graph_replays = 0;
  int programmatic_edges = 0;           // kernel-to-kernel edges of the last recording that became programmatic (PDL)
  long step_calls = 0;
  uint64_t touch_at_exit = 0;           // value of the outside-touch counter when the last step() returned
  void drop_graphs();
  explicit Solver(const cp::SolverParameter& p);
  ~Solver();
  float get_learning_rate();
  void step(int iters);
  void step_profiled(std::vector<float>* fwd_ms, std::vector<float>* bwd_ms, float* update_ms);
  // SGD update of ONE layer's parameters on the side stream, behind its backward-filter: the update of the big inner-product
  // layers (fc6 alone is 62 % of CaffeNet's parameters) then runs under the backward pass of the layers below instead of
  // after it.  sgd_done marks the parameters already updated in this iteration; update_rest finishes the others.
  std::map<const Blob*, size_t> param_index;
  std::vector<char> sgd_done;
  void update_layer_early(Layer* l);
  void update_rest(cudaStream_t s);
  std::vector<int> profile_launches_fwd, profile_launches_bwd;   // kernels each layer launched in the last profiled iteration
  void test_and_store(int num_steps, std::vector<float>* score);
};

}  // namespace sn
