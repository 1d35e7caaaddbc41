// runtime.cpp — see runtime.hpp.  Host logic only; all arithmetic is in the CUDA kernels.
#include "runtime.hpp"

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <sstream>
#include <atomic>
#include <functional>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <sched.h>
#include <ctype.h>

namespace sn {

error_callback_t g_error_callback = nullptr;

void fatal(const char* fmt, ...) {
  char buf[2048];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  fprintf(stderr, "libccaffe(b200): FATAL: %s\n", buf);
  fflush(stderr);
  if (g_error_callback) g_error_callback(buf);
  abort();   // reference behaviour: glog CHECK / LOG(FATAL) aborts the process (SURVEY §8b "Errors")
}

#define SN_CHECK(cond, ...)            \
  do {                                 \
    if (!(cond)) sn::fatal(__VA_ARGS__); \
  } while (0)

// ------------------------------------------------------------------------------------------
// Context
// ------------------------------------------------------------------------------------------
Context& raw_ctx() {
  static Context c;
  return c;
}
Context& ctx() {
  Context& c = raw_ctx();
  if (!c.initialised) c.init();
  return c;
}
// Keep the thread that drives this GPU (and the copy threads / pinned buffers it creates) on the GPU's NUMA node: the
// input path moves every minibatch through host memory twice (callback write, DMA read).  Only when the process may
// run on several nodes and the GPU's node is one of them; SPARKNET_NUMA_PIN=0 turns it off.
static void numa_pin(Context& c) {
  char bus[32] = {0};
  if (cudaDeviceGetPCIBusId(bus, (int)sizeof(bus), c.device) != cudaSuccess) { cudaGetLastError(); return; }
  for (char* p = bus; *p; ++p) *p = (char)tolower(*p);
  char path[128];
  snprintf(path, sizeof(path), "/sys/bus/pci/devices/%s/numa_node", bus);
  FILE* f = fopen(path, "r");
  if (!f) return;
  int node = -1;
  if (fscanf(f, "%d", &node) != 1) node = -1;
  fclose(f);
  c.numa_node = node;
  const char* e = getenv("SPARKNET_NUMA_PIN");
  if (node < 0 || (e && atoi(e) == 0)) return;
  snprintf(path, sizeof(path), "/sys/devices/system/node/node%d/cpulist", node);
  f = fopen(path, "r");
  if (!f) return;
  char list[4096] = {0};
  const bool ok = fgets(list, sizeof(list), f) != nullptr;
  fclose(f);
  if (!ok) return;
  cpu_set_t node_cpus, cur, both;
  CPU_ZERO(&node_cpus);
  for (char* p = list; *p && *p != '\n';) {                 // "0-31,64-95"
    char* end = nullptr;
    long a = strtol(p, &end, 10), b = a;
    if (end == p) break;
    if (*end == '-') { p = end + 1; b = strtol(p, &end, 10); }
    for (long k = a; k <= b && k < CPU_SETSIZE; ++k) CPU_SET((int)k, &node_cpus);
    p = (*end == ',') ? end + 1 : end;
  }
  if (sched_getaffinity(0, sizeof(cur), &cur) != 0) return;
  CPU_AND(&both, &cur, &node_cpus);
  if (CPU_COUNT(&both) == 0 || CPU_COUNT(&both) == CPU_COUNT(&cur)) return;   // not allowed there / already confined to it
  if (sched_setaffinity(0, sizeof(both), &both) == 0) c.numa_pinned = true;
}
void Context::init() {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0)
    fatal("no CUDA device available (%s). This library has no CPU path.", cudaGetErrorString(e));
  SN_CUDA(cudaSetDevice(device));
  numa_pin(*this);
  SN_CUDA(cudaStreamCreateWithFlags(&compute, cudaStreamNonBlocking));
  SN_CUDA(cudaStreamCreateWithFlags(&copy, cudaStreamNonBlocking));
  SN_CUDA(cudaStreamCreateWithFlags(&side, cudaStreamNonBlocking));
  SN_CUDA(cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming));
  SN_CUDA(cudaEventCreateWithFlags(&ev_join, cudaEventDisableTiming));
  initialised = true;
}
void bump_config_epoch();
static int g_ws_sel = 0;            // which scratch buffer get_workspace hands out: 1 while launches go to the side stream
float* get_workspace(size_t nfloats) {
  Context& c = ctx();
  const int k = g_ws_sel;
  if (nfloats > c.workspace_floats[k]) {
    SN_CUDA(cudaStreamSynchronize(c.compute));
    SN_CUDA(cudaStreamSynchronize(c.side));
    if (c.workspace[k]) SN_CUDA(cudaFree(c.workspace[k]));
    size_t want = std::max(nfloats, c.workspace_floats[k] * 2);
    SN_CUDA(cudaMalloc(&c.workspace[k], want * sizeof(float)));
    c.workspace_floats[k] = want;
    // recorded iterations (Solver::step) carry the old pointer in their split-reduction and bias-partial launches:
    // they must be recorded again
    bump_config_epoch();
  }
  return c.workspace[k];
}
// Backward-filter on a side stream.  Inside solver_step the weight gradient of a layer is a leaf of the backward pass:
// nothing reads it before the SGD update.  Its kernels (64 registers per thread, L2 / tensor bound) are queued on a second
// stream that forks after the layer's top gradient is packed and joins in front of the update, so the compute stream goes
// straight on to backward-data and to the pooling / LRN / pack kernels of the layers below (HBM bound, small CTAs), which
// share the SMs with the backward-filter CTAs.  Own scratch buffer (the two streams' split-reduction slices and bias
// partials are live at the same time).  SPARKNET_WGRAD_STREAM=0 keeps everything on the compute stream.
static bool g_side_allowed = false;   // set by Solver::step around its iterations
static cudaStream_t side_begin() {
  Context& c = ctx();
  static const bool enabled = !(getenv("SPARKNET_WGRAD_STREAM") && atoi(getenv("SPARKNET_WGRAD_STREAM")) == 0);
  if (!enabled || !g_side_allowed) return c.compute;
  SN_CUDA(cudaEventRecord(c.ev_fork, c.compute));
  SN_CUDA(cudaStreamWaitEvent(c.side, c.ev_fork, 0));
  c.side_used = true;
  g_ws_sel = 1;
  return c.side;
}
static void side_end() { g_ws_sel = 0; }
static void side_join() {
  Context& c = ctx();
  if (!c.side_used) return;
  SN_CUDA(cudaEventRecord(c.ev_join, c.side));
  SN_CUDA(cudaStreamWaitEvent(c.compute, c.ev_join, 0));
  c.side_used = false;
}
void count_launch(int n) { ctx().launches += n; }

// ------------------------------------------------------------------------------------------
// Storage / Blob
// ------------------------------------------------------------------------------------------
float* alloc_guarded(size_t count, float** raw) {
  SN_CUDA(cudaMalloc(raw, (std::max<size_t>(count, 1) + 2 * Storage::kGuardFloats) * sizeof(float)));
  SN_CUDA(cudaMemset(*raw, 0, Storage::kGuardFloats * sizeof(float)));
  SN_CUDA(cudaMemset(*raw + Storage::kGuardFloats + std::max<size_t>(count, 1), 0, Storage::kGuardFloats * sizeof(float)));
  return *raw + Storage::kGuardFloats;
}
Storage::~Storage() {
  if (dev_raw && owns_dev) cudaFree(dev_raw);
  if (host && owns_host) cudaFreeHost(host);
}
void Storage::ensure_dev() {
  if (!dev) {
    dev = alloc_guarded(count, &dev_raw);
    owns_dev = true;
  }
}
void Storage::ensure_host() {
  if (!host) SN_CUDA(cudaMallocHost(&host, std::max<size_t>(count, 1) * sizeof(float)));
}
// Anything that reads or writes a Storage OUTSIDE Solver::step (get_data, forward(), averaging, weight IO ...)
// advances this counter; a solver_step call that finds it unchanged since the previous call returned knows that
// host-side state and device contents are exactly as that call left them (Solver::step replays from iteration one).
static uint64_t g_outside_touch = 0;
static bool g_in_step = false;
static inline void note_touch() { if (!g_in_step) ++g_outside_touch; }
const float* Storage::gpu() {
  Context& c = ctx();
  if (head == UNINIT || head == AT_HOST) note_touch();
  switch (head) {
    case UNINIT:
      ensure_dev();
      SN_CUDA(cudaMemsetAsync(dev, 0, count * sizeof(float), c.compute));
      head = AT_DEVICE;
      break;
    case AT_HOST:
      ensure_dev();
      SN_CUDA(cudaMemcpyAsync(dev, host, count * sizeof(float), cudaMemcpyHostToDevice, c.compute));
      head = SYNCED;
      break;
    default: break;
  }
  return dev;
}
float* Storage::mutable_gpu() {
  note_touch();
  gpu();
  head = AT_DEVICE;
  ++gen;
  return dev;
}
float* Storage::mutable_cpu() {
  Context& c = ctx();
  note_touch();
  ensure_host();
  switch (head) {
    case UNINIT:
      memset(host, 0, count * sizeof(float));
      break;
    case AT_DEVICE:
      if (arena) { arena->download(); break; }   // the whole arena in one copy: the other weight blobs become SYNCED
      SN_CUDA(cudaMemcpyAsync(host, dev, count * sizeof(float), cudaMemcpyDeviceToHost, c.compute));
      SN_CUDA(cudaStreamSynchronize(c.compute));
      break;
    case SYNCED:
      // a previous upload from this host buffer may still be in flight; the caller may write next
      SN_CUDA(cudaStreamSynchronize(c.compute));
      break;
    default: break;
  }
  head = AT_HOST;
  ++gen;
  return host;
}
// Members whose `from` side is newest move; a run of them — with SYNCED members in between, whose bytes are the same on
// both sides — is one copy.  A member whose OTHER side is newer ends a run (copying over it would lose data).
static void arena_move(ParamArena& a, Storage::Head from, bool to_host) {
  Context& c = ctx();
  const size_t n = a.members.size();
  size_t i = 0;
  while (i < n) {
    if (a.members[i]->head != from && a.members[i]->head != Storage::SYNCED) { ++i; continue; }
    size_t j = i, first = n, last = n;
    while (j < n && (a.members[j]->head == from || a.members[j]->head == Storage::SYNCED)) {
      if (a.members[j]->head == from) { if (first == n) first = j; last = j; }
      ++j;
    }
    if (first != n) {                               // [first, last]: SYNCED members in between ride along (same bytes both sides)
      const size_t lo = a.offsets[first], hi = a.offsets[last] + a.members[last]->count;
      if (to_host) SN_CUDA(cudaMemcpyAsync(a.host + lo, a.dev + lo, (hi - lo) * sizeof(float), cudaMemcpyDeviceToHost, c.compute));
      else SN_CUDA(cudaMemcpyAsync(a.dev + lo, a.host + lo, (hi - lo) * sizeof(float), cudaMemcpyHostToDevice, c.compute));
      ++(to_host ? c.param_d2h_copies : c.param_h2d_copies);
      for (size_t k = first; k <= last; ++k) a.members[k]->head = Storage::SYNCED;
    }
    i = j;
  }
}
void ParamArena::download() {
  arena_move(*this, Storage::AT_DEVICE, true);
  SN_CUDA(cudaStreamSynchronize(ctx().compute));
}
void ParamArena::upload() { arena_move(*this, Storage::AT_HOST, false); }
void Blob::reshape(const std::vector<int>& s) {
  shape = s;
  const size_t c = count();
  if (!data || data->count != c) {
    data = std::make_shared<Storage>(c);
    diff = std::make_shared<Storage>(c);
  }
}

// ------------------------------------------------------------------------------------------
// Fillers (caffe/include/caffe/filler.hpp:31-46,48-63,65-95,144-166) — host side, one-time.
// The RNG stream differs from boost's (unpinned in the reference); tests copy weights instead.
// ------------------------------------------------------------------------------------------
static std::mt19937& host_rng() {
  static std::mt19937 g(std::random_device{}());
  return g;
}
static void fill_blob(const cp::FillerParameter& f, Blob* b) {
  const size_t n = b->count();
  SN_CHECK(n > 0, "filler: empty blob");
  float* d = b->data->mutable_cpu();
  if (f.type == "constant") {
    for (size_t i = 0; i < n; ++i) d[i] = f.value;
  } else if (f.type == "gaussian") {
    SN_CHECK(f.sparse == -1, "sparse gaussian filler is not on the supported path");
    std::normal_distribution<float> dist(f.mean, f.std);
    for (size_t i = 0; i < n; ++i) d[i] = dist(host_rng());
  } else if (f.type == "xavier") {
    const int fan_in = (int)(n / (size_t)b->num()), fan_out = (int)(n / (size_t)b->channels());
    float nn = (float)fan_in;
    if (f.variance_norm == 2) nn = (fan_in + fan_out) / 2.f;
    else if (f.variance_norm == 1) nn = (float)fan_out;
    const float scale = std::sqrt(3.f / nn);
    std::uniform_real_distribution<float> dist(-scale, scale);
    for (size_t i = 0; i < n; ++i) d[i] = dist(host_rng());
  } else if (f.type == "uniform") {
    std::uniform_real_distribution<float> dist(f.min, f.max);
    for (size_t i = 0; i < n; ++i) d[i] = dist(host_rng());
  } else {
    fatal("filler type '%s' is not on the supported path", f.type.c_str());
  }
}

Layer::~Layer() {
  if (w_shadow && !w_shadow_in_arena) cudaFree(w_shadow);
}
const float* Layer::weights_tf32() {
  Storage* st = blobs[0]->data.get();
  const float* w = st->gpu();
  if (!w_shadow) {
    const Net* owner = net && net->shares_with ? net->shares_with : net;   // adopted weights: their net's shadow arena
    if (owner && owner->arena_s && w >= owner->arena_w && w < owner->arena_w + owner->arena_floats) {
      w_shadow = owner->arena_s + (w - owner->arena_w);
      w_shadow_in_arena = true;
    } else {
      SN_CUDA(cudaMalloc(&w_shadow, st->count * sizeof(float)));
    }
  }
  if (w_shadow_gen != st->gen) {
    tc_round_to_tf32(ctx().compute, st->count, w, w_shadow);
    w_shadow_gen = st->gen;
  }
  return w_shadow;
}
void Layer::set_loss_weights(const std::vector<Blob*>& top) {   // layer.hpp:405-427
  loss.assign(top.size(), 0.f);
  for (size_t i = 0; i < lp.loss_weight.size() && i < top.size(); ++i) {
    const float w = lp.loss_weight[i];
    if (w == 0.f) continue;
    loss[i] = w;
    float* d = top[i]->diff->mutable_cpu();
    for (size_t k = 0; k < top[i]->count(); ++k) d[k] = w;
  }
}

using std::string;
using std::vector;
static cudaStream_t S() { return ctx().compute; }

// ------------------------------------------------------------------------------------------
// JavaDataLayer — reference caffe/src/caffe/layers/java_data_layer.cpp:10-44: callback into a
// malloc'ed buffer, then a blocking pageable H2D on the next gpu_data() (syncedmem.cpp:61-69).
// Here: the callback fills PINNED memory, the upload runs on a side stream into one of two
// device buffers, and the compute stream only waits on an event; the upload of iteration i+1
// overlaps the compute of iteration i.  Callback order and count are exactly the reference's
// (one call per JavaData layer per forward pass, data before label).
// ------------------------------------------------------------------------------------------
// What a byte-input JavaData layer hands to the one convolution that reads its blob, so that the convolution's pack kernel
// widens the pixels itself (conv_flat.cu s2d_pack_kernel<true>) and the fp32 minibatch is never materialised inside solver_step.
struct ByteFeed { const uint8_t* bytes; bool is_signed; const float* mean; int mean_div, mean_period; float scale; const Storage* src; uint64_t gen; };
struct JavaDataLayer : Layer {
  java_callback_t cb = nullptr;
  std::function<bool()> consumer_takes_bytes;                  // set by Net::Net when a single convolution reads the top blob
  std::function<void(const ByteFeed&)> consumer_feed;
  bool fp32_stale = false;                                     // the fp32 blob does not hold the current minibatch yet (see materialize)
  int stale_slot = 0;
  // after solver_step: make the fp32 blob current again for whoever reads it next (get_data, forward(), tests)
  void materialize() {
    if (!fp32_stale) return;
    widen_bytes(ctx().compute, count, dev_b[stale_slot], byte_mode == 1, mean_dev, mean_div, mean_period, byte_scale, devbuf[stale_slot]);
    fp32_stale = false;
  }
  vector<int> shape, java_shape;
  size_t count = 0;
  float* pinned[2] = {nullptr, nullptr};
  float* devbuf[2] = {nullptr, nullptr};
  float* devraw[2] = {nullptr, nullptr};
  float* pool_raw = nullptr;
  cudaEvent_t h2d_done[2], consumed[2];
  bool consumed_valid[2] = {false, false};
  int parity = 0;
  // resident mode (sparknet_set_resident_inputs): a pool of minibatches already in HBM
  float* pool = nullptr;
  int pool_n = 0, pool_i = 0;
  // In-call prefetch (Solver::step): a helper thread runs the callbacks of the NEXT iterations while this thread enqueues
  // the current one.  Order and count are the reference's: data then label, exactly one pair per iteration, all of
  // them inside the solver_step call (SURVEY 8b "Prefetching is allowed only inside those bounds").
  //   filled   batches whose callback has returned (helper -> caller)
  //   enqueued batches whose H2D copy has been enqueued and whose h2d_done event is recorded (caller -> helper)
  // Byte-pixel input (sparknet_set_train_data_callback_bytes): the callback fills batch*C*H*W BYTES — what the reference
  // holds before its callback widens them on the host (ByteImage.java:86 cropInto: `buffer[i] = red[...]`, a signed Java
  // byte to float; Net.scala:194-221) — a quarter of the host traffic and of the H2D copy; the device widens them into
  // the same fp32 NCHW blob: x = ((float)b - mean[(i / mean_div) % mean_period]) * scale.
  int byte_mode = 0;                            // 0 fp32 callback, 1 signed bytes (Java), 2 unsigned bytes
  uint8_t* pinned_b[2] = {nullptr, nullptr};
  uint8_t* dev_b[2] = {nullptr, nullptr};
  float* mean_dev = nullptr;
  int mean_div = 1, mean_period = 0;
  float byte_scale = 1.f;
  void set_byte_mode(int mode, const float* mean, int mean_count, float scale) {
    if (mode == 0) { byte_mode = 0; return; }
    for (int i = 0; i < 2; ++i) {
      if (!pinned_b[i]) SN_CUDA(cudaMallocHost(&pinned_b[i], count));
      if (!dev_b[i]) SN_CUDA(cudaMalloc(&dev_b[i], (count + 15) / 16 * 16));
    }
    if (mean_dev) { SN_CUDA(cudaStreamSynchronize(ctx().compute)); cudaFree(mean_dev); mean_dev = nullptr; }
    mean_period = 0; mean_div = 1;
    if (mean && mean_count > 0) {
      const size_t per_image = count / (size_t)shape[0];
      const size_t hw = shape.size() >= 3 ? per_image / (size_t)shape[1] : 1;
      if ((size_t)mean_count == per_image) { mean_div = 1; mean_period = mean_count; }            // mean image (C*H*W)
      else if (shape.size() >= 3 && mean_count == shape[1]) { mean_div = (int)hw; mean_period = mean_count; }   // per channel
      else fatal("byte input: the mean must have C*H*W = %zu or C = %d values, got %d", per_image, shape.size() >= 2 ? shape[1] : 1, mean_count);
      SN_CUDA(cudaMalloc(&mean_dev, (size_t)mean_count * sizeof(float)));
      SN_CUDA(cudaMemcpy(mean_dev, mean, (size_t)mean_count * sizeof(float), cudaMemcpyHostToDevice));
    }
    byte_scale = scale;
    byte_mode = mode;
  }
  // host-side cost of the input path since the last sparknet_input_stats call
  double cb_ns = 0, wait_ns = 0;
  long cb_calls = 0;
  struct PrefetchSync { std::mutex m; std::condition_variable cv; };
  PrefetchSync* pf = nullptr;                   // non-null while a prefetch thread serves this layer
  long pf_filled = 0, pf_enqueued = 0;
  int pf_base_parity = 0;
  // helper thread: produce batch k of this call into pinned[(base parity + k) % 2]
  void prefetch_fill(long k) {
    const int slot = (pf_base_parity + (int)k) & 1;
    {
      std::unique_lock<std::mutex> lk(pf->m);
      pf->cv.wait(lk, [&] { return pf_enqueued + 2 > k; });       // batch k-2 (same slot) has been enqueued
    }
    SN_CUDA(cudaEventSynchronize(h2d_done[slot]));                // ... and its copy has left the pinned buffer
    cb(byte_mode ? (void*)pinned_b[slot] : (void*)pinned[slot], shape[0], (int)shape.size() - 1, java_shape.data());
    {
      std::lock_guard<std::mutex> lk(pf->m);
      pf_filled = k + 1;
    }
    pf->cv.notify_all();
  }
  using Layer::Layer;
  ~JavaDataLayer() override {
    for (int i = 0; i < 2; ++i) {
      if (pinned[i]) cudaFreeHost(pinned[i]);
      if (pinned_b[i]) cudaFreeHost(pinned_b[i]);
      if (dev_b[i]) cudaFree(dev_b[i]);
      if (devraw[i]) cudaFree(devraw[i]);
      if (count) { cudaEventDestroy(h2d_done[i]); cudaEventDestroy(consumed[i]); }
    }
    if (pool_raw) cudaFree(pool_raw);
    if (mean_dev) cudaFree(mean_dev);
  }
  const char* type() const override { return "JavaData"; }
  void setup(const vector<Blob*>&, const vector<Blob*>& top) override {
    SN_CHECK(lp.has_java_data_param, "JavaData layer %s needs java_data_param", lp.name.c_str());
    for (int64_t d : lp.java_data_shape.dim) shape.push_back((int)d);
    SN_CHECK(!shape.empty(), "JavaData shape needs a batch dimension");
    java_shape.assign(shape.begin() + 1, shape.end());
    top[0]->reshape(shape);
    count = top[0]->count();
    SN_CHECK(count > 0, "batch_size, channels, height, and width must be specified and positive in java_data_param");
    for (int i = 0; i < 2; ++i) {
      SN_CUDA(cudaMallocHost(&pinned[i], count * sizeof(float)));
      memset(pinned[i], 0, count * sizeof(float));
      devbuf[i] = alloc_guarded(count, &devraw[i]);
      SN_CUDA(cudaEventCreateWithFlags(&h2d_done[i], cudaEventDisableTiming));
      SN_CUDA(cudaEventCreateWithFlags(&consumed[i], cudaEventDisableTiming));
    }
    // the top blob's device pointer is swapped between devbuf[0/1] (or pool slots)
    top[0]->data = std::make_shared<Storage>(count, devbuf[0]);
    SN_CUDA(cudaMemset(devbuf[0], 0, count * sizeof(float)));
    top[0]->data->head = Storage::AT_DEVICE;
  }
  void forward(const vector<Blob*>&, const vector<Blob*>& top) override {
    Context& c = ctx();
    Storage* st = top[0]->data.get();
    if (pool) {                                   // resident inputs: no callback, no copy
      st->dev = pool + (size_t)(pool_i % pool_n) * count;
      st->head = Storage::AT_DEVICE;
      ++st->gen;                                  // new contents: consumers that cache a repacked copy must redo it
      ++pool_i;
      return;
    }
    SN_CHECK(cb, "JavaDataLayer needs to be initalized by calling SetCallback");
    const int p = parity, q = parity ^ 1;
    // everything enqueued so far on the compute stream is the last reader of buffer q
    SN_CUDA(cudaEventRecord(consumed[q], c.compute));
    consumed_valid[q] = true;
    if (pf) {                                     // the helper thread runs the callback: wait for this batch
      std::unique_lock<std::mutex> lk(pf->m);
      pf->cv.wait(lk, [&] { return pf_filled > pf_enqueued; });
    } else {
      const auto t0 = std::chrono::steady_clock::now();
      SN_CUDA(cudaEventSynchronize(h2d_done[p]));   // pinned[p] free again (no-op the first time)
      const auto t1 = std::chrono::steady_clock::now();
      const bool in_step = g_in_step;
      g_in_step = false;                            // a callback that reaches back into the ABI (get_data ...) counts as outside
      cb(byte_mode ? (void*)pinned_b[p] : (void*)pinned[p], shape[0], (int)shape.size() - 1, java_shape.data());
      g_in_step = in_step;
      const auto t2 = std::chrono::steady_clock::now();
      wait_ns += std::chrono::duration<double, std::nano>(t1 - t0).count();
      cb_ns += std::chrono::duration<double, std::nano>(t2 - t1).count();
      ++cb_calls;
    }
    if (consumed_valid[p]) SN_CUDA(cudaStreamWaitEvent(c.copy, consumed[p], 0));
    if (byte_mode) SN_CUDA(cudaMemcpyAsync(dev_b[p], pinned_b[p], count, cudaMemcpyHostToDevice, c.copy));
    else SN_CUDA(cudaMemcpyAsync(devbuf[p], pinned[p], count * sizeof(float), cudaMemcpyHostToDevice, c.copy));
    SN_CUDA(cudaEventRecord(h2d_done[p], c.copy));
    SN_CUDA(cudaStreamWaitEvent(c.compute, h2d_done[p], 0));
    const bool lazy = byte_mode && g_in_step && consumer_takes_bytes && consumer_takes_bytes();
    if (byte_mode && !lazy) widen_bytes(c.compute, count, dev_b[p], byte_mode == 1, mean_dev, mean_div, mean_period, byte_scale, devbuf[p]);
    fp32_stale = lazy;
    stale_slot = p;
    st->dev = devbuf[p];
    st->head = Storage::AT_DEVICE;
    ++st->gen;
    if (lazy) consumer_feed(ByteFeed{dev_b[p], byte_mode == 1, mean_dev, mean_div, mean_period, byte_scale, st, st->gen});
    parity = q;
    if (pf) {
      { std::lock_guard<std::mutex> lk(pf->m); ++pf_enqueued; }
      pf->cv.notify_all();
    }
  }
  void backward(const vector<Blob*>&, const vector<bool>&, const vector<Blob*>&) override {}
  // Pull `n` minibatches through the callback now and keep them in HBM.
  void make_resident(int n) {
    if (pool_raw) { cudaFree(pool_raw); pool_raw = nullptr; pool = nullptr; }
    pool_n = n; pool_i = 0;
    if (n <= 0) return;
    SN_CHECK(cb, "set the data callback before sparknet_set_resident_inputs");
    pool = alloc_guarded((size_t)n * count, &pool_raw);
  }
  void fill_resident_slot(int i) {
    cb(byte_mode ? (void*)pinned_b[0] : (void*)pinned[0], shape[0], (int)shape.size() - 1, java_shape.data());
    if (byte_mode) {
      SN_CUDA(cudaMemcpy(dev_b[0], pinned_b[0], count, cudaMemcpyHostToDevice));
      widen_bytes(ctx().compute, count, dev_b[0], byte_mode == 1, mean_dev, mean_div, mean_period, byte_scale, pool + (size_t)i * count);
      SN_CUDA(cudaStreamSynchronize(ctx().compute));
    } else {
      SN_CUDA(cudaMemcpy(pool + (size_t)i * count, pinned[0], count * sizeof(float), cudaMemcpyHostToDevice));
    }
  }
};

// ------------------------------------------------------------------------------------------
// Convolution — reference conv_layer.{cpp:8-71,cu:7-56} + base_conv_layer.cpp:12-376 (N x im2col +
// group x sgemm + bias gemm forward; gemv + im2col + 2 sgemm + col2im backward).  Here: one
// implicit GEMM per group and direction, bias (+ in-place ReLU) in the epilogue.
// ------------------------------------------------------------------------------------------
struct ConvolutionLayer : Layer {
  ConvGeom g;
  bool bias_term = true, fuse_relu = false;
  bool gate_relu = false;     // the bottom is the output of an in-place ReLU whose backward is applied while this layer writes dx
  using Layer::Layer;
  const char* type() const override { return "Convolution"; }
  void setup(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    const cp::ConvolutionParameter& c = lp.convolution_param;
    SN_CHECK(bottom[0]->shape.size() == 4, "Convolution %s: input must be 4-D (N,C,H,W)", lp.name.c_str());
    SN_CHECK(c.axis == 1, "Convolution %s: only channel axis 1 is supported", lp.name.c_str());
    if (c.has_kernel_h || c.has_kernel_w) {
      SN_CHECK(c.kernel_size.empty(), "Either kernel_size or kernel_h/w should be specified; not both.");
      g.kh = (int)c.kernel_h; g.kw = (int)c.kernel_w;
    } else {
      SN_CHECK(!c.kernel_size.empty(), "Convolution %s: kernel_size must be specified", lp.name.c_str());
      g.kh = (int)c.kernel_size[0];
      g.kw = (int)(c.kernel_size.size() > 1 ? c.kernel_size[1] : c.kernel_size[0]);
    }
    SN_CHECK(g.kh > 0 && g.kw > 0, "Filter dimensions must be nonzero.");
    if (c.has_stride_h || c.has_stride_w) { g.sh = (int)c.stride_h; g.sw = (int)c.stride_w; }
    else if (c.stride.empty()) { g.sh = g.sw = 1; }
    else { g.sh = (int)c.stride[0]; g.sw = (int)(c.stride.size() > 1 ? c.stride[1] : c.stride[0]); }
    if (c.has_pad_h || c.has_pad_w) { g.ph = (int)c.pad_h; g.pw = (int)c.pad_w; }
    else if (c.pad.empty()) { g.ph = g.pw = 0; }
    else { g.ph = (int)c.pad[0]; g.pw = (int)(c.pad.size() > 1 ? c.pad[1] : c.pad[0]); }
    SN_CHECK(g.sh > 0 && g.sw > 0, "stride must be positive");
    g.N = bottom[0]->num(); g.C = bottom[0]->channels(); g.H = bottom[0]->height(); g.W = bottom[0]->width();
    g.Cout = (int)c.num_output;
    g.group = (int)c.group;
    SN_CHECK(g.Cout > 0, "num_output must be positive");
    SN_CHECK(g.C % g.group == 0 && g.Cout % g.group == 0, "channels and num_output must be multiples of group");
    bias_term = c.bias_term;
    g.Ho = (g.H + 2 * g.ph - g.kh) / g.sh + 1;
    g.Wo = (g.W + 2 * g.pw - g.kw) / g.sw + 1;
    if (blobs.empty()) {
      blobs.push_back(std::make_shared<Blob>());
      blobs[0]->reshape({g.Cout, g.C / g.group, g.kh, g.kw});
      fill_blob(c.weight_filler, blobs[0].get());
      if (bias_term) {
        blobs.push_back(std::make_shared<Blob>());
        blobs[1]->reshape({g.Cout});
        fill_blob(c.bias_filler, blobs[1].get());
      }
    }
    param_propagate_down.assign(blobs.size(), true);
    top[0]->reshape({g.N, g.Cout, g.Ho, g.Wo});
  }
  // The window convolution that reads this layer's output (through nothing but an in-place ReLU this layer applies
  // itself) gets its packed, tf32-rounded, channels-last input straight from this layer's epilogue: one pack launch and
  // one pass over the tensor less per iteration.  Set by Net::Net; SPARKNET_FUSE_PACK=0 turns it off.
  ConvolutionLayer* flat_consumer = nullptr;
  bool takes_flat_input() { return net->precision == PREC_TF32 && !use_s2d() && use_flat(1); }
  FlatOut flat_target() {
    if (!xflat) xflat = flat_alloc(g.C);
    const FlatGeom f = tc_flat_geom(g);
    return FlatOut{xflat, f.lead, f.Gp, f.Wp, g.C};
  }
  void flat_written(Blob* src) { xflat_src = src->data.get(); xflat_gen = src->data->gen; }
  void forward(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    const float* x = bottom[0]->data->gpu();
    const float* b = bias_term ? blobs[1]->data->gpu() : nullptr;
    float* y = top[0]->data->mutable_gpu();
    FlatOut fo{nullptr, 0, 0, 0, 0};
    const bool feed_next = flat_consumer && flat_consumer->takes_flat_input();
    if (feed_next) fo = flat_consumer->flat_target();
    bool fed = false;
    if (net->precision == PREC_TF32 && use_s2d()) {
      tcw_conv_fwd(S(), gs, packed_input(bottom[0]), packed_filter(false), b, y, fuse_relu, feed_next ? &fo : nullptr);
      fed = feed_next;
    } else if (net->precision == PREC_TF32 && use_flat(1)) {
      if (use_window()) { tcw_conv_fwd(S(), g, packed_input(bottom[0]), packed_filter(false), b, y, fuse_relu, feed_next ? &fo : nullptr); fed = feed_next; }
      else tc_conv_fwd_flat(S(), g, packed_input(bottom[0]), packed_filter(false), b, y, fuse_relu);
    } else if (net->precision == PREC_TF32) tc_conv_fwd(S(), g, x, weights_tf32(), b, y, fuse_relu);
    else ffma_conv_fwd(S(), g, x, blobs[0]->data->gpu(), b, y, fuse_relu);
    if (fed) flat_consumer->flat_written(top[0]);
  }
  // ---- flat-grid operands (gemm_tc.cu "flat-grid convolution"): packed copies owned by the layer ----
  float *xflat = nullptr, *dyflat = nullptr, *wflat[2] = {nullptr, nullptr};
  uint64_t xflat_gen = ~0ull, wflat_gen[2] = {~0ull, ~0ull};
  const Storage* xflat_src = nullptr;
  int flat_ok = -1;
  // SPARKNET_CONV_FLAT: bit 0 forward, bit 1 backward-filter, bit 2 backward-data (default 7; 0 = gather kernels only)
  bool use_flat(int bit) {
    if (flat_ok < 0) {
      const char* e = getenv("SPARKNET_CONV_FLAT");
      flat_ok = tc_flat_supported(g) ? (e ? atoi(e) : 7) : 0;
    }
    return (flat_ok & bit) != 0;
  }
  // SPARKNET_FLAT_WINDOW=0 falls back from the window kernels (conv_flat.cu) to the chunk-per-tap flat kernels
  int window_ok = -1;
  bool use_window() {
    if (window_ok < 0) {
      const char* e = getenv("SPARKNET_FLAT_WINDOW");
      window_ok = (tcw_supported(g) && !(e && atoi(e) == 0)) ? 1 : 0;
    }
    return window_ok != 0;
  }
  // strided layers folded to stride 1 (space to depth): gs is the equivalent stride-1 layer the window kernels run
  ConvGeom gs;
  int s2d_ok = -1;
  float* dws = nullptr;                               // filter gradient in the folded layout
  // s2d_ok: 1 = space to depth (stride folded away), 2 = channel padding (opt-in, SPARKNET_CONV_CPAD=1): a stride-1 first
  // layer with fewer than 16 input channels (cifar conv1: C = 3) runs on the window kernels over a grid whose rows carry
  // 16 channels, 13 of them zero.  Correct (tests/test_gpu_tc.py) but not faster on B200 at cifar sizes: forward 44 us
  // either way, backward 105 us against 73 us on the gather kernels (pack + window + slice reduction + unpack are four
  // latency-bound launches), so the gather kernels stay the default.
  bool use_s2d() {
    if (s2d_ok < 0) {
      const char* e = getenv("SPARKNET_CONV_S2D");
      s2d_ok = (tcw_s2d_supported(g) && !(e && atoi(e) == 0)) ? 1 : 0;
      if (s2d_ok) gs = tcw_s2d_geom(g);
      const char* ep = getenv("SPARKNET_CONV_CPAD");
      if (!s2d_ok && ep && atoi(ep) == 1 && g.sh == 1 && g.sw == 1 && g.group == 1 && g.C < 16 && g.Cout % 16 == 0) {
        ConvGeom gp = g;
        gp.C = 16;
        if (tcw_supported(gp) && tcw_wgrad_supported(gp)) { gs = gp; s2d_ok = 2; }
      }
    }
    return s2d_ok != 0;
  }
  bool cpad() { return use_s2d() && s2d_ok == 2; }
  const ConvGeom& fg() { return use_s2d() ? gs : g; } // geometry of the flat grid in use
  float* flat_alloc(int channels) {
    float* p = nullptr;
    const size_t bytes = (size_t)channels * tc_flat_rows(fg()) * sizeof(float);
    SN_CUDA(cudaMalloc(&p, bytes));
    SN_CUDA(cudaMemsetAsync(p, 0, bytes, S()));      // the bands between images stay zero for good
    return p;
  }
  // byte-pixel input handed over by the JavaData layer below (ByteFeed): valid for exactly that generation of the blob
  ByteFeed byte_feed{};
  bool takes_bytes() { return net->precision == PREC_TF32 && use_s2d() && !cpad(); }
  const float* packed_input(Blob* bottom) {
    const float* x = bottom->data->gpu();
    if (!xflat) xflat = flat_alloc(fg().C);
    if (xflat_src != bottom->data.get() || xflat_gen != bottom->data->gen) {
      if (takes_bytes() && byte_feed.bytes && byte_feed.src == bottom->data.get() && byte_feed.gen == bottom->data->gen)
        tcw_s2d_pack_input_bytes(S(), g, byte_feed.bytes, byte_feed.is_signed, byte_feed.mean, byte_feed.mean_div, byte_feed.mean_period,
                                 byte_feed.scale, xflat);
      else if (cpad()) tc_flat_pack(S(), gs, false, x, xflat, nullptr, false, g.C);
      else if (use_s2d()) tcw_s2d_pack_input(S(), g, x, xflat);
      else tc_flat_pack(S(), g, false, x, xflat);
      xflat_src = bottom->data.get();
      xflat_gen = bottom->data->gen;
    }
    return xflat;
  }
  const float* packed_filter(bool dgrad) {
    Storage* st = blobs[0]->data.get();
    const float* w = st->gpu();
    const int i = dgrad ? 1 : 0;
    if (!wflat[i]) SN_CUDA(cudaMalloc(&wflat[i], tc_flat_filter_floats(fg(), dgrad) * sizeof(float)));
    if (wflat_gen[i] != st->gen) {
      if (cpad()) tc_flat_pack_filter(S(), gs, false, w, wflat[i], g.C);
      else if (use_s2d()) tcw_s2d_pack_filter(S(), g, w, wflat[i]);
      else tc_flat_pack_filter(S(), g, dgrad, w, wflat[i]);
      wflat_gen[i] = st->gen;
    }
    return wflat[i];
  }
  // Solver::step, start of an iteration: the stale filter copies this layer is about to need go into the net-wide batch
  // (one launch for all layers); packed_filter() then finds them current.  Anything missed here is packed lazily as before.
  void queue_filter_packs(FilterPackBatch& b, bool propagate_down) {
    if (net->precision != PREC_TF32) return;
    Storage* st = blobs[0]->data.get();
    auto add = [&](bool dgrad, int mode) {
      const int i = dgrad ? 1 : 0;
      if (b.n >= FilterPackBatch::kMax || (wflat[i] && wflat_gen[i] == st->gen)) return;
      if (!wflat[i]) SN_CUDA(cudaMalloc(&wflat[i], tc_flat_filter_floats(fg(), dgrad) * sizeof(float)));
      b.job[b.n++] = tcw_filter_pack_job(g, mode, st->gpu(), wflat[i]);
      wflat_gen[i] = st->gen;
    };
    if (cpad()) {
      if (b.n < FilterPackBatch::kMax && !(wflat[0] && wflat_gen[0] == st->gen)) {
        if (!wflat[0]) SN_CUDA(cudaMalloc(&wflat[0], tc_flat_filter_floats(gs, false) * sizeof(float)));
        b.job[b.n++] = tcw_filter_pack_job(gs, 0, st->gpu(), wflat[0], g.C);
        wflat_gen[0] = st->gen;
      }
      return;
    }
    if (use_s2d()) { add(false, 2); return; }
    if (use_flat(1)) add(false, 0);
    if (use_flat(4) && propagate_down) add(true, 1);
  }
  ~ConvolutionLayer() override {
    cudaFree(xflat); cudaFree(dyflat); cudaFree(wflat[0]); cudaFree(wflat[1]); cudaFree(dws);
  }
  void backward(const vector<Blob*>& top, const vector<bool>& pd, const vector<Blob*>& bottom) override {
    const float* dy = top[0]->diff->gpu();
    const bool acc = !net->overwrite_param_diffs;
    const bool tf32 = net->precision == PREC_TF32;
    const bool flat_w = tf32 && use_flat(2) && param_propagate_down[0], flat_d = tf32 && use_flat(4) && pd[0];
    const bool want_db = bias_term && param_propagate_down[1];
    if (tf32 && use_s2d() && param_propagate_down[0]) {
      // folded layer: pack the top gradient on the folded grid (+ bias gradient), window backward-filter, unfold dW
      if (!dyflat) dyflat = flat_alloc(gs.Cout);
      tc_flat_pack(S(), gs, true, dy, dyflat, want_db ? blobs[1]->diff->mutable_gpu() : nullptr, acc);
      if (!dws) SN_CUDA(cudaMalloc(&dws, (size_t)gs.Cout * gs.C * gs.kh * gs.kw * sizeof(float)));
      const float* xin = packed_input(bottom[0]);                 // (current since the forward pass; before the fork if not)
      float* dwp = blobs[0]->diff->mutable_gpu();
      cudaStream_t ws = side_begin();
      tcw_conv_wgrad(ws, gs, xin, dyflat, dws, false);
      if (cpad()) tcw_cpad_unpack_wgrad(ws, g, gs.C, dws, dwp, acc);
      else tcw_s2d_unpack_wgrad(ws, g, dws, dwp, acc);
      if (ws != S() && !pd[0] && net->on_param_grads_ready) net->on_param_grads_ready(this);   // (this path runs backward-data, if any, afterwards)
      side_end();
      if (pd[0]) {
        float* dx = bottom[0]->diff->mutable_gpu();
        tc_conv_dgrad(S(), g, dy, blobs[0]->data->gpu(), dx);
        if (gate_relu) relu_bwd(S(), bottom[0]->count(), dx, bottom[0]->data->gpu(), dx, 0.f);
      }
      return;
    }
    if (flat_w || flat_d) {
      if (!dyflat) dyflat = flat_alloc(g.Cout);
      // one pass over the top gradient packs it and sums the bias gradient
      tc_flat_pack(S(), g, true, dy, dyflat, want_db ? blobs[1]->diff->mutable_gpu() : nullptr, acc);
    } else if (want_db) {
      conv_bias_grad(S(), g.N, g.Cout, g.Ho * g.Wo, dy, blobs[1]->diff->mutable_gpu(), acc);
    }
    // Backward-filter is a leaf (the side stream runs it, side_begin); backward-data is on the critical path of the layers
    // below.  Large layers, whose kernels each fill the GPU: backward-data first, so that backward-filter then shares the SMs
    // with the pooling / LRN / pack kernels that follow.  Small layers (LeNet: neither kernel fills the GPU): backward-filter
    // is released first and runs beside backward-data.  Measured both ways on the four nets.
    const bool wgrad_first = 2.0 * g.N * g.Ho * g.Wo * (double)g.Cout * g.Cg() * g.kh * g.kw < 4e8;
    auto run_dgrad = [&] {
      if (pd[0]) {
        const float* w = blobs[0]->data->gpu();
        float* dx = bottom[0]->diff->mutable_gpu();
        const float* gate = gate_relu ? bottom[0]->data->gpu() : nullptr;       // y of the in-place ReLU below (y > 0 <=> x > 0)
        if (flat_d && use_window()) { tcw_conv_dgrad(S(), g, dyflat, packed_filter(true), dx, gate); gate = nullptr; }
        else if (flat_d) tc_conv_dgrad_flat(S(), g, dyflat, packed_filter(true), dx);
        else if (net->precision == PREC_TF32) tc_conv_dgrad(S(), g, dy, w, dx);
        else ffma_conv_dgrad(S(), g, dy, w, dx);
        if (gate) relu_bwd(S(), bottom[0]->count(), dx, gate, dx, 0.f);          // paths without a gating epilogue
      }
    };
    auto run_wgrad = [&] {
      if (param_propagate_down[0]) {
        const float* x = bottom[0]->data->gpu();
        float* dw = blobs[0]->diff->mutable_gpu();
        const float* xin = flat_w ? packed_input(bottom[0]) : nullptr;
        cudaStream_t ws = side_begin();
        if (flat_w && use_window() && tcw_wgrad_supported(g)) tcw_conv_wgrad(ws, g, xin, dyflat, dw, acc);
        else if (flat_w) tc_conv_wgrad_flat(ws, g, xin, dyflat, dw, acc);
        else if (net->precision == PREC_TF32) tc_conv_wgrad(ws, g, x, dy, dw, acc);
        else ffma_conv_wgrad(ws, g, x, dy, dw, acc);
        if (ws != S() && !wgrad_first && net->on_param_grads_ready) net->on_param_grads_ready(this);
        side_end();
      }
    };
    if (wgrad_first) { run_wgrad(); run_dgrad(); } else { run_dgrad(); run_wgrad(); }
  }
};

// InnerProduct — reference inner_product_layer.{cpp:10-117,cu:10-56}
struct InnerProductLayer : Layer {
  int M = 0, N = 0, K = 0;
  bool bias_term = true, fuse_relu = false;
  using Layer::Layer;
  const char* type() const override { return "InnerProduct"; }
  void setup(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    const cp::InnerProductParameter& p = lp.inner_product_param;
    N = (int)p.num_output;
    SN_CHECK(N > 0, "InnerProduct %s: num_output must be positive", lp.name.c_str());
    bias_term = p.bias_term;
    const int axis = p.axis < 0 ? p.axis + (int)bottom[0]->shape.size() : p.axis;
    K = (int)bottom[0]->count(axis, (int)bottom[0]->shape.size());
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
  }
  // ---- all-TMA path (conv_flat.cu tcw_matmul*): operands that are not weights get a tf32-rounded copy first ----
  float *x_r = nullptr, *dy_r = nullptr;
  uint64_t x_r_gen = ~0ull;
  const Storage* x_r_src = nullptr;
  int tma_ok = -1;
  bool use_tma() {                                     // SPARKNET_IP_TMA=0 keeps the gather kernels of gemm_tc.cu
    if (tma_ok < 0) {
      const char* e = getenv("SPARKNET_IP_TMA");
      tma_ok = (e && atoi(e) == 0) ? 0 : 1;
    }
    return tma_ok != 0;
  }
  const float* rounded_input(Blob* bottom) {
    const float* x = bottom->data->gpu();
    if (!x_r) SN_CUDA(cudaMalloc(&x_r, (size_t)M * K * sizeof(float)));
    if (x_r_src != bottom->data.get() || x_r_gen != bottom->data->gen) {
      tc_round_to_tf32(S(), (size_t)M * K, x, x_r);
      x_r_src = bottom->data.get();
      x_r_gen = bottom->data->gen;
    }
    return x_r;
  }
  ~InnerProductLayer() override { cudaFree(x_r); cudaFree(dy_r); }
  void forward(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    const float* x = bottom[0]->data->gpu();
    const float* b = bias_term ? blobs[1]->data->gpu() : nullptr;
    float* y = top[0]->data->mutable_gpu();
    if (net->precision == PREC_TF32 && use_tma() && tcw_matmul_supported(M, N, K, x, blobs[0]->data->gpu(), false, y))
      tcw_matmul(S(), M, N, K, rounded_input(bottom[0]), weights_tf32(), false, b, fuse_relu, y);
    else if (net->precision == PREC_TF32) tc_ip_fwd(S(), M, N, K, x, weights_tf32(), b, y, fuse_relu);
    else ffma_ip_fwd(S(), M, N, K, x, blobs[0]->data->gpu(), b, y, fuse_relu);
  }
  void backward(const vector<Blob*>& top, const vector<bool>& pd, const vector<Blob*>& bottom) override {
    const float* dy = top[0]->diff->gpu();
    const bool acc = !net->overwrite_param_diffs;
    const bool tf32 = net->precision == PREC_TF32;
    const float* x = bottom[0]->data->gpu();
    const bool tma_w = tf32 && use_tma() && param_propagate_down[0] && tcw_matmul_wgrad_supported(M, N, K, dy, x, blobs[0]->diff->gpu());
    const bool tma_d = tf32 && use_tma() && pd[0] && tcw_matmul_supported(M, K, N, dy, blobs[0]->data->gpu(), true, bottom[0]->diff->gpu());
    bool bias_done = false;
    if (tma_w || tma_d) {
      if (!dy_r) SN_CUDA(cudaMalloc(&dy_r, (size_t)M * N * sizeof(float)));
      const bool want_db = bias_term && param_propagate_down[1];
      float* db = want_db ? blobs[1]->diff->mutable_gpu() : nullptr;
      if (N % 4 == 0 && ((uintptr_t)dy & 15) == 0 && (!db || ((uintptr_t)db & 15) == 0)) {
        ip_round_bias(S(), M, N, dy, dy_r, db, acc);      // rounding pass and bias gradient share one read of dy
        bias_done = want_db;
      } else {
        tc_round_to_tf32(S(), (size_t)M * N, dy, dy_r);
      }
    }
    if (bias_term && param_propagate_down[1] && !bias_done) ip_bias_grad(S(), M, N, dy, blobs[1]->diff->mutable_gpu(), acc);
    const bool wgrad_first = 2.0 * M * (double)N * K < 4e8;          // see ConvolutionLayer::backward
    auto run_dgrad = [&] {
      if (pd[0]) {
        float* dx = bottom[0]->diff->mutable_gpu();
        if (tma_d) tcw_matmul(S(), M, K, N, dy_r, weights_tf32(), true, nullptr, false, dx);
        else if (tf32) tc_ip_dgrad(S(), M, N, K, dy, weights_tf32(), dx);
        else ffma_ip_dgrad(S(), M, N, K, dy, blobs[0]->data->gpu(), dx);
      }
    };
    auto run_wgrad = [&] {
      if (param_propagate_down[0]) {
        float* dw = blobs[0]->diff->mutable_gpu();
        const float* xr = tma_w ? rounded_input(bottom[0]) : nullptr;
        cudaStream_t ws = side_begin();
        if (tma_w) tcw_matmul_wgrad(ws, M, N, K, dy_r, xr, dw, acc);
        else if (tf32) tc_ip_wgrad(ws, M, N, K, x, dy, dw, acc);
        else ffma_ip_wgrad(ws, M, N, K, x, dy, dw, acc);
        if (ws != S() && !wgrad_first && net->on_param_grads_ready) net->on_param_grads_ready(this);
        side_end();
      }
    };
    if (wgrad_first) { run_wgrad(); run_dgrad(); } else { run_dgrad(); run_wgrad(); }
  }
};

// Pooling — reference pooling_layer.{cpp:17-307,cu:11-380}
struct PoolingLayer : Layer {
  PoolGeom g;
  int method = 0;
  int* mask = nullptr;
  bool gate_relu = false;     // the bottom is the output of an in-place ReLU whose backward is folded into this layer's
  using Layer::Layer;
  ~PoolingLayer() override { if (mask) cudaFree(mask); }
  const char* type() const override { return "Pooling"; }
  void setup(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    const cp::PoolingParameter& p = lp.pooling_param;
    method = p.pool;
    SN_CHECK(method == 0 || method == 1, "Pooling %s: only MAX and AVE are on the supported path", lp.name.c_str());
    SN_CHECK(bottom[0]->shape.size() == 4, "Pooling %s: input must have 4 axes", lp.name.c_str());
    g.N = bottom[0]->num(); g.C = bottom[0]->channels(); g.H = bottom[0]->height(); g.W = bottom[0]->width();
    if (p.global_pooling) { g.kh = g.H; g.kw = g.W; }
    else if (p.has_kernel_size) { g.kh = g.kw = (int)p.kernel_size; }
    else {
      SN_CHECK(p.has_kernel_h && p.has_kernel_w, "For non-square filters both kernel_h and kernel_w are required.");
      g.kh = (int)p.kernel_h; g.kw = (int)p.kernel_w;
    }
    SN_CHECK(g.kh > 0 && g.kw > 0, "Filter dimensions cannot be zero.");
    if (p.has_pad_h || p.has_pad_w) { g.ph = (int)p.pad_h; g.pw = (int)p.pad_w; } else { g.ph = g.pw = (int)p.pad; }
    if (p.has_stride_h || p.has_stride_w) { g.sh = (int)p.stride_h; g.sw = (int)p.stride_w; } else { g.sh = g.sw = (int)p.stride; }
    if (g.ph || g.pw) SN_CHECK(g.ph < g.kh && g.pw < g.kw, "pad must be smaller than the kernel");
    g.PH = (int)std::ceil((float)(g.H + 2 * g.ph - g.kh) / g.sh) + 1;   // pooling_layer.cpp:90-93 (ceil)
    g.PW = (int)std::ceil((float)(g.W + 2 * g.pw - g.kw) / g.sw) + 1;
    if (g.ph || g.pw) {                                                 // :94-105
      if ((g.PH - 1) * g.sh >= g.H + g.ph) --g.PH;
      if ((g.PW - 1) * g.sw >= g.W + g.pw) --g.PW;
    }
    top[0]->reshape({g.N, g.C, g.PH, g.PW});
    if (method == 0 && !mask) SN_CUDA(cudaMalloc(&mask, top[0]->count() * sizeof(int)));
  }
  void forward(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    const float* x = bottom[0]->data->gpu();
    float* y = top[0]->data->mutable_gpu();
    if (method == 0) pool_max_fwd(S(), g, x, y, mask); else pool_ave_fwd(S(), g, x, y);
  }
  void backward(const vector<Blob*>& top, const vector<bool>& pd, const vector<Blob*>& bottom) override {
    if (!pd[0]) return;
    const float* dy = top[0]->diff->gpu();
    float* dx = bottom[0]->diff->mutable_gpu();
    if (method == 0) pool_max_bwd(S(), g, dy, mask, dx, gate_relu ? top[0]->data->gpu() : nullptr);
    else pool_ave_bwd(S(), g, dy, dx);
  }
};

// ReLU — reference relu_layer.{cpp:9-36,cu:9-58}.  When the producer's epilogue already applied it
// (fused_by_producer), forward is a no-op; backward is unchanged (tests y > 0, same predicate in place).
struct ReLULayer : Layer {
  bool fused_by_producer = false;
  bool bwd_fused_by_consumer = false;   // a max-pooling consumer applies dx = dy * (y > 0) while it scatters (PoolingLayer::gate_relu)
  using Layer::Layer;
  const char* type() const override { return "ReLU"; }
  void setup(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    if (top[0] != bottom[0]) top[0]->reshape(bottom[0]->shape);
  }
  void forward(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    if (fused_by_producer) return;
    const float* x = bottom[0]->data->gpu();
    relu_fwd(S(), bottom[0]->count(), x, top[0]->data->mutable_gpu(), lp.relu_negative_slope);
  }
  void backward(const vector<Blob*>& top, const vector<bool>& pd, const vector<Blob*>& bottom) override {
    if (!pd[0] || bwd_fused_by_consumer) return;
    const float* dy = top[0]->diff->gpu();
    const float* x = bottom[0]->data->gpu();
    relu_bwd(S(), bottom[0]->count(), dy, x, bottom[0]->diff->mutable_gpu(), lp.relu_negative_slope);
  }
};

// Dropout — reference dropout_layer.{cpp:11-65,cu:10-67}
struct DropoutLayer : Layer {
  float ratio = 0.5f, scale = 2.f;
  uint8_t* mask = nullptr;
  size_t n = 0;
  bool injected = false;
  uint64_t seed = 0, offset = 0;
  const unsigned long long* offset_slot = nullptr;   // inside Solver::step: the counter is read from device memory
  void advance() { offset += (n + 3) / 4; }
  using Layer::Layer;
  ~DropoutLayer() override { if (mask) cudaFree(mask); }
  const char* type() const override { return "Dropout"; }
  void setup(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    ratio = lp.dropout_ratio;
    SN_CHECK(ratio > 0.f && ratio < 1.f, "dropout_ratio must be in (0,1)");
    scale = (float)(1. / (1. - ratio));
    if (top[0] != bottom[0]) top[0]->reshape(bottom[0]->shape);
    n = bottom[0]->count();
    if (!mask) SN_CUDA(cudaMalloc(&mask, n));
  }
  void forward(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    const float* x = bottom[0]->data->gpu();
    float* y = top[0]->data->mutable_gpu();
    if (phase == cp::TRAIN) {
      dropout_fwd(S(), n, x, y, mask, ratio, scale, !injected, seed, offset, offset_slot);
      advance();
    } else if (x != y) {
      SN_CUDA(cudaMemcpyAsync(y, x, n * sizeof(float), cudaMemcpyDeviceToDevice, S()));
    }
  }
  void backward(const vector<Blob*>& top, const vector<bool>& pd, const vector<Blob*>& bottom) override {
    if (!pd[0]) return;
    const float* dy = top[0]->diff->gpu();
    float* dx = bottom[0]->diff->mutable_gpu();
    if (phase == cp::TRAIN) dropout_bwd(S(), n, dy, mask, dx, scale);
    else if (dx != dy) SN_CUDA(cudaMemcpyAsync(dx, dy, n * sizeof(float), cudaMemcpyDeviceToDevice, S()));
  }
};

// LRN — reference lrn_layer.{cpp:9-247,cu:9-190}
struct LRNLayer : Layer {
  LrnGeom g;
  int region = 0;
  ConvolutionLayer* flat_consumer = nullptr;   // see ConvolutionLayer::flat_consumer
  using Layer::Layer;
  const char* type() const override { return "LRN"; }
  void setup(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    const cp::LRNParameter& p = lp.lrn_param;
    g.size = (int)p.local_size;
    SN_CHECK(g.size % 2 == 1, "LRN only supports odd values for local_size");
    g.alpha = p.alpha; g.beta = p.beta; g.k = p.k; region = p.norm_region;
    SN_CHECK(bottom[0]->shape.size() == 4, "LRN %s: input must have 4 axes", lp.name.c_str());
    g.N = bottom[0]->num(); g.C = bottom[0]->channels(); g.H = bottom[0]->height(); g.W = bottom[0]->width();
    top[0]->reshape(bottom[0]->shape);
  }
  void forward(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    const float* x = bottom[0]->data->gpu();
    float* y = top[0]->data->mutable_gpu();
    if (region == 0) {
      FlatOut fo{nullptr, 0, 0, 0, 0};
      const bool feed_next = flat_consumer && flat_consumer->takes_flat_input();
      if (feed_next) fo = flat_consumer->flat_target();
      lrn_across_fwd(S(), g, x, y, feed_next ? &fo : nullptr);
      if (feed_next) flat_consumer->flat_written(top[0]);
    } else {
      lrn_within_fwd(S(), g, x, y);
    }
  }
  void backward(const vector<Blob*>& top, const vector<bool>& pd, const vector<Blob*>& bottom) override {
    if (!pd[0]) return;
    const float* dy = top[0]->diff->gpu();
    const float* x = bottom[0]->data->gpu();
    float* dx = bottom[0]->diff->mutable_gpu();
    if (region == 0) lrn_across_bwd(S(), g, x, dy, dx);
    else lrn_within_bwd(S(), g, x, dy, dx);
  }
};

// SoftmaxWithLoss — reference softmax_loss_layer.{cpp:11-119,cu:11-120}, softmax_layer.{cpp,cu}, loss_layer.cpp:8-23
struct SoftmaxWithLossLayer : Layer {
  int rows = 0, classes = 0;
  float* prob = nullptr;
  float* row_scratch = nullptr;
  using Layer::Layer;
  ~SoftmaxWithLossLayer() override { if (prob) cudaFree(prob); if (row_scratch) cudaFree(row_scratch); }
  const char* type() const override { return "SoftmaxWithLoss"; }
  bool is_loss() const override { return true; }
  void setup(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    if (lp.loss_weight.empty()) lp.loss_weight.push_back(1.f);
    SN_CHECK(bottom.size() == 2, "SoftmaxWithLoss needs (prediction, label) bottoms");
    const int axis = lp.softmax_axis < 0 ? lp.softmax_axis + (int)bottom[0]->shape.size() : lp.softmax_axis;
    rows = (int)bottom[0]->count(0, axis);
    classes = bottom[0]->shape[(size_t)axis];
    const size_t inner = bottom[0]->count(axis + 1, (int)bottom[0]->shape.size());
    SN_CHECK(inner == 1, "SoftmaxWithLoss %s: spatial softmax (inner_num > 1) is not on the supported path", lp.name.c_str());
    SN_CHECK((size_t)rows == bottom[1]->count(), "Number of labels must match number of predictions");
    SN_CHECK(!lp.loss_param.has_ignore_label && lp.loss_param.normalize, "ignore_label / normalize=false are not on the supported path");
    if (!prob) {
      SN_CUDA(cudaMalloc(&prob, (size_t)rows * classes * sizeof(float)));
      SN_CUDA(cudaMalloc(&row_scratch, (size_t)rows * sizeof(float)));
    }
    top[0]->reshape({});
  }
  void forward(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    softmax_loss_fwd(S(), rows, classes, bottom[0]->data->gpu(), bottom[1]->data->gpu(), prob,
                     top[0]->data->mutable_gpu(), row_scratch, net->loss_log);
  }
  void backward(const vector<Blob*>& top, const vector<bool>& pd, const vector<Blob*>& bottom) override {
    SN_CHECK(pd.size() < 2 || !pd[1], "SoftmaxWithLoss Layer cannot backpropagate to label inputs.");
    if (!pd[0]) return;
    softmax_loss_bwd(S(), rows, classes, prob, bottom[1]->data->gpu(), top[0]->diff->gpu(), bottom[0]->diff->mutable_gpu());
  }
};

// Accuracy (top-1) — reference accuracy_layer.cpp:13-103 (CPU only there; on the device here)
struct AccuracyLayer : Layer {
  int rows = 0, classes = 0;
  float* row_scratch = nullptr;
  using Layer::Layer;
  ~AccuracyLayer() override { if (row_scratch) cudaFree(row_scratch); }
  const char* type() const override { return "Accuracy"; }
  void setup(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    SN_CHECK(lp.accuracy_param.top_k == 1, "Accuracy %s: only top_k = 1 is on the supported path", lp.name.c_str());
    SN_CHECK(!lp.accuracy_param.has_ignore_label, "Accuracy ignore_label is not on the supported path");
    const int axis = lp.accuracy_param.axis < 0 ? lp.accuracy_param.axis + (int)bottom[0]->shape.size() : lp.accuracy_param.axis;
    rows = (int)bottom[0]->count(0, axis);
    classes = bottom[0]->shape[(size_t)axis];
    SN_CHECK(bottom[0]->count(axis + 1, (int)bottom[0]->shape.size()) == 1, "Accuracy: inner_num > 1 not supported");
    SN_CHECK((size_t)rows == bottom[1]->count(), "Number of labels must match number of predictions");
    if (!row_scratch) SN_CUDA(cudaMalloc(&row_scratch, (size_t)rows * sizeof(float)));
    top[0]->reshape({});
  }
  void forward(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    accuracy_top1(S(), rows, classes, bottom[0]->data->gpu(), bottom[1]->data->gpu(), top[0]->data->mutable_gpu(), row_scratch);
  }
  void backward(const vector<Blob*>&, const vector<bool>&, const vector<Blob*>&) override {}
};

// Split — reference split_layer.cpp:10-49: tops alias the bottom's data; backward sums the diffs.
struct SplitLayer : Layer {
  using Layer::Layer;
  const char* type() const override { return "Split"; }
  void setup(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    for (Blob* t : top) { t->reshape(bottom[0]->shape); t->share_data(*bottom[0]); }
  }
  void forward(const vector<Blob*>& bottom, const vector<Blob*>& top) override {
    for (Blob* t : top) t->share_data(*bottom[0]);
  }
  void backward(const vector<Blob*>& top, const vector<bool>& pd, const vector<Blob*>& bottom) override {
    if (!pd[0]) return;
    const size_t n = bottom[0]->count();
    float* bd = bottom[0]->diff->mutable_gpu();
    if (top.size() == 1) {
      SN_CUDA(cudaMemcpyAsync(bd, top[0]->diff->gpu(), n * sizeof(float), cudaMemcpyDeviceToDevice, S()));
      return;
    }
    split_add(S(), n, top[0]->diff->gpu(), top[1]->diff->gpu(), bd);
    for (size_t k = 2; k < top.size(); ++k) axpy(S(), n, 1.f, top[k]->diff->gpu(), bd);
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
  fatal("layer type '%s' (layer '%s') is outside the supported hot path", t.c_str(), lp.name.c_str());
}

// ------------------------------------------------------------------------------------------
// InsertSplits — reference caffe/src/caffe/util/insert_splits.cpp:12-137
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
    const cp::LayerParameter& lp = layers[(size_t)i];
    for (int j = 0; j < (int)lp.bottom.size(); ++j) {
      auto it = last_top.find(lp.bottom[(size_t)j]);
      SN_CHECK(it != last_top.end(), "Unknown bottom blob '%s' (layer '%s', bottom index %d)", lp.bottom[(size_t)j].c_str(), lp.name.c_str(), j);
      bottom_src[{i, j}] = it->second;
      ++top_count[it->second];
    }
    for (int j = 0; j < (int)lp.top.size(); ++j) last_top[lp.top[(size_t)j]] = {i, j};
    const int last_loss = std::min((int)lp.loss_weight.size(), (int)lp.top.size());
    for (int j = 0; j < last_loss; ++j) {
      const PI t = last_top[lp.top[(size_t)j]];
      top_lw[t] = lp.loss_weight[(size_t)j];
      if (top_lw[t]) ++top_count[t];
    }
  }
  vector<cp::LayerParameter> out;
  for (int i = 0; i < (int)layers.size(); ++i) {
    cp::LayerParameter lp = layers[(size_t)i];
    for (int j = 0; j < (int)lp.bottom.size(); ++j) {
      const PI t = bottom_src[{i, j}];
      if (top_count[t] > 1)
        lp.bottom[(size_t)j] = split_blob_name(layers[(size_t)t.first].name, lp.bottom[(size_t)j], t.second, split_idx[t]++);
    }
    vector<cp::LayerParameter> splits;
    for (int j = 0; j < (int)lp.top.size(); ++j) {
      const PI t{i, j};
      const int sc = top_count[t];
      if (sc > 1) {
        cp::LayerParameter sp;
        sp.bottom.push_back(lp.top[(size_t)j]);
        sp.name = split_layer_name(lp.name, lp.top[(size_t)j], j);
        sp.type = "Split";
        const float lw = top_lw[t];
        for (int k = 0; k < sc; ++k) {
          sp.top.push_back(split_blob_name(lp.name, lp.top[(size_t)j], j, k));
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
// Net — reference caffe/src/caffe/net.cpp:40-284 (Init), 385-562 (Append*), 565-581, 635-645,
// 737-766 (ShareTrainedLayersWith)
// ------------------------------------------------------------------------------------------
Net::Net(const cp::NetParameter& in, int phase_, uint64_t seed, const Net* share_from) : phase(phase_), rng_seed(seed), shares_with(share_from) {
  cp::NetState st = in.state;
  st.phase = phase_;
  vector<cp::LayerParameter> param = insert_splits(cp::filter_net(in, st));
  name = in.name;
  SN_CHECK(in.input.empty(), "net-level `input` blobs are not on the supported path (use JavaData layers)");
  std::map<string, int> name_to_idx;
  std::set<string> available;
  const int L = (int)param.size();
  bottom_vecs.resize((size_t)L); top_vecs.resize((size_t)L); bottom_ids.resize((size_t)L); top_ids.resize((size_t)L);
  bottom_need_backward.resize((size_t)L);
  for (int li = 0; li < L; ++li) {
    cp::LayerParameter lp = param[(size_t)li];
    if (!lp.has_phase) { lp.phase = phase; lp.has_phase = true; }
    std::shared_ptr<Layer> layer = create_layer(lp);
    layer->phase = lp.phase;
    layer->net = this;
    if (share_from) {                                  // adopt the other net's parameters before setup(): no allocation, no filler
      for (size_t k = 0; k < share_from->layers.size(); ++k)
        if (share_from->layer_names[k] == lp.name && !share_from->layers[k]->blobs.empty()) layer->blobs = share_from->layers[k]->blobs;
    }
    const bool adopted = !layer->blobs.empty();
    layers.push_back(layer);
    layer_names.push_back(lp.name);
    bool need_backward = false;
    for (int b = 0; b < (int)lp.bottom.size(); ++b) {
      const string& bn = lp.bottom[(size_t)b];
      SN_CHECK(available.count(bn), "Unknown bottom blob '%s' (layer '%s', bottom index %d)", bn.c_str(), lp.name.c_str(), b);
      const int id = name_to_idx[bn];
      bottom_vecs[(size_t)li].push_back(blobs[(size_t)id].get());
      bottom_ids[(size_t)li].push_back(id);
      available.erase(bn);
      const bool pd = lp.propagate_down.empty() ? true : (bool)lp.propagate_down[(size_t)b];
      bottom_need_backward[(size_t)li].push_back(blob_need_backward[(size_t)id] && pd);
      need_backward |= blob_need_backward[(size_t)id];
    }
    for (int t = 0; t < (int)lp.top.size(); ++t) {
      const string& tn = lp.top[(size_t)t];
      if (t < (int)lp.bottom.size() && lp.bottom[(size_t)t] == tn) {   // in-place computation
        const int id = name_to_idx[tn];
        top_vecs[(size_t)li].push_back(blobs[(size_t)id].get());
        top_ids[(size_t)li].push_back(id);
      } else {
        SN_CHECK(!name_to_idx.count(tn), "Top blob '%s' produced by multiple sources.", tn.c_str());
        blobs.push_back(std::make_shared<Blob>());
        const int id = (int)blobs.size() - 1;
        blob_names.push_back(tn);
        blob_need_backward.push_back(false);
        name_to_idx[tn] = id;
        top_vecs[(size_t)li].push_back(blobs[(size_t)id].get());
        top_ids[(size_t)li].push_back(id);
      }
      available.insert(tn);
    }
    layer->setup(bottom_vecs[(size_t)li], top_vecs[(size_t)li]);
    layer->set_loss_weights(top_vecs[(size_t)li]);
    const int nparam = (int)layer->blobs.size();
    SN_CHECK((int)lp.param.size() <= nparam, "Too many params specified for layer %s", lp.name.c_str());
    for (int p = 0; p < nparam; ++p) {
      cp::ParamSpec spec;
      if (p < (int)lp.param.size()) spec = lp.param[(size_t)p];
      SN_CHECK(spec.name.empty(), "named (shared) params are not on the supported path");
      const bool pnb = spec.lr_mult != 0.f;
      need_backward |= pnb;
      layer->param_propagate_down[(size_t)p] = pnb;
      if (adopted) continue;                           // owned, placed and updated by the net they were adopted from
      learnable_params.push_back(layer->blobs[(size_t)p].get());
      params_lr.push_back(spec.lr_mult);
      params_decay.push_back(spec.decay_mult);
    }
    layer_need_backward.push_back(need_backward);
    if (need_backward)
      for (int id : top_ids[(size_t)li]) blob_need_backward[(size_t)id] = true;
    if (DropoutLayer* d = dynamic_cast<DropoutLayer*>(layer.get())) d->seed = rng_seed * 0x9E3779B97F4A7C15ull + (uint64_t)li;
  }
  // which blobs contribute to the loss (net.cpp:191-247)
  std::set<string> under_loss, skip_bp;
  for (int li = L - 1; li >= 0; --li) {
    bool contributes = false, skip_pd = true;
    for (int t = 0; t < (int)top_vecs[(size_t)li].size(); ++t) {
      const string& bn = blob_names[(size_t)top_ids[(size_t)li][(size_t)t]];
      if ((t < (int)layers[(size_t)li]->loss.size() && layers[(size_t)li]->loss[(size_t)t]) || under_loss.count(bn)) contributes = true;
      if (!skip_bp.count(bn)) skip_pd = false;
      if (contributes && !skip_pd) break;
    }
    if (layer_need_backward[(size_t)li] && skip_pd) {
      layer_need_backward[(size_t)li] = false;
      for (size_t b = 0; b < bottom_need_backward[(size_t)li].size(); ++b) bottom_need_backward[(size_t)li][b] = false;
    }
    if (!contributes) layer_need_backward[(size_t)li] = false;
    for (int b = 0; b < (int)bottom_vecs[(size_t)li].size(); ++b) {
      const string& bn = blob_names[(size_t)bottom_ids[(size_t)li][(size_t)b]];
      if (contributes) under_loss.insert(bn); else bottom_need_backward[(size_t)li][(size_t)b] = false;
      if (!bottom_need_backward[(size_t)li][(size_t)b]) skip_bp.insert(bn);
    }
  }
  if (in.force_backward) {   // net.cpp:248-265
    for (int li = 0; li < L; ++li) {
      layer_need_backward[(size_t)li] = true;
      for (size_t b = 0; b < bottom_need_backward[(size_t)li].size(); ++b) {
        const bool allow = !(layers[(size_t)li]->is_loss() && b == 1) && string(layers[(size_t)li]->type()) != "Accuracy";
        bottom_need_backward[(size_t)li][b] = bottom_need_backward[(size_t)li][b] || allow;
        const int id = bottom_ids[(size_t)li][b];
        blob_need_backward[(size_t)id] = blob_need_backward[(size_t)id] || bottom_need_backward[(size_t)li][b];
      }
      for (size_t p = 0; p < layers[(size_t)li]->param_propagate_down.size(); ++p) layers[(size_t)li]->param_propagate_down[p] = true;
    }
  }
  for (const string& bn : available) {   // std::set => name order (net.cpp:266-274)
    output_blobs.push_back(blobs[(size_t)name_to_idx[bn]].get());
    output_blob_indices.push_back(name_to_idx[bn]);
  }
  // Producer/ReLU fusion: Convolution or InnerProduct immediately followed by an in-place ReLU
  // (negative_slope 0) on its top applies the ReLU in its epilogue.
  for (int li = 0; li + 1 < L; ++li) {
    ReLULayer* r = dynamic_cast<ReLULayer*>(layers[(size_t)li + 1].get());
    if (!r || r->lp.relu_negative_slope != 0.f) continue;
    if (bottom_vecs[(size_t)li + 1][0] != top_vecs[(size_t)li + 1][0] || bottom_vecs[(size_t)li + 1][0] != top_vecs[(size_t)li][0]) continue;
    if (ConvolutionLayer* c = dynamic_cast<ConvolutionLayer*>(layers[(size_t)li].get())) { c->fuse_relu = true; r->fused_by_producer = true; }
    else if (InnerProductLayer* ip = dynamic_cast<InnerProductLayer*>(layers[(size_t)li].get())) { ip->fuse_relu = true; r->fused_by_producer = true; }
  }
  // ReLU-backward / max-pool fusion: an in-place ReLU (slope 0) whose blob is read by exactly one later layer, a MAX
  // pooling layer, hands its backward to that layer (pool_max_bwd gate).  SPARKNET_FUSE_RELU_POOL=0 turns it off.
  {
    const char* e = getenv("SPARKNET_FUSE_RELU_POOL");
    const bool enabled = !(e && atoi(e) == 0);
    for (int li = 0; enabled && li < L; ++li) {
      ReLULayer* r = dynamic_cast<ReLULayer*>(layers[(size_t)li].get());
      if (!r || r->lp.relu_negative_slope != 0.f || bottom_vecs[(size_t)li][0] != top_vecs[(size_t)li][0]) continue;
      Blob* blob = top_vecs[(size_t)li][0];
      int consumers = 0, consumer = -1;
      bool rewritten = false;
      for (int lj = li + 1; lj < L; ++lj) {
        for (Blob* b : bottom_vecs[(size_t)lj]) if (b == blob) { ++consumers; consumer = lj; }
        for (Blob* t : top_vecs[(size_t)lj]) if (t == blob) rewritten = true;      // another in-place layer on the same blob
      }
      if (consumers != 1 || rewritten) continue;
      if (ConvolutionLayer* cl = dynamic_cast<ConvolutionLayer*>(layers[(size_t)consumer].get())) {
        // ... or a convolution: its backward-data epilogue gates the gradient (conv_flat.cu WinEpi::gate).  SPARKNET_FUSE_RELU_CONV=0 turns it off.
        const char* ec = getenv("SPARKNET_FUSE_RELU_CONV");
        if (ec && atoi(ec) == 0) continue;
        if (bottom_vecs[(size_t)consumer].size() != 1) continue;
        cl->gate_relu = true;
        r->bwd_fused_by_consumer = true;
        continue;
      }
      PoolingLayer* pl = dynamic_cast<PoolingLayer*>(layers[(size_t)consumer].get());
      if (!pl || pl->method != 0) continue;
      pl->gate_relu = true;
      r->bwd_fused_by_consumer = true;
    }
  }
  // A JavaData layer whose blob is read by exactly one convolution can hand it byte pixels (JavaDataLayer byte input)
  for (int li = 0; li < L; ++li) {
    JavaDataLayer* jd = dynamic_cast<JavaDataLayer*>(layers[(size_t)li].get());
    if (!jd) continue;
    Blob* xb = top_vecs[(size_t)li][0];
    ConvolutionLayer* only = nullptr;
    int readers = 0;
    for (int lj = li + 1; lj < L; ++lj)
      for (Blob* b : bottom_vecs[(size_t)lj])
        if (b == xb) { ++readers; only = dynamic_cast<ConvolutionLayer*>(layers[(size_t)lj].get()); }
    if (readers != 1 || !only) continue;
    jd->consumer_takes_bytes = [only] { return only->takes_bytes(); };
    jd->consumer_feed = [only](const ByteFeed& f) { only->byte_feed = f; };
  }
  // Producer-side packing: a window convolution whose input blob is computed by another convolution or an
  // across-channel LRN (and rewritten by nothing but in-place ReLUs the producer applies itself) has its flat grid
  // written by that producer.  SPARKNET_FUSE_PACK=0 turns it off.
  {
    const char* e = getenv("SPARKNET_FUSE_PACK");
    const bool enabled = !(e && atoi(e) == 0);
    for (int lj = 0; enabled && lj < L; ++lj) {
      ConvolutionLayer* consumer = dynamic_cast<ConvolutionLayer*>(layers[(size_t)lj].get());
      if (!consumer || bottom_vecs[(size_t)lj].size() != 1) continue;
      Blob* xb = bottom_vecs[(size_t)lj][0];
      int prod = -1;
      bool clean = true;
      for (int li = 0; li < lj; ++li)
        for (Blob* t : top_vecs[(size_t)li]) {
          if (t != xb) continue;
          const bool inplace = !bottom_vecs[(size_t)li].empty() && bottom_vecs[(size_t)li][0] == xb;
          if (!inplace) { prod = li; continue; }
          ReLULayer* r = dynamic_cast<ReLULayer*>(layers[(size_t)li].get());
          if (!(r && r->fused_by_producer)) clean = false;
        }
      if (prod < 0 || !clean) continue;
      if (ConvolutionLayer* p = dynamic_cast<ConvolutionLayer*>(layers[(size_t)prod].get())) p->flat_consumer = consumer;
      else if (LRNLayer* n = dynamic_cast<LRNLayer*>(layers[(size_t)prod].get())) { if (n->region == 0) n->flat_consumer = consumer; }
    }
  }
  // Parameter arena: one contiguous device allocation for all weights and one for all gradients;
  // every blob starts on a 1024-float boundary (lets one SGD launch and one all-reduce cover the net).
  size_t off = 0;
  for (Blob* b : learnable_params) {
    param_offsets.push_back(off);
    off += (b->count() + 1023) / 1024 * 1024;
  }
  arena_floats = off;
  if (arena_floats) {
    SN_CUDA(cudaMalloc(&arena_w, arena_floats * sizeof(float)));
    SN_CUDA(cudaMalloc(&arena_g, arena_floats * sizeof(float)));
    SN_CUDA(cudaMemset(arena_w, 0, arena_floats * sizeof(float)));
    SN_CUDA(cudaMemset(arena_g, 0, arena_floats * sizeof(float)));
    SN_CUDA(cudaMalloc(&arena_s, arena_floats * sizeof(float)));
    SN_CUDA(cudaMemset(arena_s, 0, arena_floats * sizeof(float)));
    owns_arena = true;
    // pinned mirror of the weight arena: the fillers' host copies move into it, every weight blob's host pointer
    // (what get_data hands out) points into it from now on
    param_arena.dev = arena_w;
    param_arena.floats = arena_floats;
    SN_CUDA(cudaMallocHost(&param_arena.host, arena_floats * sizeof(float)));
    memset(param_arena.host, 0, arena_floats * sizeof(float));
    for (size_t i = 0; i < learnable_params.size(); ++i) {
      Blob* b = learnable_params[i];
      SN_CHECK(b->data->dev == nullptr && b->diff->dev == nullptr, "param blob touched before arena placement");
      b->data->dev = arena_w + param_offsets[i];
      b->data->owns_dev = false;
      b->diff->dev = arena_g + param_offsets[i];
      b->diff->owns_dev = false;
      b->diff->head = Storage::AT_DEVICE;   // zeroed above
      Storage* st = b->data.get();
      float* slot = param_arena.host + param_offsets[i];
      if (st->host) {
        memcpy(slot, st->host, st->count * sizeof(float));
        if (st->owns_host) SN_CUDA(cudaFreeHost(st->host));
      }
      st->host = slot;
      st->owns_host = false;
      st->arena = &param_arena;
      param_arena.members.push_back(st);
      param_arena.offsets.push_back(param_offsets[i]);
    }
  }
}
void Net::params_to_device() {
  if (shares_with) { const_cast<Net*>(shares_with)->params_to_device(); return; }
  if (param_arena.host) param_arena.upload();
}
Net::~Net() {
  // weight blobs may outlive this net in a net that adopted them (they are destroyed with it, right after):
  // detach them from the arena and its mirror before both go away
  for (Storage* st : param_arena.members) { st->arena = nullptr; st->host = nullptr; st->dev = nullptr; }
  layers.clear();
  blobs.clear();
  if (owns_arena) {
    cudaFree(arena_w);
    cudaFree(arena_g);
    cudaFree(arena_s);
    if (param_arena.host) cudaFreeHost(param_arena.host);
  }
}
// after sgd_update wrote the rounded weights into the shadow arena: layers whose shadow lives there need no re-rounding
void Net::mark_shadows_current() {
  for (auto& l : layers)
    if (l->w_shadow_in_arena && !l->blobs.empty()) l->w_shadow_gen = l->blobs[0]->data->gen;
}
void Net::prepack_filters() {
  FilterPackBatch b{};
  for (size_t i = 0; i < layers.size(); ++i)
    if (ConvolutionLayer* c = dynamic_cast<ConvolutionLayer*>(layers[i].get()))
      c->queue_filter_packs(b, layer_need_backward[i] && bottom_need_backward[i][0]);
  tcw_filter_pack_batch(ctx().compute, b);
}
void Net::forward_from_to(int start, int end) {
  for (int i = start; i <= end; ++i) layers[(size_t)i]->forward(bottom_vecs[(size_t)i], top_vecs[(size_t)i]);
}
void Net::backward_from_to(int start, int end) {
  for (int i = start; i >= end; --i)
    if (layer_need_backward[(size_t)i]) layers[(size_t)i]->backward(top_vecs[(size_t)i], bottom_need_backward[(size_t)i], bottom_vecs[(size_t)i]);
}
void Net::share_trained_layers_with(const Net& other) {
  for (size_t i = 0; i < other.layers.size(); ++i) {
    const string& src = other.layer_names[i];
    size_t t = 0;
    while (t != layer_names.size() && layer_names[t] != src) ++t;
    if (t == layer_names.size()) continue;
    SN_CHECK(layers[t]->blobs.size() == other.layers[i]->blobs.size(), "Incompatible number of blobs for layer %s", src.c_str());
    for (size_t j = 0; j < layers[t]->blobs.size(); ++j) {
      SN_CHECK(layers[t]->blobs[j]->shape == other.layers[i]->blobs[j]->shape, "Cannot share param %zu weights from layer '%s'", j, src.c_str());
      layers[t]->blobs[j] = other.layers[i]->blobs[j];
    }
  }
}
float Net::total_loss() {
  float l = 0.f;
  for (size_t li = 0; li < layers.size(); ++li)
    for (size_t t = 0; t < top_vecs[li].size() && t < layers[li]->loss.size(); ++t) {
      if (layers[li]->loss[t] == 0.f) continue;
      Blob* b = top_vecs[li][t];
      // read back without disturbing the head state of the blob
      std::vector<float> d(b->count()), w(b->count());
      SN_CUDA(cudaMemcpyAsync(d.data(), b->data->gpu(), d.size() * sizeof(float), cudaMemcpyDeviceToHost, S()));
      SN_CUDA(cudaMemcpyAsync(w.data(), b->diff->gpu(), w.size() * sizeof(float), cudaMemcpyDeviceToHost, S()));
      SN_CUDA(cudaStreamSynchronize(S()));
      for (size_t k = 0; k < d.size(); ++k) l += d[k] * w[k];
    }
  return l;
}

// ------------------------------------------------------------------------------------------
// Solver — reference caffe/src/caffe/solver.cpp:44-190,193-282,413-444 and
// caffe/src/caffe/solvers/sgd_solver.cpp:27-63,66-78,102-116,145-239
// ------------------------------------------------------------------------------------------
Solver::Solver(const cp::SolverParameter& p) : param(p) {
  SN_CHECK(param.type == "SGD", "only the SGD solver is on the path (libccaffe/ccaffe.cpp:131 hard-codes SGDSolver)");
  SN_CHECK(param.has_net_param || param.has_train_net_param, "SolverParameter must embed the net (net_param); file paths are not supported");
  SN_CHECK(param.regularization_type == "L2", "only L2 regularization is on the supported path");
  SN_CHECK(param.clip_gradients < 0, "clip_gradients is not on the supported path");
  SN_CHECK(param.iter_size == 1, "iter_size != 1 is not on the supported path");
  uint64_t seed = param.random_seed >= 0 ? (uint64_t)param.random_seed : (uint64_t)std::random_device{}();
  if (param.random_seed >= 0) host_rng().seed((uint32_t)param.random_seed);   // solver.cpp:52-54
  const cp::NetParameter& np = param.has_train_net_param ? param.train_net_param : param.net_param;
  net.reset(new Net(np, cp::TRAIN, seed));
  // the TEST net adopts the TRAIN net's parameter blobs while it is built (net.cpp:737-766): no second set of arenas
  if (param.has_net_param) test_net.reset(new Net(param.net_param, cp::TEST, seed + 1, net.get()));
  else if (!param.test_net_param.empty()) test_net.reset(new Net(param.test_net_param[0], cp::TEST, seed + 1, net.get()));
  const size_t n = net->arena_floats;
  if (n) {
    SN_CUDA(cudaMalloc(&arena_h, n * sizeof(float)));
    SN_CUDA(cudaMemset(arena_h, 0, n * sizeof(float)));
    std::vector<SgdChunk> table(n / 1024);
    for (size_t i = 0; i < net->learnable_params.size(); ++i) {
      const size_t c0 = net->param_offsets[i] / 1024, c1 = c0 + (net->learnable_params[i]->count() + 1023) / 1024;
      for (size_t c = c0; c < c1; ++c) table[c] = SgdChunk{net->params_lr[i], param.weight_decay * net->params_decay[i]};
    }
    SN_CUDA(cudaMalloc(&chunk_table, table.size() * sizeof(SgdChunk)));
    SN_CUDA(cudaMemcpy(chunk_table, table.data(), table.size() * sizeof(SgdChunk), cudaMemcpyHostToDevice));
  }
  for (size_t i = 0; i < net->learnable_params.size(); ++i) param_index[net->learnable_params[i]] = i;
  SN_CUDA(cudaMalloc(&loss_hist, (kLossHist + 1) * sizeof(float)));      // last slot: the loss of a replayed iteration
  SN_CUDA(cudaMemset(loss_hist, 0, (kLossHist + 1) * sizeof(float)));
  SN_CUDA(cudaMalloc(&scalars_dev, sizeof(StepScalars)));
  SN_CUDA(cudaMemset(scalars_dev, 0, sizeof(StepScalars)));
  SN_CUDA(cudaMallocHost(&loss_hist_host, kLossHist * sizeof(float)));
  memset(loss_hist_host, 0, kLossHist * sizeof(float));
}
Solver::~Solver() {
  test_net.reset();
  net.reset();
  if (arena_h) cudaFree(arena_h);
  if (chunk_table) cudaFree(chunk_table);
  if (loss_hist) cudaFree(loss_hist);
  if (loss_hist_host) cudaFreeHost(loss_hist_host);
  drop_graphs();
  if (scalars_dev) cudaFree(scalars_dev);
}
// Configuration epoch: bumped by every ABI call that can change WHICH kernels an iteration launches or with what
// arguments (precision, resident inputs, injected masks, development switches, callbacks); recordings of an older
// epoch are discarded.
static uint64_t g_config_epoch = 1;
void bump_config_epoch() { ++g_config_epoch; }
// Programmatic dependent launch inside a recorded iteration: every kernel-to-kernel edge of the captured graph becomes a
// programmatic edge, so the next kernel's CTAs are scheduled, and run their prologue (barrier set-up, TMEM allocation,
// descriptor prefetch: 8-12 us for a tcgen05 kernel), while the last CTAs of the previous kernel drain.  Correctness rests
// on every kernel of this library executing griddepcontrol.wait before its first global-memory access (SN_GRID_DEP_WAIT):
// the wait returns when the previous kernel has completed and flushed, and by induction so have all before it.  Edges that
// touch a non-kernel node (memcpy, memset) stay full dependencies.  SPARKNET_PDL=0 leaves the graph as captured.
static int make_edges_programmatic(cudaGraph_t graph) {
  static const bool enabled = !(getenv("SPARKNET_PDL") && atoi(getenv("SPARKNET_PDL")) == 0);
  if (!enabled) return 0;
  size_t n = 0;
  static const bool debug = getenv("SPARKNET_PDL_DEBUG") != nullptr;
  cudaError_t e0 = cudaGraphGetEdges_v2(graph, nullptr, nullptr, nullptr, &n);
  if (debug) fprintf(stderr, "sparknet pdl: %zu edges (%s)\n", n, cudaGetErrorString(e0));
  if (e0 != cudaSuccess || n == 0) { cudaGetLastError(); return 0; }
  std::vector<cudaGraphNode_t> from(n), to(n);
  std::vector<cudaGraphEdgeData> ed(n);
  if (cudaGraphGetEdges_v2(graph, from.data(), to.data(), ed.data(), &n) != cudaSuccess) { cudaGetLastError(); return 0; }
  // a kernel may have ONE programmatic predecessor; the captured iteration is a chain, anything else is left alone
  // ... and a kernel with several successors (the fork to the backward-filter stream) keeps full dependencies
  std::map<cudaGraphNode_t, int> in_degree, out_degree;
  for (size_t i = 0; i < n; ++i) { ++in_degree[to[i]]; ++out_degree[from[i]]; }
  int converted = 0;
  for (size_t i = 0; i < n; ++i) {
    cudaGraphNodeType tf, tt;
    if (cudaGraphNodeGetType(from[i], &tf) != cudaSuccess || cudaGraphNodeGetType(to[i], &tt) != cudaSuccess) { cudaGetLastError(); continue; }
    if (tf != cudaGraphNodeTypeKernel || tt != cudaGraphNodeTypeKernel || in_degree[to[i]] != 1 || out_degree[from[i]] != 1) continue;
    if (ed[i].type == cudaGraphDependencyTypeProgrammatic) continue;
    cudaGraphEdgeData prog{};
    prog.from_port = cudaGraphKernelNodePortProgrammatic;
    prog.to_port = 0;
    prog.type = cudaGraphDependencyTypeProgrammatic;
    cudaError_t er = cudaGraphRemoveDependencies_v2(graph, &from[i], &to[i], &ed[i], 1);
    if (er != cudaSuccess) { if (debug) fprintf(stderr, "sparknet pdl: remove edge %zu: %s\n", i, cudaGetErrorString(er)); cudaGetLastError(); continue; }
    er = cudaGraphAddDependencies_v2(graph, &from[i], &to[i], &prog, 1);
    if (er != cudaSuccess) {
      if (debug) fprintf(stderr, "sparknet pdl: add programmatic edge %zu: %s\n", i, cudaGetErrorString(er));
      cudaGetLastError();
      SN_CUDA(cudaGraphAddDependencies_v2(graph, &from[i], &to[i], &ed[i], 1));     // put the full dependency back
      continue;
    }
    ++converted;
  }
  if (debug) fprintf(stderr, "sparknet pdl: %d of %zu edges programmatic\n", converted, n);
  return converted;
}
void Solver::drop_graphs() {
  for (GraphEntry& g : graphs) if (g.exec) cudaGraphExecDestroy(g.exec);
  graphs.clear();
  eager_iters = 0;
}
void Solver::update_layer_early(Layer* l) {
  // the layer's blobs sit next to each other in the arena (weights, then bias, each on a 1024-float boundary)
  size_t lo = ~(size_t)0, hi = 0;
  for (auto& b : l->blobs) {
    auto it = param_index.find(b.get());
    if (it == param_index.end() || sgd_done[it->second]) return;
    const size_t off = net->param_offsets[it->second], len = (b->count() + 1023) / 1024 * 1024;
    lo = std::min(lo, off);
    hi = std::max(hi, off + len);
  }
  size_t covered = 0;
  for (auto& b : l->blobs) covered += (b->count() + 1023) / 1024 * 1024;
  if (hi <= lo || covered != hi - lo) return;                     // not contiguous: leave it to update_rest
  sgd_update(ctx().side, hi - lo, net->arena_w + lo, net->arena_g + lo, arena_h + lo, net->arena_s + lo, chunk_table + lo / 1024, 0.f,
             param.momentum, &scalars_dev->rate);
  for (auto& b : l->blobs) sgd_done[param_index[b.get()]] = 1;
}
void Solver::update_rest(cudaStream_t s) {
  const size_t n = net->learnable_params.size();
  size_t i = 0;
  while (i < n) {
    if (sgd_done[i]) { ++i; continue; }
    size_t j = i;
    while (j < n && !sgd_done[j]) ++j;
    const size_t lo = net->param_offsets[i], hi = j < n ? net->param_offsets[j] : net->arena_floats;
    sgd_update(s, hi - lo, net->arena_w + lo, net->arena_g + lo, arena_h + lo, net->arena_s + lo, chunk_table + lo / 1024, 0.f, param.momentum,
               &scalars_dev->rate);
    i = j;
  }
  std::fill(sgd_done.begin(), sgd_done.end(), 0);
}
float Solver::get_learning_rate() {   // sgd_solver.cpp:27-63, same expression types
  const string& pol = param.lr_policy;
  float rate = 0.f;
  if (pol == "fixed") rate = param.base_lr;
  else if (pol == "step") { current_step = iter / param.stepsize; rate = (float)(param.base_lr * std::pow(param.gamma, current_step)); }
  else if (pol == "exp") rate = (float)(param.base_lr * std::pow(param.gamma, iter));
  else if (pol == "inv") rate = param.base_lr * std::pow(1.f + param.gamma * iter, -param.power);
  else if (pol == "multistep") {
    if (current_step < (int)param.stepvalue.size() && iter >= param.stepvalue[(size_t)current_step]) current_step++;
    rate = (float)(param.base_lr * std::pow(param.gamma, current_step));
  } else if (pol == "poly") rate = param.base_lr * std::pow(1.f - ((float)iter / (float)param.max_iter), param.power);
  else if (pol == "sigmoid") rate = param.base_lr * (1.f / (1.f + std::exp(-param.gamma * ((float)iter - (float)param.stepsize))));
  else fatal("Unknown learning rate policy: %s", pol.c_str());
  return rate;
}
void Solver::step(int iters) {
  const int stop_iter = iter + iters;
  const int L = (int)net->layers.size();
  // nothing touched any blob since the previous call returned: iteration one may replay too
  uint64_t touch_seen = step_calls > 0 ? touch_at_exit : ~0ull;
  ++step_calls;
  g_in_step = true;
  g_side_allowed = true;
  {
    static const bool early = !(getenv("SPARKNET_SGD_EARLY") && atoi(getenv("SPARKNET_SGD_EARLY")) == 0);
    sgd_done.assign(net->learnable_params.size(), 0);
    if (early) net->on_param_grads_ready = [this](Layer* l) { update_layer_early(l); };
  }
  // every learnable blob's device copy must be current before the fused update touches the arena: host-written weights
  // go up in one copy of the pinned arena mirror
  // (a blob whose device copy is already the newest keeps its generation: its tf32 shadow / packed copies stay valid)
  net->params_to_device();
  for (Blob* b : net->learnable_params) {
    if (b->data->head != Storage::AT_DEVICE) b->data->mutable_gpu();
    b->diff->mutable_gpu();
  }
  net->overwrite_param_diffs = true;     // Net::ClearParamDiffs (net.cpp:990-1008) folded into the wgrad epilogues
  // In-call prefetch of the data callbacks (see JavaDataLayer): worth a thread when several iterations follow
  std::vector<JavaDataLayer*> feeds;
  for (auto& l : net->layers)
    if (JavaDataLayer* d = dynamic_cast<JavaDataLayer*>(l.get())) feeds.push_back(d);
  // opt-in (SPARKNET_PREFETCH=1): measured on B200 it changes nothing while the GPU, not the host, bounds the step
  // (CaffeNet: callback fill 3 ms + launches 0.5 ms per 4.2 ms step), and callbacks then arrive on a non-caller thread
  bool prefetch = iters >= 2 && !feeds.empty() && getenv("SPARKNET_PREFETCH") && atoi(getenv("SPARKNET_PREFETCH")) == 1;
  for (JavaDataLayer* d : feeds) if (d->pool || !d->cb) prefetch = false;
  JavaDataLayer::PrefetchSync sync;
  std::thread helper;
  if (prefetch) {
    const int device = ctx().device;
    for (JavaDataLayer* d : feeds) { d->pf = &sync; d->pf_filled = d->pf_enqueued = 0; d->pf_base_parity = d->parity; }
    helper = std::thread([&feeds, iters, device] {
      cudaSetDevice(device);
      for (long k = 0; k < iters; ++k)
        for (JavaDataLayer* d : feeds) d->prefetch_fill(k);      // data callback, then label callback: one pair per iteration
    });
  }
  // ---- CUDA-graph replay (see runtime.hpp Solver::GraphEntry) ----
  int D = 0;                                                   // leading JavaData layers run eagerly every iteration
  while (D < L && dynamic_cast<JavaDataLayer*>(net->layers[(size_t)D].get())) ++D;
  std::vector<DropoutLayer*> drops;
  for (auto& l : net->layers)
    if (DropoutLayer* d = dynamic_cast<DropoutLayer*>(l.get())) drops.push_back(d);
  bool use_graph = graph_mode >= 0 ? graph_mode == 1 : !(getenv("SPARKNET_GRAPH") && atoi(getenv("SPARKNET_GRAPH")) == 0);
  if (param.display || prefetch || (int)feeds.size() != D || (int)drops.size() > StepScalars::kMaxDropout || !net->arena_floats) use_graph = false;
  for (size_t i = 0; i < drops.size(); ++i) drops[i]->offset_slot = &scalars_dev->dropout_offset[i];
  cudaStream_t cs = ctx().compute;
  while (iter < stop_iter) {
    // recordings of an older configuration (or of an older scratch workspace, see get_workspace) are stale
    if (graphs_epoch != g_config_epoch) { drop_graphs(); graphs_epoch = g_config_epoch; }
    for (int i = 0; i < D; ++i) net->layers[(size_t)i]->forward(net->bottom_vecs[(size_t)i], net->top_vecs[(size_t)i]);
    // this iteration's scalars -> device memory (one tiny launch, skipped while nothing changed)
    StepScalars sc{};
    sc.rate = get_learning_rate();
    sc.n_dropout = (unsigned)drops.size();
    for (size_t i = 0; i < drops.size(); ++i) sc.dropout_offset[i] = drops[i]->offset;
    if (!scalars_published || memcmp(&sc, &scalars_last, sizeof(sc)) != 0) {
      publish_step_scalars(cs, scalars_dev, sc);
      scalars_last = sc;
      scalars_published = true;
    }
    const float rate = sc.rate;
    GraphEntry* entry = nullptr;
    // eager whenever something outside this loop touched a blob since the last iteration (always true for the first
    // iteration after get_data / setWeights / forward() / averaging ...; a data callback that calls back into the ABI)
    const bool dirty = g_outside_touch != touch_seen;
    touch_seen = g_outside_touch;
    bool replay = use_graph && !dirty && eager_iters >= 2;
    if (replay) {
      std::vector<const void*> key;
      for (int i = 0; i < D; ++i) key.push_back(net->top_vecs[(size_t)i][0]->data->dev);
      for (GraphEntry& g : graphs) if (g.key == key) entry = &g;
      if (!entry) {
        if (graphs.size() >= 64) drop_graphs();                // input buffers keep changing: start over
        // record: the host walks the layers exactly as in an eager iteration, the launches go into the graph
        const long l0 = ctx().launches;
        cudaGraph_t graph = nullptr;
        SN_CUDA(cudaStreamBeginCapture(cs, cudaStreamCaptureModeRelaxed));
        net->loss_log = loss_hist + kLossHist;
        net->prepack_filters();
        net->forward_from_to(D, L - 1);
        net->backward_from_to(L - 1, 0);
        side_join();                                           // the backward-filter stream rejoins before the update (and the end of the capture)
        update_rest(cs);
        cudaError_t e = cudaStreamEndCapture(cs, &graph);
        GraphEntry g;
        if (e == cudaSuccess && graph) {
          programmatic_edges = make_edges_programmatic(graph);
          e = cudaGraphInstantiate(&g.exec, graph, 0);
          cudaGraphDestroy(graph);
        }
        if (e != cudaSuccess || !g.exec) {
          cudaGetLastError();
          fprintf(stderr, "sparknet: recording the iteration into a CUDA graph failed (%s); continuing without graphs\n", cudaGetErrorString(e));
          use_graph = false; graph_mode = 0; replay = false;
        } else {
          g.key = key;
          g.launches = ctx().launches - l0;
          ctx().launches = l0;                                 // nothing ran yet; the replay below counts them
          graphs.push_back(std::move(g));
          entry = &graphs.back();
          for (DropoutLayer* d : drops) d->offset -= (d->n + 3) / 4;   // the recording pass advanced them; the replay does it below
        }
      }
    }
    if (replay && entry) {
      SN_CUDA(cudaGraphLaunch(entry->exec, cs));
      ctx().launches += entry->launches;
      ++graph_replays;
      for (DropoutLayer* d : drops) d->advance();
      SN_CUDA(cudaMemcpyAsync(loss_hist_host + (iter % kLossHist), loss_hist + kLossHist, sizeof(float), cudaMemcpyDeviceToHost, cs));
    } else {
      net->loss_log = loss_hist + (iter % kLossHist);
      net->prepack_filters();
      net->forward_from_to(D, L - 1);
      net->backward_from_to(L - 1, 0);
      side_join();
      if (param.display && iter % param.display == 0) {
        const float loss = net->total_loss();
        fprintf(stderr, "Iteration %d, loss = %g, lr = %g\n", iter, loss, rate);
      }
      if (net->arena_floats) update_rest(cs);
      // this iteration's loss goes to the pinned host ring too: 4 bytes, asynchronous, read after the caller's next sync
      SN_CUDA(cudaMemcpyAsync(loss_hist_host + (iter % kLossHist), loss_hist + (iter % kLossHist), sizeof(float), cudaMemcpyDeviceToHost, cs));
      ++eager_iters;
    }
    for (Blob* b : net->learnable_params) ++b->data->gen;    // weights changed on the device
    net->mark_shadows_current();                              // ... and sgd_update rounded them into the shadow arena
    ++iter;
  }
  for (DropoutLayer* d : drops) d->offset_slot = nullptr;
  for (JavaDataLayer* d : feeds) d->materialize();              // byte input: the fp32 blob of the last minibatch, once per call
  if (prefetch) {
    helper.join();
    for (JavaDataLayer* d : feeds) d->pf = nullptr;
  }
  net->overwrite_param_diffs = false;
  net->loss_log = nullptr;
  g_in_step = false;
  g_side_allowed = false;
  net->on_param_grads_ready = nullptr;
  touch_at_exit = g_outside_touch;
}
// One training iteration with CUDA events around every layer's forward and backward and around the
// update (the per-layer timing of `caffe time`, caffe/tools/caffe.cpp:290-375, on the device).
void Solver::step_profiled(std::vector<float>* fwd_ms, std::vector<float>* bwd_ms, float* update_ms) {
  const int L = (int)net->layers.size();
  std::vector<cudaEvent_t> ev((size_t)(2 * L + 3));
  for (auto& e : ev) SN_CUDA(cudaEventCreate(&e));
  cudaStream_t s = ctx().compute;
  // (a blob whose device copy is already the newest keeps its generation: its tf32 shadow / packed copies stay valid)
  net->params_to_device();
  for (Blob* b : net->learnable_params) {
    if (b->data->head != Storage::AT_DEVICE) b->data->mutable_gpu();
    b->diff->mutable_gpu();
  }
  net->overwrite_param_diffs = true;
  net->loss_log = loss_hist + (iter % kLossHist);
  profile_launches_fwd.assign((size_t)L, 0);
  profile_launches_bwd.assign((size_t)L, 0);
  for (int i = 0; i < L; ++i) {
    SN_CUDA(cudaEventRecord(ev[(size_t)i], s));
    const long l0 = ctx().launches;
    net->layers[(size_t)i]->forward(net->bottom_vecs[(size_t)i], net->top_vecs[(size_t)i]);
    profile_launches_fwd[(size_t)i] = (int)(ctx().launches - l0);
  }
  for (int i = L - 1; i >= 0; --i) {
    SN_CUDA(cudaEventRecord(ev[(size_t)(L + (L - 1 - i))], s));
    const long l0 = ctx().launches;
    if (net->layer_need_backward[(size_t)i])
      net->layers[(size_t)i]->backward(net->top_vecs[(size_t)i], net->bottom_need_backward[(size_t)i], net->bottom_vecs[(size_t)i]);
    profile_launches_bwd[(size_t)i] = (int)(ctx().launches - l0);
  }
  SN_CUDA(cudaEventRecord(ev[(size_t)(2 * L)], s));
  const float rate = get_learning_rate();
  if (net->arena_floats) sgd_update(s, net->arena_floats, net->arena_w, net->arena_g, arena_h, net->arena_s, chunk_table, rate, param.momentum);
  SN_CUDA(cudaEventRecord(ev[(size_t)(2 * L + 1)], s));
  SN_CUDA(cudaMemcpyAsync(loss_hist_host + (iter % kLossHist), loss_hist + (iter % kLossHist), sizeof(float), cudaMemcpyDeviceToHost, s));
  SN_CUDA(cudaStreamSynchronize(s));
  for (Blob* b : net->learnable_params) ++b->data->gen;
  net->mark_shadows_current();
  ++iter;
  net->overwrite_param_diffs = false;
  net->loss_log = nullptr;
  fwd_ms->assign((size_t)L, 0.f);
  bwd_ms->assign((size_t)L, 0.f);
  for (int i = 0; i < L; ++i) SN_CUDA(cudaEventElapsedTime(&(*fwd_ms)[(size_t)i], ev[(size_t)i], ev[(size_t)i + 1]));
  for (int i = L - 1; i >= 0; --i) {
    const size_t k = (size_t)(L + (L - 1 - i));
    SN_CUDA(cudaEventElapsedTime(&(*bwd_ms)[(size_t)i], ev[k], ev[k + 1]));
  }
  SN_CUDA(cudaEventElapsedTime(update_ms, ev[(size_t)(2 * L)], ev[(size_t)(2 * L + 1)]));
  for (auto& e : ev) cudaEventDestroy(e);
}
void Solver::test_and_store(int num_steps, std::vector<float>* score) {   // solver.cpp:413-444
  SN_CHECK(test_net != nullptr, "solver_test: the solver has no test net");
  net->params_to_device();
  test_net->precision = net->precision;
  test_net->share_trained_layers_with(*net);
  const int L = (int)test_net->layers.size();
  size_t total = 0;
  for (Blob* r : test_net->output_blobs) total += r->count();
  std::vector<float> sums(total, 0.f), tmp;
  for (int i = 0; i < num_steps; ++i) {
    test_net->forward_from_to(0, L - 1);
    size_t idx = 0;
    for (Blob* r : test_net->output_blobs) {
      tmp.resize(r->count());
      SN_CUDA(cudaMemcpyAsync(tmp.data(), r->data->gpu(), tmp.size() * sizeof(float), cudaMemcpyDeviceToHost, S()));
      SN_CUDA(cudaStreamSynchronize(S()));
      for (float v : tmp) sums[idx++] += v;
    }
  }
  score->assign(sums.begin(), sums.end());
}

}  // namespace sn
#include "caffe_text.hpp"
#include "ccaffe_abi.inc"
