/* ccaffe.h — C ABI of the B200-native libccaffe.so.
 *
 * The first block is the reference's own boundary, symbol for symbol: libccaffe/ccaffe.h:5-77
 * (implemented by the reference in libccaffe/ccaffe.cpp:22-304 and bound from the JVM by
 * src/main/java/libs/CaffeLibrary.java:8-76 through JNA).  Names, argument order and types are
 * unchanged so Net.scala / ProtoLoader.scala keep working without modification.
 *
 * The second block ("extensions") is an ABI superset used by the in-box multi-GPU driver, the
 * tests and bench.py.  Nothing in the first block depends on it.
 *
 * Conventions kept from the reference: opaque state handle owned by the library; blob handles are
 * valid for the life of the state; get_data/get_diff return a host pointer that shows the current
 * values and becomes the authoritative copy (the caller may write through it; the next device use
 * re-uploads); errors abort the process after printing to stderr (the reference's glog CHECK).
 * There is no CPU compute path: set_mode_cpu() is accepted and ignored.
 */
#ifndef SPARKNET_B200_CCAFFE_H_
#define SPARKNET_B200_CCAFFE_H_

#include <stdbool.h>

#define DTYPE float

#ifdef __cplusplus
extern "C" {
#endif

typedef void (*error_callback_t)(const char* msg);                                /* ccaffe.h:6 */
/* caffe/include/caffe/data_layers.hpp:23-25; num_dims/shape exclude the batch dimension */
typedef void (*java_callback_t)(void* data, int batch_size, int num_dims, int* shape);

struct caffenet_state;                                                            /* ccaffe.h:10 */
typedef struct caffenet_state caffenet_state;

/* ---- reference ABI ------------------------------------------------------------------------ */
void init_logging(const char* log_filename, int log_verbosity);                   /* ccaffe.h:13, ccaffe.cpp:33-37 */
void set_basepath(const char* path);                                              /* :15, cpp:39-41 */
int get_int_size(void);                                                           /* :17, cpp:43-45 */
int get_dtype_size(void);                                                         /* :18, cpp:47-49 */

/* LevelDB/LMDB writers: out of the hot path; exported, abort with a message */
void create_db(caffenet_state* state, char* db_name, int name_len, char* db_type, int db_type_len);      /* :20 */
void write_to_db(caffenet_state* state, char* image, int label, int channels, int height, int width, char* key_str); /* :21 */
void commit_db_txn(caffenet_state* state);                                        /* :22 */
void close_db(caffenet_state* state);                                             /* :23 */
void save_mean_image(caffenet_state* state, float* mean_image, int channels, int height, int width,
                     char* filename, int filename_len);                           /* :25, cpp:83-97 */

caffenet_state* create_state(void);                                               /* :27, cpp:99-110 */
void destroy_state(caffenet_state* state);                                        /* :28, cpp:112-126 */

void load_solver_from_protobuf(caffenet_state* state, const char* solver_param, int solver_param_len); /* :30, cpp:128-134 */
void load_net_from_protobuf(caffenet_state* state, const char* net_param, int net_param_len);          /* :31, cpp:136-140 */

/* 0 = ok, -1 = layer_idx invalid, -2 = layer is not a JavaData layer */
int set_train_data_callback(caffenet_state* state, int layer_idx, java_callback_t callback);           /* :35, cpp:197-212 */
int set_test_data_callback(caffenet_state* state, int layer_idx, java_callback_t callback);            /* :36, cpp:214-216 */

void forward(caffenet_state* state);                                              /* :38, cpp:218-222 */
void backward(caffenet_state* state);                                             /* :39, cpp:224-228 */

int solver_step(caffenet_state* state, int step);                                 /* :41, cpp:230-233 */
void solver_test(caffenet_state* state, int num_steps);                           /* :42, cpp:235-238 */
int num_test_scores(caffenet_state* state);                                       /* :44, cpp:174-177 */
DTYPE get_test_score(caffenet_state* state, int test_score_layer_idx);            /* :45, cpp:240-243 */

int num_layers(caffenet_state* state);                                            /* :47, cpp:142-144 */
const char* layer_name(caffenet_state* state, int layer_idx);                     /* :48, cpp:146-148 */
int num_layer_weights(caffenet_state* state, int layer_idx);                      /* :49, cpp:150-152 */

int num_data_blobs(caffenet_state* state);                                        /* :51, cpp:166-168 */
void* get_data_blob(caffenet_state* state, int blob_idx);                         /* :52, cpp:154-156 */
const char* get_data_blob_name(caffenet_state* state, int blob_idx);              /* :53, cpp:158-160 */
void* get_weight_blob(caffenet_state* state, int layer_idx, int weight_idx);      /* :54, cpp:162-164 */

DTYPE* get_data(void* blob);                                                      /* :56, cpp:181-183 */
DTYPE* get_diff(void* blob);                                                      /* :57, cpp:185-187 */
int get_num_axes(void* blob);                                                     /* :59, cpp:189-191 */
int get_axis_shape(void* blob, int axis);                                         /* :60, cpp:193-195 */

void set_global_error_callback(error_callback_t c);                               /* :62, cpp:245-247 */

void set_mode_cpu(void);                                                          /* :64, cpp:253-255 — accepted, ignored */
void set_mode_gpu(void);                                                          /* :65, cpp:249-251 */
void set_device(int gpu_id);                                                      /* :66, cpp:257-259 */

void load_weights_from_file(caffenet_state* state, const char* filename);         /* :68, cpp:261-263 */
void save_weights_to_file(caffenet_state* state, const char* filename);           /* :69, cpp:265-269 */
void restore_solver_from_file(caffenet_state* state, const char* filename);       /* :70, cpp:271-273 */

bool parse_net_prototxt(caffenet_state* state, const char* filename);             /* :72, cpp:290-292 */
bool parse_solver_prototxt(caffenet_state* state, const char* filename);          /* :73, cpp:294-296 */
int get_prototxt_len(caffenet_state* state);                                      /* :75, cpp:298-300 */
char* get_prototxt_data(caffenet_state* state);                                   /* :76, cpp:302-304 */

/* ---- extensions (no counterpart in libccaffe/ccaffe.h) --------------------------------------- */
const char* sparknet_version(void);
/* 0 = fp32 FFMA validation kernels, 1 = TF32 tcgen05/TMEM kernels.  Returns the mode in effect. */
int sparknet_set_precision(caffenet_state* state, int mode);
int sparknet_get_precision(caffenet_state* state);
int sparknet_tc_available(void);
/* cudaStream_t all kernels of this library are launched on. */
void* sparknet_compute_stream(caffenet_state* state);
/* Device address / length in floats of the contiguous weight arena (all learnable blobs in layer
 * order, each starting on a 1024-float boundary, padding zero).  The device copy becomes authoritative. */
int sparknet_param_arena(caffenet_state* state, void** ptr, long* count);
int sparknet_get_iter(caffenet_state* state);
/* Loss of the most recent forward pass (reads one float back; synchronises the stream). */
float sparknet_last_loss(caffenet_state* state);
/* Copies the losses of the last n solver iterations (oldest first) kept on the device; returns count. */
int sparknet_get_loss_history(caffenet_state* state, float* out, int n);
/* Model + solver-state snapshot in Caffe's binaryproto formats (Solver::Snapshot, solver.cpp:447-470 /
 * sgd_solver.cpp:255-272; the reference triggers it from SolverParameter.snapshot, here it is a call): writes the
 * NetParameter with every layer's blobs to model_filename and SolverState {iter, learned_net = model_filename,
 * current_step, history} to state_filename; restore_solver_from_file reads the pair back. */
int sparknet_snapshot(caffenet_state* state, const char* model_filename, const char* state_filename);
/* memcpy on `threads` host threads, for data callbacks that fill the library's pinned staging buffer (no reference
 * counterpart: the JVM callback of Net.scala:194-221 copies element by element). */
void sparknet_host_copy(void* dst, const void* src, long bytes, int threads);
/* n bytes -> n floats on `threads` host threads (signed Java bytes or 0..255): the widening the reference's callback
 * does per image (src/main/java/libs/ByteImage.java:86-104), for fp32 callbacks whose pixels are held as bytes. */
void sparknet_host_widen(void* dst, const void* src, long n, int is_signed, int threads);
/* Make Dropout layer `layer_idx` of the train net use `mask` (count values, 0 = drop, non-zero = keep)
 * instead of drawing one; NULL restores drawing.  0 ok, -1 bad index, -2 not Dropout, -3 bad count. */
int sparknet_inject_dropout_mask(caffenet_state* state, int layer_idx, const unsigned* mask, int count);
/* Blocks until all queued device work of this process is done. */
void sparknet_sync(caffenet_state* state);
/* Learnable parameters: count, and bulk copies in (layer, blob) order, tightly packed. */
long sparknet_num_params(caffenet_state* state);
int sparknet_get_weights_flat(caffenet_state* state, float* out, long count);
int sparknet_set_weights_flat(caffenet_state* state, const float* in, long count);
/* Momentum history of learnable blob `param_idx` (testing aid). */
int sparknet_get_history(caffenet_state* state, int param_idx, float* out, long count);
/* int32 argmax mask of MAX-pooling layer `layer_idx` (testing aid). */
int sparknet_get_pool_mask(caffenet_state* state, int layer_idx, int* out, long count);
/* Pull n minibatches through the train callbacks now and keep them resident in HBM; later passes
 * cycle through them without calling back or copying.  n = 0 restores callback mode. */
int sparknet_set_resident_inputs(caffenet_state* state, int n);
/* Number of kernels launched by this process so far. */
long sparknet_kernel_launches(caffenet_state* state);
/* Byte-pixel data callback.  The reference keeps images as Java bytes and widens them to float on the host inside its
 * callback (src/main/java/libs/ByteImage.java:86-104 cropInto; src/main/scala/libs/Net.scala:194-221 makeImageCallback;
 * mean subtraction src/main/scala/apps/ImageNetApp.scala:158-175).  With this entry the callback (same type, same call
 * pattern and pairing as set_train_data_callback, libccaffe/ccaffe.h:33) fills batch*C*H*W BYTES in NCHW order and the
 * device widens them into the layer's fp32 blob: x = ((float)b - mean[..]) * scale, b read as a signed Java byte
 * (is_signed != 0) or as 0..255; mean = NULL, C*H*W values (mean image) or C values (per channel), copied by the call.
 * The blob is bit-identical to what the fp32 callback path holds for the same pixels.  Returns 0, -1 bad layer index,
 * -2 not a JavaData layer.  set_train_data_callback on the same layer switches back to fp32 buffers. */
int sparknet_set_train_data_callback_bytes(caffenet_state* state, int layer_idx, java_callback_t callback, int is_signed,
                                           const float* mean, int mean_count, float scale);
/* Host-side cost of the input path since the previous call: out[0] = callbacks run, out[1] = mean ms inside a callback,
 * out[2] = mean ms the solver thread waited for a pinned staging buffer before a callback. */
int sparknet_input_stats(caffenet_state* state, double* out3);
/* Net::ForwardFromTo / BackwardFromTo (caffe/src/caffe/net.cpp:565-581, 635-645), which libccaffe/ccaffe.h does not
 * expose: forward over layers [start, end]; backward from layer `start` down to layer `end`.  0 ok, -1 bad range. */
int sparknet_forward_range(caffenet_state* state, int start, int end);
int sparknet_backward_range(caffenet_state* state, int start, int end);
/* Grow the per-process scratch workspace (split-reduction slices, bias partials) to at least nfloats now; returns its
 * size in floats.  Growing reallocates, so iterations recorded for replay are recorded again afterwards. */
long sparknet_reserve_workspace(long nfloats);
/* Bulk copies between the weight arena and its pinned host mirror so far.  All get_weight_blob handles of a net share
 * one mirror: the first get_data after device-side updates moves the whole arena in ONE device-to-host copy (the other
 * blobs' get_data calls then copy nothing), and the next device use after host writes is ONE host-to-device copy —
 * what keeps Net.scala:132-172 getWeights / setWeights cheap for the unchanged JVM driver. */
void sparknet_param_copy_counts(caffenet_state* state, long* d2h, long* h2d);
/* NUMA placement of this process: writes "gpu_node=<n> pinned=<0|1> cpus=<count>" into buf. */
int sparknet_numa_info(char* buf, int n);
/* Test hook for the tensor-core GEMM core: D[M][N] = A x B^T, host pointers.  a_mn/b_mn: 0 = operand
 * stored [rows][R], 1 = stored [R][rows]; tma_a: feed A through TMA. */
int sparknet_test_tc_gemm(int a_mn, int b_mn, int tma_a, int M, int N, int R, const float* A, const float* B, float* D);
/* Development switches of the window convolution kernels (no reference counterpart): what = 1 forces the
 * number of accumulators per CTA (0 = the planner's choice); what = 2 turns per-CTA clock stamps on / off, which
 * sparknet_debug_read_stamps copies out (8 per CTA of the most recent window launch). */
void sparknet_debug_set(int what, int value);
int sparknet_debug_read_stamps(long long* out, int max_ctas);
/* Development aid: clocks per TF32 tcgen05.mma (128*cg x n_mma x 8) issued back to back on resident tiles. */
float sparknet_debug_mma_rate(int cg, int n_mma, int iters, int shift_rows, int nacc, int ctas, float* issue_only);
/* CUDA-event stopwatch on the library's compute stream (bench.py times the hot path with it). */
void sparknet_timer_start(caffenet_state* state);
float sparknet_timer_stop_ms(caffenet_state* state);
/* Run ONE training iteration with events around every layer (consumes one callback pair).  Fills
 * per-layer forward/backward milliseconds (n = capacity) and the update time; returns the layer count. */
int sparknet_profile_step(caffenet_state* state, float* fwd_ms, float* bwd_ms, float* update_ms, int n);
/* Kernels each layer launched (forward / backward) in the iteration sparknet_profile_step ran last; 0 = the layer's work
 * rode in another layer's kernel.  Returns the layer count. */
int sparknet_profile_launches(caffenet_state* state, int* fwd, int* bwd, int n);
/* out[i] = (((r0[i] + r1[i]) + ...) / n in IEEE fp32, replica order; all pointers are device pointers
 * on the current device (single-GPU emulation of the cross-GPU exchange; the arithmetic is the same). */
int sparknet_average_replicas_device(const void* const* replicas, int n, long count, void* out);
/* Same, enqueued on the compute stream and not synchronised (the multi-GPU driver gathers the
 * replicas with NCCL on that stream first). */
int sparknet_average_replicas_async(const void* const* replicas, int n, long count, void* out);

/* solver_step records the steady-state iteration into a CUDA graph and replays it (the first iteration of every
 * call runs eagerly).  mode 1 on, 0 off, -1 = SPARKNET_GRAPH environment variable (default on); returns the previous
 * mode.  sparknet_graph_replays = iterations that ran as replays so far. */
int sparknet_set_graph_mode(caffenet_state* state, int mode);
long sparknet_graph_replays(caffenet_state* state);
/* Kernel-to-kernel edges of the most recent recording that run as programmatic dependent launches (the next kernel's
 * prologue overlaps the previous kernel's tail); 0 with SPARKNET_PDL=0. */
int sparknet_graph_programmatic_edges(caffenet_state* state);

/* Weight averaging over peer memory — replaces the driver's collect / reduce(WeightCollection.add) /
 * scalarDivide / broadcast / setWeights round (src/main/scala/apps/CifarApp.scala:92-99,133-134;
 * src/main/scala/libs/Net.scala:14-47,132-172; src/main/java/libs/JavaNDArray.java:84-117) for the replicas
 * of one box, one process per GPU.  Result on every replica: (((w0 + w1) + ...) + w_{n-1}) / (float)n in
 * IEEE fp32, rank order — bit-exact against the reference's sequential sum.
 *   sparknet_peer_handle_bytes   size of one handle (a cudaIpcMemHandle_t, 64 bytes)
 *   sparknet_peer_export         writes this replica's two handles (weight arena, flag block)
 *   sparknet_peer_attach         takes all replicas' handles (nranks x handle_bytes each, rank order); -3 when a
 *                                peer's memory cannot be mapped (no peer access / no IPC): the caller picks another path
 *   sparknet_average_peers       one round, enqueued on the compute stream; all replicas must call it
 *   sparknet_peer_status         synchronises; 0 = all rounds completed, 1 = a peer did not arrive in time
 *   sparknet_peer_timing         device-side clock of the rounds since its last call: mean ms waiting for the slowest
 *                                replica, mean ms of the exchange proper (all ready -> all stored); returns #rounds */
int sparknet_peer_handle_bytes(void);
int sparknet_peer_export(caffenet_state* state, void* arena_handle, void* flags_handle);
int sparknet_peer_attach(caffenet_state* state, int rank, int nranks, const void* arena_handles, const void* flags_handles);
int sparknet_average_peers(caffenet_state* state);
int sparknet_peer_status(caffenet_state* state);
int sparknet_peer_timing(caffenet_state* state, float* wait_ms, float* exchange_ms);
/* Test hook: the same kernel and flag protocol with n replicas on ONE GPU (n streams, small grids).  in / out are
 * host arrays [n][count]; every replica's arena is returned after `rounds` exchanges.  0 = ok, 1 = time-out. */
int sparknet_test_average_peers(int n, long count, const float* in, float* out, int rounds);

/* No counterpart in the reference (libccaffe/ccaffe.h has nothing like it).
 * D[M][N] = A x B^T in TF32 on the tensor cores for callers that must not allocate device memory while
 * running (inside a stream capture, in latency-critical threads, under a fixed memory budget): the
 * split-reduction scratch is a static 8 MiB device array inside the library instead of the grow-on-demand
 * workspace; no cudaMalloc/cudaFree happens in this call.  a_mn/b_mn: 0 = operand stored [rows][R],
 * 1 = stored [R][rows].  A, B, D are DEVICE pointers owned by the caller; stream is a cudaStream_t (NULL =
 * default stream); the call does not synchronise.  Returns 0, or -1 bad arguments, -2 the problem needs
 * more scratch than the static array holds (refused, nothing launched), < -9 CUDA errors.  The first call
 * on a device resolves the array's device address: make it before any stream capture.  All calls share the
 * array, so calls that can overlap on the GPU race with each other. */
int sparknet_tc_gemm_static_scratch(int a_mn, int b_mn, int M, int N, int R, const float* A, const float* B,
                                    float* D, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SPARKNET_B200_CCAFFE_H_ */
