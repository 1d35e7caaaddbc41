// caffe_proto.hpp — dependency-free decoder/encoder for the protobuf messages that cross the
// libccaffe boundary (SolverParameter / NetParameter / LayerParameter and their on-path
// sub-messages, BlobProto, SolverState).
//
// The reference links libprotobuf + protoc-generated caffe.pb.{h,cc}
// (libccaffe/CMakeLists.txt:7-13); neither exists in this image, so the wire format
// (proto2) is decoded by hand.  Field numbers and defaults follow
// caffe/src/caffe/proto/caffe.proto (line numbers given per message).  Unknown fields are
// skipped, exactly like protobuf does for fields it has no descriptor for.
//
// Host-only, header-only.  Shared by the product (libccaffe.so) and by the CPU oracle
// (oracle/), which otherwise share no code.
#pragma once
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace caffe_proto {

// ------------------------------------------------------------------------------------------
// wire reader / writer
// ------------------------------------------------------------------------------------------
struct Reader {
  const uint8_t* p;
  const uint8_t* end;
  Reader(const void* data, size_t n) : p((const uint8_t*)data), end((const uint8_t*)data + n) {}
  bool done() const { return p >= end; }
  uint64_t varint() {
    uint64_t v = 0;
    int shift = 0;
    while (true) {
      if (p >= end) throw std::runtime_error("protobuf: truncated varint");
      uint8_t c = *p++;
      v |= (uint64_t)(c & 0x7F) << shift;
      if (!(c & 0x80)) return v;
      shift += 7;
      if (shift > 63) throw std::runtime_error("protobuf: varint too long");
    }
  }
  uint32_t fixed32() {
    if (end - p < 4) throw std::runtime_error("protobuf: truncated fixed32");
    uint32_t v;
    memcpy(&v, p, 4);
    p += 4;
    return v;
  }
  uint64_t fixed64() {
    if (end - p < 8) throw std::runtime_error("protobuf: truncated fixed64");
    uint64_t v;
    memcpy(&v, p, 8);
    p += 8;
    return v;
  }
  Reader sub() {
    uint64_t n = varint();
    if ((uint64_t)(end - p) < n) throw std::runtime_error("protobuf: truncated length-delimited field");
    Reader r(p, (size_t)n);
    p += n;
    return r;
  }
  std::string str() {
    Reader r = sub();
    return std::string((const char*)r.p, (size_t)(r.end - r.p));
  }
  void skip(int wire_type) {
    switch (wire_type) {
      case 0: varint(); break;
      case 1: fixed64(); break;
      case 2: sub(); break;
      case 5: fixed32(); break;
      default: throw std::runtime_error("protobuf: unsupported wire type");
    }
  }
};

inline float as_float(uint32_t bits) {
  float f;
  memcpy(&f, &bits, 4);
  return f;
}

// Reads one scalar (or a packed run) of a repeated numeric field.
template <class T, class F>
inline void read_repeated(Reader& r, int wt, std::vector<T>& out, F one) {
  if (wt == 2) {
    Reader s = r.sub();
    while (!s.done()) out.push_back(one(s));
  } else {
    out.push_back(one(r));
  }
}
inline float rd_float(Reader& r) { return as_float(r.fixed32()); }
inline int32_t rd_i32(Reader& r) { return (int32_t)r.varint(); }
inline int64_t rd_i64(Reader& r) { return (int64_t)r.varint(); }
inline uint32_t rd_u32(Reader& r) { return (uint32_t)r.varint(); }
inline bool rd_bool(Reader& r) { return r.varint() != 0; }

struct Writer {
  std::string out;
  void varint(uint64_t v) {
    while (v >= 0x80) {
      out.push_back((char)((v & 0x7F) | 0x80));
      v >>= 7;
    }
    out.push_back((char)v);
  }
  void tag(int field, int wt) { varint(((uint64_t)field << 3) | (uint64_t)wt); }
  void i32(int field, int64_t v) { tag(field, 0); varint((uint64_t)v); }
  void f32(int field, float f) {
    tag(field, 5);
    char b[4];
    memcpy(b, &f, 4);
    out.append(b, 4);
  }
  void bytes(int field, const std::string& s) { tag(field, 2); varint(s.size()); out += s; }
  void packed_floats(int field, const float* v, size_t n) {
    tag(field, 2);
    varint(n * 4);
    out.append((const char*)v, n * 4);
  }
  void packed_i64(int field, const std::vector<int64_t>& v) {
    Writer w;
    for (int64_t x : v) w.varint((uint64_t)x);
    bytes(field, w.out);
  }
};

// ------------------------------------------------------------------------------------------
// messages (only the fields the hot path and its callers read; defaults from caffe.proto)
// ------------------------------------------------------------------------------------------
enum Phase { TRAIN = 0, TEST = 1 };

struct BlobShape {                       // caffe.proto:5-8
  std::vector<int64_t> dim;
  void parse(Reader r) {
    while (!r.done()) {
      uint64_t k = r.varint();
      int f = (int)(k >> 3), wt = (int)(k & 7);
      if (f == 1) read_repeated(r, wt, dim, rd_i64);
      else r.skip(wt);
    }
  }
};

struct BlobProto {                       // caffe.proto:10-22
  bool has_shape = false;
  BlobShape shape;
  std::vector<float> data, diff;
  int num = 0, channels = 0, height = 0, width = 0;
  bool has_legacy = false;
  void parse(Reader r) {
    while (!r.done()) {
      uint64_t k = r.varint();
      int f = (int)(k >> 3), wt = (int)(k & 7);
      switch (f) {
        case 1: num = rd_i32(r); has_legacy = true; break;
        case 2: channels = rd_i32(r); has_legacy = true; break;
        case 3: height = rd_i32(r); has_legacy = true; break;
        case 4: width = rd_i32(r); has_legacy = true; break;
        case 5:
          if (wt == 2) {  // packed: bulk copy
            Reader s = r.sub();
            size_t n = (size_t)(s.end - s.p) / 4, old = data.size();
            data.resize(old + n);
            memcpy(data.data() + old, s.p, n * 4);
          } else data.push_back(rd_float(r));
          break;
        case 6: read_repeated(r, wt, diff, rd_float); break;
        case 7: shape.parse(r.sub()); has_shape = true; break;
        default: r.skip(wt);
      }
    }
  }
};

struct FillerParameter {                 // caffe.proto:43-62
  std::string type = "constant";
  float value = 0.f, min = 0.f, max = 1.f, mean = 0.f, std = 1.f;
  int sparse = -1;
  int variance_norm = 0;                 // FAN_IN
  void parse(Reader r) {
    while (!r.done()) {
      uint64_t k = r.varint();
      int f = (int)(k >> 3), wt = (int)(k & 7);
      switch (f) {
        case 1: type = r.str(); break;
        case 2: value = rd_float(r); break;
        case 3: min = rd_float(r); break;
        case 4: max = rd_float(r); break;
        case 5: mean = rd_float(r); break;
        case 6: std = rd_float(r); break;
        case 7: sparse = rd_i32(r); break;
        case 8: variance_norm = rd_i32(r); break;
        default: r.skip(wt);
      }
    }
  }
};

struct NetState {                        // caffe.proto:259-266
  int phase = TEST;
  int level = 0;
  std::vector<std::string> stage;
  void parse(Reader r) {
    while (!r.done()) {
      uint64_t k = r.varint();
      int f = (int)(k >> 3), wt = (int)(k & 7);
      switch (f) {
        case 1: phase = rd_i32(r); break;
        case 2: level = rd_i32(r); break;
        case 3: stage.push_back(r.str()); break;
        default: r.skip(wt);
      }
    }
  }
};

struct NetStateRule {                    // caffe.proto:268-288
  bool has_phase = false, has_min_level = false, has_max_level = false;
  int phase = 0, min_level = 0, max_level = 0;
  std::vector<std::string> stage, not_stage;
  void parse(Reader r) {
    while (!r.done()) {
      uint64_t k = r.varint();
      int f = (int)(k >> 3), wt = (int)(k & 7);
      switch (f) {
        case 1: phase = rd_i32(r); has_phase = true; break;
        case 2: min_level = rd_i32(r); has_min_level = true; break;
        case 3: max_level = rd_i32(r); has_max_level = true; break;
        case 4: stage.push_back(r.str()); break;
        case 5: not_stage.push_back(r.str()); break;
        default: r.skip(wt);
      }
    }
  }
};

struct ParamSpec {                       // caffe.proto:292-315
  std::string name;
  float lr_mult = 1.f, decay_mult = 1.f;
  void parse(Reader r) {
    while (!r.done()) {
      uint64_t k = r.varint();
      int f = (int)(k >> 3), wt = (int)(k & 7);
      switch (f) {
        case 1: name = r.str(); break;
        case 3: lr_mult = rd_float(r); break;
        case 4: decay_mult = rd_float(r); break;
        default: r.skip(wt);
      }
    }
  }
};

struct ConvolutionParameter {            // caffe.proto:495-547
  uint32_t num_output = 0;
  bool bias_term = true;
  std::vector<uint32_t> pad, kernel_size, stride;
  uint32_t group = 1;
  FillerParameter weight_filler, bias_filler;
  bool has_pad_h = false, has_pad_w = false, has_kernel_h = false, has_kernel_w = false, has_stride_h = false,
       has_stride_w = false;
  uint32_t pad_h = 0, pad_w = 0, kernel_h = 0, kernel_w = 0, stride_h = 0, stride_w = 0;
  int axis = 1;
  void parse(Reader r) {
    while (!r.done()) {
      uint64_t k = r.varint();
      int f = (int)(k >> 3), wt = (int)(k & 7);
      switch (f) {
        case 1: num_output = rd_u32(r); break;
        case 2: bias_term = rd_bool(r); break;
        case 3: read_repeated(r, wt, pad, rd_u32); break;
        case 4: read_repeated(r, wt, kernel_size, rd_u32); break;
        case 5: group = rd_u32(r); break;
        case 6: read_repeated(r, wt, stride, rd_u32); break;
        case 7: weight_filler.parse(r.sub()); break;
        case 8: bias_filler.parse(r.sub()); break;
        case 9: pad_h = rd_u32(r); has_pad_h = true; break;
        case 10: pad_w = rd_u32(r); has_pad_w = true; break;
        case 11: kernel_h = rd_u32(r); has_kernel_h = true; break;
        case 12: kernel_w = rd_u32(r); has_kernel_w = true; break;
        case 13: stride_h = rd_u32(r); has_stride_h = true; break;
        case 14: stride_w = rd_u32(r); has_stride_w = true; break;
        case 16: axis = rd_i32(r); break;
        default: r.skip(wt);
      }
    }
  }
};

struct InnerProductParameter {           // caffe.proto:715-727
  uint32_t num_output = 0;
  bool bias_term = true;
  FillerParameter weight_filler, bias_filler;
  int axis = 1;
  void parse(Reader r) {
    while (!r.done()) {
      uint64_t k = r.varint();
      int f = (int)(k >> 3), wt = (int)(k & 7);
      switch (f) {
        case 1: num_output = rd_u32(r); break;
        case 2: bias_term = rd_bool(r); break;
        case 3: weight_filler.parse(r.sub()); break;
        case 4: bias_filler.parse(r.sub()); break;
        case 5: axis = rd_i32(r); break;
        default: r.skip(wt);
      }
    }
  }
};

struct PoolingParameter {                // caffe.proto:774-802
  int pool = 0;                          // MAX=0 AVE=1 STOCHASTIC=2
  uint32_t pad = 0, pad_h = 0, pad_w = 0;
  bool has_kernel_size = false, has_kernel_h = false, has_kernel_w = false;
  uint32_t kernel_size = 0, kernel_h = 0, kernel_w = 0;
  uint32_t stride = 1;
  bool has_stride_h = false, has_stride_w = false;
  uint32_t stride_h = 0, stride_w = 0;
  bool has_pad_h = false, has_pad_w = false;
  bool global_pooling = false;
  void parse(Reader r) {
    while (!r.done()) {
      uint64_t k = r.varint();
      int f = (int)(k >> 3), wt = (int)(k & 7);
      switch (f) {
        case 1: pool = rd_i32(r); break;
        case 2: kernel_size = rd_u32(r); has_kernel_size = true; break;
        case 3: stride = rd_u32(r); break;
        case 4: pad = rd_u32(r); break;
        case 5: kernel_h = rd_u32(r); has_kernel_h = true; break;
        case 6: kernel_w = rd_u32(r); has_kernel_w = true; break;
        case 7: stride_h = rd_u32(r); has_stride_h = true; break;
        case 8: stride_w = rd_u32(r); has_stride_w = true; break;
        case 9: pad_h = rd_u32(r); has_pad_h = true; break;
        case 10: pad_w = rd_u32(r); has_pad_w = true; break;
        case 12: global_pooling = rd_bool(r); break;
        default: r.skip(wt);
      }
    }
  }
};

struct LRNParameter {                    // caffe.proto:738-757
  uint32_t local_size = 5;
  float alpha = 1.f, beta = 0.75f, k = 1.f;
  int norm_region = 0;                   // ACROSS_CHANNELS=0 WITHIN_CHANNEL=1
  void parse(Reader r) {
    while (!r.done()) {
      uint64_t kk = r.varint();
      int f = (int)(kk >> 3), wt = (int)(kk & 7);
      switch (f) {
        case 1: local_size = rd_u32(r); break;
        case 2: alpha = rd_float(r); break;
        case 3: beta = rd_float(r); break;
        case 4: norm_region = rd_i32(r); break;
        case 5: k = rd_float(r); break;
        default: r.skip(wt);
      }
    }
  }
};

struct LossParameter {                   // caffe.proto:420-430
  bool has_ignore_label = false;
  int ignore_label = 0;
  bool normalize = true;
  void parse(Reader r) {
    while (!r.done()) {
      uint64_t k = r.varint();
      int f = (int)(k >> 3), wt = (int)(k & 7);
      switch (f) {
        case 1: ignore_label = rd_i32(r); has_ignore_label = true; break;
        case 2: normalize = rd_bool(r); break;
        default: r.skip(wt);
      }
    }
  }
};

struct AccuracyParameter {               // caffe.proto:435-452
  uint32_t top_k = 1;
  int axis = 1;
  bool has_ignore_label = false;
  int ignore_label = 0;
  void parse(Reader r) {
    while (!r.done()) {
      uint64_t k = r.varint();
      int f = (int)(k >> 3), wt = (int)(k & 7);
      switch (f) {
        case 1: top_k = rd_u32(r); break;
        case 2: axis = rd_i32(r); break;
        case 3: ignore_label = rd_i32(r); has_ignore_label = true; break;
        default: r.skip(wt);
      }
    }
  }
};

struct PowerParameter {                  // caffe.proto:805-810
  float power = 1.f, scale = 1.f, shift = 0.f;
  void parse(Reader r) {
    while (!r.done()) {
      uint64_t k = r.varint();
      int f = (int)(k >> 3), wt = (int)(k & 7);
      switch (f) {
        case 1: power = rd_float(r); break;
        case 2: scale = rd_float(r); break;
        case 3: shift = rd_float(r); break;
        default: r.skip(wt);
      }
    }
  }
};

struct EltwiseParameter {                // caffe.proto:601-613
  int operation = 1;                     // PROD=0 SUM=1 MAX=2
  std::vector<float> coeff;
  bool stable_prod_grad = true;
  void parse(Reader r) {
    while (!r.done()) {
      uint64_t k = r.varint();
      int f = (int)(k >> 3), wt = (int)(k & 7);
      switch (f) {
        case 1: operation = rd_i32(r); break;
        case 2: read_repeated(r, wt, coeff, rd_float); break;
        case 3: stable_prod_grad = rd_bool(r); break;
        default: r.skip(wt);
      }
    }
  }
};

struct LayerParameter {                  // caffe.proto:310-400
  std::string name, type;
  std::vector<std::string> bottom, top;
  bool has_phase = false;
  int phase = TRAIN;
  std::vector<float> loss_weight;
  std::vector<ParamSpec> param;
  std::vector<BlobProto> blobs;
  std::vector<bool> propagate_down;
  std::vector<NetStateRule> include, exclude;
  LossParameter loss_param;
  AccuracyParameter accuracy_param;
  ConvolutionParameter convolution_param;
  float dropout_ratio = 0.5f;            // DropoutParameter, caffe.proto:578-580
  EltwiseParameter eltwise_param;
  InnerProductParameter inner_product_param;
  LRNParameter lrn_param;
  PoolingParameter pooling_param;
  PowerParameter power_param;
  float relu_negative_slope = 0.f;       // ReLUParameter, caffe.proto:818-831
  int softmax_axis = 1;                  // SoftmaxParameter, caffe.proto:849-861
  bool has_java_data_param = false;
  BlobShape java_data_shape;             // JavaDataParameter (SparkNet), caffe.proto:991-993
  std::string raw;                       // the serialized message as received (Layer::ToProto copies layer_param_ verbatim)
  void parse(Reader r) {
    raw.assign((const char*)r.p, (size_t)(r.end - r.p));
    while (!r.done()) {
      uint64_t k = r.varint();
      int f = (int)(k >> 3), wt = (int)(k & 7);
      switch (f) {
        case 1: name = r.str(); break;
        case 2: type = r.str(); break;
        case 3: bottom.push_back(r.str()); break;
        case 4: top.push_back(r.str()); break;
        case 5: read_repeated(r, wt, loss_weight, rd_float); break;
        case 6: { ParamSpec p; p.parse(r.sub()); param.push_back(p); break; }
        case 7: { blobs.emplace_back(); blobs.back().parse(r.sub()); break; }
        case 8: { NetStateRule s; s.parse(r.sub()); include.push_back(s); break; }
        case 9: { NetStateRule s; s.parse(r.sub()); exclude.push_back(s); break; }
        case 10: phase = rd_i32(r); has_phase = true; break;
        case 11: read_repeated(r, wt, propagate_down, rd_bool); break;
        case 101: loss_param.parse(r.sub()); break;
        case 102: accuracy_param.parse(r.sub()); break;
        case 106: convolution_param.parse(r.sub()); break;
        case 108: {
          Reader s = r.sub();
          while (!s.done()) {
            uint64_t kk = s.varint();
            if ((kk >> 3) == 1) dropout_ratio = rd_float(s); else s.skip((int)(kk & 7));
          }
          break;
        }
        case 110: eltwise_param.parse(r.sub()); break;
        case 117: inner_product_param.parse(r.sub()); break;
        case 118: lrn_param.parse(r.sub()); break;
        case 121: pooling_param.parse(r.sub()); break;
        case 122: power_param.parse(r.sub()); break;
        case 123: {
          Reader s = r.sub();
          while (!s.done()) {
            uint64_t kk = s.varint();
            if ((kk >> 3) == 1) relu_negative_slope = rd_float(s); else s.skip((int)(kk & 7));
          }
          break;
        }
        case 125: {
          Reader s = r.sub();
          while (!s.done()) {
            uint64_t kk = s.varint();
            if ((kk >> 3) == 2) softmax_axis = rd_i32(s); else s.skip((int)(kk & 7));
          }
          break;
        }
        case 149: {
          Reader s = r.sub();
          has_java_data_param = true;
          while (!s.done()) {
            uint64_t kk = s.varint();
            if ((kk >> 3) == 1) java_data_shape.parse(s.sub()); else s.skip((int)(kk & 7));
          }
          break;
        }
        default: r.skip(wt);
      }
    }
  }
};

struct NetParameter {                    // caffe.proto:64-100
  std::string name;
  std::vector<std::string> input;
  bool force_backward = false;
  NetState state;
  bool has_state = false;
  bool debug_info = false;
  std::vector<LayerParameter> layer;
  void parse(Reader r) {
    while (!r.done()) {
      uint64_t k = r.varint();
      int f = (int)(k >> 3), wt = (int)(k & 7);
      switch (f) {
        case 1: name = r.str(); break;
        case 3: input.push_back(r.str()); break;
        case 5: force_backward = rd_bool(r); break;
        case 6: state.parse(r.sub()); has_state = true; break;
        case 7: debug_info = rd_bool(r); break;
        case 100: { layer.emplace_back(); layer.back().parse(r.sub()); break; }
        case 2: {   // V1LayerParameter (caffe.proto:1063-1160), as in the published bvlc caffemodels: only name (4) and blobs (6) matter here
          Reader s = r.sub();
          LayerParameter l;
          while (!s.done()) {
            uint64_t kk = s.varint();
            const int ff = (int)(kk >> 3), ww = (int)(kk & 7);
            if (ff == 4) l.name = s.str();
            else if (ff == 6) { l.blobs.emplace_back(); l.blobs.back().parse(s.sub()); }
            else s.skip(ww);
          }
          v1_layers.push_back(std::move(l));
          break;
        }
        default: r.skip(wt);
      }
    }
  }
  std::vector<LayerParameter> v1_layers;   // legacy `layers` entries (name + blobs only)
};

struct SolverParameter {                 // caffe.proto:102-243
  bool has_net_param = false, has_train_net_param = false;
  NetParameter net_param, train_net_param;
  std::vector<NetParameter> test_net_param;
  std::string net, train_net;
  std::vector<std::string> test_net;
  std::vector<int> test_iter;
  int test_interval = 0;
  float base_lr = 0.f;
  int display = 0, max_iter = 0;
  std::string lr_policy;
  float gamma = 0.f, power = 0.f, momentum = 0.f, weight_decay = 0.f;
  int stepsize = 0;
  std::vector<int> stepvalue;
  int snapshot = 0;
  std::string snapshot_prefix;
  bool snapshot_diff = false;
  int solver_mode = 1, device_id = 0;
  int64_t random_seed = -1;
  std::string regularization_type = "L2";
  int average_loss = 1;
  float clip_gradients = -1.f;
  int iter_size = 1;
  std::string type = "SGD";
  bool debug_info = false;
  NetState train_state;
  std::vector<NetState> test_state;
  void parse(Reader r) {
    train_state.phase = TRAIN;
    while (!r.done()) {
      uint64_t k = r.varint();
      int f = (int)(k >> 3), wt = (int)(k & 7);
      switch (f) {
        case 1: train_net = r.str(); break;
        case 2: test_net.push_back(r.str()); break;
        case 3: read_repeated(r, wt, test_iter, rd_i32); break;
        case 4: test_interval = rd_i32(r); break;
        case 5: base_lr = rd_float(r); break;
        case 6: display = rd_i32(r); break;
        case 7: max_iter = rd_i32(r); break;
        case 8: lr_policy = r.str(); break;
        case 9: gamma = rd_float(r); break;
        case 10: power = rd_float(r); break;
        case 11: momentum = rd_float(r); break;
        case 12: weight_decay = rd_float(r); break;
        case 13: stepsize = rd_i32(r); break;
        case 14: snapshot = rd_i32(r); break;
        case 15: snapshot_prefix = r.str(); break;
        case 16: snapshot_diff = rd_bool(r); break;
        case 17: solver_mode = rd_i32(r); break;
        case 18: device_id = rd_i32(r); break;
        case 20: random_seed = rd_i64(r); break;
        case 21: train_net_param.parse(r.sub()); has_train_net_param = true; break;
        case 22: { test_net_param.emplace_back(); test_net_param.back().parse(r.sub()); break; }
        case 23: debug_info = rd_bool(r); break;
        case 24: net = r.str(); break;
        case 25: net_param.parse(r.sub()); has_net_param = true; break;
        case 26: train_state.parse(r.sub()); break;
        case 27: { test_state.emplace_back(); test_state.back().parse(r.sub()); break; }
        case 29: regularization_type = r.str(); break;
        case 33: average_loss = rd_i32(r); break;
        case 34: read_repeated(r, wt, stepvalue, rd_i32); break;
        case 35: clip_gradients = rd_float(r); break;
        case 36: iter_size = rd_i32(r); break;
        case 40: type = r.str(); break;
        default: r.skip(wt);
      }
    }
  }
};

struct SolverState {                     // caffe.proto:245-255
  int iter = 0, current_step = 0;
  std::string learned_net;
  std::vector<BlobProto> history;
  void parse(Reader r) {
    while (!r.done()) {
      uint64_t k = r.varint();
      int f = (int)(k >> 3), wt = (int)(k & 7);
      switch (f) {
        case 1: iter = rd_i32(r); break;
        case 2: learned_net = r.str(); break;
        case 3: { history.emplace_back(); history.back().parse(r.sub()); break; }
        case 4: current_step = rd_i32(r); break;
        default: r.skip(wt);
      }
    }
  }
};

// ------------------------------------------------------------------------------------------
// Net filtering by phase (reference: caffe/src/caffe/net.cpp:287-368 FilterNet/StateMeetsRule)
// ------------------------------------------------------------------------------------------
inline bool state_meets_rule(const NetState& state, const NetStateRule& rule) {
  if (rule.has_phase && rule.phase != state.phase) return false;
  if (rule.has_min_level && state.level < rule.min_level) return false;
  if (rule.has_max_level && state.level > rule.max_level) return false;
  for (const std::string& s : rule.stage) {
    bool has = false;
    for (const std::string& t : state.stage) has |= (s == t);
    if (!has) return false;
  }
  for (const std::string& s : rule.not_stage) {
    for (const std::string& t : state.stage)
      if (s == t) return false;
  }
  return true;
}

inline std::vector<LayerParameter> filter_net(const NetParameter& param, const NetState& state) {
  std::vector<LayerParameter> out;
  for (const LayerParameter& lp : param.layer) {
    if (!lp.include.empty() && !lp.exclude.empty())
      throw std::runtime_error("layer " + lp.name + ": specify either include rules or exclude rules; not both");
    bool included = lp.include.empty();
    for (size_t j = 0; included && j < lp.exclude.size(); ++j)
      if (state_meets_rule(state, lp.exclude[j])) included = false;
    for (size_t j = 0; !included && j < lp.include.size(); ++j)
      if (state_meets_rule(state, lp.include[j])) included = true;
    if (included) out.push_back(lp);
  }
  return out;
}

}  // namespace caffe_proto
