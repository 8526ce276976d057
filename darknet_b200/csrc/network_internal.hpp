// Internal definitions shared by the host runtime files (network.cu: load / plan / inference, network_train.cu:
// training step).  Not part of the ABI.
#pragma once
#include <string>
#include <utility>
#include <vector>

#include "yc_common.cuh"

enum LType { L_CONV = 0, L_MAXPOOL = 1, L_UPSAMPLE = 2, L_ROUTE = 3, L_SHORTCUT = 4, L_YOLO = 5, L_GYOLO = 6 };

inline int activation_id(const std::string& s) {  // get_activation, activations.c:60-105 (subset on the path)
  if (s == "logistic") return YC_LOGISTIC;
  if (s == "relu") return YC_RELU;
  if (s == "linear") return YC_LINEAR;
  if (s == "leaky") return YC_LEAKY;
  if (s == "swish") return YC_SWISH;
  if (s == "mish") return YC_MISH;
  return -1;
}

struct NLayer {
  int type = -1;
  int c = 0, h = 0, w = 0, out_c = 0, out_h = 0, out_w = 0;
  // conv
  int n = 0, size = 1, stride = 1, pad = 0, dilation = 1, bn = 0, act = YC_LINEAR;
  std::vector<float> weights, biases, scales, rmean, rvar;  // host fp32, reference layouts
  void* w_packed = nullptr;
  float* bias_dev = nullptr;
  float* scale_dev = nullptr;
  yc_conv_plan* plan = nullptr;
  // route / shortcut
  std::vector<int> inputs;
  int groups = 1, group_id = 0;
  int from = -1;
  // yolo
  std::vector<int> mask;
  std::vector<float> anchors;
  int classes = 0, total = 0, nms_kind = 0;
  float scale_x_y = 1.f, beta_nms = 0.6f;
  float* dec_dev = nullptr;   // decoded fp32 NCHW (the reference's l->output_gpu)
  float* dec_host = nullptr;  // pinned (the reference's l->output, yolo_layer.cpp:62-66)
  float* anchors_dev = nullptr;
  int* mask_dev = nullptr;
  // storage of the NHWC output
  bool alias = false;       // a view of another layer's storage (single-input route)
  int alias_src = -1, alias_coff = 0;
  int parent = -1, parent_coff = 0;  // placed inside `parent`'s concat buffer
  int fused_into = -1;      // conv whose epilogue also performs shortcut `fused_into`
  int fused_conv = -1;      // shortcut computed by conv `fused_conv`
  void* buf = nullptr;      // owned allocation (root buffers only)
  int own_ldc = 0;
  int dtype = YC_BF16;
  std::vector<std::pair<int, int>> copies;  // route fallback: (input layer, dst channel offset)
  int consumers = 0;
};

struct Op {
  int layer;
  int kind;  // 0 conv, 1 maxpool, 2 upsample, 3 shortcut, 4 route copy, 5 yolo decode, 6 input convert
  yc_tensor a, b, y;
  int i0 = 0, i1 = 0, i2 = 0, i3 = 0;
  float f0 = 0.f;
};

struct yc_network {
  int batch = 1, w = 0, h = 0, c = 0, flags = YC_NET_DEFAULT;
  std::vector<NLayer> layers;
  std::vector<Op> ops;
  bool finalized = false;
  cudaStream_t stream = nullptr;
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t graph_exec = nullptr;
  float* input_dev = nullptr;   // fp32 NCHW staging (net->input_state_gpu)
  float* input_pinned = nullptr;  // net->input_pinned_cpu
  void* input_nhwc = nullptr;   // bf16 NHWC, ldc 8
  const float* graph_input = nullptr;
  std::vector<int> yolo_layers;
  int n_conv = 0;
  // detection extraction workspace
  float* det_regions = nullptr;
  int* det_counts = nullptr;
  float* det_dense = nullptr;
  int* det_img_counts = nullptr;
  float* det_host = nullptr;
  int* det_img_counts_host = nullptr;
  int det_cap = 0, det_dense_cap = 0, det_R = 0;
  std::vector<float> det_rows_out;
};


#define NET_FAIL(code, ...)   \
  do {                        \
    set_error(__VA_ARGS__);   \
    return code;              \
  } while (0)


#define NET_CUDA(call)                                                                     \
  do {                                                                                     \
    cudaError_t _e = (call);                                                               \
    if (_e != cudaSuccess) {                                                               \
      set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e));     \
      return YC_ERR_CUDA;                                                                  \
    }                                                                                      \
  } while (0)


// resolve a layer to its physical storage: (root layer that owns a buffer, channel offset, pitch)
struct Loc { int root; int coff; };
inline Loc locate(const yc_network* net, int i) {
  int coff = 0;
  for (;;) {
    const NLayer& L = net->layers[i];
    if (L.alias) { coff += L.alias_coff; i = L.alias_src; continue; }
    if (L.parent >= 0) { coff += L.parent_coff; i = L.parent; continue; }
    return Loc{i, coff};
  }
}

inline yc_tensor view_of(const yc_network* net, int i) {
  const NLayer& L = net->layers[i];
  const Loc loc = locate(net, i);
  const NLayer& R = net->layers[loc.root];
  yc_tensor t;
  const int es = R.dtype == YC_F32 ? 4 : 2;
  t.ptr = (char*)R.buf + (size_t)loc.coff * es;
  t.n = net->batch; t.h = L.out_h; t.w = L.out_w; t.c = L.out_c;
  t.ldc = R.own_ldc;
  t.dtype = R.dtype;
  return t;
}

