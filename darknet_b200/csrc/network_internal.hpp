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
  bool pair_s2 = false;  // inference: 3x3 stride-2 conv of 32 channels read through the pixel-pair view (network.cu)
  int pair_first = -1;   // inference: this 1x1 conv is computed by layer pair_first's kernel as its second output
  int pair_second = -1;  // ... and this one also computes layer pair_second
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
  bool head_decoded = false;  // head conv whose epilogue writes the following [yolo] layer's decoded output (no output of its own)
  int fused_conv = -1;      // shortcut computed by conv `fused_conv`
  void* buf = nullptr;      // owned allocation (root buffers only)
  int own_ldc = 0;
  int dtype = YC_BF16;
  std::vector<std::pair<int, int>> copies;  // route fallback: (input layer, dst channel offset)
  int consumers = 0;
  // per-layer training gates (parser.cpp:1362-1366; honoured by the backward / update loops, network_kernels.cu:140-143,
  // 231-241)
  int onlyforward = 0, dont_update = 0, burnin_update = 0, stopbackward = 0, train_only_bn = 0;
  // ---- training state (network_train.cu; allocated only with YC_NET_TRAIN) ----
  // yolo hyper-parameters (ParseYolo parser.cpp:312-415)
  int max_boxes = 90, iou_loss = 2 /*MSE*/, iou_thresh_kind = 0, focal_loss = 0;
  float ignore_thresh = .5f, truth_thresh = 1.f, iou_thresh = 1.f, iou_normalizer = .75f, cls_normalizer = 1.f,
        label_smooth_eps = 0.f, max_delta = 3.402823466e+38F, uc_normalizer = 1.f;
  float* w_master = nullptr;     // fp32 OIHW (l->weights_gpu)
  float* rmean_dev = nullptr;    // l->rolling_mean_gpu
  float* rvar_dev = nullptr;
  float* mean_dev = nullptr;     // l->mean_gpu
  float* var_dev = nullptr;
  float* stats_dev = nullptr;    // [2][n]: sum, sum of squares accumulated by the conv epilogue
  float* ws_dev = nullptr;       // 5n floats of batchnorm-backward workspace
  float* dw_packed = nullptr;    // l->weight_updates_gpu, [n][tap][ldw] fp32
  size_t bucket_floats = 0;      // dw_packed + bias_upd + scale_upd, contiguous (one all-reduce per layer)
  float* bias_upd = nullptr;     // l->bias_updates_gpu
  float* scale_upd = nullptr;    // l->scale_updates_gpu
  float* adam_m = nullptr;       // [net] adam=1: first / second moments, laid out like the gradient bucket (l->m_gpu, l->v_gpu,
  float* adam_v = nullptr;       //   bias_m/v, scale_m/v of convolutional_layer.cpp:589-600)
  void* w_packed_d = nullptr;    // data-gradient operand packing
  void* z_buf = nullptr;         // bf16 NHWC pre-batchnorm conv output (l->x_gpu)
  int z_ldc = 0;
  void* delta_buf = nullptr;     // bf16 NHWC l->delta_gpu (inside the network's delta slab)
  int delta_ldc = 0;
  void* dil_buf = nullptr;       // zero-inserted dz of strided layers
  int dil_h = 0, dil_w = 0;
  unsigned char* idx_buf = nullptr;  // maxpool winners (l->indexes_gpu)
  yc_wgrad_plan* wplan = nullptr;
  yc_conv_plan* dplan = nullptr;
  float* ydelta = nullptr;       // yolo: fp32 NCHW l->delta_gpu
  float* ystats = nullptr;       // yolo: 8 floats (cost, count, ...)
};

struct Op {
  int layer;
  int kind;  // 0 conv, 1 maxpool, 2 upsample, 3 shortcut, 4 route copy, 5 yolo decode, 6 input convert, 7 stem operand, 8 direct stem conv,
             // 9 several stride-1 max-pools of one tensor (SPP): outputs y, b, y3 with (size, pad) = (i0, i1), (i2, i3), (i4, i5)
  yc_tensor a, b, y, y3;
  int i4 = 0, i5 = 0, n_out = 1;
  int i0 = 0, i1 = 0, i2 = 0, i3 = 0;
  float f0 = 0.f;
};

struct yc_network {
  int batch = 1, w = 0, h = 0, c = 0, flags = YC_NET_DEFAULT;
  std::vector<NLayer> layers;
  std::string cfg_text;  // what the layers were built from (yc_network_resize re-derives the shapes from it)
  std::vector<Op> ops;
  bool finalized = false;
  cudaStream_t stream = nullptr;
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t graph_exec = nullptr;
  float* input_dev = nullptr;   // fp32 NCHW staging (net->input_state_gpu)
  float* input_pinned = nullptr;  // net->input_pinned_cpu
  // input prefetch (yc_network_prefetch): second pinned slot, device staging buffer, copy stream
  float* input_pinned2 = nullptr;
  float* input_dev2 = nullptr;
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t copy_done = nullptr, stage_free = nullptr;
  const float* staged_host = nullptr;    // batch whose H2D into input_dev2 has been started
  const float* resident_host = nullptr;  // batch already moved into input_dev, not yet consumed by a forward
  unsigned char* u8_pinned = nullptr;  // u8 HWC staging of yc_network_predict_u8 (allocated on first use)
  unsigned char* u8_dev = nullptr;
  unsigned char* u8_src_pinned = nullptr;  // frames of another size than the network's (yc_network_predict_u8_resized),
  unsigned char* u8_src_dev = nullptr;     //   grown on demand
  size_t u8_src_bytes = 0;
  yc::SgdItem* sgd_items = nullptr;  // device table of the one-launch update (built on the first yc_network_update)
  int sgd_n = 0, sgd_blocks = 0;
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
  // ---- training (network_train.cu) ----
  float lr = .001f, momentum = .9f, decay = .0001f;  // [net] learning_rate / momentum / decay (parser.cpp:1116-1140)
  int adam = 0;                                       // [net] adam / B1 / B2 / eps (parser.cpp:995-1001)
  float B1 = .9f, B2 = .999f, adam_eps = .000001f;
  int subdivisions = 1;
  // learning-rate schedule (ParseNetOptions parser.cpp:944-995, GetCurrLr network.cpp:32-84)
  int policy = 0;  // 0 constant 1 step 2 exp 3 poly 4 steps 5 sigmoid 6 random 7 sgdr
  int burn_in = 0, step = 1, max_iter = 0, sgdr_cycle = 0, sgdr_mult = 2;
  float lr_min = .00001f, scale = 1.f, gamma = 1.f, power = 4.f;
  std::vector<float> steps, scales;
  long long curr_iter = 0;  // net->curr_iter: completed updates
  bool any_gate = false;    // some layer has dont_update / burnin_update / train_only_bn: the update table is per call
  int train_flags = 0;
  bool stem = false;  // layer 0 runs as a 1x1 conv over the 32-column im2col operand built by yc_stem_im2col
  bool stem_pair = false;  // inference: the stem conv runs on pixel pairs (network.cu)
  bool stem_direct = false;  // inference: layer 0 is computed from the fp32 NCHW input by one kernel (no operand, no plan)
  bool stem_tc = false;      //   ... by yc_stem_conv_tc (tcgen05) rather than yc_stem_conv_direct (mma.sync)
  void* delta_slab = nullptr;
  size_t delta_slab_bytes = 0;
  float* truth_dev = nullptr;     // [batch][max_boxes*5] (*net->truth_gpu)
  float* truth_pinned = nullptr;
  int truth_floats = 0;
  float* cost_host = nullptr;     // pinned: 8 floats per yolo layer
  std::vector<Op> bwd_ops;
  long long seen = 0;
  int layers_loaded = 0;          // layers [0, layers_loaded) got their parameters from the weights file
  void* dp = nullptr;             // data-parallel context (NCCL), network_dp.cu
};

// network_train.cu
int yc_train_finalize(yc_network* net);
void yc_train_free(yc_network* net);
// network_dp.cu: called by the backward pass right after layer i's weight gradient has been enqueued on `compute`
// cudaMemcpy from PAGEABLE host memory returns once the data sits in the driver's staging buffer; the DMA into device
// memory may still be in flight on the legacy default stream.  The network's own stream is non-blocking, so a kernel
// launched there next (weight packing) is NOT ordered after that DMA: it read stale device memory for a random large
// layer of a freshly loaded network (found in round 1: outputs differed between two instances of the same net).
inline cudaError_t h2d_done() { return cudaStreamSynchronize(cudaStreamLegacy); }

int yc_dp_layer_ready(yc_network* net, int layer, cudaStream_t compute);
int yc_dp_backward_done(yc_network* net, cudaStream_t compute);
int yc_dp_reduce_pending(yc_network* net, cudaStream_t compute);
void yc_dp_update_done(yc_network* net);
// conv_fprop_sm100.cu: yc_conv_plan_create with a rectangular kernel / stride / padding along W
extern "C" int yc_conv_plan_create_rect(yc_conv_plan** out_plan, const yc_conv_desc* d, int kw, int stride_w, int pad_l,
                                        int pad_r, const yc_tensor* x, const void* w_packed, const float* bias,
                                        const float* scale, const yc_tensor* residual, const yc_tensor* y,
                                        unsigned plan_flags);
void yc_dp_free(yc_network* net);


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

