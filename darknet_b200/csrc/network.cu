// Whole-network host runtime (C++): the B200-native counterpart of the reference's parser.cpp (cfg + .weights),
// network.cpp (NetworkPredict / GetNetworkBoxes / FuseConvBatchNorm) and network_kernels.cu
// (ForwardNetworkGpu / NetworkPredictGpu).  No `struct layer` crosses into the kernels: this file translates the
// parsed network into POD descriptors and calls the launch ABI of include/yc_b200.h.
//
// Design (differs from the reference on purpose):
//   * device tensors are NHWC bf16 (heads fp32); the NCHW fp32 layout exists only at the seam (network input,
//     YOLO outputs, yc_network_layer_output).
//   * every conv is ONE kernel (implicit GEMM + bias/BN + activation + residual add), see conv_fprop_sm100.cu.
//   * [route] costs nothing: single-input routes (incl. groups/group_id slices, route_layer.c:125-142) are views;
//     multi-input routes own a concat buffer and their producers write straight into their channel slice.
//   * [shortcut] (shortcut_layer.c:191-204) is fused into the producing conv's epilogue when that conv has no other
//     consumer.
//   * the whole forward (input layout conversion ... YOLO decode) is captured once into a CUDA graph.
//   * all layer outputs stay allocated (like the reference, which keeps l->output_gpu per layer): yolov4-608 at
//     batch 64 is 22 GB of bf16, well inside 180 GB of HBM3e.
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <fstream>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

#include "cfg_parser.hpp"
#include "network_internal.hpp"
#include "yc_common.cuh"

namespace yc {
int launch_yolo_extract(const float* decoded, int n, int h, int w, int anchors, int classes, const float* anchor_wh,
                        const int* mask, int net_w, int net_h, float thresh, int gaussian, float* rows,
                        int rows_per_image, int* counts, int R, float tag, cudaStream_t stream);
int launch_yolo_gather(const float* regions, const int* counts, int heads, int n, int cap, int R, float* dense,
                       int* img_counts, int dense_cap, cudaStream_t stream);
}  // namespace yc

using namespace yc;

namespace {

// ---- cfg -> layers (ParseNetworkCfg, parser.cpp:1076-1519; shape rules cited per layer) --------------------
// w_over / h_over > 0 replace [net] width / height (yc_network_resize).
int build_layers(yc_network* net, const std::string& text, int batch, int w_over = 0, int h_over = 0) {
  std::vector<Section> secs;
  std::string err;
  if (!parse_cfg_text(text, secs, err)) NET_FAIL(YC_ERR_PARSE, "%s", err.c_str());
  const Section& n0 = secs[0];
  net->subdivisions = std::max(1, n0.get_int("subdivisions", 1));
  // ParseNetOptions (parser.cpp:926-928): the network's batch is the MINI-batch, [net] batch / subdivisions; an explicit
  // `batch` argument is the mini-batch itself (the reference's inference path forces it to 1, parser.cpp:1114-1115)
  net->batch = batch > 0 ? batch : std::max(1, n0.get_int("batch", 1) / net->subdivisions);
  net->w = w_over > 0 ? w_over : n0.get_int("width", 0);
  net->h = h_over > 0 ? h_over : n0.get_int("height", 0);
  net->c = n0.get_int("channels", 0);
  if (!(net->w > 0 && net->h > 0 && net->c > 0)) NET_FAIL(YC_ERR_PARSE, "[net] needs width/height/channels");
  net->lr = n0.get_float("learning_rate", .001f);   // ParseNetOptions parser.cpp:1116-1140
  net->momentum = n0.get_float("momentum", .9f);
  net->decay = n0.get_float("decay", .0001f);
  net->adam = n0.get_int("adam", 0);                 // parser.cpp:995-1001
  if (net->adam) {
    net->B1 = n0.get_float("B1", .9f);
    net->B2 = n0.get_float("B2", .999f);
    net->adam_eps = n0.get_float("eps", .000001f);
  }
  {  // learning-rate schedule (parser.cpp:944-995, 1028)
    const std::string pol = n0.get_str("policy", "constant");
    const char* names[] = {"constant", "step", "exp", "poly", "steps", "sigmoid", "random", "sgdr"};
    net->policy = 0;
    for (int k = 0; k < 8; ++k)
      if (pol == names[k]) net->policy = k;
    net->lr_min = n0.get_float("learning_rate_min", .00001f);
    net->burn_in = n0.get_int("burn_in", 0);
    net->power = n0.get_float("power", 4.f);
    net->step = n0.get_int("step", 1);
    net->scale = n0.get_float("scale", 1.f);
    net->gamma = n0.get_float("gamma", 1.f);
    net->sgdr_cycle = n0.get_int("sgdr_cycle", 0);
    net->sgdr_mult = n0.get_int("sgdr_mult", 2);
    net->steps = parse_float_list(n0.get_str("steps", ""));
    net->scales = parse_float_list(n0.get_str("scales", ""));
    if (net->policy == 4 && (net->steps.empty() || net->scales.size() < net->steps.size()))
      NET_FAIL(YC_ERR_PARSE, "STEPS policy must have steps and scales in cfg file");  // parser.cpp:963-964
    net->max_iter = n0.get_int("max_batches", 0);  // the reference's driver derives max_iter from max_epoch; see yc_network_set_schedule
  }
  int c = net->c, h = net->h, w = net->w;
  for (size_t si = 1; si < secs.size(); ++si) {
    const Section& s = secs[si];
    const int idx = (int)si - 1;
    NLayer L;
    L.c = c; L.h = h; L.w = w;
    if (s.name == "convolutional" || s.name == "conv") {  // ParseConv parser.cpp:179-242
      L.type = L_CONV;
      L.n = s.get_int("filters", 1);
      L.size = s.get_int("size", 1);
      int sx = s.get_int("stride_x", -1), sy = s.get_int("stride_y", -1);
      L.stride = s.get_int("stride", 1);
      if (sx >= 1 && sy >= 1 && sx != sy) NET_FAIL(YC_ERR_ARG, "layer %d: stride_x != stride_y is not supported", idx);
      if (sx >= 1 && sy >= 1) L.stride = sx;
      L.dilation = L.size == 1 ? 1 : s.get_int("dilation", 1);
      const int pad = s.get_int("pad", 0);
      L.pad = pad ? L.size / 2 : s.get_int("padding", 0);
      if (s.get_int("groups", 1) != 1) NET_FAIL(YC_ERR_ARG, "layer %d: grouped convolution is outside the hot path", idx);
      if (s.get_int("xnor", 0) || s.get_int("binary", 0)) NET_FAIL(YC_ERR_ARG, "layer %d: xnor/binary conv is outside the hot path", idx);
      L.bn = s.get_int("batch_normalize", 0);
      L.act = activation_id(s.get_str("activation", "logistic"));
      if (L.act < 0) NET_FAIL(YC_ERR_ARG, "layer %d: activation '%s' is outside the hot path", idx, s.get_str("activation", "").c_str());
      L.out_c = L.n;
      const int eff = L.dilation * (L.size - 1) + 1;  // convolutional_layer.cpp:87-95
      L.out_h = (h + 2 * L.pad - eff) / L.stride + 1;
      L.out_w = (w + 2 * L.pad - eff) / L.stride + 1;
      L.inputs = {idx - 1};
      const size_t nw = (size_t)L.n * L.c * L.size * L.size;
      L.weights.assign(nw, 0.f);
      L.biases.assign(L.n, 0.f);
      if (L.bn) { L.scales.assign(L.n, 1.f); L.rmean.assign(L.n, 0.f); L.rvar.assign(L.n, 1.f); }
      net->n_conv++;
    } else if (s.name == "maxpool" || s.name == "max") {  // ParseMaxpool parser.cpp:640-659
      L.type = L_MAXPOOL;
      L.stride = s.get_int("stride", 1);
      L.size = s.get_int("size", L.stride);
      L.pad = s.get_int("padding", L.size - 1);
      if (s.get_int("maxpool_depth", 0)) NET_FAIL(YC_ERR_ARG, "layer %d: maxpool_depth is outside the hot path", idx);
      L.out_c = c;
      L.out_h = (h + L.pad - L.size) / L.stride + 1;  // maxpool_layer.cpp:62-63
      L.out_w = (w + L.pad - L.size) / L.stride + 1;
      L.inputs = {idx - 1};
    } else if (s.name == "upsample") {  // parser.cpp:819-830
      L.type = L_UPSAMPLE;
      L.stride = s.get_int("stride", 2);
      L.scale_x_y = s.get_float("scale", 1.f);
      L.out_c = c; L.out_h = h * L.stride; L.out_w = w * L.stride;
      L.inputs = {idx - 1};
    } else if (s.name == "route") {  // ParseRoute parser.cpp:832-896
      L.type = L_ROUTE;
      std::vector<int> ids = parse_int_list(s.get_str("layers", ""));
      if (ids.empty()) NET_FAIL(YC_ERR_PARSE, "layer %d: route needs layers=", idx);
      L.groups = s.get_int("groups", 1);
      L.group_id = s.get_int("group_id", 0);
      int oc = 0;
      for (int& i : ids) {
        if (i < 0) i += idx;
        if (i < 0 || i >= idx) NET_FAIL(YC_ERR_PARSE, "layer %d: route input %d out of range", idx, i);
        const NLayer& src = net->layers[i];
        if (src.out_h != net->layers[ids[0]].out_h || src.out_w != net->layers[ids[0]].out_w)
          NET_FAIL(YC_ERR_ARG, "layer %d: route inputs differ in spatial size", idx);
        oc += src.out_c / L.groups;
      }
      L.inputs = ids;
      L.out_c = oc; L.out_h = net->layers[ids[0]].out_h; L.out_w = net->layers[ids[0]].out_w;
    } else if (s.name == "shortcut") {  // ParseShortcut parser.cpp:720-779 (single `from`)
      L.type = L_SHORTCUT;
      int f = s.get_int("from", -1);
      if (f < 0) f += idx;
      if (f < 0 || f >= idx) NET_FAIL(YC_ERR_PARSE, "layer %d: shortcut from out of range", idx);
      const NLayer& src = net->layers[f];
      if (src.out_c != c || src.out_h != h || src.out_w != w)  // shortcut_layer.c:151-164 asserts equal dims
        NET_FAIL(YC_ERR_ARG, "layer %d: shortcut source %d has a different shape", idx, f);
      L.from = f;
      L.act = activation_id(s.get_str("activation", "linear"));
      if (L.act < 0) NET_FAIL(YC_ERR_ARG, "layer %d: shortcut activation outside the hot path", idx);
      L.inputs = {idx - 1, f};
      L.out_c = c; L.out_h = h; L.out_w = w;
    } else if (s.name == "yolo" || s.name == "Gaussian_yolo") {  // ParseYolo :312-415, ParseGaussianYolo :443-557
      L.type = s.name == "yolo" ? L_YOLO : L_GYOLO;
      L.classes = s.get_int("classes", 20);
      L.total = s.get_int("num", 1);
      L.mask = parse_int_list(s.get_str("mask", ""));
      if (L.mask.empty()) for (int i = 0; i < L.total; ++i) L.mask.push_back(i);
      L.anchors = parse_float_list(s.get_str("anchors", ""));
      L.anchors.resize((size_t)2 * L.total, 0.5f);  // FillYoloLayer default biases .5 (yolo_layer.cpp:44-47)
      L.scale_x_y = s.get_float("scale_x_y", 1.f);
      L.beta_nms = s.get_float("beta_nms", 0.6f);
      const std::string nk = s.get_str("nms_kind", "default");
      L.nms_kind = nk == "diounms" ? 1 : 0;
      {  // loss hyper-parameters (ParseYolo parser.cpp:337-392)
        auto kind = [](const std::string& k, int dflt) {
          if (k == "iou") return 0; if (k == "giou") return 1; if (k == "mse") return 2; if (k == "diou") return 3;
          if (k == "ciou") return 4; return dflt;
        };
        L.max_boxes = s.get_int("max", 90);
        L.iou_loss = kind(s.get_str("iou_loss", "mse"), 0);
        L.iou_thresh_kind = kind(s.get_str("iou_thresh_kind", "iou"), 0);
        L.focal_loss = s.get_int("focal_loss", 0);
        L.ignore_thresh = s.get_float("ignore_thresh", .5f);
        L.truth_thresh = s.get_float("truth_thresh", 1.f);
        L.iou_thresh = s.get_float("iou_thresh", 1.f);
        L.iou_normalizer = s.get_float("iou_normalizer", .75f);
        L.cls_normalizer = s.get_float("cls_normalizer", 1.f);
        L.label_smooth_eps = s.get_float("label_smooth_eps", 0.f);
        L.max_delta = s.get_float("max_delta", 3.402823466e+38F);
        L.uc_normalizer = s.get_float("uc_normalizer", 1.f);  // ParseGaussianYolo parser.cpp:471
        const std::string yp = s.get_str("yolo_point", "center");
        if (L.type == L_GYOLO && yp != "center")
          NET_FAIL(YC_ERR_ARG, "layer %d: yolo_point=%s is not supported (center only)", idx, yp.c_str());
      }
      const int E = (L.type == L_YOLO ? 5 : 9) + L.classes;
      if ((int)L.mask.size() * E != c)
        NET_FAIL(YC_ERR_ARG, "layer %d: filters=%d does not correspond to classes/mask (want %d) (parser.cpp:323-330)", idx, c,
                 (int)L.mask.size() * E);
      L.n = (int)L.mask.size();
      L.out_c = c; L.out_h = h; L.out_w = w;
      L.inputs = {idx - 1};
      net->yolo_layers.push_back(idx);
    } else {
      NET_FAIL(YC_ERR_ARG, "layer %d: [%s] is outside the conv/YOLO hot path this library replaces", idx, s.name.c_str());
    }
    // training gates, parsed for every layer type (parser.cpp:1362-1366)
    L.onlyforward = s.get_int("onlyforward", 0);
    L.dont_update = s.get_int("dont_update", 0);
    L.burnin_update = s.get_int("burnin_update", 0);
    L.stopbackward = s.get_int("stopbackward", 0);
    L.train_only_bn = s.get_int("train_only_bn", 0);
    if (idx == 0 && L.type != L_CONV) NET_FAIL(YC_ERR_ARG, "the first layer must be convolutional");
    if (L.out_h <= 0 || L.out_w <= 0) NET_FAIL(YC_ERR_ARG, "layer %d: empty output", idx);
    net->layers.push_back(std::move(L));
    c = net->layers.back().out_c; h = net->layers.back().out_h; w = net->layers.back().out_w;
  }
  if (net->layers.empty()) NET_FAIL(YC_ERR_PARSE, "cfg has no layers");
  {  // SetTrainOnlyBn (parser.cpp:1062-1074): a train_only_bn mark applies to every EARLIER layer as well
    int mark = 0;
    for (int i = (int)net->layers.size() - 1; i >= 0; --i) {
      if (net->layers[i].train_only_bn) mark = net->layers[i].train_only_bn;
      if (mark) net->layers[i].train_only_bn = mark;
    }
  }
  for (const NLayer& L : net->layers)
    if (L.dont_update || L.burnin_update || L.train_only_bn) net->any_gate = true;
  return YC_OK;
}

// ---- .weights (LoadWeightsUpTo parser.cpp:1778-1844, LoadConvolutionalWeights :1695-1759) -------------------
int load_weights_file(yc_network* net, const char* path) {
  FILE* fp = fopen(path, "rb");
  if (!fp) NET_FAIL(YC_ERR_IO, "cannot open weights file %s", path);
  int hdr[3];
  uint64_t seen = 0;
  if (!(fread(hdr, sizeof(int), 3, fp) == 3 && fread(&seen, sizeof(uint64_t), 1, fp) == 1)) {
    fclose(fp);
    NET_FAIL(YC_ERR_IO, "weights file %s: bad header", path);
  }
  net->seen = (long long)seen;  // *net->seen (parser.cpp:1797-1806); curr_iter = seen / (batch * subdivisions)
  net->curr_iter = (long long)(seen / (uint64_t)std::max(1, net->batch * net->subdivisions));
  // LoadWeightsUpTo stops quietly at the end of the file (parser.cpp:1837): a backbone file such as yolov4.conv.137
  // fills the first layers and leaves the rest at their initial values.  A file that ends INSIDE a layer is an error.
  for (size_t i = 0; i < net->layers.size(); ++i) {
    NLayer& L = net->layers[i];
    if (L.type != L_CONV) continue;
    const size_t got = fread(L.biases.data(), 4, L.n, fp);
    if (got == 0 && feof(fp)) break;  // clean end at a layer boundary
    bool ok = got == (size_t)L.n;
    if (ok && L.bn) {
      ok = fread(L.scales.data(), 4, L.n, fp) == (size_t)L.n && fread(L.rmean.data(), 4, L.n, fp) == (size_t)L.n &&
           fread(L.rvar.data(), 4, L.n, fp) == (size_t)L.n;
    }
    if (ok) ok = fread(L.weights.data(), 4, L.weights.size(), fp) == L.weights.size();
    if (!ok) {
      fclose(fp);
      NET_FAIL(YC_ERR_IO, "weights file %s is truncated inside layer %zu", path, i);
    }
    net->layers_loaded = (int)i + 1;
  }
  fclose(fp);
  return YC_OK;
}

int free_device_state(yc_network* net) {
  yc_train_free(net);
  if (net->graph_exec) { cudaGraphExecDestroy(net->graph_exec); net->graph_exec = nullptr; }
  if (net->graph) { cudaGraphDestroy(net->graph); net->graph = nullptr; }
  for (NLayer& L : net->layers) {
    if (L.plan) { yc_conv_plan_destroy(L.plan); L.plan = nullptr; }
    if (L.w_packed) { cudaFree(L.w_packed); L.w_packed = nullptr; }
    if (L.bias_dev) { cudaFree(L.bias_dev); L.bias_dev = nullptr; }
    if (L.scale_dev) { cudaFree(L.scale_dev); L.scale_dev = nullptr; }
    if (L.buf) { cudaFree(L.buf); L.buf = nullptr; }
    if (L.dec_dev) { cudaFree(L.dec_dev); L.dec_dev = nullptr; }
    if (L.dec_host) { cudaFreeHost(L.dec_host); L.dec_host = nullptr; }
    if (L.anchors_dev) { cudaFree(L.anchors_dev); L.anchors_dev = nullptr; }
    if (L.mask_dev) { cudaFree(L.mask_dev); L.mask_dev = nullptr; }
    L.alias = false; L.alias_src = -1; L.alias_coff = 0; L.parent = -1; L.parent_coff = 0;
    L.fused_into = -1; L.fused_conv = -1; L.copies.clear(); L.consumers = 0; L.head_decoded = false;
    L.pair_first = -1; L.pair_second = -1;
  }
  if (net->input_dev) { cudaFree(net->input_dev); net->input_dev = nullptr; }
  if (net->input_pinned) { cudaFreeHost(net->input_pinned); net->input_pinned = nullptr; }
  if (net->u8_pinned) { cudaFreeHost(net->u8_pinned); net->u8_pinned = nullptr; }
  if (net->u8_src_pinned) { cudaFreeHost(net->u8_src_pinned); net->u8_src_pinned = nullptr; }
  if (net->u8_src_dev) { cudaFree(net->u8_src_dev); net->u8_src_dev = nullptr; }
  net->u8_src_bytes = 0;
  if (net->copy_stream) { cudaStreamSynchronize(net->copy_stream); cudaStreamDestroy(net->copy_stream); net->copy_stream = nullptr; }
  if (net->copy_done) { cudaEventDestroy(net->copy_done); net->copy_done = nullptr; }
  if (net->stage_free) { cudaEventDestroy(net->stage_free); net->stage_free = nullptr; }
  if (net->input_pinned2) { cudaFreeHost(net->input_pinned2); net->input_pinned2 = nullptr; }
  if (net->input_dev2) { cudaFree(net->input_dev2); net->input_dev2 = nullptr; }
  net->staged_host = net->resident_host = nullptr;
  if (net->u8_dev) { cudaFree(net->u8_dev); net->u8_dev = nullptr; }
  if (net->sgd_items) { cudaFree(net->sgd_items); net->sgd_items = nullptr; }
  if (net->input_nhwc) { cudaFree(net->input_nhwc); net->input_nhwc = nullptr; }
  if (net->det_regions) { cudaFree(net->det_regions); net->det_regions = nullptr; }
  if (net->det_counts) { cudaFree(net->det_counts); net->det_counts = nullptr; }
  if (net->det_dense) { cudaFree(net->det_dense); net->det_dense = nullptr; }
  if (net->det_img_counts) { cudaFree(net->det_img_counts); net->det_img_counts = nullptr; }
  if (net->det_host) { cudaFreeHost(net->det_host); net->det_host = nullptr; }
  if (net->det_img_counts_host) { cudaFreeHost(net->det_img_counts_host); net->det_img_counts_host = nullptr; }
  net->ops.clear();
  net->finalized = false;
  return YC_OK;
}

int run_op(yc_network* net, const Op& op, cudaStream_t st) {
  NLayer* L = op.layer >= 0 ? &net->layers[op.layer] : nullptr;
  switch (op.kind) {
    case 0: return yc_conv_fprop(L->plan, st);
    case 1: return yc_maxpool_fwd(&op.a, &op.y, op.i0, op.i1, op.i2, op.i3, st);
    case 2: return yc_upsample_fwd(&op.a, &op.y, op.i0, op.f0, st);
    case 9: {
      const yc_tensor* ys[3] = {&op.y, &op.b, &op.y3};
      const int sizes[3] = {op.i0, op.i2, op.i4}, pads[3] = {op.i1, op.i3, op.i5};
      return yc_maxpool_s1_multi(&op.a, op.n_out, ys, sizes, pads, st);
    }
    case 3: return yc_shortcut_fwd(&op.a, &op.b, &op.y, op.i0, st);
    case 4: return yc_route_copy(&op.a, &op.y, st);
    case 5: return yc_yolo_decode(&op.a, L->dec_dev, op.i0, op.i1, op.f0, op.i2, st);
    case 6: return yc_nchw_to_nhwc(net->graph_input, &op.y, st);
    case 7: return yc_stem_im2col(net->graph_input, net->c, net->h, net->w, op.i0, op.i1, op.i2, &op.y, st);
    case 8: return (op.i1 ? yc_stem_conv_tc : yc_stem_conv_direct)(net->graph_input, net->batch, net->h, net->w, L->w_packed, L->bias_dev, op.i0, &op.y, st);
  }
  return YC_ERR_ARG;
}

int run_ops(yc_network* net, cudaStream_t st) {
  for (const Op& op : net->ops) {
    int rc = run_op(net, op, st);
    if (rc) return rc;
  }
  return YC_OK;
}

}  // namespace

// ================================================================================================
extern "C" int yc_network_finalize(yc_network* net) {
  YC_REQUIRE(net, "yc_network_finalize: null network");
  if (yc_device_count() <= 0) {
    set_error("yc_network_finalize: no CUDA device; this library has no CPU fallback");
    return YC_ERR_NO_DEVICE;
  }
  free_device_state(net);
  const int nl = (int)net->layers.size();
  if (net->flags & YC_NET_TRAIN) {  // training keeps batchnorm live and every layer's own output; ops run un-graphed
    net->flags &= ~(YC_NET_FUSE_BN | YC_NET_FUSE_SHORTCUT | YC_NET_CUDA_GRAPH | YC_NET_PULL_HEADS);
    // Default training semantics = the reference's GPU path, which this library replaces: the data gradient ADDS to the
    // previous layer's delta (cudnnConvolutionBackwardData beta = 1, convolutional_kernels.cu:735-738) and the batchnorm
    // shift is trained (backward_bias_gpu, batchnorm_layer.cpp:330).  YC_NET_TRAIN_REF_CPU selects the CPU path's
    // behaviour instead (col2im overwrites, no gradient for the shift) — the mode the CPU oracle is
    // compared in.
    if (net->flags & YC_NET_TRAIN_REF_CPU) net->flags &= ~(YC_NET_DGRAD_ACCUMULATE | YC_NET_BN_BIAS_GRAD);
    else net->flags |= YC_NET_DGRAD_ACCUMULATE | YC_NET_BN_BIAS_GRAD;
  }
  const int flags = net->flags;
  auto& Ls = net->layers;
  if (!net->stream) NET_CUDA(cudaStreamCreateWithFlags(&net->stream, cudaStreamNonBlocking));

  // ---- consumers ------------------------------------------------------------------------------
  for (int i = 0; i < nl; ++i)
    for (int in : Ls[i].inputs)
      if (in >= 0) Ls[in].consumers++;

  // ---- aliases: single-input routes are views (route_layer.c:125-142 with groups/group_id) ----------
  for (int i = 0; i < nl; ++i) {
    NLayer& L = Ls[i];
    if (L.type == L_ROUTE && L.inputs.size() == 1 && (flags & YC_NET_ELIDE_ROUTE)) {
      const NLayer& S = Ls[L.inputs[0]];
      const int part = S.out_c / L.groups;
      if ((part * L.group_id) % 8 == 0 && S.type != L_YOLO && S.type != L_GYOLO) {
        L.alias = true;
        L.alias_src = L.inputs[0];
        L.alias_coff = part * L.group_id;
      }
    }
  }
  // ---- shortcut fusion --------------------------------------------------------------------------
  if (flags & YC_NET_FUSE_SHORTCUT) {
    for (int i = 1; i < nl; ++i) {
      NLayer& S = Ls[i];
      if (S.type != L_SHORTCUT) continue;
      NLayer& C = Ls[i - 1];
      if (C.type == L_CONV && C.consumers == 1 && S.from != i - 1) {
        C.fused_into = i;
        S.fused_conv = i - 1;
      }
    }
  }
  // ---- concat placement ---------------------------------------------------------------------------
  for (int i = 0; i < nl; ++i) {
    NLayer& R = Ls[i];
    if (R.type != L_ROUTE || R.alias) continue;
    int coff = 0;
    for (int in : R.inputs) {
      NLayer& S = Ls[in];
      const int part = S.out_c / R.groups;
      // physical producer whose whole output is this slice, not yet placed elsewhere
      int p = in;
      bool whole = R.groups == 1;
      while (whole && Ls[p].alias) {
        if (Ls[p].alias_coff != 0 || Ls[p].out_c != Ls[Ls[p].alias_src].out_c) { whole = false; break; }
        p = Ls[p].alias_src;
      }
      if (whole && Ls[p].fused_conv >= 0) { /* shortcut computed by a conv: the shortcut layer is the physical one */ }
      const bool placeable = (flags & YC_NET_ELIDE_ROUTE) && whole && Ls[p].parent < 0 && Ls[p].fused_into < 0 &&
                             Ls[p].type != L_YOLO && Ls[p].type != L_GYOLO && coff % 8 == 0 && p != i;
      if (placeable) {
        Ls[p].parent = i;
        Ls[p].parent_coff = coff;
      } else {
        R.copies.push_back({in, coff});
      }
      coff += part;
    }
  }
  // ---- allocate root buffers -----------------------------------------------------------------------
  for (int i = 0; i < nl; ++i) {
    NLayer& L = Ls[i];
    if (L.type == L_YOLO || L.type == L_GYOLO) continue;
    if (L.alias || L.parent >= 0 || L.fused_into >= 0) continue;
    // head convs (feeding a YOLO layer) keep fp32 so thresholds see unrounded logits
    const bool head = (i + 1 < nl) && (Ls[i + 1].type == L_YOLO || Ls[i + 1].type == L_GYOLO) && L.type == L_CONV;
    L.dtype = head ? YC_F32 : YC_BF16;
    L.own_ldc = round_up(L.out_c, head ? 4 : 8);
    const size_t bytes = (size_t)net->batch * L.out_h * L.out_w * L.own_ldc * (head ? 4 : 2);
    NET_CUDA(cudaMalloc(&L.buf, bytes));
    NET_CUDA(cudaMemset(L.buf, 0, bytes));
  }
  // input staging
  const size_t in_elems = (size_t)net->batch * net->c * net->h * net->w;
  NET_CUDA(cudaMalloc(&net->input_dev, in_elems * 4));
  NET_CUDA(cudaHostAlloc(&net->input_pinned, in_elems * 4, cudaHostAllocDefault));
  // Stem as im2col + 1x1 (see yc_stem_im2col): the input conversion writes the first conv's GEMM operand directly.
  const bool stem = (flags & YC_NET_STEM_IM2COL) && nl > 0 && Ls[0].type == L_CONV && Ls[0].size > 1 &&
                    Ls[0].dilation == 1 && Ls[0].c * Ls[0].size * Ls[0].size <= 32;
  net->stem = stem;
  net->stem_pair = false;  // decided once the buffers are placed (needs layer 0's output pitch)
  const int in_ldc = stem ? 32 : round_up(net->c, 8);
  const int in_h = stem ? Ls[0].out_h : net->h, in_w = stem ? Ls[0].out_w : net->w, in_c = stem ? 32 : net->c;
  NET_CUDA(cudaMalloc(&net->input_nhwc, (size_t)net->batch * in_h * in_w * in_ldc * 2));
  NET_CUDA(cudaMemset(net->input_nhwc, 0, (size_t)net->batch * in_h * in_w * in_ldc * 2));

  if (flags & YC_NET_TRAIN) {
    int rc = yc_train_finalize(net);
    if (rc) return rc;
    net->finalized = true;
    return YC_OK;
  }
  const unsigned plan_flags = (((unsigned)flags >> YC_NET_PLAN_SHIFT) & 0x7fffu) | ((flags & YC_NET_PDL) ? (unsigned)YC_PLAN_PDL : 0u);
  // The 3 -> 8..32 channel 3x3 / stride 1 / pad 1 first layer of the YOLO backbones (folded batchnorm) is computed straight
  // from the fp32 NCHW input, no operand in global memory at all: by yc_stem_conv_tc (tcgen05; the default), by
  // yc_stem_conv_direct under YC_NET_STEM_DIRECT (mma.sync, measured slower), or — YC_NET_STEM_OPERAND — through the round-1
  // path (im2col operand built by the input conversion + a 1x1 layer; A/B and parity coverage).
  const bool stem_kernel_ok = stem && Ls[0].c == 3 && Ls[0].size == 3 && Ls[0].stride == 1 && Ls[0].pad == 1 &&
                              Ls[0].n % 8 == 0 && Ls[0].n <= 32 && Ls[0].fused_into < 0 &&
                              (!Ls[0].bn || (flags & YC_NET_FUSE_BN)) && view_of(net, 0).dtype == YC_BF16 &&
                              (Ls[0].act == YC_MISH || Ls[0].act == YC_LEAKY || Ls[0].act == YC_LINEAR);
  net->stem_tc = stem_kernel_ok && !(flags & (YC_NET_STEM_DIRECT | YC_NET_STEM_OPERAND)) && net->w % 4 == 0;
  net->stem_direct = stem_kernel_ok && ((flags & YC_NET_STEM_DIRECT) || net->stem_tc);
  if (!net->stem_direct && stem && Ls[0].n % 8 == 0 && 2 * Ls[0].n <= 256 && Ls[0].out_w % 2 == 0 && Ls[0].fused_into < 0 &&
      view_of(net, 0).ldc == Ls[0].n && view_of(net, 0).dtype == YC_BF16 && !(flags & YC_NET_NO_STEM_PAIR))
    net->stem_pair = true;
  // ---- pairs of 1x1 convolutions that read the same tensor (the two branches of a CSP block: conv, route back,
  // conv) become ONE GEMM with both filters stacked along out_c: the input is read once and the tile is twice as wide.
  // The second layer's output is produced when the first one runs (its buffer has no other writer).
  if (!(flags & YC_NET_NO_PAIR_FUSE)) {
    auto in_view = [&](int i) { return i == 0 ? yc_tensor{nullptr, 0, 0, 0, 0, 0, 0} : view_of(net, i - 1); };
    auto eligible = [&](int i) {
      const NLayer& L = Ls[i];
      return i > 0 && L.type == L_CONV && L.size == 1 && L.stride == 1 && L.pad == 0 && L.fused_into < 0 &&
             L.pair_first < 0 && L.pair_second < 0 && view_of(net, i).dtype == YC_BF16;
    };
    for (int a = 1; a < nl; ++a) {
      if (!eligible(a) || Ls[a].n % 32 != 0) continue;
      const yc_tensor va = in_view(a);
      for (int b = a + 1; b < std::min(nl, a + 4); ++b) {
        if (!eligible(b)) continue;
        const yc_tensor vb = in_view(b);
        if (vb.ptr != va.ptr || vb.c != va.c || vb.ldc != va.ldc || vb.dtype != va.dtype) continue;
        if (Ls[b].act != Ls[a].act || Ls[b].bn != Ls[a].bn) continue;
        Ls[a].pair_second = b;
        Ls[b].pair_first = a;
        break;
      }
    }
  }
  // ---- parameters: fuse BN (FuseConvBatchNorm network.cpp:647-682), pack, upload ----------------------
  for (int i = 0; i < nl; ++i) {
    NLayer& L = Ls[i];
    if (L.type != L_CONV) continue;
    const size_t fs = (size_t)L.size * L.size * L.c;
    std::vector<float> w = L.weights, b = L.biases, sc;
    if (L.bn) {
      if (flags & YC_NET_FUSE_BN) {
        for (int f = 0; f < L.n; ++f) {
          const float sd = (float)sqrt((double)(L.rvar[f] + 0.00001f));
          b[f] -= L.scales[f] * L.rmean[f] / sd;
          const float m = L.scales[f] / sd;
          for (size_t k = 0; k < fs; ++k) w[f * fs + k] *= m;
        }
      } else {
        // inference-mode BN from rolling stats (batchnorm_layer.cpp:226-237): (x-mean)/sqrt(var+1e-6)*gamma + beta
        sc.resize(L.n);
        for (int f = 0; f < L.n; ++f) {
          const float inv = 1.f / sqrtf(L.rvar[f] + .000001f);
          sc[f] = L.scales[f] * inv;
          b[f] = L.biases[f] - L.rmean[f] * sc[f];
        }
      }
    }
    int pc = L.c, ps = L.size;  // geometry the packed weights are laid out for
    if (stem && i == 0) {       // OIHW [n][c][kh][kw] -> [n][32] with column (kh*size+kw)*c + ci, as yc_stem_im2col writes
      std::vector<float> w2((size_t)L.n * 32, 0.f);
      for (int f = 0; f < L.n; ++f)
        for (int ci = 0; ci < L.c; ++ci)
          for (int t = 0; t < L.size * L.size; ++t) w2[(size_t)f * 32 + t * L.c + ci] = w[((size_t)f * L.c + ci) * L.size * L.size + t];
      w.swap(w2);
      pc = 32; ps = 1;
    }
    // 3x3 stride-2 pad-1 conv of 32 dense channels (layer 1 of the YOLO backbones): read the input as pixel pairs
    // [N][H][W/2][64].  Output column wo needs input columns 2wo-1, 2wo, 2wo+1 = the second half of pair wo-1 and both
    // halves of pair wo: a 3x2 kernel over pairs, stride 1 / left pad 1 along W, with zero weights for the unused
    // half.  6 TMA loads of 128-byte rows per tile instead of 9 of 64-byte rows (the request rate bounds this layer).
    L.pair_s2 = false;
    if (i > 0 && !(stem && i == 0) && L.size == 3 && L.stride == 2 && L.pad == 1 && L.dilation == 1 && L.c == 32 &&
        L.w % 2 == 0 && view_of(net, i - 1).ldc == 32 && view_of(net, i - 1).dtype == YC_BF16 &&
        !(flags & YC_NET_NO_PAIR_S2)) {
      L.pair_s2 = true;
      std::vector<float> w2((size_t)L.n * 6 * 64, 0.f);  // [n][kh][pw][64]
      for (int f = 0; f < L.n; ++f)
        for (int ci = 0; ci < 32; ++ci)
          for (int kh = 0; kh < 3; ++kh) {
            const float* src = &w[((size_t)f * 32 + ci) * 9 + kh * 3];
            float* dst = &w2[((size_t)f * 6 + kh * 2) * 64];
            dst[32 + ci] = src[0];       // pair wo-1, second pixel  <- kw = 0
            dst[64 + ci] = src[1];       // pair wo,   first pixel   <- kw = 1
            dst[64 + 32 + ci] = src[2];  // pair wo,   second pixel  <- kw = 2
          }
      w.swap(w2);
      pc = 6 * 64; ps = 1;  // packs as [n][1][384] == [n][6 taps][64]
    }
    int pn = L.n;  // output channels of the packed operand
    if (stem && i == 0 && net->stem_pair) {
      // Pixel pairs: the stem operand is read as [N][H][W/2][64] (two pixels' im2col rows = one 128-byte row) and
      // multiplied by the block-diagonal [2n][64] filter; an output row is then the 2n channels of the same two
      // pixels, which is exactly their NHWC layout.  Half the tiles, full-width TMA rows (64-byte rows are
      // request-rate bound), at the price of MMAs on the zero blocks (the layer is nowhere near tensor-bound).
      std::vector<float> w3((size_t)2 * L.n * 64, 0.f);
      for (int f = 0; f < L.n; ++f)
        for (int k = 0; k < 32; ++k) {
          w3[(size_t)f * 64 + k] = w[(size_t)f * 32 + k];
          w3[(size_t)(L.n + f) * 64 + 32 + k] = w[(size_t)f * 32 + k];
        }
      w.swap(w3);
      pc = 64; pn = 2 * L.n;
      b.resize(pn);
      for (int f = 0; f < L.n; ++f) b[L.n + f] = b[f];
      if (!sc.empty()) {
        sc.resize(pn);
        for (int f = 0; f < L.n; ++f) sc[L.n + f] = sc[f];
      }
    }
    float* w_dev = nullptr;
    NET_CUDA(cudaMalloc(&w_dev, w.size() * 4));
    NET_CUDA(cudaMemcpy(w_dev, w.data(), w.size() * 4, cudaMemcpyHostToDevice));
    NET_CUDA(h2d_done());
    // a fused pair shares the first layer's buffers: [first's rows | second's rows]
    const int pall = L.pair_second >= 0 ? pn + Ls[L.pair_second].n : pn;
    void* w_dst;
    float *b_dst, *s_dst = nullptr;
    if (L.pair_first >= 0) {
      NLayer& A = Ls[L.pair_first];
      w_dst = (char*)A.w_packed + yc_conv_packed_bytes(A.n, pc, ps);
      b_dst = A.bias_dev + A.n;
      if (!sc.empty()) s_dst = A.scale_dev + A.n;
    } else {
      NET_CUDA(cudaMalloc(&L.w_packed, yc_conv_packed_bytes(pall, pc, ps)));
      NET_CUDA(cudaMalloc(&L.bias_dev, pall * 4));
      if (!sc.empty()) NET_CUDA(cudaMalloc(&L.scale_dev, pall * 4));
      w_dst = L.w_packed; b_dst = L.bias_dev; s_dst = L.scale_dev;
    }
    int rc = yc_conv_pack(w_dev, pn, pc, ps, w_dst, net->stream);
    if (rc) { cudaFree(w_dev); return rc; }
    NET_CUDA(cudaStreamSynchronize(net->stream));
    NET_CUDA(cudaFree(w_dev));
    NET_CUDA(cudaMemcpy(b_dst, b.data(), pn * 4, cudaMemcpyHostToDevice));
    NET_CUDA(h2d_done());
    if (!sc.empty()) {
      NET_CUDA(cudaMemcpy(s_dst, sc.data(), pn * 4, cudaMemcpyHostToDevice));
      NET_CUDA(h2d_done());
    }
  }

  // ---- op list -------------------------------------------------------------------------------------
  if (!net->stem_direct) {
    Op op;
    op.layer = -1; op.kind = stem ? 7 : 6;
    op.y = yc_tensor{net->input_nhwc, net->batch, in_h, in_w, in_c, in_ldc, YC_BF16};
    if (stem) { op.i0 = Ls[0].size; op.i1 = Ls[0].stride; op.i2 = Ls[0].pad; }
    net->ops.push_back(op);
  }
  auto input_view = [&](int i) -> yc_tensor {
    if (i == 0) return yc_tensor{net->input_nhwc, net->batch, in_h, in_w, in_c, in_ldc, YC_BF16};
    return view_of(net, i - 1);
  };
  for (int i = 0; i < nl; ++i) {
    NLayer& L = Ls[i];
    Op op;
    op.layer = i;
    switch (L.type) {
      case L_CONV: {
        if (L.pair_first >= 0) break;  // computed by layer pair_first's kernel (second output tensor)
        if (i == 0 && net->stem_direct) {  // input conversion + conv in one kernel, no plan
          op.kind = 8; op.i0 = L.act; op.i1 = net->stem_tc ? 1 : 0; op.y = view_of(net, 0);
          net->ops.push_back(op);
          break;
        }
        yc_conv_desc d{L.size, L.stride, L.pad, L.dilation, L.act, YC_LINEAR};
        if (stem && i == 0) { d.ksize = 1; d.stride = 1; d.pad = 0; }
        yc_tensor x = input_view(i);
        yc_tensor y, res;
        if (stem && i == 0 && net->stem_pair) {  // pixel pairs: [N][H][W/2][64] -> [N][H][W/2][2n]
          x.w /= 2; x.c = 64; x.ldc = 64;
        }
        const yc_tensor* resp = nullptr;
        if (L.fused_into >= 0) {
          const NLayer& S = Ls[L.fused_into];
          y = view_of(net, L.fused_into);
          res = view_of(net, S.from);
          resp = &res;
          d.residual_activation = S.act;
        } else {
          y = view_of(net, i);
        }
        if (stem && i == 0 && net->stem_pair) { y.w /= 2; y.c *= 2; y.ldc *= 2; }
        int rc;
        if (L.pair_s2) {
          x.w /= 2; x.c = 64; x.ldc = 64;
          rc = yc_conv_plan_create_rect(&L.plan, &d, 2, 1, 1, 0, &x, L.w_packed, L.bias_dev, L.scale_dev, resp, &y,
                                        plan_flags);
        } else if (L.pair_second >= 0) {
          const yc_tensor y2 = view_of(net, L.pair_second);
          rc = yc_conv_plan_create_ex(&L.plan, &d, &x, L.w_packed, L.bias_dev, L.scale_dev, nullptr, &y, &y2, plan_flags);
        } else {
          rc = yc_conv_plan_create_ex(&L.plan, &d, &x, L.w_packed, L.bias_dev, L.scale_dev, resp, &y, nullptr, plan_flags);
        }
        if (rc) return rc;
        op.kind = 0;
        net->ops.push_back(op);
        break;
      }
      case L_MAXPOOL:
        op.kind = 1; op.a = input_view(i); op.y = view_of(net, i);
        op.i0 = L.size; op.i1 = L.stride; op.i2 = L.pad; op.i3 = (flags & YC_NET_REF_CPU_MAXPOOL) ? 1 : 0;
        net->ops.push_back(op);
        break;
      case L_UPSAMPLE:
        op.kind = 2; op.a = input_view(i); op.y = view_of(net, i); op.i0 = L.stride; op.f0 = L.scale_x_y;
        net->ops.push_back(op);
        break;
      case L_SHORTCUT:
        if (L.fused_conv >= 0) break;  // computed by the conv's epilogue
        op.kind = 3; op.a = input_view(i); op.b = view_of(net, L.from); op.y = view_of(net, i); op.i0 = L.act;
        net->ops.push_back(op);
        break;
      case L_ROUTE:
        if (L.alias) break;
        for (auto& cp : L.copies) {
          const NLayer& S = Ls[cp.first];
          const int part = S.out_c / L.groups;
          Op c2; c2.layer = i; c2.kind = 4;
          c2.a = view_of(net, cp.first);
          c2.a.ptr = (char*)c2.a.ptr + (size_t)part * L.group_id * (c2.a.dtype == YC_F32 ? 4 : 2);
          c2.a.c = part;
          c2.y = view_of(net, i);
          c2.y.ptr = (char*)c2.y.ptr + (size_t)cp.second * 2;
          c2.y.c = part;
          if (c2.a.dtype != YC_BF16) NET_FAIL(YC_ERR_ARG, "layer %d: route from an fp32 head output is not supported", i);
          net->ops.push_back(c2);
        }
        break;
      case L_YOLO:
      case L_GYOLO: {
        const size_t elems = (size_t)net->batch * L.out_c * L.out_h * L.out_w;
        NET_CUDA(cudaMalloc(&L.dec_dev, elems * 4));
        NET_CUDA(cudaHostAlloc(&L.dec_host, elems * 4, cudaHostAllocDefault));
        memset(L.dec_host, 0, elems * 4);
        NET_CUDA(cudaMalloc(&L.anchors_dev, L.anchors.size() * 4));
        NET_CUDA(cudaMemcpy(L.anchors_dev, L.anchors.data(), L.anchors.size() * 4, cudaMemcpyHostToDevice));
    NET_CUDA(h2d_done());
        NET_CUDA(cudaMalloc(&L.mask_dev, L.mask.size() * 4));
        NET_CUDA(cudaMemcpy(L.mask_dev, L.mask.data(), L.mask.size() * 4, cudaMemcpyHostToDevice));
    NET_CUDA(h2d_done());
        // ForwardYoloLayerGpu inside the head convolution's epilogue: when the layer before is a convolution with an fp32
        // output that nobody else reads, its plan decodes and writes this layer's NCHW output itself — no NHWC head tensor,
        // no decode kernel.  Opt-in (YC_PLAN_HEAD_DECODE): measured neutral on the step and it moves decode time into the
        // conv kernels (profiles/r02_ab_head_decode.txt)
        if ((plan_flags & YC_PLAN_HEAD_DECODE) && i > 0 && Ls[i - 1].type == L_CONV && Ls[i - 1].plan &&
            Ls[i - 1].consumers == 1 && Ls[i - 1].fused_into < 0 && Ls[i - 1].pair_first < 0 && Ls[i - 1].pair_second < 0 &&
            L.inputs.size() == 1 && L.inputs[0] == i - 1 && view_of(net, i - 1).dtype == YC_F32 &&
            yc_conv_plan_info(Ls[i - 1].plan, 5) != 2 &&
            yc_conv_plan_set_yolo_decode(Ls[i - 1].plan, L.dec_dev, L.out_c / L.n, L.scale_x_y, L.type == L_GYOLO ? 1 : 0) == YC_OK) {
          Ls[i - 1].head_decoded = true;
          break;
        }
        op.kind = 5; op.a = input_view(i); op.i0 = L.n; op.i1 = L.classes; op.f0 = L.scale_x_y;
        op.i2 = L.type == L_GYOLO ? 1 : 0;
        net->ops.push_back(op);
        break;
      }
    }
  }
  // ---- the SPP block: consecutive stride-1 max-pools that read the SAME tensor (route elision makes the route layers
  // between them views) become one launch that stages the input once (yc_maxpool_s1_multi)
  {
    std::vector<Op> merged;
    for (size_t k = 0; k < net->ops.size(); ++k) {
      Op op = net->ops[k];
      auto poolable = [&](const Op& o) {
        return o.kind == 1 && o.i3 == 0 && yc_maxpool_s1_multi_eligible(&o.a, &o.y, o.i0, o.i1, o.i2);
      };
      auto same_input = [&](const Op& o) {
        return o.a.ptr == op.a.ptr && o.a.n == op.a.n && o.a.h == op.a.h && o.a.w == op.a.w && o.a.c == op.a.c &&
               o.a.ldc == op.a.ldc && o.a.dtype == op.a.dtype;
      };
      if (poolable(op)) {
        size_t e = k + 1;
        while (e < net->ops.size() && e - k < 3 && poolable(net->ops[e]) && same_input(net->ops[e])) ++e;
        if (e - k >= 2) {
          Op m = op;  // op.layer stays the first pool's: per-layer timing reports the block there
          m.kind = 9;
          m.n_out = (int)(e - k);
          m.i0 = op.i0; m.i1 = op.i2;
          m.b = net->ops[k + 1].y; m.i2 = net->ops[k + 1].i0; m.i3 = net->ops[k + 1].i2;
          if (e - k == 3) { m.y3 = net->ops[k + 2].y; m.i4 = net->ops[k + 2].i0; m.i5 = net->ops[k + 2].i2; }
          merged.push_back(m);
          k = e - 1;
          continue;
        }
      }
      merged.push_back(op);
    }
    net->ops.swap(merged);
  }
  NET_CUDA(cudaDeviceSynchronize());  // memsets / uploads of the set-up ran on the legacy stream: finish them
  net->finalized = true;
  return YC_OK;
}

static int load_common(yc_network** out, const std::string& text, const char* weights_path, int batch, int flags) {
  YC_REQUIRE(out, "yc_network_load: null output");
  yc_network* net = new yc_network();
  net->flags = flags;
  net->cfg_text = text;
  int rc = build_layers(net, text, batch);
  if (!rc && weights_path && weights_path[0]) {
    rc = load_weights_file(net, weights_path);
    if (!rc) rc = yc_network_finalize(net);
  }
  if (rc) {
    yc_network_free(net);
    return rc;
  }
  *out = net;
  return YC_OK;
}

extern "C" int yc_network_load_text(yc_network** out, const char* cfg_text, const char* weights_path, int batch,
                                    int flags) {
  YC_REQUIRE(cfg_text, "yc_network_load_text: null cfg");
  return load_common(out, cfg_text, weights_path, batch, flags);
}

extern "C" int yc_network_load(yc_network** out, const char* cfg_path, const char* weights_path, int batch, int flags) {
  YC_REQUIRE(cfg_path, "yc_network_load: null cfg path");
  std::ifstream f(cfg_path);
  if (!f) {
    set_error("Couldn't open file: %s", cfg_path);  // file_error utils.cpp
    return YC_ERR_IO;
  }
  std::stringstream ss;
  ss << f.rdbuf();
  return load_common(out, ss.str(), weights_path, batch, flags);
}

extern "C" int yc_network_free(yc_network* net) {
  if (!net) return YC_OK;
  free_device_state(net);
  if (net->stream) cudaStreamDestroy(net->stream);
  delete net;
  return YC_OK;
}

int yc_train_pull_params(yc_network* net);  // network_train.cu

// ResizeNetwork (network.cpp:255-383; detector.cpp:150-182 calls it every ten iterations when a yolo layer has
// random=1): every layer's shape is re-derived for a new input size, the parameters stay.  The reference resizes layer
// by layer and reallocates each output; here the shapes are re-derived from the cfg text with the new size, the
// spatially independent state (parameters, rolling statistics, the momentum-carrying gradient buckets) takes a round
// trip through host memory, and plans / activation buffers / graph are rebuilt by yc_network_finalize.  A data-parallel
// context survives (it holds the communicator and per-layer events, no buffer pointers); every rank must resize alike.
extern "C" int yc_network_resize(yc_network* net, int w, int h) {
  YC_REQUIRE(net && w > 0 && h > 0, "yc_network_resize: bad argument");
  YC_REQUIRE(!net->cfg_text.empty(), "yc_network_resize: the network was not built from a cfg");
  if (w == net->w && h == net->h) return YC_OK;
  const bool was_final = net->finalized, train = (net->flags & YC_NET_TRAIN) != 0;
  const size_t nl = net->layers.size();
  std::vector<std::vector<float>> buckets(nl), adam_m(nl), adam_v(nl);
  if (was_final && train) {
    int rc = yc_train_pull_params(net);
    if (rc) return rc;
    for (size_t i = 0; i < nl; ++i) {
      const NLayer& L = net->layers[i];
      if (L.type != L_CONV || !L.dw_packed || !L.bucket_floats) continue;
      buckets[i].resize(L.bucket_floats);
      NET_CUDA(cudaMemcpyAsync(buckets[i].data(), L.dw_packed, L.bucket_floats * 4, cudaMemcpyDeviceToHost, net->stream));
      if (L.adam_m && L.adam_v) {  // Adam moments travel like the momentum carried in the bucket
        adam_m[i].resize(L.bucket_floats);
        adam_v[i].resize(L.bucket_floats);
        NET_CUDA(cudaMemcpyAsync(adam_m[i].data(), L.adam_m, L.bucket_floats * 4, cudaMemcpyDeviceToHost, net->stream));
        NET_CUDA(cudaMemcpyAsync(adam_v[i].data(), L.adam_v, L.bucket_floats * 4, cudaMemcpyDeviceToHost, net->stream));
      }
    }
    NET_CUDA(cudaStreamSynchronize(net->stream));
  }
  void* dp = net->dp;
  net->dp = nullptr;  // yc_train_free would drop the communicator
  free_device_state(net);
  // scalars that the host may have changed since the cfg was parsed
  const float lr = net->lr, momentum = net->momentum, decay = net->decay;
  const int max_iter = net->max_iter, batch = net->batch, old_w = net->w, old_h = net->h;
  std::vector<NLayer> old = std::move(net->layers);
  const std::vector<int> old_yolo = net->yolo_layers;
  net->layers.clear();
  net->yolo_layers.clear();
  net->n_conv = 0;
  int rc = build_layers(net, net->cfg_text, batch, w, h);
  if (!rc && net->layers.size() != nl) { set_error("yc_network_resize: layer count changed"); rc = YC_ERR_ARG; }
  if (rc) {  // e.g. a size at which some layer's output is empty: the network keeps its old shape (not finalized)
    net->layers = std::move(old);
    net->yolo_layers = old_yolo;
    net->n_conv = 0;
    for (const NLayer& L : net->layers) net->n_conv += L.type == L_CONV;
    net->w = old_w; net->h = old_h;
    net->dp = dp;
    net->lr = lr; net->momentum = momentum; net->decay = decay; net->max_iter = max_iter;
    return rc;
  }
  net->lr = lr; net->momentum = momentum; net->decay = decay; net->max_iter = max_iter;
  for (size_t i = 0; i < nl; ++i) {
    NLayer& L = net->layers[i];
    NLayer& O = old[i];
    if (L.type != L_CONV) continue;
    L.weights.swap(O.weights); L.biases.swap(O.biases); L.scales.swap(O.scales);
    L.rmean.swap(O.rmean); L.rvar.swap(O.rvar);
  }
  net->dp = dp;
  if (!was_final) return YC_OK;
  rc = yc_network_finalize(net);
  if (rc) return rc;
  if (train) {
    for (size_t i = 0; i < nl; ++i) {
      NLayer& L = net->layers[i];
      if (buckets[i].empty()) continue;
      if (!L.dw_packed || L.bucket_floats != buckets[i].size()) {
        set_error("yc_network_resize: layer %zu changed its gradient bucket", i);
        return YC_ERR_ARG;
      }
      NET_CUDA(cudaMemcpyAsync(L.dw_packed, buckets[i].data(), L.bucket_floats * 4, cudaMemcpyHostToDevice, net->stream));
      if (!adam_m[i].empty() && L.adam_m && L.adam_v) {
        NET_CUDA(cudaMemcpyAsync(L.adam_m, adam_m[i].data(), L.bucket_floats * 4, cudaMemcpyHostToDevice, net->stream));
        NET_CUDA(cudaMemcpyAsync(L.adam_v, adam_v[i].data(), L.bucket_floats * 4, cudaMemcpyHostToDevice, net->stream));
      }
    }
    NET_CUDA(cudaStreamSynchronize(net->stream));
  }
  return YC_OK;
}

extern "C" int yc_network_set_conv_weights(yc_network* net, int layer, const float* weights, const float* biases,
                                           const float* scales, const float* rmean, const float* rvar) {
  YC_REQUIRE(net && layer >= 0 && layer < (int)net->layers.size(), "yc_network_set_conv_weights: bad layer");
  NLayer& L = net->layers[layer];
  YC_REQUIRE(L.type == L_CONV, "yc_network_set_conv_weights: layer %d is not convolutional", layer);
  if (weights) memcpy(L.weights.data(), weights, L.weights.size() * 4);
  if (biases) memcpy(L.biases.data(), biases, L.n * 4);
  if (L.bn) {
    if (scales) memcpy(L.scales.data(), scales, L.n * 4);
    if (rmean) memcpy(L.rmean.data(), rmean, L.n * 4);
    if (rvar) memcpy(L.rvar.data(), rvar, L.n * 4);
  }
  net->finalized = false;
  return YC_OK;
}


// ---- weights on disk / initialisation / learning-rate schedule ---------------------------------------------------
// In a training network the current values live on the device (master weights in fp32): bring them back into the
// host-side copies the loader / saver / finalize work from (PullConvolutionalLayer, convolutional_kernels.cu:817-836).
int yc_train_pull_params(yc_network* net);  // network_train.cu

// LoadWeights (parser.cpp:1846-1850) on an already parsed network; call yc_network_finalize after.
extern "C" int yc_network_load_weights(yc_network* net, const char* path) {
  YC_REQUIRE(net && path, "yc_network_load_weights: null argument");
  int rc = load_weights_file(net, path);
  net->finalized = false;
  return rc;
}

// SaveWeightsUpto (parser.cpp:1590-1643) + SaveConvolutionalWeights (:1545-1560): header major, minor, revision (int32)
// and seen (uint64), then per conv layer biases, [scales, rolling_mean, rolling_variance], weights (OIHW fp32).
extern "C" int yc_network_save_weights(yc_network* net, const char* path, int cutoff) {
  YC_REQUIRE(net && path, "yc_network_save_weights: null argument");
  if ((net->flags & YC_NET_TRAIN) && net->finalized) {
    int rc = yc_train_pull_params(net);
    if (rc) return rc;
  }
  FILE* fp = fopen(path, "wb");
  if (!fp) NET_FAIL(YC_ERR_IO, "cannot open %s for writing", path);
  const int hdr[3] = {0, 2, 5};  // MAJOR_VERSION, MINOR_VERSION, PATCH_VERSION (parser.cpp:1603-1609)
  const uint64_t seen = (uint64_t)net->seen;
  bool ok = fwrite(hdr, sizeof(int), 3, fp) == 3 && fwrite(&seen, sizeof(uint64_t), 1, fp) == 1;
  const int n = cutoff > 0 ? std::min<int>(cutoff, (int)net->layers.size()) : (int)net->layers.size();
  for (int i = 0; ok && i < n; ++i) {
    const NLayer& L = net->layers[i];
    if (L.type != L_CONV) continue;
    ok = fwrite(L.biases.data(), 4, L.n, fp) == (size_t)L.n;
    if (ok && L.bn)
      ok = fwrite(L.scales.data(), 4, L.n, fp) == (size_t)L.n && fwrite(L.rmean.data(), 4, L.n, fp) == (size_t)L.n &&
           fwrite(L.rvar.data(), 4, L.n, fp) == (size_t)L.n;
    if (ok) ok = fwrite(L.weights.data(), 4, L.weights.size(), fp) == L.weights.size();
  }
  ok = (fclose(fp) == 0) && ok;
  if (!ok) NET_FAIL(YC_ERR_IO, "short write to %s", path);
  return YC_OK;
}

// FillConvLayer's initialisation (convolutional_layer.cpp:478-493): W = sqrt(2 / (k*k*c)) * RandUniform(-1, 1) with
// RandUniform = (float)rand() / RAND_MAX * 2 - 1 (utils.cpp:368-383), biases 0, scales 1, rolling statistics 0.  With
// srand(seed) and the same layer order the values are the reference's own, bit for bit, for every conv layer parsed
// before the first [yolo] section (FillYoloLayer re-seeds from the clock, yolo_layer.cpp:86).  first_layer: layers
// below it keep their values (what loading a backbone file and training the rest does).
extern "C" int yc_network_init_weights(yc_network* net, unsigned seed, int first_layer) {
  YC_REQUIRE(net, "yc_network_init_weights: null network");
  srand(seed);
  for (size_t i = 0; i < net->layers.size(); ++i) {
    NLayer& L = net->layers[i];
    if (L.type != L_CONV) continue;
    const float scale = (float)sqrt(2. / (L.size * L.size * L.c));
    if ((int)i < first_layer) {
      for (size_t k = 0; k < L.weights.size(); ++k) (void)rand();  // keep the sequence aligned with a full initialisation
      continue;
    }
    for (size_t k = 0; k < L.weights.size(); ++k) L.weights[k] = scale * (((float)rand() / RAND_MAX * 2.f) + -1.f);
    std::fill(L.biases.begin(), L.biases.end(), 0.f);
    if (L.bn) {
      std::fill(L.scales.begin(), L.scales.end(), 1.f);
      std::fill(L.rmean.begin(), L.rmean.end(), 0.f);
      std::fill(L.rvar.begin(), L.rvar.end(), 0.f);
    }
  }
  net->finalized = false;
  return YC_OK;
}

// GetCurrLr (network.cpp:32-84).  max_iter <= 0 takes the value set by yc_network_set_schedule / [net] max_batches.
extern "C" float yc_network_current_rate(const yc_network* net, long long iter, int max_iter) {
  if (!net) return 0.f;
  if (max_iter <= 0) max_iter = net->max_iter;
  if (iter < net->burn_in) return net->lr * (float)pow((float)iter / net->burn_in, net->power);
  switch (net->policy) {
    case 0: return net->lr;
    case 1: return net->lr * (float)pow(net->scale, iter / net->step);
    case 4: {  // STEPS: the step positions are FRACTIONS of max_iter in this reference (network.cpp:49-58)
      float lr = net->lr;
      for (size_t i = 0; i < net->steps.size(); ++i) {
        if (max_iter * net->steps[i] > iter) return lr;
        lr *= net->scales[i];
      }
      return lr;
    }
    case 2: return net->lr * (float)pow(net->gamma, iter);
    case 3: return net->lr * (float)pow(1 - (float)iter / max_iter, net->power);
    case 6: return net->lr;  // RANDOM draws from rand(): not reproducible; the base rate is returned
    case 5: return (float)(net->lr * (1. / (1. + exp(net->gamma * (iter - net->step)))));
    case 7: {
      int last_iter = 0;
      int cycle = net->sgdr_cycle > 0 ? net->sgdr_cycle : max_iter;  // parser.cpp:957 (a zero cycle would never end)
      if (cycle <= 0) return net->lr;
      while (last_iter + cycle < iter) {
        last_iter += cycle;
        cycle *= net->sgdr_mult;
      }
      return net->lr_min + 0.5f * (net->lr - net->lr_min) * (1.0f + (float)cos((float)(iter - last_iter) * M_PI / cycle));
    }
  }
  return net->lr;
}

// net->max_iter and net->curr_iter of the reference are owned by its training driver (detector.cpp); here the caller
// sets them.  iter < 0 keeps the current iteration count (it advances by one per yc_network_update).
extern "C" int yc_network_set_schedule(yc_network* net, int max_iter, long long iter) {
  YC_REQUIRE(net, "yc_network_set_schedule: null network");
  if (max_iter > 0) net->max_iter = max_iter;
  if (iter >= 0) net->curr_iter = iter;
  return YC_OK;
}

extern "C" int yc_network_info(const yc_network* net, int what) {
  if (!net) return -1;
  switch (what) {
    case 0: return (int)net->layers.size();
    case 1: return net->batch;
    case 2: return net->w;
    case 3: return net->h;
    case 4: return net->c;
    case 5: return net->n_conv;
    case 6: return (int)net->ops.size();
    case 7: return (int)net->yolo_layers.size();
    case 8: return net->subdivisions;
    case 9: return (int)std::min<long long>(net->curr_iter, 0x7fffffff);
    case 10: return (int)std::min<long long>(net->seen, 0x7fffffff);
    case 11: return net->layers_loaded;
    case 12: return net->flags;
  }
  return -1;
}

extern "C" int yc_network_layer_info(const yc_network* net, int layer, int what) {
  if (!net || layer < 0 || layer >= (int)net->layers.size()) return -1;
  const NLayer& L = net->layers[layer];
  switch (what) {
    case 0: return L.type;
    case 1: return L.out_c;
    case 2: return L.out_h;
    case 3: return L.out_w;
    case 4: return (L.fused_into >= 0 || L.head_decoded) ? 0 : 1;
    case 5: return L.c;
    case 6: return L.size;
    case 7: return L.stride;
    case 8: return L.act;
    case 9: return L.bn;
    case 10: return L.plan ? yc_conv_plan_info(L.plan, 0) : -1;
    case 11: return L.plan ? yc_conv_plan_info(L.plan, 2) : -1;
    case 12: return L.plan ? yc_conv_plan_info(L.plan, 3) : -1;
    case 13: return L.plan ? yc_conv_plan_info(L.plan, 7) : (L.pair_first >= 0 ? yc_conv_plan_info(net->layers[L.pair_first].plan, 7) : -1);
    case 14: return L.pair_first;   // >= 0: this conv is computed by that layer's kernel (fused 1x1 pair)
    case 15: return L.pair_second;
    case 16: return L.classes;
    case 17: return L.total;
    case 18: return L.nms_kind;
    case 19: return L.pad;
    case 20: return L.dilation;
    case 21: return L.groups;
    case 22: return L.group_id;
    case 23: return L.from;
    case 24: return L.n;
    case 25: return (int)L.inputs.size();
    case 26: return L.max_boxes;
  }
  return -1;
}

extern "C" int yc_network_layer_input(const yc_network* net, int layer, int k) {
  if (!net || layer < 0 || layer >= (int)net->layers.size()) return -1;
  const NLayer& L = net->layers[layer];
  return (k >= 0 && k < (int)L.inputs.size()) ? L.inputs[k] : -1;
}

extern "C" float yc_network_layer_float(const yc_network* net, int layer, int what) {
  if (!net || layer < 0 || layer >= (int)net->layers.size()) return 0.f;
  const NLayer& L = net->layers[layer];
  switch (what) {
    case 0: return L.scale_x_y;
    case 1: return L.beta_nms;
    case 2: return L.ignore_thresh;
    case 3: return L.truth_thresh;
    case 4: return L.iou_thresh;
    case 5: return L.iou_normalizer;
    case 6: return L.cls_normalizer;
    case 7: return L.label_smooth_eps;
    case 8: return L.max_delta;
    case 9: return L.uc_normalizer;
  }
  return 0.f;
}

extern "C" int yc_network_layer_array(const yc_network* net, int layer, int what, float* dst, int capacity) {
  YC_REQUIRE(net && dst && layer >= 0 && layer < (int)net->layers.size(), "yc_network_layer_array: bad argument");
  const NLayer& L = net->layers[layer];
  int n = 0;
  if (what == 0) {
    n = (int)L.anchors.size();
    YC_REQUIRE(n <= capacity, "yc_network_layer_array: capacity %d < %d", capacity, n);
    for (int i = 0; i < n; ++i) dst[i] = L.anchors[i];
  } else if (what == 1) {
    n = (int)L.mask.size();
    YC_REQUIRE(n <= capacity, "yc_network_layer_array: capacity %d < %d", capacity, n);
    for (int i = 0; i < n; ++i) dst[i] = (float)L.mask[i];
  } else {
    YC_REQUIRE(false, "yc_network_layer_array: unknown array %d", what);
  }
  return n;
}

extern "C" double yc_network_flops(const yc_network* net) {
  double f = 0;
  if (!net) return 0;
  for (const NLayer& L : net->layers)
    if (L.type == L_CONV) f += 2.0 * (double)L.n * L.c * L.size * L.size * L.out_h * L.out_w;
  return f;
}

static int ensure_graph(yc_network* net, const float* input_dev) {
  if (net->graph_exec && net->graph_input == input_dev) return YC_OK;
  if (net->graph_exec) { cudaGraphExecDestroy(net->graph_exec); net->graph_exec = nullptr; }
  if (net->graph) { cudaGraphDestroy(net->graph); net->graph = nullptr; }
  net->graph_input = input_dev;
  NET_CUDA(cudaStreamBeginCapture(net->stream, cudaStreamCaptureModeThreadLocal));
  int rc = run_ops(net, net->stream);
  cudaError_t e = cudaStreamEndCapture(net->stream, &net->graph);
  if (rc) return rc;
  if (e != cudaSuccess) {
    set_error("cudaStreamEndCapture: %s", cudaGetErrorString(e));
    return YC_ERR_CUDA;
  }
  NET_CUDA(cudaGraphInstantiate(&net->graph_exec, net->graph, 0));
  return YC_OK;
}

extern "C" int yc_network_forward(yc_network* net, const float* input_dev, void* stream) {
  YC_REQUIRE(net && input_dev, "yc_network_forward: null argument");
  YC_REQUIRE(!(net->flags & YC_NET_TRAIN), "yc_network_forward: network was loaded for training (use yc_network_forward_backward)");
  if (!net->finalized) {
    int rc = yc_network_finalize(net);
    if (rc) return rc;
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (net->flags & YC_NET_CUDA_GRAPH) {
    int rc = ensure_graph(net, input_dev);
    if (rc) return rc;
    NET_CUDA(cudaGraphLaunch(net->graph_exec, st));
    count_launch((int)net->ops.size());
    return YC_OK;
  }
  net->graph_input = input_dev;
  return run_ops(net, st);
}

static float* predict_tail(yc_network* net);  // forward on net->input_dev + D2H of the heads

// staging buffer -> the graph's input, on the network's stream, once the H2D on the copy stream has landed (~0.1 ms D2D)
static int consume_staged(yc_network* net) {
  const size_t in_bytes = (size_t)net->batch * net->c * net->h * net->w * 4;
  NET_CUDA(cudaStreamWaitEvent(net->stream, net->copy_done, 0));
  NET_CUDA(cudaMemcpyAsync(net->input_dev, net->input_dev2, in_bytes, cudaMemcpyDeviceToDevice, net->stream));
  NET_CUDA(cudaEventRecord(net->stage_free, net->stream));
  net->resident_host = net->staged_host;
  net->staged_host = nullptr;
  return YC_OK;
}

extern "C" float* yc_network_predict(yc_network* net, const float* input_host) {
  if (!net || !input_host) {
    set_error("yc_network_predict: null argument");
    return nullptr;
  }
  if (!net->finalized && yc_network_finalize(net)) return nullptr;
  const size_t in_bytes = (size_t)net->batch * net->c * net->h * net->w * 4;
  // network_kernels.cu:513-514: memcpy into the pinned staging buffer, then H2D
  // (a caller that produced its batch directly in yc_network_input_buffer() skips the staging copy)
  if (net->resident_host == input_host) {  // announced two calls ago and already moved into the graph's input
    net->resident_host = nullptr;
    return predict_tail(net);
  }
  if (net->staged_host == input_host) {  // announced with yc_network_prefetch: its H2D ran on the copy stream
    if (consume_staged(net)) return nullptr;
    net->resident_host = nullptr;
    return predict_tail(net);
  }
  net->resident_host = nullptr;
  if (input_host != net->input_pinned) memcpy(net->input_pinned, input_host, in_bytes);
  if (cudaMemcpyAsync(net->input_dev, net->input_pinned, in_bytes, cudaMemcpyHostToDevice, net->stream) != cudaSuccess) {
    set_error("yc_network_predict: H2D failed: %s", cudaGetErrorString(cudaGetLastError()));
    return nullptr;
  }
  return predict_tail(net);
}

// Input double buffering for a stream of batches: call yc_network_prefetch(next) right BEFORE yc_network_predict(current).
// The H2D copy of `next` runs on a separate stream (copy engine) into a device staging buffer while the forward of
// `current` computes; a later yc_network_predict(net, next) finds it on the device.  input_host should be yc_network_input_buffer_slot(net, 0|1)
// (pinned); any other pointer is first copied into slot 1.
static int ensure_prefetch(yc_network* net) {
  const size_t in_bytes = (size_t)net->batch * net->c * net->h * net->w * 4;
  if (!net->input_pinned2) NET_CUDA(cudaHostAlloc(&net->input_pinned2, in_bytes, cudaHostAllocDefault));
  if (!net->input_dev2) NET_CUDA(cudaMalloc(&net->input_dev2, in_bytes));
  if (!net->copy_stream) NET_CUDA(cudaStreamCreateWithFlags(&net->copy_stream, cudaStreamNonBlocking));
  if (!net->copy_done) NET_CUDA(cudaEventCreateWithFlags(&net->copy_done, cudaEventDisableTiming));
  if (!net->stage_free) {
    NET_CUDA(cudaEventCreateWithFlags(&net->stage_free, cudaEventDisableTiming));
    NET_CUDA(cudaEventRecord(net->stage_free, net->stream));
  }
  return YC_OK;
}
extern "C" float* yc_network_input_buffer_slot(yc_network* net, int slot) {
  if (!net || slot < 0 || slot > 1) return nullptr;
  if (!net->finalized && yc_network_finalize(net)) return nullptr;
  if (slot == 0) return net->input_pinned;
  if (ensure_prefetch(net)) return nullptr;
  return net->input_pinned2;
}
extern "C" int yc_network_prefetch(yc_network* net, const float* input_host) {
  YC_REQUIRE(net && input_host, "yc_network_prefetch: null argument");
  if (!net->finalized) {
    int rc = yc_network_finalize(net);
    if (rc) return rc;
  }
  int rc = ensure_prefetch(net);
  if (rc) return rc;
  const size_t in_bytes = (size_t)net->batch * net->c * net->h * net->w * 4;
  const float* src = input_host;
  if (input_host != net->input_pinned && input_host != net->input_pinned2) {
    memcpy(net->input_pinned2, input_host, in_bytes);
    src = net->input_pinned2;
  }
  // a batch staged earlier and not consumed yet moves into the graph's input first (the previous forward has finished:
  // yc_network_predict is synchronous), so that the staging buffer is free for this copy
  if (net->staged_host) {
    rc = consume_staged(net);
    if (rc) return rc;
  }
  NET_CUDA(cudaStreamWaitEvent(net->copy_stream, net->stage_free, 0));
  NET_CUDA(cudaMemcpyAsync(net->input_dev2, src, in_bytes, cudaMemcpyHostToDevice, net->copy_stream));
  NET_CUDA(cudaEventRecord(net->copy_done, net->copy_stream));
  net->staged_host = input_host;
  return YC_OK;
}

// Mat2Image + NetworkPredict (yolo_core.cpp:110-112, visualize.cpp:26-54) for a batch of u8 HWC images already at the
// network's resolution: H2D of the BYTES (4x less than fp32), /255 and HWC->NCHW on the device, then the same forward.
static int ensure_u8(yc_network* net) {
  const size_t bytes = (size_t)net->batch * net->c * net->h * net->w;
  if (!net->u8_pinned) NET_CUDA(cudaHostAlloc(&net->u8_pinned, bytes, cudaHostAllocDefault));
  if (!net->u8_dev) NET_CUDA(cudaMalloc(&net->u8_dev, bytes));
  return YC_OK;
}
extern "C" unsigned char* yc_network_input_buffer_u8(yc_network* net) {
  if (!net || (!net->finalized && yc_network_finalize(net)) || ensure_u8(net)) return nullptr;
  return net->u8_pinned;
}
extern "C" float* yc_network_predict_u8(yc_network* net, const unsigned char* input_host_hwc, int swap_rb) {
  if (!net || !input_host_hwc) {
    set_error("yc_network_predict_u8: null argument");
    return nullptr;
  }
  if (!net->finalized && yc_network_finalize(net)) return nullptr;
  if (ensure_u8(net)) return nullptr;
  const size_t bytes = (size_t)net->batch * net->c * net->h * net->w;
  if (input_host_hwc != net->u8_pinned) memcpy(net->u8_pinned, input_host_hwc, bytes);
  if (cudaMemcpyAsync(net->u8_dev, net->u8_pinned, bytes, cudaMemcpyHostToDevice, net->stream) != cudaSuccess) {
    set_error("yc_network_predict_u8: H2D failed: %s", cudaGetErrorString(cudaGetLastError()));
    return nullptr;
  }
  if (yc_u8hwc_to_nchw(net->u8_dev, net->batch, net->c, net->h, net->w, swap_rb, net->input_dev, net->stream)) return nullptr;
  return predict_tail(net);
}

// ProcImage (yolo_core.cpp:104-112) from the camera frame on: cv::resize to the network's size, cvtColor(RGB2BGR),
// Mat2Image, NetworkPredict — the H2D copy carries the frames as they are (u8, any size), the rest runs on the device.
extern "C" float* yc_network_predict_u8_resized(yc_network* net, const unsigned char* frames_host_hwc, int src_h, int src_w,
                                                int swap_rb) {
  if (!net || !frames_host_hwc || src_h <= 0 || src_w <= 0) {
    set_error("yc_network_predict_u8_resized: bad arguments");
    return nullptr;
  }
  if (!net->finalized && yc_network_finalize(net)) return nullptr;
  if (src_h == net->h && src_w == net->w) return yc_network_predict_u8(net, frames_host_hwc, swap_rb);
  const size_t bytes = (size_t)net->batch * net->c * src_h * src_w;
  if (bytes > net->u8_src_bytes) {  // staging for frames of this size (grown, never shrunk)
    if (cudaStreamSynchronize(net->stream) != cudaSuccess) return nullptr;
    if (net->u8_src_pinned) cudaFreeHost(net->u8_src_pinned);
    if (net->u8_src_dev) cudaFree(net->u8_src_dev);
    net->u8_src_pinned = nullptr; net->u8_src_dev = nullptr; net->u8_src_bytes = 0;
    if (cudaHostAlloc(&net->u8_src_pinned, bytes, cudaHostAllocDefault) != cudaSuccess ||
        cudaMalloc(&net->u8_src_dev, bytes) != cudaSuccess) {
      set_error("yc_network_predict_u8_resized: staging allocation of %zu bytes failed: %s", bytes, cudaGetErrorString(cudaGetLastError()));
      return nullptr;
    }
    net->u8_src_bytes = bytes;
  }
  memcpy(net->u8_src_pinned, frames_host_hwc, bytes);
  if (cudaMemcpyAsync(net->u8_src_dev, net->u8_src_pinned, bytes, cudaMemcpyHostToDevice, net->stream) != cudaSuccess) {
    set_error("yc_network_predict_u8_resized: H2D failed: %s", cudaGetErrorString(cudaGetLastError()));
    return nullptr;
  }
  if (yc_resize_u8_to_nchw(net->u8_src_dev, net->batch, net->c, src_h, src_w, swap_rb, net->input_dev, net->h, net->w, net->stream))
    return nullptr;
  return predict_tail(net);
}

static float* predict_tail(yc_network* net) {
  if (yc_network_forward(net, net->input_dev, net->stream)) return nullptr;
  float* last = nullptr;
  if (net->flags & YC_NET_PULL_HEADS) {
    for (int i : net->yolo_layers) {  // yolo_layer.cpp:856 cuda_pull_array_async(l->output_gpu, l->output, ...)
      NLayer& L = net->layers[i];
      const size_t bytes = (size_t)net->batch * L.out_c * L.out_h * L.out_w * 4;
      if (cudaMemcpyAsync(L.dec_host, L.dec_dev, bytes, cudaMemcpyDeviceToHost, net->stream) != cudaSuccess) {
        set_error("yc_network_predict: D2H failed: %s", cudaGetErrorString(cudaGetLastError()));
        return nullptr;
      }
    }
  }
  if (cudaStreamSynchronize(net->stream) != cudaSuccess) {
    set_error("yc_network_predict: %s", cudaGetErrorString(cudaGetLastError()));
    return nullptr;
  }
  // GetNetworkOutput (network.cpp:390-398): output of the last layer
  const NLayer& Z = net->layers.back();
  if (Z.type == L_YOLO || Z.type == L_GYOLO) last = Z.dec_host;
  if (!last) last = net->input_pinned;  // non-YOLO tail: callers use yc_network_layer_output
  return last;
}

extern "C" float* yc_network_input_buffer(yc_network* net) {
  if (!net) return nullptr;
  if (!net->finalized && yc_network_finalize(net)) return nullptr;
  return net->input_pinned;
}

extern "C" const float* yc_network_yolo_output(yc_network* net, int layer) {
  if (!net || layer < 0 || layer >= (int)net->layers.size()) return nullptr;
  return net->layers[layer].dec_host;
}

extern "C" int yc_network_layer_output(yc_network* net, int layer, float* dst_host) {
  YC_REQUIRE(net && dst_host && layer >= 0 && layer < (int)net->layers.size(), "yc_network_layer_output: bad argument");
  YC_REQUIRE(net->finalized, "yc_network_layer_output: network not finalized");
  NLayer& L = net->layers[layer];
  YC_REQUIRE(!L.head_decoded, "yc_network_layer_output: layer %d is a head convolution whose results go, decoded, straight "
             "into the [yolo] layer after it (yc_network_layer_info(layer, 4) == 0; YC_PLAN_HEAD_DECODE)",
             layer);
  const size_t elems = (size_t)net->batch * L.out_c * L.out_h * L.out_w;
  if (L.type == L_YOLO || L.type == L_GYOLO) {
    NET_CUDA(cudaMemcpyAsync(dst_host, L.dec_dev, elems * 4, cudaMemcpyDeviceToHost, net->stream));
    NET_CUDA(cudaStreamSynchronize(net->stream));
    return YC_OK;
  }
  YC_REQUIRE(L.fused_into < 0, "layer %d has no output of its own: it is fused with shortcut %d (load without YC_NET_FUSE_SHORTCUT)",
             layer, L.fused_into);
  yc_tensor v = view_of(net, layer);
  float* tmp = nullptr;
  NET_CUDA(cudaMalloc(&tmp, elems * 4));
  int rc = yc_nhwc_to_nchw(&v, tmp, net->stream);
  if (!rc) {
    cudaError_t e = cudaMemcpyAsync(dst_host, tmp, elems * 4, cudaMemcpyDeviceToHost, net->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(net->stream);
    if (e != cudaSuccess) { set_error("yc_network_layer_output: %s", cudaGetErrorString(e)); rc = YC_ERR_CUDA; }
  }
  cudaFree(tmp);
  return rc;
}

extern "C" int yc_network_profile(yc_network* net, const float* input_dev, int iters, float* ms_per_layer) {
  YC_REQUIRE(net && input_dev && ms_per_layer && iters > 0, "yc_network_profile: bad argument");
  if (!net->finalized) {
    int rc = yc_network_finalize(net);
    if (rc) return rc;
  }
  const int nl = (int)net->layers.size();
  for (int i = 0; i <= nl; ++i) ms_per_layer[i] = 0.f;
  net->graph_input = input_dev;
  const size_t nops = net->ops.size();
  std::vector<cudaEvent_t> ev(nops + 1);
  for (auto& e : ev) NET_CUDA(cudaEventCreate(&e));
  int rc = YC_OK;
  for (int it = 0; it < iters && !rc; ++it) {
    NET_CUDA(cudaEventRecord(ev[0], net->stream));
    for (size_t k = 0; k < nops && !rc; ++k) {
      rc = run_op(net, net->ops[k], net->stream);
      NET_CUDA(cudaEventRecord(ev[k + 1], net->stream));
    }
    NET_CUDA(cudaStreamSynchronize(net->stream));
    for (size_t k = 0; k < nops && !rc; ++k) {
      float ms = 0.f;
      NET_CUDA(cudaEventElapsedTime(&ms, ev[k], ev[k + 1]));
      const int li = net->ops[k].layer < 0 ? nl : net->ops[k].layer;
      ms_per_layer[li] += ms / iters;
    }
  }
  for (auto& e : ev) cudaEventDestroy(e);
  return rc;
}

// The forward's kernels of one class replayed back to back as a CUDA graph, the way they run inside the real step
// (yc_network_profile brackets every launch with events: each kernel then starts on an idle GPU and the stream drains in
// between, which costs the many short layers 5-10 us each).  which: 0 = the convolution launches only, 1 = every other
// kernel, 2 = all of them.  The buffers hold whatever the last forward left: the kernels' timing does not depend on the
// values.  ms_per_replay = mean over `iters` replays, CUDA events on the network's stream around the whole batch of
// replays.
extern "C" int yc_network_time_ops(yc_network* net, const float* input_dev, int which, int iters, float* ms_per_replay,
                                   int* launches) {
  YC_REQUIRE(net && input_dev && ms_per_replay && iters > 0 && which >= 0 && which <= 2, "yc_network_time_ops: bad argument");
  if (!net->finalized) {
    int rc = yc_network_finalize(net);
    if (rc) return rc;
  }
  net->graph_input = input_dev;
  cudaGraph_t g = nullptr;
  cudaGraphExec_t ge = nullptr;
  int n = 0;
  NET_CUDA(cudaStreamBeginCapture(net->stream, cudaStreamCaptureModeThreadLocal));
  int rc = YC_OK;
  for (size_t k = 0; k < net->ops.size() && !rc; ++k) {
    const bool conv = net->ops[k].kind == 0 || net->ops[k].kind == 8;
    if (which == 2 || (which == 0) == conv) {
      rc = run_op(net, net->ops[k], net->stream);
      ++n;
    }
  }
  cudaError_t e = cudaStreamEndCapture(net->stream, &g);
  if (rc) { if (g) cudaGraphDestroy(g); return rc; }
  if (e != cudaSuccess) NET_FAIL(YC_ERR_CUDA, "yc_network_time_ops: cudaStreamEndCapture: %s", cudaGetErrorString(e));
  if (launches) *launches = n;
  if (n == 0) { cudaGraphDestroy(g); *ms_per_replay = 0.f; return YC_OK; }
  NET_CUDA(cudaGraphInstantiate(&ge, g, 0));
  cudaEvent_t e0, e1;
  NET_CUDA(cudaEventCreate(&e0));
  NET_CUDA(cudaEventCreate(&e1));
  for (int i = 0; i < 2; ++i) NET_CUDA(cudaGraphLaunch(ge, net->stream));  // warm-up
  NET_CUDA(cudaEventRecord(e0, net->stream));
  for (int i = 0; i < iters; ++i) NET_CUDA(cudaGraphLaunch(ge, net->stream));
  NET_CUDA(cudaEventRecord(e1, net->stream));
  NET_CUDA(cudaStreamSynchronize(net->stream));
  float ms = 0.f;
  NET_CUDA(cudaEventElapsedTime(&ms, e0, e1));
  *ms_per_replay = ms / iters;
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaGraphExecDestroy(ge);
  cudaGraphDestroy(g);
  return YC_OK;
}

// ---- detections on the host: GetNetworkBoxes / NmsSort --------------------------------------------------
namespace {

inline int entry_index(const NLayer& L, int batch, int location, int entry, int E) {  // EntryIndex yolo_layer.cpp:380-386
  const int wh = L.out_w * L.out_h;
  const int n = location / wh, loc = location % wh;
  return batch * (L.out_c * wh) + n * wh * E + entry * wh + loc;
}

struct HBox { float x, y, w, h; };
inline float overlap1(float x1, float w1, float x2, float w2) {  // Box::Overlap box.cpp:36-46
  const float l1 = x1 - w1 / 2, l2 = x2 - w2 / 2;
  const float left = l1 > l2 ? l1 : l2;
  const float r1 = x1 + w1 / 2, r2 = x2 + w2 / 2;
  const float right = r1 < r2 ? r1 : r2;
  return right - left;
}
inline float box_intersect(const HBox& a, const HBox& b) {  // box.cpp:48-57
  const float w = overlap1(a.x, a.w, b.x, b.w), h = overlap1(a.y, a.h, b.y, b.h);
  if (w < 0 || h < 0) return 0;
  return w * h;
}
inline float box_iou(const HBox& a, const HBox& b) {  // box.cpp:59-72
  const float I = box_intersect(a, b);
  const float U = a.w * a.h + b.w * b.h - I;
  if (fabsf(I) < FLT_EPSILON || fabsf(U) < FLT_EPSILON) return 0;
  return I / U;
}
inline float box_diou(const HBox& a, const HBox& b, float beta) {  // Box::Diou box.cpp:97-115
  const float top = fminf(a.y - a.h / 2, b.y - b.h / 2), bot = fmaxf(a.y + a.h / 2, b.y + b.h / 2);
  const float left = fminf(a.x - a.w / 2, b.x - b.w / 2), right = fmaxf(a.x + a.w / 2, b.x + b.w / 2);
  const float w = right - left, h = bot - top;
  const float c = w * w + h * h;
  const float iou = box_iou(a, b);
  if (fabsf(c) < FLT_EPSILON) return iou;
  const float d = (a.x - b.x) * (a.x - b.x) + (a.y - b.y) * (a.y - b.y);
  const float diou_term = powf(d / c, beta);
  return iou - diou_term;
}

int nms_comparator(const void* pa, const void* pb) {  // NmsComparator box.cpp:372-389
  const yc_detection* a = (const yc_detection*)pa;
  const yc_detection* b = (const yc_detection*)pb;
  float diff = 0;
  if (b->sort_class >= 0) diff = a->prob[b->sort_class] - b->prob[b->sort_class];
  else diff = a->objectness - b->objectness;
  if (diff < 0) return 1;
  if (diff > 0) return -1;
  return 0;
}

}  // namespace

// skip_empty: classes nobody scored change nothing but the array order in the reference (every prob[k]==0 is
// skipped, box.cpp:405); the throughput path, whose rows do not move, skips their sort.
static void nms_sort_impl(yc_detection* dets, int total, int classes, float thresh, int nms_kind, float beta,
                          bool skip_empty) {
  for (int k = 0; k < classes; ++k) {
    if (skip_empty) {
      bool any = false;
      for (int i = 0; i < total && !any; ++i) any = fabsf(dets[i].prob[k]) >= FLT_EPSILON;
      if (!any) continue;
    }
    for (int i = 0; i < total; ++i) dets[i].sort_class = k;
    qsort(dets, total, sizeof(yc_detection), nms_comparator);
    for (int i = 0; i < total; ++i) {
      if (fabsf(dets[i].prob[k]) < FLT_EPSILON) continue;
      const HBox a{dets[i].x, dets[i].y, dets[i].w, dets[i].h};
      for (int j = i + 1; j < total; ++j) {
        const HBox b{dets[j].x, dets[j].y, dets[j].w, dets[j].h};
        if (nms_kind == 0 && box_iou(a, b) > thresh) dets[j].prob[k] = 0.f;
        else if (nms_kind == 1 && box_diou(a, b, beta) > thresh) dets[j].prob[k] = 0.f;
      }
    }
  }
}

extern "C" void yc_nms_sort(yc_detection* dets, int total, int classes, float thresh, int nms_kind, float beta) {
  nms_sort_impl(dets, total, classes, thresh, nms_kind, beta, false);
}

extern "C" yc_detection* yc_network_get_boxes(yc_network* net, int b, float thresh, int* num) {
  if (!net || b < 0 || b >= net->batch || !(net->flags & YC_NET_PULL_HEADS)) {
    set_error("yc_network_get_boxes: bad argument (needs YC_NET_PULL_HEADS)");
    return nullptr;
  }
  // NumDetections (network.cpp:432-448)
  int count = 0;
  int classes = 0;
  for (int li : net->yolo_layers) {
    const NLayer& L = net->layers[li];
    const bool g = L.type == L_GYOLO;
    const int E = (g ? 9 : 5) + L.classes, wh = L.out_w * L.out_h;
    classes = L.classes;
    for (int n = 0; n < L.n; ++n)
      for (int i = 0; i < wh; ++i)
        if (L.dec_host[entry_index(L, b, n * wh + i, g ? 8 : 4, E)] > thresh) ++count;
  }
  yc_detection* dets = (yc_detection*)calloc(count > 0 ? count : 1, sizeof(yc_detection));  // MakeNetworkBoxes :450-471
  for (int i = 0; i < count; ++i) dets[i].prob = (float*)calloc(classes, sizeof(float));
  int k = 0;
  for (int li : net->yolo_layers) {  // FillNetworkBoxes :473-496
    const NLayer& L = net->layers[li];
    const bool g = L.type == L_GYOLO;
    const int E = (g ? 9 : 5) + L.classes, wh = L.out_w * L.out_h;
    const float* pred = L.dec_host;
    for (int n = 0; n < L.n; ++n)
      for (int i = 0; i < wh; ++i) {
        const int loc = n * wh + i;
        const float obj = pred[entry_index(L, b, loc, g ? 8 : 4, E)];
        if (obj <= thresh) continue;
        const int bi = entry_index(L, b, loc, 0, E);
        const int col = i % L.out_w, row = i / L.out_w;
        yc_detection& d = dets[k++];
        const int m = L.mask[n];
        float scale = 1.f;
        if (g) {  // GetGaussianYoloBox gaussian_yolo_layer.cpp:151-160
          d.w = (float)(exp((double)pred[bi + 4 * wh]) * L.anchors[2 * m] / net->w);
          d.h = (float)(exp((double)pred[bi + 6 * wh]) * L.anchors[2 * m + 1] / net->h);
          d.x = (col + pred[bi]) / L.out_w;
          d.y = (row + pred[bi + 2 * wh]) / L.out_h;
          d.uc[0] = pred[bi + 1 * wh]; d.uc[1] = pred[bi + 3 * wh]; d.uc[2] = pred[bi + 5 * wh]; d.uc[3] = pred[bi + 7 * wh];
        } else {  // GetYoloBox yolo_layer.cpp:139-148
          d.x = (col + pred[bi]) / L.out_w;
          d.y = (row + pred[bi + wh]) / L.out_h;
          d.w = (float)(exp((double)pred[bi + 2 * wh]) * L.anchors[2 * m] / net->w);
          d.h = (float)(exp((double)pred[bi + 3 * wh]) * L.anchors[2 * m + 1] / net->h);
        }
        d.objectness = obj;
        d.classes = L.classes;
        for (int j = 0; j < L.classes; ++j) {
          const float cls = pred[entry_index(L, b, loc, (g ? 9 : 5) + j, E)];
          float prob;
          if (g) {
            const float uc_avg = (float)((d.uc[0] + d.uc[1] + d.uc[2] + d.uc[3]) / 4.0);
            prob = (float)(obj * cls * (1.0 - uc_avg));
          } else {
            prob = obj * cls;
          }
          d.prob[j] = (prob > thresh) ? prob : 0;
        }
        (void)scale;
      }
  }
  if (num) *num = count;
  return dets;
}

extern "C" void yc_free_detections(yc_detection* dets, int n) {
  if (!dets) return;
  for (int i = 0; i < n; ++i) free(dets[i].prob);
  free(dets);
}

extern "C" int yc_network_detect(yc_network* net, float thresh, float nms_thresh, const float** rows, int* counts,
                                 int* total) {
  YC_REQUIRE(net && rows && counts && total && net->finalized, "yc_network_detect: bad argument");
  const int heads = (int)net->yolo_layers.size();
  YC_REQUIRE(heads > 0, "yc_network_detect: network has no YOLO layer");
  const int classes = net->layers[net->yolo_layers[0]].classes;
  const int R = 7 + classes;
  const int B = net->batch;
  if (!net->det_regions) {
    int max_cells = 0;
    for (int li : net->yolo_layers) max_cells = std::max(max_cells, net->layers[li].n * net->layers[li].out_h * net->layers[li].out_w);
    net->det_cap = std::min(max_cells, 16384);  // rows kept per (head, image); grown to max_cells when an image has more
    net->det_dense_cap = B * 256;
    net->det_R = R;
    NET_CUDA(cudaMalloc(&net->det_regions, (size_t)heads * B * net->det_cap * R * 4));
    NET_CUDA(cudaMalloc(&net->det_counts, (size_t)heads * B * 4));
    NET_CUDA(cudaMalloc(&net->det_dense, (size_t)net->det_dense_cap * R * 4));
    NET_CUDA(cudaMalloc(&net->det_img_counts, (size_t)B * 4));
    NET_CUDA(cudaHostAlloc(&net->det_host, (size_t)net->det_dense_cap * R * 4, cudaHostAllocDefault));
    NET_CUDA(cudaHostAlloc(&net->det_img_counts_host, (size_t)(heads + 1) * B * 4, cudaHostAllocDefault));
  }
  cudaStream_t st = net->stream;
  int rc = YC_OK;
  long long tot = 0;
  for (int attempt = 0; attempt < 2; ++attempt) {
    for (int hd = 0; hd < heads; ++hd) {
      const int li = net->yolo_layers[hd];
      NLayer& L = net->layers[li];
      rc = launch_yolo_extract(L.dec_dev, B, L.out_h, L.out_w, L.n, L.classes, L.anchors_dev, L.mask_dev, net->w, net->h,
                               thresh, L.type == L_GYOLO, net->det_regions + (size_t)hd * B * net->det_cap * R,
                               net->det_cap, net->det_counts + (size_t)hd * B, R, (float)li, st);
      if (rc) return rc;
    }
    rc = launch_yolo_gather(net->det_regions, net->det_counts, heads, B, net->det_cap, R, net->det_dense,
                            net->det_img_counts, net->det_dense_cap, st);
    if (rc) return rc;
    NET_CUDA(cudaMemcpyAsync(net->det_img_counts_host, net->det_img_counts, (size_t)B * 4, cudaMemcpyDeviceToHost, st));
    NET_CUDA(cudaMemcpyAsync(net->det_img_counts_host + B, net->det_counts, (size_t)heads * B * 4, cudaMemcpyDeviceToHost, st));
    NET_CUDA(cudaStreamSynchronize(st));
    tot = 0;
    for (int b = 0; b < B; ++b) tot += net->det_img_counts_host[b];
    int worst = 0;
    for (int i = 0; i < heads * B; ++i) worst = std::max(worst, net->det_img_counts_host[B + i]);
    if (worst <= net->det_cap) break;
    // an image has more candidates in one head than the per-head regions hold (the reference has no limit: GetNetworkBoxes
    // sizes its array from the count, network.cpp:432-470): regrow them to the number of cells, which no count can
    // exceed, and extract again
    int max_cells = 0;
    for (int li : net->yolo_layers) max_cells = std::max(max_cells, net->layers[li].n * net->layers[li].out_h * net->layers[li].out_w);
    YC_REQUIRE(attempt == 0 && max_cells > net->det_cap, "yc_network_detect: %d candidates in one head exceed its %d cells", worst, max_cells);
    cudaFree(net->det_regions);
    net->det_regions = nullptr;
    net->det_cap = max_cells;
    NET_CUDA(cudaMalloc(&net->det_regions, (size_t)heads * B * net->det_cap * R * 4));
  }
  if (tot > net->det_dense_cap) {  // grow the dense buffers and gather again (the per-head regions are intact)
    cudaFree(net->det_dense);
    cudaFreeHost(net->det_host);
    net->det_dense = nullptr; net->det_host = nullptr;
    net->det_dense_cap = (int)std::min<long long>(tot + tot / 2, (long long)heads * B * net->det_cap);
    NET_CUDA(cudaMalloc(&net->det_dense, (size_t)net->det_dense_cap * R * 4));
    NET_CUDA(cudaHostAlloc(&net->det_host, (size_t)net->det_dense_cap * R * 4, cudaHostAllocDefault));
    rc = launch_yolo_gather(net->det_regions, net->det_counts, heads, B, net->det_cap, R, net->det_dense,
                            net->det_img_counts, net->det_dense_cap, st);
    if (rc) return rc;
  }
  if (tot > 0) {
    NET_CUDA(cudaMemcpyAsync(net->det_host, net->det_dense, (size_t)tot * R * 4, cudaMemcpyDeviceToHost, st));
    NET_CUDA(cudaStreamSynchronize(st));
  }
  for (int b = 0; b < B; ++b) counts[b] = net->det_img_counts_host[b];
  *total = (int)tot;
  *rows = net->det_host;
  if (nms_thresh > 0 && tot > 0) {
    const NLayer& L0 = net->layers[net->yolo_layers[0]];
    const int nms_kind = L0.nms_kind;
    const float beta = L0.beta_nms;
    std::vector<int> offs(B + 1, 0);
    for (int b = 0; b < B; ++b) offs[b + 1] = offs[b] + counts[b];
    float* base = net->det_host;
    auto work = [&](int b0, int b1) {
      std::vector<yc_detection> d;
      std::vector<int> order;
      for (int b = b0; b < b1; ++b) {
        const int n = counts[b];
        if (n <= 1) continue;
        d.resize(n);
        for (int i = 0; i < n; ++i) {
          float* r = base + (size_t)(offs[b] + i) * R;
          d[i].x = r[1]; d[i].y = r[2]; d[i].w = r[3]; d[i].h = r[4]; d[i].objectness = r[5];
          d[i].classes = classes; d[i].prob = r + 7; d[i].sort_class = -1;
        }
        nms_sort_impl(d.data(), n, classes, nms_thresh, nms_kind, beta, true);  // probs are zeroed in place in the rows
      }
    };
    const int nthreads = std::max(1, std::min<int>(B, (int)std::thread::hardware_concurrency()));
    if (nthreads == 1 || tot < 256) {
      work(0, B);
    } else {
      std::vector<std::thread> th;
      const int per = (B + nthreads - 1) / nthreads;
      for (int t = 0; t < nthreads; ++t) {
        const int b0 = t * per, b1 = std::min(B, b0 + per);
        if (b0 < b1) th.emplace_back(work, b0, b1);
      }
      for (auto& t : th) t.join();
    }
  }
  return YC_OK;
}
