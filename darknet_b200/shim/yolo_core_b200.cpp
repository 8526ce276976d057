// libyolo_core_b200.so — the reference's own entry points, with the reference's own signatures, on top of libyc_b200.so.
//
// Compiled AGAINST the reference's public header (src/yolo_core.h: struct layer, Network, NetworkState, Detection) so
// that the symbols below can be dropped into a yolo-core build in place of the originals:
//
//   library API (yolo_core.h:625-663, box.h:93-94)       replaced here by
//   ----------------------------------------------------  -----------------------------------------------------------
//   bool LoadNetwork(Network*, cfg, weights, train, clear) yc_network_load + a `Network` whose layers[] mirror the plan
//   void FreeNetwork(Network*)                             yc_network_free
//   float* NetworkPredict(Network*, float* input)          yc_network_predict         (network.cpp:412-430)
//   float* NetworkPredictGpu(Network*, float*)             the same                   (network_kernels.cu:502-522)
//   void ForwardNetworkGpu(Network*, NetworkState)         yc_network_forward         (network_kernels.cu:45-114)
//   Detection* GetNetworkBoxes(Network*, thresh, int*)     yc_network_get_boxes       (network.cpp:498-503)
//   void FreeDetections(Detection*, int)                   (network.cpp:505-516)
//   void NmsSort(Detection*, total, classes, thresh, ...)  yc_nms_sort                (box.cpp:393-419)
//   void FuseConvBatchNorm(Network*)                       no-op: folded at load      (network.cpp:647-682)
//   void ResizeNetwork(Network*, w, h)  [C++ linkage]      yc_network_resize          (network.cpp:255-383)
//
//   per-layer operator seam (yolo_core.h:157-159: void (*forward_gpu)(layer*, NetworkState))
//   ----------------------------------------------------  -----------------------------------------------------------
//   ForwardConvolutionalLayerGpu   (convolutional_kernels.cu:252)   yc_conv_fprop        on a plan keyed by layer*
//   ForwardMaxpoolLayerGpu         (maxpool_layer_kernels.cu:145)   yc_maxpool_fwd
//   ForwardUpsampleLayerGpu        (upsample_layer.c:106)           yc_upsample_fwd
//   ForwardShortcutLayerGpu        (shortcut_layer.c:191)           yc_shortcut_fwd
//   ForwardRouteLayerGpu           (route_layer.c:125)              yc_nchw_to_nhwc into concat slices
//   ForwardYoloLayerGpu / ForwardGaussianYoloLayerGpu (yolo_layer.cpp:836, gaussian_yolo_layer.cpp:934)  yc_yolo_decode
//
// The per-layer trampolines keep the reference's contract at the seam — state.input and l->output_gpu are device fp32
// NCHW — and convert to the library's NHWC bf16 layout on the way in and out (two extra passes per layer: they exist
// so that a single reference layer can be swapped and checked; the fast path is the whole-network replacement above,
// which converts only at the network input and the YOLO heads).  Per-layer state (packed weights, plan, staging
// tensors) hangs off a side table keyed by layer*; nothing is added to struct layer.
//
// Error convention of the reference (utils.cpp:120-125, dark_cuda.c:58-83): print and exit.  The trampolines are void,
// so a failing yc_* call is converted into exactly that.
#include <dlfcn.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <unordered_map>
#include <vector>

#include "yolo_core.h"  // the reference's header (include path: <reference>/src)

#include "../../include/yc_b200.h"

#define SHIM_API extern "C" __attribute__((visibility("default")))

namespace {

[[noreturn]] void die(const char* what) {  // error() in utils.cpp: message + exit
  fprintf(stderr, "yolo_core_b200: %s failed: %s\n", what, yc_last_error());
  exit(EXIT_FAILURE);
}
#define YC_OR_DIE(call, what) \
  do {                        \
    if ((call) != 0) die(what); \
  } while (0)

std::mutex g_mu;
std::unordered_map<const Network*, yc_network*> g_nets;

yc_network* handle_of(const Network* net) {
  std::lock_guard<std::mutex> lk(g_mu);
  auto it = g_nets.find(net);
  return it == g_nets.end() ? nullptr : it->second;
}

// the stream the host program's other layers run on: get_cuda_stream() of the reference when this library is loaded
// into a yolo-core process (dark_cuda.c:128-152), else one stream of our own
void* shim_stream() {
  static void* st = nullptr;
  static bool init = false;
  if (!init) {
    init = true;
    typedef void* (*get_stream_fn)();
    get_stream_fn f = (get_stream_fn)dlsym(RTLD_DEFAULT, "get_cuda_stream");
    if (f) st = f();
    else YC_OR_DIE(yc_stream_create(&st), "yc_stream_create");
  }
  return st;
}

LAYER_TYPE ref_type(int t) {
  switch (t) {
    case 0: return CONVOLUTIONAL;
    case 1: return MAXPOOL;
    case 2: return UPSAMPLE;
    case 3: return ROUTE;
    case 4: return SHORTCUT;
    case 5: return YOLO;
    default: return GAUSSIAN_YOLO;
  }
}

// ---- per-layer state ------------------------------------------------------------------------------------------
struct LayerState {
  void* w_packed = nullptr;
  float* bias = nullptr;
  float* scale = nullptr;
  yc_conv_plan* plan = nullptr;
  void* x = nullptr;   // NHWC staging of the input(s)
  void* x2 = nullptr;  // shortcut: second operand
  void* y = nullptr;   // NHWC staging of the output
  yc_tensor xv{}, x2v{}, yv{};
};
std::unordered_map<const layer*, LayerState> g_layers;
std::unordered_map<const layer*, float*> g_owned_out;  // l->output_gpu buffers yc_shim_install allocated
int g_ref_cpu_maxpool = 0;  // yc_shim_set_ref_cpu_maxpool

LayerState& state_of(const layer* l, bool* fresh) {
  std::lock_guard<std::mutex> lk(g_mu);
  auto it = g_layers.find(l);
  *fresh = it == g_layers.end();
  return g_layers[l];
}

int pad8(int c) { return (c + 7) / 8 * 8; }

yc_tensor make_view(void* p, int n, int h, int w, int c, int ldc, int dtype) {
  yc_tensor t;
  t.ptr = p; t.n = n; t.h = h; t.w = w; t.c = c; t.ldc = ldc; t.dtype = dtype;
  return t;
}

void* dev_alloc(size_t bytes) {
  void* p = yc_malloc(bytes);
  if (!p) die("yc_malloc");
  return p;
}

void stage_io(LayerState& s, const layer* l, int out_dtype) {
  const int ldx = pad8(l->c), ldy = out_dtype == YC_F32 ? (l->out_c + 3) / 4 * 4 : pad8(l->out_c);
  s.x = dev_alloc((size_t)l->batch * l->h * l->w * ldx * 2);
  s.y = dev_alloc((size_t)l->batch * l->out_h * l->out_w * ldy * (out_dtype == YC_F32 ? 4 : 2));
  s.xv = make_view(s.x, l->batch, l->h, l->w, l->c, ldx, YC_BF16);
  s.yv = make_view(s.y, l->batch, l->out_h, l->out_w, l->out_c, ldy, out_dtype);
}

bool feeds_yolo(const layer* l, const NetworkState& st) {
  if (!st.net || st.index + 1 >= st.net->n) return false;
  const LAYER_TYPE t = st.net->layers[st.index + 1].type;
  return t == YOLO || t == GAUSSIAN_YOLO;
}

}  // namespace

// ================================== per-layer trampolines ======================================================
SHIM_API void ForwardConvolutionalLayerGpu(layer* l, NetworkState state) {
  if (state.train) {
    fprintf(stderr, "yolo_core_b200: ForwardConvolutionalLayerGpu(train) is served by yc_network_train_datum "
                    "(whole-network replacement), not by the per-layer trampoline\n");
    exit(EXIT_FAILURE);
  }
  void* st = shim_stream();
  bool fresh;
  LayerState& s = state_of(l, &fresh);
  if (fresh) {
    if (l->groups > 1 || l->xnor || l->binary) {
      fprintf(stderr, "yolo_core_b200: grouped / xnor convolutions are outside the hot path\n");
      exit(EXIT_FAILURE);
    }
    const int n = l->n, c = l->c, k = l->size;
    const size_t nw = (size_t)n * c * k * k;
    // parameters: host fp32 in the reference's layouts (l->weights OIHW, l->biases, BN: scales / rolling stats).
    // A layer that still has batchnorm (FuseConvBatchNorm not run) is normalised from the rolling statistics in the
    // epilogue: scale = gamma / sqrt(var + 1e-6), bias = beta - mean * scale (batchnorm_layer.cpp:226-237).
    std::vector<float> b(l->biases, l->biases + n), sc;
    if (l->batch_normalize) {
      sc.resize(n);
      for (int f = 0; f < n; ++f) {
        sc[f] = l->scales[f] / sqrtf(l->rolling_variance[f] + .000001f);
        b[f] = l->biases[f] - l->rolling_mean[f] * sc[f];
      }
    }
    float* w_dev = (float*)dev_alloc(nw * 4);
    YC_OR_DIE(yc_push(w_dev, l->weights, nw * 4, st), "yc_push(weights)");
    s.w_packed = dev_alloc(yc_conv_packed_bytes(n, c, k));
    YC_OR_DIE(yc_conv_pack(w_dev, n, c, k, s.w_packed, st), "yc_conv_pack");
    s.bias = (float*)dev_alloc((size_t)n * 4);
    YC_OR_DIE(yc_push(s.bias, b.data(), (size_t)n * 4, st), "yc_push(bias)");
    if (!sc.empty()) {
      s.scale = (float*)dev_alloc((size_t)n * 4);
      YC_OR_DIE(yc_push(s.scale, sc.data(), (size_t)n * 4, st), "yc_push(scale)");
    }
    YC_OR_DIE(yc_stream_sync(st), "yc_stream_sync");
    yc_free(w_dev);
    stage_io(s, l, feeds_yolo(l, state) ? YC_F32 : YC_BF16);  // head outputs stay fp32 (thresholds see unrounded logits)
    yc_conv_desc d;
    d.ksize = k; d.stride = l->stride; d.pad = l->pad; d.dilation = l->dilation > 0 ? l->dilation : 1;
    d.activation = (int)l->activation;  // YC_* ids == enum ACTIVATION
    d.residual_activation = YC_LINEAR;
    YC_OR_DIE(yc_conv_plan_create(&s.plan, &d, &s.xv, s.w_packed, s.bias, s.scale, nullptr, &s.yv), "yc_conv_plan_create");
  }
  YC_OR_DIE(yc_nchw_to_nhwc(state.input, &s.xv, st), "yc_nchw_to_nhwc");
  YC_OR_DIE(yc_conv_fprop(s.plan, st), "yc_conv_fprop");
  YC_OR_DIE(yc_nhwc_to_nchw(&s.yv, l->output_gpu, st), "yc_nhwc_to_nchw");
}

SHIM_API void ForwardMaxpoolLayerGpu(layer* l, NetworkState state) {
  void* st = shim_stream();
  bool fresh;
  LayerState& s = state_of(l, &fresh);
  if (fresh) stage_io(s, l, YC_BF16);
  YC_OR_DIE(yc_nchw_to_nhwc(state.input, &s.xv, st), "yc_nchw_to_nhwc");
  YC_OR_DIE(yc_maxpool_fwd(&s.xv, &s.yv, l->size, l->stride, l->pad, g_ref_cpu_maxpool, st), "yc_maxpool_fwd");
  YC_OR_DIE(yc_nhwc_to_nchw(&s.yv, l->output_gpu, st), "yc_nhwc_to_nchw");
}

SHIM_API void ForwardUpsampleLayerGpu(layer* l, NetworkState state) {
  void* st = shim_stream();
  bool fresh;
  LayerState& s = state_of(l, &fresh);
  if (fresh) stage_io(s, l, YC_BF16);
  YC_OR_DIE(yc_nchw_to_nhwc(state.input, &s.xv, st), "yc_nchw_to_nhwc");
  YC_OR_DIE(yc_upsample_fwd(&s.xv, &s.yv, l->stride, l->scale, st), "yc_upsample_fwd");
  YC_OR_DIE(yc_nhwc_to_nchw(&s.yv, l->output_gpu, st), "yc_nhwc_to_nchw");
}

SHIM_API void ForwardShortcutLayerGpu(layer* l, NetworkState state) {
  void* st = shim_stream();
  bool fresh;
  LayerState& s = state_of(l, &fresh);
  if (fresh) {
    stage_io(s, l, YC_BF16);
    s.x2 = dev_alloc((size_t)l->batch * l->h * l->w * s.xv.ldc * 2);
    s.x2v = s.xv;
    s.x2v.ptr = s.x2;
  }
  const layer* from = &state.net->layers[l->input_layers ? l->input_layers[0] : l->index];
  YC_OR_DIE(yc_nchw_to_nhwc(state.input, &s.xv, st), "yc_nchw_to_nhwc");
  YC_OR_DIE(yc_nchw_to_nhwc(from->output_gpu, &s.x2v, st), "yc_nchw_to_nhwc");
  YC_OR_DIE(yc_shortcut_fwd(&s.xv, &s.x2v, &s.yv, (int)l->activation, st), "yc_shortcut_fwd");
  YC_OR_DIE(yc_nhwc_to_nchw(&s.yv, l->output_gpu, st), "yc_nhwc_to_nchw");
}

SHIM_API void ForwardRouteLayerGpu(layer* l, NetworkState state) {
  void* st = shim_stream();
  bool fresh;
  LayerState& s = state_of(l, &fresh);
  if (fresh) {
    const int ld = pad8(l->out_c);
    s.y = dev_alloc((size_t)l->batch * l->out_h * l->out_w * ld * 2);
    s.yv = make_view(s.y, l->batch, l->out_h, l->out_w, l->out_c, ld, YC_BF16);
  }
  // route_layer.c:125-142: input i contributes its group_id-th 1/groups channel slice; slices are concatenated
  int coff = 0;
  for (int i = 0; i < l->n; ++i) {
    const layer* in = &state.net->layers[l->input_layers[i]];
    const int part = in->out_c / l->groups;
    const size_t plane = (size_t)in->out_h * in->out_w;
    for (int b = 0; b < l->batch; ++b) {
      const float* src = in->output_gpu + ((size_t)b * in->out_c + (size_t)part * l->group_id) * plane;
      yc_tensor dst = make_view((char*)s.y + ((size_t)b * l->out_h * l->out_w * s.yv.ldc + coff) * 2, 1, l->out_h, l->out_w,
                                part, s.yv.ldc, YC_BF16);
      YC_OR_DIE(yc_nchw_to_nhwc(src, &dst, st), "yc_nchw_to_nhwc");
    }
    coff += part;
  }
  YC_OR_DIE(yc_nhwc_to_nchw(&s.yv, l->output_gpu, st), "yc_nhwc_to_nchw");
}

static void forward_yolo(layer* l, NetworkState state, int gaussian) {
  void* st = shim_stream();
  bool fresh;
  LayerState& s = state_of(l, &fresh);
  if (fresh) {
    const int ld = (l->c + 3) / 4 * 4;
    s.x = dev_alloc((size_t)l->batch * l->h * l->w * ld * 4);
    s.xv = make_view(s.x, l->batch, l->h, l->w, l->c, ld, YC_F32);
  }
  YC_OR_DIE(yc_nchw_to_nhwc(state.input, &s.xv, st), "yc_nchw_to_nhwc");
  YC_OR_DIE(yc_yolo_decode(&s.xv, l->output_gpu, l->n, l->classes, l->scale_x_y, gaussian, st), "yc_yolo_decode");
  if (!state.train && l->output) {  // yolo_layer.cpp:856: cuda_pull_array_async(l->output_gpu, l->output, ...) + sync
    YC_OR_DIE(yc_pull(l->output, l->output_gpu, (size_t)l->batch * l->outputs * 4, st), "yc_pull");
    YC_OR_DIE(yc_stream_sync(st), "yc_stream_sync");
  }
}
SHIM_API void ForwardYoloLayerGpu(layer* l, NetworkState state) { forward_yolo(l, state, 0); }
SHIM_API void ForwardGaussianYoloLayerGpu(layer* l, NetworkState state) { forward_yolo(l, state, 1); }

// Install the trampolines into a reference-parsed layer (what FillConvLayer & co. do at convolutional_layer.cpp:601-603,
// maxpool_layer.cpp:85-86, shortcut_layer.c:61-62, route_layer.c:37-38, upsample_layer.c:35-36, yolo_layer.cpp:57-58) and
// give it a device output buffer if the parser that built it had none (CPU-only build).  Returns 0 if the layer type is
// not on the hot path.
SHIM_API int yc_shim_install(layer* l) {
  switch (l->type) {
    case CONVOLUTIONAL: l->forward_gpu = ForwardConvolutionalLayerGpu; break;
    case MAXPOOL: l->forward_gpu = ForwardMaxpoolLayerGpu; break;
    case UPSAMPLE: l->forward_gpu = ForwardUpsampleLayerGpu; break;
    case SHORTCUT: l->forward_gpu = ForwardShortcutLayerGpu; break;
    case ROUTE: l->forward_gpu = ForwardRouteLayerGpu; break;
    case YOLO: l->forward_gpu = ForwardYoloLayerGpu; break;
    case GAUSSIAN_YOLO: l->forward_gpu = ForwardGaussianYoloLayerGpu; break;
    default: return 0;
  }
  if (!l->output_gpu) {
    l->output_gpu = (float*)dev_alloc((size_t)l->batch * l->outputs * 4);
    std::lock_guard<std::mutex> lk(g_mu);
    g_owned_out[l] = l->output_gpu;
  }
  return 1;
}

// 1: ForwardMaxpoolLayerGpu reproduces the border artefact of the reference's CPU *inference* routine
// (forward_maxpool_layer_avx, gemm.c:2127-2150) so that a trampoline run can be compared bit for bit with the CPU path;
// 0 (default): the semantics of the GPU kernel it replaces (maxpool_layer_kernels.cu:58-101).
SHIM_API void yc_shim_set_ref_cpu_maxpool(int on) { g_ref_cpu_maxpool = on ? 1 : 0; }

SHIM_API float* yc_shim_layer_output_gpu(layer* l) { return l->output_gpu; }

// Drop the side-table entry of a layer (free_layer's counterpart for what the trampolines allocated).
SHIM_API void yc_shim_release(layer* l) {
  std::lock_guard<std::mutex> lk(g_mu);
  auto oo = g_owned_out.find(l);
  if (oo != g_owned_out.end()) {
    if (l->output_gpu == oo->second) l->output_gpu = nullptr;
    yc_free(oo->second);
    g_owned_out.erase(oo);
  }
  auto it = g_layers.find(l);
  if (it == g_layers.end()) return;
  LayerState& s = it->second;
  if (s.plan) yc_conv_plan_destroy(s.plan);
  for (void* p : {s.w_packed, (void*)s.bias, (void*)s.scale, s.x, s.x2, s.y})
    if (p) yc_free(p);
  g_layers.erase(it);
}

// Call l->forward_gpu(l, state) the way ForwardNetworkGpu does (network_kernels.cu:62-73): NetworkState by value,
// state.input = the previous layer's device output.  For bindings that cannot build a by-value struct (ctypes tests).
SHIM_API void yc_shim_call_forward_gpu(Network* net, int index, float* input_dev, int train) {
  NetworkState state;
  memset(&state, 0, sizeof(state));
  state.net = net;
  state.index = index;
  state.input = input_dev;
  state.train = train;
  layer* l = &net->layers[index];
  l->forward_gpu(l, state);
  YC_OR_DIE(yc_stream_sync(shim_stream()), "yc_stream_sync");
}

// ================================== library API ==================================================================
SHIM_API bool LoadNetwork(Network* net, char const* model_file, char const* weights_file, bool train, bool clear) {
  printf(" Try to load model: %s, weights: %s, clear = %d \n", model_file, weights_file ? weights_file : "(null)", clear);
  if (yc_set_device(yc_get_device() >= 0 ? yc_get_device() : 0)) {
    fprintf(stderr, "yolo_core_b200: %s\n", yc_last_error());
    return false;
  }
  yc_network* h = nullptr;
  // the reference forces batch 1 for inference (parser.cpp:1114-1115); training keeps [net] batch / subdivisions
  const int flags = train ? (YC_NET_DEFAULT | YC_NET_TRAIN) : YC_NET_DEFAULT;
  if (yc_network_load(&h, model_file, weights_file, train ? 0 : 1, flags)) {
    fprintf(stderr, "yolo_core_b200: LoadNetwork: %s\n", yc_last_error());
    return false;
  }
  if (!weights_file && yc_network_finalize(h)) {
    fprintf(stderr, "yolo_core_b200: LoadNetwork: %s\n", yc_last_error());
    yc_network_free(h);
    return false;
  }
  memset(net, 0, sizeof(*net));
  net->n = yc_network_info(h, 0);
  net->batch = yc_network_info(h, 1);
  net->w = yc_network_info(h, 2);
  net->h = yc_network_info(h, 3);
  net->c = yc_network_info(h, 4);
  net->inputs = net->w * net->h * net->c;
  net->subdiv = train ? yc_network_train_info(h, 1) : 1;
  net->train = train;
  net->layers = (layer*)calloc(net->n, sizeof(layer));
  for (int i = 0; i < net->n; ++i) {
    layer* l = &net->layers[i];
    l->type = ref_type(yc_network_layer_info(h, i, 0));
    l->index = i;
    l->batch = net->batch;
    l->out_c = yc_network_layer_info(h, i, 1);
    l->out_h = yc_network_layer_info(h, i, 2);
    l->out_w = yc_network_layer_info(h, i, 3);
    l->c = yc_network_layer_info(h, i, 5);
    l->outputs = l->out_c * l->out_h * l->out_w;
    l->size = yc_network_layer_info(h, i, 6);
    l->stride = yc_network_layer_info(h, i, 7);
    l->activation = (ACTIVATION)yc_network_layer_info(h, i, 8);
    l->batch_normalize = train ? yc_network_layer_info(h, i, 9) : 0;  // folded at load, like FuseConvBatchNorm leaves it
    l->pad = yc_network_layer_info(h, i, 19);
    l->dilation = yc_network_layer_info(h, i, 20);
    l->groups = yc_network_layer_info(h, i, 21);
    l->group_id = yc_network_layer_info(h, i, 22);
    l->n = yc_network_layer_info(h, i, 24);
    const int nin = yc_network_layer_info(h, i, 25);
    const int in0 = yc_network_layer_input(h, i, 0);
    if (in0 >= 0) { l->h = net->layers[in0].out_h; l->w = net->layers[in0].out_w; }
    else { l->h = net->h; l->w = net->w; }
    l->inputs = l->h * l->w * l->c;
    if (l->type == ROUTE || l->type == SHORTCUT) {
      const int first = l->type == SHORTCUT ? 1 : 0;  // shortcut: inputs = {previous, from}; the reference lists `from`
      l->n = nin - first;
      l->input_layers = (int*)calloc(l->n > 0 ? l->n : 1, sizeof(int));
      l->input_sizes = (int*)calloc(l->n > 0 ? l->n : 1, sizeof(int));
      for (int k = first; k < nin; ++k) {
        const int src = yc_network_layer_input(h, i, k);
        l->input_layers[k - first] = src;
        l->input_sizes[k - first] = net->layers[src].outputs;
      }
      if (l->type == SHORTCUT) l->index = l->input_layers[0];  // shortcut_layer.c:41: l->index names the `from` layer
    }
    if (l->type == UPSAMPLE) l->scale = yc_network_layer_float(h, i, 0);
    if (l->type == YOLO || l->type == GAUSSIAN_YOLO) {
      l->classes = yc_network_layer_info(h, i, 16);
      l->total = yc_network_layer_info(h, i, 17);
      l->nms_kind = (NMS_KIND)yc_network_layer_info(h, i, 18);
      l->max_boxes = yc_network_layer_info(h, i, 26);
      l->scale_x_y = yc_network_layer_float(h, i, 0);
      l->beta_nms = yc_network_layer_float(h, i, 1);
      l->ignore_thresh = yc_network_layer_float(h, i, 2);
      l->truth_thresh = yc_network_layer_float(h, i, 3);
      l->biases = (float*)calloc((size_t)2 * l->total, sizeof(float));
      yc_network_layer_array(h, i, 0, l->biases, 2 * l->total);
      std::vector<float> m(l->n > 0 ? l->n : 1);
      yc_network_layer_array(h, i, 1, m.data(), (int)m.size());
      l->mask = (int*)calloc(m.size(), sizeof(int));
      for (size_t k = 0; k < m.size(); ++k) l->mask[k] = (int)m[k];
      l->output = train ? nullptr : (float*)yc_network_yolo_output(h, i);  // library-owned pinned buffer (yolo_layer.cpp:62-66)
    }
    switch (l->type) {
      case CONVOLUTIONAL: l->forward_gpu = ForwardConvolutionalLayerGpu; break;
      case MAXPOOL: l->forward_gpu = ForwardMaxpoolLayerGpu; break;
      case UPSAMPLE: l->forward_gpu = ForwardUpsampleLayerGpu; break;
      case SHORTCUT: l->forward_gpu = ForwardShortcutLayerGpu; break;
      case ROUTE: l->forward_gpu = ForwardRouteLayerGpu; break;
      case YOLO: l->forward_gpu = ForwardYoloLayerGpu; break;
      default: l->forward_gpu = ForwardGaussianYoloLayerGpu; break;
    }
  }
  net->outputs = net->layers[net->n - 1].outputs;
  if (clear) {
    net->seen = 0;
    net->curr_iter = 0;
  }
  std::lock_guard<std::mutex> lk(g_mu);
  g_nets[net] = h;
  return true;
}

SHIM_API void FreeNetwork(Network* net) {
  yc_network* h = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_nets.find(net);
    if (it != g_nets.end()) { h = it->second; g_nets.erase(it); }
  }
  if (h) yc_network_free(h);
  if (net->layers) {
    for (int i = 0; i < net->n; ++i) {
      layer* l = &net->layers[i];
      yc_shim_release(l);
      free(l->input_layers);
      free(l->input_sizes);
      free(l->mask);
      if (l->type == YOLO || l->type == GAUSSIAN_YOLO) free(l->biases);
    }
    free(net->layers);
    net->layers = nullptr;
  }
}

// ResizeNetwork (network.h:34, network.cpp:255-383) — C++ linkage like the original; yc_shim_resize_network is the same
// entry for C callers.  The library re-derives every shape (yc_network_resize); the mirror's spatial fields follow, and
// per-layer trampoline state of the old shapes is dropped.
__attribute__((visibility("default"))) void ResizeNetwork(Network* net, int w, int h) {
  yc_network* hn = handle_of(net);
  if (!hn) {
    fprintf(stderr, "yolo_core_b200: ResizeNetwork: not a network of this library\n");
    exit(EXIT_FAILURE);
  }
  YC_OR_DIE(yc_network_resize(hn, w, h), "yc_network_resize");
  net->w = yc_network_info(hn, 2);
  net->h = yc_network_info(hn, 3);
  net->inputs = net->w * net->h * net->c;
  for (int i = 0; i < net->n; ++i) {
    layer* l = &net->layers[i];
    yc_shim_release(l);
    l->out_h = yc_network_layer_info(hn, i, 2);
    l->out_w = yc_network_layer_info(hn, i, 3);
    l->outputs = l->out_c * l->out_h * l->out_w;
    const int in0 = yc_network_layer_input(hn, i, 0);
    if (in0 >= 0) { l->h = net->layers[in0].out_h; l->w = net->layers[in0].out_w; }
    else { l->h = net->h; l->w = net->w; }
    l->inputs = l->h * l->w * l->c;
    if (l->type == ROUTE || l->type == SHORTCUT)
      for (int k = 0; k < l->n; ++k) l->input_sizes[k] = net->layers[l->input_layers[k]].outputs;
    if ((l->type == YOLO || l->type == GAUSSIAN_YOLO) && !net->train)
      l->output = (float*)yc_network_yolo_output(hn, i);  // the pinned head buffers were reallocated
  }
  net->outputs = net->layers[net->n - 1].outputs;
}
SHIM_API void yc_shim_resize_network(Network* net, int w, int h) { ResizeNetwork(net, w, h); }

SHIM_API void FuseConvBatchNorm(Network* net) { (void)net; /* folded when the network was planned (csrc/network.cu) */ }

SHIM_API float* NetworkPredict(Network* net, float* input) {
  yc_network* h = handle_of(net);
  if (!h) {
    fprintf(stderr, "yolo_core_b200: NetworkPredict on a network that LoadNetwork did not create\n");
    exit(EXIT_FAILURE);
  }
  float* out = yc_network_predict(h, input);
  if (!out) die("NetworkPredict");
  return out;
}
SHIM_API float* NetworkPredictGpu(Network* net, float* input) { return NetworkPredict(net, input); }

SHIM_API void ForwardNetworkGpu(Network* net, NetworkState state) {
  yc_network* h = handle_of(net);
  if (!h) {
    fprintf(stderr, "yolo_core_b200: ForwardNetworkGpu on a network that LoadNetwork did not create\n");
    exit(EXIT_FAILURE);
  }
  YC_OR_DIE(yc_network_forward(h, state.input, shim_stream()), "ForwardNetworkGpu");
}

SHIM_API Detection* GetNetworkBoxes(Network* net, float thresh, int* num) {
  yc_network* h = handle_of(net);
  int n = 0;
  yc_detection* d = h ? yc_network_get_boxes(h, 0, thresh, &n) : nullptr;
  if (!d) die("GetNetworkBoxes");
  const layer* last = &net->layers[net->n - 1];
  Detection* dets = (Detection*)calloc(n > 0 ? n : 1, sizeof(Detection));  // MakeNetworkBoxes network.cpp:450-471
  for (int i = 0; i < n; ++i) {
    dets[i].bbox.x = d[i].x; dets[i].bbox.y = d[i].y; dets[i].bbox.w = d[i].w; dets[i].bbox.h = d[i].h;
    dets[i].classes = d[i].classes;
    dets[i].prob = d[i].prob;  // ownership moves to the Detection (freed by FreeDetections)
    d[i].prob = nullptr;
    dets[i].objectness = d[i].objectness;
    dets[i].sort_class = d[i].sort_class;
    if (last->type == GAUSSIAN_YOLO) {
      dets[i].uc = (float*)calloc(4, sizeof(float));
      memcpy(dets[i].uc, d[i].uc, 4 * sizeof(float));
    }
    dets[i].points = YOLO_CENTER;
  }
  free(d);  // the array shell only: every prob pointer has been handed over
  if (num) *num = n;
  return dets;
}

SHIM_API void FreeDetections(Detection* dets, int n) {  // network.cpp:505-516
  for (int i = 0; i < n; ++i) {
    free(dets[i].prob);
    if (dets[i].uc) free(dets[i].uc);
    if (dets[i].mask) free(dets[i].mask);
  }
  free(dets);
}

// NmsSort (box.cpp:393-419) on the reference's Detection array, final array order included: yc_nms_sort runs on a
// mirror that shares the prob arrays, then the Detection structs are permuted the way the qsorts moved the mirror.
// (box.h declares it outside the extern "C" block: C++ linkage, like the original)
__attribute__((visibility("default"))) void NmsSort(Detection* dets, int total, int classes, float thresh, NMS_KIND nms_kind,
                                                    float beta) {
  if (total <= 0) return;
  std::vector<yc_detection> m(total);
  std::unordered_map<const float*, int> where;
  for (int i = 0; i < total; ++i) {
    m[i].x = dets[i].bbox.x; m[i].y = dets[i].bbox.y; m[i].w = dets[i].bbox.w; m[i].h = dets[i].bbox.h;
    m[i].classes = classes;
    m[i].prob = dets[i].prob;
    m[i].objectness = dets[i].objectness;
    m[i].sort_class = dets[i].sort_class;
    where[dets[i].prob] = i;
  }
  yc_nms_sort(m.data(), total, classes, thresh, nms_kind == DIOU_NMS ? 1 : 0, beta);
  std::vector<Detection> old(dets, dets + total);
  for (int i = 0; i < total; ++i) {
    dets[i] = old[where[m[i].prob]];
    dets[i].sort_class = m[i].sort_class;
  }
}

// C-linkage alias for bindings that cannot name a mangled symbol
SHIM_API void yc_shim_nms_sort(Detection* dets, int total, int classes, float thresh, int nms_kind, float beta) {
  NmsSort(dets, total, classes, thresh, (NMS_KIND)nms_kind, beta);
}

SHIM_API int CudaGetDeviceCount() { return yc_device_count(); }
SHIM_API void cuda_set_device(int n) { YC_OR_DIE(yc_set_device(n), "cuda_set_device"); }

// sizes of the reference structs this library was compiled against: a binding checks them against its own headers
SHIM_API int yc_shim_sizeof(int what) {
  switch (what) {
    case 0: return (int)sizeof(layer);
    case 1: return (int)sizeof(Network);
    case 2: return (int)sizeof(NetworkState);
    case 3: return (int)sizeof(Detection);
  }
  return -1;
}
