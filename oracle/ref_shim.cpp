// TEST INFRASTRUCTURE — not product code.
//
// Thin C-ABI shim over the UNMODIFIED reference CPU path (yolo-core, compiled from
// /root/reference/src without -DGPU by oracle/Makefile into oracle/_ref/libyolo_ref.so).
// It lets tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg
// drive the reference's own LoadNetwork / NetworkPredict / GetNetworkBoxes / NmsSort /
// TrainNetworkDatum through ctypes and read back per-layer buffers.
//
// Nothing here re-implements arithmetic: every function forwards to a reference entry point
// (cited per function).  The reference headers are included from where they lie under
// /root/reference/src at build time; no reference source is copied into this repo.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <chrono>
#include <string>
#include <vector>

#include "yolo_core.h"      // reference: struct layer / Network / NetworkState, LIB_API
#include "network.h"        // reference: ForwardNetwork, TrainNetworkDatum ...
#include "parser.h"         // reference: ParseNetworkCfg, LoadWeights, SaveWeights
#include "box.h"            // reference: Box, Detection, NmsSort
#include "convolutional_layer.h"
#include "batchnorm_layer.h"
#include "activations.h"
#include "blas.h"
#include "gemm.h"
#include "im2col.h"

extern "C" {

// ---- stubs for the OpenCV / data-loader symbols of files we do not build -----------------
// (image.cpp, data.cpp, image_opencv.cpp are excluded: they need OpenCV C++ headers)
void free_data(data d) { (void)d; }
}
std::vector<std::string> GetList(std::string filename) {   // data.cpp:36 (Metadata only; unused here)
  (void)filename; return std::vector<std::string>();
}
void get_next_batch(data d, int n, int offset, float* X, float* y) {
  (void)d; (void)n; (void)offset; (void)X; (void)y;
  fprintf(stderr, "ref_shim: get_next_batch is not available in the oracle build\n");
  abort();
}

extern "C" {

struct RefNet {
  Network net;
  int train;
};

// LoadNetwork — reference parser.cpp:1852-1875 (ParseNetworkCfg + LoadWeights + FuseConvBatchNorm)
void* ref_load(const char* cfg, const char* weights, int train) {
  RefNet* r = (RefNet*)calloc(1, sizeof(RefNet));
  r->train = train;
  bool ok = LoadNetwork(&r->net, cfg, (weights && weights[0]) ? weights : nullptr, train != 0, false);
  if (!ok) { fprintf(stderr, "ref_load: LoadNetwork failed\n"); }
  return r;
}

// ParseNetworkCfg(train) + LoadWeights WITHOUT FuseConvBatchNorm: used for batch>1 forward
// with BN in rolling-stat mode and for training steps.
void* ref_load_nofuse(const char* cfg, const char* weights, int train) {
  RefNet* r = (RefNet*)calloc(1, sizeof(RefNet));
  r->train = train;
  bool ok = ParseNetworkCfg(&r->net, cfg, train != 0);
  if (ok && weights && weights[0]) ok = LoadWeights(&r->net, weights);
  if (!ok) { fprintf(stderr, "ref_load_nofuse: failed\n"); }
  return r;
}

void ref_free(void* h) {
  RefNet* r = (RefNet*)h;
  FreeNetwork(&r->net);
  free(r);
}

int ref_num_layers(void* h) { return ((RefNet*)h)->net.n; }
// raw pointers to the reference's own structs (tests of the drop-in boundary hand them to libyolo_core_b200.so's
// trampolines: `layer*` parsed by the reference, `forward_gpu` installed by the shim) and their sizes as compiled here
void* ref_network_ptr(void* h) { return &((RefNet*)h)->net; }
void* ref_layer_ptr(void* h, int i) { return &((RefNet*)h)->net.layers[i]; }
int ref_sizeof(int what) {
  switch (what) {
    case 0: return (int)sizeof(layer);
    case 1: return (int)sizeof(Network);
    case 2: return (int)sizeof(NetworkState);
    case 3: return (int)sizeof(Detection);
  }
  return -1;
}
int ref_net_dim(void* h, int which) {
  Network* n = &((RefNet*)h)->net;
  switch (which) { case 0: return n->w; case 1: return n->h; case 2: return n->c; case 3: return n->batch;
                   case 4: return n->inputs; case 5: return n->outputs; case 6: return n->subdiv; }
  return -1;
}

// per-layer integer attributes, read straight from struct layer (yolo_core.h:149-437)
int ref_layer_int(void* h, int i, const char* key) {
  layer* l = &((RefNet*)h)->net.layers[i];
#define K(name) if (!strcmp(key, #name)) return (int)l->name;
  K(type) K(activation) K(batch_normalize) K(batch) K(inputs) K(outputs) K(nweights) K(nbiases)
  K(h) K(w) K(c) K(out_h) K(out_w) K(out_c) K(n) K(groups) K(group_id) K(size) K(stride)
  K(stride_x) K(stride_y) K(dilation) K(pad) K(index) K(classes) K(total) K(max_boxes) K(truths)
#undef K
  return -999999;
}

// per-layer float buffers: returns pointer + element count (borrowed; owned by the layer)
float* ref_layer_buf(void* h, int i, const char* key, long long* count) {
  RefNet* r = (RefNet*)h;
  layer* l = &r->net.layers[i];
  long long nb = (long long)l->batch;
  float* p = nullptr; long long c = 0;
  if (!strcmp(key, "output"))            { p = l->output; c = nb * l->outputs; }
  else if (!strcmp(key, "delta"))        { p = l->delta; c = nb * l->outputs; }
  else if (!strcmp(key, "weights"))      { p = l->weights; c = l->nweights; }
  else if (!strcmp(key, "biases"))       { p = l->biases; c = l->n; }
  else if (!strcmp(key, "scales"))       { p = l->scales; c = l->n; }
  else if (!strcmp(key, "rolling_mean")) { p = l->rolling_mean; c = l->n; }
  else if (!strcmp(key, "rolling_variance")) { p = l->rolling_variance; c = l->n; }
  else if (!strcmp(key, "mean"))         { p = l->mean; c = l->n; }
  else if (!strcmp(key, "variance"))     { p = l->variance; c = l->n; }
  else if (!strcmp(key, "weight_updates")) { p = l->weight_updates; c = l->nweights; }
  else if (!strcmp(key, "bias_updates")) { p = l->bias_updates; c = l->n; }
  else if (!strcmp(key, "scale_updates")) { p = l->scale_updates; c = l->n; }
  else if (!strcmp(key, "x"))            { p = l->x; c = nb * l->outputs; }
  else if (!strcmp(key, "x_norm"))       { p = l->x_norm; c = nb * l->outputs; }
  else if (!strcmp(key, "activation_input")) { p = l->activation_input; c = nb * l->outputs; }
  else if (!strcmp(key, "anchors"))      { p = l->biases; c = 2 * l->total; }
  else if (!strcmp(key, "cost"))         { p = l->cost; c = 1; }
  if (count) *count = p ? c : 0;
  return p;
}

// NetworkPredict — reference network.cpp:412-430 (CPU branch: ForwardNetwork, train=0)
float* ref_predict(void* h, float* input) {
  return NetworkPredict(&((RefNet*)h)->net, input);
}

// Forward only with an explicit train flag and optional truth (reference network.cpp:101-114)
void ref_forward(void* h, float* input, float* truth, int train) {
  RefNet* r = (RefNet*)h;
  NetworkState state = {0};
  state.net = &r->net;
  state.index = 0;
  state.input = input;
  state.truth = truth;
  state.train = train;
  state.delta = 0;
  state.workspace = r->net.workspace;
  ForwardNetwork(&r->net, state);
}

// forward + backward without update: mirrors TrainNetworkDatum (network.cpp:188-208) minus
// UpdateNetwork, so weight_updates / deltas can be inspected.  Returns GetNetworkCost.
float ref_forward_backward(void* h, float* x, float* y) {
  RefNet* r = (RefNet*)h;
  NetworkState state = {0};
  state.net = &r->net;
  state.index = 0;
  state.input = x;
  state.truth = y;
  state.train = 1;
  state.delta = 0;
  state.workspace = r->net.workspace;
  ForwardNetwork(&r->net, state);
  BackwardNetwork(&r->net, state);
  return GetNetworkCost(&r->net);
}

// TrainNetworkDatum — reference network.cpp:188-208 (forward, backward, update when seen%subdiv==0)
float ref_train_datum(void* h, float* x, float* y) {
  return TrainNetworkDatum(&((RefNet*)h)->net, x, y);
}
void ref_update(void* h) { UpdateNetwork(&((RefNet*)h)->net); }
void ref_set_hyper(void* h, float lr, float momentum, float decay) {
  Network* n = &((RefNet*)h)->net;
  n->lr = lr; n->momentum = momentum; n->decay = decay;
}
void ref_save_weights(void* h, const char* path) { SaveWeights(&((RefNet*)h)->net, path); }

// GetNetworkBoxes + NmsSort — reference network.cpp:498-503, box.cpp:393-419.
// Output rows: [x y w h objectness best_class best_prob] followed by `classes` probs per det
// when probs != NULL.  Returns the number of detections (before dropping all-zero ones).
int ref_get_boxes(void* h, float thresh, float nms_thresh, int nms_kind, float beta,
                  float* out7, float* probs, int max_out) {
  RefNet* r = (RefNet*)h;
  int n = 0;
  Detection* dets = GetNetworkBoxes(&r->net, thresh, &n);
  layer* last = &r->net.layers[r->net.n - 1];
  int classes = last->classes;
  if (nms_thresh > 0) NmsSort(dets, n, classes, nms_thresh, (NMS_KIND)nms_kind, beta);
  int m = n < max_out ? n : max_out;
  for (int i = 0; i < m; ++i) {
    float* o = out7 + 7 * (size_t)i;
    o[0] = dets[i].bbox.x; o[1] = dets[i].bbox.y; o[2] = dets[i].bbox.w; o[3] = dets[i].bbox.h;
    o[4] = dets[i].objectness;
    int best = -1; float bp = 0;
    for (int c = 0; c < classes; ++c) if (dets[i].prob[c] > bp) { bp = dets[i].prob[c]; best = c; }
    o[5] = (float)best; o[6] = bp;
    if (probs) memcpy(probs + (size_t)i * classes, dets[i].prob, sizeof(float) * classes);
  }
  FreeDetections(dets, n);
  return n;
}

// ForwardScaleChannelsLayer / BackwardScaleChannelsLayer (scale_channels_layer.c:72-131) and reorg_cpu (blas.c:12-46)
// on caller-provided arrays: a two-layer Network is put together by hand (layers[0] = the `from` layer, layers[1] =
// the [scale_channels] layer) and the reference's own functions run on it.
}
#include "scale_channels_layer.h"
extern "C" {
void ref_scale_channels(const float* from_out, const float* scale_in, int B, int C, int H, int W, int act, float* out,
                        float* delta /* in: l->delta, out: after gradient_array; NULL = forward only */,
                        float* d_scale_in, float* d_from) {
  Network net;
  memset(&net, 0, sizeof(net));
  layer ls[2];
  memset(ls, 0, sizeof(ls));
  net.n = 2;
  net.layers = ls;
  ls[0].output = (float*)from_out;
  ls[0].delta = d_from;
  ls[0].out_w = W; ls[0].out_h = H; ls[0].out_c = C;
  layer* l = &ls[1];
  l->type = SCALE_CHANNELS;
  l->batch = B; l->out_c = C; l->out_h = H; l->out_w = W;
  l->outputs = C * H * W;
  l->index = 0;
  l->scale_wh = 0;
  l->activation = (ACTIVATION)act;
  l->output = out;
  l->delta = delta;
  NetworkState st;
  memset(&st, 0, sizeof(st));
  st.net = &net;
  st.index = 1;
  st.input = (float*)scale_in;
  st.delta = d_scale_in;
  ForwardScaleChannelsLayer(l, st);
  if (delta) BackwardScaleChannelsLayer(l, st);
}
void ref_reorg(const float* x, int ow, int oh, int oc, int batch, int stride, int forward, float* out) {
  reorg_cpu((float*)x, ow, oh, oc, batch, stride, forward, out);
}

// One layer's own CPU forward in training mode on a caller-provided input (used for the loss half of
// ForwardYoloLayer / ForwardGaussianYoloLayer, yolo_layer.cpp:388-772, gaussian_yolo_layer.cpp:486-857, on a head tensor
// that another implementation produced): l->forward(l, state) with state.input = input, state.truth = truth,
// state.train = 1; copies l->output and l->delta out and returns *l->cost.
float ref_layer_forward_train(void* h, int i, const float* input, const float* truth, float* out, float* delta) {
  RefNet* r = (RefNet*)h;
  layer* l = &r->net.layers[i];
  NetworkState st;
  memset(&st, 0, sizeof(st));
  st.net = &r->net;
  st.index = i;
  st.input = (float*)input;
  st.truth = (float*)truth;
  st.train = 1;
  if (l->delta) memset(l->delta, 0, sizeof(float) * (size_t)l->outputs * l->batch);  // network.cpp:108-109
  l->forward(l, st);
  const size_t n = (size_t)l->outputs * l->batch;
  if (out) memcpy(out, l->output, n * sizeof(float));
  if (delta && l->delta) memcpy(delta, l->delta, n * sizeof(float));
  return l->cost ? *l->cost : 0.f;
}

// GetCurrLr (network.cpp:32-84) at a given iteration / max_iter (both are owned by the reference's training driver)
float ref_curr_lr(void* h, long long iter, int max_iter) {
  Network* n = &((RefNet*)h)->net;
  n->curr_iter = (int)iter;
  n->max_iter = max_iter;
  return GetCurrLr(n);
}

// ResizeNetwork (network.cpp:255-383) of the loaded reference network
void ref_resize(void* h, int w, int hh) { ResizeNetwork(&((RefNet*)h)->net, w, hh); }

// NmsSort — reference box.cpp:393-419 on caller-provided detections.  boxes: [n][4] (x y w h), obj: [n],
// probs: [n][classes] (zeroed in place); order_out[i] = original index of the detection that ends at
// position i of the reference's array after its qsorts.
void ref_nms(float* boxes, float* obj, float* probs, int n, int classes, float thresh, int nms_kind, float beta,
             int* order_out) {
  Detection* dets = (Detection*)calloc(n > 0 ? n : 1, sizeof(Detection));
  for (int i = 0; i < n; ++i) {
    dets[i].bbox.x = boxes[4 * i]; dets[i].bbox.y = boxes[4 * i + 1];
    dets[i].bbox.w = boxes[4 * i + 2]; dets[i].bbox.h = boxes[4 * i + 3];
    dets[i].classes = classes;
    dets[i].objectness = obj[i];
    dets[i].prob = probs + (size_t)i * classes;
  }
  NmsSort(dets, n, classes, thresh, (NMS_KIND)nms_kind, beta);
  for (int i = 0; i < n; ++i) order_out[i] = (int)((dets[i].prob - probs) / classes);
  free(dets);
}

// wall-clock timing of K NetworkPredict calls using the reference's own timer granularity
// (std::chrono steady_clock, same clock as GetTimePoint utils.cpp:437-444).  Returns seconds.
double ref_time_predict(void* h, float* input, int iters) {
  RefNet* r = (RefNet*)h;
  auto t0 = std::chrono::steady_clock::now();
  for (int i = 0; i < iters; ++i) NetworkPredict(&r->net, input);
  auto t1 = std::chrono::steady_clock::now();
  return std::chrono::duration<double>(t1 - t0).count();
}

// ---- direct access to the reference's L1 primitives, for pinning oracle/yc_oracle.c ------
// gemm — gemm.c:97-101 → gemm_cpu gemm.c:2967-3005
void ref_gemm(int TA, int TB, int M, int N, int K, float ALPHA, float* A, int lda, float* B, int ldb,
              float BETA, float* C, int ldc) {
  gemm(TA, TB, M, N, K, ALPHA, A, lda, B, ldb, BETA, C, ldc);
}
// im2col_cpu_ext — im2col.c:56-104
void ref_im2col(float* im, int c, int h, int w, int k, int pad, int stride, int dil, float* col) {
  im2col_cpu_ext(im, c, h, w, k, k, pad, pad, stride, stride, dil, dil, col);
}
// activate_array — activations.c:139-168 (and mish/swish variants :170-197)
void ref_activate(float* x, int n, int act, float* saved_input, float* out) {
  if ((ACTIVATION)act == MISH) activate_array_mish(x, n, saved_input, out);
  else if ((ACTIVATION)act == SWISH) activate_array_swish(x, n, saved_input, out);
  else { if (out != x) memcpy(out, x, sizeof(float) * n); activate_array(out, n, (ACTIVATION)act); }
}
int ref_activation_id(const char* name) { return (int)get_activation((char*)name); }

}  // extern "C"
