// Training step of the whole network (C++ host runtime): the B200-native counterpart of the reference's
// ForwardBackwardNetworkGpu / BackwardNetworkGpu / UpdateNetworkGpu / TrainNetworkDatumGpu
// (network_kernels.cu:45-293) and of the per-layer backward_gpu / update_gpu entries they dispatch to.
//
// Same buffers as inference for the activations (NHWC bf16, routes elided by concat placement), plus per layer:
//   conv : fp32 master weights (OIHW, the reference layout), the two bf16 operand packings, the pre-batchnorm
//          output z (l->x_gpu), batch statistics, fp32 gradient accumulators;  delta: NHWC bf16 (l->delta_gpu).
// One step = zero the deltas -> forward (conv + fused statistics -> finalize -> normalise+activate ... YOLO decode
// -> YOLO loss, whose kernel also adds its delta into the head conv's delta) -> backward in reverse layer order
// (BN/activation backward in place, tcgen05 wgrad, dgrad through the forward kernel, scatter/gather passes of the
// memory-bound layers).  Semantics follow the reference CPU path: dgrad OVERWRITES the previous
// layer's delta (col2im.c:70) unless YC_NET_DGRAD_ACCUMULATE, the batchnorm shift gets no gradient unless
// YC_NET_BN_BIAS_GRAD (B8), weight gradients accumulate until the update scales them by the momentum (B7).
#include <math.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include "network_internal.hpp"

using namespace yc;

namespace {

yc_tensor in_view(const yc_network* net, int i) {
  if (i == 0 && net->stem)
    return yc_tensor{net->input_nhwc, net->batch, net->layers[0].out_h, net->layers[0].out_w, 32, 32, YC_BF16};
  if (i == 0) return yc_tensor{net->input_nhwc, net->batch, net->h, net->w, net->c, round_up(net->c, 8), YC_BF16};
  return view_of(net, i - 1);
}
yc_tensor delta_view(const yc_network* net, int i) {
  const NLayer& L = net->layers[i];
  return yc_tensor{L.delta_buf, net->batch, L.out_h, L.out_w, L.out_c, L.delta_ldc, YC_BF16};
}
yc_tensor z_view(const yc_network* net, int i) {
  const NLayer& L = net->layers[i];
  return yc_tensor{L.z_buf, net->batch, L.out_h, L.out_w, L.out_c, L.z_ldc, YC_BF16};
}

// Geometry the GEMM kernels see for conv layer i: the layer's own, or — for the stem (net->stem, layer 0) — a 1x1
// convolution over the 32 im2col columns, with weights / gradients kept in that [n][32] order (column = tap*c + ci).
struct Geo { int c, size, stride, pad; };
Geo geo(const yc_network* net, int i) {
  const NLayer& L = net->layers[i];
  if (i == 0 && net->stem) return Geo{32, 1, 1, 0};
  return Geo{L.c, L.size, L.stride, L.pad};
}
// OIHW [n][c][k][k] <-> stem order [n][32]
void stem_permute(const NLayer& L, const float* src, float* dst, bool to_stem) {
  const int kk = L.size * L.size;
  if (to_stem) std::fill(dst, dst + (size_t)L.n * 32, 0.f);
  for (int f = 0; f < L.n; ++f)
    for (int ci = 0; ci < L.c; ++ci)
      for (int t = 0; t < kk; ++t) {
        const size_t o = ((size_t)f * L.c + ci) * kk + t, q = (size_t)f * 32 + t * L.c + ci;
        if (to_stem) dst[q] = src[o];
        else dst[o] = src[q];
      }
}

template <typename T>
int dev_alloc(T** p, size_t count, bool zero = true) {
  NET_CUDA(cudaMalloc((void**)p, std::max<size_t>(count, 1) * sizeof(T)));
  if (zero) NET_CUDA(cudaMemset(*p, 0, std::max<size_t>(count, 1) * sizeof(T)));
  return YC_OK;
}
int upload(float** p, const std::vector<float>& v) {
  int rc = dev_alloc(p, v.size(), false);
  if (rc) return rc;
  NET_CUDA(cudaMemcpy(*p, v.data(), v.size() * 4, cudaMemcpyHostToDevice));
    NET_CUDA(h2d_done());
  return YC_OK;
}

}  // namespace

void yc_train_free(yc_network* net) {
  for (NLayer& L : net->layers) {
    L.bias_upd = L.scale_upd = nullptr;  // inside the dw_packed bucket
    float** fp[] = {&L.w_master, &L.rmean_dev, &L.rvar_dev, &L.mean_dev, &L.var_dev, &L.stats_dev, &L.ws_dev,
                    &L.dw_packed, &L.ydelta, &L.ystats, &L.adam_m, &L.adam_v};
    for (float** p : fp)
      if (*p) { cudaFree(*p); *p = nullptr; }
    void** vp[] = {&L.w_packed_d, &L.z_buf, &L.dil_buf};
    for (void** p : vp)
      if (*p) { cudaFree(*p); *p = nullptr; }
    if (L.idx_buf) { cudaFree(L.idx_buf); L.idx_buf = nullptr; }
    if (L.wplan) { yc_conv_wgrad_plan_destroy(L.wplan); L.wplan = nullptr; }
    if (L.dplan) { yc_conv_plan_destroy(L.dplan); L.dplan = nullptr; }
    L.delta_buf = nullptr;
  }
  if (net->delta_slab) { cudaFree(net->delta_slab); net->delta_slab = nullptr; }
  if (net->truth_dev) { cudaFree(net->truth_dev); net->truth_dev = nullptr; }
  if (net->truth_pinned) { cudaFreeHost(net->truth_pinned); net->truth_pinned = nullptr; }
  if (net->cost_host) { cudaFreeHost(net->cost_host); net->cost_host = nullptr; }
  yc_dp_free(net);
}

// Called by yc_network_finalize (after the activation buffers and the input staging exist) when YC_NET_TRAIN is set.
int yc_train_finalize(yc_network* net) {
  const int nl = (int)net->layers.size();
  auto& Ls = net->layers;
  const int B = net->batch;
  // ---- delta slab: one allocation, zeroed by one memset per step (network.cpp:108-109 zeroes per layer) ----------
  size_t off = 0;
  std::vector<size_t> offs(nl, 0);
  for (int i = 0; i < nl; ++i) {
    NLayer& L = Ls[i];
    if (L.type == L_YOLO || L.type == L_GYOLO) continue;
    L.delta_ldc = L.type == L_CONV ? yc_conv_dgrad_channels(L.out_c) : round_up(L.out_c, 8);
    offs[i] = off;
    off += ((size_t)B * L.out_h * L.out_w * L.delta_ldc * 2 + 255) & ~(size_t)255;
  }
  net->delta_slab_bytes = off;
  NET_CUDA(cudaMalloc(&net->delta_slab, std::max<size_t>(off, 256)));
  NET_CUDA(cudaMemset(net->delta_slab, 0, std::max<size_t>(off, 256)));
  for (int i = 0; i < nl; ++i)
    if (Ls[i].type != L_YOLO && Ls[i].type != L_GYOLO) Ls[i].delta_buf = (char*)net->delta_slab + offs[i];

  for (int i = 0; i < nl; ++i) {
    NLayer& L = Ls[i];
    switch (L.type) {
      case L_CONV: {
        const Geo G = geo(net, i);
        int rc;
        if (i == 0 && net->stem) {
          std::vector<float> w2((size_t)L.n * 32);
          stem_permute(L, L.weights.data(), w2.data(), true);
          rc = upload(&L.w_master, w2);
        } else {
          rc = upload(&L.w_master, L.weights);
        }
        if (!rc) rc = upload(&L.bias_dev, L.biases);
        if (!rc && L.bn) rc = upload(&L.scale_dev, L.scales);
        if (!rc && L.bn) rc = upload(&L.rmean_dev, L.rmean);
        if (!rc && L.bn) rc = upload(&L.rvar_dev, L.rvar);
        if (!rc && L.bn) rc = dev_alloc(&L.mean_dev, L.n);
        if (!rc && L.bn) rc = dev_alloc(&L.var_dev, L.n);
        if (!rc && L.bn) rc = dev_alloc(&L.stats_dev, 2 * (size_t)L.n);
        if (!rc) rc = dev_alloc(&L.ws_dev, yc_bn_bwd_workspace_floats(L.n));  // zero-filled: holds the reduction ticket
        if (!rc) {
          // one contiguous gradient bucket per layer: [weight_updates | bias_updates | scale_updates] — what the
          // data-parallel all-reduce sends in a single call (network_dp.cu)
          const size_t nw = yc_conv_wgrad_bytes(L.n, G.c, G.size) / 4;
          L.bucket_floats = nw + (size_t)round_up(L.n, 4) * 2;
          rc = dev_alloc(&L.dw_packed, L.bucket_floats);
          L.bias_upd = L.dw_packed + nw;
          L.scale_upd = L.bias_upd + round_up(L.n, 4);
          if (!rc && net->adam) rc = dev_alloc(&L.adam_m, L.bucket_floats);  // zero-filled like the reference's calloc
          if (!rc && net->adam) rc = dev_alloc(&L.adam_v, L.bucket_floats);
        }
        if (rc) return rc;
        NET_CUDA(cudaMalloc(&L.w_packed, yc_conv_packed_bytes(L.n, G.c, G.size)));
        NET_CUDA(cudaMalloc(&L.w_packed_d, yc_conv_dgrad_packed_bytes(L.n, G.c, G.size)));
        rc = yc_conv_pack(L.w_master, L.n, G.c, G.size, L.w_packed, net->stream);
        if (!rc) rc = yc_conv_pack_dgrad(L.w_master, L.n, G.c, G.size, L.w_packed_d, net->stream);
        if (rc) return rc;
        const yc_tensor x = in_view(net, i);
        const yc_tensor y = view_of(net, i);
        yc_conv_desc d{G.size, G.stride, G.pad, L.dilation, L.bn ? YC_LINEAR : L.act, YC_LINEAR};
        if (L.bn) {
          L.z_ldc = round_up(L.out_c, 8);
          const size_t zb = (size_t)B * L.out_h * L.out_w * L.z_ldc * 2;
          NET_CUDA(cudaMalloc(&L.z_buf, zb));
          NET_CUDA(cudaMemset(L.z_buf, 0, zb));
          const yc_tensor z = z_view(net, i);
          rc = yc_conv_plan_create(&L.plan, &d, &x, L.w_packed, nullptr, nullptr, nullptr, &z);
          if (!rc) rc = yc_conv_plan_set_stats(L.plan, L.stats_dev, L.stats_dev + L.n);
          // mean / variance / rolling averages come out of the same kernel (its last CTA): no yc_bn_finalize launch
          if (!rc) rc = yc_conv_plan_set_bn_finalize(L.plan, (long long)B * L.out_h * L.out_w, L.mean_dev, L.var_dev, L.rmean_dev, L.rvar_dev);
        } else {
          rc = yc_conv_plan_create(&L.plan, &d, &x, L.w_packed, L.bias_dev, nullptr, nullptr, &y);
        }
        if (rc) return rc;
        // backward plans: the GEMMs read dz = this layer's delta after the in-place BN/activation backward
        yc_tensor dz = delta_view(net, i);
        dz.c = L.delta_ldc;
        yc_conv_desc fd{G.size, G.stride, G.pad, L.dilation, YC_LINEAR, YC_LINEAR};
        rc = yc_conv_wgrad_plan_create(&L.wplan, &fd, &x, &dz, L.n, L.dw_packed);
        if (rc) return rc;
        if (i > 0) {
          yc_tensor dz1 = dz;
          if (L.stride > 1) {
            rc = yc_conv_dgrad_dilated_size(&fd, L.h, L.w, &L.dil_h, &L.dil_w);
            if (rc) return rc;
            const size_t db = (size_t)B * L.dil_h * L.dil_w * L.delta_ldc * 2;
            NET_CUDA(cudaMalloc(&L.dil_buf, db));
            NET_CUDA(cudaMemset(L.dil_buf, 0, db));
            dz1 = yc_tensor{L.dil_buf, B, L.dil_h, L.dil_w, L.delta_ldc, L.delta_ldc, YC_BF16};
          }
          const yc_tensor dx = delta_view(net, i - 1);
          rc = yc_conv_dgrad_plan_create(&L.dplan, &fd, &dz1, L.w_packed_d, &dx,
                                         (net->flags & YC_NET_DGRAD_ACCUMULATE) ? 1 : 0);
          if (rc) return rc;
        }
        break;
      }
      case L_MAXPOOL: {
        const size_t ib = (size_t)B * L.out_h * L.out_w * round_up(L.out_c, 8);
        NET_CUDA(cudaMalloc(&L.idx_buf, ib));
        NET_CUDA(cudaMemset(L.idx_buf, 255, ib));
        break;
      }
      case L_YOLO:
      case L_GYOLO: {
        const size_t elems = (size_t)B * L.out_c * L.out_h * L.out_w;
        NET_CUDA(cudaMalloc(&L.dec_dev, elems * 4));
        NET_CUDA(cudaHostAlloc(&L.dec_host, elems * 4, cudaHostAllocDefault));
        memset(L.dec_host, 0, elems * 4);
        NET_CUDA(cudaMalloc(&L.anchors_dev, L.anchors.size() * 4));
        NET_CUDA(cudaMemcpy(L.anchors_dev, L.anchors.data(), L.anchors.size() * 4, cudaMemcpyHostToDevice));
    NET_CUDA(h2d_done());
        NET_CUDA(cudaMalloc(&L.mask_dev, L.mask.size() * 4));
        NET_CUDA(cudaMemcpy(L.mask_dev, L.mask.data(), L.mask.size() * 4, cudaMemcpyHostToDevice));
    NET_CUDA(h2d_done());
        int rc = dev_alloc(&L.ydelta, elems);
        if (!rc) rc = dev_alloc(&L.ystats, yc_yolo_loss_stats_floats(B, L.out_h, L.out_w));
        if (rc) return rc;
        net->truth_floats = std::max(net->truth_floats, L.max_boxes * 5);
        if (i == 0 || Ls[i - 1].type != L_CONV) NET_FAIL(YC_ERR_ARG, "layer %d: a [yolo] layer must follow its head conv", i);
        break;
      }
      default: break;
    }
  }
  if (net->truth_floats == 0) NET_FAIL(YC_ERR_ARG, "training needs at least one [yolo] layer");
  NET_CUDA(cudaMalloc(&net->truth_dev, (size_t)B * net->truth_floats * 4));
  NET_CUDA(cudaHostAlloc(&net->truth_pinned, (size_t)B * net->truth_floats * 4, cudaHostAllocDefault));
  NET_CUDA(cudaHostAlloc(&net->cost_host, net->yolo_layers.size() * 8 * 4, cudaHostAllocDefault));
  NET_CUDA(cudaStreamSynchronize(net->stream));
  return YC_OK;
}

namespace {

// Optional per-category device timing of one step (yc_network_train_profile): an event after every operator; the
// time since the previous event belongs to that operator's category.
enum { CAT_FPROP = 0, CAT_WGRAD = 1, CAT_DGRAD = 2, CAT_BN = 3, CAT_OTHER = 4, CAT_COUNT = 5 };
struct Prof {
  std::vector<cudaEvent_t> ev;
  std::vector<int> cat;
};
Prof* g_prof = nullptr;
inline void tick(int cat, cudaStream_t st) {
  if (!g_prof) return;
  cudaEvent_t e;
  cudaEventCreate(&e);
  cudaEventRecord(e, st);
  g_prof->ev.push_back(e);
  g_prof->cat.push_back(cat);
}

// ForwardNetworkGpu with state.train (network_kernels.cu:45-114)
int train_forward(yc_network* net, const float* input_dev, const float* truth_dev, cudaStream_t st) {
  const int nl = (int)net->layers.size();
  auto& Ls = net->layers;
  const int B = net->batch;
  NET_CUDA(cudaMemsetAsync(net->delta_slab, 0, net->delta_slab_bytes, st));
  {
    const yc_tensor in = in_view(net, 0);
    int rc = net->stem ? yc_stem_im2col(input_dev, net->c, net->h, net->w, Ls[0].size, Ls[0].stride, Ls[0].pad, &in, st)
                       : yc_nchw_to_nhwc(input_dev, &in, st);
    if (rc) return rc;
    tick(CAT_OTHER, st);
  }
  for (int i = 0; i < nl; ++i) {
    NLayer& L = Ls[i];
    int rc = YC_OK;
    switch (L.type) {
      case L_CONV: {
        // (the statistics kernel OVERWRITES sum / sum of squares from its ordered partials: no memset per layer)
        rc = yc_conv_fprop(L.plan, st);
        tick(CAT_FPROP, st);
        if (!rc && L.bn) {
          const yc_tensor z = z_view(net, i), y = view_of(net, i);
          if (!rc) rc = yc_bn_act_fwd(&z, L.mean_dev, L.var_dev, L.scale_dev, L.bias_dev, L.act, &y, st);
          tick(CAT_BN, st);
        }
        break;
      }
      case L_MAXPOOL: {
        const yc_tensor x = in_view(net, i), y = view_of(net, i);
        rc = yc_maxpool_fwd_train(&x, &y, L.idx_buf, L.size, L.stride, L.pad, st);
        break;
      }
      case L_UPSAMPLE: {
        const yc_tensor x = in_view(net, i), y = view_of(net, i);
        rc = yc_upsample_fwd(&x, &y, L.stride, L.scale_x_y, st);
        break;
      }
      case L_SHORTCUT: {
        const yc_tensor a = in_view(net, i), b = view_of(net, L.from), y = view_of(net, i);
        rc = yc_shortcut_fwd(&a, &b, &y, L.act, st);
        break;
      }
      case L_ROUTE: {
        if (L.alias) break;
        for (auto& cp : L.copies) {
          const NLayer& S = Ls[cp.first];
          const int part = S.out_c / L.groups;
          yc_tensor a = view_of(net, cp.first);
          a.ptr = (char*)a.ptr + (size_t)part * L.group_id * 2;
          a.c = part;
          yc_tensor y = view_of(net, i);
          y.ptr = (char*)y.ptr + (size_t)cp.second * 2;
          y.c = part;
          rc = yc_route_copy(&a, &y, st);
          if (rc) return rc;
        }
        break;
      }
      case L_YOLO:
      case L_GYOLO: {
        const yc_tensor x = in_view(net, i);
        rc = yc_yolo_decode(&x, L.dec_dev, L.n, L.classes, L.scale_x_y, L.type == L_GYOLO ? 1 : 0, st);
        if (rc) return rc;
        yc_yolo_loss_desc d;
        d.anchors = L.n; d.total = L.total; d.classes = L.classes; d.max_boxes = L.max_boxes;
        d.net_w = net->w; d.net_h = net->h; d.iou_loss = L.iou_loss; d.iou_thresh_kind = L.iou_thresh_kind;
        d.focal_loss = L.focal_loss; d.ignore_thresh = L.ignore_thresh; d.truth_thresh = L.truth_thresh;
        d.iou_thresh = L.iou_thresh; d.iou_normalizer = L.iou_normalizer; d.cls_normalizer = L.cls_normalizer;
        d.label_smooth_eps = L.label_smooth_eps; d.max_delta = L.max_delta;
        d.gaussian = L.type == L_GYOLO ? 1 : 0; d.uc_normalizer = L.uc_normalizer;
        const yc_tensor pd = delta_view(net, i - 1);
        // BackwardYoloLayerGpu (yolo_layer.cpp:884-888) is folded into the loss kernel: the head conv's delta was
        // zeroed at the start of the step and has no other contributor
        rc = yc_yolo_loss(&d, L.dec_dev, truth_dev, L.anchors_dev, L.mask_dev, B, L.out_h, L.out_w, L.ydelta, &pd,
                          L.ystats, st);
        break;
      }
      default: rc = YC_ERR_ARG;
    }
    if (rc) return rc;
    if (L.type != L_CONV) tick(CAT_OTHER, st);
  }
  return YC_OK;
}

// BackwardNetworkGpu (network_kernels.cu:116-243): last layer to first; state.delta = layers[i-1].delta_gpu
int train_backward(yc_network* net, cudaStream_t st) {
  const int nl = (int)net->layers.size();
  auto& Ls = net->layers;
  for (int i = nl - 1; i >= 0; --i) {
    NLayer& L = Ls[i];
    // network_kernels.cu:140-143: stopbackward = 1 ends the backward pass at this layer for good, stopbackward = k > 1
    // until iteration k; :158 onlyforward layers are skipped
    if (L.stopbackward == 1 || L.stopbackward > net->curr_iter) break;
    if (L.onlyforward) continue;
    int rc = YC_OK;
    switch (L.type) {
      case L_CONV: {
        const yc_tensor dy = delta_view(net, i);
        yc_tensor dz = dy;
        dz.c = L.delta_ldc;
        if (L.bn) {
          const yc_tensor z = z_view(net, i);
          rc = yc_bn_bwd(&dy, &z, L.mean_dev, L.var_dev, L.scale_dev, L.bias_dev, L.act, L.scale_upd,
                         (net->flags & YC_NET_BN_BIAS_GRAD) ? L.bias_upd : nullptr, L.ws_dev, &dz, st);
        } else {
          const yc_tensor y = view_of(net, i);
          rc = yc_bias_act_bwd(&dy, L.act == YC_LINEAR ? nullptr : &y, L.act, L.bias_upd, L.ws_dev, &dz, st);
        }
        tick(CAT_BN, st);
        if (!rc) rc = yc_conv_wgrad(L.wplan, st);
        tick(CAT_WGRAD, st);
        if (!rc && net->dp) rc = yc_dp_layer_ready(net, i, st);  // all-reduce of this layer's bucket overlaps the rest
        if (!rc && i > 0) {
          if (L.stride > 1) {
            const yc_tensor up{L.dil_buf, net->batch, L.dil_h, L.dil_w, L.delta_ldc, L.delta_ldc, YC_BF16};
            rc = yc_conv_dgrad_dilate(&dz, L.stride, &up, st);
          }
          if (!rc) rc = yc_conv_fprop(L.dplan, st);
          tick(CAT_DGRAD, st);
        }
        break;
      }
      case L_MAXPOOL: {
        const yc_tensor dy = delta_view(net, i), dx = delta_view(net, i - 1);
        rc = yc_maxpool_bwd(&dy, L.idx_buf, &dx, L.size, L.stride, L.pad, st);
        break;
      }
      case L_UPSAMPLE: {
        const yc_tensor dy = delta_view(net, i), dx = delta_view(net, i - 1);
        rc = yc_upsample_bwd(&dy, &dx, L.stride, L.scale_x_y, st);
        break;
      }
      case L_SHORTCUT: {
        const yc_tensor dy = delta_view(net, i), y = view_of(net, i), dp = delta_view(net, i - 1), df = delta_view(net, L.from);
        rc = yc_shortcut_bwd(&dy, L.act == YC_LINEAR ? nullptr : &y, L.act, &dp, &df, st);
        break;
      }
      case L_ROUTE: {  // BackwardRouteLayer (route_layer.c:106-122)
        int offset = 0;
        for (int in : L.inputs) {
          const NLayer& S = Ls[in];
          const int part = S.out_c / L.groups;
          if (S.type != L_YOLO && S.type != L_GYOLO) {
            yc_tensor src = delta_view(net, i);
            src.ptr = (char*)src.ptr + (size_t)offset * 2;
            src.c = part;
            yc_tensor dst = delta_view(net, in);
            dst.ptr = (char*)dst.ptr + (size_t)part * L.group_id * 2;
            dst.c = part;
            if (offset % 8 || (part * L.group_id) % 8) NET_FAIL(YC_ERR_ARG, "layer %d: route slice not 16-byte aligned", i);
            rc = yc_axpy_view(&src, &dst, st);
            if (rc) return rc;
          }
          offset += part;
        }
        break;
      }
      default: break;  // yolo: folded into the forward's loss kernel
    }
    if (rc) return rc;
    if (L.type != L_CONV) tick(CAT_OTHER, st);
  }
  if (net->dp) return yc_dp_backward_done(net, st);
  return YC_OK;
}

int pull_costs(yc_network* net, cudaStream_t st, float* cost) {
  int k = 0;
  for (int li : net->yolo_layers)
    NET_CUDA(cudaMemcpyAsync(net->cost_host + 8 * (k++), net->layers[li].ystats, 32, cudaMemcpyDeviceToHost, st));
  NET_CUDA(cudaStreamSynchronize(st));
  double sum = 0;  // GetNetworkCost network.cpp:146-160: mean over the layers that own a cost
  for (size_t j = 0; j < net->yolo_layers.size(); ++j) sum += net->cost_host[8 * j];
  if (cost) *cost = (float)(sum / std::max<size_t>(1, net->yolo_layers.size()));
  return YC_OK;
}

}  // namespace

// ForwardBackwardNetworkGpu on device-resident inputs (network_kernels.cu:245-293), asynchronous on `stream`.
extern "C" int yc_network_forward_backward(yc_network* net, const float* input_dev_nchw, const float* truth_dev,
                                           void* stream) {
  YC_REQUIRE(net && input_dev_nchw && truth_dev, "yc_network_forward_backward: null argument");
  YC_REQUIRE(net->flags & YC_NET_TRAIN, "yc_network_forward_backward: network was not loaded with YC_NET_TRAIN");
  if (!net->finalized) {
    int rc = yc_network_finalize(net);
    if (rc) return rc;
  }
  // NULL = the network's own stream, as in yc_network_update (round 2: NULL used to mean the legacy default stream HERE and
  // the network's stream THERE, so a caller passing NULL to both — bench.py did, through torch's default stream — had the
  // update racing with the backward pass it follows; found through a NaN cost in the bench's host-fed leg)
  cudaStream_t st = stream ? (cudaStream_t)stream : net->stream;
  int rc = train_forward(net, input_dev_nchw, truth_dev, st);
  if (!rc) rc = train_backward(net, st);
  if (!rc) net->seen += net->batch;
  return rc;
}

// TrainNetworkDatum (network.cpp:188-208 / TrainNetworkDatumGpu network_kernels.cu:269-293): host x and truth,
// forward + backward, returns GetNetworkCost in *cost.  Synchronous.
extern "C" int yc_network_train_datum(yc_network* net, const float* x_host, const float* truth_host, float* cost) {
  YC_REQUIRE(net && x_host && truth_host, "yc_network_train_datum: null argument");
  YC_REQUIRE(net->flags & YC_NET_TRAIN, "yc_network_train_datum: network was not loaded with YC_NET_TRAIN");
  if (!net->finalized) {
    int rc = yc_network_finalize(net);
    if (rc) return rc;
  }
  const size_t in_bytes = (size_t)net->batch * net->c * net->h * net->w * 4;
  const size_t t_bytes = (size_t)net->batch * net->truth_floats * 4;
  if (x_host != net->input_pinned) memcpy(net->input_pinned, x_host, in_bytes);
  memcpy(net->truth_pinned, truth_host, t_bytes);
  NET_CUDA(cudaMemcpyAsync(net->input_dev, net->input_pinned, in_bytes, cudaMemcpyHostToDevice, net->stream));
  NET_CUDA(cudaMemcpyAsync(net->truth_dev, net->truth_pinned, t_bytes, cudaMemcpyHostToDevice, net->stream));
  int rc = yc_network_forward_backward(net, net->input_dev, net->truth_dev, net->stream);
  if (rc) return rc;
  return pull_costs(net, net->stream, cost);
}

// UpdateNetworkGpu (network_kernels.cu:214-243) -> UpdateConvolutionalLayerGpu per layer.  lr/momentum/decay < 0 take
// the cfg's [net] values; batch = net->batch * subdivisions (network.cpp:118).
extern "C" int yc_network_update(yc_network* net, float lr, float momentum, float decay, void* stream) {
  YC_REQUIRE(net && (net->flags & YC_NET_TRAIN) && net->finalized, "yc_network_update: not a finalized training network");
  // TrainNetwork (network.cpp:227-231) advances curr_iter BEFORE UpdateNetworkGpu: the k-th update of a run sees iter = k
  // in GetCurrLr, in the update gates and as Adam's t (l->t = iter, network_kernels.cu:229).
  const long long iter = net->curr_iter + 1;
  if (lr < 0) lr = yc_network_current_rate(net, iter, 0);  // UpdateNetworkGpu: lr = GetCurrLr(net)
  if (momentum < 0) momentum = net->momentum;
  if (decay < 0) decay = net->decay;
  const int world = std::max(1, net->train_flags);  // data-parallel world size (network_dp.cu)
  const int batch = net->batch * net->subdivisions * world;
  // Under data parallelism every replica holds the all-reduced (global) gradient sum in its accumulator.  The next
  // step's all-reduce sums the accumulators again, so the momentum carry-over is stored divided by the world size:
  // sum_r (momentum/world * G + g_r) = momentum * G + sum_r g_r, the single-GPU recurrence on the global batch.
  const float carry = momentum / (float)world;
  cudaStream_t st = stream ? (cudaStream_t)stream : net->stream;
  if (net->dp) {  // gradients not yet exchanged (subdivisions > 1 or overlap switched off): one grouped all-reduce now
    int rc = yc_dp_reduce_pending(net, st);
    if (rc) return rc;
  }
  // one launch for every layer's weights, biases and scales (the same arithmetic as yc_sgd_update_conv per layer:
  // 61 launches of a few microseconds each on yolov4-tiny otherwise).  The table is built once — or, when some layer
  // carries an update gate, per call from the layers that update at this iteration (network_kernels.cu:231-241).
  if (!net->sgd_items || net->any_gate) {
    std::vector<yc::SgdItem> items;
    int blocks = 0;
    auto push = [&](yc::SgdItem it, long long elements) {
      it.first_block = blocks;
      (void)elements;
      it.blocks = yc::sgd_item_blocks(it);
      blocks += it.blocks;
      items.push_back(it);
    };
    for (NLayer& L : net->layers) {
      if (L.type != L_CONV) continue;
      if (L.burnin_update && (long long)L.burnin_update * net->burn_in > iter) continue;
      if (L.train_only_bn) continue;
      if (!(L.dont_update < iter) && L.dont_update) continue;  // the reference updates when l->dont_update < iter
      const Geo G = geo(net, (int)(&L - net->layers.data()));
      const int taps = G.size * G.size;
      // Adam moments mirror the gradient bucket: same offsets as dw_packed / bias_upd / scale_upd
      auto mom = [&](float* base, const float* upd) -> float* { return base ? base + (upd - L.dw_packed) : nullptr; };
      yc::SgdItem w{L.w_master, L.dw_packed, (__nv_bfloat16*)L.w_packed, (__nv_bfloat16*)L.w_packed_d, L.n, G.c, taps,
                    yc_conv_wgrad_ld(G.c), yc_conv_padded_in_c(G.c), yc_conv_padded_in_c(yc_conv_dgrad_channels(L.n)), 0, 0,
                    L.adam_m, L.adam_v};
      push(w, (long long)L.n * taps * G.c);
      if (L.bias_dev && L.bias_upd)
        push(yc::SgdItem{L.bias_dev, L.bias_upd, nullptr, nullptr, L.n, 0, 0, 0, 0, 0, 0, 0, mom(L.adam_m, L.bias_upd),
                         mom(L.adam_v, L.bias_upd)}, L.n);
      if (L.bn && L.scale_dev && L.scale_upd)
        push(yc::SgdItem{L.scale_dev, L.scale_upd, nullptr, nullptr, L.n, 0, 0, 0, 0, 0, 0, 0, mom(L.adam_m, L.scale_upd),
                         mom(L.adam_v, L.scale_upd)}, L.n);
    }
    if (net->sgd_items) {  // gated networks only: the previous update may still be reading the table
      YC_CUDA(cudaStreamSynchronize(st));
      cudaFree(net->sgd_items);
      net->sgd_items = nullptr;
    }
    if (!items.empty()) {
      YC_CUDA(cudaMalloc(&net->sgd_items, items.size() * sizeof(yc::SgdItem)));
      YC_CUDA(cudaMemcpy(net->sgd_items, items.data(), items.size() * sizeof(yc::SgdItem), cudaMemcpyHostToDevice));
      h2d_done();
    }
    net->sgd_n = (int)items.size();
    net->sgd_blocks = blocks;
  }
  net->curr_iter = iter;
  const yc::UpdRule rule{lr / batch, decay * batch, carry, net->adam, (int)std::min<long long>(iter, 0x7fffffff), net->B1, net->B2,
                         net->adam_eps, lr};
  int rc = yc::sgd_update_multi(net->sgd_items, net->sgd_n, net->sgd_blocks, rule, st);
  if (net->dp) yc_dp_update_done(net);  // the reduced gradients are consumed: the next backward pass starts a new step
  // layers without batchnorm (the YOLO heads) add their bias in the conv epilogue, from the plan's own padded copy: it
  // follows the update (found by tests/test_resize_gpu.py: a re-planned network saw the new biases, the running one did not)
  for (NLayer& L : net->layers)
    if (!rc && L.type == L_CONV && !L.bn && L.plan && L.bias_dev && L.bias_upd) rc = yc_conv_plan_set_bias(L.plan, L.bias_dev, st);
  return rc;
}

// PullConvolutionalLayer for every conv layer (convolutional_kernels.cu:817-836): device master copies -> the host-side
// parameter vectors that yc_network_save_weights / yc_network_finalize read.
int yc_train_pull_params(yc_network* net) {
  cudaStream_t st = net->stream;
  for (size_t i = 0; i < net->layers.size(); ++i) {
    NLayer& L = net->layers[i];
    if (L.type != L_CONV || !L.w_master) continue;
    if (i == 0 && net->stem) {
      std::vector<float> h((size_t)L.n * 32);
      NET_CUDA(cudaMemcpyAsync(h.data(), L.w_master, h.size() * 4, cudaMemcpyDeviceToHost, st));
      NET_CUDA(cudaStreamSynchronize(st));
      stem_permute(L, h.data(), L.weights.data(), false);
    } else {
      NET_CUDA(cudaMemcpyAsync(L.weights.data(), L.w_master, L.weights.size() * 4, cudaMemcpyDeviceToHost, st));
    }
    NET_CUDA(cudaMemcpyAsync(L.biases.data(), L.bias_dev, (size_t)L.n * 4, cudaMemcpyDeviceToHost, st));
    if (L.bn) {
      NET_CUDA(cudaMemcpyAsync(L.scales.data(), L.scale_dev, (size_t)L.n * 4, cudaMemcpyDeviceToHost, st));
      NET_CUDA(cudaMemcpyAsync(L.rmean.data(), L.rmean_dev, (size_t)L.n * 4, cudaMemcpyDeviceToHost, st));
      NET_CUDA(cudaMemcpyAsync(L.rvar.data(), L.rvar_dev, (size_t)L.n * 4, cudaMemcpyDeviceToHost, st));
    }
  }
  NET_CUDA(cudaStreamSynchronize(st));
  return YC_OK;
}

// Device time of one forward+backward(+update) by operator category, averaged over iters:
// ms[0] conv fprop, [1] wgrad, [2] dgrad (+ zero insertion), [3] batchnorm / activation passes, [4] everything else
// (input conversion, pooling, routes, YOLO decode + loss, memsets), [5] SGD update, [6] total.
extern "C" int yc_network_train_profile(yc_network* net, const float* input_dev_nchw, const float* truth_dev, int iters,
                                        float* ms) {
  YC_REQUIRE(net && input_dev_nchw && truth_dev && ms && iters > 0 && (net->flags & YC_NET_TRAIN) && net->finalized,
             "yc_network_train_profile: bad argument");
  for (int i = 0; i < 7; ++i) ms[i] = 0.f;
  cudaStream_t st = net->stream;
  int rc = YC_OK;
  const long long iter0 = net->curr_iter, seen0 = net->seen;  // a timing pass is not a training iteration
  // ... nor may it leave traces in the training state: the passes below accumulate gradients, scale the accumulators by the
  // momentum (update with lr = 0) and move the rolling batchnorm statistics — snapshot both, restore afterwards
  std::vector<float*> snap_b(net->layers.size(), nullptr), snap_r(net->layers.size(), nullptr);
  for (size_t i = 0; i < net->layers.size(); ++i) {
    NLayer& L = net->layers[i];
    if (L.type != L_CONV) continue;
    if (L.dw_packed && L.bucket_floats) {
      NET_CUDA(cudaMalloc(&snap_b[i], L.bucket_floats * 4));
      NET_CUDA(cudaMemcpyAsync(snap_b[i], L.dw_packed, L.bucket_floats * 4, cudaMemcpyDeviceToDevice, st));
    }
    if (L.bn && L.rmean_dev && L.rvar_dev) {
      NET_CUDA(cudaMalloc(&snap_r[i], (size_t)2 * L.n * 4));
      NET_CUDA(cudaMemcpyAsync(snap_r[i], L.rmean_dev, (size_t)L.n * 4, cudaMemcpyDeviceToDevice, st));
      NET_CUDA(cudaMemcpyAsync(snap_r[i] + L.n, L.rvar_dev, (size_t)L.n * 4, cudaMemcpyDeviceToDevice, st));
    }
  }
  for (int it = 0; it < iters && !rc; ++it) {
    Prof prof;
    g_prof = &prof;
    tick(CAT_OTHER, st);
    rc = train_forward(net, input_dev_nchw, truth_dev, st);
    if (!rc) rc = train_backward(net, st);
    const size_t n_fb = prof.ev.size();
    if (!rc) rc = yc_network_update(net, 0.f, -1.f, 0.f, st);  // lr = 0, decay = 0: timing only, weights unchanged
    tick(5, st);
    g_prof = nullptr;
    cudaStreamSynchronize(st);
    for (size_t k = 1; k < prof.ev.size(); ++k) {
      float t = 0.f;
      cudaEventElapsedTime(&t, prof.ev[k - 1], prof.ev[k]);
      const int c = k >= n_fb ? 5 : prof.cat[k];
      ms[c] += t / iters;
      ms[6] += t / iters;
    }
    for (cudaEvent_t e : prof.ev) cudaEventDestroy(e);
  }
  g_prof = nullptr;
  net->curr_iter = iter0;
  net->seen = seen0;
  for (size_t i = 0; i < net->layers.size(); ++i) {
    NLayer& L = net->layers[i];
    if (snap_b[i]) cudaMemcpyAsync(L.dw_packed, snap_b[i], L.bucket_floats * 4, cudaMemcpyDeviceToDevice, st);
    if (snap_r[i]) {
      cudaMemcpyAsync(L.rmean_dev, snap_r[i], (size_t)L.n * 4, cudaMemcpyDeviceToDevice, st);
      cudaMemcpyAsync(L.rvar_dev, snap_r[i] + L.n, (size_t)L.n * 4, cudaMemcpyDeviceToDevice, st);
    }
  }
  cudaStreamSynchronize(st);
  for (float* q : snap_b) if (q) cudaFree(q);
  for (float* q : snap_r) if (q) cudaFree(q);
  if (net->dp) yc_dp_update_done(net);
  return rc;
}

extern "C" int yc_network_train_info(const yc_network* net, int what) {
  if (!net) return -1;
  switch (what) {
    case 0: return net->truth_floats;
    case 1: return net->subdivisions;
    case 2: return (int)(net->delta_slab_bytes >> 20);
  }
  return -1;
}

static int pull_nchw(yc_network* net, const yc_tensor v, size_t elems, float* dst_host, const char* who) {
  float* tmp = nullptr;
  NET_CUDA(cudaMalloc(&tmp, elems * 4));
  int rc = yc_nhwc_to_nchw(&v, tmp, net->stream);
  if (!rc) {
    cudaError_t e = cudaMemcpyAsync(dst_host, tmp, elems * 4, cudaMemcpyDeviceToHost, net->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(net->stream);
    if (e != cudaSuccess) { set_error("%s: %s", who, cudaGetErrorString(e)); rc = YC_ERR_CUDA; }
  }
  cudaFree(tmp);
  return rc;
}

// cuda_pull_array(l->delta_gpu): layer delta as fp32 NCHW on the host
extern "C" int yc_network_layer_delta(yc_network* net, int layer, float* dst_host) {
  YC_REQUIRE(net && dst_host && layer >= 0 && layer < (int)net->layers.size() && (net->flags & YC_NET_TRAIN) && net->finalized,
             "yc_network_layer_delta: bad argument");
  NLayer& L = net->layers[layer];
  const size_t elems = (size_t)net->batch * L.out_c * L.out_h * L.out_w;
  if (L.type == L_YOLO || L.type == L_GYOLO) {
    NET_CUDA(cudaMemcpyAsync(dst_host, L.ydelta, elems * 4, cudaMemcpyDeviceToHost, net->stream));
    NET_CUDA(cudaStreamSynchronize(net->stream));
    return YC_OK;
  }
  return pull_nchw(net, delta_view(net, layer), elems, dst_host, "yc_network_layer_delta");
}

// cuda_pull_array(l->x_gpu): the convolution output before batchnorm (batchnorm_layer.cpp:104 keeps it for backward)
extern "C" int yc_network_layer_x(yc_network* net, int layer, float* dst_host) {
  YC_REQUIRE(net && dst_host && layer >= 0 && layer < (int)net->layers.size() && (net->flags & YC_NET_TRAIN) && net->finalized,
             "yc_network_layer_x: bad argument");
  NLayer& L = net->layers[layer];
  YC_REQUIRE(L.type == L_CONV && L.bn && L.z_buf, "yc_network_layer_x: not a batch-normalised convolution");
  return pull_nchw(net, z_view(net, layer), (size_t)net->batch * L.out_c * L.out_h * L.out_w, dst_host, "yc_network_layer_x");
}

// PullConvolutionalLayer (convolutional_kernels.cu:817-836) by name: weights, biases, scales, rolling_mean,
// rolling_variance, mean, variance, weight_updates, bias_updates, scale_updates (conv); cost, count (yolo).
// Returns the number of floats written (<= capacity) or a negative status.
extern "C" int yc_network_get_param(yc_network* net, int layer, const char* key, float* dst_host, int capacity) {
  YC_REQUIRE(net && key && dst_host && layer >= 0 && layer < (int)net->layers.size() && (net->flags & YC_NET_TRAIN) && net->finalized,
             "yc_network_get_param: bad argument");
  NLayer& L = net->layers[layer];
  const std::string k = key;
  const float* src = nullptr;
  size_t count = 0;
  float* tmp = nullptr;
  if (L.type == L_CONV && layer == 0 && net->stem && (k == "weights" || k == "weight_updates")) {
    // the stem keeps its weights / gradients in im2col column order [n][32]: hand them out as OIHW
    const size_t nw = (size_t)L.n * L.c * L.size * L.size, ns = (size_t)L.n * 32;
    if ((size_t)capacity < nw) {
      set_error("yc_network_get_param: capacity %d < %zu", capacity, nw);
      return YC_ERR_ARG;
    }
    std::vector<float> h(ns);
    const float* from = L.w_master;
    if (k == "weight_updates") {
      NET_CUDA(cudaMalloc(&tmp, ns * 4));
      int rc = yc_conv_wgrad_unpack(L.dw_packed, L.n, 32, 1, tmp, net->stream);
      if (rc) { cudaFree(tmp); return rc; }
      from = tmp;
    }
    cudaError_t e = cudaMemcpyAsync(h.data(), from, ns * 4, cudaMemcpyDeviceToHost, net->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(net->stream);
    if (tmp) cudaFree(tmp);
    if (e != cudaSuccess) {
      set_error("yc_network_get_param: %s", cudaGetErrorString(e));
      return YC_ERR_CUDA;
    }
    stem_permute(L, h.data(), dst_host, false);
    return (int)nw;
  }
  if (L.type == L_CONV) {
    const size_t nw = (size_t)L.n * L.c * L.size * L.size;
    if (k == "weights") { src = L.w_master; count = nw; }
    else if (k == "biases") { src = L.bias_dev; count = L.n; }
    else if (k == "bias_updates") { src = L.bias_upd; count = L.n; }
    else if (k == "weight_updates") {
      NET_CUDA(cudaMalloc(&tmp, nw * 4));
      int rc = yc_conv_wgrad_unpack(L.dw_packed, L.n, L.c, L.size, tmp, net->stream);
      if (rc) { cudaFree(tmp); return rc; }
      src = tmp; count = nw;
    } else if (L.bn) {
      if (k == "scales") { src = L.scale_dev; count = L.n; }
      else if (k == "rolling_mean") { src = L.rmean_dev; count = L.n; }
      else if (k == "rolling_variance") { src = L.rvar_dev; count = L.n; }
      else if (k == "mean") { src = L.mean_dev; count = L.n; }
      else if (k == "variance") { src = L.var_dev; count = L.n; }
      else if (k == "scale_updates") { src = L.scale_upd; count = L.n; }
    }
  } else if (L.type == L_YOLO || L.type == L_GYOLO) {
    if (k == "cost") { src = L.ystats; count = 1; }
    else if (k == "stats") { src = L.ystats; count = 8; }
  }
  if (!src) {
    set_error("yc_network_get_param: layer %d has no '%s'", layer, key);
    return YC_ERR_ARG;
  }
  if ((size_t)capacity < count) {
    if (tmp) cudaFree(tmp);
    set_error("yc_network_get_param: capacity %d < %zu", capacity, count);
    return YC_ERR_ARG;
  }
  cudaError_t e = cudaMemcpyAsync(dst_host, src, count * 4, cudaMemcpyDeviceToHost, net->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(net->stream);
  if (tmp) cudaFree(tmp);
  if (e != cudaSuccess) {
    set_error("yc_network_get_param: %s", cudaGetErrorString(e));
    return YC_ERR_CUDA;
  }
  return (int)count;
}
