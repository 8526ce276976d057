// YOLO / Gaussian-YOLO loss and deltas on the device.
//
// Replaces the training half of ForwardYoloLayer (yolo_layer.cpp:409-772) / ForwardGaussianYoloLayer
// (gaussian_yolo_layer.cpp:486-857) and BackwardYoloLayerGpu (:884-888).  In the reference this is ONE host thread between
// a device->host copy of the head and a host->device copy of the deltas; here the head never leaves HBM and the work is
// spread as follows:
//
//   objectness pass : thread per (image, anchor, cell), one image per blockIdx.y so that the image's truth list sits in
//                     shared memory: best IoU over the truths -> objectness delta (ignore_thresh / truth_thresh)
//   assignment pass : block per image.  Every (truth, candidate anchor) pair becomes a RECORD in shared memory, computed
//                     in parallel (thread per truth).  The reference walks the truths in order and several truths may
//                     land on the same (anchor, cell): box deltas then accumulate, class deltas depend on what an
//                     earlier truth left.  That order is restored by a resolve step: the first record of every
//                     location owns it and applies all records of that location in record order.  Different locations
//                     proceed in parallel; nothing depends on thread scheduling.
//   finish pass     : block per (image, 32 cells): box deltas / number of positive classes (averages_yolo_deltas), the
//                     sums of squares of the cost, and prev.delta += delta straight into the head conv's NHWC bf16 delta
//   cost pass       : one block adds the per-image and per-block partial sums IN INDEX ORDER (no float atomics: two runs
//                     give the same bits) and forms *(l->cost)
//
// The arithmetic of every formula follows the reference's CPU code operation by operation (its float / double mix
// included), because the deltas are compared with it at 2e-5; what is this file's own is how the formulas are organised:
// one edge representation per box pair, the IoU-family gradient as a template over the loss kind with every shared
// sub-expression formed once, and per-edge derivative tables instead of sixteen named scalars.
//
// Layouts: `out` / `delta` are fp32 NCHW [B][n*(E)][H][W] like l->output / l->delta (EntryIndex, yolo_layer.cpp:380-386);
// truth is [B][max_boxes][5] = (x, y, w, h, class id), ended by the first entry with a valid class and x == 0.
#include <float.h>
#include <math.h>
#include <algorithm>

#include "yc_common.cuh"

namespace yc {

enum LossKind { K_IOU = 0, K_GIOU = 1, K_MSE = 2, K_DIOU = 3, K_CIOU = 4 };  // IOU_LOSS, box.h:6-13

struct Box4 { float x, y, w, h; };
struct Edges { float l, r, t, b; };
__device__ __forceinline__ Edges edges_of(const Box4& q) {
  return Edges{q.x - q.w / 2.0f, q.x + q.w / 2.0f, q.y - q.h / 2.0f, q.y + q.h / 2.0f};
}
__device__ __forceinline__ float sq(float a) { return a * a; }

// ---- overlap measures of a box pair (box.cpp:36-128), all from one evaluation of intersection and union ------------
struct Overlap {
  float inter, uni, iou;
  float enc_w, enc_h;  // smallest enclosing box
};
__device__ __forceinline__ float span_overlap(float c1, float e1, float c2, float e2) {  // Box::Overlap
  const float lo = fmaxf(c1 - e1 / 2, c2 - e2 / 2), hi = fminf(c1 + e1 / 2, c2 + e2 / 2);
  return hi - lo;
}
__device__ __forceinline__ Overlap overlap_of(const Box4& a, const Box4& b) {
  Overlap o;
  const float w = span_overlap(a.x, a.w, b.x, b.w), h = span_overlap(a.y, a.h, b.y, b.h);
  o.inter = (w < 0 || h < 0) ? 0.f : w * h;
  o.uni = a.w * a.h + b.w * b.h - o.inter;
  o.iou = (fabsf(o.inter) < FLT_EPSILON || fabsf(o.uni) < FLT_EPSILON) ? 0.f : o.inter / o.uni;
  const Edges ea = edges_of(a), eb = edges_of(b);
  o.enc_w = fmaxf(ea.r, eb.r) - fminf(ea.l, eb.l);
  o.enc_h = fmaxf(ea.b, eb.b) - fminf(ea.t, eb.t);
  return o;
}
__device__ __forceinline__ float iou_of(const Box4& a, const Box4& b) { return overlap_of(a, b).iou; }
__device__ __forceinline__ float giou_of(const Overlap& o) {  // Box::Giou
  const float c = o.enc_w * o.enc_h;
  return fabsf(c) < FLT_EPSILON ? o.iou : o.iou - (c - o.uni) / c;
}
__device__ __forceinline__ float diou_of(const Overlap& o, const Box4& a, const Box4& b, float beta) {  // Box::Diou
  const float c = sq(o.enc_w) + sq(o.enc_h);
  if (fabsf(c) < FLT_EPSILON) return o.iou;
  const float d = sq(a.x - b.x) + sq(a.y - b.y);
  return o.iou - (float)pow((double)(d / c), (double)beta);
}
__device__ __forceinline__ float ciou_of(const Overlap& o, const Box4& a, const Box4& b) {  // Box::Ciou
  const float c = sq(o.enc_w) + sq(o.enc_h);
  if (fabsf(c) < FLT_EPSILON) return o.iou;
  const float d = (sq(a.x - b.x) + sq(a.y - b.y)) / c;
  const float ar_gt = b.w / b.h, ar_pred = a.w / a.h;
  const float ar_loss = (float)(4 / (M_PI * M_PI) * sq(atan((double)ar_gt) - atan((double)ar_pred)));
  const float alpha = (float)(ar_loss / (1 - o.iou + ar_loss + 0.000001));
  return o.iou - (d + alpha * ar_loss);
}
__device__ float measure_of(const Box4& a, const Box4& b, int kind) {  // Box::Iou(b1, b2, iou_type), box.cpp:136-151
  if (kind == K_MSE) return sqrtf(sq(a.x - b.x) + sq(a.y - b.y) + sq(a.w - b.w) + sq(a.h - b.h));
  const Overlap o = overlap_of(a, b);
  switch (kind) {
    case K_GIOU: return giou_of(o);
    case K_DIOU: return diou_of(o, a, b, 0.6f);
    case K_CIOU: return ciou_of(o, a, b);
  }
  return o.iou;
}

// ---- gradient of IoU / GIoU / DIoU / CIoU with respect to (x, y, w, h) of the prediction (Box::DxIou, box.cpp:153-351) --
// Edge order of the tables: 0 top, 1 bottom, 2 left, 3 right.
template <int KIND>
__device__ float4 iou_family_grad(const Box4& pred, const Box4& gt) {
  const Edges pe = edges_of(pred), ge = edges_of(gt);
  // a prediction with negative width / height has its edges swapped: the gradient is taken on the sorted edges
  const float pt = fminf(pe.t, pe.b), pb = fmaxf(pe.t, pe.b), pl = fminf(pe.l, pe.r), pr = fmaxf(pe.l, pe.r);
  const float pw = pr - pl, ph = pb - pt;
  const float area_p = ph * pw, area_g = (ge.b - ge.t) * (ge.r - ge.l);
  const float ih = fminf(pb, ge.b) - fmaxf(pt, ge.t), iw = fminf(pr, ge.r) - fmaxf(pl, ge.l);
  const float inter = iw * ih, uni = area_p + area_g - inter;
  const bool disjoint = iw <= 0 || ih <= 0;
  // d(area_p), d(inter) per edge; d(uni) = their difference
  const float d_area[4] = {-1 * pw, pw, -1 * ph, ph};
  const float d_inter[4] = {pt > ge.t ? (-1 * iw) : 0, pb < ge.b ? iw : 0, pl > ge.l ? (-1 * ih) : 0, pr < ge.r ? ih : 0};
  float d_uni[4], g[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    d_uni[e] = d_area[e] - d_inter[e];
    g[e] = uni > 0 ? ((uni * d_inter[e]) - (inter * d_uni[e])) / (uni * uni) : 0.f;
  }
  // the reference re-assigns the pairs one after the other (box.cpp:215-218): for swapped edges BOTH end up with the
  // second one's value
  g[0] = pe.t < pe.b ? g[0] : g[1];
  g[1] = pe.t < pe.b ? g[1] : g[0];
  g[2] = pe.l < pe.r ? g[2] : g[3];
  g[3] = pe.l < pe.r ? g[3] : g[2];
  if (KIND == K_GIOU) {
    const float cw = fmaxf(pr, ge.r) - fminf(pl, ge.l), ch = fmaxf(pb, ge.b) - fminf(pt, ge.t), c = cw * ch;
    const float d_enc[4] = {pt < ge.t ? (-1 * cw) : 0, pb > ge.b ? cw : 0, pl < ge.l ? (-1 * ch) : 0, pr > ge.r ? ch : 0};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float term = ((c * d_uni[e]) - (uni * d_enc[e])) / (c * c);  // formed once: added when c > 0, alone when disjoint
      if (c > 0) g[e] += term;
      if (disjoint) g[e] = term;
    }
  }
  float dx = g[2] + g[3], dy = g[0] + g[1], dw = g[3] - g[2], dh = g[1] - g[0];
  if (KIND == K_DIOU || KIND == K_CIOU) {
    // centre-distance penalty S / C with C the squared diagonal of the enclosing box (of the UNSORTED prediction edges)
    const float cw = fmaxf(pe.r, ge.r) - fminf(pe.l, ge.l), ch = fmaxf(pe.b, ge.b) - fminf(pe.t, ge.t);
    const float c = sq(cw) + sq(ch);
    const float s = sq(pred.x - gt.x) + sq(pred.y - gt.y);
    const float dcw_dx = (pr > ge.r ? 1.f : 0.f) - (pl < ge.l ? 1.f : 0.f);
    const float dcw_dw = (pr > ge.r ? 0.5f : 0.f) - (pl < ge.l ? -0.5f : 0.f);
    const float dch_dy = (pb > ge.b ? 1.f : 0.f) - (pt < ge.t ? 1.f : 0.f);
    const float dch_dh = (pb > ge.b ? 0.5f : 0.f) - (pt < ge.t ? -0.5f : 0.f);
    const float zero = 0.f;  // the reference's dCh_dx, dCw_dy, dCh_dw, dCw_dh: kept in the sums (inf * 0 must stay NaN)
    float tx = (2 * (gt.x - pred.x) * c - (2 * cw * dcw_dx + 2 * ch * zero) * s) / sq(c);
    float ty = (2 * (gt.y - pred.y) * c - (2 * cw * zero + 2 * ch * dch_dy) * s) / sq(c);
    float tw = (2 * cw * dcw_dw + 2 * ch * zero) * s / sq(c);
    float th = (2 * cw * zero + 2 * ch * dch_dh) * s / sq(c);
    if (KIND == K_CIOU) {  // aspect-ratio term
      const float ar_gt = gt.w / gt.h, ar_pred = pred.w / pred.h;
      const double dat = atan((double)ar_gt) - atan((double)ar_pred);
      const float ar_loss = (float)(4 / (M_PI * M_PI) * (dat * dat));
      const float alpha = (float)(ar_loss / (1 - inter / uni + ar_loss + 0.000001));
      tw = tw + alpha * (float)(8 / (M_PI * M_PI) * dat * pred.h);
      th = th + alpha * (float)(-8 / (M_PI * M_PI) * dat * pred.w);
    }
    if (c > 0) { dx += tx; dy += ty; dw += tw; dh += th; }
    if (disjoint) { dx = tx; dy = ty; dw = tw; dh = th; }
  }
  return make_float4(dx, dy, dw, dh);
}
__device__ float4 iou_family_grad_rt(const Box4& pred, const Box4& gt, int kind) {
  switch (kind) {
    case K_GIOU: return iou_family_grad<K_GIOU>(pred, gt);
    case K_DIOU: return iou_family_grad<K_DIOU>(pred, gt);
    case K_CIOU: return iou_family_grad<K_CIOU>(pred, gt);
  }
  return iou_family_grad<K_IOU>(pred, gt);
}

// ---- the head tensor ---------------------------------------------------------------------------------------------------
struct Head {
  int B, H, W, n, total, classes, max_boxes, net_w, net_h;
  int iou_loss, iou_thresh_kind, focal_loss;
  float ignore_thresh, truth_thresh, iou_thresh, iou_normalizer, cls_normalizer, label_smooth_eps, max_delta;
  int gaussian;  // 1: entries mu_x, s_x, mu_y, s_y, mu_w, s_w, mu_h, s_h, objectness, classes (gaussian_yolo_layer.cpp)
  float uc_normalizer;
  const float* out;     // decoded head, fp32 NCHW
  const float* truth;   // [B][max_boxes*5]
  const float* anchors; // l->biases, 2*total
  const int* mask;      // n
  float* delta;         // fp32 NCHW, shape of out
  float* stats;         // [0..8) public: cost, count, sum(1-iou), sum(1-giou), |delta|^2, |delta w/o box|^2, sum iou, class
                        // count; then per-image partials [B][4] and per-block partials of the finish pass [blocks][2]
  int finish_blocks;

  __device__ __forceinline__ int wh() const { return W * H; }
  __device__ __forceinline__ int box_entries() const { return gaussian ? 8 : 4; }
  __device__ __forceinline__ int entries() const { return box_entries() + 1 + classes; }
  // element (image b, anchor slot a, entry e, cell) — EntryIndex yolo_layer.cpp:380-386
  __device__ __forceinline__ long long at(int b, int a, int e, int cell) const {
    return ((long long)(b * n + a) * entries() + e) * wh() + cell;
  }
  __device__ __forceinline__ float* img_partial(int b) const { return stats + 8 + 4 * b; }
  __device__ __forceinline__ float* blk_partial(int k) const { return stats + 8 + 4 * B + 2 * k; }
};

// GetYoloBox (yolo_layer.cpp:139-148) / GetGaussianYoloBox (gaussian_yolo_layer.cpp:151-160, yolo_point = center) for
// anchor `anchor` (absolute index) at cell (col, row); `base` = index of the cell's first entry
__device__ __forceinline__ Box4 predicted_box(const Head& p, long long base, int anchor, int col, int row) {
  const int s = p.wh();
  const int ex = 0, ey = p.gaussian ? 2 : 1, ew = p.gaussian ? 4 : 2, eh = p.gaussian ? 6 : 3;
  Box4 q;
  q.x = (col + p.out[base + ex * s]) / p.W;
  q.y = (row + p.out[base + ey * s]) / p.H;
  q.w = expf(p.out[base + ew * s]) * p.anchors[2 * anchor] / p.net_w;
  q.h = expf(p.out[base + eh * s]) * p.anchors[2 * anchor + 1] / p.net_h;
  return q;
}
__device__ __forceinline__ float sanitize(float v) { return (isnan(v) || isinf(v)) ? 0.f : v; }       // fix_nan_inf
__device__ __forceinline__ float clamp_abs(float v, float m) { return v > m ? m : (v < -m ? -m : v); }  // clip_value

// Box delta of one (truth, location) pair, ACCUMULATED into delta (delta_yolo_box yolo_layer.cpp:172-262,
// delta_gaussian_yolo_box gaussian_yolo_layer.cpp:198-402 with yolo_point = center).  Returns (iou, giou) for the layer
// statistics.  Executed by one thread.
__device__ float2 box_delta(const Head& p, const Box4& truth, long long base, int anchor, int col, int row) {
  const int s = p.wh();
  const float scale = 2 - truth.w * truth.h;
  Box4 pred = predicted_box(p, base, anchor, col, row);
  const Overlap o = overlap_of(pred, truth);
  float2 r = make_float2(o.iou, giou_of(o));
  if (pred.w == 0) pred.w = 1.0f;
  if (pred.h == 0) pred.h = 1.0f;
  // regression targets in the head's own parametrisation
  const float target[4] = {truth.x * p.W - col, truth.y * p.H - row, logf(truth.w * p.net_w / p.anchors[2 * anchor]),
                           logf(truth.h * p.net_h / p.anchors[2 * anchor + 1])};
  if (!p.gaussian) {
    float d[4];
    if (p.iou_loss == K_MSE) {
#pragma unroll
      for (int k = 0; k < 4; ++k) d[k] = scale * (target[k] - p.out[base + k * s]) * p.iou_normalizer;
    } else {
      const float4 g = iou_family_grad_rt(pred, truth, p.iou_loss);
      d[0] = g.x; d[1] = g.y;
      d[2] = g.z * expf(p.out[base + 2 * s]);
      d[3] = g.w * expf(p.out[base + 3 * s]);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        d[k] = sanitize(d[k] * p.iou_normalizer);
        if (p.max_delta != FLT_MAX) d[k] = clamp_abs(d[k], p.max_delta);
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) p.delta[base + k * s] += d[k];
    return r;
  }
  // Gaussian head: negative log-likelihood of the target under N(mu, sigma) per coordinate (sigma_const 0.3, eps 1e-9);
  // the reference evaluates it in a float / double mix that is kept here
  const float sigma_const = 0.3f;
  const float epsi = (float)pow(10.0, -9.0);
  float dmu[4], dsg[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float mu = p.out[base + (2 * k) * s], sg = p.out[base + (2 * k + 1) * s];
    const float z = (target[k] - mu) / sg;
    const float z2 = (float)pow((double)z, 2.0);
    const float normal = (float)(exp((double)z2 * (-1. / 2.)) / (sqrt(M_PI * 2.0) * (double)(sg + sigma_const)));
    const float t = (float)((1. / 2.) * 1. / (double)(normal + epsi) * (double)normal * (double)scale);
    dmu[k] = (float)((double)(t * z) * (1. / (double)sg));
    dsg[k] = (float)((double)t * ((double)(z2 / sg) - 1. / (double)(sg + sigma_const)));
  }
  if (p.iou_loss == K_MSE) {
    r.y = r.x;                // the Gaussian layer keeps ONE overlap figure for its statistics: IoU with mse ...
  } else {                    // ... GIoU otherwise (gaussian_yolo_layer.cpp:352); IoU-family gradient for the means
    r.x = r.y;
    const float4 g = iou_family_grad_rt(pred, truth, p.iou_loss);
    dmu[0] = g.x; dmu[1] = g.y;
    dmu[2] = g.z * expf(p.out[base + 4 * s]);
    dmu[3] = g.w * expf(p.out[base + 6 * s]);
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    dmu[k] = sanitize(dmu[k] * p.iou_normalizer);
    dsg[k] = sanitize(dsg[k] * p.uc_normalizer);
    if (p.max_delta != FLT_MAX) {
      dmu[k] = clamp_abs(dmu[k], p.max_delta);
      dsg[k] = clamp_abs(dsg[k], p.max_delta);
    }
    p.delta[base + (2 * k) * s] += dmu[k];
    p.delta[base + (2 * k + 1) * s] += dsg[k];
  }
  return r;
}

// Class deltas of one (truth, location) pair (delta_yolo_class yolo_layer.cpp:285-357, DeltaGaussianYoloClass
// gaussian_yolo_layer.cpp:429-461); cls0 = index of the cell's first class entry.  One thread.
__device__ void class_delta(const Head& p, long long cls0, int class_id) {
  const int s = p.wh();
  const float smooth = p.label_smooth_eps;
  const float on = smooth ? 1 * (1 - smooth) + 0.5f * smooth : 1.f;
  const float off = smooth ? 0 * (1 - smooth) + 0.5f * smooth : 0.f;
  // a location an earlier truth has already marked only gets this truth's class refreshed.  [yolo] looks at this class'
  // delta, [Gaussian_yolo] at the first class' delta.
  const bool marked = p.gaussian ? (p.delta[cls0] != 0.f) : (p.delta[cls0 + (long long)s * class_id] != 0.f);
  if (marked) {
    const float r = on - p.out[cls0 + (long long)s * class_id];
    if (p.gaussian || (!isnan(r) && !isinf(r))) p.delta[cls0 + (long long)s * class_id] = r;
    return;
  }
  if (p.focal_loss && !p.gaussian) {
    const float pt = p.out[cls0 + (long long)s * class_id] + 0.000000000000001F;
    const float grad = -(1 - pt) * (2 * pt * logf(pt) + pt - 1);
    for (int c = 0; c < p.classes; ++c) {
      float d = ((c == class_id) ? 1 : 0) - p.out[cls0 + (long long)s * c];
      d *= 0.5f * grad;
      p.delta[cls0 + (long long)s * c] = d;
    }
    return;
  }
  for (int c = 0; c < p.classes; ++c) {
    const float r = (c == class_id ? on : off) - p.out[cls0 + (long long)s * c];
    if (p.gaussian || (!isnan(r) && !isinf(r))) p.delta[cls0 + (long long)s * c] = r;
  }
}

constexpr int kMaxTruthSmem = 256;  // truth boxes staged per image (the cfgs use max = 90 .. 200)

// the truth list of image b into shared memory; returns the number of entries before the terminator
__device__ int stage_truths(const Head& p, int b, float* sh /* [max_boxes*5] */, int* sh_end) {
  const float* tr = p.truth + (long long)b * p.max_boxes * 5;
  if (threadIdx.x == 0) *sh_end = p.max_boxes;
  for (int i = threadIdx.x; i < p.max_boxes * 5; i += blockDim.x) sh[i] = tr[i];
  __syncthreads();
  for (int t = threadIdx.x; t < p.max_boxes; t += blockDim.x) {
    const int cls = (int)sh[t * 5 + 4];
    if (cls < p.classes && cls >= 0 && !sh[t * 5]) atomicMin(sh_end, t);  // first valid-class entry with x == 0 ends the list
  }
  __syncthreads();
  return *sh_end;
}
__device__ __forceinline__ bool truth_valid(const Head& p, const float* sh, int t) {
  const int cls = (int)sh[t * 5 + 4];
  return cls < p.classes && cls >= 0;
}

// ---- objectness pass (yolo_layer.cpp:431-522) ----------------------------------------------------------------------------
__global__ void __launch_bounds__(128) yolo_objectness_kernel(Head p) {
  __shared__ float tr[kMaxTruthSmem * 5];
  __shared__ int end_s;
  const int b = blockIdx.y;
  const int n_truth = stage_truths(p, b, tr, &end_s);
  const int s = p.wh();
  const int be = p.box_entries();
  for (int t0 = blockIdx.x * blockDim.x + threadIdx.x; t0 < p.n * s; t0 += gridDim.x * blockDim.x) {
    const int a = t0 / s, cell = t0 - a * s;
    const int col = cell % p.W, row = cell / p.W;
    const long long base = p.at(b, a, 0, cell);
    const Box4 pred = predicted_box(p, base, p.mask[a], col, row);
    float best_iou = 0, best_match = 0;
    int best_t = 0;
    int confident = -1;  // compare_yolo_class(..., 0.25): some class score above 0.25; evaluated once, on demand
    for (int t = 0; t < n_truth; ++t) {
      if (!truth_valid(p, tr, t)) continue;
      const Box4 g{tr[t * 5], tr[t * 5 + 1], tr[t * 5 + 2], tr[t * 5 + 3]};
      if (confident < 0) {
        confident = 0;
        for (int c = 0; c < p.classes; ++c)
          if (p.out[base + (long long)(be + 1 + c) * s] > 0.25f) { confident = 1; break; }
      }
      const float iou = iou_of(pred, g);
      if (iou > best_match && confident == 1) best_match = iou;
      if (iou > best_iou) { best_iou = iou; best_t = t; }
    }
    const float obj = p.out[base + (long long)be * s];
    p.delta[base + (long long)be * s] = best_match > p.ignore_thresh ? 0.f : p.cls_normalizer * (0 - obj);
    if (best_iou > p.truth_thresh) {  // (truth_thresh = 1 in every cfg of the path: this never fires there)
      p.delta[base + (long long)be * s] = p.cls_normalizer * (1 - obj);
      const Box4 g{tr[best_t * 5], tr[best_t * 5 + 1], tr[best_t * 5 + 2], tr[best_t * 5 + 3]};
      class_delta(p, base + (long long)(be + 1) * s, (int)tr[best_t * 5 + 4]);
      box_delta(p, g, base, p.mask[a], col, row);
    }
  }
}

// ---- assignment pass (yolo_layer.cpp:524-676) -----------------------------------------------------------------------------
// record (t, k): truth t with its best anchor (k = 0) or with anchor k - 1 when that one passes iou_thresh (k >= 1)
struct Rec {
  int loc;     // slot * wh + cell inside the image, -1: not applicable
  short t, anchor;
};
__global__ void __launch_bounds__(128) yolo_assign_kernel(Head p) {
  extern __shared__ unsigned char smem_[];
  float* tr = reinterpret_cast<float*>(smem_);                       // [max_boxes*5]
  Rec* recs = reinterpret_cast<Rec*>(tr + p.max_boxes * 5);            // [max_boxes*(total+1)]
  float* rstat = reinterpret_cast<float*>(recs + p.max_boxes * (p.total + 1));     // (iou, giou) per applied record
  __shared__ int end_s;
  const int b = blockIdx.x;
  const int n_truth = stage_truths(p, b, tr, &end_s);
  const int s = p.wh();
  const int per = p.total + 1;
  const int n_rec = n_truth * per;
  // records: thread per truth
  for (int t = threadIdx.x; t < n_truth; t += blockDim.x) {
    for (int k = 0; k < per; ++k) recs[t * per + k].loc = -1;
    if (!truth_valid(p, tr, t)) continue;
    const Box4 g{tr[t * 5], tr[t * 5 + 1], tr[t * 5 + 2], tr[t * 5 + 3]};
    const int col = (int)(g.x * p.W), row = (int)(g.y * p.H);
    const Box4 shape{0, 0, g.w, g.h};
    float best = 0;
    int best_a = 0;
    for (int a = 0; a < p.total; ++a) {  // best anchor by shape IoU
      const Box4 an{0, 0, p.anchors[2 * a] / p.net_w, p.anchors[2 * a + 1] / p.net_h};
      const float iou = iou_of(an, shape);
      if (iou > best) { best = iou; best_a = a; }
    }
    for (int k = 0; k < per; ++k) {
      const int a = k == 0 ? best_a : k - 1;
      if (k > 0) {
        if (a == best_a || !(p.iou_thresh < 1.0f)) continue;
        const Box4 an{0, 0, p.anchors[2 * a] / p.net_w, p.anchors[2 * a + 1] / p.net_h};
        if (!(measure_of(an, shape, p.iou_thresh_kind) > p.iou_thresh)) continue;
      }
      int slot = -1;
      for (int m = 0; m < p.n; ++m)
        if (p.mask[m] == a) { slot = m; break; }  // int_index: is this anchor one of this head's?
      if (slot < 0) continue;
      recs[t * per + k] = Rec{slot * s + row * p.W + col, (short)t, (short)a};
    }
  }
  __syncthreads();
  // resolve: the first record of a location applies every record of that location, in record order
  for (int r = threadIdx.x; r < n_rec; r += blockDim.x) {
    rstat[2 * r] = rstat[2 * r + 1] = 0.f;
    const int loc = recs[r].loc;
    if (loc < 0) continue;
    bool owner = true;
    for (int q = 0; q < r && owner; ++q) owner = recs[q].loc != loc;
    if (!owner) continue;
    const int slot = loc / s, cell = loc - slot * s;
    const int col = cell % p.W, row = cell / p.W;
    const long long base = p.at(b, slot, 0, cell);
    const int be = p.box_entries();
    for (int q = r; q < n_rec; ++q) {
      if (recs[q].loc != loc) continue;
      const int t = recs[q].t;
      const Box4 g{tr[t * 5], tr[t * 5 + 1], tr[t * 5 + 2], tr[t * 5 + 3]};
      const float2 m = box_delta(p, g, base, recs[q].anchor, col, row);
      rstat[2 * q] = m.x;
      rstat[2 * q + 1] = m.y;
      p.delta[base + (long long)be * s] = p.cls_normalizer * (1 - p.out[base + (long long)be * s]);
      class_delta(p, base + (long long)(be + 1) * s, (int)tr[t * 5 + 4]);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {  // this image's statistics, summed in record order
    float count = 0, iou_loss = 0, giou_loss = 0, iou_sum = 0;
    for (int r = 0; r < n_rec; ++r) {
      if (recs[r].loc < 0) continue;
      count += 1;
      iou_sum += rstat[2 * r];
      iou_loss += 1 - rstat[2 * r];
      giou_loss += 1 - rstat[2 * r + 1];
    }
    float* o = p.img_partial(b);
    o[0] = count; o[1] = iou_loss; o[2] = giou_loss; o[3] = iou_sum;
  }
}

// ---- finish pass: averages_yolo_deltas (:264-283, :679-696), sums of squares of the cost (:708-733), BackwardYoloLayer's
// prev.delta += delta (:774-777) into the head conv's NHWC bf16 delta.  Block per (image, 32 cells): all n*E entries of the
// cells staged in shared memory, so the fp32 NCHW planes move 128 bytes per warp and the NHWC delta is updated row by row.
__global__ void __launch_bounds__(256) yolo_finish_tiled_kernel(Head p, __nv_bfloat16* __restrict__ prev, long long prev_ld) {
  extern __shared__ float tile[];  // [n*E][33] then int positives[n][32]
  const int s = p.wh();
  const int be = p.box_entries(), E = p.entries(), CH = p.n * E;
  int* positives = reinterpret_cast<int*>(tile + CH * 33);
  const int tiles = (s + 31) / 32;
  const int b = blockIdx.x / tiles, cell0 = (blockIdx.x % tiles) * 32;
  float* dl = p.delta + (long long)b * CH * s;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const bool in_range = cell0 + lane < s;
  for (int ch = warp; ch < CH; ch += 8) tile[ch * 33 + lane] = in_range ? dl[(long long)ch * s + cell0 + lane] : 0.f;
  __syncthreads();
  for (int t = threadIdx.x; t < p.n * 32; t += 256) {
    const int a = t >> 5, l = t & 31;
    int k = 0;
    for (int c = 0; c < p.classes; ++c)
      if (tile[(a * E + be + 1 + c) * 33 + l] > 0) ++k;
    positives[t] = k;
  }
  __syncthreads();
  float sum_all = 0, sum_rest = 0;
  for (int ch = warp; ch < CH; ch += 8) {
    const int a = ch / E, e = ch - a * E;
    float d = tile[ch * 33 + lane];
    const int k = positives[a * 32 + lane];
    if (e < be && k > 0 && in_range) {
      d /= k;
      dl[(long long)ch * s + cell0 + lane] = d;
      tile[ch * 33 + lane] = d;
    }
    sum_all += d * d;  // cells past the image hold zeros
    if (e >= be) sum_rest += d * d;
  }
  __syncthreads();
  if (prev) {
    for (int l = warp; l < 32 && cell0 + l < s; l += 8) {
      __nv_bfloat16* pv = prev + ((long long)b * s + cell0 + l) * prev_ld;
      for (int c2 = lane * 2; c2 < CH; c2 += 64) {
        if (c2 + 1 < CH) {
          __nv_bfloat162* q = reinterpret_cast<__nv_bfloat162*>(pv + c2);
          const float2 f = __bfloat1622float2(*q);
          *q = __floats2bfloat162_rn(f.x + tile[c2 * 33 + l], f.y + tile[(c2 + 1) * 33 + l]);
        } else {
          pv[c2] = __float2bfloat16_rn(__bfloat162float(pv[c2]) + tile[c2 * 33 + l]);
        }
      }
    }
  }
  __shared__ float part[2][8];
  for (int off = 16; off; off >>= 1) {
    sum_all += __shfl_xor_sync(0xffffffffu, sum_all, off);
    sum_rest += __shfl_xor_sync(0xffffffffu, sum_rest, off);
  }
  if (lane == 0) { part[0][warp] = sum_all; part[1][warp] = sum_rest; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0, q = 0;
    for (int w = 0; w < 8; ++w) { a += part[0][w]; q += part[1][w]; }
    float* o = p.blk_partial(blockIdx.x);
    o[0] = a; o[1] = q;
  }
}
// the same without the shared-memory tile, for heads whose n*E entries do not fit it: thread per (image, anchor, cell)
__global__ void __launch_bounds__(128) yolo_finish_kernel(Head p, __nv_bfloat16* __restrict__ prev, long long prev_ld) {
  const int s = p.wh();
  const int be = p.box_entries(), E = p.entries();
  const long long total = (long long)p.B * p.n * s;
  float sum_all = 0, sum_rest = 0;
  for (long long t0 = blockIdx.x * (long long)blockDim.x + threadIdx.x; t0 < total; t0 += (long long)gridDim.x * blockDim.x) {
    const int cell = (int)(t0 % s);
    const int a = (int)((t0 / s) % p.n);
    const int b = (int)(t0 / ((long long)s * p.n));
    const long long base = p.at(b, a, 0, cell);
    int k = 0;
    for (int c = 0; c < p.classes; ++c)
      if (p.delta[base + (long long)(be + 1 + c) * s] > 0) ++k;
    __nv_bfloat16* pv = prev ? prev + ((long long)b * s + cell) * prev_ld + a * E : nullptr;
    for (int e = 0; e < E; ++e) {
      float d = p.delta[base + (long long)e * s];
      if (e < be && k > 0) {
        d /= k;
        p.delta[base + (long long)e * s] = d;
      }
      sum_all += d * d;
      if (e >= be) sum_rest += d * d;
      if (pv) pv[e] = __float2bfloat16_rn(__bfloat162float(pv[e]) + d);
    }
  }
  __shared__ float part[2][4];
  for (int off = 16; off; off >>= 1) {
    sum_all += __shfl_xor_sync(0xffffffffu, sum_all, off);
    sum_rest += __shfl_xor_sync(0xffffffffu, sum_rest, off);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { part[0][warp] = sum_all; part[1][warp] = sum_rest; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0, q = 0;
    for (int w = 0; w < 4; ++w) { a += part[0][w]; q += part[1][w]; }
    float* o = p.blk_partial(blockIdx.x);
    o[0] = a; o[1] = q;
  }
}

// ---- cost pass: partial sums in index order, then *(l->cost) (yolo_layer.cpp:735-754, gaussian_yolo_layer.cpp:835) --------
__global__ void __launch_bounds__(256) yolo_cost_kernel(Head p) {
  __shared__ float acc[256][2];
  float a = 0, q = 0;
  for (int k = threadIdx.x; k < p.finish_blocks; k += 256) {
    a += p.blk_partial(k)[0];
    q += p.blk_partial(k)[1];
  }
  acc[threadIdx.x][0] = a;
  acc[threadIdx.x][1] = q;
  __syncthreads();
  if (threadIdx.x != 0) return;
  float sum_all = 0, sum_rest = 0;
  for (int k = 0; k < 256; ++k) { sum_all += acc[k][0]; sum_rest += acc[k][1]; }
  float count = 0, iou_loss = 0, giou_loss = 0, iou_sum = 0;
  for (int b = 0; b < p.B; ++b) {
    const float* o = p.img_partial(b);
    count += o[0]; iou_loss += o[1]; giou_loss += o[2]; iou_sum += o[3];
  }
  float* st = p.stats;
  st[1] = count; st[2] = iou_loss; st[3] = giou_loss; st[4] = sum_all; st[5] = sum_rest; st[6] = iou_sum; st[7] = count;
  if (p.iou_loss == K_MSE || p.gaussian) {
    st[0] = sum_all;
  } else {
    const float n = count == 0 ? 1 : count;
    st[0] = p.iou_normalizer * ((p.iou_loss == K_GIOU ? giou_loss : iou_loss) / n) + p.cls_normalizer * sum_rest;
  }
}

}  // namespace yc

using namespace yc;

// floats of the `stats` buffer yc_yolo_loss needs for a head of n images and h x w cells: the 8 public values, one
// 4-float partial per image and one 2-float partial per block of the finish pass
extern "C" size_t yc_yolo_loss_stats_floats(int n, int h, int w) {
  const size_t tiles = (size_t)n * (((size_t)h * w + 31) / 32);
  return 8 + 4 * (size_t)n + 2 * std::max<size_t>(tiles, (size_t)148 * 16);
}

extern "C" int yc_yolo_loss(const yc_yolo_loss_desc* d, const float* out_nchw, const float* truth, const float* anchor_wh,
                            const int* mask, int n, int h, int w, float* delta_nchw, const yc_tensor* prev_delta,
                            float* stats, void* stream) {
  YC_REQUIRE(d && out_nchw && truth && anchor_wh && mask && delta_nchw && stats && n > 0 && h > 0 && w > 0,
             "yc_yolo_loss: bad arguments");
  YC_REQUIRE(d->anchors > 0 && d->total >= d->anchors && d->classes > 0 && d->max_boxes > 0, "yc_yolo_loss: bad descriptor");
  YC_REQUIRE(d->max_boxes <= kMaxTruthSmem, "yc_yolo_loss: more than %d truth boxes per image", kMaxTruthSmem);
  const int E = (d->gaussian ? 9 : 5) + d->classes;
  if (prev_delta)
    YC_REQUIRE(prev_delta->dtype == YC_BF16 && prev_delta->n == n && prev_delta->h == h && prev_delta->w == w &&
                   prev_delta->ldc >= d->anchors * E,
               "yc_yolo_loss: prev_delta must be the head conv's NHWC bf16 delta");
  YC_REQUIRE(!(d->gaussian && d->focal_loss), "yc_yolo_loss: [Gaussian_yolo] has no focal loss");
  cudaStream_t st = (cudaStream_t)stream;
  Head p;
  p.B = n; p.H = h; p.W = w; p.n = d->anchors; p.total = d->total; p.classes = d->classes; p.max_boxes = d->max_boxes;
  p.net_w = d->net_w; p.net_h = d->net_h; p.iou_loss = d->iou_loss; p.iou_thresh_kind = d->iou_thresh_kind;
  p.focal_loss = d->focal_loss; p.ignore_thresh = d->ignore_thresh; p.truth_thresh = d->truth_thresh;
  p.iou_thresh = d->iou_thresh; p.iou_normalizer = d->iou_normalizer; p.cls_normalizer = d->cls_normalizer;
  p.label_smooth_eps = d->label_smooth_eps; p.max_delta = d->max_delta;
  p.gaussian = d->gaussian ? 1 : 0; p.uc_normalizer = d->uc_normalizer;
  p.out = out_nchw; p.truth = truth; p.anchors = anchor_wh; p.mask = mask; p.delta = delta_nchw; p.stats = stats;
  const size_t elems = (size_t)n * d->anchors * E * h * w;
  YC_CUDA(cudaMemsetAsync(delta_nchw, 0, elems * 4, st));
  const int cells = d->anchors * h * w;
  const dim3 grid_obj((unsigned)std::max(1, std::min(ceil_div(cells, 128), std::max(1, sm_count() * 16 / n))), (unsigned)n);
  yolo_objectness_kernel<<<grid_obj, 128, 0, st>>>(p);
  YC_LAUNCH_CHECK();
  const size_t assign_smem = (size_t)d->max_boxes * 5 * 4 + (size_t)d->max_boxes * (d->total + 1) * (sizeof(Rec) + sizeof(float2));
  YC_REQUIRE(assign_smem <= 48 * 1024, "yc_yolo_loss: max_boxes * (num + 1) too large for the assignment pass");
  yolo_assign_kernel<<<n, 128, assign_smem, st>>>(p);
  YC_LAUNCH_CHECK();
  const size_t tile_bytes = (size_t)d->anchors * E * 33 * 4 + (size_t)d->anchors * 32 * 4;
  __nv_bfloat16* pv = prev_delta ? (__nv_bfloat16*)prev_delta->ptr : nullptr;
  const long long pld = prev_delta ? prev_delta->ldc : 0;
  if (tile_bytes <= 48 * 1024 && (!prev_delta || prev_delta->ldc % 2 == 0)) {
    p.finish_blocks = n * ((h * w + 31) / 32);
    yolo_finish_tiled_kernel<<<p.finish_blocks, 256, tile_bytes, st>>>(p, pv, pld);
  } else {
    p.finish_blocks = (int)std::max<long long>(1, std::min<long long>(ceil_div_ll((long long)n * cells, 128), (long long)148 * 16));
    yolo_finish_kernel<<<p.finish_blocks, 128, 0, st>>>(p, pv, pld);
  }
  YC_LAUNCH_CHECK();
  yolo_cost_kernel<<<1, 256, 0, st>>>(p);
  YC_LAUNCH_CHECK();
  return YC_OK;
}
