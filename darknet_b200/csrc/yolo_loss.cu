// YOLO-layer loss and deltas on the device.
//
// Replaces the training half of ForwardYoloLayer (yolo_layer.cpp:409-772: delta_yolo_box, delta_yolo_class,
// averages_yolo_deltas, compare_yolo_class; IoU math box.cpp:36-351) and BackwardYoloLayerGpu (:884-888).  In the
// reference this runs on ONE host thread between a device->host copy of the head and a host->device copy of the
// deltas (yolo_layer.cpp:856-880); here the head never leaves HBM:
//   1. yolo_loss_cells   : thread per (image, anchor, cell)   — best IoU over the truths, objectness delta
//   2. yolo_loss_truths  : one warp per image, truths in order — box / objectness / class deltas of the
//                          responsible anchors (order-dependent in the reference, so kept sequential per image;
//                          the class loop runs across the lanes)
//   3. yolo_loss_finish  : thread per (image, anchor, cell)   — averages_yolo_deltas, sum of squares for the cost,
//                          and prev.delta += delta written straight into the head conv's NHWC bf16 delta
// Layouts: `out` / `delta` are fp32 NCHW [B][n*(5+classes)][H][W] exactly like l->output / l->delta
// (EntryIndex, yolo_layer.cpp:380-386); truth is [B][max_boxes][5] = (x, y, w, h, class id).
#include <float.h>
#include <math.h>
#include <algorithm>

#include "yc_common.cuh"

namespace yc {

enum { IOU_IOU = 0, IOU_GIOU = 1, IOU_MSE = 2, IOU_DIOU = 3, IOU_CIOU = 4 };  // IOU_LOSS, box.h:6-13

struct BoxF { float x, y, w, h; };
struct DxRep { float dt, db, dl, dr; };
struct IousF { float iou, giou, diou, ciou; };

#define YSQ(a) ((a) * (a))

__device__ __forceinline__ float b_overlap(float x1, float w1, float x2, float w2) {  // Box::Overlap box.cpp:36-46
  const float l1 = x1 - w1 / 2, l2 = x2 - w2 / 2;
  const float left = l1 > l2 ? l1 : l2;
  const float r1 = x1 + w1 / 2, r2 = x2 + w2 / 2;
  const float right = r1 < r2 ? r1 : r2;
  return right - left;
}
__device__ __forceinline__ float b_intersect(const BoxF& a, const BoxF& b) {  // box.cpp:48-57
  const float w = b_overlap(a.x, a.w, b.x, b.w), h = b_overlap(a.y, a.h, b.y, b.h);
  if (w < 0 || h < 0) return 0;
  return w * h;
}
__device__ __forceinline__ float b_union(const BoxF& a, const BoxF& b) { return a.w * a.h + b.w * b.h - b_intersect(a, b); }
__device__ __forceinline__ float b_iou(const BoxF& a, const BoxF& b) {  // box.cpp:64-72
  const float I = b_intersect(a, b), U = b_union(a, b);
  if (fabsf(I) < FLT_EPSILON || fabsf(U) < FLT_EPSILON) return 0;
  return I / U;
}
struct AbsF { float left, right, top, bottom; };
__device__ __forceinline__ AbsF b_abs(const BoxF& b) {  // box.cpp:353-359
  return AbsF{b.x - b.w / 2.0f, b.x + b.w / 2.0f, b.y - b.h / 2.0f, b.y + b.h / 2.0f};
}
__device__ __forceinline__ AbsF b_abs2(const BoxF& a, const BoxF& b) {  // box.cpp:361-370
  const AbsF p = b_abs(a), q = b_abs(b);
  return AbsF{fminf(p.left, q.left), fmaxf(p.right, q.right), fminf(p.top, q.top), fmaxf(p.bottom, q.bottom)};
}
__device__ float b_giou(const BoxF& a, const BoxF& b) {  // box.cpp:113-128
  const AbsF ab = b_abs2(a, b);
  const float w = ab.right - ab.left, h = ab.bottom - ab.top;
  const float c = w * h;
  const float iou = b_iou(a, b);
  if (fabsf(c) < FLT_EPSILON) return iou;
  const float u = b_union(a, b);
  return iou - (c - u) / c;
}
__device__ float b_diou(const BoxF& a, const BoxF& b, float beta) {  // box.cpp:96-111
  const AbsF ab = b_abs2(a, b);
  const float w = ab.right - ab.left, h = ab.bottom - ab.top;
  const float c = w * w + h * h;
  const float iou = b_iou(a, b);
  if (fabsf(c) < FLT_EPSILON) return iou;
  const float d = YSQ(a.x - b.x) + YSQ(a.y - b.y);
  return iou - (float)pow((double)(d / c), (double)beta);
}
__device__ float b_ciou(const BoxF& a, const BoxF& b) {  // box.cpp:74-94
  const AbsF ab = b_abs2(a, b);
  const float w = ab.right - ab.left, h = ab.bottom - ab.top;
  const float c = w * w + h * h;
  const float iou = b_iou(a, b);
  if (fabsf(c) < FLT_EPSILON) return iou;
  const float u = YSQ(a.x - b.x) + YSQ(a.y - b.y);
  const float d = u / c;
  const float ar_gt = b.w / b.h, ar_pred = a.w / a.h;
  const float ar_loss = (float)(4 / YSQ(M_PI) * YSQ(atan((double)ar_gt) - atan((double)ar_pred)));
  const float alpha = (float)(ar_loss / (1 - iou + ar_loss + 0.000001));
  return iou - (d + alpha * ar_loss);
}
__device__ float b_rmse(const BoxF& a, const BoxF& b) {
  return sqrtf(YSQ(a.x - b.x) + YSQ(a.y - b.y) + YSQ(a.w - b.w) + YSQ(a.h - b.h));
}
__device__ float b_iou_kind(const BoxF& a, const BoxF& b, int kind) {  // box.cpp:136-151
  switch (kind) {
    case IOU_GIOU: return b_giou(a, b);
    case IOU_MSE: return b_rmse(a, b);
    case IOU_DIOU: return b_diou(a, b, 0.6f);
    case IOU_CIOU: return b_ciou(a, b);
  }
  return b_iou(a, b);
}

// Box::DxIou (box.cpp:153-351): gradient of IoU / GIoU / DIoU / CIoU w.r.t. (x, y, w, h) of the prediction
__device__ DxRep b_dx_iou(const BoxF& pred, const BoxF& gt, int iou_type) {
  const AbsF ap = b_abs(pred);
  const float pred_t = fminf(ap.top, ap.bottom), pred_b = fmaxf(ap.top, ap.bottom);
  const float pred_l = fminf(ap.left, ap.right), pred_r = fmaxf(ap.left, ap.right);
  const AbsF ag = b_abs(gt);
  const float X = (pred_b - pred_t) * (pred_r - pred_l);
  const float Xhat = (ag.bottom - ag.top) * (ag.right - ag.left);
  const float Ih = fminf(pred_b, ag.bottom) - fmaxf(pred_t, ag.top);
  const float Iw = fminf(pred_r, ag.right) - fmaxf(pred_l, ag.left);
  const float I = Iw * Ih;
  const float U = X + Xhat - I;
  const float S = YSQ(pred.x - gt.x) + YSQ(pred.y - gt.y);
  const float giou_Cw = fmaxf(pred_r, ag.right) - fminf(pred_l, ag.left);
  const float giou_Ch = fmaxf(pred_b, ag.bottom) - fminf(pred_t, ag.top);
  const float giou_C = giou_Cw * giou_Ch;
  const float dX_wrt_t = -1 * (pred_r - pred_l), dX_wrt_b = pred_r - pred_l;
  const float dX_wrt_l = -1 * (pred_b - pred_t), dX_wrt_r = pred_b - pred_t;
  const float dI_wrt_t = pred_t > ag.top ? (-1 * Iw) : 0, dI_wrt_b = pred_b < ag.bottom ? Iw : 0;
  const float dI_wrt_l = pred_l > ag.left ? (-1 * Ih) : 0, dI_wrt_r = pred_r < ag.right ? Ih : 0;
  const float dU_wrt_t = dX_wrt_t - dI_wrt_t, dU_wrt_b = dX_wrt_b - dI_wrt_b;
  const float dU_wrt_l = dX_wrt_l - dI_wrt_l, dU_wrt_r = dX_wrt_r - dI_wrt_r;
  const float dC_wrt_t = pred_t < ag.top ? (-1 * giou_Cw) : 0, dC_wrt_b = pred_b > ag.bottom ? giou_Cw : 0;
  const float dC_wrt_l = pred_l < ag.left ? (-1 * giou_Ch) : 0, dC_wrt_r = pred_r > ag.right ? giou_Ch : 0;
  float p_dt = 0, p_db = 0, p_dl = 0, p_dr = 0;
  if (U > 0) {
    p_dt = ((U * dI_wrt_t) - (I * dU_wrt_t)) / (U * U);
    p_db = ((U * dI_wrt_b) - (I * dU_wrt_b)) / (U * U);
    p_dl = ((U * dI_wrt_l) - (I * dU_wrt_l)) / (U * U);
    p_dr = ((U * dI_wrt_r) - (I * dU_wrt_r)) / (U * U);
  }
  // sequential re-assignments exactly as in the reference (box.cpp:215-218)
  p_dt = ap.top < ap.bottom ? p_dt : p_db;
  p_db = ap.top < ap.bottom ? p_db : p_dt;
  p_dl = ap.left < ap.right ? p_dl : p_dr;
  p_dr = ap.left < ap.right ? p_dr : p_dl;
  if (iou_type == IOU_GIOU) {
    if (giou_C > 0) {
      p_dt += ((giou_C * dU_wrt_t) - (U * dC_wrt_t)) / (giou_C * giou_C);
      p_db += ((giou_C * dU_wrt_b) - (U * dC_wrt_b)) / (giou_C * giou_C);
      p_dl += ((giou_C * dU_wrt_l) - (U * dC_wrt_l)) / (giou_C * giou_C);
      p_dr += ((giou_C * dU_wrt_r) - (U * dC_wrt_r)) / (giou_C * giou_C);
    }
    if (Iw <= 0 || Ih <= 0) {
      p_dt = ((giou_C * dU_wrt_t) - (U * dC_wrt_t)) / (giou_C * giou_C);
      p_db = ((giou_C * dU_wrt_b) - (U * dC_wrt_b)) / (giou_C * giou_C);
      p_dl = ((giou_C * dU_wrt_l) - (U * dC_wrt_l)) / (giou_C * giou_C);
      p_dr = ((giou_C * dU_wrt_r) - (U * dC_wrt_r)) / (giou_C * giou_C);
    }
  }
  const float Ct = fminf(pred.y - pred.h / 2, gt.y - gt.h / 2), Cb = fmaxf(pred.y + pred.h / 2, gt.y + gt.h / 2);
  const float Cl = fminf(pred.x - pred.w / 2, gt.x - gt.w / 2), Cr = fmaxf(pred.x + pred.w / 2, gt.x + gt.w / 2);
  const float Cw = Cr - Cl, Ch = Cb - Ct;
  const float C = YSQ(Cw) + YSQ(Ch);
  const float dCt_dy = pred_t < ag.top ? 1 : 0, dCt_dh = pred_t < ag.top ? -0.5f : 0;
  const float dCb_dy = pred_b > ag.bottom ? 1 : 0, dCb_dh = pred_b > ag.bottom ? 0.5f : 0;
  const float dCl_dx = pred_l < ag.left ? 1 : 0, dCl_dw = pred_l < ag.left ? -0.5f : 0;
  const float dCr_dx = pred_r > ag.right ? 1 : 0, dCr_dw = pred_r > ag.right ? 0.5f : 0;
  const float dCw_dx = dCr_dx - dCl_dx, dCw_dy = 0, dCw_dw = dCr_dw - dCl_dw, dCw_dh = 0;
  const float dCh_dx = 0, dCh_dy = dCb_dy - dCt_dy, dCh_dw = 0, dCh_dh = dCb_dh - dCt_dh;
  float p_dx = p_dl + p_dr;
  float p_dy = p_dt + p_db;
  float p_dw = (p_dr - p_dl);
  float p_dh = (p_db - p_dt);
  if (iou_type == IOU_DIOU) {
    if (C > 0) {
      p_dx += (2 * (gt.x - pred.x) * C - (2 * Cw * dCw_dx + 2 * Ch * dCh_dx) * S) / YSQ(C);
      p_dy += (2 * (gt.y - pred.y) * C - (2 * Cw * dCw_dy + 2 * Ch * dCh_dy) * S) / YSQ(C);
      p_dw += (2 * Cw * dCw_dw + 2 * Ch * dCh_dw) * S / YSQ(C);
      p_dh += (2 * Cw * dCw_dh + 2 * Ch * dCh_dh) * S / YSQ(C);
    }
    if (Iw <= 0 || Ih <= 0) {
      p_dx = (2 * (gt.x - pred.x) * C - (2 * Cw * dCw_dx + 2 * Ch * dCh_dx) * S) / YSQ(C);
      p_dy = (2 * (gt.y - pred.y) * C - (2 * Cw * dCw_dy + 2 * Ch * dCh_dy) * S) / YSQ(C);
      p_dw = (2 * Cw * dCw_dw + 2 * Ch * dCh_dw) * S / YSQ(C);
      p_dh = (2 * Cw * dCw_dh + 2 * Ch * dCh_dh) * S / YSQ(C);
    }
  }
  if (iou_type == IOU_CIOU) {
    const float ar_gt = gt.w / gt.h, ar_pred = pred.w / pred.h;
    const double dat = atan((double)ar_gt) - atan((double)ar_pred);
    const float ar_loss = (float)(4 / YSQ(M_PI) * YSQ(dat));
    const float alpha = (float)(ar_loss / (1 - I / U + ar_loss + 0.000001));
    const float ar_dw = (float)(8 / YSQ(M_PI) * dat * pred.h);
    const float ar_dh = (float)(-8 / YSQ(M_PI) * dat * pred.w);
    if (C > 0) {
      p_dx += (2 * (gt.x - pred.x) * C - (2 * Cw * dCw_dx + 2 * Ch * dCh_dx) * S) / YSQ(C);
      p_dy += (2 * (gt.y - pred.y) * C - (2 * Cw * dCw_dy + 2 * Ch * dCh_dy) * S) / YSQ(C);
      p_dw += (2 * Cw * dCw_dw + 2 * Ch * dCh_dw) * S / YSQ(C) + alpha * ar_dw;
      p_dh += (2 * Cw * dCw_dh + 2 * Ch * dCh_dh) * S / YSQ(C) + alpha * ar_dh;
    }
    if (Iw <= 0 || Ih <= 0) {
      p_dx = (2 * (gt.x - pred.x) * C - (2 * Cw * dCw_dx + 2 * Ch * dCh_dx) * S) / YSQ(C);
      p_dy = (2 * (gt.y - pred.y) * C - (2 * Cw * dCw_dy + 2 * Ch * dCh_dy) * S) / YSQ(C);
      p_dw = (2 * Cw * dCw_dw + 2 * Ch * dCh_dw) * S / YSQ(C) + alpha * ar_dw;
      p_dh = (2 * Cw * dCw_dh + 2 * Ch * dCh_dh) * S / YSQ(C) + alpha * ar_dh;
    }
  }
  return DxRep{p_dx, p_dy, p_dw, p_dh};
}

struct LossParams {
  int B, H, W, n, total, classes, max_boxes, net_w, net_h;
  int iou_loss, iou_thresh_kind, focal_loss;
  float ignore_thresh, truth_thresh, iou_thresh, iou_normalizer, cls_normalizer, label_smooth_eps, max_delta;
  int gaussian;         // 1: [Gaussian_yolo] entry layout mu_x,s_x,mu_y,s_y,mu_w,s_w,mu_h,s_h,obj,classes (gaussian_yolo_layer.cpp)
  float uc_normalizer;
  const float* out;     // decoded head, fp32 NCHW
  const float* truth;   // [B][max_boxes*5]
  const float* biases;  // anchors, 2*total
  const int* mask;      // n
  float* delta;         // fp32 NCHW, same shape as out
  float* stats;         // [0] cost [1] count [2] tot_iou_loss [3] tot_giou_loss [4] sum delta^2 [5] sum delta^2 w/o box
                        // [6] tot_iou [7] class_count
};

__device__ __forceinline__ int box_entries(const LossParams& p) { return p.gaussian ? 8 : 4; }  // obj follows, then classes
// yolo_layer.cpp:380-386 / EntryGaussianIndex gaussian_yolo_layer.cpp:476-484
__device__ __forceinline__ int entry_index(const LossParams& p, int b, int location, int entry) {
  const int wh = p.W * p.H;
  const int E = box_entries(p) + 1 + p.classes;
  const int n = location / wh, loc = location % wh;
  return b * (p.n * E * wh) + n * wh * E + entry * wh + loc;
}
__device__ __forceinline__ BoxF get_yolo_box(const float* x, const float* biases, int n, int index, int i, int j, int lw,
                                             int lh, int w, int h, int stride) {  // yolo_layer.cpp:139-148
  BoxF b;
  b.x = (i + x[index + 0 * stride]) / lw;
  b.y = (j + x[index + 1 * stride]) / lh;
  b.w = expf(x[index + 2 * stride]) * biases[2 * n] / w;
  b.h = expf(x[index + 3 * stride]) * biases[2 * n + 1] / h;
  return b;
}
__device__ __forceinline__ float fix_nan_inf(float v) { return (isnan(v) || isinf(v)) ? 0.f : v; }
__device__ __forceinline__ float clip_value(float v, float m) { return v > m ? m : (v < -m ? -m : v); }

// delta_yolo_box (yolo_layer.cpp:172-262), accumulate = 1; executed by ONE thread
__device__ IousF delta_yolo_box(const LossParams& p, const BoxF& truth, int n, int index, int i, int j, float scale,
                                float iou_normalizer) {
  const int stride = p.W * p.H;
  const float* x = p.out;
  float* delta = p.delta;
  IousF all;
  BoxF pred = get_yolo_box(x, p.biases, n, index, i, j, p.W, p.H, p.net_w, p.net_h, stride);
  all.iou = b_iou(pred, truth);
  all.giou = b_giou(pred, truth);
  all.diou = b_diou(pred, truth, 0.6f);
  all.ciou = b_ciou(pred, truth);
  if (pred.w == 0) pred.w = 1.0f;
  if (pred.h == 0) pred.h = 1.0f;
  if (p.iou_loss == IOU_MSE) {
    const float tx = (truth.x * p.W - i);
    const float ty = (truth.y * p.H - j);
    const float tw = logf(truth.w * p.net_w / p.biases[2 * n]);
    const float th = logf(truth.h * p.net_h / p.biases[2 * n + 1]);
    delta[index + 0 * stride] += scale * (tx - x[index + 0 * stride]) * iou_normalizer;
    delta[index + 1 * stride] += scale * (ty - x[index + 1 * stride]) * iou_normalizer;
    delta[index + 2 * stride] += scale * (tw - x[index + 2 * stride]) * iou_normalizer;
    delta[index + 3 * stride] += scale * (th - x[index + 3 * stride]) * iou_normalizer;
  } else {
    const DxRep g = b_dx_iou(pred, truth, p.iou_loss);
    float dx = g.dt, dy = g.db, dw = g.dl, dh = g.dr;  // "dt, db, dl, dr are already dx, dy, dw, dh" (:229-232)
    dw *= expf(x[index + 2 * stride]);
    dh *= expf(x[index + 3 * stride]);
    dx *= iou_normalizer; dy *= iou_normalizer; dw *= iou_normalizer; dh *= iou_normalizer;
    dx = fix_nan_inf(dx); dy = fix_nan_inf(dy); dw = fix_nan_inf(dw); dh = fix_nan_inf(dh);
    if (p.max_delta != FLT_MAX) {
      dx = clip_value(dx, p.max_delta); dy = clip_value(dy, p.max_delta);
      dw = clip_value(dw, p.max_delta); dh = clip_value(dh, p.max_delta);
    }
    delta[index + 0 * stride] += dx;
    delta[index + 1 * stride] += dy;
    delta[index + 2 * stride] += dw;
    delta[index + 3 * stride] += dh;
  }
  return all;
}

// GetGaussianYoloBox (gaussian_yolo_layer.cpp:151-178), yolo_point = center
__device__ __forceinline__ BoxF get_gaussian_box(const float* x, const float* biases, int n, int index, int i, int j, int lw,
                                                 int lh, int w, int h, int stride) {
  BoxF b;
  b.w = expf(x[index + 4 * stride]) * biases[2 * n] / w;
  b.h = expf(x[index + 6 * stride]) * biases[2 * n + 1] / h;
  b.x = (i + x[index + 0 * stride]) / lw;
  b.y = (j + x[index + 2 * stride]) / lh;
  return b;
}
__device__ __forceinline__ BoxF get_box(const LossParams& p, int n, int index, int i, int j) {
  const int wh = p.W * p.H;
  return p.gaussian ? get_gaussian_box(p.out, p.biases, n, index, i, j, p.W, p.H, p.net_w, p.net_h, wh)
                    : get_yolo_box(p.out, p.biases, n, index, i, j, p.W, p.H, p.net_w, p.net_h, wh);
}

// delta_gaussian_yolo_box (gaussian_yolo_layer.cpp:198-402), accumulate = 1, yolo_point = center; ONE thread.
// The reference mixes float and double (the literals 1./2., M_PI, pow): the same promotions are kept.
__device__ float delta_gaussian_box(const LossParams& p, const BoxF& truth, int n, int index, int i, int j, float scale,
                                    float iou_normalizer) {
  const int stride = p.W * p.H;
  const float* x = p.out;
  float* delta = p.delta;
  BoxF pred = get_gaussian_box(x, p.biases, n, index, i, j, p.W, p.H, p.net_w, p.net_h, stride);
  float iou = b_iou(pred, truth);
  const float giou = b_giou(pred, truth);
  if (pred.w == 0) pred.w = 1.0f;
  if (pred.h == 0) pred.h = 1.0f;
  const float sigma_const = 0.3f;
  const float epsi = (float)pow(10.0, -9.0);
  const float t4[4] = {truth.x * p.W - i, truth.y * p.H - j, logf(truth.w * p.net_w / p.biases[2 * n]),
                       logf(truth.h * p.net_h / p.biases[2 * n + 1])};
  float dl[4], du[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float mu = x[index + (2 * k) * stride], sg = x[index + (2 * k + 1) * stride];
    const float d = t4[k] - mu;
    const float in_exp = d / sg;
    const float in_exp_2 = (float)pow((double)in_exp, 2.0);
    const float normal = (float)(exp((double)in_exp_2 * (-1. / 2.)) / (sqrt(M_PI * 2.0) * (double)(sg + sigma_const)));
    const float temp = (float)((1. / 2.) * 1. / (double)(normal + epsi) * (double)normal * (double)scale);
    dl[k] = (float)((double)(temp * in_exp) * (1. / (double)sg));
    du[k] = (float)((double)temp * ((double)(in_exp_2 / sg) - 1. / (double)(sg + sigma_const)));
  }
  if (p.iou_loss != IOU_MSE) {
    iou = giou;
    const DxRep g = b_dx_iou(pred, truth, p.iou_loss);
    dl[0] = g.dt; dl[1] = g.db;
    dl[2] = g.dl * expf(x[index + 4 * stride]);
    dl[3] = g.dr * expf(x[index + 6 * stride]);
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    dl[k] = fix_nan_inf(dl[k] * iou_normalizer);
    du[k] = fix_nan_inf(du[k] * p.uc_normalizer);
    if (p.max_delta != FLT_MAX) {
      dl[k] = clip_value(dl[k], p.max_delta);
      du[k] = clip_value(du[k], p.max_delta);
    }
    delta[index + (2 * k) * stride] += dl[k];
    delta[index + (2 * k + 1) * stride] += du[k];
  }
  return iou;
}

// DeltaGaussianYoloClass (gaussian_yolo_layer.cpp:429-461): the "already marked" test looks at the FIRST class entry
// (delta[index]), no focal loss, no NaN filter.  Lanes split the class loop like delta_yolo_class.
__device__ float delta_gaussian_class(const LossParams& p, int index, int class_id, int lane, int nl) {
  const int stride = p.W * p.H;
  const float* output = p.out;
  float* delta = p.delta;
  const bool marked = delta[index] != 0.f;
  if (nl > 1) __syncwarp();
  if (marked) {
    float y_true = 1;
    if (p.label_smooth_eps) y_true = y_true * (1 - p.label_smooth_eps) + 0.5f * p.label_smooth_eps;
    if (lane == 0) delta[index + stride * class_id] = y_true - output[index + stride * class_id];
    if (nl > 1) __syncwarp();
    return output[index + stride * class_id];
  }
  for (int n = lane; n < p.classes; n += nl) {
    float y_true = ((n == class_id) ? 1 : 0);
    if (p.label_smooth_eps) y_true = y_true * (1 - p.label_smooth_eps) + 0.5f * p.label_smooth_eps;
    delta[index + stride * n] = y_true - output[index + stride * n];
  }
  if (nl > 1) __syncwarp();
  return output[index + stride * class_id];
}

// delta_yolo_class (yolo_layer.cpp:285-357); the class loop is strided over `nl` cooperating lanes (lane, nl) =
// (0, 1) for a single thread.  Returns the avg_cat contribution.
__device__ float delta_yolo_class(const LossParams& p, int index, int class_id, int lane, int nl) {
  const int stride = p.W * p.H;
  const float* output = p.out;
  float* delta = p.delta;
  if (delta[index + stride * class_id]) {  // this class was already marked by an earlier truth
    float y_true = 1;
    if (p.label_smooth_eps) y_true = y_true * (1 - p.label_smooth_eps) + 0.5f * p.label_smooth_eps;
    const float r = y_true - output[index + stride * class_id];
    if (nl > 1) __syncwarp();
    if (lane == 0 && !isnan(r) && !isinf(r)) delta[index + stride * class_id] = r;
    if (nl > 1) __syncwarp();
    return output[index + stride * class_id];
  }
  if (nl > 1) __syncwarp();
  if (p.focal_loss) {
    const float alpha = 0.5f;
    const int ti = index + stride * class_id;
    const float pt = output[ti] + 0.000000000000001F;
    const float grad = -(1 - pt) * (2 * pt * logf(pt) + pt - 1);
    for (int n = lane; n < p.classes; n += nl) {
      float d = (((n == class_id) ? 1 : 0) - output[index + stride * n]);
      d *= alpha * grad;
      delta[index + stride * n] = d;
    }
  } else {
    for (int n = lane; n < p.classes; n += nl) {
      float y_true = ((n == class_id) ? 1 : 0);
      if (p.label_smooth_eps) y_true = y_true * (1 - p.label_smooth_eps) + 0.5f * p.label_smooth_eps;
      const float r = y_true - output[index + stride * n];
      if (!isnan(r) && !isinf(r)) delta[index + stride * n] = r;
    }
  }
  if (nl > 1) __syncwarp();
  return output[index + stride * class_id];
}

// ---- 1. per (image, anchor, cell): objectness delta from the best IoU over the truths (yolo_layer.cpp:431-522) ----
__global__ void yolo_loss_cells_kernel(LossParams p) {
  const int wh = p.W * p.H;
  const long long total = (long long)p.B * p.n * wh;
  for (long long t0 = blockIdx.x * (long long)blockDim.x + threadIdx.x; t0 < total; t0 += (long long)gridDim.x * blockDim.x) {
    const int loc = (int)(t0 % wh);
    const int n = (int)((t0 / wh) % p.n);
    const int b = (int)(t0 / ((long long)wh * p.n));
    const int i = loc % p.W, j = loc / p.W;
    const int box_index = entry_index(p, b, n * wh + loc, 0);
    const int obj_index = entry_index(p, b, n * wh + loc, box_entries(p));
    const int class_index = entry_index(p, b, n * wh + loc, box_entries(p) + 1);
    const BoxF pred = get_box(p, p.mask[n], box_index, i, j);
    float best_match_iou = 0, best_iou = 0;
    int best_t = 0;
    int class_match = -1;  // compare_yolo_class(..., 0.25f) does not depend on the truth: evaluated once, on demand
    const float* tr = p.truth + (long long)b * p.max_boxes * 5;
    for (int t = 0; t < p.max_boxes; ++t) {
      const BoxF truth{tr[t * 5 + 0], tr[t * 5 + 1], tr[t * 5 + 2], tr[t * 5 + 3]};
      const int class_id = (int)tr[t * 5 + 4];
      if (class_id >= p.classes || class_id < 0) continue;
      if (!truth.x) break;
      if (class_match < 0) {
        class_match = 0;
        for (int c = 0; c < p.classes; ++c)
          if (p.out[class_index + wh * c] > 0.25f) { class_match = 1; break; }
      }
      const float iou = b_iou(pred, truth);
      if (iou > best_match_iou && class_match == 1) best_match_iou = iou;
      if (iou > best_iou) { best_iou = iou; best_t = t; }
    }
    const float o = p.out[obj_index];
    float d = p.cls_normalizer * (0 - o);
    if (best_match_iou > p.ignore_thresh) d = 0;
    p.delta[obj_index] = d;
    if (best_iou > p.truth_thresh) {
      p.delta[obj_index] = p.cls_normalizer * (1 - o);
      const int class_id = (int)tr[best_t * 5 + 4];
      const BoxF truth{tr[best_t * 5 + 0], tr[best_t * 5 + 1], tr[best_t * 5 + 2], tr[best_t * 5 + 3]};
      if (p.gaussian) {
        delta_gaussian_class(p, class_index, class_id, 0, 1);
        delta_gaussian_box(p, truth, p.mask[n], box_index, i, j, (2 - truth.w * truth.h), p.iou_normalizer);
      } else {
        delta_yolo_class(p, class_index, class_id, 0, 1);
        delta_yolo_box(p, truth, p.mask[n], box_index, i, j, (2 - truth.w * truth.h), p.iou_normalizer);
      }
    }
  }
}

// ---- 2. one warp per image: the truths in order (yolo_layer.cpp:524-676) -------------------------------------------
__global__ void yolo_loss_truths_kernel(LossParams p) {
  const int b = blockIdx.x;
  const int lane = threadIdx.x;
  const int wh = p.W * p.H;
  const float* tr = p.truth + (long long)b * p.max_boxes * 5;
  float tot_iou = 0, tot_iou_loss = 0, tot_giou_loss = 0;
  int count = 0;
  for (int t = 0; t < p.max_boxes; ++t) {
    const BoxF truth{tr[t * 5 + 0], tr[t * 5 + 1], tr[t * 5 + 2], tr[t * 5 + 3]};
    const int class_id = (int)tr[t * 5 + 4];
    if (class_id >= p.classes || class_id < 0) continue;
    if (!truth.x) break;
    float best_iou = 0;
    int best_n = 0;
    const int i = (int)(truth.x * p.W);
    const int j = (int)(truth.y * p.H);
    BoxF truth_shift = truth;
    truth_shift.x = truth_shift.y = 0;
    for (int n = 0; n < p.total; ++n) {
      BoxF pred{0, 0, p.biases[2 * n] / p.net_w, p.biases[2 * n + 1] / p.net_h};
      const float iou = b_iou(pred, truth_shift);
      if (iou > best_iou) { best_iou = iou; best_n = n; }
    }
    for (int pass = 0; pass <= p.total; ++pass) {
      // pass 0: the best anchor (:563-605); passes 1..total: the other anchors over iou_thresh (:607-676)
      int n_abs, mask_n = -1;
      if (pass == 0) {
        n_abs = best_n;
      } else {
        n_abs = pass - 1;
        if (n_abs == best_n || !(p.iou_thresh < 1.0f)) continue;
      }
      for (int m = 0; m < p.n; ++m)
        if (p.mask[m] == n_abs) { mask_n = m; break; }  // int_index
      if (mask_n < 0) continue;
      if (pass > 0) {
        BoxF pred{0, 0, p.biases[2 * n_abs] / p.net_w, p.biases[2 * n_abs + 1] / p.net_h};
        const float iou = b_iou_kind(pred, truth_shift, p.iou_thresh_kind);
        if (!(iou > p.iou_thresh)) continue;
      }
      const int loc = mask_n * wh + j * p.W + i;
      const int box_index = entry_index(p, b, loc, 0);
      const int obj_index = entry_index(p, b, loc, box_entries(p));
      const int class_index = entry_index(p, b, loc, box_entries(p) + 1);
      __syncwarp();
      IousF all{0, 0, 0, 0};
      if (lane == 0) {
        if (p.gaussian) all.iou = all.giou = delta_gaussian_box(p, truth, n_abs, box_index, i, j, (2 - truth.w * truth.h), p.iou_normalizer);
        else all = delta_yolo_box(p, truth, n_abs, box_index, i, j, (2 - truth.w * truth.h), p.iou_normalizer);
        p.delta[obj_index] = p.cls_normalizer * (1 - p.out[obj_index]);
        tot_iou += all.iou;
        tot_iou_loss += 1 - all.iou;
        tot_giou_loss += 1 - all.giou;
        ++count;
      }
      __syncwarp();
      if (p.gaussian) delta_gaussian_class(p, class_index, class_id, lane, 32);
      else delta_yolo_class(p, class_index, class_id, lane, 32);
    }
  }
  if (lane == 0 && count > 0) {
    atomicAdd(&p.stats[1], (float)count);
    atomicAdd(&p.stats[2], tot_iou_loss);
    atomicAdd(&p.stats[3], tot_giou_loss);
    atomicAdd(&p.stats[6], tot_iou);
    atomicAdd(&p.stats[7], (float)count);
  }
}

// ---- 3. per (image, anchor, cell): averages_yolo_deltas (:264-283, :679-696), the two sums of squares of the cost
// (:708-733), and BackwardYoloLayer's prev.delta += delta (:774-777) into the head conv's NHWC bf16 delta -----------
__global__ void yolo_loss_finish_kernel(LossParams p, __nv_bfloat16* __restrict__ prev, long long prev_ld) {
  const int wh = p.W * p.H;
  const int BE = box_entries(p);
  const int E = BE + 1 + p.classes;
  const long long total = (long long)p.B * p.n * wh;
  float s_all = 0, s_nobox = 0;
  for (long long t0 = blockIdx.x * (long long)blockDim.x + threadIdx.x; t0 < total; t0 += (long long)gridDim.x * blockDim.x) {
    const int loc = (int)(t0 % wh);
    const int n = (int)((t0 / wh) % p.n);
    const int b = (int)(t0 / ((long long)wh * p.n));
    const int base = entry_index(p, b, n * wh + loc, 0);
    int in_box = 0;
    for (int c = 0; c < p.classes; ++c)
      if (p.delta[base + (BE + 1 + c) * wh] > 0) ++in_box;
    __nv_bfloat16* pv = prev ? prev + ((long long)b * wh + loc) * prev_ld + n * E : nullptr;
    for (int e = 0; e < E; ++e) {
      float d = p.delta[base + e * wh];
      if (e < BE && in_box > 0) {
        d /= in_box;
        p.delta[base + e * wh] = d;
      }
      s_all += d * d;
      if (e >= BE) s_nobox += d * d;
      if (pv) pv[e] = __float2bfloat16_rn(__bfloat162float(pv[e]) + d);
    }
  }
  __shared__ float sh[2][32];
  for (int off = 16; off; off >>= 1) {
    s_all += __shfl_xor_sync(0xffffffffu, s_all, off);
    s_nobox += __shfl_xor_sync(0xffffffffu, s_nobox, off);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { sh[0][warp] = s_all; sh[1][warp] = s_nobox; }
  __syncthreads();
  if (warp == 0) {
    const int nw = blockDim.x >> 5;
    float a = lane < nw ? sh[0][lane] : 0.f, q = lane < nw ? sh[1][lane] : 0.f;
    for (int off = 16; off; off >>= 1) {
      a += __shfl_xor_sync(0xffffffffu, a, off);
      q += __shfl_xor_sync(0xffffffffu, q, off);
    }
    if (lane == 0) {
      atomicAdd(&p.stats[4], a);
      atomicAdd(&p.stats[5], q);
    }
  }
}

// The same, tiled: one block per (image, 32 cells) stages all n*E delta entries of those cells in shared memory, so the
// fp32 NCHW planes are read and written 128 bytes per warp and the NHWC bf16 head delta is updated row by row
// (n*E consecutive channels per cell) instead of one 2-byte access per thread at a pitch of prev_ld.
__global__ void __launch_bounds__(256) yolo_loss_finish_tiled_kernel(LossParams p, __nv_bfloat16* __restrict__ prev,
                                                                     long long prev_ld) {
  extern __shared__ float tile[];  // [n*E][33] then int cnt[n][32]
  const int wh = p.W * p.H;
  const int BE = box_entries(p);
  const int E = BE + 1 + p.classes;
  const int CH = p.n * E;
  int* cnt = reinterpret_cast<int*>(tile + CH * 33);
  const int tiles = (wh + 31) / 32;
  const int b = blockIdx.x / tiles, loc0 = (blockIdx.x % tiles) * 32;
  float* dl = p.delta + (long long)b * CH * wh;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const bool loc_ok = loc0 + lane < wh;
  for (int ch = warp; ch < CH; ch += 8) tile[ch * 33 + lane] = loc_ok ? dl[(long long)ch * wh + loc0 + lane] : 0.f;
  __syncthreads();
  for (int t = threadIdx.x; t < p.n * 32; t += 256) {
    const int a = t >> 5, l = t & 31;
    int in_box = 0;
    for (int c = 0; c < p.classes; ++c)
      if (tile[(a * E + BE + 1 + c) * 33 + l] > 0) ++in_box;
    cnt[t] = in_box;
  }
  __syncthreads();
  float s_all = 0, s_nobox = 0;
  for (int ch = warp; ch < CH; ch += 8) {
    const int a = ch / E, e = ch - a * E;
    float d = tile[ch * 33 + lane];
    const int in_box = cnt[a * 32 + lane];
    if (e < BE && in_box > 0 && loc_ok) {
      d /= in_box;
      dl[(long long)ch * wh + loc0 + lane] = d;
      tile[ch * 33 + lane] = d;
    }
    s_all += d * d;  // cells past wh hold zeros
    if (e >= BE) s_nobox += d * d;
  }
  __syncthreads();
  if (prev) {
    for (int l = warp; l < 32 && loc0 + l < wh; l += 8) {
      __nv_bfloat16* pv = prev + ((long long)b * wh + loc0 + l) * prev_ld;
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
  __shared__ float sh[2][8];
  for (int off = 16; off; off >>= 1) {
    s_all += __shfl_xor_sync(0xffffffffu, s_all, off);
    s_nobox += __shfl_xor_sync(0xffffffffu, s_nobox, off);
  }
  if (lane == 0) { sh[0][warp] = s_all; sh[1][warp] = s_nobox; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0, q = 0;
    for (int w = 0; w < 8; ++w) { a += sh[0][w]; q += sh[1][w]; }
    atomicAdd(&p.stats[4], a);
    atomicAdd(&p.stats[5], q);
  }
}

// *(l->cost) (yolo_layer.cpp:735-754)
__global__ void yolo_loss_cost_kernel(LossParams p) {
  float count = p.stats[1];
  if (count == 0) count = 1;
  const float classification_loss = p.cls_normalizer * p.stats[5];
  if (p.iou_loss == IOU_MSE || p.gaussian) {  // Gaussian: *(l->cost) = |delta|^2 always (gaussian_yolo_layer.cpp:835)
    p.stats[0] = p.stats[4];
  } else {
    const float tot = p.iou_loss == IOU_GIOU ? p.stats[3] : p.stats[2];
    p.stats[0] = p.iou_normalizer * (tot / count) + classification_loss;
  }
}

}  // namespace yc

using namespace yc;

extern "C" int yc_yolo_loss(const yc_yolo_loss_desc* d, const float* out_nchw, const float* truth, const float* anchor_wh,
                            const int* mask, int n, int h, int w, float* delta_nchw, const yc_tensor* prev_delta,
                            float* stats, void* stream) {
  YC_REQUIRE(d && out_nchw && truth && anchor_wh && mask && delta_nchw && stats && n > 0 && h > 0 && w > 0,
             "yc_yolo_loss: bad arguments");
  YC_REQUIRE(d->anchors > 0 && d->total >= d->anchors && d->classes > 0 && d->max_boxes > 0, "yc_yolo_loss: bad descriptor");
  const int E = (d->gaussian ? 9 : 5) + d->classes;
  if (prev_delta)
    YC_REQUIRE(prev_delta->dtype == YC_BF16 && prev_delta->n == n && prev_delta->h == h && prev_delta->w == w &&
                   prev_delta->ldc >= d->anchors * E,
               "yc_yolo_loss: prev_delta must be the head conv's NHWC bf16 delta");
  cudaStream_t st = (cudaStream_t)stream;
  LossParams p;
  p.B = n; p.H = h; p.W = w; p.n = d->anchors; p.total = d->total; p.classes = d->classes; p.max_boxes = d->max_boxes;
  p.net_w = d->net_w; p.net_h = d->net_h; p.iou_loss = d->iou_loss; p.iou_thresh_kind = d->iou_thresh_kind;
  p.focal_loss = d->focal_loss; p.ignore_thresh = d->ignore_thresh; p.truth_thresh = d->truth_thresh;
  p.iou_thresh = d->iou_thresh; p.iou_normalizer = d->iou_normalizer; p.cls_normalizer = d->cls_normalizer;
  p.label_smooth_eps = d->label_smooth_eps; p.max_delta = d->max_delta;
  p.gaussian = d->gaussian ? 1 : 0; p.uc_normalizer = d->uc_normalizer;
  YC_REQUIRE(!(p.gaussian && p.focal_loss), "yc_yolo_loss: [Gaussian_yolo] has no focal loss");
  p.out = out_nchw; p.truth = truth; p.biases = anchor_wh; p.mask = mask; p.delta = delta_nchw; p.stats = stats;
  const size_t elems = (size_t)n * d->anchors * E * h * w;
  YC_CUDA(cudaMemsetAsync(delta_nchw, 0, elems * 4, st));
  YC_CUDA(cudaMemsetAsync(stats, 0, 8 * sizeof(float), st));
  const long long cells = (long long)n * d->anchors * h * w;
  const int grid = (int)std::max<long long>(1, std::min<long long>(ceil_div_ll(cells, 128), (long long)sm_count() * 16));
  yolo_loss_cells_kernel<<<grid, 128, 0, st>>>(p);
  YC_LAUNCH_CHECK();
  yolo_loss_truths_kernel<<<n, 32, 0, st>>>(p);
  YC_LAUNCH_CHECK();
  const size_t tile_bytes = (size_t)d->anchors * E * 33 * 4 + (size_t)d->anchors * 32 * 4;
  if (tile_bytes <= 48 * 1024 && (!prev_delta || prev_delta->ldc % 2 == 0)) {
    yolo_loss_finish_tiled_kernel<<<n * ((h * w + 31) / 32), 256, tile_bytes, st>>>(
        p, prev_delta ? (__nv_bfloat16*)prev_delta->ptr : nullptr, prev_delta ? prev_delta->ldc : 0);
  } else {
    yolo_loss_finish_kernel<<<grid, 128, 0, st>>>(p, prev_delta ? (__nv_bfloat16*)prev_delta->ptr : nullptr,
                                                  prev_delta ? prev_delta->ldc : 0);
  }
  YC_LAUNCH_CHECK();
  yolo_loss_cost_kernel<<<1, 1, 0, st>>>(p);
  YC_LAUNCH_CHECK();
  return YC_OK;
}
