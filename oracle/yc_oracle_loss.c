/* TEST INFRASTRUCTURE — CPU restatement of the reference's YOLO loss (training half of ForwardYoloLayer,
 * yolo_layer.cpp:409-772) and of the IoU-family math it uses (box.cpp:36-351).  Plain C, serial, fp32 like the
 * reference.  Pinned against the UNMODIFIED reference by tests/test_oracle_train_pin.py.  Used by the training
 * walker oracle/pytrain.py; never by the product. */
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "yc_oracle.h"

typedef struct { float x, y, w, h; } obox;
typedef struct { float dt, db, dl, dr; } odx;
typedef struct { float iou, giou, diou, ciou; } oious;
enum { K_IOU = 0, K_GIOU = 1, K_MSE = 2, K_DIOU = 3, K_CIOU = 4 };   /* IOU_LOSS box.h:6-13 */
#define SQ(a) ((a) * (a))

static float o_overlap(float x1, float w1, float x2, float w2) {       /* box.cpp:36-46 */
  float l1 = x1 - w1 / 2, l2 = x2 - w2 / 2;
  float left = l1 > l2 ? l1 : l2;
  float r1 = x1 + w1 / 2, r2 = x2 + w2 / 2;
  float right = r1 < r2 ? r1 : r2;
  return right - left;
}
static float o_intersect(obox a, obox b) {                              /* box.cpp:48-57 */
  float w = o_overlap(a.x, a.w, b.x, b.w), h = o_overlap(a.y, a.h, b.y, b.h);
  if (w < 0 || h < 0) return 0;
  return w * h;
}
static float o_union(obox a, obox b) { return a.w * a.h + b.w * b.h - o_intersect(a, b); }
static float o_iou(obox a, obox b) {                                    /* box.cpp:64-72 */
  float I = o_intersect(a, b), U = o_union(a, b);
  if (fabsf(I) < FLT_EPSILON || fabsf(U) < FLT_EPSILON) return 0;
  return I / U;
}
static void o_abs2(obox a, obox b, float* left, float* right, float* top, float* bottom) {  /* box.cpp:353-370 */
  *left = fminf(a.x - a.w / 2.0f, b.x - b.w / 2.0f);
  *right = fmaxf(a.x + a.w / 2.0f, b.x + b.w / 2.0f);
  *top = fminf(a.y - a.h / 2.0f, b.y - b.h / 2.0f);
  *bottom = fmaxf(a.y + a.h / 2.0f, b.y + b.h / 2.0f);
}
static float o_giou(obox a, obox b) {                                   /* box.cpp:113-128 */
  float l, r, t, bo;
  o_abs2(a, b, &l, &r, &t, &bo);
  float w = r - l, h = bo - t, c = w * h, iou = o_iou(a, b);
  if (fabsf(c) < FLT_EPSILON) return iou;
  return iou - (c - o_union(a, b)) / c;
}
static float o_diou(obox a, obox b, float beta) {                       /* box.cpp:96-111 */
  float l, r, t, bo;
  o_abs2(a, b, &l, &r, &t, &bo);
  float w = r - l, h = bo - t, c = w * w + h * h, iou = o_iou(a, b);
  if (fabsf(c) < FLT_EPSILON) return iou;
  float d = SQ(a.x - b.x) + SQ(a.y - b.y);
  return iou - (float)pow(d / c, beta);
}
static float o_ciou(obox a, obox b) {                                   /* box.cpp:74-94 */
  float l, r, t, bo;
  o_abs2(a, b, &l, &r, &t, &bo);
  float w = r - l, h = bo - t, c = w * w + h * h, iou = o_iou(a, b);
  if (fabsf(c) < FLT_EPSILON) return iou;
  float u = SQ(a.x - b.x) + SQ(a.y - b.y);
  float d = u / c;
  float ar_gt = b.w / b.h, ar_pred = a.w / a.h;
  float ar_loss = 4 / SQ(M_PI) * SQ(atan(ar_gt) - atan(ar_pred));
  float alpha = ar_loss / (1 - iou + ar_loss + 0.000001);
  return iou - (d + alpha * ar_loss);
}
static float o_iou_kind(obox a, obox b, int kind) {                     /* box.cpp:136-151 */
  switch (kind) {
    case K_GIOU: return o_giou(a, b);
    case K_MSE: return sqrtf(SQ(a.x - b.x) + SQ(a.y - b.y) + SQ(a.w - b.w) + SQ(a.h - b.h));
    case K_DIOU: return o_diou(a, b, 0.6f);
    case K_CIOU: return o_ciou(a, b);
  }
  return o_iou(a, b);
}

static odx o_dx_iou(obox pred, obox gt, int iou_type) {                 /* Box::DxIou box.cpp:153-351 */
  float a_top = pred.y - pred.h / 2.0f, a_bottom = pred.y + pred.h / 2.0f;
  float a_left = pred.x - pred.w / 2.0f, a_right = pred.x + pred.w / 2.0f;
  float pred_t = fminf(a_top, a_bottom), pred_b = fmaxf(a_top, a_bottom);
  float pred_l = fminf(a_left, a_right), pred_r = fmaxf(a_left, a_right);
  float g_top = gt.y - gt.h / 2.0f, g_bottom = gt.y + gt.h / 2.0f, g_left = gt.x - gt.w / 2.0f, g_right = gt.x + gt.w / 2.0f;
  float X = (pred_b - pred_t) * (pred_r - pred_l);
  float Xhat = (g_bottom - g_top) * (g_right - g_left);
  float Ih = fminf(pred_b, g_bottom) - fmaxf(pred_t, g_top);
  float Iw = fminf(pred_r, g_right) - fmaxf(pred_l, g_left);
  float I = Iw * Ih;
  float U = X + Xhat - I;
  float S = SQ(pred.x - gt.x) + SQ(pred.y - gt.y);
  float giou_Cw = fmaxf(pred_r, g_right) - fminf(pred_l, g_left);
  float giou_Ch = fmaxf(pred_b, g_bottom) - fminf(pred_t, g_top);
  float giou_C = giou_Cw * giou_Ch;
  float dX_wrt_t = -1 * (pred_r - pred_l), dX_wrt_b = pred_r - pred_l;
  float dX_wrt_l = -1 * (pred_b - pred_t), dX_wrt_r = pred_b - pred_t;
  float dI_wrt_t = pred_t > g_top ? (-1 * Iw) : 0, dI_wrt_b = pred_b < g_bottom ? Iw : 0;
  float dI_wrt_l = pred_l > g_left ? (-1 * Ih) : 0, dI_wrt_r = pred_r < g_right ? Ih : 0;
  float dU_wrt_t = dX_wrt_t - dI_wrt_t, dU_wrt_b = dX_wrt_b - dI_wrt_b;
  float dU_wrt_l = dX_wrt_l - dI_wrt_l, dU_wrt_r = dX_wrt_r - dI_wrt_r;
  float dC_wrt_t = pred_t < g_top ? (-1 * giou_Cw) : 0, dC_wrt_b = pred_b > g_bottom ? giou_Cw : 0;
  float dC_wrt_l = pred_l < g_left ? (-1 * giou_Ch) : 0, dC_wrt_r = pred_r > g_right ? giou_Ch : 0;
  float p_dt = 0, p_db = 0, p_dl = 0, p_dr = 0;
  if (U > 0) {
    p_dt = ((U * dI_wrt_t) - (I * dU_wrt_t)) / (U * U);
    p_db = ((U * dI_wrt_b) - (I * dU_wrt_b)) / (U * U);
    p_dl = ((U * dI_wrt_l) - (I * dU_wrt_l)) / (U * U);
    p_dr = ((U * dI_wrt_r) - (I * dU_wrt_r)) / (U * U);
  }
  p_dt = a_top < a_bottom ? p_dt : p_db;
  p_db = a_top < a_bottom ? p_db : p_dt;
  p_dl = a_left < a_right ? p_dl : p_dr;
  p_dr = a_left < a_right ? p_dr : p_dl;
  if (iou_type == K_GIOU) {
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
  float Ct = fminf(pred.y - pred.h / 2, gt.y - gt.h / 2), Cb = fmaxf(pred.y + pred.h / 2, gt.y + gt.h / 2);
  float Cl = fminf(pred.x - pred.w / 2, gt.x - gt.w / 2), Cr = fmaxf(pred.x + pred.w / 2, gt.x + gt.w / 2);
  float Cw = Cr - Cl, Ch = Cb - Ct;
  float C = SQ(Cw) + SQ(Ch);
  float dCt_dy = pred_t < g_top ? 1 : 0, dCt_dh = pred_t < g_top ? -0.5 : 0;
  float dCb_dy = pred_b > g_bottom ? 1 : 0, dCb_dh = pred_b > g_bottom ? 0.5 : 0;
  float dCl_dx = pred_l < g_left ? 1 : 0, dCl_dw = pred_l < g_left ? -0.5 : 0;
  float dCr_dx = pred_r > g_right ? 1 : 0, dCr_dw = pred_r > g_right ? 0.5 : 0;
  float dCw_dx = dCr_dx - dCl_dx, dCw_dy = 0, dCw_dw = dCr_dw - dCl_dw, dCw_dh = 0;
  float dCh_dx = 0, dCh_dy = dCb_dy - dCt_dy, dCh_dw = 0, dCh_dh = dCb_dh - dCt_dh;
  float p_dx = p_dl + p_dr, p_dy = p_dt + p_db, p_dw = (p_dr - p_dl), p_dh = (p_db - p_dt);
  if (iou_type == K_DIOU) {
    if (C > 0) {
      p_dx += (2 * (gt.x - pred.x) * C - (2 * Cw * dCw_dx + 2 * Ch * dCh_dx) * S) / SQ(C);
      p_dy += (2 * (gt.y - pred.y) * C - (2 * Cw * dCw_dy + 2 * Ch * dCh_dy) * S) / SQ(C);
      p_dw += (2 * Cw * dCw_dw + 2 * Ch * dCh_dw) * S / SQ(C);
      p_dh += (2 * Cw * dCw_dh + 2 * Ch * dCh_dh) * S / SQ(C);
    }
    if (Iw <= 0 || Ih <= 0) {
      p_dx = (2 * (gt.x - pred.x) * C - (2 * Cw * dCw_dx + 2 * Ch * dCh_dx) * S) / SQ(C);
      p_dy = (2 * (gt.y - pred.y) * C - (2 * Cw * dCw_dy + 2 * Ch * dCh_dy) * S) / SQ(C);
      p_dw = (2 * Cw * dCw_dw + 2 * Ch * dCh_dw) * S / SQ(C);
      p_dh = (2 * Cw * dCw_dh + 2 * Ch * dCh_dh) * S / SQ(C);
    }
  }
  if (iou_type == K_CIOU) {
    float ar_gt = gt.w / gt.h, ar_pred = pred.w / pred.h;
    float ar_loss = 4 / SQ(M_PI) * SQ(atan(ar_gt) - atan(ar_pred));
    float alpha = ar_loss / (1 - I / U + ar_loss + 0.000001);
    float ar_dw = 8 / SQ(M_PI) * (atan(ar_gt) - atan(ar_pred)) * pred.h;
    float ar_dh = -8 / SQ(M_PI) * (atan(ar_gt) - atan(ar_pred)) * pred.w;
    if (C > 0) {
      p_dx += (2 * (gt.x - pred.x) * C - (2 * Cw * dCw_dx + 2 * Ch * dCh_dx) * S) / SQ(C);
      p_dy += (2 * (gt.y - pred.y) * C - (2 * Cw * dCw_dy + 2 * Ch * dCh_dy) * S) / SQ(C);
      p_dw += (2 * Cw * dCw_dw + 2 * Ch * dCh_dw) * S / SQ(C) + alpha * ar_dw;
      p_dh += (2 * Cw * dCw_dh + 2 * Ch * dCh_dh) * S / SQ(C) + alpha * ar_dh;
    }
    if (Iw <= 0 || Ih <= 0) {
      p_dx = (2 * (gt.x - pred.x) * C - (2 * Cw * dCw_dx + 2 * Ch * dCh_dx) * S) / SQ(C);
      p_dy = (2 * (gt.y - pred.y) * C - (2 * Cw * dCw_dy + 2 * Ch * dCh_dy) * S) / SQ(C);
      p_dw = (2 * Cw * dCw_dw + 2 * Ch * dCh_dw) * S / SQ(C) + alpha * ar_dw;
      p_dh = (2 * Cw * dCw_dh + 2 * Ch * dCh_dh) * S / SQ(C) + alpha * ar_dh;
    }
  }
  odx r = {p_dx, p_dy, p_dw, p_dh};
  return r;
}

typedef struct {
  int B, H, W, n, total, classes, max_boxes, net_w, net_h, iou_loss, iou_thresh_kind, focal_loss;
  float ignore_thresh, truth_thresh, iou_thresh, iou_normalizer, cls_normalizer, label_smooth_eps, max_delta;
  const float* out; const float* truth; const float* biases; const int* mask; float* delta;
} oloss;

static int o_entry(const oloss* p, int b, int location, int entry) {      /* EntryIndex yolo_layer.cpp:380-386 */
  int wh = p->W * p->H;
  int n = location / wh, loc = location % wh;
  return b * (p->n * (5 + p->classes) * wh) + n * wh * (5 + p->classes) + entry * wh + loc;
}
static obox o_get_box(const float* x, const float* biases, int n, int index, int i, int j, int lw, int lh, int w, int h,
                      int stride) {                                       /* GetYoloBox yolo_layer.cpp:139-148 */
  obox b;
  b.x = (i + x[index + 0 * stride]) / lw;
  b.y = (j + x[index + 1 * stride]) / lh;
  b.w = expf(x[index + 2 * stride]) * biases[2 * n] / w;
  b.h = expf(x[index + 3 * stride]) * biases[2 * n + 1] / h;
  return b;
}
static float o_fix(float v) { return (isnan(v) || isinf(v)) ? 0 : v; }
static float o_clip(float v, float m) { return v > m ? m : (v < -m ? -m : v); }

static oious o_delta_box(const oloss* p, obox truth, int n, int index, int i, int j, float scale, float iou_normalizer) {
  int stride = p->W * p->H;                                               /* delta_yolo_box yolo_layer.cpp:172-262 */
  const float* x = p->out;
  float* delta = p->delta;
  oious all;
  obox pred = o_get_box(x, p->biases, n, index, i, j, p->W, p->H, p->net_w, p->net_h, stride);
  all.iou = o_iou(pred, truth); all.giou = o_giou(pred, truth); all.diou = o_diou(pred, truth, 0.6f); all.ciou = o_ciou(pred, truth);
  if (pred.w == 0) pred.w = 1.0;
  if (pred.h == 0) pred.h = 1.0;
  if (p->iou_loss == K_MSE) {
    float tx = (truth.x * p->W - i), ty = (truth.y * p->H - j);
    float tw = logf(truth.w * p->net_w / p->biases[2 * n]), th = logf(truth.h * p->net_h / p->biases[2 * n + 1]);
    delta[index + 0 * stride] += scale * (tx - x[index + 0 * stride]) * iou_normalizer;
    delta[index + 1 * stride] += scale * (ty - x[index + 1 * stride]) * iou_normalizer;
    delta[index + 2 * stride] += scale * (tw - x[index + 2 * stride]) * iou_normalizer;
    delta[index + 3 * stride] += scale * (th - x[index + 3 * stride]) * iou_normalizer;
  } else {
    odx g = o_dx_iou(pred, truth, p->iou_loss);
    float dx = g.dt, dy = g.db, dw = g.dl, dh = g.dr;
    dw *= expf(x[index + 2 * stride]);
    dh *= expf(x[index + 3 * stride]);
    dx *= iou_normalizer; dy *= iou_normalizer; dw *= iou_normalizer; dh *= iou_normalizer;
    dx = o_fix(dx); dy = o_fix(dy); dw = o_fix(dw); dh = o_fix(dh);
    if (p->max_delta != FLT_MAX) { dx = o_clip(dx, p->max_delta); dy = o_clip(dy, p->max_delta); dw = o_clip(dw, p->max_delta); dh = o_clip(dh, p->max_delta); }
    delta[index + 0 * stride] += dx; delta[index + 1 * stride] += dy;
    delta[index + 2 * stride] += dw; delta[index + 3 * stride] += dh;
  }
  return all;
}
static void o_delta_class(const oloss* p, int index, int class_id) {     /* delta_yolo_class yolo_layer.cpp:285-357 */
  int stride = p->W * p->H;
  const float* output = p->out;
  float* delta = p->delta;
  if (delta[index + stride * class_id]) {
    float y_true = 1;
    if (p->label_smooth_eps) y_true = y_true * (1 - p->label_smooth_eps) + 0.5 * p->label_smooth_eps;
    float r = y_true - output[index + stride * class_id];
    if (!isnan(r) && !isinf(r)) delta[index + stride * class_id] = r;
    return;
  }
  if (p->focal_loss) {
    float alpha = 0.5;
    int ti = index + stride * class_id;
    float pt = output[ti] + 0.000000000000001F;
    float grad = -(1 - pt) * (2 * pt * logf(pt) + pt - 1);
    for (int n = 0; n < p->classes; ++n) {
      delta[index + stride * n] = (((n == class_id) ? 1 : 0) - output[index + stride * n]);
      delta[index + stride * n] *= alpha * grad;
    }
  } else {
    for (int n = 0; n < p->classes; ++n) {
      float y_true = ((n == class_id) ? 1 : 0);
      if (p->label_smooth_eps) y_true = y_true * (1 - p->label_smooth_eps) + 0.5 * p->label_smooth_eps;
      float r = y_true - output[index + stride * n];
      if (!isnan(r) && !isinf(r)) delta[index + stride * n] = r;
    }
  }
}

/* ForwardYoloLayer, training half (yolo_layer.cpp:409-772).  out = decoded head [B][n*(5+classes)][H][W]; delta is
 * overwritten; iparams = {n, total, classes, max_boxes, net_w, net_h, iou_loss, iou_thresh_kind, focal_loss};
 * fparams = {ignore_thresh, truth_thresh, iou_thresh, iou_normalizer, cls_normalizer, label_smooth_eps, max_delta}.
 * Returns *(l->cost). */
float yo_yolo_loss(const float* out, const float* truth, const float* biases, const int* mask, int B, int H, int W,
                   const int* iparams, const float* fparams, float* delta) {
  oloss P;
  oloss* p = &P;
  P.B = B; P.H = H; P.W = W; P.n = iparams[0]; P.total = iparams[1]; P.classes = iparams[2]; P.max_boxes = iparams[3];
  P.net_w = iparams[4]; P.net_h = iparams[5]; P.iou_loss = iparams[6]; P.iou_thresh_kind = iparams[7]; P.focal_loss = iparams[8];
  P.ignore_thresh = fparams[0]; P.truth_thresh = fparams[1]; P.iou_thresh = fparams[2]; P.iou_normalizer = fparams[3];
  P.cls_normalizer = fparams[4]; P.label_smooth_eps = fparams[5]; P.max_delta = fparams[6];
  P.out = out; P.truth = truth; P.biases = biases; P.mask = mask; P.delta = delta;
  const int wh = W * H, E = 5 + P.classes;
  const long long total = (long long)B * P.n * E * wh;
  memset(delta, 0, sizeof(float) * total);
  float tot_iou_loss = 0, tot_giou_loss = 0;
  int count = 0;
  for (int b = 0; b < B; ++b) {
    const float* tr = truth + (long long)b * P.max_boxes * 5;
    for (int j = 0; j < H; ++j) for (int i = 0; i < W; ++i) for (int n = 0; n < P.n; ++n) {
      int loc = n * wh + j * W + i;
      int box_index = o_entry(p, b, loc, 0), obj_index = o_entry(p, b, loc, 4), class_index = o_entry(p, b, loc, 5);
      obox pred = o_get_box(out, biases, mask[n], box_index, i, j, W, H, P.net_w, P.net_h, wh);
      float best_match_iou = 0, best_iou = 0;
      int best_t = 0;
      for (int t = 0; t < P.max_boxes; ++t) {
        obox tb = {tr[t * 5], tr[t * 5 + 1], tr[t * 5 + 2], tr[t * 5 + 3]};
        int class_id = tr[t * 5 + 4];
        if (class_id >= P.classes || class_id < 0) continue;
        if (!tb.x) break;
        int match = 0;
        for (int c = 0; c < P.classes; ++c) if (out[class_index + wh * c] > 0.25f) { match = 1; break; }
        float iou = o_iou(pred, tb);
        if (iou > best_match_iou && match == 1) best_match_iou = iou;
        if (iou > best_iou) { best_iou = iou; best_t = t; }
      }
      delta[obj_index] = P.cls_normalizer * (0 - out[obj_index]);
      if (best_match_iou > P.ignore_thresh) delta[obj_index] = 0;
      if (best_iou > P.truth_thresh) {
        delta[obj_index] = P.cls_normalizer * (1 - out[obj_index]);
        int class_id = tr[best_t * 5 + 4];
        o_delta_class(p, class_index, class_id);
        obox tb = {tr[best_t * 5], tr[best_t * 5 + 1], tr[best_t * 5 + 2], tr[best_t * 5 + 3]};
        o_delta_box(p, tb, mask[n], box_index, i, j, (2 - tb.w * tb.h), P.iou_normalizer);
      }
    }
    for (int t = 0; t < P.max_boxes; ++t) {
      obox tb = {tr[t * 5], tr[t * 5 + 1], tr[t * 5 + 2], tr[t * 5 + 3]};
      int class_id = tr[t * 5 + 4];
      if (class_id >= P.classes || class_id < 0) continue;
      if (!tb.x) break;
      float best_iou = 0;
      int best_n = 0;
      int i = (tb.x * W), j = (tb.y * H);
      obox ts = tb; ts.x = ts.y = 0;
      for (int n = 0; n < P.total; ++n) {
        obox pr = {0, 0, biases[2 * n] / P.net_w, biases[2 * n + 1] / P.net_h};
        float iou = o_iou(pr, ts);
        if (iou > best_iou) { best_iou = iou; best_n = n; }
      }
      for (int pass = 0; pass <= P.total; ++pass) {
        int n_abs = pass == 0 ? best_n : pass - 1, mask_n = -1;
        if (pass > 0 && (n_abs == best_n || !(P.iou_thresh < 1.0f))) continue;
        for (int m = 0; m < P.n; ++m) if (mask[m] == n_abs) { mask_n = m; break; }
        if (mask_n < 0) continue;
        if (pass > 0) {
          obox pr = {0, 0, biases[2 * n_abs] / P.net_w, biases[2 * n_abs + 1] / P.net_h};
          if (!(o_iou_kind(pr, ts, P.iou_thresh_kind) > P.iou_thresh)) continue;
        }
        int loc = mask_n * wh + j * W + i;
        int box_index = o_entry(p, b, loc, 0), obj_index = o_entry(p, b, loc, 4), class_index = o_entry(p, b, loc, 5);
        oious all = o_delta_box(p, tb, n_abs, box_index, i, j, (2 - tb.w * tb.h), P.iou_normalizer);
        tot_iou_loss += 1 - all.iou;
        tot_giou_loss += 1 - all.giou;
        delta[obj_index] = P.cls_normalizer * (1 - out[obj_index]);
        o_delta_class(p, class_index, class_id);
        ++count;
      }
    }
    for (int j = 0; j < H; ++j) for (int i = 0; i < W; ++i) for (int n = 0; n < P.n; ++n) {   /* averages_yolo_deltas */
      int loc = n * wh + j * W + i;
      int box_index = o_entry(p, b, loc, 0), class_index = o_entry(p, b, loc, 5);
      int in_box = 0;
      for (int c = 0; c < P.classes; ++c) if (delta[class_index + wh * c] > 0) in_box++;
      if (in_box > 0) for (int e = 0; e < 4; ++e) delta[box_index + e * wh] /= in_box;
    }
  }
  if (count == 0) count = 1;
  double s_all = 0, s_nobox = 0;
  for (int b = 0; b < B; ++b) for (int n = 0; n < P.n; ++n) for (int e = 0; e < E; ++e) for (int l = 0; l < wh; ++l) {
    float d = delta[o_entry(p, b, n * wh + l, e)];
    s_all += (double)d * d;
    if (e >= 4) s_nobox += (double)d * d;
  }
  float classification_loss = P.cls_normalizer * (float)s_nobox;
  if (P.iou_loss == K_MSE) return (float)s_all;
  float tot = P.iou_loss == K_GIOU ? tot_giou_loss : tot_iou_loss;
  return P.iou_normalizer * (tot / count) + classification_loss;
}
