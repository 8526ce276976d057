/* TEST INFRASTRUCTURE — CPU restatement of the reference algorithm (see yc_oracle.h header).
 * Written from the reference's behaviour, loop structure chosen for clarity; not product code. */
#include "yc_oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ---- activations: activations.h:63-121 ------------------------------------------------------- */
static float softplus_t(float x, float t) {          /* activations.h:106-113 */
  if (x > t) return x;
  if (x < -t) return expf(x);
  return logf(expf(x) + 1);
}
static float tanh_ref(float x) { return 2 / (1 + expf(-2 * x)) - 1; } /* activations.h:100-103 */

float yo_activate1(float x, int act) {
  switch (act) {
    case YO_LOGISTIC: return 1.f / (1.f + expf(-x));             /* activations.h:80-83 */
    case YO_RELU:     return x * (x > 0);
    case YO_LINEAR:   return x;
    case YO_LEAKY:    return (x > 0) ? x : .1f * x;              /* activations.h:98 */
    case YO_SWISH:    return x * (1.f / (1.f + expf(-x)));       /* activations.c:170-182 */
    case YO_MISH:     return x * tanh_ref(softplus_t(x, 20.f));  /* activations.c:185-197 */
  }
  return x;
}
void yo_activate(float* x, long long n, int act) {
  for (long long i = 0; i < n; ++i) x[i] = yo_activate1(x[i], act);
}

/* gradients.  leaky/logistic/linear/relu take the POST-activation output (activations.h:177-192,
 * convolutional_layer.cpp:1327); mish/swish take the saved PRE-activation (activations.c:401-452). */
void yo_gradient(const float* v, long long n, int act, float* delta) {
  for (long long i = 0; i < n; ++i) {
    float x = v[i], g = 1.f;
    switch (act) {
      case YO_LOGISTIC: g = (1 - x) * x; break;
      case YO_RELU:     g = (x > 0); break;
      case YO_LINEAR:   g = 1; break;
      case YO_LEAKY:    g = (x > 0) ? 1.f : .1f; break;
      case YO_SWISH: {  /* activations.c:401-410: swish + sigmoid*(1-swish), sigmoid recomputed */
        float s = 1.f / (1.f + expf(-x)); float sw = x * s; g = sw + s * (1 - sw); break; }
      case YO_MISH: {   /* activations.c:426-452 */
        float sp = softplus_t(x, 20.f);
        float grad_sp = 1 - exp(-sp);
        float tsp = tanh(sp);
        float grad_tsp = (1 - tsp * tsp) * grad_sp;
        g = x * grad_tsp + tsp; break; }
    }
    delta[i] *= g;
  }
}

/* ---- im2col / col2im / gemm ------------------------------------------------------------------ */
void yo_im2col(const float* im, int c, int h, int w, int k, int pad, int stride, int dil, float* col) {
  int oh = (h + 2 * pad - (dil * (k - 1) + 1)) / stride + 1;
  int ow = (w + 2 * pad - (dil * (k - 1) + 1)) / stride + 1;
  for (int ch = 0; ch < c; ++ch)
    for (int kr = 0; kr < k; ++kr)
      for (int kc = 0; kc < k; ++kc)
        for (int y = 0; y < oh; ++y) {
          int iy = -pad + kr * dil + y * stride;
          for (int x = 0; x < ow; ++x) {
            int ix = -pad + kc * dil + x * stride;
            float v = 0;
            if (iy >= 0 && iy < h && ix >= 0 && ix < w) v = im[((long long)ch * h + iy) * w + ix];
            *col++ = v;
          }
        }
}
void yo_col2im(const float* col, int c, int h, int w, int k, int pad, int stride, int dil, float* im) {
  int oh = (h + 2 * pad - (dil * (k - 1) + 1)) / stride + 1;
  int ow = (w + 2 * pad - (dil * (k - 1) + 1)) / stride + 1;
  memset(im, 0, sizeof(float) * (size_t)c * h * w);           /* col2im.c:70 */
  for (int ch = 0; ch < c; ++ch)
    for (int kr = 0; kr < k; ++kr)
      for (int kc = 0; kc < k; ++kc)
        for (int y = 0; y < oh; ++y) {
          int iy = -pad + kr * dil + y * stride;
          for (int x = 0; x < ow; ++x) {
            int ix = -pad + kc * dil + x * stride;
            float v = *col++;
            if (iy >= 0 && iy < h && ix >= 0 && ix < w) im[((long long)ch * h + iy) * w + ix] += v;
          }
        }
}
void yo_gemm(int TA, int TB, int M, int N, int K, float alpha, const float* A, int lda, const float* B,
             int ldb, float beta, float* C, int ldc) {
  for (int i = 0; i < M; ++i)
    for (int j = 0; j < N; ++j) C[(size_t)i * ldc + j] *= beta;            /* gemm.c:2972-2982 */
  if (!TA && !TB) {                                                         /* gemm_nn gemm.c:798-859 */
    for (int i = 0; i < M; ++i)
      for (int k = 0; k < K; ++k) {
        float a = alpha * A[(size_t)i * lda + k];
        for (int j = 0; j < N; ++j) C[(size_t)i * ldc + j] += a * B[(size_t)k * ldb + j];
      }
  } else if (!TA && TB) {                                                   /* gemm_nt gemm.c:2914-2930 */
    for (int i = 0; i < M; ++i)
      for (int j = 0; j < N; ++j) {
        float s = 0;
        for (int k = 0; k < K; ++k) s += alpha * A[(size_t)i * lda + k] * B[(size_t)j * ldb + k];
        C[(size_t)i * ldc + j] += s;
      }
  } else if (TA && !TB) {                                                   /* gemm_tn gemm.c:2932-2947 */
    for (int i = 0; i < M; ++i)
      for (int k = 0; k < K; ++k) {
        float a = alpha * A[(size_t)k * lda + i];
        for (int j = 0; j < N; ++j) C[(size_t)i * ldc + j] += a * B[(size_t)k * ldb + j];
      }
  } else {                                                                  /* gemm_tt gemm.c:2949-2965 */
    for (int i = 0; i < M; ++i)
      for (int j = 0; j < N; ++j) {
        float s = 0;
        for (int k = 0; k < K; ++k) s += alpha * A[(size_t)i + (size_t)k * lda] * B[(size_t)k + (size_t)j * ldb];
        C[(size_t)i * ldc + j] += s;
      }
  }
}

/* ---- conv forward ------------------------------------------------------------------------------ */
void yo_fuse_bn(float* w, float* b, const float* scales, const float* mean, const float* var, int n,
                int per_filter) {
  for (int f = 0; f < n; ++f) {
    float sd = sqrt(var[f] + 0.00001f);                        /* network.cpp:662 */
    b[f] -= scales[f] * mean[f] / sd;
    for (int i = 0; i < per_filter; ++i) w[(size_t)f * per_filter + i] *= scales[f] / sd;
  }
}

void yo_conv_forward(const float* x, const float* w, const float* bias, const float* scales,
                     const float* rmean, const float* rvar, int B, int C, int H, int W, int N, int k,
                     int stride, int pad, int dil, int act, float* out, float* pre) {
  int oh = (H + 2 * pad - (dil * (k - 1) + 1)) / stride + 1;
  int ow = (W + 2 * pad - (dil * (k - 1) + 1)) / stride + 1;
  int S = oh * ow, K = k * k * C;
  float* col = (float*)malloc(sizeof(float) * (size_t)K * S);
  memset(out, 0, sizeof(float) * (size_t)B * N * S);            /* convolutional_layer.cpp:1134 */
  for (int b = 0; b < B; ++b) {
    const float* im = x + (size_t)b * C * H * W;
    const float* bmat = im;
    if (!(k == 1 && stride == 1 && pad == 0)) {                 /* :1243-1257 (size==1 shortcut) */
      yo_im2col(im, C, H, W, k, pad, stride, dil, col);
      bmat = col;
    }
    yo_gemm(0, 0, N, S, K, 1.f, w, K, bmat, S, 1.f, out + (size_t)b * N * S, S);   /* :1259 */
  }
  for (int b = 0; b < B; ++b)
    for (int f = 0; f < N; ++f) {
      float* o = out + ((size_t)b * N + f) * S;
      if (scales) {                                             /* batchnorm_layer.cpp:231-237, blas.c:204-218 */
        float m = rmean[f], sd = sqrt(rvar[f] + .000001f);
        for (int i = 0; i < S; ++i) o[i] = (o[i] - m) / sd;
        for (int i = 0; i < S; ++i) o[i] *= scales[f];
      }
      for (int i = 0; i < S; ++i) o[i] += bias[f];              /* add_bias convolutional_layer.cpp:916-929 */
    }
  if (pre) memcpy(pre, out, sizeof(float) * (size_t)B * N * S);
  yo_activate(out, (long long)B * N * S, act);
  free(col);
}

/* ---- memory-bound layers ---------------------------------------------------------------------- */
void yo_maxpool(const float* x, int B, int C, int H, int W, int size, int stride, int pad, float* out,
                int* idx, int infer_avx) {
  int oh = (H + pad - size) / stride + 1;                       /* maxpool_layer.cpp:62-63 */
  int ow = (W + pad - size) / stride + 1;
  int off = -pad / 2;                                           /* :258-259 */
  for (int b = 0; b < B; ++b)
    for (int c = 0; c < C; ++c)
      for (int i = 0; i < oh; ++i) {
        int j = 0;
        /* Inference path on AVX hosts: forward_maxpool_layer_avx (gemm.c:2108-2215), taken when
         * !train (maxpool_layer.cpp:249-253).  Its stride-1 vector loop (gemm.c:2127-2150) tests tap
         * validity for lane 0 only and SKIPS the whole 8-wide tap otherwise, so in the first block
         * (j=0) outputs 1..7 never see the taps left of the window centre.  Restated literally. */
        if (infer_avx && stride == 1) {
          for (j = 0; j < ow - 8 - (size - 1); j += 8) {
            float mx[8];
            for (int l = 0; l < 8; ++l) mx[l] = -FLT_MAX;
            for (int n = 0; n < size; ++n)
              for (int m = 0; m < size; ++m) {
                int ch = off + i * stride + n, cw = off + j * stride + m;
                if (!(ch >= 0 && ch < H && cw >= 0 && cw < W)) continue;
                const float* s = x + cw + W * (ch + H * (c + b * C));
                for (int l = 0; l < 8; ++l) mx[l] = s[l] > mx[l] ? s[l] : mx[l];
              }
            for (int l = 0; l < 8; ++l) out[j + l + ow * (i + oh * (c + C * b))] = mx[l];
          }
        }
        for (; j < ow; ++j) {
          float mx = -FLT_MAX; int mi = -1;
          for (int n = 0; n < size; ++n)
            for (int m = 0; m < size; ++m) {
              int ch = off + i * stride + n, cw = off + j * stride + m;
              int index = cw + W * (ch + H * (c + b * C));
              int valid = (ch >= 0 && ch < H && cw >= 0 && cw < W);
              float v = valid ? x[index] : -FLT_MAX;
              if (v > mx) { mx = v; mi = index; }               /* strict > keeps first max :286-287 */
            }
          int o = j + ow * (i + oh * (c + C * b));
          out[o] = mx;
          if (idx) idx[o] = mi;
        }
      }
}

void yo_upsample(const float* x, int B, int C, int H, int W, int stride, float scale, float* out) {
  for (int b = 0; b < B; ++b)
    for (int k = 0; k < C; ++k)
      for (int j = 0; j < H * stride; ++j)
        for (int i = 0; i < W * stride; ++i) {
          size_t in = ((size_t)b * C + k) * H * W + (size_t)(j / stride) * W + i / stride;
          size_t o = ((size_t)b * C + k) * H * W * stride * stride + (size_t)j * W * stride + i;
          out[o] = scale * x[in];                               /* blas.c:399 */
        }
}

void yo_shortcut(const float* in, const float* add, long long n, int act, float* out) {
  for (long long i = 0; i < n; ++i) out[i] = in[i] + add[i];    /* shortcut_layer.c:156-159 */
  yo_activate(out, n, act);
}

void yo_reorg(const float* x, int ow, int oh, int oc, int batch, int stride, int forward, float* out) {
  int ic = oc / (stride * stride);                              /* blas.c:12-46 */
  for (int b = 0; b < batch; ++b)
    for (int k = 0; k < oc; ++k)
      for (int j = 0; j < oh; ++j)
        for (int i = 0; i < ow; ++i) {
          int in_index = i + ow * (j + oh * (k + oc * b));
          int c2 = k % ic, offset = k / ic;
          int w2 = i * stride + offset % stride, h2 = j * stride + offset / stride;
          int out_index = w2 + ow * stride * (h2 + oh * stride * (c2 + ic * b));
          if (forward) out[out_index] = x[in_index]; else out[in_index] = x[out_index];
        }
}

void yo_scale_channels(const float* in, const float* scale, int B, int C, int HW, float* out) {
  long long n = (long long)B * C * HW;                          /* scale_channels_layer.c:91-97 */
  for (long long i = 0; i < n; ++i) out[i] = scale[i / HW] * in[i];
}

/* ForwardScaleChannelsLayer incl. the activation (scale_channels_layer.c:72-97, scale_wh = 0):
 * out = act(scale_in[b,c] * from_out[b,c,h,w]) */
void yo_scale_channels_forward(const float* from_out, const float* scale_in, int B, int C, int HW, int act, float* out) {
  yo_scale_channels(from_out, scale_in, B, C, HW, out);
  yo_activate(out, (long long)B * C * HW, act);
}

/* BackwardScaleChannelsLayer (scale_channels_layer.c:99-131, scale_wh = 0): gradient_array on the OUTPUT modifies
 * delta in place; d_scale_in[b,c] += sum_hw delta * from_out; d_from += scale_in[b,c] * delta. */
void yo_scale_channels_backward(float* delta, const float* out, int act, const float* from_out, const float* scale_in,
                                int B, int C, int HW, float* d_scale_in, float* d_from) {
  long long n = (long long)B * C * HW;
  yo_gradient(out, n, act, delta);
  for (long long i = 0; i < n; ++i) {
    d_scale_in[i / HW] += delta[i] * from_out[i];
    d_from[i] += scale_in[i / HW] * delta[i];
  }
}

/* ---- YOLO head -------------------------------------------------------------------------------- */
static float logistic1(float x) { return 1.f / (1.f + expf(-x)); }

void yo_yolo_decode(const float* in, int B, int n, int classes, int H, int W, float sxy, float* out) {
  int S = H * W, E = 5 + classes;
  for (int b = 0; b < B; ++b)
    for (int a = 0; a < n; ++a)
      for (int e = 0; e < E; ++e) {
        const float* s = in + (((size_t)b * n + a) * E + e) * S;
        float* d = out + (((size_t)b * n + a) * E + e) * S;
        for (int i = 0; i < S; ++i) {
          float v = s[i];
          if (e < 2) { v = logistic1(v); v = v * sxy + (float)(-0.5 * (sxy - 1)); }  /* :399-401, blas.c scal_add */
          else if (e >= 4) v = logistic1(v);                                          /* :402-404 */
          d[i] = v;
        }
      }
}

void yo_gaussian_yolo_decode(const float* in, int B, int n, int classes, int H, int W, float sxy,
                             float* out) {
  int S = H * W, E = 9 + classes;
  for (int b = 0; b < B; ++b)
    for (int a = 0; a < n; ++a)
      for (int e = 0; e < E; ++e) {
        const float* s = in + (((size_t)b * n + a) * E + e) * S;
        float* d = out + (((size_t)b * n + a) * E + e) * S;
        int sig = (e <= 3) || e == 5 || e == 7 || e >= 8;       /* gaussian_yolo_layer.cpp:497-516 */
        int aff = (e == 0) || (e == 2);
        for (int i = 0; i < S; ++i) {
          float v = s[i];
          if (sig) v = logistic1(v);
          if (aff) v = v * sxy + (float)(-0.5 * (sxy - 1));
          d[i] = v;
        }
      }
}

int yo_yolo_detections(const float* pred, int b, int n, int classes, int H, int W, const float* anchors,
                       const int* mask, int net_w, int net_h, float thresh, float* out, int max_out) {
  int S = H * W, E = 5 + classes, cnt = 0;
  const float* p = pred + (size_t)b * n * E * S;
  for (int a = 0; a < n; ++a)
    for (int i = 0; i < S; ++i) {
      const float* q = p + (size_t)a * E * S + i;               /* EntryIndex yolo_layer.cpp:380-386 */
      float obj = q[4 * S];
      if (obj <= thresh) continue;                              /* :807-809 */
      if (cnt >= max_out) return cnt;
      int col = i % W, row = i / W;
      float* o = out + (size_t)cnt * E;
      o[0] = (col + q[0]) / W;                                  /* GetYoloBox :139-148 */
      o[1] = (row + q[S]) / H;
      o[2] = exp(q[2 * S]) * anchors[2 * mask[a]] / net_w;
      o[3] = exp(q[3 * S]) * anchors[2 * mask[a] + 1] / net_h;
      o[4] = obj;
      for (int j = 0; j < classes; ++j) {
        float pr = obj * q[(size_t)(5 + j) * S];
        o[5 + j] = (pr > thresh) ? pr : 0;                      /* :820-826 */
      }
      ++cnt;
    }
  return cnt;
}

/* box.cpp:36-72 */
static float overlap1(float x1, float w1, float x2, float w2) {
  float l1 = x1 - w1 / 2, l2 = x2 - w2 / 2;
  float left = l1 > l2 ? l1 : l2;
  float r1 = x1 + w1 / 2, r2 = x2 + w2 / 2;
  float right = r1 < r2 ? r1 : r2;
  return right - left;
}
float yo_box_iou(const float* a, const float* b) {
  float w = overlap1(a[0], a[2], b[0], b[2]);
  float h = overlap1(a[1], a[3], b[1], b[3]);
  float I = (w < 0 || h < 0) ? 0 : w * h;
  float U = a[2] * a[3] + b[2] * b[3] - I;
  if (fabsf(I) < FLT_EPSILON || fabsf(U) < FLT_EPSILON) return 0;
  return I / U;
}
static float box_diou(const float* a, const float* b, float beta) {     /* box.cpp:97-111 */
  float al = a[0] - a[2] / 2, ar = a[0] + a[2] / 2, at = a[1] - a[3] / 2, ab = a[1] + a[3] / 2;
  float bl = b[0] - b[2] / 2, br = b[0] + b[2] / 2, bt = b[1] - b[3] / 2, bb = b[1] + b[3] / 2;
  float w = (ar > br ? ar : br) - (al < bl ? al : bl);
  float h = (ab > bb ? ab : bb) - (at < bt ? at : bt);
  float c = w * w + h * h;
  float iou = yo_box_iou(a, b);
  if (fabsf(c) < FLT_EPSILON) return iou;
  float d = (a[0] - b[0]) * (a[0] - b[0]) + (a[1] - b[1]) * (a[1] - b[1]);
  return iou - (float)pow(d / c, beta);
}

void yo_nms_sort(float* dets, int total, int classes, float thresh, int nms_kind, float beta) {
  int E = 5 + classes;
  float* tmp = (float*)malloc(sizeof(float) * E);
  for (int k = 0; k < classes; ++k) {
    /* stable insertion sort by prob[k] descending (reference: qsort, box.cpp:402) */
    for (int i = 1; i < total; ++i) {
      memcpy(tmp, dets + (size_t)i * E, sizeof(float) * E);
      int j = i - 1;
      while (j >= 0 && dets[(size_t)j * E + 5 + k] < tmp[5 + k]) {
        memcpy(dets + (size_t)(j + 1) * E, dets + (size_t)j * E, sizeof(float) * E);
        --j;
      }
      memcpy(dets + (size_t)(j + 1) * E, tmp, sizeof(float) * E);
    }
    for (int i = 0; i < total; ++i) {
      if (fabsf(dets[(size_t)i * E + 5 + k]) < FLT_EPSILON) continue;    /* box.cpp:405 */
      const float* a = dets + (size_t)i * E;
      for (int j = i + 1; j < total; ++j) {
        float* b = dets + (size_t)j * E;
        if (nms_kind == 0 && yo_box_iou(a, b) > thresh) b[5 + k] = 0.f;
        else if (nms_kind == 1 && box_diou(a, b, beta) > thresh) b[5 + k] = 0.f;
      }
    }
  }
  free(tmp);
}

/* ---- training: conv forward with batch statistics, conv backward -------------------------------- */
/* ForwardConvolutionalLayer with state.train (convolutional_layer.cpp:1128-1305) + ForwardBatchnormLayer
 * (batchnorm_layer.cpp:206-238): z = conv(x,w); if bn: mean/var (blas.c:164-195, unbiased), rolling .9/.1,
 * xhat=(z-mean)/sqrt(var+1e-6) (blas.c:214), y = act(gamma*xhat+beta); else y = act(z + bias).
 * z_out receives the pre-batchnorm conv output (l->x), xnorm_out the normalised values (l->x_norm), pre_out
 * the pre-activation (l->activation_input); any may be NULL. */
void yo_conv_forward_train(const float* x, const float* w, const float* biases, const float* scales,
                           float* rolling_mean, float* rolling_var, float* mean, float* var, int B, int C, int H,
                           int W, int N, int k, int stride, int pad, int dil, int act, int bn, float* out,
                           float* z_out, float* xnorm_out, float* pre_out) {
  int oh = (H + 2 * pad - (dil * (k - 1) + 1)) / stride + 1;
  int ow = (W + 2 * pad - (dil * (k - 1) + 1)) / stride + 1;
  int S = oh * ow, K = k * k * C;
  float* col = (float*)malloc(sizeof(float) * (size_t)K * S);
  memset(out, 0, sizeof(float) * (size_t)B * N * S);
  for (int b = 0; b < B; ++b) {
    yo_im2col(x + (size_t)b * C * H * W, C, H, W, k, pad * dil, stride, dil, col);
    yo_gemm(0, 0, N, S, K, 1.f, w, K, col, S, 1.f, out + (size_t)b * N * S, S);
  }
  free(col);
  if (bn) {
    for (int f = 0; f < N; ++f) {                                  /* mean_cpu blas.c:164-177 */
      float m = 0;
      for (int b = 0; b < B; ++b) for (int i = 0; i < S; ++i) m += out[((size_t)b * N + f) * S + i];
      mean[f] = m * (1.f / (B * S));
    }
    for (int f = 0; f < N; ++f) {                                  /* variance_cpu blas.c:179-195 */
      float v = 0;
      for (int b = 0; b < B; ++b) for (int i = 0; i < S; ++i) {
        float d = out[((size_t)b * N + f) * S + i] - mean[f];
        v += d * d;
      }
      var[f] = v * (1.f / (B * S - 1));
    }
    for (int f = 0; f < N; ++f) {                                  /* batchnorm_layer.cpp:221-224 */
      if (rolling_mean) rolling_mean[f] = .9f * rolling_mean[f] + .1f * mean[f];
      if (rolling_var) rolling_var[f] = .9f * rolling_var[f] + .1f * var[f];
    }
    if (z_out) memcpy(z_out, out, sizeof(float) * (size_t)B * N * S);
    for (int b = 0; b < B; ++b) for (int f = 0; f < N; ++f) for (int i = 0; i < S; ++i) {
      size_t idx = ((size_t)b * N + f) * S + i;
      out[idx] = (out[idx] - mean[f]) / (sqrt(var[f] + .000001f));  /* normalize_cpu blas.c:214 */
    }
    if (xnorm_out) memcpy(xnorm_out, out, sizeof(float) * (size_t)B * N * S);
    for (int b = 0; b < B; ++b) for (int f = 0; f < N; ++f) for (int i = 0; i < S; ++i) {
      size_t idx = ((size_t)b * N + f) * S + i;
      out[idx] = out[idx] * scales[f] + biases[f];
    }
  } else {
    for (int b = 0; b < B; ++b) for (int f = 0; f < N; ++f) for (int i = 0; i < S; ++i)
      out[((size_t)b * N + f) * S + i] += biases[f];
    if (z_out) memcpy(z_out, out, sizeof(float) * (size_t)B * N * S);
  }
  if (pre_out) memcpy(pre_out, out, sizeof(float) * (size_t)B * N * S);
  yo_activate(out, (long long)B * N * S, act);
}

/* BackwardConvolutionalLayer (convolutional_layer.cpp:1307-1380) + BackwardBatchnormLayer
 * (batchnorm_layer.cpp:92-166, 240-255).  delta [B,N,oh,ow] is modified in place exactly like l->delta
 * (activation gradient, then the batchnorm backward), weight_updates / bias_updates / scale_updates are
 * ACCUMULATED into, dx (may be NULL) is OVERWRITTEN (col2im.c:70).  out = l->output, pre = saved
 * pre-activation (mish/swish), z = l->x, xnorm = l->x_norm. */
void yo_conv_backward(const float* x, const float* w, const float* scales, const float* mean, const float* var,
                      const float* out, const float* pre, const float* z, const float* xnorm, float* delta, int B,
                      int C, int H, int W, int N, int k, int stride, int pad, int dil, int act, int bn,
                      float* weight_updates, float* bias_updates, float* scale_updates, float* dx) {
  int oh = (H + 2 * pad - (dil * (k - 1) + 1)) / stride + 1;
  int ow = (W + 2 * pad - (dil * (k - 1) + 1)) / stride + 1;
  int S = oh * ow, K = k * k * C;
  long long tot = (long long)B * N * S;
  if (act == YO_MISH || act == YO_SWISH) yo_gradient(pre, tot, act, delta);
  else yo_gradient(out, tot, act, delta);
  if (bn) {
    for (int f = 0; f < N; ++f) {                                  /* backward_scale_cpu :92-110 */
      float sum = 0;
      for (int b = 0; b < B; ++b) for (int i = 0; i < S; ++i) {
        size_t idx = ((size_t)b * N + f) * S + i;
        sum += delta[idx] * xnorm[idx];
      }
      scale_updates[f] += sum;
    }
    for (int b = 0; b < B; ++b) for (int f = 0; f < N; ++f) for (int i = 0; i < S; ++i)   /* scale_bias */
      delta[((size_t)b * N + f) * S + i] *= scales[f];
    float* md = (float*)calloc(N, sizeof(float));
    float* vd = (float*)calloc(N, sizeof(float));
    for (int f = 0; f < N; ++f) {                                  /* mean_delta_cpu :112-129 */
      for (int b = 0; b < B; ++b) for (int i = 0; i < S; ++i) md[f] += delta[((size_t)b * N + f) * S + i];
      md[f] *= (-1. / sqrt(var[f] + .00001f));
    }
    for (int f = 0; f < N; ++f) {                                  /* variance_delta_cpu :130-147 */
      for (int b = 0; b < B; ++b) for (int i = 0; i < S; ++i) {
        size_t idx = ((size_t)b * N + f) * S + i;
        vd[f] += delta[idx] * (z[idx] - mean[f]);
      }
      vd[f] *= -.5 * pow(var[f] + .00001f, (float)(-3. / 2.));
    }
    for (int b = 0; b < B; ++b) for (int f = 0; f < N; ++f) for (int i = 0; i < S; ++i) {  /* normalize_delta_cpu */
      size_t idx = ((size_t)b * N + f) * S + i;
      delta[idx] = delta[idx] * 1. / (sqrt(var[f]) + .00001f) + vd[f] * 2. * (z[idx] - mean[f]) / (S * B) +
                   md[f] / (S * B);
    }
    free(md); free(vd);
  } else {
    for (int b = 0; b < B; ++b) for (int f = 0; f < N; ++f) {      /* backward_bias */
      float s = 0;
      for (int i = 0; i < S; ++i) s += delta[((size_t)b * N + f) * S + i];
      bias_updates[f] += s;
    }
  }
  float* col = (float*)malloc(sizeof(float) * (size_t)K * S);
  for (int b = 0; b < B; ++b) {
    const float* a = delta + (size_t)b * N * S;
    yo_im2col(x + (size_t)b * C * H * W, C, H, W, k, pad * dil, stride, dil, col);
    yo_gemm(0, 1, N, K, S, 1.f, a, S, col, S, 1.f, weight_updates, K);           /* :1358 */
    if (dx) {
      yo_gemm(1, 0, K, S, N, 1.f, w, K, a, S, 0.f, col, S);                       /* :1366 */
      yo_col2im(col, C, H, W, k, pad * dil, stride, dil, dx + (size_t)b * C * H * W);
    }
  }
  free(col);
}
