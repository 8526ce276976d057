// Backward passes of the memory-bound layers + the SGD update, NHWC bf16 deltas, fp32 math, 16-byte accesses.
// Replaces BackwardMaxpoolLayerGpu (maxpool_layer_kernels.cu:103-143), BackwardUpsampleLayerGpu
// (upsample_layer.c:121-133), BackwardRouteLayerGpu (route_layer.c:144-160), BackwardShortcutLayerGpu
// (shortcut_layer.c:206-220, blas_kernels.cu:980-1018) and UpdateConvolutionalLayerGpu
// (convolutional_kernels.cu:865-921); semantics follow the CPU twins cited per kernel.
#include <float.h>
#include <algorithm>

#include "yc_common.cuh"

namespace yc {

struct BV {  // bf16 NHWC view
  __nv_bfloat16* ptr;
  int n, h, w, c;
  long long ldc;
};
static inline BV bv(const yc_tensor* t) { return BV{(__nv_bfloat16*)t->ptr, t->n, t->h, t->w, t->c, t->ldc}; }

__device__ __forceinline__ void bld8(const BV& t, long long pix, int c0, float (&v)[8]) {
  const __nv_bfloat16* p = t.ptr + pix * t.ldc + c0;
  if (c0 + 8 <= t.c) {
    const uint4 q = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 f = __bfloat1622float2(h[e]);
      v[2 * e] = f.x;
      v[2 * e + 1] = f.y;
    }
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = (c0 + j < t.c) ? __bfloat162float(p[j]) : 0.f;
  }
}
__device__ __forceinline__ void bst8(const BV& t, long long pix, int c0, const float (&v)[8]) {
  __nv_bfloat16* p = t.ptr + pix * t.ldc + c0;
  if (c0 + 8 <= t.c) {
    uint4 q;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&q);
#pragma unroll
    for (int e = 0; e < 4; ++e) h[e] = __floats2bfloat162_rn(v[2 * e], v[2 * e + 1]);
    *reinterpret_cast<uint4*>(p) = q;
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (c0 + j < t.c) p[j] = __float2bfloat16_rn(v[j]);
  }
}

// ---- maxpool, training forward: also records which tap of the window won (ForwardMaxpoolLayer,
// maxpool_layer.cpp:258-288: strict '>' keeps the first maximum in row-major scan order; the reference stores the
// flat input index in l->indexes, we store the tap number n*size+m, one byte per output element) ----------------
__global__ void maxpool_train_kernel(BV x, BV y, uint8_t* __restrict__ idx, int size, int stride, int pad) {
  const int groups = (x.c + 7) / 8;
  const long long total = (long long)y.n * y.h * y.w * groups;
  const int off = -pad / 2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i % groups);
    const long long pix = i / groups;
    const int ow = (int)(pix % y.w);
    const int oh = (int)((pix / y.w) % y.h);
    const long long b = pix / ((long long)y.w * y.h);
    float m[8];
    int mi[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { m[j] = -FLT_MAX; mi[j] = 255; }
    for (int n = 0; n < size; ++n) {
      const int ih = off + oh * stride + n;
      if (ih < 0 || ih >= x.h) continue;
      for (int k = 0; k < size; ++k) {
        const int iw = off + ow * stride + k;
        if (iw < 0 || iw >= x.w) continue;
        float v[8];
        bld8(x, (b * x.h + ih) * x.w + iw, g * 8, v);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (v[j] > m[j]) { m[j] = v[j]; mi[j] = n * size + k; }
      }
    }
    bst8(y, pix, g * 8, m);
    uint8_t* ip = idx + pix * (long long)(groups * 8) + g * 8;
    uint2 q;
    q.x = (uint32_t)mi[0] | ((uint32_t)mi[1] << 8) | ((uint32_t)mi[2] << 16) | ((uint32_t)mi[3] << 24);
    q.y = (uint32_t)mi[4] | ((uint32_t)mi[5] << 8) | ((uint32_t)mi[6] << 16) | ((uint32_t)mi[7] << 24);
    *reinterpret_cast<uint2*>(ip) = q;
  }
}

// BackwardMaxpoolLayer (maxpool_layer.cpp:311-323): state.delta[indexes[i]] += l->delta[i], as a gather over the
// windows that contain each input pixel (no atomics, deterministic)
__global__ void maxpool_bwd_kernel(BV dy, const uint8_t* __restrict__ idx, BV dx, int size, int stride, int pad) {
  const int groups = (dx.c + 7) / 8;
  const long long total = (long long)dx.n * dx.h * dx.w * groups;
  const int off = -pad / 2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i % groups);
    const long long pix = i / groups;
    const int iw = (int)(pix % dx.w);
    const int ih = (int)((pix / dx.w) % dx.h);
    const long long b = pix / ((long long)dx.w * dx.h);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    bool any = false;
    // windows (oh, ow) with ih = off + oh*stride + n, 0 <= n < size
    for (int n = 0; n < size; ++n) {
      const int th = ih - off - n;
      if (th < 0 || th % stride) continue;
      const int oh = th / stride;
      if (oh >= dy.h) continue;
      for (int k = 0; k < size; ++k) {
        const int tw = iw - off - k;
        if (tw < 0 || tw % stride) continue;
        const int ow = tw / stride;
        if (ow >= dy.w) continue;
        const long long opix = (b * dy.h + oh) * dy.w + ow;
        const uint2 q = *reinterpret_cast<const uint2*>(idx + opix * (long long)(groups * 8) + g * 8);
        const int tap = n * size + k;
        float d[8];
        bld8(dy, opix, g * 8, d);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint32_t w32 = j < 4 ? q.x : q.y;
          const int t = (w32 >> (8 * (j & 3))) & 255;
          if (t == tap) { acc[j] += d[j]; any = true; }
        }
      }
    }
    if (any) {
      float v[8];
      bld8(dx, pix, g * 8, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] += acc[j];
      bst8(dx, pix, g * 8, v);
    }
  }
}

// The same for the 2x2 / stride-2 windows of the tiny nets (size == stride, windows start at 0): every input pixel
// lies in exactly one window, so one thread owns a window and its four input pixels; all loads are issued before
// the first use (the general kernel chains idx -> dy -> dx).
__global__ void __launch_bounds__(256) maxpool_bwd_2x2_kernel(BV dy, const uint8_t* __restrict__ idx, BV dx) {
  const int groups = dx.c >> 3;  // launcher guarantees c % 8 == 0
  const long long total = (long long)dy.n * dy.h * dy.w * groups;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i % groups);
    const long long opix = i / groups;
    const int ow = (int)(opix % dy.w);
    const int oh = (int)((opix / dy.w) % dy.h);
    const long long b = opix / ((long long)dy.w * dy.h);
    const uint2 q = *reinterpret_cast<const uint2*>(idx + opix * (long long)dx.c + g * 8);
    const uint4 dq = *reinterpret_cast<const uint4*>(dy.ptr + opix * dy.ldc + g * 8);
    uint4 xq[4];
    __nv_bfloat16* xp[4];
    bool ok[4];
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int ih = oh * 2 + (t >> 1), iw = ow * 2 + (t & 1);
      ok[t] = ih < dx.h && iw < dx.w;
      xp[t] = dx.ptr + ((b * dx.h + ih) * dx.w + iw) * dx.ldc + g * 8;
      xq[t] = ok[t] ? *reinterpret_cast<const uint4*>(xp[t]) : make_uint4(0, 0, 0, 0);
    }
    const __nv_bfloat162* dh = reinterpret_cast<const __nv_bfloat162*>(&dq);
    float d[8];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 f = __bfloat1622float2(dh[e]);
      d[2 * e] = f.x;
      d[2 * e + 1] = f.y;
    }
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      if (!ok[t]) continue;
      __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&xq[t]);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float2 f = __bfloat1622float2(h[e]);
        const uint32_t w32 = e < 2 ? q.x : q.y;
        const int t0 = (w32 >> (16 * (e & 1))) & 255, t1 = (w32 >> (16 * (e & 1) + 8)) & 255;
        if (t0 == t) f.x += d[2 * e];
        if (t1 == t) f.y += d[2 * e + 1];
        h[e] = __floats2bfloat162_rn(f.x, f.y);
      }
      *reinterpret_cast<uint4*>(xp[t]) = xq[t];
    }
  }
}

// BackwardUpsampleLayer (upsample_layer.c:92-103 -> upsample_cpu forward=0, blas.c:382-406):
// state.delta[in] += scale * sum of the stride^2 children of l->delta
__global__ void upsample_bwd_kernel(BV dy, BV dx, int stride, float scale) {
  const int groups = (dx.c + 7) / 8;
  const long long total = (long long)dx.n * dx.h * dx.w * groups;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i % groups);
    const long long pix = i / groups;
    const int iw = (int)(pix % dx.w);
    const int ih = (int)((pix / dx.w) % dx.h);
    const long long b = pix / ((long long)dx.w * dx.h);
    float v[8];
    bld8(dx, pix, g * 8, v);
    for (int r = 0; r < stride; ++r)
      for (int s = 0; s < stride; ++s) {
        float d[8];
        bld8(dy, (b * dy.h + ih * stride + r) * dy.w + iw * stride + s, g * 8, d);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] += scale * d[j];
      }
    bst8(dx, pix, g * 8, v);
  }
}

// dst += src on views: BackwardRouteLayer's axpy_cpu per input (route_layer.c:106-122) and BackwardShortcutCpu's two
// accumulations (blas.c:101-129)
__global__ void axpy_view_kernel(BV src, BV dst) {
  const int groups = (dst.c + 7) / 8;
  const long long total = (long long)dst.n * dst.h * dst.w * groups;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i % groups);
    const long long pix = i / groups;
    float a[8], d[8];
    bld8(src, pix, g * 8, a);
    bld8(dst, pix, g * 8, d);
#pragma unroll
    for (int j = 0; j < 8; ++j) d[j] += a[j];
    bst8(dst, pix, g * 8, d);
  }
}

// BackwardShortcutLayer (shortcut_layer.c:176-188): dy *= act'(y) in place, then prev.delta += dy, from.delta += dy
__global__ void shortcut_bwd_kernel(BV dy, BV y, int act, BV dprev, BV dfrom) {
  const int groups = (dy.c + 7) / 8;
  const long long total = (long long)dy.n * dy.h * dy.w * groups;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i % groups);
    const long long pix = i / groups;
    float d[8];
    bld8(dy, pix, g * 8, d);
    if (act != YC_LINEAR) {
      float o[8];
      bld8(y, pix, g * 8, o);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float gr = 1.f;
        if (act == YC_LEAKY) gr = o[j] > 0.f ? 1.f : .1f;
        else if (act == YC_RELU) gr = o[j] > 0.f ? 1.f : 0.f;
        else if (act == YC_LOGISTIC) gr = (1.f - o[j]) * o[j];
        d[j] *= gr;
      }
      bst8(dy, pix, g * 8, d);
    }
    float v[8];
    if (dprev.ptr) {
      bld8(dprev, pix, g * 8, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] += d[j];
      bst8(dprev, pix, g * 8, v);
    }
    if (dfrom.ptr) {
      bld8(dfrom, pix, g * 8, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] += d[j];
      bst8(dfrom, pix, g * 8, v);
    }
  }
}

}  // namespace yc

namespace yc {

// One parameter of UpdateConvolutionalLayerGpu (convolutional_kernels.cu:865-922).  Weight arrays first get the GPU path's
// scrub (:881-882): reset_nan_and_inf on the gradient (NaN / Inf -> 0) and fix_nan_and_inf on the weight (-> 1 / (index + 1),
// index = the element's OIHW position; blas_kernels.cu:1157-1201).  Then SGD — dW -= decay*batch*W (weights only);
// W += (lr/batch)*dW; dW *= momentum — or adam_update_gpu (blas_kernels.cu:99-134), which decays biases and scales too:
//   m = B1 m; v = B2 v; d -= decay*batch*w; m += (1-B1) d; d = d*d; v += (1-B2) d;
//   w += rate * (m / (1 - B1^t)) / (sqrt(v / (1 - B2^t)) + eps); d = 0.
// bc1 / bc2 = 1 - B^t are evaluated by the caller once per thread (device powf like the reference's adam_kernel).
__device__ __forceinline__ void upd_elem(const UpdRule& r, float bc1, float bc2, float& w, float& g, float* m, float* v,
                                         bool is_weight, int oihw_index) {
  if (is_weight) {
    if (isnan(g) || isinf(g)) g = 0.f;
    if (isnan(w) || isinf(w)) w = 1.0f / (fabsf((float)oihw_index) + 1);
  }
  if (r.adam) {
    float mm = *m * r.B1, vv = *v * r.B2;
    g = fmaf(-r.decay_batch, w, g);
    mm = fmaf(1.f - r.B1, g, mm);
    g = g * g;
    vv = fmaf(1.f - r.B2, g, vv);
    const float mhat = mm / bc1, vhat = vv / bc2;
    w = w + r.rate * mhat / (sqrtf(vhat) + r.eps);
    *m = mm;
    *v = vv;
    g = 0.f;
  } else {
    if (is_weight) g += -r.decay_batch * w;
    w += r.lr_over_batch * g;
    g *= r.momentum;
  }
}

// UpdateConvolutionalLayer (convolutional_layer.cpp:1382-1400), SGD with momentum and weight decay:
//   dW -= decay*batch*W ; W += (lr/batch)*dW ; dW *= momentum
// in one pass over the packed gradient; also refreshes the two bf16 operand packings the GEMMs read
// (forward [n][tap][c_p], data-gradient [c][flipped tap][n_p]) so no separate repack launch is needed.
__global__ void sgd_weights_kernel(float* __restrict__ w_oihw, float* __restrict__ dwp, int out_c, int in_c, int taps,
                                   int ldw, float lr_over_batch, float decay_batch, float momentum,
                                   __nv_bfloat16* __restrict__ wp_f, int in_c_p, __nv_bfloat16* __restrict__ wp_d,
                                   int out_c_p) {
  const long long total = (long long)out_c * taps * in_c;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % in_c);
    const int tap = (int)((i / in_c) % taps);
    const int n = (int)(i / ((long long)in_c * taps));
    const long long gi = ((long long)n * taps + tap) * ldw + c;
    const long long wi = ((long long)n * in_c + c) * taps + tap;
    float w = w_oihw[wi];
    float g = dwp[gi];
    const UpdRule r{lr_over_batch, decay_batch, momentum, 0, 0, 0.f, 0.f, 0.f, 0.f};
    upd_elem(r, 1.f, 1.f, w, g, nullptr, nullptr, true, (int)wi);
    w_oihw[wi] = w;
    dwp[gi] = g;
    const __nv_bfloat16 wb = __float2bfloat16_rn(w);
    if (wp_f) wp_f[((long long)n * taps + tap) * in_c_p + c] = wb;
    if (wp_d) wp_d[((long long)c * taps + (taps - 1 - tap)) * out_c_p + n] = wb;
  }
}
// biases / scales: p += (lr/batch)*dp ; dp *= momentum   (no decay)
__global__ void sgd_vector_kernel(float* __restrict__ p, float* __restrict__ dp, int n, float lr_over_batch,
                                  float momentum) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float g = dp[i];
  p[i] += lr_over_batch * g;
  dp[i] = g * momentum;
}
constexpr int kSgdTN = 16, kSgdTC = 32, kSgdMaxTaps = 9;  // tile of the weights branch of sgd_multi_kernel
// The whole network's update in ONE launch (yc_network_update): a table of items, 1024 elements per block, the block
// finds its item by bisection of the block prefix.  Items with taps == 0 are plain vectors (biases, scales).
__global__ void __launch_bounds__(256) sgd_multi_kernel(const SgdItem* __restrict__ items, int n_items, const UpdRule rule) {
  float bc1 = 1.f, bc2 = 1.f;
  if (rule.adam) {
    bc1 = 1.f - powf(rule.B1, (float)rule.t);
    bc2 = 1.f - powf(rule.B2, (float)rule.t);
  }
  int lo = 0, hi = n_items - 1;
  while (lo < hi) {  // last item whose first_block <= blockIdx.x
    const int mid = (lo + hi + 1) >> 1;
    if (items[mid].first_block <= (int)blockIdx.x) lo = mid; else hi = mid - 1;
  }
  const SgdItem it = items[lo];
  const long long base = (long long)((int)blockIdx.x - it.first_block) * 1024;
  if (it.taps == 0) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const long long i = base + k * 256 + threadIdx.x;
      if (i >= it.out_c) break;
      float g = it.dw[i], w = it.w[i];
      upd_elem(rule, bc1, bc2, w, g, it.m ? it.m + i : nullptr, it.v ? it.v + i : nullptr, false, 0);
      it.w[i] = w;
      it.dw[i] = g;
    }
    return;
  }
  if (it.taps <= kSgdMaxTaps) {
    // ---- weights, tiled: 16 filters x 32 input channels x all taps per block, every global access coalesced -----------
    // The master weights are OIHW ([n][c][tap]: what the reference's files and get_param hold), the gradient and the
    // forward packing are [n][tap][c], the data-gradient packing is [c][flipped tap][n]: a thread-per-element kernel in
    // any one of these orders is strided by `taps` floats in the second and scattered in the third (1.08 ms for the 62 M
    // parameters of Gaussian-yolov3 against a 0.2 ms floor).  Here the tile goes through shared memory: OIHW segments of
    // 32*taps contiguous floats in and out, packed rows of 32 channels, and 16 consecutive filters per (c, tap) row of
    // the data-gradient packing.
    __shared__ float Ws[kSgdTN][kSgdTC * kSgdMaxTaps + 1];
    __shared__ __nv_bfloat16 Bd[kSgdTC][kSgdMaxTaps][kSgdTN];
    const int taps = it.taps;
    const int c_tiles = (it.in_c + kSgdTC - 1) / kSgdTC;
    const int tile = (int)blockIdx.x - it.first_block;
    const int n0 = (tile / c_tiles) * kSgdTN, c0 = (tile % c_tiles) * kSgdTC;
    const int nw = min(kSgdTN, it.out_c - n0), cw = min(kSgdTC, it.in_c - c0);
    const int seg = cw * taps;  // contiguous OIHW floats of one filter inside this tile
    for (int idx = threadIdx.x; idx < nw * seg; idx += blockDim.x) {
      const int nl = idx / seg, r = idx - nl * seg;
      Ws[nl][r] = it.w[((long long)(n0 + nl) * it.in_c + c0) * taps + r];
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < nw * taps * kSgdTC; idx += blockDim.x) {
      const int cl = idx % kSgdTC, t = (idx / kSgdTC) % taps, nl = idx / (kSgdTC * taps);
      if (cl >= cw) continue;
      const int n = n0 + nl, c = c0 + cl;
      float* gp = it.dw + ((long long)n * taps + t) * it.ldw + c;
      float ww = Ws[nl][cl * taps + t];
      float gg = *gp;
      const long long go = gp - it.dw;
      upd_elem(rule, bc1, bc2, ww, gg, it.m ? it.m + go : nullptr, it.v ? it.v + go : nullptr, true,
               (int)(((long long)n * it.in_c + c) * taps + t));
      *gp = gg;
      Ws[nl][cl * taps + t] = ww;
      const __nv_bfloat16 wb = __float2bfloat16_rn(ww);
      if (it.wp_f) it.wp_f[((long long)n * taps + t) * it.in_c_p + c] = wb;
      Bd[cl][taps - 1 - t][nl] = wb;
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < nw * seg; idx += blockDim.x) {
      const int nl = idx / seg, r = idx - nl * seg;
      it.w[((long long)(n0 + nl) * it.in_c + c0) * taps + r] = Ws[nl][r];
    }
    if (it.wp_d) {
      for (int idx = threadIdx.x; idx < cw * taps * kSgdTN; idx += blockDim.x) {
        const int nl = idx % kSgdTN, t2 = (idx / kSgdTN) % taps, cl = idx / (kSgdTN * taps);
        if (nl < nw) it.wp_d[((long long)(c0 + cl) * taps + t2) * it.out_c_p + n0 + nl] = Bd[cl][t2][nl];
      }
    }
    return;
  }
  const long long total = (long long)it.out_c * it.taps * it.in_c;
  float w[4], g[4];
  long long gi[4], wi[4];
  int cc[4], tt[4], nn[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {  // loads of the four elements first
    const long long i = base + k * 256 + threadIdx.x;
    const bool ok = i < total;
    const long long ii = ok ? i : 0;
    cc[k] = (int)(ii % it.in_c);
    tt[k] = (int)((ii / it.in_c) % it.taps);
    nn[k] = ok ? (int)(ii / ((long long)it.in_c * it.taps)) : -1;
    gi[k] = ((long long)max(nn[k], 0) * it.taps + tt[k]) * it.ldw + cc[k];
    wi[k] = ((long long)max(nn[k], 0) * it.in_c + cc[k]) * it.taps + tt[k];
    w[k] = it.w[wi[k]];
    g[k] = it.dw[gi[k]];
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    if (nn[k] < 0) continue;
    float gg = g[k], ww = w[k];
    upd_elem(rule, bc1, bc2, ww, gg, it.m ? it.m + gi[k] : nullptr, it.v ? it.v + gi[k] : nullptr, true, (int)wi[k]);
    it.w[wi[k]] = ww;
    it.dw[gi[k]] = gg;
    const __nv_bfloat16 wb = __float2bfloat16_rn(ww);
    if (it.wp_f) it.wp_f[((long long)nn[k] * it.taps + tt[k]) * it.in_c_p + cc[k]] = wb;
    if (it.wp_d) it.wp_d[((long long)cc[k] * it.taps + (it.taps - 1 - tt[k])) * it.out_c_p + nn[k]] = wb;
  }
}

}  // namespace yc

using namespace yc;

static bool okv(const yc_tensor* t) {
  return t && t->ptr && t->dtype == YC_BF16 && t->n > 0 && t->h > 0 && t->w > 0 && t->c > 0 && t->ldc >= t->c &&
         t->ldc % 8 == 0 && ((uintptr_t)t->ptr & 15) == 0;
}
static int tl_grid(long long work) {
  return (int)std::max<long long>(1, std::min<long long>(ceil_div_ll(work, 256), (long long)sm_count() * 16));
}

extern "C" int yc_maxpool_fwd_train(const yc_tensor* x, const yc_tensor* y, unsigned char* idx, int size, int stride,
                                    int pad, void* stream) {
  YC_REQUIRE(okv(x) && okv(y) && idx && size >= 1 && size <= 15 && stride >= 1 && pad >= 0, "yc_maxpool_fwd_train: bad arguments");
  YC_REQUIRE(y->n == x->n && y->c == x->c && y->h == (x->h + pad - size) / stride + 1 &&
                 y->w == (x->w + pad - size) / stride + 1 && ((uintptr_t)idx & 7) == 0,
             "yc_maxpool_fwd_train: output shape mismatch");
  const long long work = (long long)y->n * y->h * y->w * ((y->c + 7) / 8);
  maxpool_train_kernel<<<tl_grid(work), 256, 0, (cudaStream_t)stream>>>(bv(x), bv(y), idx, size, stride, pad);
  YC_LAUNCH_CHECK();
  return YC_OK;
}
extern "C" int yc_maxpool_bwd(const yc_tensor* dy, const unsigned char* idx, const yc_tensor* dx, int size, int stride,
                              int pad, void* stream) {
  YC_REQUIRE(okv(dy) && okv(dx) && idx && size >= 1 && size <= 15 && stride >= 1, "yc_maxpool_bwd: bad arguments");
  YC_REQUIRE(dy->n == dx->n && dy->c == dx->c && dy->h == (dx->h + pad - size) / stride + 1 &&
                 dy->w == (dx->w + pad - size) / stride + 1,
             "yc_maxpool_bwd: shape mismatch");
  if (size == 2 && stride == 2 && pad / 2 == 0 && dx->c % 8 == 0 && dx->ldc % 8 == 0 && dy->ldc % 8 == 0) {
    const long long w2 = (long long)dy->n * dy->h * dy->w * (dx->c / 8);
    maxpool_bwd_2x2_kernel<<<tl_grid(w2), 256, 0, (cudaStream_t)stream>>>(bv(dy), idx, bv(dx));
    YC_LAUNCH_CHECK();
    return YC_OK;
  }
  const long long work = (long long)dx->n * dx->h * dx->w * ((dx->c + 7) / 8);
  maxpool_bwd_kernel<<<tl_grid(work), 256, 0, (cudaStream_t)stream>>>(bv(dy), idx, bv(dx), size, stride, pad);
  YC_LAUNCH_CHECK();
  return YC_OK;
}
extern "C" int yc_upsample_bwd(const yc_tensor* dy, const yc_tensor* dx, int stride, float scale, void* stream) {
  YC_REQUIRE(okv(dy) && okv(dx) && stride >= 1 && dy->n == dx->n && dy->c == dx->c && dy->h == dx->h * stride &&
                 dy->w == dx->w * stride,
             "yc_upsample_bwd: bad arguments");
  const long long work = (long long)dx->n * dx->h * dx->w * ((dx->c + 7) / 8);
  upsample_bwd_kernel<<<tl_grid(work), 256, 0, (cudaStream_t)stream>>>(bv(dy), bv(dx), stride, scale);
  YC_LAUNCH_CHECK();
  return YC_OK;
}
extern "C" int yc_axpy_view(const yc_tensor* src, const yc_tensor* dst, void* stream) {
  YC_REQUIRE(okv(src) && okv(dst) && src->n == dst->n && src->h == dst->h && src->w == dst->w && src->c == dst->c,
             "yc_axpy_view: bad arguments");
  const long long work = (long long)dst->n * dst->h * dst->w * ((dst->c + 7) / 8);
  axpy_view_kernel<<<tl_grid(work), 256, 0, (cudaStream_t)stream>>>(bv(src), bv(dst));
  YC_LAUNCH_CHECK();
  return YC_OK;
}
extern "C" int yc_shortcut_bwd(const yc_tensor* dy, const yc_tensor* y, int activation, const yc_tensor* dprev,
                               const yc_tensor* dfrom, void* stream) {
  YC_REQUIRE(okv(dy) && (activation == YC_LINEAR || okv(y)) && (!dprev || okv(dprev)) && (!dfrom || okv(dfrom)),
             "yc_shortcut_bwd: bad arguments");
  YC_REQUIRE(activation == YC_LINEAR || activation == YC_LEAKY || activation == YC_RELU || activation == YC_LOGISTIC,
             "yc_shortcut_bwd: activation outside the hot path");
  BV none{nullptr, 0, 0, 0, 0, 0};
  const long long work = (long long)dy->n * dy->h * dy->w * ((dy->c + 7) / 8);
  shortcut_bwd_kernel<<<tl_grid(work), 256, 0, (cudaStream_t)stream>>>(bv(dy), activation == YC_LINEAR ? bv(dy) : bv(y),
                                                                       activation, dprev ? bv(dprev) : none,
                                                                       dfrom ? bv(dfrom) : none);
  YC_LAUNCH_CHECK();
  return YC_OK;
}
// blocks a table entry needs: tiles of 16 filters x 32 input channels for weights (up to 3x3 filters), 1024 elements per
// block for vectors and larger filters
int yc::sgd_item_blocks(const SgdItem& it) {
  if (it.taps > 0 && it.taps <= kSgdMaxTaps) return ceil_div(it.out_c, kSgdTN) * ceil_div(it.in_c, kSgdTC);
  const long long elements = it.taps > 0 ? (long long)it.out_c * it.taps * it.in_c : it.out_c;
  return (int)((elements + 1023) / 1024);
}
int yc::sgd_update_multi(const SgdItem* items_dev, int n_items, int total_blocks, const UpdRule& rule, cudaStream_t st) {
  if (n_items <= 0 || total_blocks <= 0) return YC_OK;
  sgd_multi_kernel<<<total_blocks, 256, 0, st>>>(items_dev, n_items, rule);
  YC_LAUNCH_CHECK();
  return YC_OK;
}
extern "C" int yc_sgd_update_conv(float* w_oihw, float* dw_packed, int out_c, int in_c, int ksize, float lr, int batch,
                                  float momentum, float decay, void* w_packed_fwd, void* w_packed_dgrad, float* biases,
                                  float* bias_updates, float* scales, float* scale_updates, void* stream) {
  YC_REQUIRE(w_oihw && dw_packed && out_c > 0 && in_c > 0 && ksize > 0 && batch > 0, "yc_sgd_update_conv: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  const int taps = ksize * ksize;
  const long long total = (long long)out_c * taps * in_c;
  const int in_c_p = yc_conv_padded_in_c(in_c);
  const int out_c_p = yc_conv_padded_in_c(yc_conv_dgrad_channels(out_c));
  sgd_weights_kernel<<<tl_grid(total), 256, 0, st>>>(w_oihw, dw_packed, out_c, in_c, taps, yc_conv_wgrad_ld(in_c),
                                                     lr / batch, decay * batch, momentum, (__nv_bfloat16*)w_packed_fwd,
                                                     in_c_p, (__nv_bfloat16*)w_packed_dgrad, out_c_p);
  YC_LAUNCH_CHECK();
  if (biases && bias_updates) {
    sgd_vector_kernel<<<ceil_div(out_c, 128), 128, 0, st>>>(biases, bias_updates, out_c, lr / batch, momentum);
    YC_LAUNCH_CHECK();
  }
  if (scales && scale_updates) {
    sgd_vector_kernel<<<ceil_div(out_c, 128), 128, 0, st>>>(scales, scale_updates, out_c, lr / batch, momentum);
    YC_LAUNCH_CHECK();
  }
  return YC_OK;
}
