// Training-mode batchnorm / bias / activation-gradient passes around the conv GEMMs (NHWC bf16, fp32 math).
//
// Replaces, for a convolutional layer in training (convolutional_kernels.cu:252-553, 555-815 and their CPU twins):
//   forward : fast_mean_gpu / fast_variance_gpu (fused into the conv epilogue, conv_fprop_sm100.cu), the rolling
//             update, normalize_gpu, scale_bias_gpu, add_bias_gpu, activate_array_ongpu   -> bn_finalize + bn_act_fwd
//   backward: gradient_array_ongpu, backward_scale_gpu, scale_bias, fast_mean_delta_gpu, fast_variance_delta_gpu,
//             normalize_delta_gpu (batchnorm_layer.cpp:240-255)                            -> bn_bwd_reduce + bn_bwd_apply
//             gradient_array_ongpu + backward_bias_gpu for layers without batchnorm       -> bias_act_bwd
// Numerical semantics follow the reference CPU path.
// Each pass reads/writes every tensor element once with 16-byte accesses; reductions go registers -> shared-memory
// atomics -> one global atomic per (block, channel).
#include <algorithm>
#include <mutex>

#include "yc_common.cuh"

#ifndef YC_BN_MINB
#define YC_BN_MINB 2  // resident blocks per SM the three row kernels are compiled for (3 caps them at 80 registers: spills)
#endif
namespace yc {

struct TView {
  void* ptr;
  long long rows;  // n*h*w
  int c;
  long long ldc;
  int f32;
};
static inline TView tv(const yc_tensor* t) {
  return TView{t->ptr, (long long)t->n * t->h * t->w, t->c, t->ldc, t->dtype == YC_F32};
}

__device__ __forceinline__ void ld8(const TView& t, long long row, int c0, float (&v)[8]) {
  if (t.f32) {
    const float* p = reinterpret_cast<const float*>(t.ptr) + row * t.ldc + c0;
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = (c0 + j < t.c) ? p[j] : 0.f;
    return;
  }
  const __nv_bfloat16* p = reinterpret_cast<const __nv_bfloat16*>(t.ptr) + row * t.ldc + c0;
  if (c0 + 8 <= t.ldc) {  // whole 16-byte group lies inside the pixel pitch (pad lanes are readable)
    const uint4 q = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 f = __bfloat1622float2(h[e]);
      v[2 * e] = f.x;
      v[2 * e + 1] = f.y;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (c0 + j >= t.c) v[j] = 0.f;
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = (c0 + j < t.c) ? __bfloat162float(p[j]) : 0.f;
  }
}
// pad_zero: the tensor owns lanes [c, ldc) and wants zeros there (delta buffers feeding the GEMMs)
__device__ __forceinline__ void st8(const TView& t, long long row, int c0, const float (&v)[8], bool own_pad) {
  if (t.f32) {
    float* p = reinterpret_cast<float*>(t.ptr) + row * t.ldc + c0;
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (c0 + j < t.c) p[j] = v[j];
    return;
  }
  __nv_bfloat16* p = reinterpret_cast<__nv_bfloat16*>(t.ptr) + row * t.ldc + c0;
  if (c0 + 8 <= t.c || (own_pad && c0 + 8 <= t.ldc)) {
    uint4 q;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&q);
#pragma unroll
    for (int e = 0; e < 4; ++e)
      h[e] = __floats2bfloat162_rn(c0 + 2 * e < t.c ? v[2 * e] : 0.f, c0 + 2 * e + 1 < t.c ? v[2 * e + 1] : 0.f);
    *reinterpret_cast<uint4*>(p) = q;
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (c0 + j < t.c) p[j] = __float2bfloat16_rn(v[j]);
  }
}

// Two tensors x two rows: all four 16-byte loads are issued BEFORE any conversion.  (ld8 converts inside its own
// branch, so back-to-back ld8 calls leave one load in flight per thread: measured 1.1 TB/s on the BN passes.)
__device__ __forceinline__ void cvt8(const uint4& q, int c0, int c, float (&v)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const float2 f = __bfloat1622float2(h[e]);
    v[2 * e] = f.x;
    v[2 * e + 1] = f.y;
  }
#pragma unroll
  for (int j = 0; j < 8; ++j)
    if (c0 + j >= c) v[j] = 0.f;
}
__device__ __forceinline__ uint4 ldq(const TView& t, long long row, int c0, bool pred) {
  uint4 q = make_uint4(0u, 0u, 0u, 0u);
  if (pred) q = __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(t.ptr) + row * t.ldc + c0));
  return q;
}
__device__ __forceinline__ bool fast_view(const TView& t, int c0) { return !t.f32 && c0 + 8 <= t.ldc; }
// rows r0 (always) and r1 (if two) of a and b; absent rows read as zeros
__device__ __forceinline__ void ld8_2x2(const TView& a, const TView& b, long long r0, long long r1, bool two, int c0,
                                        float (&a0)[8], float (&b0)[8], float (&a1)[8], float (&b1)[8]) {
  if (fast_view(a, c0) && fast_view(b, c0)) {
    const uint4 qa0 = ldq(a, r0, c0, true), qb0 = ldq(b, r0, c0, true);
    const uint4 qa1 = ldq(a, r1, c0, two), qb1 = ldq(b, r1, c0, two);
    cvt8(qa0, c0, a.c, a0);
    cvt8(qb0, c0, b.c, b0);
    cvt8(qa1, c0, a.c, a1);
    cvt8(qb1, c0, b.c, b1);
  } else {
    ld8(a, r0, c0, a0);
    ld8(b, r0, c0, b0);
    if (two) {
      ld8(a, r1, c0, a1);
      ld8(b, r1, c0, b1);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) a1[j] = b1[j] = 0.f;
    }
  }
}
// one tensor x four rows
__device__ __forceinline__ void ld8_1x4(const TView& a, const long long (&r)[4], const bool (&ok)[4], int c0, float (&v)[4][8]) {
  if (fast_view(a, c0)) {
    uint4 q[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) q[k] = ldq(a, r[k], c0, ok[k]);
#pragma unroll
    for (int k = 0; k < 4; ++k) cvt8(q[k], c0, a.c, v[k]);
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (ok[k]) ld8(a, r[k], c0, v[k]);
      else {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[k][j] = 0.f;
      }
    }
  }
}

// gradient of the activation at pre-activation u (activations.h:152-178, activations.c:412-452).  The reference
// evaluates leaky/relu/logistic on the post-activation output; those are the same functions of u.
__device__ __forceinline__ float act_grad(float u, int act) {
  switch (act) {
    case YC_LINEAR: return 1.f;
    case YC_LEAKY: return u > 0.f ? 1.f : .1f;
    case YC_RELU: return u > 0.f ? 1.f : 0.f;
    case YC_LOGISTIC: {
      const float y = 1.f / (1.f + expf(-u));
      return (1.f - y) * y;
    }
    case YC_SWISH: {  // gradient_array_swish: swish + sigmoid*(1-swish)
      const float s = 1.f / (1.f + expf(-u));
      const float y = u * s;
      return y + s * (1.f - y);
    }
    case YC_MISH: {  // gradient_array_mish (activations.c:426-452)
      const float sp = u > 20.f ? u : (u < -20.f ? expf(u) : logf(expf(u) + 1.f));
      const float grad_sp = 1.f - expf(-sp);
      const float tsp = tanhf(sp);
      const float grad_tsp = (1.f - tsp * tsp) * grad_sp;
      return u * grad_tsp + tsp;
    }
  }
  return 1.f;
}
// gradient evaluated on the post-activation output, as gradient_array(l->output, ...) does for layers without a
// saved pre-activation (activations.h:152-178)
__device__ __forceinline__ float act_grad_out(float y, int act) {
  switch (act) {
    case YC_LEAKY: return y > 0.f ? 1.f : .1f;
    case YC_RELU: return y > 0.f ? 1.f : 0.f;
    case YC_LOGISTIC: return (1.f - y) * y;
  }
  return 1.f;
}
__device__ __forceinline__ float act_fwd_exact(float x, int act) {  // training forward: libm-accurate forms
  switch (act) {
    case YC_LINEAR: return x;
    case YC_LEAKY: return x > 0.f ? x : .1f * x;
    case YC_RELU: return x > 0.f ? x : 0.f;
    case YC_LOGISTIC: return 1.f / (1.f + expf(-x));
    case YC_SWISH: return x / (1.f + expf(-x));
    case YC_MISH: {
      const float sp = x > 20.f ? x : (x < -20.f ? expf(x) : logf(expf(x) + 1.f));
      return x * (2.f / (1.f + expf(-2.f * sp)) - 1.f);
    }
  }
  return x;
}

// ---- per-channel finalize of the statistics the conv epilogue accumulated --------------------------------
// mean_cpu / variance_cpu (blas.c:164-195: unbiased 1/(N-1)) + rolling update .9/.1 (batchnorm_layer.cpp:221-224)
__global__ void bn_finalize_kernel(const float* __restrict__ sum, const float* __restrict__ sq, float count,
                                   float* __restrict__ mean, float* __restrict__ var, float* __restrict__ rmean,
                                   float* __restrict__ rvar, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float m = sum[i] / count;
  float v = (sq[i] - count * m * m) / (count - 1.f);
  v = fmaxf(v, 0.f);
  mean[i] = m;
  var[i] = v;
  if (rmean) rmean[i] = .9f * rmean[i] + .1f * m;
  if (rvar) rvar[i] = .9f * rvar[i] + .1f * v;
}

// y = act( gamma * (z - mean)/sqrt(var + 1e-6) + beta )   (normalize_cpu blas.c:214, scale_bias, add_bias)
// thread = (row slot, 8-channel group): the per-channel constants live in registers, the thread walks rows with a
// stride of gridDim*slots, two rows per iteration so that both 16-byte loads are in flight before the math.
template <int ACT>
__global__ void __launch_bounds__(256, YC_BN_MINB) bn_act_fwd_kernel(TView z, const float* __restrict__ mean, const float* __restrict__ var,
                                  const float* __restrict__ gamma, const float* __restrict__ beta, TView y,
                                  int y_own_pad) {
  constexpr int act = ACT;  // compile-time: with a runtime switch unrolled 32x the kernel was 64 KB of code
  const int groups = (z.c + 7) / 8;
  const int tg = threadIdx.x % groups;  // blockDim.x is a multiple of groups (host)
  const int slots = blockDim.x / groups;
  const int slot = threadIdx.x / groups;
  float A[8], Bc[8];  // y = act(A*z + Bc): the SAME expression the backward kernels evaluate, so leaky' sees the same sign
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = tg * 8 + j;
    const bool ok = c < z.c;
    A[j] = ok ? gamma[c] * (1.f / sqrtf(var[c] + .000001f)) : 0.f;
    Bc[j] = ok ? beta[c] - mean[c] * A[j] : 0.f;
  }
  const long long step = (long long)gridDim.x * slots;
  long long row = (long long)blockIdx.x * slots + slot;
  if (fast_view(z, tg * 8)) {
    // software pipeline: the next four rows are requested before the current four are converted and stored
    uint4 q[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) q[k] = ldq(z, row + k * step, tg * 8, row + k * step < z.rows);
    for (; row < z.rows; row += 4 * step) {
      uint4 nq[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) nq[k] = ldq(z, row + (4 + k) * step, tg * 8, row + (4 + k) * step < z.rows);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const long long r = row + k * step;
        if (r < z.rows) {
          float v[8];
          cvt8(q[k], tg * 8, z.c, v);
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = act_fwd_exact(fmaf(A[j], v[j], Bc[j]), act);
          st8(y, r, tg * 8, v, y_own_pad != 0);
        }
        q[k] = nq[k];
      }
    }
  }
  for (; row < z.rows; row += 4 * step) {
    const long long r[4] = {row, row + step, row + 2 * step, row + 3 * step};
    const bool ok[4] = {true, r[1] < z.rows, r[2] < z.rows, r[3] < z.rows};
    float v[4][8];
    ld8_1x4(z, r, ok, tg * 8, v);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (!ok[k]) continue;
#pragma unroll
      for (int j = 0; j < 8; ++j) v[k][j] = act_fwd_exact(fmaf(A[j], v[k][j], Bc[j]), act);
      st8(y, r[k], tg * 8, v[k], y_own_pad != 0);
    }
  }
}

// ---- backward reductions ------------------------------------------------------------------------------------
// MODE 0 (batchnorm):  g = dy * act'(u), u = gamma*xhat + beta;  S1[c] += sum g ; S2[c] += sum g*(z - mean)
// MODE 1 (bias only):  g = dy * act'(y) evaluated on the layer OUTPUT y (leaky/relu/logistic/linear): S1 only
// thread = (row slot, 8-channel group); registers -> shared atomics -> global atomics.
template <int MODE, int ACT>
__global__ void __launch_bounds__(256, YC_BN_MINB) bwd_reduce_kernel(TView dy, TView z, const float* __restrict__ mean, const float* __restrict__ var,
                                  const float* __restrict__ gamma, const float* __restrict__ beta,
                                  float* __restrict__ S1, float* __restrict__ S2, float* __restrict__ P,
                                  unsigned* __restrict__ ticket, float* __restrict__ scale_updates = nullptr,
                                  float* __restrict__ bias_updates = nullptr, float* __restrict__ coef = nullptr,
                                  float m_rows = 0.f) {
  constexpr int act = ACT;
  // Reproducible sums (no float atomics): thread sums -> one shared-memory entry per (warp | row slot, channel) -> the
  // block's partial = entries added in index order -> P[block][2][C8] -> the block that draws the last ticket adds the
  // partials in block order.  MODE 0 overwrites S1 / S2, MODE 1 adds the total to S1 (bias_updates accumulate, B7).
  extern __shared__ float sh[];  // [entries][2][groups*8]
  const int groups = (dy.c + 7) / 8;
  const int C8 = groups * 8;
  const int tg = threadIdx.x % groups;  // blockDim.x is a multiple of groups (host)
  const int slots = blockDim.x / groups;
  const int slot = threadIdx.x / groups;
  float a1[8], a2[8];
  float mu[8], A[8], Bc[8];  // u = A*x + Bc
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    a1[j] = a2[j] = 0.f;
    const int c = tg * 8 + j;
    const bool ok = c < dy.c;
    mu[j] = (MODE == 0 && ok) ? mean[c] : 0.f;
    A[j] = (MODE == 0 && ok) ? gamma[c] * (1.f / sqrtf(var[c] + .000001f)) : 1.f;
    Bc[j] = ((ok && beta) ? beta[c] : 0.f) - mu[j] * A[j];
  }
  if (slot < slots) {
    const long long step = (long long)gridDim.x * slots;
    long long row = (long long)blockIdx.x * slots + slot;
    if (fast_view(dy, tg * 8) && fast_view(z, tg * 8)) {
      // software pipeline: the loads of the NEXT two rows are issued before the math of the current two, so a
      // thread always has four 16-byte loads in flight (load latency and math overlap instead of alternating)
      uint4 qd[2], qz[2];
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const long long r = row + k * step;
        qd[k] = ldq(dy, r, tg * 8, r < dy.rows);
        qz[k] = ldq(z, r, tg * 8, r < dy.rows);
      }
      for (; row < dy.rows; row += 2 * step) {
        uint4 nd[2], nz[2];
#pragma unroll
        for (int k = 0; k < 2; ++k) {
          const long long r = row + (2 + k) * step;
          nd[k] = ldq(dy, r, tg * 8, r < dy.rows);
          nz[k] = ldq(z, r, tg * 8, r < dy.rows);
        }
#pragma unroll
        for (int k = 0; k < 2; ++k) {
          float d[8], x[8];
          cvt8(qd[k], tg * 8, dy.c, d);  // absent rows: d = 0, so they add nothing
          cvt8(qz[k], tg * 8, z.c, x);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            if (MODE == 0) {
              const float g = d[j] * act_grad(fmaf(A[j], x[j], Bc[j]), act);
              a1[j] += g;
              a2[j] = fmaf(g, x[j] - mu[j], a2[j]);
            } else {
              a1[j] = fmaf(d[j], act_grad_out(x[j], act), a1[j]);
            }
          }
          qd[k] = nd[k];
          qz[k] = nz[k];
        }
      }
    }
    for (; row < dy.rows; row += 2 * step) {
      const long long row2 = row + step;
      const bool two = row2 < dy.rows;
      float d[8], x[8], d2[8], x2[8];
      ld8_2x2(dy, z, row, row2, two, tg * 8, d, x, d2, x2);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (MODE == 0) {
          const float xc = x[j] - mu[j], xc2 = x2[j] - mu[j];
          const float g = d[j] * act_grad(fmaf(A[j], x[j], Bc[j]), act);
          const float g2 = d2[j] * act_grad(fmaf(A[j], x2[j], Bc[j]), act);   // d2 = 0 when the row is absent
          a1[j] += g + g2;
          a2[j] += g * xc + g2 * xc2;
        } else {
          a1[j] += d[j] * act_grad_out(x[j], act) + d2[j] * act_grad_out(x2[j], act);
        }
      }
    }
    // lanes l and l + groups of a warp own the same channels when groups divides 32: fold them with shuffles first,
    // so that `groups` lanes per warp (not 32) issue the shared-memory atomics (32-way same-address conflicts before)
    const bool fold = groups < 32 && (32 % groups) == 0 && (blockDim.x & 31) == 0;
    if (fold) {
      for (int off = 16; off >= groups; off >>= 1) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          a1[j] += __shfl_xor_sync(0xffffffffu, a1[j], off);
          if (MODE == 0) a2[j] += __shfl_xor_sync(0xffffffffu, a2[j], off);
        }
      }
    }
    const int e = fold ? (threadIdx.x >> 5) : slot;
    if (!fold || (threadIdx.x & 31) < groups) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        sh[(e * 2) * C8 + tg * 8 + j] = a1[j];
        if (MODE == 0) sh[(e * 2 + 1) * C8 + tg * 8 + j] = a2[j];
      }
    }
  }
  __syncthreads();
  {
    const bool fold = groups < 32 && (32 % groups) == 0 && (blockDim.x & 31) == 0;
    const int entries = fold ? (int)(blockDim.x >> 5) : slots;
    float* mine = P + (size_t)blockIdx.x * 2 * C8;
    for (int i = threadIdx.x; i < C8; i += blockDim.x) {
      float s1 = 0.f, s2 = 0.f;
      for (int e = 0; e < entries; ++e) {
        s1 += sh[(e * 2) * C8 + i];
        if (MODE == 0) s2 += sh[(e * 2 + 1) * C8 + i];
      }
      mine[i] = s1;
      if (MODE == 0) mine[C8 + i] = s2;
    }
  }
  __threadfence();
  __syncthreads();
  __shared__ unsigned is_last;
  if (threadIdx.x == 0) is_last = atomicAdd(ticket, 1u) == gridDim.x - 1 ? 1u : 0u;
  __syncthreads();
  if (is_last) {
    __threadfence();
    for (int i = threadIdx.x; i < dy.c; i += blockDim.x) {
      float s1 = 0.f, s2 = 0.f;
      for (unsigned b = 0; b < gridDim.x; ++b) {
        s1 += __ldcg(P + (size_t)b * 2 * C8 + i);
        if (MODE == 0) s2 += __ldcg(P + (size_t)b * 2 * C8 + C8 + i);
      }
      if (MODE == 0) {
        S1[i] = s1;
        S2[i] = s2;
        if (coef) {  // the per-channel coefficients right here (what bn_bwd_coef_kernel computes: one launch less per layer)
          const float v = var[i], ga = gamma[i];
          if (scale_updates) scale_updates[i] += s2 / sqrtf(v + .000001f);
          if (bias_updates) bias_updates[i] += s1;
          const double md = (double)(ga * s1) * (-1. / sqrt((double)(v + .00001f)));
          const double vd = (double)(ga * s2) * (-.5 * pow((double)(v + .00001f), (double)(float)(-3. / 2.)));
          coef[i] = (float)((double)ga / (sqrt((double)v) + .00001));
          coef[dy.c + i] = (float)(vd * 2. / m_rows);
          coef[2 * dy.c + i] = (float)(md / m_rows);
        }
      } else {
        S1[i] += s1;
      }
    }
    if (threadIdx.x == 0) *ticket = 0u;  // ready for the next launch
  }
}

// per-channel coefficients of normalize_delta_cpu (batchnorm_layer.cpp:112-166) from the two sums:
//   scale_updates += S2 / sqrt(var + 1e-6)                                   (backward_scale_cpu on x_norm)
//   mean_delta     = gamma*S1 * (-1/sqrt(var + 1e-5))
//   variance_delta = gamma*S2 * (-.5 * (var + 1e-5)^(-3/2))
//   dz = gamma*g /(sqrt(var) + 1e-5) + variance_delta*2*(z - mean)/m + mean_delta/m
// coef[0][c] = gamma/(sqrt(var)+1e-5), coef[1][c] = 2*variance_delta/m, coef[2][c] = mean_delta/m
__global__ void bn_bwd_coef_kernel(const float* __restrict__ S1, const float* __restrict__ S2,
                                   const float* __restrict__ var, const float* __restrict__ gamma, float m,
                                   float* __restrict__ scale_updates, float* __restrict__ bias_updates,
                                   float* __restrict__ coef, int n) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n) return;
  const float v = var[c], ga = gamma[c];
  if (scale_updates) scale_updates[c] += S2[c] / sqrtf(v + .000001f);
  if (bias_updates) bias_updates[c] += S1[c];
  const double md = (double)(ga * S1[c]) * (-1. / sqrt((double)(v + .00001f)));
  const double vd = (double)(ga * S2[c]) * (-.5 * pow((double)(v + .00001f), (double)(float)(-3. / 2.)));
  coef[c] = (float)((double)ga / (sqrt((double)v) + .00001));
  coef[n + c] = (float)(vd * 2. / m);
  coef[2 * n + c] = (float)(md / m);
}

template <int ACT>
__global__ void __launch_bounds__(256, YC_BN_MINB) bn_bwd_apply_kernel(TView dy, TView z, const float* __restrict__ mean, const float* __restrict__ var,
                                    const float* __restrict__ gamma, const float* __restrict__ beta,
                                    const float* __restrict__ coef, TView dz, int groups) {
  constexpr int act = ACT;
  // thread = (row slot, 8-channel group of dz's pitch); groups beyond the data channels only write zeros
  const int n = dy.c;
  const int tg = threadIdx.x % groups;  // blockDim.x is a multiple of groups (host)
  const int slots = blockDim.x / groups;
  const int slot = threadIdx.x / groups;
  const bool live = tg * 8 < n;
  // per-channel constants folded to five registers each: u = A*x + Bc (the pre-activation), dz = g*c0 + c1*x + c2'
  float A[8], Bc[8], c0[8], c1[8], c2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = tg * 8 + j;
    const bool ok = c < n;
    const float mu = ok ? mean[c] : 0.f;
    A[j] = ok ? gamma[c] * (1.f / sqrtf(var[c] + .000001f)) : 0.f;
    Bc[j] = ok ? beta[c] - mu * A[j] : 0.f;
    c0[j] = ok ? coef[c] : 0.f;
    c1[j] = ok ? coef[n + c] : 0.f;
    c2[j] = ok ? coef[2 * n + c] - c1[j] * mu : 0.f;
  }
  const long long step = (long long)gridDim.x * slots;
  long long row = (long long)blockIdx.x * slots + slot;
  if (live && fast_view(dy, tg * 8) && fast_view(z, tg * 8)) {
    // software pipeline (see bwd_reduce_kernel): next two rows' loads before the current rows' math and stores
    uint4 qd[2], qz[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const long long r = row + k * step;
      qd[k] = ldq(dy, r, tg * 8, r < dy.rows);
      qz[k] = ldq(z, r, tg * 8, r < dy.rows);
    }
    for (; row < dy.rows; row += 2 * step) {
      uint4 nd[2], nz[2];
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const long long r = row + (2 + k) * step;
        nd[k] = ldq(dy, r, tg * 8, r < dy.rows);
        nz[k] = ldq(z, r, tg * 8, r < dy.rows);
      }
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const long long r = row + k * step;
        if (r < dy.rows) {
          float d[8], x[8];
          cvt8(qd[k], tg * 8, dy.c, d);
          cvt8(qz[k], tg * 8, z.c, x);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float gg = d[j] * act_grad(fmaf(A[j], x[j], Bc[j]), act);
            d[j] = fmaf(gg, c0[j], fmaf(c1[j], x[j], c2[j]));
          }
          st8(dz, r, tg * 8, d, true);
        }
        qd[k] = nd[k];
        qz[k] = nz[k];
      }
    }
  }
  for (; row < dy.rows; row += 2 * step) {
    const long long row2 = row + step;
    const bool two = row2 < dy.rows;
    float d[8] = {0, 0, 0, 0, 0, 0, 0, 0}, x[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    float d2[8] = {0, 0, 0, 0, 0, 0, 0, 0}, x2[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (live) {
      ld8_2x2(dy, z, row, row2, two, tg * 8, d, x, d2, x2);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float gg = d[j] * act_grad(fmaf(A[j], x[j], Bc[j]), act);
        const float gg2 = d2[j] * act_grad(fmaf(A[j], x2[j], Bc[j]), act);
        d[j] = fmaf(gg, c0[j], fmaf(c1[j], x[j], c2[j]));      // channels >= n: all coefficients are 0
        d2[j] = fmaf(gg2, c0[j], fmaf(c1[j], x2[j], c2[j]));
      }
    }
    st8(dz, row, tg * 8, d, true);
    if (two) st8(dz, row2, tg * 8, d2, true);
  }
}

// no batchnorm: dz = dy * act'(y) on the layer output y; bias_updates += column sums of dz
// (gradient_array + backward_bias, convolutional_layer.cpp:1327-1336)
template <int ACT>
__global__ void act_bwd_apply_kernel(TView dy, TView u, TView dz, int groups) {
  constexpr int act = ACT;
  const long long total = dy.rows * groups;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i % groups);
    const long long row = i / groups;
    float d[8] = {0, 0, 0, 0, 0, 0, 0, 0}, x[8];
    if (g * 8 < dy.c) {
      ld8(dy, row, g * 8, d);
      if (act != YC_LINEAR) {
        ld8(u, row, g * 8, x);
#pragma unroll
        for (int j = 0; j < 8; ++j) d[j] *= act_grad_out(x[j], act);
      }
    }
    st8(dz, row, g * 8, d, true);
  }
}

// grid for the row-walking kernels: at most ONE wave of resident blocks (YC_BN_MINB per SM).  Every block pays a
// dependent chain of per-channel constant loads before its first row; with 8 blocks per SM the mid-size layers ran
// 2-4 waves of that chain (26^2 x 256 x 64 images: 42 us for 44 MB), with one wave the software-pipelined row loop
// amortises it.
static int row_grid(long long rows, int slots, int per_iter = 2) {
  return (int)std::max<long long>(1, std::min<long long>(ceil_div_ll(rows, (long long)slots * per_iter), (long long)sm_count() * YC_BN_MINB));
}
// instantiate a launch for the activation id (compile-time in the kernels)
#define YC_ACT_DISPATCH(act_id, ...)                                                  \
  switch (act_id) {                                                                   \
    case YC_LINEAR: { constexpr int A = YC_LINEAR; __VA_ARGS__; } break;              \
    case YC_LEAKY: { constexpr int A = YC_LEAKY; __VA_ARGS__; } break;                \
    case YC_RELU: { constexpr int A = YC_RELU; __VA_ARGS__; } break;                  \
    case YC_LOGISTIC: { constexpr int A = YC_LOGISTIC; __VA_ARGS__; } break;          \
    case YC_SWISH: { constexpr int A = YC_SWISH; __VA_ARGS__; } break;                \
    case YC_MISH: { constexpr int A = YC_MISH; __VA_ARGS__; } break;                  \
    default: set_error("unsupported activation %d", (int)(act_id)); return YC_ERR_ARG; \
  }
// dynamic shared memory of bwd_reduce_kernel: one entry per warp when the channel groups fold inside a warp, else one
// per row slot; each entry is [2][groups*8] floats
static size_t reduce_smem(int groups, int threads) {
  const bool fold = groups < 32 && (32 % groups) == 0 && (threads & 31) == 0;
  const int entries = fold ? threads / 32 : threads / groups;
  return sizeof(float) * 2 * (size_t)groups * 8 * entries;
}
static int reduce_threads(int groups) {
  int t = (256 / groups) * groups;
  if (t < groups) t = groups;  // more than 256 groups (c > 2048): refused by the callers (launch bounds are 256)
  return t;
}

}  // namespace yc

using namespace yc;

static bool ok_view(const yc_tensor* t) {
  return t && t->ptr && t->n > 0 && t->h > 0 && t->w > 0 && t->c > 0 && t->ldc >= t->c &&
         ((uintptr_t)t->ptr & 15) == 0 && t->ldc % (t->dtype == YC_F32 ? 4 : 8) == 0;
}
static int ew_grid(long long work) {
  long long blocks = ceil_div_ll(work, 256);
  return (int)std::max<long long>(1, std::min<long long>(blocks, (long long)sm_count() * 16));
}

// floats of the workspace yc_bn_bwd / yc_bias_act_bwd need for c channels: S1, S2, three coefficient vectors, the
// ticket, and one [2][ceil8(c)] partial per block of the reduction grid (at most one wave of resident blocks)
extern "C" size_t yc_bn_bwd_workspace_floats(int c) {
  const size_t c8 = (size_t)round_up(c, 8);
  return 5 * (size_t)c + 4 + (size_t)sm_count() * YC_BN_MINB * 2 * c8;
}

extern "C" int yc_bn_finalize(const float* sum, const float* sumsq, long long count, float* mean, float* var,
                              float* rolling_mean, float* rolling_var, int n, void* stream) {
  YC_REQUIRE(sum && sumsq && mean && var && n > 0 && count > 1, "yc_bn_finalize: bad arguments");
  bn_finalize_kernel<<<ceil_div(n, 128), 128, 0, (cudaStream_t)stream>>>(sum, sumsq, (float)count, mean, var,
                                                                         rolling_mean, rolling_var, n);
  YC_LAUNCH_CHECK();
  return YC_OK;
}

extern "C" int yc_bn_act_fwd(const yc_tensor* z, const float* mean, const float* var, const float* gamma,
                             const float* beta, int activation, const yc_tensor* y, void* stream) {
  YC_REQUIRE(ok_view(z) && ok_view(y) && mean && var && gamma && beta, "yc_bn_act_fwd: bad arguments");
  YC_REQUIRE(z->n == y->n && z->h == y->h && z->w == y->w && z->c == y->c, "yc_bn_act_fwd: shape mismatch");
  YC_REQUIRE(z->c <= 2048, "yc_bn_act_fwd: more than 2048 channels");
  const TView zv = tv(z), yv = tv(y);
  const int own = y->ldc == round_up(y->c, 8);
  const int groups = (z->c + 7) / 8;
  const int threads = reduce_threads(groups);
  YC_ACT_DISPATCH(activation, (bn_act_fwd_kernel<A><<<row_grid(zv.rows, threads / groups, 4), threads, 0, (cudaStream_t)stream>>>(
                                  zv, mean, var, gamma, beta, yv, own)));
  YC_LAUNCH_CHECK();
  return YC_OK;
}

extern "C" int yc_bn_bwd(const yc_tensor* dy, const yc_tensor* z, const float* mean, const float* var,
                         const float* gamma, const float* beta, int activation, float* scale_updates,
                         float* bias_updates, float* workspace, const yc_tensor* dz, void* stream) {
  YC_REQUIRE(ok_view(dy) && ok_view(z) && ok_view(dz) && mean && var && gamma && beta && workspace,
             "yc_bn_bwd: bad arguments");
  YC_REQUIRE(dy->n == z->n && dy->h == z->h && dy->w == z->w && dy->c == z->c && dz->n == dy->n && dz->h == dy->h &&
                 dz->w == dy->w && dz->c >= dy->c && dz->dtype == YC_BF16,
             "yc_bn_bwd: shape mismatch");
  YC_REQUIRE(dz->ldc <= 2048, "yc_bn_bwd: more than 2048 channels");
  cudaStream_t st = (cudaStream_t)stream;
  const int n = dy->c;
  const TView dv = tv(dy), zv = tv(z);
  TView ov = tv(dz);
  float* S1 = workspace;          // [n]
  float* S2 = workspace + n;      // [n]
  float* coef = workspace + 2 * n;  // [3n]
  unsigned* ticket = reinterpret_cast<unsigned*>(workspace + 5 * n);  // zero between launches (the kernel resets it)
  float* P = workspace + 5 * n + 4;  // [grid][2][C8] block partials
  const int groups = (n + 7) / 8;
  const int threads = reduce_threads(groups);
  const int slots = threads / groups;
  const int grid = row_grid(dv.rows, slots, 2);
  // (the block that draws the last ticket also turns the two sums into the coefficients of the apply pass and adds the
  // scale / bias gradients: bn_bwd_coef_kernel's arithmetic, without its launch)
  YC_ACT_DISPATCH(activation, (bwd_reduce_kernel<0, A><<<grid, threads, reduce_smem(groups, threads), st>>>(
                                  dv, zv, mean, var, gamma, beta, S1, S2, P, ticket, scale_updates, bias_updates, coef, (float)dv.rows)));
  YC_LAUNCH_CHECK();
  ov.c = dy->c;
  const int agroups = (dz->ldc >= (long long)groups * 8) ? dz->ldc / 8 : groups;
  const int athreads = reduce_threads(agroups);
  YC_ACT_DISPATCH(activation, (bn_bwd_apply_kernel<A><<<row_grid(dv.rows, athreads / agroups), athreads, 0, st>>>(
                                  dv, zv, mean, var, gamma, beta, coef, ov, agroups)));
  YC_LAUNCH_CHECK();
  return YC_OK;
}

extern "C" int yc_bias_act_bwd(const yc_tensor* dy, const yc_tensor* y, int activation, float* bias_updates,
                               float* workspace, const yc_tensor* dz, void* stream) {
  YC_REQUIRE(ok_view(dy) && ok_view(dz) && (activation == YC_LINEAR || ok_view(y)) && (!bias_updates || workspace),
             "yc_bias_act_bwd: bad arguments (bias_updates needs a workspace of yc_bn_bwd_workspace_floats(c) floats)");
  YC_REQUIRE(activation == YC_LINEAR || activation == YC_LEAKY || activation == YC_RELU || activation == YC_LOGISTIC,
             "yc_bias_act_bwd: mish/swish need the saved pre-activation, which only batch-normalised layers keep");
  YC_REQUIRE(dz->n == dy->n && dz->h == dy->h && dz->w == dy->w && dz->c >= dy->c && dz->dtype == YC_BF16,
             "yc_bias_act_bwd: shape mismatch");
  cudaStream_t st = (cudaStream_t)stream;
  const int n = dy->c;
  const TView dv = tv(dy);
  const TView uv = activation == YC_LINEAR ? dv : tv(y);
  TView ov = tv(dz);
  ov.c = n;
  const int groups = (n + 7) / 8;
  if (bias_updates) {
    const int threads = reduce_threads(groups);
    const int slots = threads / groups;
    const int grid = row_grid(dv.rows, slots, 2);
    unsigned* ticket = reinterpret_cast<unsigned*>(workspace + 5 * n);  // same layout as yc_bn_bwd's workspace
    float* P = workspace + 5 * n + 4;
    YC_ACT_DISPATCH(activation, (bwd_reduce_kernel<1, A><<<grid, threads, reduce_smem(groups, threads), st>>>(
                                    dv, uv, nullptr, nullptr, nullptr, nullptr, bias_updates, nullptr, P, ticket)));
    YC_LAUNCH_CHECK();
  }
  const int agroups = (dz->ldc >= (long long)groups * 8) ? dz->ldc / 8 : groups;
  YC_ACT_DISPATCH(activation, (act_bwd_apply_kernel<A><<<ew_grid(dv.rows * agroups), 256, 0, st>>>(dv, uv, ov, agroups)));
  YC_LAUNCH_CHECK();
  return YC_OK;
}

// ---- yc_bias_act_bwd_noalloc: the same operator for callers that cannot allocate (CUDA-graph capture, serving loops) ----
// No counterpart in the reference.  The reduction's scratch (the ticket
// and one [2][ceil8(c)] partial per block) lives in this file-scope device array instead of a caller-allocated
// workspace: every GPU that loads the library holds YC_NOALLOC_SCRATCH_FLOATS * 4 bytes (2.4 MB) until exit.
// All calls on a device share the array: calls that overlap on the GPU (two streams) race.
#define YC_NOALLOC_MAX_C 1024                       /* channels */
#define YC_NOALLOC_MAX_BLOCKS 296                   /* blocks of the reduction grid (148 SMs x YC_BN_MINB) */
#define YC_NOALLOC_SCRATCH_FLOATS (64 + (size_t)YC_NOALLOC_MAX_BLOCKS * 2 * YC_NOALLOC_MAX_C)
static __device__ __align__(256) float g_noalloc_scratch[YC_NOALLOC_SCRATCH_FLOATS];

// In host code `g_noalloc_scratch` is only a handle; the device address comes from cudaGetSymbolAddress, asked once per
// device with that device current.
static float* noalloc_scratch_ptr() {
  constexpr int kMaxDev = 64;
  static float* ptrs[kMaxDev] = {};
  static std::mutex mu;
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDev) {
    set_error("yc_bias_act_bwd_noalloc: no current CUDA device");
    return nullptr;
  }
  std::lock_guard<std::mutex> g(mu);
  if (!ptrs[dev]) {
    float* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) {
      set_error("yc_bias_act_bwd_noalloc: cudaGetSymbolAddress(g_noalloc_scratch) on device %d -> %s", dev,
                cudaGetErrorString(e));
      return nullptr;
    }
    ptrs[dev] = p;
  }
  return ptrs[dev];
}

extern "C" size_t yc_bias_act_bwd_noalloc_max_channels(void) { return YC_NOALLOC_MAX_C; }

extern "C" int yc_bias_act_bwd_noalloc(const yc_tensor* dy, const yc_tensor* y, int activation, float* bias_updates,
                                       const yc_tensor* dz, void* stream) {
  YC_REQUIRE(ok_view(dy) && ok_view(dz) && (activation == YC_LINEAR || ok_view(y)) && bias_updates,
             "yc_bias_act_bwd_noalloc: bad arguments");
  YC_REQUIRE(activation == YC_LINEAR || activation == YC_LEAKY || activation == YC_RELU || activation == YC_LOGISTIC,
             "yc_bias_act_bwd_noalloc: mish/swish need the saved pre-activation, which only batch-normalised layers keep");
  YC_REQUIRE(dz->n == dy->n && dz->h == dy->h && dz->w == dy->w && dz->c >= dy->c && dz->dtype == YC_BF16,
             "yc_bias_act_bwd_noalloc: shape mismatch");
  YC_REQUIRE(dy->c <= YC_NOALLOC_MAX_C, "yc_bias_act_bwd_noalloc: %d channels exceed the static scratch's maximum of %d",
             dy->c, YC_NOALLOC_MAX_C);
  cudaStream_t st = (cudaStream_t)stream;
  const int n = dy->c;
  const TView dv = tv(dy);
  const TView uv = activation == YC_LINEAR ? dv : tv(y);
  TView ov = tv(dz);
  ov.c = n;
  const int groups = (n + 7) / 8;
  const int threads = reduce_threads(groups);
  const int slots = threads / groups;
  const int grid = row_grid(dv.rows, slots, 2);
  YC_REQUIRE(grid <= YC_NOALLOC_MAX_BLOCKS,
             "yc_bias_act_bwd_noalloc: a reduction grid of %d blocks exceeds the static scratch's maximum of %d", grid,
             YC_NOALLOC_MAX_BLOCKS);
  float* scratch = noalloc_scratch_ptr();
  if (!scratch) return YC_ERR_CUDA;
  unsigned* ticket = reinterpret_cast<unsigned*>(scratch);
  float* P = scratch + 64;
  // the array keeps the last call's contents: the ticket must start at zero (a call that was aborted would leave it
  // set); the partials are overwritten before they are read
  YC_CUDA(cudaMemsetAsync(ticket, 0, sizeof(unsigned), st));
  YC_ACT_DISPATCH(activation, (bwd_reduce_kernel<1, A><<<grid, threads, reduce_smem(groups, threads), st>>>(
                                  dv, uv, nullptr, nullptr, nullptr, nullptr, bias_updates, nullptr, P, ticket)));
  YC_LAUNCH_CHECK();
  const int agroups = (dz->ldc >= (long long)groups * 8) ? dz->ldc / 8 : groups;
  YC_ACT_DISPATCH(activation, (act_bwd_apply_kernel<A><<<ew_grid(dv.rows * agroups), 256, 0, st>>>(dv, uv, ov, agroups)));
  YC_LAUNCH_CHECK();
  return YC_OK;
}
