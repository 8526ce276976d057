// Shared host-side plumbing for libyc_b200: status codes, last-error string, launch counter.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdint.h>

#include "../../include/yc_b200.h"

namespace yc {

void set_error(const char* fmt, ...);
void count_launch(int n = 1);

#define YC_CUDA(call)                                                                          \
  do {                                                                                         \
    cudaError_t _e = (call);                                                                   \
    if (_e != cudaSuccess) {                                                                   \
      yc::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e));     \
      return YC_ERR_CUDA;                                                                      \
    }                                                                                          \
  } while (0)

#define YC_LAUNCH_CHECK()                                                                      \
  do {                                                                                         \
    cudaError_t _e = cudaPeekAtLastError();                                                    \
    if (_e != cudaSuccess) {                                                                   \
      yc::set_error("%s:%d: kernel launch -> %s", __FILE__, __LINE__, cudaGetErrorString(_e)); \
      return YC_ERR_CUDA;                                                                      \
    }                                                                                          \
    yc::count_launch();                                                                        \
  } while (0)

#define YC_REQUIRE(cond, ...)          \
  do {                                 \
    if (!(cond)) {                     \
      yc::set_error(__VA_ARGS__);      \
      return YC_ERR_ARG;               \
    }                                  \
  } while (0)

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
static inline long long ceil_div_ll(long long a, long long b) { return (a + b - 1) / b; }
static inline int round_up(int a, int b) { return ceil_div(a, b) * b; }

int sm_count();

// one entry of the whole-network SGD update (train_layers.cu: sgd_multi_kernel; built by yc_network_update)
struct SgdItem {
  float* w;               // fp32 master: OIHW weights, or the bias / scale vector
  float* dw;              // gradient accumulator (packed [n][tap][ldw] for weights)
  __nv_bfloat16* wp_f;    // bf16 forward packing [n][tap][in_c_p] (or null)
  __nv_bfloat16* wp_d;    // bf16 data-gradient packing [c][flipped tap][out_c_p] (or null)
  int out_c, in_c, taps, ldw, in_c_p, out_c_p;  // taps == 0: plain vector of out_c elements
  int first_block, blocks;
  float* m;               // Adam moments, laid out like dw (null: SGD)
  float* v;
};
// UpdateConvolutionalLayerGpu's rule (convolutional_kernels.cu:865-922): SGD with momentum and decay, or Adam
// (adam_update_gpu, blas_kernels.cu:99-134) when adam != 0: rate is the learning rate NOT divided by the batch, t the
// iteration (l->t, network_kernels.cu:229).
struct UpdRule {
  float lr_over_batch, decay_batch, momentum;
  int adam, t;
  float B1, B2, eps, rate;
};
int sgd_item_blocks(const SgdItem& it);
int sgd_update_multi(const SgdItem* items_dev, int n_items, int total_blocks, const UpdRule& rule, cudaStream_t st);

// ---- device-side activation math shared by the conv epilogue and the elementwise kernels ------
// Formulas follow the reference (activations.h:80-121, activation_kernels.cu:56-79), fp32.
// MUFU wrappers with flush-to-zero: one SASS instruction each (the non-ftz intrinsics add range-scaling code).
__device__ __forceinline__ float ex2_ftz(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_ftz(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float act_logistic(float x) {
  // 1/(1+e^-x); the clamp keeps e^-x finite so that rcp never sees inf (activations.h:80)
  return rcp_ftz(1.f + ex2_ftz(fmaxf(x, -80.f) * -1.4426950408889634f));
}
__device__ __forceinline__ float act_mish(float x) {
  // x * tanh(softplus(x)); with n = e^x: tanh(log(1+n)) = n(n+2) / (n(n+2) + 2) = 1 - 2/q, q = n(n+2) + 2, so
  // mish(x) = x - 2x/q = x + x*r with r = 1/(-q/2).  Written on h = n/2 = 2^(x*log2e - 1) every step is one fma:
  // -(n+2) = -2h - 2, -q/2 = h*(-(n+2)) - 1.  No clamp is needed: when h (or q) overflows, r = -0 and the result is
  // x, which is what the reference's softplus threshold at +20 gives (activations.h:106-113); for x -> -inf, r -> -1
  // and the result -> 0 (absolute error <= |x| * 2^-23 where the reference has x*e^x: far below one bf16 ulp of any
  // stored value).
  const float h = ex2_ftz(fmaf(x, 1.4426950408889634f, -1.f));
  const float m = fmaf(h, -2.f, -2.f);
  return fmaf(x, rcp_ftz(fmaf(h, m, -1.f)), x);
}
// Packed fp32x2 arithmetic (sm_100 FFMA2 / FMUL2 / FADD2: one issue slot for two lanes of work).  The mish epilogue is
// issue-bound (ncu: 66 % of the issue slots, XU pipe 53 %), so two outputs per instruction is what counts.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float a, float b) {
  f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
  f32x2 d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
  f32x2 d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ float apply_act(float x, int act) {
  switch (act) {
    case YC_LINEAR: return x;
    case YC_LEAKY: return fmaxf(x, 0.1f * x);
    case YC_MISH: return act_mish(x);
    case YC_LOGISTIC: return act_logistic(x);
    case YC_SWISH: return x * act_logistic(x);
    case YC_RELU: return x > 0.f ? x : 0.f;
  }
  return x;
}

}  // namespace yc
