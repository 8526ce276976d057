// Convolution weight gradient as ONE implicit-GEMM kernel for sm_100a.
//
// Replaces the wgrad half of BackwardConvolutionalLayerGpu (convolutional_kernels.cu:555-815: per image
// im2col_gpu_ext + gemm_ongpu(0,1,...) accumulating into weight_updates_gpu; CPU twin
// convolutional_layer.cpp:1338-1358), with no im2col workspace and no per-image loop.
//
// GEMM view, per filter tap t = (r, s):
//   dW[n][t][c] += sum over output pixels m of  dY[m][n] * X[pixel(m) + tap t][c]
//   D[M = 128 output channels][N = block_n input channels] accumulates over K = output pixels.
// Both operands are "MN-major" for the tensor core: the contiguous axis of NHWC is the channel axis, which is the
// GEMM's M (for dY) and N (for X), while K = pixels is the strided axis.  TMA drops 64-pixel x 64-channel boxes
// (128-byte rows, 128B swizzle) into shared memory — dY through a tiled map, X through an im2col-mode map that
// applies the tap offset, stride, dilation and zero padding — and tcgen05.mma reads them through MN-major
// descriptors (instruction-descriptor bits 15/16 = 1).  K is split across CTAs (split-K over pixel ranges) so that
// small filters still fill 148 SMs.  The partial sums of the splits go through a scratch buffer and are added to the
// fp32 gradient in split order by the last unit to finish (no float atomics: the step is reproducible bit for bit);
// `+=` into dw is the reference's accumulation semantics.
//
// CTA = 320 threads, persistent over work units (tile x K-split):
//   warp 0 / lane 0 : TMA producer, warp 1 / lane 0 : tcgen05.mma issuer, warps 2..9 : epilogue (TMEM -> red.global)
#include <cuda.h>
#include <algorithm>

#include "ptx.cuh"
#include "yc_common.cuh"

namespace yc {

constexpr int kWgThreads = 64 + 32 * 8;
constexpr int kWgKBlk = 64;       // pixels per pipeline stage
constexpr int kWgMaxStages = 8;
constexpr int kBoxBytes = kWgKBlk * 128;  // one 64-pixel x 64-channel bf16 box

struct WgradParams {
  int M;            // output pixels B*Ho*Wo
  int N_out, C_in;  // real channel counts
  int taps, ksize, dil, stride, pad;
  int Ho, Wo;
  int block_n;      // input channels per tile (multiple of 64)
  int mma_n;        // N of the MMA instruction (multiple of 16, <= block_n)
  int n_tiles;      // ceil(N_out / 128)
  int c_tiles;      // ceil(C_in / block_n)
  int tap_group;    // taps per work unit (1, 3 = one kernel row, or 9): they share the unit's dY boxes and accumulate into
                    // tap_group independent TMEM regions, so the MMA stream alternates between accumulators
  int tgroups;      // taps / tap_group
  int acc_cols;     // TMEM columns per tap accumulator (mma_n rounded up to 32)
  int nbuf;         // accumulator sets in TMEM: 2 (the epilogue of unit i overlaps the main loop of unit i + 1) or 1
  int tiles;        // n_tiles * tgroups * c_tiles
  int splits, kb_total, kb_per_split;
  int stages;
  int a_boxes, b_boxes;
  uint32_t tmem_cols;
  float* dw;        // [N_out][taps][ldw] fp32
  int ldw;
  // split-K without atomics (splits > 1): every (tile, split) unit stores its partial sums to partial[split]; a second,
  // fully coalesced kernel (wgrad_reduce_kernel) adds the splits in order into dw.  The sum order is fixed whatever the
  // scheduling, so two runs of a training step give bit-identical gradients.  (A first version let the last-arriving
  // epilogue warp do the ordered sum: row-per-lane reads of the partials made wgrad 2.8x slower on yolov4-tiny.)
  float* partial;             // [splits][N_out * taps * ldw] or nullptr (splits == 1: a single writer per element)
  long long partial_stride;   // floats per split
};

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}

// MN-major shared-memory descriptor, 128B swizzle: rows (K index) of 128 B = 64 MN-elements; 8-row groups SBO
// apart; the next 64 MN-elements LBO apart (cute::UMMA make_umma_desc<Major::MN>: ((8,n),(8,k)):((1,LBO),(8,SBO))
// in 16-byte units).
__device__ __forceinline__ uint64_t make_desc_mn_sw128(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;  // SWIZZLE_128B
  return d;
}

__global__ void __launch_bounds__(kWgThreads, 1)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap tmap_dy, const __grid_constant__ CUtensorMap tmap_x,
                  const WgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t a_bytes = p.a_boxes * kBoxBytes;
  const uint32_t stage_bytes = (p.a_boxes + p.tap_group * p.b_boxes) * kBoxBytes;
  const uint32_t bar_base = smem_base + p.stages * stage_bytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kWgMaxStages + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * kWgMaxStages + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * kWgMaxStages + 2 + a); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * kWgMaxStages + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    ptx::prefetch_tmap(&tmap_dy);
    ptx::prefetch_tmap(&tmap_x);
    for (int s = 0; s < p.stages; ++s) {
      ptx::mbar_init(full_bar(s), 1);
      ptx::mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      ptx::mbar_init(tfull_bar(a), 1);
      ptx::mbar_init(tempty_bar(a), 8);  // one arrival per epilogue warp
    }
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_slot, p.tmem_cols);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  const int units = p.tiles * p.splits;

  if (warp == 0) {
    // ===================== TMA producer =====================
    // whole warp converged, one elected lane issues: coordinates / descriptors stay in uniform registers (the
    // single-lane form pays an R2UR chain per TMA / MMA instruction; see conv_fprop_sm100.cu)
    const bool leader = ptx::elect_one();
    int stage = 0;
    uint32_t phase = 0;
    const int hw = p.Ho * p.Wo;
    const uint32_t tx_bytes = stage_bytes;
    for (int u = blockIdx.x; u < units; u += gridDim.x) {
      const int tile = u % p.tiles, split = u / p.tiles;
      const int ct = tile % p.c_tiles;
      const int tap0 = ((tile / p.c_tiles) % p.tgroups) * p.tap_group;
      const int nt = tile / (p.c_tiles * p.tgroups);
      const int kb0 = split * p.kb_per_split;
      const int kb1 = min(p.kb_total, kb0 + p.kb_per_split);
      for (int kb = kb0; kb < kb1; ++kb) {
        const int m0 = kb * kWgKBlk;
        const int img = m0 / hw;
        const int rem = m0 - img * hw;
        const int ho = rem / p.Wo;
        const int base_h = ho * p.stride - p.pad;
        const int base_w = (rem - ho * p.Wo) * p.stride - p.pad;
        ptx::mbar_wait(empty_bar(stage), phase ^ 1u);
        const uint32_t sa = smem_base + stage * stage_bytes;
        if (leader) {
          ptx::mbar_arrive_expect_tx(full_bar(stage), tx_bytes);
          for (int j = 0; j < p.a_boxes; ++j)
            ptx::tma_load_2d(sa + j * kBoxBytes, &tmap_dy, full_bar(stage), nt * 128 + j * 64, m0);
          for (int t = 0; t < p.tap_group; ++t) {  // the X boxes of every tap of the group (the dY boxes are shared)
            const int tap = tap0 + t;
            const int r = tap / p.ksize, s = tap - r * p.ksize;
            for (int j = 0; j < p.b_boxes; ++j)
              ptx::tma_load_im2col_4d(sa + a_bytes + (t * p.b_boxes + j) * kBoxBytes, &tmap_x, full_bar(stage),
                                      ct * p.block_n + j * 64, base_w, base_h, img, (uint16_t)(s * p.dil), (uint16_t)(r * p.dil));
          }
        }
        if (++stage == p.stages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    const bool leader = ptx::elect_one();
    // both operands MN-major: instruction-descriptor bits 15 (A) and 16 (B)
    const uint32_t idesc = ptx::make_idesc_bf16(128, p.mma_n) | (1u << 15) | (1u << 16);
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int u = blockIdx.x; u < units; u += gridDim.x, ++it) {
      const int split = u / p.tiles;
      const int kb0 = split * p.kb_per_split;
      const int kb1 = min(p.kb_total, kb0 + p.kb_per_split);
      const int acc = p.nbuf == 2 ? (it & 1) : 0;
      const uint32_t acc_phase = (p.nbuf == 2 ? (it >> 1) : it) & 1;
      ptx::mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
      ptx::tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * (p.tap_group * p.acc_cols);
      const uint32_t b_tap_bytes = p.b_boxes * kBoxBytes;
      for (int kb = kb0; kb < kb1; ++kb) {
        ptx::mbar_wait(full_bar(stage), phase);
        ptx::tc_fence_after();
        const uint32_t sa = smem_base + stage * stage_bytes;
        const uint64_t adesc = make_desc_mn_sw128(sa, kBoxBytes, 1024);
        const uint64_t bdesc = make_desc_mn_sw128(sa + a_bytes, kBoxBytes, 1024);
        if (leader) {
#pragma unroll
          for (int k = 0; k < kWgKBlk / 16; ++k) {
            // advancing K by 16 pixels = 16 rows of 128 B = 2048 B -> +128 in the (addr >> 4) field; the taps of the group
            // go to independent accumulators, so consecutive MMAs do not depend on each other
            for (int t = 0; t < p.tap_group; ++t)
              ptx::mma_bf16_ss(d_tmem + t * p.acc_cols, adesc + 128u * k, bdesc + (uint64_t)((t * b_tap_bytes) >> 4) + 128u * k,
                               idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          ptx::mma_commit(empty_bar(stage));
          if (kb == kb1 - 1) ptx::mma_commit(tfull_bar(acc));
        }
        __syncwarp();
        if (++stage == p.stages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp >= 2) {
    // ===================== epilogue: TMEM -> red.global.add =====================
    const int quad = warp & 3;
    const int half = (warp - 2) >> 2;  // 0/1: alternate 32-column groups
    int it = 0;
    for (int u = blockIdx.x; u < units; u += gridDim.x, ++it) {
      const int tile = u % p.tiles, split = u / p.tiles;
      const int ct = tile % p.c_tiles;
      const int tap0 = ((tile / p.c_tiles) % p.tgroups) * p.tap_group;
      const int nt = tile / (p.c_tiles * p.tgroups);
      const int acc = p.nbuf == 2 ? (it & 1) : 0;
      const uint32_t acc_phase = (p.nbuf == 2 ? (it >> 1) : it) & 1;
      const int n = nt * 128 + quad * 32 + lane;
      const bool row_ok = n < p.N_out;
      const int groups = p.mma_n >> 5;  // whole 32-column groups
      const int tail16 = (p.mma_n & 31) ? 1 : 0;
      ptx::mbar_wait(tfull_bar(acc), acc_phase);
      ptx::tc_fence_after();
      for (int tt = 0; tt < p.tap_group; ++tt) {
      const int tap = tap0 + tt;
      const uint32_t t_row = tmem_base + ((uint32_t)(quad * 32) << 16) + acc * (p.tap_group * p.acc_cols) + tt * p.acc_cols;
      float* rowp = p.dw + ((long long)n * p.taps + tap) * p.ldw + ct * p.block_n;
      for (int g = half; g < groups + tail16; g += 2) {
        uint32_t rr[32];
        const bool is_tail = g == groups;
        if (!is_tail) {
          ptx::tmem_ld_32x32b_x32(t_row + g * 32, rr);
        } else {
          uint32_t r16[16];
          ptx::tmem_ld_32x32b_x16(t_row + g * 32, r16);
#pragma unroll
          for (int j = 0; j < 16; ++j) rr[j] = r16[j];
#pragma unroll
          for (int j = 16; j < 32; ++j) rr[j] = 0u;
        }
        ptx::tmem_ld_wait();
        if (row_ok) {
          const int cbase = ct * p.block_n + g * 32;
          const int lim = is_tail ? 16 : 32;
          float* dst = p.partial ? p.partial + (long long)split * p.partial_stride + (rowp - p.dw) : rowp;
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            if (j < lim && cbase + j < p.ldw) {
              if (p.partial)
                *reinterpret_cast<float4*>(dst + g * 32 + j) = make_float4(__uint_as_float(rr[j]), __uint_as_float(rr[j + 1]),
                                                                           __uint_as_float(rr[j + 2]), __uint_as_float(rr[j + 3]));
              else  // one split: this unit is the only writer of these elements in this launch
                red_add_v4(dst + g * 32 + j, __uint_as_float(rr[j]), __uint_as_float(rr[j + 1]), __uint_as_float(rr[j + 2]),
                           __uint_as_float(rr[j + 3]));
            }
          }
        }
      }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(tempty_bar(acc));
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

// dw[i] += partial[0][i] + partial[1][i] + ... in split order; 16-byte accesses, one pass
__global__ void __launch_bounds__(256) wgrad_reduce_kernel(const float4* __restrict__ partial, long long stride4, int splits,
                                                           float4* __restrict__ dw, long long n4) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 a = dw[i];
    // split order, eight partials requested before the first is added (same order of additions as a plain loop; the
    // plain loop's chain of dependent loads made this pass 0.65 ms per yolov4-tiny step for 0.5 GB)
    int s = 0;
    for (; s + 8 <= splits; s += 8) {
      float4 v[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = __ldcs(partial + (long long)(s + k) * stride4 + i);
#pragma unroll
      for (int k = 0; k < 8; ++k) { a.x += v[k].x; a.y += v[k].y; a.z += v[k].z; a.w += v[k].w; }
    }
    for (; s < splits; ++s) {
      const float4 v = __ldcs(partial + (long long)s * stride4 + i);
      a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
    }
    dw[i] = a;
  }
}

// The same for small filters with MANY splits (a 64 x 9 x 64 gradient has 99 of them, the stem 296: the plain kernel is one
// thread per element walking all splits, i.e. latency-bound chains on a handful of blocks — 25-64 us per layer for a few
// MB).  Block = 32 elements x 8 split groups: group g adds the splits [g*per, (g+1)*per) in order, the eight group sums
// are then added in group order on top of dw.  A fixed order again, so the step stays reproducible bit for bit.
__global__ void __launch_bounds__(256) wgrad_reduce_grouped_kernel(const float4* __restrict__ partial, long long stride4, int splits,
                                                                   float4* __restrict__ dw, long long n4) {
  __shared__ float4 sh[8][32];
  const int e = threadIdx.x & 31, g = threadIdx.x >> 5;
  const int per = (splits + 7) / 8;
  const int s0 = g * per, s1 = min(splits, s0 + per);
  for (long long base = (long long)blockIdx.x * 32; base < n4; base += (long long)gridDim.x * 32) {
    const long long i = base + e;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < n4) {
      int s = s0;
      for (; s + 4 <= s1; s += 4) {
        float4 v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = __ldcs(partial + (long long)(s + k) * stride4 + i);
#pragma unroll
        for (int k = 0; k < 4; ++k) { a.x += v[k].x; a.y += v[k].y; a.z += v[k].z; a.w += v[k].w; }
      }
      for (; s < s1; ++s) {
        const float4 v = __ldcs(partial + (long long)s * stride4 + i);
        a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
      }
    }
    sh[g][e] = a;
    __syncthreads();
    if (g == 0 && i < n4) {
      float4 t = dw[i];
#pragma unroll
      for (int k = 0; k < 8; ++k) { t.x += sh[k][e].x; t.y += sh[k][e].y; t.z += sh[k][e].z; t.w += sh[k][e].w; }
      dw[i] = t;
    }
    __syncthreads();
  }
}

// dW packed fp32 [n][tap][ldw] -> OIHW fp32 (the reference's weight_updates layout), for pull / inspection
__global__ void wgrad_unpack_kernel(const float* __restrict__ dwp, int out_c, int in_c, int taps, int ldw,
                                    float* __restrict__ dw_oihw) {
  const long long total = (long long)out_c * in_c * taps;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int tap = (int)(i % taps);
    const int c = (int)((i / taps) % in_c);
    const int n = (int)(i / ((long long)taps * in_c));
    dw_oihw[i] = dwp[((long long)n * taps + tap) * ldw + c];
  }
}
__global__ void wgrad_pack_kernel(const float* __restrict__ dw_oihw, int out_c, int in_c, int taps, int ldw,
                                  float* __restrict__ dwp) {
  const long long total = (long long)out_c * taps * ldw;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % ldw);
    const int tap = (int)((i / ldw) % taps);
    const int n = (int)(i / ((long long)ldw * taps));
    dwp[i] = c < in_c ? dw_oihw[((long long)n * in_c + c) * taps + tap] : 0.f;
  }
}

}  // namespace yc

using namespace yc;

struct yc_wgrad_plan {
  CUtensorMap tmap_dy, tmap_x;
  WgradParams p;
  int grid = 0;
  size_t smem = 0;
  float* partial = nullptr;      // plan-owned split-K scratch (splits > 1)
};

namespace {
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn w_encode_tiled = nullptr;
EncodeIm2colFn w_encode_im2col = nullptr;
int load_entry_points() {
  if (w_encode_tiled && w_encode_im2col) return YC_OK;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qr;
  YC_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr));
  YC_REQUIRE(fn && qr == cudaDriverEntryPointSuccess, "cuTensorMapEncodeTiled not available");
  w_encode_tiled = (EncodeTiledFn)fn;
  YC_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &fn, cudaEnableDefault, &qr));
  YC_REQUIRE(fn && qr == cudaDriverEntryPointSuccess, "cuTensorMapEncodeIm2col not available");
  w_encode_im2col = (EncodeIm2colFn)fn;
  return YC_OK;
}
}  // namespace

// plan-time switch (A/B measurements, parity coverage of both schedules): 0 = one tap per work unit as in round 1
static int g_wgrad_tap_groups = 1;
extern "C" int yc_conv_wgrad_set_tap_groups(int on) {
  const int old = g_wgrad_tap_groups;
  g_wgrad_tap_groups = on ? 1 : 0;
  return old;
}

extern "C" int yc_conv_wgrad_ld(int in_c) { return round_up(in_c, 32); }
extern "C" size_t yc_conv_wgrad_bytes(int out_c, int in_c, int ksize) {
  return (size_t)out_c * ksize * ksize * yc_conv_wgrad_ld(in_c) * 4;
}

extern "C" int yc_conv_wgrad_plan_create(yc_wgrad_plan** out_plan, const yc_conv_desc* d, const yc_tensor* x,
                                         const yc_tensor* dy, int out_c, float* dw_packed) {
  YC_REQUIRE(out_plan && d && x && dy && dw_packed && out_c > 0, "yc_conv_wgrad_plan_create: null argument");
  YC_REQUIRE(x->dtype == YC_BF16 && dy->dtype == YC_BF16, "wgrad operands must be bf16 NHWC");
  YC_REQUIRE(x->ldc % 8 == 0 && ((uintptr_t)x->ptr & 15) == 0 && dy->ldc % 8 == 0 && ((uintptr_t)dy->ptr & 15) == 0,
             "wgrad views must be 16-byte aligned");
  const int eff = d->dilation * (d->ksize - 1) + 1;
  const int Ho = (x->h + 2 * d->pad - eff) / d->stride + 1;
  const int Wo = (x->w + 2 * d->pad - eff) / d->stride + 1;
  YC_REQUIRE(dy->n == x->n && dy->h == Ho && dy->w == Wo && dy->c >= out_c, "wgrad: dy shape mismatch");
  YC_REQUIRE((long long)x->n * Ho * Wo < (1LL << 31), "M too large");
  int rc = load_entry_points();
  if (rc) return rc;

  yc_wgrad_plan* pl = new yc_wgrad_plan();
  WgradParams& p = pl->p;
  p.M = x->n * Ho * Wo;
  p.N_out = out_c;
  p.C_in = x->c;
  p.taps = d->ksize * d->ksize;
  p.ksize = d->ksize; p.dil = d->dilation; p.stride = d->stride; p.pad = d->pad;
  p.Ho = Ho; p.Wo = Wo;
  const int C = x->c;
  if (C <= 256) p.block_n = round_up(C, 64);
  else if (C % 256 == 0) p.block_n = 256;
  else if (C % 192 == 0) p.block_n = 192;
  else p.block_n = 128;
  p.mma_n = C < 64 ? round_up(C, 16) : p.block_n;
  p.n_tiles = ceil_div(out_c, 128);
  p.c_tiles = ceil_div(C, p.block_n);
  // taps per work unit: a whole 3x3 filter or one kernel row share the unit's dY boxes (one third / one ninth of the dY
  // traffic and of the TMA instructions, and 3 or 9 independent accumulators for the MMA stream to alternate between) as
  // far as the accumulators fit the 512 TMEM columns — twice (double-buffered against the epilogue) or, failing that, once
  p.acc_cols = round_up(p.mma_n, 32);
  p.tap_group = 1;
  p.nbuf = 2;
  if (p.taps == 9 && g_wgrad_tap_groups) {
    if (2 * 9 * p.acc_cols <= 512) p.tap_group = 9;
    else if (2 * 3 * p.acc_cols <= 512) p.tap_group = 3;
    else if (3 * p.acc_cols <= 512) { p.tap_group = 3; p.nbuf = 1; }
  }
  p.tgroups = p.taps / p.tap_group;
  p.tiles = p.n_tiles * p.tgroups * p.c_tiles;
  p.kb_total = ceil_div(p.M, kWgKBlk);
  const int sms = sm_count();
  int splits = std::max(1, ceil_div(2 * sms, p.tiles));
  splits = std::min(splits, std::max(1, p.kb_total / 8));
  p.kb_per_split = ceil_div(p.kb_total, splits);
  p.splits = ceil_div(p.kb_total, p.kb_per_split);
  // M = 128 always.  out_c <= 64: only the first 64-channel box of dz is loaded; the MMA's second 64-row group (LBO =
  // one box further) then reads the B region of the stage -> accumulator rows 64..127 hold garbage that the epilogue
  // never stores (row_ok).  Saves one third of the TMA rows of the small layers, which are TMA-row-rate bound.
  p.a_boxes = out_c > 64 ? 2 : 1;
  p.b_boxes = p.block_n / 64;
  const uint32_t stage_bytes = (p.a_boxes + p.tap_group * p.b_boxes) * kBoxBytes;
  p.stages = std::min<int>(kWgMaxStages, (227 * 1024 - 1024 - 256) / stage_bytes);
  YC_REQUIRE(p.stages >= 2, "wgrad tile does not fit shared memory");
  pl->smem = (size_t)p.stages * stage_bytes + 1024 + 256;
  uint32_t cols = 32;
  while (cols < (uint32_t)(p.nbuf * p.tap_group * p.acc_cols)) cols <<= 1;
  p.tmem_cols = cols;
  p.dw = dw_packed;
  p.ldw = yc_conv_wgrad_ld(C);
  p.partial = nullptr;
  p.partial_stride = (long long)out_c * p.taps * p.ldw;
  if (p.splits > 1) {
    // zero-filled once: the epilogue rewrites the same elements at every launch, the padding columns it never writes
    // stay zero
    cudaError_t e = cudaMalloc(&pl->partial, (size_t)p.splits * p.partial_stride * sizeof(float));
    if (e == cudaSuccess) e = cudaMemset(pl->partial, 0, (size_t)p.splits * p.partial_stride * sizeof(float));
    if (e == cudaSuccess) e = cudaStreamSynchronize(cudaStreamLegacy);
    if (e != cudaSuccess) {
      set_error("yc_conv_wgrad_plan_create: split-K scratch: %s", cudaGetErrorString(e));
      yc_conv_wgrad_plan_destroy(pl);
      return YC_ERR_ALLOC;
    }
    p.partial = pl->partial;
  }

  CUresult cr;
  {
    cuuint64_t dims[2] = {(cuuint64_t)dy->c, (cuuint64_t)p.M};
    cuuint64_t strides[1] = {(cuuint64_t)dy->ldc * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)kWgKBlk};
    cuuint32_t estr[2] = {1, 1};
    cr = w_encode_tiled(&pl->tmap_dy, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dy->ptr, dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  if (cr != CUDA_SUCCESS) {
    set_error("yc_conv_wgrad_plan_create: dy tensor map encode failed (CUresult %d)", (int)cr);
    delete pl;
    return YC_ERR_CUDA;
  }
  {
    cuuint64_t dims[4] = {(cuuint64_t)x->c, (cuuint64_t)x->w, (cuuint64_t)x->h, (cuuint64_t)x->n};
    cuuint64_t strides[3] = {(cuuint64_t)x->ldc * 2, (cuuint64_t)x->ldc * 2 * x->w, (cuuint64_t)x->ldc * 2 * x->w * x->h};
    int lower[2] = {-d->pad, -d->pad};
    int upper[2] = {d->pad - (d->ksize - 1) * d->dilation, d->pad - (d->ksize - 1) * d->dilation};
    cuuint32_t estr[4] = {1, (cuuint32_t)d->stride, (cuuint32_t)d->stride, 1};
    cr = w_encode_im2col(&pl->tmap_x, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, x->ptr, dims, strides, lower, upper, 64,
                         (cuuint32_t)kWgKBlk, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  if (cr != CUDA_SUCCESS) {
    set_error("yc_conv_wgrad_plan_create: x tensor map encode failed (CUresult %d, c=%d ldc=%d)", (int)cr, x->c, x->ldc);
    delete pl;
    return YC_ERR_CUDA;
  }
  pl->grid = std::min(p.tiles * p.splits, sms);
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(conv_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(max dynamic smem) failed: %s", cudaGetErrorString(e));
      delete pl;
      return YC_ERR_CUDA;
    }
    attr_set = true;
  }
  *out_plan = pl;
  return YC_OK;
}

extern "C" int yc_conv_wgrad_plan_destroy(yc_wgrad_plan* pl) {
  if (!pl) return YC_OK;
  if (pl->partial) cudaFree(pl->partial);
  delete pl;
  return YC_OK;
}

extern "C" int yc_conv_wgrad(const yc_wgrad_plan* pl, void* stream) {
  YC_REQUIRE(pl, "yc_conv_wgrad: null plan");
  conv_wgrad_kernel<<<pl->grid, kWgThreads, pl->smem, (cudaStream_t)stream>>>(pl->tmap_dy, pl->tmap_x, pl->p);
  YC_LAUNCH_CHECK();
  if (pl->p.partial) {  // ldw is a multiple of 32 floats: the gradient is a whole number of float4
    const long long n4 = pl->p.partial_stride / 4;
    if (pl->p.splits >= 16) {  // many splits of a small filter: split groups in parallel (see the kernel)
      const int blocks = (int)std::max<long long>(1, std::min<long long>(ceil_div_ll(n4, 32), (long long)sm_count() * 8));
      wgrad_reduce_grouped_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const float4*>(pl->p.partial), n4,
                                                                           pl->p.splits, reinterpret_cast<float4*>(pl->p.dw), n4);
      YC_LAUNCH_CHECK();
      return YC_OK;
    }
    // small filters: 64-thread blocks so that even a 64 x 9 x 64 gradient (9216 float4) spreads over all SMs
    const int rthreads = n4 < 148LL * 256 ? 64 : 256;
    const int blocks = (int)std::max<long long>(1, std::min<long long>(ceil_div_ll(n4, rthreads), (long long)sm_count() * 8));
    wgrad_reduce_kernel<<<blocks, rthreads, 0, (cudaStream_t)stream>>>(reinterpret_cast<const float4*>(pl->p.partial), n4, pl->p.splits,
                                                                   reinterpret_cast<float4*>(pl->p.dw), n4);
    YC_LAUNCH_CHECK();
  }
  return YC_OK;
}

extern "C" int yc_conv_wgrad_plan_info(const yc_wgrad_plan* pl, int what) {
  if (!pl) return -1;
  switch (what) {
    case 0: return pl->p.block_n;
    case 1: return pl->p.splits;
    case 2: return pl->p.stages;
    case 3: return pl->p.tiles;
    case 4: return pl->grid;
    case 5: return pl->p.kb_per_split;
    case 6: return pl->p.tap_group;
    case 7: return pl->p.nbuf;
  }
  return -1;
}

extern "C" int yc_conv_wgrad_unpack(const float* dw_packed, int out_c, int in_c, int ksize, float* dw_oihw, void* stream) {
  YC_REQUIRE(dw_packed && dw_oihw && out_c > 0 && in_c > 0 && ksize > 0, "yc_conv_wgrad_unpack: bad arguments");
  const long long total = (long long)out_c * in_c * ksize * ksize;
  const int blocks = (int)std::min<long long>(ceil_div_ll(total, 256), 148 * 8);
  wgrad_unpack_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(dw_packed, out_c, in_c, ksize * ksize, yc_conv_wgrad_ld(in_c), dw_oihw);
  YC_LAUNCH_CHECK();
  return YC_OK;
}
extern "C" int yc_conv_wgrad_pack(const float* dw_oihw, int out_c, int in_c, int ksize, float* dw_packed, void* stream) {
  YC_REQUIRE(dw_packed && dw_oihw && out_c > 0 && in_c > 0 && ksize > 0, "yc_conv_wgrad_pack: bad arguments");
  const long long total = (long long)out_c * ksize * ksize * yc_conv_wgrad_ld(in_c);
  const int blocks = (int)std::min<long long>(ceil_div_ll(total, 256), 148 * 8);
  wgrad_pack_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(dw_oihw, out_c, in_c, ksize * ksize, yc_conv_wgrad_ld(in_c), dw_packed);
  YC_LAUNCH_CHECK();
  return YC_OK;
}
