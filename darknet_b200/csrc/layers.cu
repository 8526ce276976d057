// Memory-bound layers of the hot path on NHWC views: layout conversion at the seam, maxpool,
// upsample, shortcut, route copy, scale_channels, reorg.  One pass each, 128-bit accesses along the
// channel axis, grids sized as a multiple of the SM count with grid-stride loops.
// Replaces maxpool_layer_kernels.cu and the corresponding pieces of blas_kernels.cu (cited per kernel).
#include <float.h>
#include <algorithm>

#include "yc_common.cuh"

namespace yc {

struct View {
  void* ptr;
  int n, h, w, c;
  long long ldc;
  int f32;
};
static inline View mk(const yc_tensor* t) { return View{t->ptr, t->n, t->h, t->w, t->c, t->ldc, t->dtype == YC_F32}; }

static inline int grid_for(long long work, int threads) {
  long long blocks = ceil_div_ll(work, threads);
  long long cap = (long long)sm_count() * 16;
  return (int)std::max<long long>(1, std::min(blocks, cap));
}

// ---- 8-channel vector load/store on a view (bf16: one 16-byte access; fp32: two) ------------------
struct V8 { float v[8]; };

__device__ __forceinline__ V8 load8(const View& t, long long pix, int c0) {
  V8 r;
  if (t.f32) {
    const float* p = reinterpret_cast<const float*>(t.ptr) + pix * t.ldc + c0;
    if (c0 + 8 <= t.c && (t.ldc & 3) == 0) {
      const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
      r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w; r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) r.v[j] = (c0 + j < t.c) ? p[j] : 0.f;
    }
  } else {
    const __nv_bfloat16* p = reinterpret_cast<const __nv_bfloat16*>(t.ptr) + pix * t.ldc + c0;
    if (c0 + 8 <= t.c) {
      const uint4 q = *reinterpret_cast<const uint4*>(p);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 f = __bfloat1622float2(h[e]);
        r.v[2 * e] = f.x; r.v[2 * e + 1] = f.y;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) r.v[j] = (c0 + j < t.c) ? __bfloat162float(p[j]) : 0.f;
    }
  }
  return r;
}
__device__ __forceinline__ void store8(const View& t, long long pix, int c0, const V8& r) {
  if (t.f32) {
    float* p = reinterpret_cast<float*>(t.ptr) + pix * t.ldc + c0;
    if (c0 + 8 <= t.c && (t.ldc & 3) == 0) {
      *reinterpret_cast<float4*>(p) = make_float4(r.v[0], r.v[1], r.v[2], r.v[3]);
      *reinterpret_cast<float4*>(p + 4) = make_float4(r.v[4], r.v[5], r.v[6], r.v[7]);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) if (c0 + j < t.c) p[j] = r.v[j];
    }
  } else {
    __nv_bfloat16* p = reinterpret_cast<__nv_bfloat16*>(t.ptr) + pix * t.ldc + c0;
    if (c0 + 8 <= t.c) {
      uint4 q;
      __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&q);
#pragma unroll
      for (int e = 0; e < 4; ++e) h[e] = __floats2bfloat162_rn(r.v[2 * e], r.v[2 * e + 1]);
      *reinterpret_cast<uint4*>(p) = q;
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) if (c0 + j < t.c) p[j] = __float2bfloat16_rn(r.v[j]);
    }
  }
}

// ---- NCHW fp32 <-> NHWC view ------------------------------------------------------------------------
// thread = (pixel, 8-channel group), pixel fastest across threads so the NCHW side is coalesced.
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ src, View dst, int real_c) {
  // dst.c may be rounded up to the (owned) pitch so that padded channels are written as zeros with
  // the same 16-byte store; real_c is the number of channels present in src.
  const long long hw = (long long)dst.h * dst.w;
  const int groups = (dst.c + 7) / 8;
  const long long total = (long long)dst.n * hw * groups;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long s = i % hw;
    const int g = (int)((i / hw) % groups);
    const long long b = i / (hw * groups);
    V8 r;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = g * 8 + j;
      r.v[j] = c < real_c ? src[(b * real_c + c) * hw + s] : 0.f;
    }
    store8(dst, b * hw + s, g * 8, r);
  }
}
// ---- stem: NCHW fp32 image -> im2col rows [b][oh][ow][32] bf16 -------------------------------------------------
// The first convolution of a detector has 3 input channels: c*k*k = 27 <= 32.  Its A operand is built HERE, while the
// fp32 image is converted anyway: row = output pixel, column k = (kh*size + kw)*c + ci, zero padding written as zeros,
// columns 27..31 zero.  The convolution then runs as a 1x1 layer with 32 input channels on the tiled TMA map (64-byte
// rows, HBM-bound) instead of nine im2col-mode tap loads of 16-byte rows per pixel.  thread = output pixel (adjacent
// lanes = adjacent pixels: the NCHW reads are coalesced and every input element is reused size^2 times out of L1).
// CT/ST > 0: channel count and kernel size known at compile time (3, 3 for every detector cfg), so the column -> (tap,
// channel) split costs nothing; CT = 0: generic.
template <int CT, int ST>
__global__ void __launch_bounds__(256) stem_im2col_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int B,
                                                          int C_, int H, int W, int OH, int OW, int size_, int stride, int pad) {
  // Each lane builds one 64-byte operand row; written directly that is four 16-byte stores at a 64-byte lane stride
  // (two partial-sector requests per 32-byte sector at the L2).  The warp's 2 KB go through shared memory instead
  // (16-byte chunks XOR-swizzled so that both sides are conflict-free) and leave as four 512-byte contiguous stores.
  __shared__ uint4 tr[256 * 4];
  const int C = CT > 0 ? CT : C_, size = ST > 0 ? ST : size_;
  const long long total = (long long)B * OH * OW;
  const int kk = C * size * size;
  const long long plane = (long long)H * W;
  const int lane = threadIdx.x & 31;
  uint4* wbuf = tr + (threadIdx.x >> 5) * 128;
  for (long long base = blockIdx.x * (long long)blockDim.x + (threadIdx.x - lane); base < total;
       base += (long long)gridDim.x * blockDim.x) {
    const long long i = base + lane;
    uint4 out[4];
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out);
    if (i < total) {
      const int ow = (int)(i % OW);
      const int oh = (int)((i / OW) % OH);
      const long long b = i / ((long long)OW * OH);
      const float* img = src + b * C * plane;
      const int ih0 = oh * stride - pad, iw0 = ow * stride - pad;
#pragma unroll
      for (int k = 0; k < 32; ++k) {
        float v = 0.f;
        if (k < kk) {
          const int tap = k / C, ci = k - tap * C;
          const int kh = tap / size, kw = tap - kh * size;
          const int ih = ih0 + kh, iw = iw0 + kw;
          if ((unsigned)ih < (unsigned)H && (unsigned)iw < (unsigned)W) v = __ldg(img + ci * plane + (long long)ih * W + iw);
        }
        o[k] = __float2bfloat16(v);
      }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) wbuf[4 * lane + (q ^ ((lane >> 1) & 3))] = out[q];
    __syncwarp();
    uint4* d = reinterpret_cast<uint4*>(dst + base * 32);
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int c = lane + 32 * r, pl = c >> 2, q = c & 3;
      if (base + pl < total) d[c] = wbuf[4 * pl + (q ^ ((pl >> 1) & 3))];
    }
    __syncwarp();
  }
}

// Four output pixels per thread for the usual stem (3 channels, 3x3, stride 1, pad 1, W % 4 == 0): per (channel, kernel
// row) one aligned float4 plus the two neighbours instead of twelve scalar loads, bounds tests once per row, and the
// warp's 8 KB of operand rows leave through shared memory as sixteen 512-byte contiguous stores.  Same values and
// rounding as stem_im2col_kernel: bit-identical operand.
__global__ void __launch_bounds__(128) stem_im2col_s1x4_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                                               int B, int H, int W) {
  __shared__ uint4 tr[4 * 512];  // per warp: 128 pixels x 4 chunks of 16 B, XOR-swizzled within groups of 8 chunks
  const int lane = threadIdx.x & 31;
  uint4* wbuf = tr + (threadIdx.x >> 5) * 512;
  const int wq = W >> 2;
  const long long total = (long long)B * H * wq, plane = (long long)H * W;
  for (long long base = blockIdx.x * (long long)blockDim.x + (threadIdx.x - lane); base < total;
       base += (long long)gridDim.x * blockDim.x) {
    const long long i = base + lane;
    if (i < total) {
      const int ow0 = (int)(i % wq) * 4;
      const int oh = (int)((i / wq) % H);
      const long long b = i / ((long long)wq * H);
      const float* img = src + b * 3 * plane;
      float v[3][3][6];  // [ci][kh][iw - (ow0 - 1)]
#pragma unroll
      for (int ci = 0; ci < 3; ++ci)
#pragma unroll
        for (int kh = 0; kh < 3; ++kh) {
          const int ih = oh + kh - 1;
          float4 m = make_float4(0.f, 0.f, 0.f, 0.f);
          float l = 0.f, r = 0.f;
          if ((unsigned)ih < (unsigned)H) {
            const float* rp = img + ci * plane + (long long)ih * W + ow0;
            m = __ldg(reinterpret_cast<const float4*>(rp));
            if (ow0 > 0) l = __ldg(rp - 1);
            if (ow0 + 4 < W) r = __ldg(rp + 4);
          }
          v[ci][kh][0] = l; v[ci][kh][1] = m.x; v[ci][kh][2] = m.y; v[ci][kh][3] = m.z; v[ci][kh][4] = m.w; v[ci][kh][5] = r;
        }
#pragma unroll
      for (int px = 0; px < 4; ++px) {
        uint4 out[4];
        __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out);
#pragma unroll
        for (int k = 0; k < 32; ++k) {
          const int tap = k / 3, ci = k - tap * 3, kh = tap / 3, kw = tap - kh * 3;
          o[k] = __float2bfloat16(k < 27 ? v[ci][kh < 3 ? kh : 0][px + kw] : 0.f);
        }
        const int P = lane * 4 + px;
#pragma unroll
        for (int q = 0; q < 4; ++q) wbuf[8 * (P >> 1) + ((4 * (P & 1) + q) ^ (lane & 7))] = out[q];
      }
    }
    __syncwarp();
    uint4* d = reinterpret_cast<uint4*>(dst + base * 4 * 32);
    const long long left = (total - base) * 4;  // pixels this warp may write
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const int c = lane + 32 * r, P = c >> 2, q = c & 3;
      if (P < left) d[c] = wbuf[8 * (P >> 1) + ((4 * (P & 1) + q) ^ ((P >> 2) & 7))];
    }
    __syncwarp();
  }
}

// ---- Mat2Image on the device (visualize.cpp:26-54): u8 HWC -> fp32 NCHW in [0,1] ----------------------------------
// image->data[k*w*h + y*w + x] = data[y*step + x*c + k] / 255.0f; swap_rb folds the preceding cvtColor(RGB2BGR)
// (yolo_core.cpp:110) by reading channel c-1-k.  thread = pixel: 3 contiguous bytes in, one coalesced store per plane.
__global__ void u8hwc_to_nchw_kernel(const unsigned char* __restrict__ src, float* __restrict__ dst, int B, int C, int H, int W,
                                     int swap) {
  const long long hw = (long long)H * W, total = (long long)B * hw;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long b = i / hw, s = i - b * hw;
    const unsigned char* px = src + i * C;
    for (int k = 0; k < C; ++k) dst[(b * C + k) * hw + s] = (float)px[swap ? C - 1 - k : k] / 255.0f;
  }
}

__global__ void nhwc_to_nchw_kernel(View src, float* __restrict__ dst) {
  const long long hw = (long long)src.h * src.w;
  const int groups = (src.c + 7) / 8;
  const long long total = (long long)src.n * hw * groups;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long s = i % hw;
    const int g = (int)((i / hw) % groups);
    const long long b = i / (hw * groups);
    const V8 r = load8(src, b * hw + s, g * 8);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = g * 8 + j;
      if (c < src.c) dst[(b * src.c + c) * hw + s] = r.v[j];
    }
  }
}

// ---- maxpool: forward_maxpool_layer_kernel (maxpool_layer_kernels.cu:58-101) ----------------------
// thread = (output pixel, 8-channel group), channel group fastest -> every window tap is one
// contiguous 16-byte read per thread, fully coalesced across the warp.
__global__ void maxpool_kernel(View x, View y, int size, int stride, int pad, int quirk) {
  const int groups = (x.c + 7) / 8;
  const long long total = (long long)y.n * y.h * y.w * groups;
  const int off = -pad / 2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i % groups);
    long long pix = i / groups;
    const int ow = (int)(pix % y.w);
    const int oh = (int)((pix / y.w) % y.h);
    const long long b = pix / ((long long)y.w * y.h);
    // reference CPU-inference artefact (gemm.c:2127-2150): in 8-wide blocks j0 < out_w-8-(size-1) a tap is
    // dropped for the whole block when it is out of range for the block's FIRST column.
    const int j0 = (ow / 8) * 8;
    const bool qblock = quirk && stride == 1 && j0 < y.w - 8 - (size - 1);
    V8 m;
#pragma unroll
    for (int j = 0; j < 8; ++j) m.v[j] = -FLT_MAX;
    for (int n = 0; n < size; ++n) {
      const int ih = off + oh * stride + n;
      if (ih < 0 || ih >= x.h) continue;
      for (int k = 0; k < size; ++k) {
        const int iw = off + ow * stride + k;
        if (iw < 0 || iw >= x.w) continue;
        if (qblock && (off + j0 + k < 0)) continue;
        const V8 v = load8(x, (b * x.h + ih) * x.w + iw, g * 8);
#pragma unroll
        for (int j = 0; j < 8; ++j) m.v[j] = v.v[j] > m.v[j] ? v.v[j] : m.v[j];
      }
    }
    store8(y, pix, g * 8, m);
  }
}

// Stride-1 windows on small maps (the SPP block: 5/9/13 at 19x19 or 13x13, maxpool_layer_kernels.cu:58-101 with
// stride 1, pad size-1): one block stages an (image, 16-channel) slab in shared memory, takes the row maxima into a
// second slab and the column maxima of those into the output: 2*size shared-memory taps per output instead of size^2
// 16-byte global loads (13x13: 26 vs 169; the direct kernel ran at 84 GB/s).  max is a selection, so it runs on the
// bf16 values themselves (__hmax2) and the result is identical to the direct kernel's.
__device__ __forceinline__ uint4 max8_bf16(const uint4& a, const uint4& b) {
  uint4 r;
  const __nv_bfloat162* pa = reinterpret_cast<const __nv_bfloat162*>(&a);
  const __nv_bfloat162* pb = reinterpret_cast<const __nv_bfloat162*>(&b);
  __nv_bfloat162* pr = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
  for (int e = 0; e < 4; ++e) pr[e] = __hmax2(pa[e], pb[e]);
  return r;
}
constexpr int kPoolG = 2;  // 8-channel groups per block: 32 contiguous bytes (one sector) per pixel
__global__ void maxpool_s1_slab_kernel(View x, View y, int size, int off) {
  extern __shared__ uint4 pool_sm[];
  const int hw = x.h * x.w, items = hw * kPoolG;
  uint4* in = pool_sm;
  uint4* rows = pool_sm + items;
  const int chunks = (x.c / 8 + kPoolG - 1) / kPoolG;
  const int b = blockIdx.x / chunks, g0 = (blockIdx.x % chunks) * kPoolG;
  const int gn = min(kPoolG, x.c / 8 - g0);
  const __nv_bfloat16* xp = reinterpret_cast<const __nv_bfloat16*>(x.ptr) + (long long)b * hw * x.ldc + g0 * 8;
  __nv_bfloat16* yp = reinterpret_cast<__nv_bfloat16*>(y.ptr) + (long long)b * hw * y.ldc + g0 * 8;
  for (int i = threadIdx.x; i < items; i += blockDim.x) {
    const int pix = i / kPoolG, g = i % kPoolG;
    if (g < gn) in[i] = *reinterpret_cast<const uint4*>(xp + (long long)pix * x.ldc + g * 8);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < items; i += blockDim.x) {
    const int pix = i / kPoolG, g = i % kPoolG;
    if (g >= gn) continue;
    const int ow = pix % x.w, r0 = pix - ow;
    const int k0 = max(0, ow + off), k1 = min(x.w - 1, ow + off + size - 1);
    uint4 m = in[(r0 + k0) * kPoolG + g];
    for (int k = k0 + 1; k <= k1; ++k) m = max8_bf16(m, in[(r0 + k) * kPoolG + g]);
    rows[i] = m;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < items; i += blockDim.x) {
    const int pix = i / kPoolG, g = i % kPoolG;
    if (g >= gn) continue;
    const int oh = pix / x.w, ow = pix - oh * x.w;
    const int n0 = max(0, oh + off), n1 = min(x.h - 1, oh + off + size - 1);
    uint4 m = rows[(n0 * x.w + ow) * kPoolG + g];
    for (int n = n0 + 1; n <= n1; ++n) m = max8_bf16(m, rows[(n * x.w + ow) * kPoolG + g]);
    *reinterpret_cast<uint4*>(yp + (long long)pix * y.ldc + g * 8) = m;
  }
}

// The SPP block as ONE launch: its three stride-1 max-pools (5 / 9 / 13) read the same tensor, so the slab is staged once
// and the separable row / column passes run once per window into the three outputs (each result identical to
// maxpool_s1_slab_kernel's: max is a selection).  Three launches of ~25 us on 24 MB tensors were launch / ramp bound.
struct PoolMulti {
  View y[3];
  int size[3], off[3];
  int n;
};
__global__ void maxpool_s1_multi_kernel(View x, PoolMulti pm) {
  extern __shared__ uint4 pool_sm[];
  const int hw = x.h * x.w, items = hw * kPoolG;
  uint4* in = pool_sm;
  uint4* rows = pool_sm + items;
  const int chunks = (x.c / 8 + kPoolG - 1) / kPoolG;
  const int b = blockIdx.x / chunks, g0 = (blockIdx.x % chunks) * kPoolG;
  const int gn = min(kPoolG, x.c / 8 - g0);
  const __nv_bfloat16* xp = reinterpret_cast<const __nv_bfloat16*>(x.ptr) + (long long)b * hw * x.ldc + g0 * 8;
  for (int i = threadIdx.x; i < items; i += blockDim.x) {
    const int pix = i / kPoolG, g = i % kPoolG;
    if (g < gn) in[i] = *reinterpret_cast<const uint4*>(xp + (long long)pix * x.ldc + g * 8);
  }
  for (int o = 0; o < pm.n; ++o) {
    const int size = pm.size[o], off = pm.off[o];
    __syncthreads();  // the slab is loaded / the previous window's column pass has read `rows`
    for (int i = threadIdx.x; i < items; i += blockDim.x) {
      const int pix = i / kPoolG, g = i % kPoolG;
      if (g >= gn) continue;
      const int ow = pix % x.w, r0 = pix - ow;
      const int k0 = max(0, ow + off), k1 = min(x.w - 1, ow + off + size - 1);
      uint4 m = in[(r0 + k0) * kPoolG + g];
      for (int k = k0 + 1; k <= k1; ++k) m = max8_bf16(m, in[(r0 + k) * kPoolG + g]);
      rows[i] = m;
    }
    __syncthreads();
    __nv_bfloat16* yp = reinterpret_cast<__nv_bfloat16*>(pm.y[o].ptr) + (long long)b * hw * pm.y[o].ldc + g0 * 8;
    const long long ldy = pm.y[o].ldc;
    for (int i = threadIdx.x; i < items; i += blockDim.x) {
      const int pix = i / kPoolG, g = i % kPoolG;
      if (g >= gn) continue;
      const int oh = pix / x.w, ow = pix - oh * x.w;
      const int n0 = max(0, oh + off), n1 = min(x.h - 1, oh + off + size - 1);
      uint4 m = rows[(n0 * x.w + ow) * kPoolG + g];
      for (int n = n0 + 1; n <= n1; ++n) m = max8_bf16(m, rows[(n * x.w + ow) * kPoolG + g]);
      *reinterpret_cast<uint4*>(yp + (long long)pix * ldy + g * 8) = m;
    }
  }
}

// ---- upsample: upsample_kernel (blas_kernels.cu:1121-1155), forward, nearest ----------------------
__global__ void upsample_kernel(View x, View y, int stride, float scale) {
  const int groups = (x.c + 7) / 8;
  const long long total = (long long)y.n * y.h * y.w * groups;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i % groups);
    const long long pix = i / groups;
    const int ow = (int)(pix % y.w);
    const int oh = (int)((pix / y.w) % y.h);
    const long long b = pix / ((long long)y.w * y.h);
    V8 v = load8(x, (b * x.h + oh / stride) * x.w + ow / stride, g * 8);
    if (scale != 1.f) {
#pragma unroll
      for (int j = 0; j < 8; ++j) v.v[j] *= scale;
    }
    store8(y, pix, g * 8, v);
  }
}

// ---- shortcut: shortcut_singlelayer_simple_kernel (blas_kernels.cu:941-978) + activation ----------
__global__ void shortcut_kernel(View a, View b, View y, int act) {
  const int groups = (y.c + 7) / 8;
  const long long total = (long long)y.n * y.h * y.w * groups;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i % groups);
    const long long pix = i / groups;
    V8 va = load8(a, pix, g * 8);
    const V8 vb = load8(b, pix, g * 8);
#pragma unroll
    for (int j = 0; j < 8; ++j) va.v[j] = apply_act(va.v[j] + vb.v[j], act);
    store8(y, pix, g * 8, va);
  }
}

// ---- route copy: simple_copy_kernel (blas_kernels.cu:520-525) on views ------------------------------
__global__ void copy_kernel(View s, View d) {
  const int groups = (d.c + 7) / 8;
  const long long total = (long long)d.n * d.h * d.w * groups;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i % groups);
    const long long pix = i / groups;
    store8(d, pix, g * 8, load8(s, pix, g * 8));
  }
}

// ---- scale_channels: scale_channels_kernel (blas_kernels.cu:1318-1346) -----------------------------
__global__ void scale_channels_kernel(View x, View s, View y, int act) {
  const int groups = (y.c + 7) / 8;
  const long long hw = (long long)y.h * y.w;
  const long long total = (long long)y.n * hw * groups;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i % groups);
    const long long pix = i / groups;
    V8 v = load8(x, pix, g * 8);
    const V8 sc = load8(s, pix / hw, g * 8);
#pragma unroll
    for (int j = 0; j < 8; ++j) v.v[j] = apply_act(v.v[j] * sc.v[j], act);
    store8(y, pix, g * 8, v);
  }
}


// ---- scale_channels backward: backward_scale_channels_kernel (blas_kernels.cu:1348-1420) ------------------------------
// BackwardScaleChannelsLayer (scale_channels_layer.c:99-131): dy *= act'(y) in place; ds[n,c] += sum_hw dy * x;
// dx += s[n,c] * dy.  One block per (image, 8-channel group): every thread walks the pixels with stride blockDim,
// 16-byte accesses along C, the per-channel sums are reduced with warp shuffles (the reference's warp-shuffle channel
// sum) and then across the eight warps through shared memory; the block owns its eight ds entries, so no atomics and a
// fixed summation order.
__device__ __forceinline__ float sc_act_grad_out(float y, int act) {  // gradient_array on the output (activations.h:152-178)
  switch (act) {
    case YC_LEAKY: return y > 0.f ? 1.f : .1f;
    case YC_RELU: return y > 0.f ? 1.f : 0.f;
    case YC_LOGISTIC: return (1.f - y) * y;
  }
  return 1.f;
}
__global__ void scale_channels_bwd_kernel(View dy, View y, int act, View x, View s, View ds, View dx) {
  const int groups = (dy.c + 7) / 8;
  const int g = blockIdx.x % groups;
  const long long n = blockIdx.x / groups;
  const long long hw = (long long)dy.h * dy.w;
  const V8 sc = load8(s, n, g * 8);
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  for (long long pp = threadIdx.x; pp < hw; pp += blockDim.x) {
    const long long pix = n * hw + pp;
    V8 d = load8(dy, pix, g * 8);
    if (act != YC_LINEAR) {
      const V8 yo = load8(y, pix, g * 8);
#pragma unroll
      for (int j = 0; j < 8; ++j) d.v[j] *= sc_act_grad_out(yo.v[j], act);
      store8(dy, pix, g * 8, d);
    }
    const V8 xv = load8(x, pix, g * 8);
    V8 dxv = load8(dx, pix, g * 8);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      acc[j] = fmaf(d.v[j], xv.v[j], acc[j]);
      dxv.v[j] = fmaf(sc.v[j], d.v[j], dxv.v[j]);
    }
    store8(dx, pix, g * 8, dxv);
  }
  __shared__ float part[8][8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    float v = acc[j];
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    if (lane == 0) part[warp][j] = v;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    V8 dsv = load8(ds, n, g * 8);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float v = 0.f;
      for (int w = 0; w < (int)(blockDim.x >> 5); ++w) v += part[w][j];
      dsv.v[j] += v;
    }
    store8(ds, n, g * 8, dsv);
  }
}

// ---- reorg: reorg_kernel (blas_kernels.cu:368-402), index math of reorg_cpu (blas.c:12-46) ----------
// forward==0 (what [reorg] forward uses): y[b,k,j,i] = x[b, k%in_c, j*s + (k/in_c)/s, i*s + (k/in_c)%s]
// forward==1: the inverse scatter.  Scalar: not used by any of the five configs.
__device__ __forceinline__ float ld1(const View& t, long long pix, int c) {
  return t.f32 ? reinterpret_cast<const float*>(t.ptr)[pix * t.ldc + c]
               : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(t.ptr)[pix * t.ldc + c]);
}
__device__ __forceinline__ void st1(const View& t, long long pix, int c, float v) {
  if (t.f32) reinterpret_cast<float*>(t.ptr)[pix * t.ldc + c] = v;
  else reinterpret_cast<__nv_bfloat16*>(t.ptr)[pix * t.ldc + c] = __float2bfloat16_rn(v);
}
__global__ void reorg_kernel(View x, View y, int stride, int forward) {
  // "small" = the tensor with (c*s*s, h/s, w/s); "big" = (c, h, w).  forward==0: y small <- x big.
  const View sm = forward ? x : y;
  const View bg = forward ? y : x;
  const int in_c = sm.c / (stride * stride);
  const long long total = (long long)sm.n * sm.h * sm.w * sm.c;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(i % sm.c);
    const long long pix = i / sm.c;
    const int ii = (int)(pix % sm.w);
    const int jj = (int)((pix / sm.w) % sm.h);
    const long long b = pix / ((long long)sm.w * sm.h);
    const int c2 = k % in_c, offset = k / in_c;
    const int w2 = ii * stride + offset % stride, h2 = jj * stride + offset / stride;
    const long long bpix = (b * bg.h + h2) * bg.w + w2;
    if (forward) st1(bg, bpix, c2, ld1(sm, pix, k));
    else st1(sm, pix, k, ld1(bg, bpix, c2));
  }
}

}  // namespace yc

using namespace yc;

#define YC_CHECK_T(t, name) YC_REQUIRE((t) && (t)->ptr && (t)->n > 0 && (t)->h > 0 && (t)->w > 0 && (t)->c > 0 && (t)->ldc >= (t)->c, "%s: bad tensor", name)
static bool aligned16(const yc_tensor* t) {
  const int e = t->dtype == YC_F32 ? 4 : 8;
  return ((uintptr_t)t->ptr & 15) == 0 && t->ldc % e == 0;
}

extern "C" int yc_nchw_to_nhwc(const float* src, const yc_tensor* dst, void* stream) {
  YC_CHECK_T(dst, "yc_nchw_to_nhwc");
  YC_REQUIRE(src && aligned16(dst), "yc_nchw_to_nhwc: dst must be 16-byte aligned");
  const long long work = (long long)dst->n * dst->h * dst->w * ((dst->c + 7) / 8);
  View d = mk(dst);
  const int e = dst->dtype == YC_F32 ? 4 : 8;
  if (dst->ldc == round_up(dst->c, e)) d.c = (int)dst->ldc;  // tensor owns its pitch: zero-fill the pad lanes
  nchw_to_nhwc_kernel<<<grid_for(work, 256), 256, 0, (cudaStream_t)stream>>>(src, d, dst->c);
  YC_LAUNCH_CHECK();
  return YC_OK;
}
// Builds the first layer's GEMM operand straight from the fp32 NCHW image (see stem_im2col_kernel).  dst: bf16
// [n][oh][ow][32] with ldc 32, oh = (h + 2*pad - size)/stride + 1.  Requires c*size*size <= 32.
extern "C" int yc_stem_im2col(const float* src, int c, int h, int w, int size, int stride, int pad, const yc_tensor* dst,
                              void* stream) {
  YC_CHECK_T(dst, "yc_stem_im2col");
  YC_REQUIRE(src && c >= 1 && size >= 1 && stride >= 1 && pad >= 0 && c * size * size <= 32,
             "yc_stem_im2col: needs c*size*size <= 32");
  YC_REQUIRE(dst->dtype == YC_BF16 && dst->c == 32 && dst->ldc == 32 && aligned16(dst), "yc_stem_im2col: dst must be bf16 [.,.,.,32], pitch 32");
  YC_REQUIRE(dst->h == (h + 2 * pad - size) / stride + 1 && dst->w == (w + 2 * pad - size) / stride + 1,
             "yc_stem_im2col: output shape mismatch");
  const long long work = (long long)dst->n * dst->h * dst->w;
  if (c == 3 && size == 3 && stride == 1 && pad == 1 && w % 4 == 0 && (reinterpret_cast<uintptr_t>(src) & 15) == 0)
    stem_im2col_s1x4_kernel<<<grid_for((work + 3) / 4, 128), 128, 0, (cudaStream_t)stream>>>(src, (__nv_bfloat16*)dst->ptr, dst->n, h, w);
  else if (c == 3 && size == 3)
    stem_im2col_kernel<3, 3><<<grid_for(work, 256), 256, 0, (cudaStream_t)stream>>>(src, (__nv_bfloat16*)dst->ptr, dst->n, c, h, w,
                                                                                    dst->h, dst->w, size, stride, pad);
  else
    stem_im2col_kernel<0, 0><<<grid_for(work, 256), 256, 0, (cudaStream_t)stream>>>(src, (__nv_bfloat16*)dst->ptr, dst->n, c, h, w,
                                                                                    dst->h, dst->w, size, stride, pad);
  YC_LAUNCH_CHECK();
  return YC_OK;
}
extern "C" int yc_u8hwc_to_nchw(const unsigned char* src, int n, int c, int h, int w, int swap_rb, float* dst, void* stream) {
  YC_REQUIRE(src && dst && n > 0 && c > 0 && h > 0 && w > 0, "yc_u8hwc_to_nchw: bad arguments");
  u8hwc_to_nchw_kernel<<<grid_for((long long)n * h * w, 256), 256, 0, (cudaStream_t)stream>>>(src, dst, n, c, h, w, swap_rb ? 1 : 0);
  YC_LAUNCH_CHECK();
  return YC_OK;
}
extern "C" int yc_nhwc_to_nchw(const yc_tensor* src, float* dst, void* stream) {
  YC_CHECK_T(src, "yc_nhwc_to_nchw");
  YC_REQUIRE(dst && aligned16(src), "yc_nhwc_to_nchw: src must be 16-byte aligned");
  const long long work = (long long)src->n * src->h * src->w * ((src->c + 7) / 8);
  nhwc_to_nchw_kernel<<<grid_for(work, 256), 256, 0, (cudaStream_t)stream>>>(mk(src), dst);
  YC_LAUNCH_CHECK();
  return YC_OK;
}

extern "C" int yc_maxpool_fwd(const yc_tensor* x, const yc_tensor* y, int size, int stride, int pad, int quirk,
                              void* stream) {
  YC_CHECK_T(x, "yc_maxpool_fwd x");
  YC_CHECK_T(y, "yc_maxpool_fwd y");
  YC_REQUIRE(aligned16(x) && aligned16(y), "yc_maxpool_fwd: views must be 16-byte aligned");
  YC_REQUIRE(size >= 1 && stride >= 1 && pad >= 0, "yc_maxpool_fwd: bad geometry");
  YC_REQUIRE(y->n == x->n && y->c == x->c && y->h == (x->h + pad - size) / stride + 1 &&
                 y->w == (x->w + pad - size) / stride + 1,
             "yc_maxpool_fwd: output shape mismatch");
  // stride-1 "same" windows on bf16 maps that fit a shared-memory slab: separable row/column maxima
  const size_t slab = (size_t)x->h * x->w * kPoolG * sizeof(uint4) * 2;
  if (stride == 1 && !quirk && x->dtype == YC_BF16 && y->dtype == YC_BF16 && x->c % 8 == 0 && y->h == x->h &&
      y->w == x->w && pad / 2 < size && size - 1 - pad / 2 >= 0 && slab <= 160 * 1024) {
    if (slab > 48 * 1024)
      YC_CUDA(cudaFuncSetAttribute(maxpool_s1_slab_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    const int chunks = (x->c / 8 + kPoolG - 1) / kPoolG;
    maxpool_s1_slab_kernel<<<x->n * chunks, 256, slab, (cudaStream_t)stream>>>(mk(x), mk(y), size, -pad / 2);
    YC_LAUNCH_CHECK();
    return YC_OK;
  }
  const long long work = (long long)y->n * y->h * y->w * ((y->c + 7) / 8);
  maxpool_kernel<<<grid_for(work, 256), 256, 0, (cudaStream_t)stream>>>(mk(x), mk(y), size, stride, pad, quirk);
  YC_LAUNCH_CHECK();
  return YC_OK;
}

// true when yc_maxpool_fwd would take the shared-memory slab path for this geometry (stride 1, "same" output, bf16)
static bool pool_slab_ok(const yc_tensor* x, const yc_tensor* y, int size, int stride, int pad) {
  const size_t slab = (size_t)x->h * x->w * kPoolG * sizeof(uint4) * 2;
  return stride == 1 && x->dtype == YC_BF16 && y->dtype == YC_BF16 && x->c % 8 == 0 && y->h == x->h && y->w == x->w &&
         y->n == x->n && y->c == x->c && pad / 2 < size && size - 1 - pad / 2 >= 0 && slab <= 160 * 1024;
}
extern "C" int yc_maxpool_s1_multi(const yc_tensor* x, int n_out, const yc_tensor* const* ys, const int* sizes, const int* pads,
                                   void* stream) {
  YC_CHECK_T(x, "yc_maxpool_s1_multi x");
  YC_REQUIRE(n_out >= 1 && n_out <= 3 && ys && sizes && pads && aligned16(x), "yc_maxpool_s1_multi: 1..3 outputs");
  PoolMulti pm;
  pm.n = n_out;
  for (int o = 0; o < n_out; ++o) {
    YC_CHECK_T(ys[o], "yc_maxpool_s1_multi y");
    YC_REQUIRE(aligned16(ys[o]) && sizes[o] >= 1 && pads[o] >= 0 && pool_slab_ok(x, ys[o], sizes[o], 1, pads[o]),
               "yc_maxpool_s1_multi: output %d is not a stride-1 same-size bf16 max-pool that fits the shared-memory slab", o);
    pm.y[o] = mk(ys[o]);
    pm.size[o] = sizes[o];
    pm.off[o] = -pads[o] / 2;
  }
  const size_t slab = (size_t)x->h * x->w * kPoolG * sizeof(uint4) * 2;
  if (slab > 48 * 1024)
    YC_CUDA(cudaFuncSetAttribute(maxpool_s1_multi_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
  const int chunks = (x->c / 8 + kPoolG - 1) / kPoolG;
  maxpool_s1_multi_kernel<<<x->n * chunks, 256, slab, (cudaStream_t)stream>>>(mk(x), pm);
  YC_LAUNCH_CHECK();
  return YC_OK;
}
extern "C" int yc_maxpool_s1_multi_eligible(const yc_tensor* x, const yc_tensor* y, int size, int stride, int pad) {
  return (x && y && aligned16(x) && aligned16(y) && pool_slab_ok(x, y, size, stride, pad)) ? 1 : 0;
}

extern "C" int yc_upsample_fwd(const yc_tensor* x, const yc_tensor* y, int stride, float scale, void* stream) {
  YC_CHECK_T(x, "yc_upsample_fwd x");
  YC_CHECK_T(y, "yc_upsample_fwd y");
  YC_REQUIRE(aligned16(x) && aligned16(y), "yc_upsample_fwd: views must be 16-byte aligned");
  YC_REQUIRE(stride >= 1 && y->n == x->n && y->c == x->c && y->h == x->h * stride && y->w == x->w * stride,
             "yc_upsample_fwd: output shape mismatch");
  const long long work = (long long)y->n * y->h * y->w * ((y->c + 7) / 8);
  upsample_kernel<<<grid_for(work, 256), 256, 0, (cudaStream_t)stream>>>(mk(x), mk(y), stride, scale);
  YC_LAUNCH_CHECK();
  return YC_OK;
}

static bool same_shape(const yc_tensor* a, const yc_tensor* b) {
  return a->n == b->n && a->h == b->h && a->w == b->w && a->c == b->c;
}

extern "C" int yc_shortcut_fwd(const yc_tensor* a, const yc_tensor* b, const yc_tensor* y, int act, void* stream) {
  YC_CHECK_T(a, "yc_shortcut_fwd a");
  YC_CHECK_T(b, "yc_shortcut_fwd b");
  YC_CHECK_T(y, "yc_shortcut_fwd y");
  YC_REQUIRE(same_shape(a, b) && same_shape(a, y), "yc_shortcut_fwd: shapes differ (the reference asserts equal dims, shortcut_layer.c:151-164)");
  YC_REQUIRE(aligned16(a) && aligned16(b) && aligned16(y), "yc_shortcut_fwd: views must be 16-byte aligned");
  const long long work = (long long)y->n * y->h * y->w * ((y->c + 7) / 8);
  shortcut_kernel<<<grid_for(work, 256), 256, 0, (cudaStream_t)stream>>>(mk(a), mk(b), mk(y), act);
  YC_LAUNCH_CHECK();
  return YC_OK;
}

extern "C" int yc_route_copy(const yc_tensor* s, const yc_tensor* d, void* stream) {
  YC_CHECK_T(s, "yc_route_copy src");
  YC_CHECK_T(d, "yc_route_copy dst");
  YC_REQUIRE(same_shape(s, d) && aligned16(s) && aligned16(d), "yc_route_copy: shape/alignment mismatch");
  const long long work = (long long)d->n * d->h * d->w * ((d->c + 7) / 8);
  copy_kernel<<<grid_for(work, 256), 256, 0, (cudaStream_t)stream>>>(mk(s), mk(d));
  YC_LAUNCH_CHECK();
  return YC_OK;
}

extern "C" int yc_scale_channels_fwd(const yc_tensor* x, const yc_tensor* s, const yc_tensor* y, int act, void* stream) {
  YC_CHECK_T(x, "yc_scale_channels_fwd x");
  YC_CHECK_T(s, "yc_scale_channels_fwd s");
  YC_CHECK_T(y, "yc_scale_channels_fwd y");
  YC_REQUIRE(same_shape(x, y) && s->n == x->n && s->c == x->c && s->h == 1 && s->w == 1,
             "yc_scale_channels_fwd: scale must be [n,1,1,c]");
  YC_REQUIRE(aligned16(x) && aligned16(s) && aligned16(y), "yc_scale_channels_fwd: alignment");
  const long long work = (long long)y->n * y->h * y->w * ((y->c + 7) / 8);
  scale_channels_kernel<<<grid_for(work, 256), 256, 0, (cudaStream_t)stream>>>(mk(x), mk(s), mk(y), act);
  YC_LAUNCH_CHECK();
  return YC_OK;
}

extern "C" int yc_scale_channels_bwd(const yc_tensor* dy, const yc_tensor* y, int act, const yc_tensor* x, const yc_tensor* s,
                                     const yc_tensor* ds, const yc_tensor* dx, void* stream) {
  YC_CHECK_T(dy, "yc_scale_channels_bwd dy");
  YC_CHECK_T(x, "yc_scale_channels_bwd x");
  YC_CHECK_T(s, "yc_scale_channels_bwd s");
  YC_CHECK_T(ds, "yc_scale_channels_bwd ds");
  YC_CHECK_T(dx, "yc_scale_channels_bwd dx");
  YC_REQUIRE(act == YC_LINEAR || act == YC_LEAKY || act == YC_RELU || act == YC_LOGISTIC,
             "yc_scale_channels_bwd: activation outside the hot path");
  if (act != YC_LINEAR) {
    YC_CHECK_T(y, "yc_scale_channels_bwd y");
    YC_REQUIRE(same_shape(dy, y) && aligned16(y), "yc_scale_channels_bwd: y must have dy's shape");
  }
  YC_REQUIRE(same_shape(dy, x) && same_shape(dy, dx) && s->n == dy->n && s->c == dy->c && s->h == 1 && s->w == 1 &&
                 same_shape(s, ds),
             "yc_scale_channels_bwd: x / dx must have dy's shape, s / ds must be [n,1,1,c]");
  YC_REQUIRE(aligned16(dy) && aligned16(x) && aligned16(s) && aligned16(ds) && aligned16(dx), "yc_scale_channels_bwd: alignment");
  const int groups = (dy->c + 7) / 8;
  scale_channels_bwd_kernel<<<dy->n * groups, 256, 0, (cudaStream_t)stream>>>(mk(dy), act == YC_LINEAR ? mk(dy) : mk(y), act, mk(x),
                                                                              mk(s), mk(ds), mk(dx));
  YC_LAUNCH_CHECK();
  return YC_OK;
}

// BackwardReorgLayerGpu (reorg_layer.c:129-143): the inverse rearrangement of l->delta into state.delta, OVERWRITING it
// like reorg_cpu does.  `reverse` = l->reverse of the forward layer.
extern "C" int yc_reorg_bwd(const yc_tensor* dy, const yc_tensor* dx, int stride, int reverse, void* stream) {
  return yc_reorg_fwd(dy, dx, stride, reverse ? 0 : 1, stream);
}

extern "C" int yc_reorg_fwd(const yc_tensor* x, const yc_tensor* y, int stride, int forward, void* stream) {
  YC_CHECK_T(x, "yc_reorg_fwd x");
  YC_CHECK_T(y, "yc_reorg_fwd y");
  const yc_tensor* sm = forward ? x : y;
  const yc_tensor* bg = forward ? y : x;
  YC_REQUIRE(stride >= 1 && sm->n == bg->n && sm->c == bg->c * stride * stride && sm->h * stride == bg->h &&
                 sm->w * stride == bg->w,
             "yc_reorg_fwd: shape mismatch");
  const long long work = (long long)sm->n * sm->h * sm->w * sm->c;
  reorg_kernel<<<grid_for(work, 256), 256, 0, (cudaStream_t)stream>>>(mk(x), mk(y), stride, forward);
  YC_LAUNCH_CHECK();
  return YC_OK;
}
