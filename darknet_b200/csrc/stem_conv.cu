// First convolution of a detection net computed straight from the network input (opt-in: YC_NET_STEM_DIRECT).
//
// Replaces, for layer 0 of the YOLO backbones (3 -> 32 channels, 3x3, stride 1, pad 1): the input upload's layout step
// (network_kernels.cu:513), im2col_gpu_ext + gemm (convolutional_kernels.cu:252-553 / im2col_kernels.cu) and the bias +
// activation launches — in ONE kernel that reads the fp32 NCHW image the caller handed over (12 bytes per pixel) and
// writes the layer's NHWC bf16 output (64 bytes per pixel).  No im2col operand exists anywhere, not even in global memory:
// the default path (yc_stem_im2col + a 1x1 tcgen05 conv over the 64-byte operand rows) moves 204 bytes per pixel and sits
// at that HBM floor (0.85 ms per yolov4-608 x 64 step), this one needs 76.
//
// K = 27 is far too short for a tcgen05 tile fed by TMA, so the contraction runs on the warp-level tensor path: mma.sync
// m16n8k16 bf16 -> fp32 with the A fragments assembled in REGISTERS from a shared-memory patch of the image (3 x 18 x 34
// fp32, double-buffered with cp.async), the 32 x 32 filter matrix resident in registers, accumulators in registers.
//
// MEASURED (same box, tools/ab_layers.py direct / im2col, profiles/r02_ab_stem_direct.txt): 1.32 ms against 0.53 + 0.32 ms.
// Hiding the patch loads (first version: 1.57 ms) did not change the picture: the 11.8 M warp-level HMMA.16816 of the
// layer are what bounds it — the legacy mma.sync path of sm_100 runs at a small fraction of the tcgen05 rate.  The kernel
// therefore stays an opt-in, tested alternative (tests/test_conv_gpu.py::test_stem_conv_direct, the micro net through
// it); the way to beat the HBM floor of the default path is a tcgen05 kernel whose producer warps build the 128 x 32
// operand tile in shared memory (not built).
//
// Tile = 16 rows x 32 columns of output pixels per 256-thread block (8 warps x two rows; a row of 32 pixels = two
// m16 tiles); persistent grid over tiles.  K order = (kh*3 + kw)*3 + ci, the order yc_stem_im2col uses (and the stem
// filter packing [n][32] of network.cu), values rounded to bf16 exactly like the operand that path writes.
#include <algorithm>

#include "yc_common.cuh"

namespace yc {

constexpr int kStemRows = 16, kStemCols = 32;
constexpr int kStemPH = kStemRows + 2, kStemPW = kStemCols + 2;  // patch with a one-pixel halo
constexpr int kStemPitch = kStemPW + 1;                           // 35: odd pitch, fewer bank conflicts between kernel rows

__device__ __forceinline__ uint32_t stem_pack(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ void mma_m16n8k16_bf16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

template <int ACT>
__global__ void __launch_bounds__(256, 2)
stem_conv_direct_kernel(const float* __restrict__ x, int B, int H, int W, const __nv_bfloat16* __restrict__ wp,
                        const float* __restrict__ bias, int n_out, __nv_bfloat16* __restrict__ y, long long ldy, int tiles_h,
                        int tiles_w) {
  __shared__ float patch2[2][3 * kStemPH * kStemPitch];  // double-buffered: the next tile's patch arrives while this one is used
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, tg = lane & 3;

  // ---- per-lane constants: patch offsets of this lane's eight K indices, filter fragments, bias ------------------
  // A fragment of m16n8k16 (row = pixel, column = k): registers {0,1} hold k = 2tg, 2tg+1 (+16 per K step), {2,3} the
  // same + 8; rows g (even registers) and g + 8 (odd ones).
  int koff[2][2][2];  // [K step][half (+8)][element]: offset of (ci, kh, kw) in the patch, -1 for the padding columns k >= 27
#pragma unroll
  for (int s = 0; s < 2; ++s)
#pragma unroll
    for (int hf = 0; hf < 2; ++hf)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int k = 16 * s + 8 * hf + 2 * tg + e;
        const int tap = k / 3, ci = k - tap * 3, kh = tap / 3, kw = tap - kh * 3;
        koff[s][hf][e] = k < 27 ? (ci * kStemPH + kh) * kStemPitch + kw : -1;
      }
  // B fragment (k x n, "col" = n-major rows of the packed filter [n][32]): b0 = W[n = 8j + g][k = 16s + 2tg, +1], b1 = .. + 8
  uint32_t bw[2][4][2];
#pragma unroll
  for (int s = 0; s < 2; ++s)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = 8 * j + g;
      const uint32_t* row = reinterpret_cast<const uint32_t*>(wp + (size_t)n * 32);
      bw[s][j][0] = n < n_out ? __ldg(row + 8 * s + tg) : 0u;
      bw[s][j][1] = n < n_out ? __ldg(row + 8 * s + tg + 4) : 0u;
    }
  // accumulator columns of this lane: n = 8j + 2tg, 2tg + 1
  float bs[4][2];
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int n = 8 * j + 2 * tg + e;
      bs[j][e] = n < n_out ? __ldg(bias + n) : 0.f;
    }

  const long long plane = (long long)H * W;
  const int tiles = B * tiles_h * tiles_w;
  const int tpi = tiles_h * tiles_w;
  // the image patch of tile t (three channels, one-pixel halo, zeros outside the image = the convolution's padding) ->
  // buffer `buf`, asynchronously: warp w copies patch rows w, w + 8, ... (54 rows of 34 floats, 136 contiguous bytes each)
  auto fetch = [&](int t, int buf) {
    const int b = t / tpi;
    const int rem = t - b * tpi;
    const int th = rem / tiles_w, tw = rem - th * tiles_w;
    const int h0 = th * kStemRows, w0 = tw * kStemCols;
    const float* xb = x + (long long)b * 3 * plane;
    const uint32_t dst0 = (uint32_t)__cvta_generic_to_shared(&patch2[buf][0]);
#pragma unroll
    for (int k = 0; k < 7; ++k) {
      const int row = warp + 8 * k;  // < 56; 54 rows exist
      if (row < 3 * kStemPH) {
        const int ci = row / kStemPH, pr = row - ci * kStemPH;
        const int hh = h0 - 1 + pr;
        const bool row_ok = hh >= 0 && hh < H;
        const float* src = xb + ci * plane + (long long)(row_ok ? hh : 0) * W;
#pragma unroll
        for (int q = 0; q < 2; ++q) {
          const int pc = lane + 32 * q;
          if (pc < kStemPW) {
            const int ww = w0 - 1 + pc;
            const bool ok = row_ok && ww >= 0 && ww < W;
            const uint32_t d = dst0 + 4u * (uint32_t)(row * kStemPitch + pc);
            const float* sp = src + (ok ? ww : 0);
            const uint32_t nbytes = ok ? 4u : 0u;  // 0: the element is written as zero without reading
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(d), "l"(sp), "r"(nbytes) : "memory");
          }
        }
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  int cur = 0;
  if ((int)blockIdx.x < tiles) fetch(blockIdx.x, 0);
  for (int t = blockIdx.x; t < tiles; t += gridDim.x, cur ^= 1) {
    const int b = t / tpi;
    const int rem = t - b * tpi;
    const int th = rem / tiles_w, tw = rem - th * tiles_w;
    const int h0 = th * kStemRows, w0 = tw * kStemCols;
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();  // this tile's patch is complete for every thread; the other buffer's readers (previous tile) are done
    if (t + (int)gridDim.x < tiles) fetch(t + gridDim.x, cur ^ 1);
    const float* patch = &patch2[cur][0];
#pragma unroll 1
    for (int rr = 0; rr < 2; ++rr) {
      const int r = warp + 8 * rr;  // output row inside the tile
      const int h = h0 + r;
      if (h >= H) continue;
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) {
        const int c0 = mt * 16;
        if (w0 + c0 >= W) continue;
        const float* pb = patch + r * kStemPitch + c0;  // (kh, kw) = (0, 0) tap of pixel column c0, channel 0
        float acc[4][4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
          for (int e = 0; e < 4; ++e) acc[j][e] = 0.f;
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          uint32_t a[4];
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            const int o0 = koff[s][hf][0], o1 = koff[s][hf][1];
            const float v00 = o0 >= 0 ? pb[o0 + g] : 0.f, v01 = o1 >= 0 ? pb[o1 + g] : 0.f;           // pixel row g
            const float v10 = o0 >= 0 ? pb[o0 + g + 8] : 0.f, v11 = o1 >= 0 ? pb[o1 + g + 8] : 0.f;   // pixel row g + 8
            a[2 * hf] = stem_pack(v00, v01);
            a[2 * hf + 1] = stem_pack(v10, v11);
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) mma_m16n8k16_bf16(acc[j], a, bw[s][j][0], bw[s][j][1]);
        }
        // ---- bias, activation, bf16; quad transpose so that every lane owns eight CONSECUTIVE channels of a pixel -----
        uint32_t lo[4], hi[4];  // word j = channels 8j + 2tg, +1 of pixel g (lo) / g + 8 (hi)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float v0 = acc[j][0] + bs[j][0], v1 = acc[j][1] + bs[j][1], v2 = acc[j][2] + bs[j][0], v3 = acc[j][3] + bs[j][1];
          if (ACT == YC_MISH) { v0 = act_mish(v0); v1 = act_mish(v1); v2 = act_mish(v2); v3 = act_mish(v3); }
          else if (ACT == YC_LEAKY) { v0 = fmaxf(v0, 0.1f * v0); v1 = fmaxf(v1, 0.1f * v1); v2 = fmaxf(v2, 0.1f * v2); v3 = fmaxf(v3, 0.1f * v3); }
          lo[j] = stem_pack(v0, v1);
          hi[j] = stem_pack(v2, v3);
        }
        // lane tg sends word (tg ^ s) to lane tg ^ s and receives that lane's word tg = channels 8tg + 2(tg ^ s), +1
        uint32_t olo[4], ohi[4];
#pragma unroll
        for (int s = 0; s < 4; ++s) {
          const int src = tg ^ s;
          const uint32_t slo = src == 0 ? lo[0] : (src == 1 ? lo[1] : (src == 2 ? lo[2] : lo[3]));
          const uint32_t shi = src == 0 ? hi[0] : (src == 1 ? hi[1] : (src == 2 ? hi[2] : hi[3]));
          const uint32_t rlo = __shfl_xor_sync(0xffffffffu, slo, s), rhi = __shfl_xor_sync(0xffffffffu, shi, s);
          // received from lane p = tg ^ s: position p of this lane's eight channels
          if (src == 0) { olo[0] = rlo; ohi[0] = rhi; }
          else if (src == 1) { olo[1] = rlo; ohi[1] = rhi; }
          else if (src == 2) { olo[2] = rlo; ohi[2] = rhi; }
          else { olo[3] = rlo; ohi[3] = rhi; }
        }
        if (8 * tg < n_out) {
          const int wa = w0 + c0 + g, wb = wa + 8;
          __nv_bfloat16* row = y + ((long long)b * H + h) * W * ldy + 8 * tg;
          if (wa < W) *reinterpret_cast<uint4*>(row + (long long)wa * ldy) = make_uint4(olo[0], olo[1], olo[2], olo[3]);
          if (wb < W) *reinterpret_cast<uint4*>(row + (long long)wb * ldy) = make_uint4(ohi[0], ohi[1], ohi[2], ohi[3]);
        }
      }
    }
  }
}

}  // namespace yc

using namespace yc;

extern "C" int yc_stem_conv_direct(const float* x_nchw, int batch, int h, int w, const void* w_packed, const float* bias,
                                   int activation, const yc_tensor* y, void* stream) {
  YC_REQUIRE(x_nchw && w_packed && bias && y && y->ptr && batch > 0 && h > 0 && w > 0, "yc_stem_conv_direct: bad arguments");
  YC_REQUIRE(y->dtype == YC_BF16 && y->n == batch && y->h == h && y->w == w && y->c % 8 == 0 && y->c >= 8 && y->c <= 32 &&
                 y->ldc >= y->c && y->ldc % 8 == 0 && ((uintptr_t)y->ptr & 15) == 0,
             "yc_stem_conv_direct: the output must be an NHWC bf16 view of the input's size with 8, 16, 24 or 32 channels");
  YC_REQUIRE(activation == YC_MISH || activation == YC_LEAKY || activation == YC_LINEAR,
             "yc_stem_conv_direct: activation %d is not supported (mish / leaky / linear)", activation);
  const int th = ceil_div(h, kStemRows), tw = ceil_div(w, kStemCols);
  const long long tiles = (long long)batch * th * tw;
  YC_REQUIRE(tiles < (1ll << 31), "yc_stem_conv_direct: too many tiles");
  const int grid = (int)std::min<long long>(tiles, (long long)sm_count() * 2 * 8);
  cudaStream_t st = (cudaStream_t)stream;
  const __nv_bfloat16* wp = (const __nv_bfloat16*)w_packed;
  __nv_bfloat16* yp = (__nv_bfloat16*)y->ptr;
  if (activation == YC_MISH)
    stem_conv_direct_kernel<YC_MISH><<<grid, 256, 0, st>>>(x_nchw, batch, h, w, wp, bias, y->c, yp, y->ldc, th, tw);
  else if (activation == YC_LEAKY)
    stem_conv_direct_kernel<YC_LEAKY><<<grid, 256, 0, st>>>(x_nchw, batch, h, w, wp, bias, y->c, yp, y->ldc, th, tw);
  else
    stem_conv_direct_kernel<YC_LINEAR><<<grid, 256, 0, st>>>(x_nchw, batch, h, w, wp, bias, y->c, yp, y->ldc, th, tw);
  YC_LAUNCH_CHECK();
  return YC_OK;
}
