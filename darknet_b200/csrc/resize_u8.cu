// cv::resize(input, resize, Size(net->w, net->h)) of ProcImage (yolo_core.cpp:108) on the device, for 8-bit interleaved
// frames, bit-identical to the library's INTER_LINEAR path (OpenCV imgproc/src/resize.cpp; restated and pinned against
// cv2 4.13 in oracle/cv_resize.py / tests/golden/cv_resize.npz), optionally fused with the cvtColor(RGB2BGR) + Mat2Image
// that follow it (yolo_core.cpp:110-111, visualize.cpp:26-54): u8 HWC frame of any size -> fp32 NCHW network input.
//
// Fixed-point arithmetic of the library, per destination pixel:
//   x: fx = (float)((dx + 0.5) * scale_x - 0.5); sx = floor(fx); fx -= sx; sx < 0 -> (0, 0); sx >= sw - 1 -> (sw - 1, 0);
//      a0 = rint((1 - fx) * 2048), a1 = rint(fx * 2048)
//   y: fy, sy the same WITHOUT the clamp: the two row indices are clipped to [0, sh - 1] instead and the weights kept
//   row_k = S[y_k][sx] * a0 + S[y_k][sx + 1] * a1                      (11 fractional bits)
//   dst   = (((b0 * (row_0 >> 4)) >> 16) + ((b1 * (row_1 >> 4)) >> 16) + 2) >> 2
// and an exact 2 x 2 decimation is INTER_AREA in the library: (a + b + c + d + 2) >> 2.
// scale = 1.0 / ((double)dsize / ssize) is computed on the host in double like resize() does and handed to the kernel.
//
// Byte work bound by memory: one thread per destination pixel, 4 source pixels of <= 4 bytes each through the read-only
// path (neighbouring threads share sectors), fp32 planes or u8 pixels written coalesced along x.  Algorithmic bytes per
// destination pixel: c * (sh * sw) / (dh * dw) read + 4 c (fp32 NCHW) or c (u8) written.
#include "yc_common.cuh"

namespace yc {

struct ResizeParams {
  const unsigned char* src;
  int n, c, sh, sw, dh, dw;
  double scale_x, scale_y;
  int area2;  // exact 2 x 2 decimation
  int swap;   // fp32 output: read channel c - 1 - k
};

__device__ __forceinline__ void resize_coef(int d, double scale, int ssize, bool clamp, int& s0, int& s1, int& a0, int& a1) {
  // separate multiply and subtract, each rounded (no fused multiply-add): what the host library's double arithmetic does
  float f = (float)__dadd_rn(__dmul_rn((double)d + 0.5, scale), -0.5);
  int s = (int)floorf(f);
  f -= (float)s;
  if (clamp) {
    if (s < 0) { s = 0; f = 0.f; }
    if (s >= ssize - 1) { s = ssize - 1; f = 0.f; }
  }
  a0 = __float2int_rn(__fmul_rn(1.f - f, 2048.f));  // cvRound: to nearest, ties to even
  a1 = __float2int_rn(__fmul_rn(f, 2048.f));
  s0 = min(max(s, 0), ssize - 1);
  s1 = min(max(s + 1, 0), ssize - 1);
}

// OUT_F32: fp32 NCHW planes, value / 255.0f (Mat2Image); else u8 HWC
template <bool OUT_F32>
__global__ void resize_u8_kernel(const ResizeParams p, void* __restrict__ dst_) {
  const int dx = blockIdx.x * blockDim.x + threadIdx.x;
  const int dy = blockIdx.y * blockDim.y + threadIdx.y;
  const int b = blockIdx.z;
  if (dx >= p.dw || dy >= p.dh) return;
  const int c = p.c;
  const unsigned char* img = p.src + (long long)b * p.sh * p.sw * c;
  int v[4] = {0, 0, 0, 0};
  if (p.area2) {
    const unsigned char* r0 = img + ((long long)(2 * dy) * p.sw + 2 * dx) * c;
    const unsigned char* r1 = r0 + (long long)p.sw * c;
    for (int k = 0; k < c; ++k) v[k] = ((int)__ldg(r0 + k) + (int)__ldg(r0 + c + k) + (int)__ldg(r1 + k) + (int)__ldg(r1 + c + k) + 2) >> 2;
  } else {
    int x0, x1, a0, a1, y0, y1, b0, b1;
    resize_coef(dx, p.scale_x, p.sw, true, x0, x1, a0, a1);
    resize_coef(dy, p.scale_y, p.sh, false, y0, y1, b0, b1);
    const unsigned char* r0 = img + (long long)y0 * p.sw * c;
    const unsigned char* r1 = img + (long long)y1 * p.sw * c;
    for (int k = 0; k < c; ++k) {
      const int h0 = (int)__ldg(r0 + x0 * c + k) * a0 + (int)__ldg(r0 + x1 * c + k) * a1;
      const int h1 = (int)__ldg(r1 + x0 * c + k) * a0 + (int)__ldg(r1 + x1 * c + k) * a1;
      const int o = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
      v[k] = min(max(o, 0), 255);
    }
  }
  if (OUT_F32) {
    float* dst = (float*)dst_;
    const long long hw = (long long)p.dh * p.dw, s = (long long)dy * p.dw + dx;
    for (int k = 0; k < c; ++k) dst[((long long)b * c + k) * hw + s] = (float)v[p.swap ? c - 1 - k : k] / 255.0f;
  } else {
    unsigned char* dst = (unsigned char*)dst_ + (((long long)b * p.dh + dy) * p.dw + dx) * c;
    for (int k = 0; k < c; ++k) dst[k] = (unsigned char)v[k];
  }
}

static int resize_launch(const unsigned char* src, int n, int c, int sh, int sw, void* dst, int dh, int dw, int swap, bool f32,
                         cudaStream_t st, const char* who) {
  YC_REQUIRE(src && dst && n > 0 && c >= 1 && c <= 4 && sh > 0 && sw > 0 && dh > 0 && dw > 0, "%s: bad arguments (1..4 channels)", who);
  YC_REQUIRE(n <= 65535, "%s: at most 65535 images per call", who);
  ResizeParams p;
  p.src = src;
  p.n = n; p.c = c; p.sh = sh; p.sw = sw; p.dh = dh; p.dw = dw;
  // resize(): inv_scale = (double)dsize / ssize; scale = 1. / inv_scale
  p.scale_x = 1.0 / ((double)dw / (double)sw);
  p.scale_y = 1.0 / ((double)dh / (double)sh);
  p.area2 = (sw == 2 * dw && sh == 2 * dh) ? 1 : 0;
  p.swap = swap ? 1 : 0;
  const dim3 block(32, 8), grid(ceil_div(dw, 32), ceil_div(dh, 8), n);
  if (f32) resize_u8_kernel<true><<<grid, block, 0, st>>>(p, dst);
  else resize_u8_kernel<false><<<grid, block, 0, st>>>(p, dst);
  YC_LAUNCH_CHECK();
  return YC_OK;
}

}  // namespace yc

using namespace yc;

extern "C" int yc_resize_u8(const unsigned char* src, int n, int c, int sh, int sw, unsigned char* dst, int dh, int dw, void* stream) {
  if (sh == dh && sw == dw) {  // resize() copies
    YC_REQUIRE(src && dst && n > 0 && c > 0 && sh > 0 && sw > 0, "yc_resize_u8: bad arguments");
    YC_CUDA(cudaMemcpyAsync(dst, src, (size_t)n * sh * sw * c, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return YC_OK;
  }
  return resize_launch(src, n, c, sh, sw, dst, dh, dw, 0, false, (cudaStream_t)stream, "yc_resize_u8");
}

extern "C" int yc_resize_u8_to_nchw(const unsigned char* src, int n, int c, int sh, int sw, int swap_rb, float* dst, int dh, int dw,
                                    void* stream) {
  if (sh == dh && sw == dw) return yc_u8hwc_to_nchw(src, n, c, sh, sw, swap_rb, dst, stream);
  return resize_launch(src, n, c, sh, sw, dst, dh, dw, swap_rb, true, (cudaStream_t)stream, "yc_resize_u8_to_nchw");
}
