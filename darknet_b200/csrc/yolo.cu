// YOLO / Gaussian-YOLO head: decode and on-device detection extraction.
//
// yc_yolo_decode replaces ForwardYoloLayerGpu's copy + B*n*(2..3) strided activation launches
// (yolo_layer.cpp:836-853; Gaussian: gaussian_yolo_layer.cpp:934-961) with ONE transposing kernel that reads
// the head conv's NHWC output, applies logistic / scale_x_y per entry and writes the reference's NCHW fp32
// `l->output` layout (EntryIndex, yolo_layer.cpp:380-386).
//
// yc_yolo_extract is GetYoloDetections (yolo_layer.cpp:794-832) for EVERY batch index, on the device,
// emitting rows in the reference's scan order (anchor, then cell) so indices are comparable.
#include "yc_common.cuh"

namespace yc {

struct HeadView {
  const void* ptr;
  long long ldc;
  int f32;
};

__device__ __forceinline__ float head_ld(const HeadView& t, long long pix, int c) {
  return t.f32 ? reinterpret_cast<const float*>(t.ptr)[pix * t.ldc + c]
               : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(t.ptr)[pix * t.ldc + c]);
}

// accurate logistic for the head (few elements; feeds thresholds): expf, not __expf
__device__ __forceinline__ float logistic_acc(float x) { return 1.f / (1.f + expf(-x)); }

// 32 pixels x (up to) 256 channels per block through shared memory: the NHWC read is coalesced along channels (one
// 128-byte request per warp, eight independent requests per pixel in flight per thread), the NCHW write along
// pixels.  The tile pitch of 257 words makes both the row-wise stores and the column-wise loads conflict-free.
// Algorithmic bytes: 4 B read + 4 B written per output (fp32 head); the exact expf and division are kept because
// the decoded objectness feeds the detection threshold (parity of the detection set with the reference).
constexpr int kDecPix = 32, kDecCh = 256, kDecPitch = kDecCh + 1;
__global__ void __launch_bounds__(256) yolo_decode_kernel(HeadView x, float* __restrict__ out, int S, int C, int E,
                                                          float sxy, int gaussian) {
  extern __shared__ float tile[];  // [kDecPix][kDecPitch]
  const int b = blockIdx.z;
  const int p0 = blockIdx.x * kDecPix, c0 = blockIdx.y * kDecCh;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float bias = -0.5f * (sxy - 1.f);
  const int nch = min(kDecCh, C - c0);
  // which of this lane's eight channels (c0 + lane + 32 k) take the logistic / the scale_x_y affine
  unsigned sig_mask = 0, aff_mask = 0;
#pragma unroll
  for (int k = 0; k < kDecCh / 32; ++k) {
    const int e = (c0 + lane + 32 * k) % E;
    bool sig, aff;
    if (gaussian) {  // gaussian_yolo_layer.cpp:497-516: x,sx,y,sy,_,sw,_,sh, obj, classes
      sig = (e <= 3) || e == 5 || e == 7 || e >= 8;
      aff = (e == 0) || (e == 2);
    } else {         // yolo_layer.cpp:399-404
      sig = (e < 2) || (e >= 4);
      aff = (e < 2);
    }
    sig_mask |= (unsigned)sig << k;
    aff_mask |= (unsigned)aff << k;
  }
  for (int i = warp; i < kDecPix; i += 8) {
    const int p = p0 + i;
    if (p >= S) break;
    float v[kDecCh / 32];
#pragma unroll
    for (int k = 0; k < kDecCh / 32; ++k) {
      const int c = lane + 32 * k;
      v[k] = c < nch ? head_ld(x, (long long)b * S + p, c0 + c) : 0.f;
    }
#pragma unroll
    for (int k = 0; k < kDecCh / 32; ++k) {
      float y = v[k];
      if ((sig_mask >> k) & 1u) y = logistic_acc(y);
      if ((aff_mask >> k) & 1u) y = y * sxy + bias;
      tile[i * kDecPitch + lane + 32 * k] = y;
    }
  }
  __syncthreads();
  const int p = p0 + lane;
  if (p < S) {
    float* o = out + ((long long)b * C + c0) * S + p;
    for (int c = warp; c < nch; c += 8) o[(long long)c * S] = tile[lane * kDecPitch + c];
  }
}

// one block per image: ordered compaction by block-wide exclusive scan over (anchor, cell)
__global__ void yolo_extract_kernel(const float* __restrict__ pred, int S, int W, int H, int A, int classes, int E,
                                    const float* __restrict__ anchor_wh, const int* __restrict__ mask, int net_w,
                                    int net_h, float thresh, int gaussian, float* __restrict__ rows,
                                    int rows_per_image, int* __restrict__ counts, int R, float tag) {
  __shared__ int warp_sums[32];
  __shared__ int base_s;
  const int b = blockIdx.x;
  // row = [batch x y w h obj (tag) probs...]; probs occupy the last `classes` columns of the R-wide row
  const int pc = R - classes;
  const float* p = pred + (long long)b * A * E * S;
  const int total = A * S;
  const int obj_e = gaussian ? 8 : 4;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
  if (threadIdx.x == 0) base_s = 0;
  __syncthreads();
  for (int start = 0; start < total; start += blockDim.x) {
    const int loc = start + threadIdx.x;
    int a = 0, i = 0;
    float obj = 0.f;
    bool keep = false;
    if (loc < total) {
      a = loc / S;
      i = loc - a * S;
      obj = p[((long long)a * E + obj_e) * S + i];
      keep = obj > thresh;  // yolo_layer.cpp:807-809
    }
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    const int within = __popc(bal & ((1u << lane) - 1u));
    if (lane == 0) warp_sums[wid] = __popc(bal);
    __syncthreads();
    int woff = 0, chunk_total = 0;
    for (int w = 0; w < nw; ++w) {
      const int s = warp_sums[w];
      if (w < wid) woff += s;
      chunk_total += s;
    }
    const int base = base_s;
    const int pos = base + woff + within;
    if (keep && pos < rows_per_image) {
      const float* q = p + (long long)a * E * S + i;
      float* o = rows + ((long long)b * rows_per_image + pos) * R;
      const int col = i % W, row = i / W;
      const int m = mask[a];
      float bx, by, bw, bh, scale = 1.f;
      if (gaussian) {  // GetGaussianYoloBox gaussian_yolo_layer.cpp:151-160 ; uncertainty :893-900
        bx = (col + q[0]) / W;
        by = (row + q[2 * (long long)S]) / H;
        bw = (float)(exp((double)q[4 * (long long)S]) * anchor_wh[2 * m] / net_w);
        bh = (float)(exp((double)q[6 * (long long)S]) * anchor_wh[2 * m + 1] / net_h);
        const float uc = (q[1 * (long long)S] + q[3 * (long long)S] + q[5 * (long long)S] + q[7 * (long long)S]) / 4.f;
        scale = 1.f - uc;
      } else {         // GetYoloBox yolo_layer.cpp:139-148 (exp in double, as the reference)
        bx = (col + q[0]) / W;
        by = (row + q[(long long)S]) / H;
        bw = (float)(exp((double)q[2 * (long long)S]) * anchor_wh[2 * m] / net_w);
        bh = (float)(exp((double)q[3 * (long long)S]) * anchor_wh[2 * m + 1] / net_h);
      }
      o[0] = (float)b; o[1] = bx; o[2] = by; o[3] = bw; o[4] = bh; o[5] = obj;
      if (pc > 6) o[6] = tag;
      for (int j = 0; j < classes; ++j) {
        const float pr = obj * q[(long long)(obj_e + 1 + j) * S] * scale;
        o[pc + j] = (pr > thresh) ? pr : 0.f;  // yolo_layer.cpp:820-826
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) base_s = base + chunk_total;
    __syncthreads();
  }
  if (threadIdx.x == 0) counts[b] = base_s;
}

// Dense gather of the per-(head,image) row regions into one [total][R] array ordered by image, then
// head (the reference's order within an image: FillNetworkBoxes walks layers in net order,
// network.cpp:473-496).  counts is [heads][n]; every block recomputes the (tiny) exclusive scan.
__global__ void yolo_gather_kernel(const float* __restrict__ regions, const int* __restrict__ counts, int heads,
                                   int n, int cap, int R, float* __restrict__ dense, int* __restrict__ img_counts,
                                   int dense_cap) {
  const int b = blockIdx.x / heads, hd = blockIdx.x % heads;
  int off = 0;
  for (int bb = 0; bb <= b; ++bb)
    for (int hh = 0; hh < heads; ++hh) {
      if (bb == b && hh >= hd) break;
      off += min(counts[hh * n + bb], cap);
    }
  const int cnt = min(counts[hd * n + b], cap);
  if (hd == 0 && threadIdx.x == 0) {
    int t = 0;
    for (int hh = 0; hh < heads; ++hh) t += min(counts[hh * n + b], cap);
    img_counts[b] = t;
  }
  const float* src = regions + ((long long)(hd * n + b) * cap) * R;
  const long long total = (long long)cnt * R;
  for (long long i = threadIdx.x; i < total; i += blockDim.x) {
    const long long row = off + i / R;
    if (row < dense_cap) dense[(long long)off * R + i] = src[i];
  }
}

int launch_yolo_extract(const float* decoded, int n, int h, int w, int anchors, int classes, const float* anchor_wh,
                        const int* mask, int net_w, int net_h, float thresh, int gaussian, float* rows,
                        int rows_per_image, int* counts, int R, float tag, cudaStream_t stream) {
  const int E = (gaussian ? 9 : 5) + classes;
  yolo_extract_kernel<<<n, 256, 0, stream>>>(decoded, h * w, w, h, anchors, classes, E, anchor_wh, mask, net_w, net_h,
                                             thresh, gaussian, rows, rows_per_image, counts, R, tag);
  YC_LAUNCH_CHECK();
  return YC_OK;
}
int launch_yolo_gather(const float* regions, const int* counts, int heads, int n, int cap, int R, float* dense,
                       int* img_counts, int dense_cap, cudaStream_t stream) {
  yolo_gather_kernel<<<heads * n, 256, 0, stream>>>(regions, counts, heads, n, cap, R, dense, img_counts, dense_cap);
  YC_LAUNCH_CHECK();
  return YC_OK;
}

}  // namespace yc

using namespace yc;

extern "C" int yc_yolo_decode(const yc_tensor* x, float* out, int anchors, int classes, float sxy, int gaussian,
                              void* stream) {
  YC_REQUIRE(x && x->ptr && out, "yc_yolo_decode: null argument");
  const int E = (gaussian ? 9 : 5) + classes;
  YC_REQUIRE(x->c == anchors * E, "yc_yolo_decode: channels %d != anchors*(%d+classes)=%d (parser.cpp:323-330)", x->c,
             gaussian ? 9 : 5, anchors * E);
  const int S = x->h * x->w;
  dim3 grid(ceil_div(S, kDecPix), ceil_div(x->c, kDecCh), x->n), block(256);
  HeadView hv{x->ptr, x->ldc, x->dtype == YC_F32};
  yolo_decode_kernel<<<grid, block, kDecPix * kDecPitch * sizeof(float), (cudaStream_t)stream>>>(hv, out, S, x->c, E, sxy,
                                                                                               gaussian);
  YC_LAUNCH_CHECK();
  return YC_OK;
}

extern "C" int yc_yolo_extract(const float* decoded, int n, int h, int w, int anchors, int classes,
                               const float* anchor_wh, const int* mask, int net_w, int net_h, float thresh,
                               int gaussian, float* rows, int rows_per_image, int* counts, void* stream) {
  YC_REQUIRE(decoded && anchor_wh && mask && rows && counts && n > 0, "yc_yolo_extract: null argument");
  return launch_yolo_extract(decoded, n, h, w, anchors, classes, anchor_wh, mask, net_w, net_h, thresh, gaussian, rows,
                             rows_per_image, counts, 6 + classes, 0.f, (cudaStream_t)stream);
}
