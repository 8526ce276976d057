/* yc_b200.h — C ABI of the B200-native conv / YOLO hot path (libyc_b200.so).
 *
 * Drop-in boundary for the reference's per-layer GPU operators (yolo-core, /root/reference/src).
 * Every entry point is extern "C", takes plain pointers / sizes / POD descriptors and an explicit
 * cudaStream_t (passed as void*), returns 0 on success or a negative yc_status, and is legal inside
 * CUDA-graph capture unless marked [setup].  No torch types, no `struct layer` below this line:
 * the trampolines that translate `layer` / `NetworkState` live in csrc/network.cpp (the host-side
 * mirror of network.cpp / parser.cpp) and in INTEGRATION.md (the binding a maintainer of the
 * reference would add).
 *
 * Data layout on the device (differs from the reference on purpose; converted only at the seam):
 *   activations  NHWC, bf16 (heads: fp32), described by yc_tensor = {ptr, n, h, w, c, ldc}
 *                where ldc (elements) >= c is the pixel pitch, so a tensor can be a channel slice
 *                of a wider concat buffer ([route] becomes a view, route_layer.c:125-142).
 *   weights      bf16, [out_c][kh][kw][in_c_padded]  (K-major rows for the implicit GEMM), packed
 *                once from the reference's OIHW fp32 (convolutional_layer.cpp:446) by yc_conv_pack.
 *   host-visible buffers at the seam (network input, YOLO outputs) stay NCHW fp32 like the reference.
 *
 * Each function cites the reference interface it replaces (file:line under /root/reference/src).
 */
#ifndef YC_B200_H
#define YC_B200_H
#include <stddef.h>
#include <stdint.h>
#if defined(__GNUC__)
#define YC_API __attribute__((visibility("default")))
#else
#define YC_API
#endif
#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  YC_OK = 0,
  YC_ERR_CUDA = -1,      /* a CUDA runtime/driver call failed; see yc_last_error() */
  YC_ERR_ARG = -2,       /* invalid argument / unsupported shape */
  YC_ERR_NO_DEVICE = -3, /* no sm_100 device: there is NO CPU fallback */
  YC_ERR_ALLOC = -4,
  YC_ERR_IO = -5,
  YC_ERR_PARSE = -6
} yc_status;

/* activation ids == the reference's enum ACTIVATION (yolo_core.h:69-94) */
enum { YC_LOGISTIC = 0, YC_RELU = 1, YC_LINEAR = 4, YC_LEAKY = 8, YC_SWISH = 16, YC_MISH = 17 };
enum { YC_BF16 = 0, YC_F32 = 1 };

/* NHWC tensor view. ptr already includes the channel offset of the view. */
typedef struct yc_tensor {
  void* ptr;
  int n, h, w, c;
  int ldc;   /* pixel pitch in elements (>= c; multiple of 8 for bf16, 4 for fp32) */
  int dtype; /* YC_BF16 | YC_F32 */
} yc_tensor;

/* ---- device runtime: replaces dark_cuda.c -------------------------------------------------- */
/* error text of the last failing call on this thread (check_error, dark_cuda.c:58-83, printed and
 * exit()ed; here the caller decides). */
YC_API const char* yc_last_error(void);
/* [setup] cuda_set_device (dark_cuda.c:36-41) + capability check (sm_100 required). */
YC_API int yc_set_device(int device);
YC_API int yc_get_device(void);                      /* cuda_get_device dark_cuda.c:43-49 */
YC_API int yc_device_count(void);                    /* CudaGetDeviceCount */
YC_API int yc_sm_count(void);
/* [setup] cuda_make_array (dark_cuda.c:416-438): device allocation, zero-filled. */
YC_API void* yc_malloc(size_t bytes);
YC_API int yc_free(void* p);                         /* cuda_free dark_cuda.c:519-523 */
YC_API void* yc_malloc_host(size_t bytes);           /* pinned: cuda_make_array_pinned / cudaHostAlloc */
YC_API int yc_free_host(void* p);
YC_API int yc_push(void* dst_dev, const void* src_host, size_t bytes, void* stream); /* cuda_push_array :525-531 */
YC_API int yc_pull(void* dst_host, const void* src_dev, size_t bytes, void* stream); /* cuda_pull_array_async :549-555 */
YC_API int yc_memset(void* dst_dev, int value, size_t bytes, void* stream);
YC_API int yc_stream_sync(void* stream);             /* cudaStreamSynchronize(get_cuda_stream()) */
YC_API int yc_stream_create(void** stream);          /* get_cuda_stream dark_cuda.c:128-152 */
YC_API int yc_stream_destroy(void* stream);
/* number of kernels this library has launched since the last reset (bench.py "gpu_launches"). */
YC_API long long yc_launch_count(int reset);

/* ---- layout conversion at the seam ----------------------------------------------------------- */
/* fp32 NCHW (reference layout, network_kernels.cu:513-514 input) -> NHWC view (bf16 or fp32). */
YC_API int yc_nchw_to_nhwc(const float* src, const yc_tensor* dst, void* stream);
/* NHWC view -> fp32 NCHW (what the reference's host code and other layers read). */
YC_API int yc_nhwc_to_nchw(const yc_tensor* src, float* dst, void* stream);

/* ---- convolution forward: replaces ForwardConvolutionalLayerGpu ------------------------------ */
/* (convolutional_kernels.cu:252-553: cuDNN / im2col_gpu_ext + cublasSgemm, add_bias_gpu,
 *  activate_array_ongpu, and the fused shortcut that follows, shortcut_layer.c:191-204)
 *  y = act2( act( conv(x, w) * scale + bias ) + residual )
 *  one implicit-GEMM kernel: TMA(im2col) -> smem -> tcgen05.mma -> TMEM -> fused epilogue. */
typedef struct yc_conv_desc {
  int ksize, stride, pad, dilation; /* square kernel; pad as in convolutional_layer.cpp:87-95 */
  int activation;                   /* act  : applied to conv*scale+bias                       */
  int residual_activation;          /* act2 : applied after adding the residual (YC_LINEAR usual) */
} yc_conv_desc;

typedef struct yc_conv_plan yc_conv_plan;

/* in_c padded to the K-chunk the kernel uses for this layer (16, 32 or a multiple of 64). */
YC_API int yc_conv_padded_in_c(int in_c);
/* bytes of the packed bf16 weight buffer for (out_c, in_c, ksize). */
YC_API size_t yc_conv_packed_bytes(int out_c, int in_c, int ksize);
/* [setup] OIHW fp32 (device) -> packed bf16 (device). Replaces PushConvolutionalLayer's weight
 * upload + cuda_f32_to_f16 shadow (convolutional_kernels.cu:838-863). */
YC_API int yc_conv_pack(const float* w_oihw_dev, int out_c, int in_c, int ksize, void* w_packed_dev, void* stream);
/* [setup] build the TMA descriptors + tile schedule for fixed buffers. x/y/residual are views;
 * residual may be NULL. bias has out_c floats; scale may be NULL (BN already folded). */
YC_API int yc_conv_plan_create(yc_conv_plan** plan, const yc_conv_desc* desc, const yc_tensor* x,
                        const void* w_packed, const float* bias, const float* scale,
                        const yc_tensor* residual, const yc_tensor* y);
YC_API int yc_conv_plan_destroy(yc_conv_plan* plan);
/* launch (1 kernel). */
YC_API int yc_conv_fprop(const yc_conv_plan* plan, void* stream);
/* plan introspection for tests/bench: what=0 block_n, 1 block_k, 2 stages, 3 tiles, 4 grid */
YC_API int yc_conv_plan_info(const yc_conv_plan* plan, int what);

/* ---- memory-bound layers (128-bit vectorised NHWC) ------------------------------------------- */
/* ForwardMaxpoolLayerGpu (maxpool_layer_kernels.cu:58-101,145-201). pad = total padding
 * (parser.cpp:646 default size-1), window offset -pad/2. ref_cpu_quirk=1 reproduces the reference
 * CPU inference routine forward_maxpool_layer_avx's lane-0-only validity test (gemm.c:2127-2150),
 * needed only for bit-parity with the reference *CPU inference* path; 0 = maxpool_layer.cpp:258-288
 * == the reference GPU kernel. */
YC_API int yc_maxpool_fwd(const yc_tensor* x, const yc_tensor* y, int size, int stride, int pad,
                   int ref_cpu_quirk, void* stream);
/* ForwardUpsampleLayerGpu (upsample_layer.c:106-119, blas_kernels.cu:1121-1155): nearest x stride */
YC_API int yc_upsample_fwd(const yc_tensor* x, const yc_tensor* y, int stride, float scale, void* stream);
/* ForwardShortcutLayerGpu (shortcut_layer.c:191-204, blas_kernels.cu:941-978): y = act(a + b) */
YC_API int yc_shortcut_fwd(const yc_tensor* a, const yc_tensor* b, const yc_tensor* y, int activation, void* stream);
/* ForwardRouteLayerGpu's simple_copy_ongpu (route_layer.c:125-142): copy a view into a view. */
YC_API int yc_route_copy(const yc_tensor* src, const yc_tensor* dst, void* stream);
/* ForwardScaleChannelsLayerGpu (scale_channels_layer.c:134-146): y = act(x * s[n,c]) ; s is [n,1,1,c] */
YC_API int yc_scale_channels_fwd(const yc_tensor* x, const yc_tensor* s, const yc_tensor* y, int activation, void* stream);
/* ForwardReorgLayerGpu (reorg_layer.c:113-127, blas_kernels.cu:368-402) on the reference's NCHW
 * index math, expressed over NHWC views. forward as in reorg_cpu (blas.c:12-46). */
YC_API int yc_reorg_fwd(const yc_tensor* x, const yc_tensor* y, int stride, int forward, void* stream);

/* ---- YOLO head: replaces ForwardYoloLayerGpu ------------------------------------------------- */
/* (yolo_layer.cpp:836-859: copy + per-(b,anchor) logistic / scal_add launches.)
 * x: head conv output NHWC [n, h, w, anchors*(5+classes)] (fp32 or bf16);
 * out_nchw: fp32 [n][anchors*(5+classes)][h][w] == the reference's l->output layout (EntryIndex,
 * yolo_layer.cpp:380-386). gaussian=1: Gaussian_yolo with 9+classes entries
 * (gaussian_yolo_layer.cpp:934-961). */
YC_API int yc_yolo_decode(const yc_tensor* x, float* out_nchw, int anchors, int classes, float scale_x_y,
                   int gaussian, void* stream);

/* On-device detection extraction (GetYoloDetections for every batch index, yolo_layer.cpp:794-832):
 * scans the decoded NCHW head in the reference's order (anchor n, then cell i), keeps cells with
 * objectness > thresh and writes rows [batch, x, y, w, h, obj, prob[classes]] compactly, ordered
 * exactly like the reference within each image. *count_dev receives the number of rows per image
 * ([n] ints). rows_per_image bounds the per-image output. */
YC_API int yc_yolo_extract(const float* decoded_nchw, int n, int h, int w, int anchors, int classes,
                    const float* anchor_wh, const int* mask, int net_w, int net_h, float thresh,
                    int gaussian, float* rows_dev, int rows_per_image, int* count_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif
