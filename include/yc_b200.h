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
/* fp32 NCHW image -> the first convolution's GEMM operand: bf16 [n][oh][ow][32], column (kh*size+kw)*c + ci, zero
 * padding and columns >= c*size*size written as zeros (needs c*size*size <= 32).  The stem then runs as a 1x1
 * convolution over 32 channels; replaces im2col_ongpu (im2col_kernels.cu) for that layer. */
YC_API int yc_stem_im2col(const float* src, int c, int h, int w, int size, int stride, int pad, const yc_tensor* dst,
                          void* stream);
/* The first convolution computed straight from the fp32 NCHW network input: 3 channels -> y->c (8, 16, 24 or 32)
 * channels, 3x3 / stride 1 / pad 1, + bias + activation (mish / leaky / linear) -> NHWC bf16 view y, in one kernel that
 * keeps no operand in global memory.  Replaces, for layer 0, the input layout step (network_kernels.cu:513),
 * im2col_gpu_ext + gemm (convolutional_kernels.cu:252-553, im2col_kernels.cu) and add_bias / activate_array.
 * w_packed: bf16 [out_c][32], column (kh*3+kw)*3 + ci (= yc_conv_pack of the filter re-ordered to [n][32][1][1]);
 * inputs are rounded to bf16 like every other layer's, accumulation is fp32. */
YC_API int yc_stem_conv_direct(const float* x_nchw, int batch, int h, int w, const void* w_packed, const float* bias,
                               int activation, const yc_tensor* y, void* stream);
/* The same layer, same arguments, on the tensor cores of sm_100a: builder warps assemble each 128-pixel tile's [128][32]
 * bf16 operand in shared memory (64B-swizzled K-major rows), one lane issues tcgen05.mma against the filters resident in
 * shared memory, accumulators in TMEM, epilogue warps add bias + activation and TMA-store NHWC bf16.  No operand in global
 * memory: 12 B read + 2*ldc B written per pixel. */
YC_API int yc_stem_conv_tc(const float* x_nchw, int batch, int h, int w, const void* w_packed, const float* bias,
                           int activation, const yc_tensor* y, void* stream);
/* Mat2Image on the device (visualize.cpp:26-54): u8 HWC [n][h][w][c] -> fp32 NCHW, value/255.0f; swap_rb reads
 * channel c-1-k (the cvtColor(RGB2BGR) of yolo_core.cpp:110). */
YC_API int yc_u8hwc_to_nchw(const unsigned char* src, int n, int c, int h, int w, int swap_rb, float* dst, void* stream);
/* cv::resize(src, dst, Size(dw, dh)) — default INTER_LINEAR — for n 8-bit interleaved images (1..4 channels) on the device,
 * bit-identical to the library (ProcImage's two resizes, yolo_core.cpp:108-109; OpenCV imgproc/src/resize.cpp: 11-bit fixed-
 * point coefficients, the (b*(row>>4))>>16 vertical pass, INTER_AREA for an exact 2x2 decimation; pinned against cv2 4.13 by
 * tests/golden/cv_resize.npz).  yc_resize_u8_to_nchw fuses the cvtColor(RGB2BGR) + Mat2Image that follow (swap_rb, /255.0f,
 * fp32 NCHW).  Device pointers. */
YC_API int yc_resize_u8(const unsigned char* src, int n, int c, int sh, int sw, unsigned char* dst, int dh, int dw, void* stream);
YC_API int yc_resize_u8_to_nchw(const unsigned char* src, int n, int c, int sh, int sw, int swap_rb, float* dst_nchw, int dh,
                                int dw, void* stream);
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
/* Plan-time switches (yc_conv_plan_create_ex; yc_network_load takes them shifted left by YC_NET_PLAN_SHIFT): 0 = the
 * measured-best defaults.  They exist for A/B measurements and so that a parity run can cover every kernel variant;
 * yc_conv_plan_info(plan, 7) reports the variant a plan actually uses. */
enum {
  YC_PLAN_NO_HALO = 1,       /* never stage 3x3/s1/p1 inputs as halo patches (nine im2col tap loads instead) */
  YC_PLAN_FORCE_HALO = 2,    /* halo patches whenever legal, whatever the cost rule says */
  YC_PLAN_NO_HALO_PAIR = 4,  /* halo mode: one tile at a time instead of tile pairs */
  YC_PLAN_NO_W_RES = 8,      /* never keep the weight slabs resident in shared memory */
  YC_PLAN_W_RES_1X1 = 16,    /* keep them resident for 1x1 layers as well */
  YC_PLAN_HEAD_DECODE = 32,  /* network level, opt-in: fold the YOLO decode into the head convolution's epilogue
                                (yc_conv_plan_set_yolo_decode) instead of running it as a kernel of its own.  Measured: the three
                                launches it removes cost 246 us, the epilogues grow by 178 us (16 warps per SM doing the
                                transposed stores), the step does not move (11.22 / 11.05 against 11.06 / 11.17 ms) and the conv
                                time — the roofline's denominator — grows: off by default.  (Bit 32 was YC_PLAN_ONE_OUT_BUF, a
                                no-op since one staging buffer became the default.) */
  YC_PLAN_ONE_OUT_BUF = 32,
  YC_PLAN_COST_V1 = 64,      /* round-1 tile-width cost model */
  YC_PLAN_TWO_OUT_BUFS = 128, /* two staging buffers wherever they fit */
  YC_PLAN_GENERIC_EPI = 256, /* never use the specialised (bf16 / bias-only / mish-leaky-linear) epilogue */
  YC_PLAN_EPI_SETS_V1 = 512, /* specialised epilogue with the generic one's warp-to-tile assignment */
  YC_PLAN_NO_TILE_PAIR = 2048, /* tiled / im2col path: one tile at a time instead of tile pairs (two accumulators) */
  YC_PLAN_MISH_NEWTON = 1024, /* specialised epilogue: mish with one MUFU + a Newton reciprocal on the FMA pipe (measured slower) */
  YC_PLAN_NO_CTA_PAIR = 4096, /* never run a layer on CTA pairs (tcgen05 cta_group::2, 256-row tiles over a cluster of two) */
  YC_PLAN_CTA_PAIR_ALL = 8192, /* CTA pairs wherever legal (tiled / im2col operands, specialised epilogue), not only for 256-wide tiles */
  YC_PLAN_PDL = 16384        /* launch with programmatic stream serialization: the kernel's set-up (barriers, TMEM, tensor-map
                                prefetch) overlaps the tail of the kernel before it in the stream; it waits
                                (griddepcontrol.wait) before touching anything that kernel wrote.  yc_network_load sets it
                                for inference plans under YC_NET_PDL (opt-in: no measured gain). */
};
/* [setup] yc_conv_plan_create with plan switches and, optionally, a SECOND output tensor y2: the GEMM then has
 * y->c + y2->c output channels (weights / bias / scale of both layers concatenated along out_c, y->c a multiple of
 * 32) and columns >= y->c are stored to y2.  Two 1x1 convolutions that read the same tensor (the two branches of a
 * CSP block, e.g. yolov4 layers 2 and 4) run as one kernel and read their input once.  residual must be NULL then. */
YC_API int yc_conv_plan_create_ex(yc_conv_plan** plan, const yc_conv_desc* desc, const yc_tensor* x,
                           const void* w_packed, const float* bias, const float* scale,
                           const yc_tensor* residual, const yc_tensor* y, const yc_tensor* y2, unsigned plan_flags);
YC_API int yc_conv_plan_destroy(yc_conv_plan* plan);
/* launch (1 kernel). */
YC_API int yc_conv_fprop(const yc_conv_plan* plan, void* stream);
/* (A/B) conv plans created from now on issue the nine taps of a halo tile unrolled with constant descriptor offsets (1, the
 * default) or through the rolled loop (0).  Process-wide like yc_conv_wgrad_set_tap_groups. */
YC_API int yc_conv_fprop_set_issue_v2(int on);
/* plan introspection for tests/bench: what=0 block_n, 1 block_k, 2 stages, 3 tiles, 4 grid, 5 A-operand mode (0 tiled,
 * 1 im2col, 2 halo patch), 6 dynamic shared memory bytes, 7 variant bits: [0,2) A mode | 4 halo tile pairs | 8 resident
 * weights | 16 two outputs | 32 two staging buffers | 64 rectangular (pixel-pair) geometry  | 128 specialised epilogue | 256 tile pairs on the tiled / im2col path;
 * 8 epilogue warp sets */
YC_API int yc_conv_plan_info(const yc_conv_plan* plan, int what);

/* ---- convolution training path: replaces ForwardConvolutionalLayerGpu(train) / BackwardConvolutionalLayerGpu --- */
/* (convolutional_kernels.cu:252-553 with state.train, 555-815; CPU twins convolutional_layer.cpp:1128-1380,
 *  batchnorm_layer.cpp:92-166,206-255.)  Semantics follow the reference CPU path.
 *
 * Forward of a batch-normalised layer in training:
 *   z = conv(x, w)                       yc_conv_fprop on a plan with activation LINEAR, bias NULL, and
 *   sum_c z, sum_c z^2                   yc_conv_plan_set_stats(): the statistics of fast_mean_gpu /
 *                                        fast_variance_gpu are reduced inside the conv epilogue and WRITTEN to
 *                                        sum / sumsq by the launch (fixed summation order, no float atomics: two
 *                                        launches on the same input give the same bits)
 *   mean, var, rolling update            yc_bn_finalize (unbiased variance, rolling = .9 roll + .1 batch)
 *   y = act(gamma*(z-mean)/sqrt(var+1e-6)+beta)   yc_bn_act_fwd */
YC_API int yc_conv_plan_set_stats(yc_conv_plan* plan, float* sum, float* sumsq);
/* Optional: fold yc_bn_finalize into the statistics plan's kernel (the CTA that adds the per-CTA partials also writes
 * mean / var and updates the rolling averages; same arithmetic, one launch less per layer).  count = batch*out_h*out_w. */
YC_API int yc_conv_plan_set_bn_finalize(yc_conv_plan* plan, long long count, float* mean, float* var, float* rolling_mean,
                                        float* rolling_var);
/* A plan keeps a padded copy of the bias it was created with; after the bias changes (update_convolutional_layer on a
 * layer without batchnorm: the YOLO heads, convolutional_layer.cpp:1382-1401 axpy on l.biases) hand it the new values.
 * Asynchronous on `stream`. */
YC_API int yc_conv_plan_set_bias(yc_conv_plan* plan, const float* bias_dev, void* stream);
/* ForwardYoloLayerGpu (yolo_layer.cpp:836-853, gaussian_yolo_layer.cpp:934-961) folded into the head convolution: the
 * plan (fp32 output, linear, 1x1 / im2col operand) applies logistic / scale_x_y per entry and writes the [yolo] layer's
 * NCHW fp32 output `out_nchw` [batch][anchors*entries][h*w] itself; its own output tensor is no longer written. */
YC_API int yc_conv_plan_set_yolo_decode(yc_conv_plan* plan, float* out_nchw, int entries, float scale_x_y, int gaussian);
YC_API int yc_bn_finalize(const float* sum, const float* sumsq, long long count, float* mean, float* var,
                          float* rolling_mean, float* rolling_var, int n, void* stream);
YC_API int yc_bn_act_fwd(const yc_tensor* z, const float* mean, const float* var, const float* gamma,
                         const float* beta, int activation, const yc_tensor* y, void* stream);
/* Backward of a batch-normalised layer up to the GEMMs: gradient_array_ongpu + BackwardBatchnormLayerGpu
 * (batchnorm_layer.cpp:240-255: backward_scale, scale_bias, mean_delta, variance_delta, normalize_delta).
 * dy = l->delta (NHWC bf16), z = saved pre-batchnorm conv output; dz (bf16, channels >= dy.c, pad lanes zeroed)
 * receives the gradient the dgrad / wgrad GEMMs read.  scale_updates is accumulated into; bias_updates may be NULL
 * (the reference CPU path leaves the batchnorm shift without gradient) or receives sum(dy*act').
 * workspace: yc_bn_bwd_workspace_floats(c) floats, ZERO-FILLED once when allocated (it holds the two sums, three
 * coefficient vectors, a ticket and one partial per block: the sums are formed in a fixed order, without float atomics,
 * so the backward pass is reproducible bit for bit). */
YC_API size_t yc_bn_bwd_workspace_floats(int c);
YC_API int yc_bn_bwd(const yc_tensor* dy, const yc_tensor* z, const float* mean, const float* var,
                     const float* gamma, const float* beta, int activation, float* scale_updates,
                     float* bias_updates, float* workspace, const yc_tensor* dz, void* stream);
/* Layers without batchnorm: dz = dy * act'(y) on the layer output y (leaky/relu/logistic/linear), bias_updates +=
 * per-channel sums (gradient_array + backward_bias, convolutional_layer.cpp:1327-1336).  y may be NULL for LINEAR.
 * workspace: as for yc_bn_bwd (needed when bias_updates is given). */
YC_API int yc_bias_act_bwd(const yc_tensor* dy, const yc_tensor* y, int activation, float* bias_updates,
                           float* workspace, const yc_tensor* dz, void* stream);
/* yc_bias_act_bwd for callers that cannot allocate device memory (code under CUDA-graph capture, serving loops,
 * harnesses that forbid cudaMalloc).  NO counterpart in the reference.
 * Same arguments and results as yc_bias_act_bwd without the workspace: the reduction's scratch (ticket + one partial
 * per block) is a file-scope static __device__ array of the library (2.4 MB, held by every GPU that loads the library
 * until exit).  Limits: dy->c <= yc_bias_act_bwd_noalloc_max_channels() (1024) and a reduction grid of at most 296
 * blocks; larger arguments return YC_ERR_ARG.  bias_updates is required.  The call allocates and frees nothing.
 * Every call on a device shares the one array: calls that overlap on the GPU (different streams) race. */
YC_API size_t yc_bias_act_bwd_noalloc_max_channels(void);
YC_API int yc_bias_act_bwd_noalloc(const yc_tensor* dy, const yc_tensor* y, int activation, float* bias_updates,
                                   const yc_tensor* dz, void* stream);

/* Data gradient (gemm(1,0,...) + col2im_gpu_ext, convolutional_kernels.cu:765-802): runs through the forward
 * implicit-GEMM kernel on the flipped, channel-transposed filter.  yc_conv_dgrad_channels(out_c) is the channel
 * count (and minimum pixel pitch) the dz view must have; strided layers first zero-insert dz
 * (yc_conv_dgrad_dilate into a buffer of yc_conv_dgrad_dilated_size).  accumulate=0 overwrites dx like col2im
 * (SURVEY B6), accumulate=1 adds to it (cuDNN beta=1). Launch with yc_conv_fprop(plan). */
YC_API int yc_conv_dgrad_channels(int out_c);
YC_API size_t yc_conv_dgrad_packed_bytes(int out_c, int in_c, int ksize);
YC_API int yc_conv_pack_dgrad(const float* w_oihw_dev, int out_c, int in_c, int ksize, void* w_packed_dev, void* stream);
YC_API int yc_conv_dgrad_dilated_size(const yc_conv_desc* desc, int in_h, int in_w, int* hu, int* wu);
YC_API int yc_conv_dgrad_dilate(const yc_tensor* dz, int stride, const yc_tensor* dz_dilated, void* stream);
YC_API int yc_conv_dgrad_plan_create(yc_conv_plan** plan, const yc_conv_desc* desc, const yc_tensor* dz_s1,
                                     const void* w_packed_dgrad, const yc_tensor* dx, int accumulate);

/* Weight gradient (im2col_gpu_ext + gemm(0,1,...) per image accumulating into weight_updates_gpu,
 * convolutional_kernels.cu:741-763): one tcgen05 implicit-GEMM kernel, K = output pixels split across CTAs,
 * fp32 partial sums ADDED to dw_packed ([out_c][kh][kw][yc_conv_wgrad_ld(in_c)] fp32) — the reference's
 * accumulate-across-images/subdivisions semantics. */
typedef struct yc_wgrad_plan yc_wgrad_plan;
YC_API int yc_conv_wgrad_ld(int in_c);
YC_API size_t yc_conv_wgrad_bytes(int out_c, int in_c, int ksize);
YC_API int yc_conv_wgrad_plan_create(yc_wgrad_plan** plan, const yc_conv_desc* desc, const yc_tensor* x,
                                     const yc_tensor* dz, int out_c, float* dw_packed);
YC_API int yc_conv_wgrad_plan_destroy(yc_wgrad_plan* plan);
YC_API int yc_conv_wgrad(const yc_wgrad_plan* plan, void* stream);
/* what=0 block_n, 1 K splits, 2 stages, 3 tiles, 4 grid, 5 K blocks per split */
YC_API int yc_conv_wgrad_plan_info(const yc_wgrad_plan* plan, int what);
/* Plan-time switch for the plans created afterwards (A/B measurements; parity coverage of both schedules): 1 (default) = a
 * 3x3 layer's work units cover a kernel row or the whole filter and share the dY boxes, 0 = one tap per unit.  Returns
 * the previous setting.  yc_conv_wgrad_plan_info(plan, 6) = taps per unit, (plan, 7) = accumulator sets in TMEM. */
YC_API int yc_conv_wgrad_set_tap_groups(int on);
/* packed fp32 gradient <-> the reference's OIHW weight_updates layout */
YC_API int yc_conv_wgrad_unpack(const float* dw_packed, int out_c, int in_c, int ksize, float* dw_oihw, void* stream);
YC_API int yc_conv_wgrad_pack(const float* dw_oihw, int out_c, int in_c, int ksize, float* dw_packed, void* stream);

/* ---- memory-bound layers (128-bit vectorised NHWC) ------------------------------------------- */
/* ForwardMaxpoolLayerGpu (maxpool_layer_kernels.cu:58-101,145-201). pad = total padding
 * (parser.cpp:646 default size-1), window offset -pad/2. ref_cpu_quirk=1 reproduces the reference
 * CPU inference routine forward_maxpool_layer_avx's lane-0-only validity test (gemm.c:2127-2150),
 * needed only for bit-parity with the reference *CPU inference* path; 0 = maxpool_layer.cpp:258-288
 * == the reference GPU kernel. */
YC_API int yc_maxpool_fwd(const yc_tensor* x, const yc_tensor* y, int size, int stride, int pad,
                   int ref_cpu_quirk, void* stream);
/* Up to three stride-1 "same" max-pools of ONE bf16 tensor in one launch (the SPP block, maxpool 5 / 9 / 13 of yolov4 /
 * yolov3-spp: three forward_maxpool_layer_gpu calls on the same input, maxpool_layer_kernels.cu:58-101): the input slab is
 * staged in shared memory once.  ys[o] may be channel slices of one concat buffer.  Results are identical to yc_maxpool_fwd.
 * yc_maxpool_s1_multi_eligible: 1 when (x, y, size, stride, pad) can be one of the outputs. */
YC_API int yc_maxpool_s1_multi(const yc_tensor* x, int n_out, const yc_tensor* const* ys, const int* sizes, const int* pads,
                               void* stream);
YC_API int yc_maxpool_s1_multi_eligible(const yc_tensor* x, const yc_tensor* y, int size, int stride, int pad);
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

/* BackwardScaleChannelsLayerGpu (scale_channels_layer.c:146-160, blas_kernels.cu:1348-1420; CPU twin :99-131),
 * scale_wh = 0: dy *= act'(y) in place (y may be NULL for LINEAR); ds[n,c] += sum over h,w of dy * x; dx += s * dy,
 * where x = the `from` layer's output, s = this layer's input ([n,1,1,c]), ds / dx their deltas. */
YC_API int yc_scale_channels_bwd(const yc_tensor* dy, const yc_tensor* y, int activation, const yc_tensor* x,
                          const yc_tensor* s, const yc_tensor* ds, const yc_tensor* dx, void* stream);
/* BackwardReorgLayerGpu (reorg_layer.c:129-143): dx = the inverse rearrangement of dy (overwrites dx like reorg_cpu);
 * reverse = l->reverse of the forward layer (0 for a plain [reorg]). */
YC_API int yc_reorg_bwd(const yc_tensor* dy, const yc_tensor* dx, int stride, int reverse, void* stream);

/* ---- backward of the memory-bound layers + update ---------------------------------------------- */
/* ForwardMaxpoolLayerGpu in training: also records the winning tap (one byte per output element, [n,h,w,ceil8(c)]),
 * the equivalent of l->indexes (maxpool_layer.cpp:258-288). */
YC_API int yc_maxpool_fwd_train(const yc_tensor* x, const yc_tensor* y, unsigned char* idx, int size, int stride,
                                int pad, void* stream);
/* BackwardMaxpoolLayerGpu (maxpool_layer_kernels.cu:103-143; maxpool_layer.cpp:311-323): dx += scatter(dy) */
YC_API int yc_maxpool_bwd(const yc_tensor* dy, const unsigned char* idx, const yc_tensor* dx, int size, int stride,
                          int pad, void* stream);
/* BackwardUpsampleLayerGpu (upsample_layer.c:121-133): dx += scale * sum of children */
YC_API int yc_upsample_bwd(const yc_tensor* dy, const yc_tensor* dx, int stride, float scale, void* stream);
/* dst += src on views: BackwardRouteLayerGpu's axpy per input (route_layer.c:144-160) */
YC_API int yc_axpy_view(const yc_tensor* src, const yc_tensor* dst, void* stream);
/* BackwardShortcutLayerGpu (shortcut_layer.c:206-220): dy *= act'(y) in place; dprev += dy; dfrom += dy */
YC_API int yc_shortcut_bwd(const yc_tensor* dy, const yc_tensor* y, int activation, const yc_tensor* dprev,
                           const yc_tensor* dfrom, void* stream);
/* UpdateConvolutionalLayerGpu (convolutional_kernels.cu:865-921; convolutional_layer.cpp:1382-1400), SGD:
 * dW -= decay*batch*W; W += lr/batch*dW; dW *= momentum (same for biases/scales without decay).  w_oihw is the fp32
 * master copy in the reference layout; the bf16 operand packings (forward and data-gradient) are refreshed in the
 * same pass.  biases/scales pairs may be NULL. */
YC_API int yc_sgd_update_conv(float* w_oihw, float* dw_packed, int out_c, int in_c, int ksize, float lr, int batch,
                              float momentum, float decay, void* w_packed_fwd, void* w_packed_dgrad, float* biases,
                              float* bias_updates, float* scales, float* scale_updates, void* stream);

/* ---- YOLO head: replaces ForwardYoloLayerGpu ------------------------------------------------- */
/* (yolo_layer.cpp:836-859: copy + per-(b,anchor) logistic / scal_add launches.)
 * x: head conv output NHWC [n, h, w, anchors*(5+classes)] (fp32 or bf16);
 * out_nchw: fp32 [n][anchors*(5+classes)][h][w] == the reference's l->output layout (EntryIndex,
 * yolo_layer.cpp:380-386). gaussian=1: Gaussian_yolo with 9+classes entries
 * (gaussian_yolo_layer.cpp:934-961). */
YC_API int yc_yolo_decode(const yc_tensor* x, float* out_nchw, int anchors, int classes, float scale_x_y,
                   int gaussian, void* stream);

/* YOLO loss + deltas on the device: the training half of ForwardYoloLayer (yolo_layer.cpp:409-772, entirely on one
 * host thread in the reference) and BackwardYoloLayerGpu (:884-888).  out_nchw = decoded head (yc_yolo_decode),
 * truth = [n][max_boxes][5] (x, y, w, h, class id; the list ends at the first x == 0), anchor_wh = l->biases
 * (2*total floats), delta_nchw = l->delta (fp32, same shape as the head, overwritten).  prev_delta (optional) is the
 * head conv's NHWC bf16 delta, which receives += delta.  stats (device, yc_yolo_loss_stats_floats(n, h, w) floats; the
 * first 8 are the results): [0] *(l->cost), [1] count, [2] sum(1-iou), [3] sum(1-giou), [4] |delta|^2, [5] |delta without
 * box entries|^2, [6] sum iou, [7] class_count; the rest holds per-image and per-block partial sums, which the last
 * kernel adds in index order (no float atomics: the loss and its deltas are reproducible bit for bit). */
typedef struct yc_yolo_loss_desc {
  int anchors;  /* l->n: anchors of this head (mask length) */
  int total;    /* l->total */
  int classes;
  int max_boxes;
  int net_w, net_h;
  int iou_loss, iou_thresh_kind; /* IOU_LOSS (box.h:6-13): 0 iou, 1 giou, 2 mse, 3 diou, 4 ciou */
  int focal_loss;
  float ignore_thresh, truth_thresh, iou_thresh, iou_normalizer, cls_normalizer, label_smooth_eps, max_delta;
  int gaussian;        /* 1: [Gaussian_yolo] — entries mu_x,s_x,mu_y,s_y,mu_w,s_w,mu_h,s_h,obj,classes; Gaussian NLL box
                          gradients + uncertainty gradients (ForwardGaussianYoloLayer, gaussian_yolo_layer.cpp:486-857,
                          yolo_point = center); cost = sum delta^2 */
  float uc_normalizer; /* l->uc_normalizer (parser.cpp:471) */
} yc_yolo_loss_desc;
YC_API size_t yc_yolo_loss_stats_floats(int n, int h, int w);
YC_API int yc_yolo_loss(const yc_yolo_loss_desc* desc, const float* out_nchw, const float* truth, const float* anchor_wh,
                        const int* mask, int n, int h, int w, float* delta_nchw, const yc_tensor* prev_delta,
                        float* stats, void* stream);

/* On-device detection extraction (GetYoloDetections for every batch index, yolo_layer.cpp:794-832):
 * scans the decoded NCHW head in the reference's order (anchor n, then cell i), keeps cells with
 * objectness > thresh and writes rows [batch, x, y, w, h, obj, prob[classes]] compactly, ordered
 * exactly like the reference within each image. *count_dev receives the number of rows per image
 * ([n] ints). rows_per_image bounds the per-image output. */
YC_API int yc_yolo_extract(const float* decoded_nchw, int n, int h, int w, int anchors, int classes,
                    const float* anchor_wh, const int* mask, int net_w, int net_h, float thresh,
                    int gaussian, float* rows_dev, int rows_per_image, int* count_dev, void* stream);

/* ---- whole network: replaces LoadNetwork / NetworkPredict / GetNetworkBoxes / NmsSort ------------ */
/* Host side (csrc/network.cu, C++): cfg reader with the reference's INI semantics (parser.cpp:59-100,
 * 1076-1519), darknet .weights reader (parser.cpp:1695-1759,1778-1844), FuseConvBatchNorm
 * (network.cpp:647-682), buffer planner (NHWC bf16, [route] elided by writing producers into concat
 * slots, [shortcut] fused into the producing conv's epilogue), CUDA-graph capture of the forward. */
typedef struct yc_network yc_network;

enum {
  YC_NET_FUSE_BN = 1,        /* fold BN into W,b on the host in fp32, eps 1e-5: what LoadNetwork(train=false)
                                does (parser.cpp:1866-1867).  Off: BN applied in the conv epilogue from the
                                rolling statistics with eps 1e-6 (batchnorm_layer.cpp:226-237, blas.c:214). */
  YC_NET_FUSE_SHORTCUT = 2,  /* conv + following [shortcut] in one kernel (the conv's own output is not kept) */
  YC_NET_ELIDE_ROUTE = 4,    /* producers write straight into concat buffers; off: route = copy kernels */
  YC_NET_CUDA_GRAPH = 8,     /* replay the forward as one captured graph */
  YC_NET_PULL_HEADS = 16,    /* D2H every YOLO layer output after forward like the reference
                                (yolo_layer.cpp:856); off: heads stay on the device (use yc_network_detect) */
  YC_NET_REF_CPU_MAXPOOL = 32, /* reproduce forward_maxpool_layer_avx's artefact (see yc_maxpool_fwd) */
  YC_NET_TRAIN = 64,         /* LoadNetwork(train=true): batchnorm uses batch statistics, every layer keeps delta
                                buffers, conv layers keep fp32 master weights + gradient accumulators */
  YC_NET_DGRAD_ACCUMULATE = 128, /* data gradient ADDS to the previous layer's delta (cuDNN beta=1,
                                convolutional_kernels.cu:735-738) instead of overwriting it like col2im (col2im.c:70;
                                SURVEY B6).  Implied by YC_NET_TRAIN unless YC_NET_TRAIN_REF_CPU. */
  YC_NET_BN_BIAS_GRAD = 256, /* also accumulate the batchnorm shift's gradient into bias_updates (the reference GPU path;
                                its CPU path leaves it at zero).  Implied likewise. */
  YC_NET_STEM_IM2COL = 512,  /* inference: a first conv with c*size*size <= 32 reads an im2col operand built by the input
                                conversion (yc_stem_im2col) and runs as a 1x1 layer */
  YC_NET_TRAIN_REF_CPU = 8192, /* with YC_NET_TRAIN: the reference CPU path's semantics instead of its GPU path's — the data
                                gradient OVERWRITES the previous layer's delta (col2im.c:70) and the batchnorm shift gets no
                                gradient.  The mode the CPU oracle is compared in.  Without it YC_NET_TRAIN
                                implies YC_NET_DGRAD_ACCUMULATE | YC_NET_BN_BIAS_GRAD. */
  YC_NET_NO_STEM_PAIR = 1024, /* stem conv on single pixels instead of pixel pairs (A/B + parity coverage) */
  YC_NET_NO_PAIR_S2 = 2048,  /* 3x3 stride-2 32-channel layer on the plain im2col path instead of the pixel-pair view */
  YC_NET_NO_PAIR_FUSE = 4096, /* do not merge two 1x1 convs that read the same tensor into one GEMM */
  YC_NET_STEM_DIRECT = 16384, /* the 3->32 3x3 stem by yc_stem_conv_direct (mma.sync) instead of yc_stem_conv_tc (tcgen05, the
                                default for such a layer).  Measured slower (1.13 ms against 0.50 ms per yolov4-608 x 64 step:
                                the legacy mma.sync rate of sm_100 bounds it), kept as a tested alternative. */
  YC_NET_PDL = 32768,        /* opt-in: programmatic dependent launches for the inference conv kernels (YC_PLAN_PDL).  Measured
                                twice (round 1, round 2 with the CTA-pair build: 11.21 / 11.40 ms against 11.38 / 11.39 ms per
                                yolov4-608 x 64 step, i.e. noise): the next kernel's CTA cannot become resident next to a
                                226 KB one, and what it could overlap — its own set-up — is not what the short layers pay for */
  YC_NET_PLAN_SHIFT = 16,    /* bits [16,31): YC_PLAN_* switches handed to every conv plan */
  YC_NET_DEFAULT = 1 | 2 | 4 | 8 | 16 | 512,
  YC_NET_STEM_OPERAND = (int)0x80000000u /* the stem through the round-1 path — im2col operand written by the input conversion
                                (yc_stem_im2col) + a 1x1 layer over it — instead of the one-kernel stem (A/B, parity coverage) */
};

/* [setup] LoadNetwork (parser.cpp:1852-1875).  batch <= 0 keeps the cfg's batch (the reference forces
 * batch=1 for inference, parser.cpp:1114-1115; batch>1 is this library's extension).  weights_path may be
 * NULL (weights then come from yc_network_set_conv_weights). */
YC_API int yc_network_load(yc_network** net, const char* cfg_path, const char* weights_path, int batch, int flags);
YC_API int yc_network_load_text(yc_network** net, const char* cfg_text, const char* weights_path, int batch, int flags);
YC_API int yc_network_free(yc_network* net);                         /* FreeNetwork network.cpp:518-645 */
/* [setup] overwrite one conv layer's parameters from host fp32 arrays in the reference's layouts
 * (OIHW weights; scales/mean/var may be NULL when the layer has no BN).  Call yc_network_finalize after. */
YC_API int yc_network_set_conv_weights(yc_network* net, int layer, const float* weights, const float* biases,
                                       const float* scales, const float* rolling_mean, const float* rolling_var);
/* [setup] fuse BN, pack weights, plan buffers and conv kernels.  Called by yc_network_load when weights_path
 * is given; idempotent. */
YC_API int yc_network_finalize(yc_network* net);
/* [setup] LoadWeights (parser.cpp:1846-1850, LoadWeightsUpTo :1778-1844) on a parsed network.  A file that ends at a layer
 * boundary fills the first layers only (a backbone such as yolov4.conv.137; yc_network_info(net, 11) = layers filled) and
 * leaves the others as they are; a file that ends inside a layer is YC_ERR_IO.  The header's `seen` becomes the
 * network's image / iteration count.  Call yc_network_finalize after. */
YC_API int yc_network_load_weights(yc_network* net, const char* path);
/* SaveWeights (parser.cpp:1545-1560, 1590-1643): darknet .weights format — int32 major, minor, revision, uint64 seen, then
 * per conv layer biases, [scales, rolling_mean, rolling_variance], weights (OIHW fp32).  A training network's current
 * device values are pulled first.  cutoff > 0 saves the first `cutoff` layers only (SaveWeightsUpto). */
YC_API int yc_network_save_weights(yc_network* net, const char* path, int cutoff);
/* [setup] FillConvLayer's initialisation law (convolutional_layer.cpp:478-493) through libc rand() after srand(seed): the
 * reference's own values for every conv parsed before the first [yolo] section (which re-seeds from the clock).  Layers
 * below first_layer keep their values (a loaded backbone).  Call yc_network_finalize after. */
YC_API int yc_network_init_weights(yc_network* net, unsigned seed, int first_layer);
/* GetCurrLr (network.cpp:32-84): learning rate at iteration `iter` under the [net] policy (constant, step, steps — whose
 * positions are fractions of max_iter in this reference —, exp, poly, sigmoid, sgdr; random returns the base rate) with
 * burn-in lr * (iter / burn_in)^power.  max_iter <= 0: the value of yc_network_set_schedule / [net] max_batches. */
YC_API float yc_network_current_rate(const yc_network* net, long long iter, int max_iter);
/* net->max_iter / net->curr_iter (owned by the reference's training driver): max_iter <= 0 and iter < 0 keep the
 * current values.  The iteration count advances by one per yc_network_update. */
YC_API int yc_network_set_schedule(yc_network* net, int max_iter, long long iter);

/* ResizeNetwork (network.cpp:255-383): a new input size for an existing network.  Every layer's shape is re-derived, the
 * parameters, rolling statistics, momentum-carrying gradient buffers, seen / iteration counters and a data-parallel
 * context are kept; plans, activation buffers and the captured graph are rebuilt (the network is finalized again if it
 * was).  Pointers obtained before the call (input buffers, yolo outputs) are invalid afterwards.  A size at which some
 * layer would have an empty output fails and leaves the old shape in place (call yc_network_finalize to go on). */
YC_API int yc_network_resize(yc_network* net, int w, int h);
/* what: 0 n_layers, 1 batch, 2 w, 3 h, 4 c, 5 n_conv, 6 kernels launched per forward, 7 n_yolo, 8 subdivisions,
 * 9 current iteration, 10 images seen, 11 layers filled from the weights file, 12 effective flags */
YC_API int yc_network_info(const yc_network* net, int what);
/* per-layer: what: 0 type(0 conv,1 maxpool,2 upsample,3 route,4 shortcut,5 yolo,6 gaussian_yolo) 1 out_c 2 out_h
 * 3 out_w 4 has_output(0 if fused away) 5 in_c 6 size 7 stride 8 activation 9 batch_normalize 10 block_n 11 stages
 * 12 tiles 13 kernel variant bits (yc_conv_plan_info what=7) 14 layer whose kernel computes this conv (fused 1x1 pair;
 * -1 = its own) 15 layer this conv's kernel also computes (-1 = none)
 * 16 classes 17 num (total anchors) 18 nms_kind 19 pad 20 dilation 21 groups 22 group_id 23 shortcut `from`
 * 24 filters / anchors of this head 25 number of inputs 26 max boxes */
YC_API int yc_network_layer_info(const yc_network* net, int layer, int what);
/* k-th input layer index of a layer (route: its `layers=` list resolved; others: the previous layer), -1 past the end */
YC_API int yc_network_layer_input(const yc_network* net, int layer, int k);
/* what: 0 scale_x_y (upsample: scale) 1 beta_nms 2 ignore_thresh 3 truth_thresh 4 iou_thresh 5 iou_normalizer
 * 6 cls_normalizer 7 label_smooth_eps 8 max_delta 9 uc_normalizer */
YC_API float yc_network_layer_float(const yc_network* net, int layer, int what);
/* what: 0 anchors (l->biases, 2*num floats) 1 mask (as floats).  Returns the number of floats written (<0: error). */
YC_API int yc_network_layer_array(const yc_network* net, int layer, int what, float* dst, int capacity);
YC_API double yc_network_flops(const yc_network* net);               /* sum 2*nweights*out_h*out_w per image
                                                                        (convolutional_layer.cpp:714) */
/* NetworkPredict (network.cpp:412-430 -> NetworkPredictGpu network_kernels.cu:502-522): input = host fp32 NCHW
 * [batch*c*h*w] (borrowed), returns the host buffer of the last layer's output (layer-owned, valid until the
 * next call) or NULL on error.  Synchronous like the reference. */
YC_API float* yc_network_predict(yc_network* net, const float* input_host);
/* Mat2Image + NetworkPredict for u8 HWC images at the network's resolution (ProcImage, yolo_core.cpp:104-112, minus
 * cv::resize): the H2D copy carries bytes (4x less than fp32), the /255 and layout change run on the device; results
 * are bit-identical to yc_network_predict on the Mat2Image output.  yc_network_input_buffer_u8: pinned staging. */
YC_API float* yc_network_predict_u8(yc_network* net, const unsigned char* input_host_hwc, int swap_rb);
YC_API unsigned char* yc_network_input_buffer_u8(yc_network* net);
/* ProcImage from the frame on (yolo_core.cpp:104-112): cv::resize(frame, Size(net->w, net->h)) + cvtColor + Mat2Image +
 * NetworkPredict for a batch of u8 HWC frames of ANY size src_h x src_w: the frames cross PCIe as bytes, resize (bit-identical
 * to the library's INTER_LINEAR for 8-bit data, yc_resize_u8), channel swap, /255 and layout run in one device kernel.
 * Equals yc_network_predict_u8 on host-resized frames bit for bit. */
YC_API float* yc_network_predict_u8_resized(yc_network* net, const unsigned char* frames_host_hwc, int src_h, int src_w,
                                            int swap_rb);
/* Input double buffering for a stream of batches (the reference's NetworkPredictGpu copies and computes serially,
 * network_kernels.cu:502-522): yc_network_prefetch(next) starts the H2D copy of the NEXT batch on a separate stream
 * while the current forward runs; the later yc_network_predict(net, next) (same pointer) skips its own H2D.
 * yc_network_input_buffer_slot(net, 0|1): two pinned staging buffers to alternate between (slot 0 ==
 * yc_network_input_buffer). */
YC_API int yc_network_prefetch(yc_network* net, const float* input_host);
YC_API float* yc_network_input_buffer_slot(yc_network* net, int slot);
/* net->input_pinned_cpu (network_kernels.cu:505-513): the library's pinned staging buffer of batch*c*h*w floats.
 * A producer may fill it in place and pass it to yc_network_predict, which then skips its staging memcpy. */
YC_API float* yc_network_input_buffer(yc_network* net);
/* ForwardNetworkGpu (network_kernels.cu:45-114) on an input already resident on the device
 * (fp32 NCHW, the reference's net->input_state_gpu layout).  Asynchronous on `stream`. */
YC_API int yc_network_forward(yc_network* net, const float* input_dev_nchw, void* stream);
/* cuda_pull_array(l->output_gpu) equivalent: layer output as fp32 NCHW [batch,out_c,out_h,out_w] on the host. */
YC_API int yc_network_layer_output(yc_network* net, int layer, float* dst_host);
/* host pointer of a YOLO layer's pulled output (l->output; needs YC_NET_PULL_HEADS) */
YC_API const float* yc_network_yolo_output(yc_network* net, int layer);
/* per-layer device time of one un-graphed forward, averaged over iters (net->benchmark_layers,
 * network_kernels.cu:70-94).  ms_per_layer has n_layers entries (+1: the input conversion at index n). */
YC_API int yc_network_profile(yc_network* net, const float* input_dev_nchw, int iters, float* ms_per_layer);
/* Measurement: the forward's kernels of one class (which = 0: convolution launches, 1: all other kernels, 2: all)
 * replayed back to back as one CUDA graph, the way they run inside the step; mean ms per replay over `iters` replays
 * (CUDA events on the network's stream), number of kernels per replay in *launches.  No reference counterpart. */
YC_API int yc_network_time_ops(yc_network* net, const float* input_dev_nchw, int which, int iters, float* ms_per_replay,
                               int* launches);

/* ---- training step: replaces TrainNetworkDatumGpu / ForwardBackwardNetworkGpu / UpdateNetworkGpu -------------- */
/* (network_kernels.cu:116-293; CPU twins network.cpp:101-208.)  The network must be loaded with YC_NET_TRAIN.
 * truth = [batch][max_boxes*5] floats (x, y, w, h, class id), the reference's net->truth layout. */
/* TrainNetworkDatum: host input + truth -> forward + backward (no update); *cost = GetNetworkCost. Synchronous. */
YC_API int yc_network_train_datum(yc_network* net, const float* x_host, const float* truth_host, float* cost);
/* ForwardBackwardNetworkGpu on device-resident fp32 NCHW input and truth; asynchronous on `stream`.  In every training
 * entry point a NULL stream means the NETWORK'S OWN stream (the one yc_network_train_datum uses; a non-blocking stream:
 * it does not synchronise with the legacy default stream) — pass one and the same stream to yc_network_forward_backward
 * and yc_network_update, and record timing events on that stream. */
YC_API int yc_network_forward_backward(yc_network* net, const float* input_dev_nchw, const float* truth_dev, void* stream);
/* The tail of TrainNetwork (network.cpp:227-231): curr_iter++ and then UpdateNetworkGpu (network_kernels.cu:214-243) ->
 * UpdateConvolutionalLayerGpu on every conv layer (convolutional_kernels.cu:865-922): NaN / Inf scrub of gradient and
 * weights, then SGD with momentum and decay, or adam_update_gpu under [net] adam=1 (B1, B2, eps; t = iter), in one launch.
 * lr < 0 = GetCurrLr at the incremented iteration (yc_network_current_rate; with the default constant policy and no burn-in
 * that is [net] learning_rate), negative momentum / decay = the cfg's [net] values; batch = net batch * subdivisions
 * (x data-parallel world size).  Layers are skipped as the reference does: burnin_update * burn_in > iter, train_only_bn,
 * dont_update >= iter.  stream NULL = the network's own stream. */
YC_API int yc_network_update(yc_network* net, float lr, float momentum, float decay, void* stream);
/* net->benchmark_layers for the training step: device time by operator category (CUDA events after every operator),
 * averaged over iters.  ms[7]: conv fprop, wgrad, dgrad, batchnorm/activation passes, other layers + loss, SGD update,
 * total.  The update inside runs with lr = 0 (weights unchanged; the momentum scaling of the accumulators applies). */
YC_API int yc_network_train_profile(yc_network* net, const float* input_dev_nchw, const float* truth_dev, int iters,
                                    float* ms);
/* what: 0 truth floats per image, 1 subdivisions, 2 delta slab MiB */
YC_API int yc_network_train_info(const yc_network* net, int what);
/* cuda_pull_array(l->delta_gpu): fp32 NCHW [batch,out_c,out_h,out_w] */
YC_API int yc_network_layer_delta(yc_network* net, int layer, float* dst_host);
/* cuda_pull_array(l->x_gpu): pre-batchnorm convolution output of a batch-normalised conv layer, fp32 NCHW */
YC_API int yc_network_layer_x(yc_network* net, int layer, float* dst_host);
/* PullConvolutionalLayer by name ("weights", "biases", "scales", "rolling_mean", "rolling_variance", "mean", "variance",
 * "weight_updates", "bias_updates", "scale_updates"; yolo: "cost", "stats"), reference layouts. Returns floats written. */
YC_API int yc_network_get_param(yc_network* net, int layer, const char* key, float* dst_host, int capacity);

/* ---- data-parallel training: replaces TrainNetworks / SyncNetworks (network_kernels.cu:295-484) ------------------- */
/* One process per GPU.  Rank 0 calls yc_dp_unique_id (ncclGetUniqueId, 128 bytes) and the host plumbing hands the id
 * to every rank; every rank then calls yc_network_dp_init (collective).  From then on yc_network_forward_backward
 * all-reduces (sum, fp32) each conv layer's gradient bucket [weight_updates | bias_updates | scale_updates] over
 * NCCL / NVLink on a separate stream as soon as that layer's weight gradient is enqueued, overlapping the rest of
 * the backward pass, and yc_network_update divides by the global batch.  With subdivisions > 1 (gradients accumulate
 * over several backward passes) or after yc_network_dp_set_overlap(net, 0) the buckets stay local and are reduced in
 * one NCCL group by yc_network_dp_reduce, which yc_network_update calls itself when a reduction is pending. */
YC_API int yc_dp_unique_id(void* out128);
YC_API int yc_dp_nccl_version(void);
YC_API int yc_network_dp_init(yc_network* net, const void* unique_id128, int rank, int world);
YC_API int yc_network_dp_set_overlap(yc_network* net, int on);
YC_API int yc_network_dp_reduce(yc_network* net, void* stream);
/* Every gradient bucket must be all-reduced exactly once between two updates (the momentum term is carried divided by
 * the world size): a backward pass after this step's gradients were reduced — a second yc_network_forward_backward before
 * yc_network_update under per-layer overlap, or one after yc_network_dp_reduce — returns YC_ERR_ARG.  A caller that runs
 * backward passes WITHOUT updates on purpose (timing loops) declares the step's gradients discarded with this call. */
YC_API int yc_network_dp_discard_step(yc_network* net);
/* what: 0 world, 1 rank, 2 buckets per step, 3 MiB all-reduced per step, 4 per-layer overlap on */
YC_API int yc_network_dp_info(const yc_network* net, int what);

/* Detection == the reference's struct Detection minus unused members (yolo_core.h:598-609). */
typedef struct yc_detection {
  float x, y, w, h;        /* Box bbox */
  int classes;
  float* prob;             /* [classes] */
  float objectness;
  int sort_class;
  float uc[4];             /* Gaussian_yolo uncertainties */
} yc_detection;
/* GetNetworkBoxes (network.cpp:498-503) for batch element b (the reference is b=0 only): scans the pulled
 * host YOLO outputs exactly like GetYoloDetections (yolo_layer.cpp:794-832).  Caller frees with
 * yc_free_detections.  Needs YC_NET_PULL_HEADS. */
YC_API yc_detection* yc_network_get_boxes(yc_network* net, int b, float thresh, int* num);
YC_API void yc_free_detections(yc_detection* dets, int n);           /* FreeDetections network.cpp:505-516 */
/* NmsSort (box.cpp:393-419). nms_kind: 0 greedy, 1 DIoU. */
YC_API void yc_nms_sort(yc_detection* dets, int total, int classes, float thresh, int nms_kind, float beta);
/* Throughput path: threshold + box decode on the device for every image (yc_yolo_extract per
 * head), compact rows to the host, then NmsSort per image on host threads when nms_thresh > 0.
 * On return *rows points to a library-owned host array of [total][7 + classes] floats
 * (image, x, y, w, h, objectness, head_layer, prob[classes]) ordered by image, then head, anchor, cell — the
 * reference's order within an image — and counts[b] is the number of rows of image b. */
YC_API int yc_network_detect(yc_network* net, float thresh, float nms_thresh, const float** rows, int* counts,
                             int* total);

#ifdef __cplusplus
}
#endif
#endif
