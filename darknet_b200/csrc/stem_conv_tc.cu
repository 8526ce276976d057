// The first convolution (3 -> 8..32 channels, 3x3, stride 1, pad 1) as a tcgen05 kernel whose A operand is built in shared
// memory by the kernel itself: NO im2col operand in global memory (north_star), one launch instead of the input conversion
// + a 1x1 convolution over a 64-byte-per-pixel operand.
//
// Replaces, for layer 0: the input upload's layout step (network_kernels.cu:513), im2col_gpu_ext + gemm
// (convolutional_kernels.cu:252-553, im2col_kernels.cu) and add_bias / activate_array.
//
// Tile = 8 rows x 16 columns of output pixels = the 128 rows of one MMA.  Per CTA (persistent, one per SM, 26 warps):
//   warp 8      one lane TMA-loads each tile's input patch — box {24 w, 10 h, 3 c} fp32 of the NCHW image starting one
//               row above and four columns left of the tile; the engine zero-fills whatever lies outside the image (=
//               the padding) — into an 8-deep ring of 2880-byte slots;
//   warps 0-7   "builders", two groups of four taking tiles round robin: thread p of a group owns pixel (p / 16, p % 16):
//               27 shared-memory loads from the patch, rounded to bf16 and written as the pixel's 64-byte K-major row —
//               column (kh*3+kw)*3+ci, 27 of 32 used — in the 64B-swizzled layout the tensor core reads (generic-proxy
//               stores + fence.proxy.async + an mbarrier the 128 builders of the group arrive on);
//   warp 9      one elected lane issues two tcgen05.mma (M128 x N32 x K16 each) per tile against the 32 x 32 filter
//               matrix resident in shared memory, accumulators in TMEM (8 buffers of 32 columns), tcgen05.commit frees
//               the A slot and publishes the accumulator;
//   warps 10-25 epilogue, four sets of four (one warp per TMEM lane quadrant) taking tiles round robin: tcgen05.ld ->
//               + bias -> activation (packed fp32 pairs) -> bf16 -> swizzled staging -> TMA store of a {32 ch, 16 w, 2 h}
//               box (clipped at the image border and at the channel count: partial tiles and concat slices need no
//               special case).
// HBM traffic: 12 B read + 64 B written per pixel (the operand path: 204 B).  The layer is then bound by its activation:
// 757 M mish values x 2 MUFU on yolov4-608 x 64 = 0.37 ms of XU-pipe time at 16 per clock per SM.
// The input map needs a row pitch that is a multiple of 16 bytes: widths that are not a multiple of 4 go to the mma.sync
// kernel (yc_stem_conv_direct, bit-identical results); the reference's network sizes are multiples of 32.
#include <cuda.h>
#include <algorithm>

#include "ptx.cuh"
#include "yc_common.cuh"

namespace yc {

constexpr int kTcBuilders = 128;    // threads that build one tile: one pixel each
constexpr int kTcBuildGroups = 2;   // groups of four builder warps (warps 0-7), tiles round robin
constexpr int kTcLoadWarp = 4 * kTcBuildGroups, kTcMmaWarp = kTcLoadWarp + 1, kTcEpiWarp0 = kTcLoadWarp + 2;
constexpr int kTcEpiSets = 4;       // sets of four epilogue warps, tiles round robin
constexpr int kTcThreads = (kTcEpiWarp0 + 4 * kTcEpiSets) * 32;  // 832
constexpr int kTcStagesX = 8;       // input patches in flight
// 18 x 10 pixels are needed (one around the tile).  The engine wants the innermost start coordinate 16-byte aligned (measured:
// a start of -1 or +3 floats is an illegal-instruction fault, -4 is fine), so the patch starts 4 columns left of the tile.
constexpr int kTcPatchLeft = 4;
constexpr int kTcPatchW = 24, kTcPatchH = 10;
constexpr int kTcPatchBytes = kTcPatchW * kTcPatchH * 3 * 4;     // 2880
constexpr int kTcPatchStride = 2944;                             // 128-byte aligned slots
constexpr int kTcStagesA = 8;       // A slots of 128 rows x 64 B
constexpr int kTcAcc = 8;           // TMEM accumulators of 32 columns
constexpr int kTcTileH = 8, kTcTileW = 16;

struct StemTcParams {
  const __nv_bfloat16* wp;    // bf16 [n_out][32], column (kh*3+kw)*3+ci
  const float* bias;          // [n_out]
  int B, H, W, n_out, act;
  int tiles_h, tiles_w, tiles;
};

__device__ __forceinline__ uint32_t tc_pack(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

// (image, tile row, tile column) of a tile index that advances by a fixed step: one division pair up front, then carries only
// (a lone thread feeds the TMA engine; two integer divisions per tile made it the slowest stage of the pipeline)
struct TcTileWalk {
  int b, th, tw, db, dth, dtw;
  __device__ __forceinline__ TcTileWalk(int t0, int step, int tiles_h, int tiles_w) {
    const int tpi = tiles_h * tiles_w;
    b = t0 / tpi;
    int rem = t0 - b * tpi;
    th = rem / tiles_w;
    tw = rem - th * tiles_w;
    db = step / tpi;
    rem = step - db * tpi;
    dth = rem / tiles_w;
    dtw = rem - dth * tiles_w;
  }
  __device__ __forceinline__ void next(int tiles_h, int tiles_w) {
    tw += dtw;
    if (tw >= tiles_w) { tw -= tiles_w; ++th; }
    th += dth;
    if (th >= tiles_h) { th -= tiles_h; ++b; }
    b += db;
  }
};

__device__ __forceinline__ float tc_lds(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
  return v;
}

// (896 = 28 warps: the register file is handed out per four warps, so 26 warps cost what 28 do; 65536 / 896 = 73 -> 72)
template <int ACT>
__global__ void __launch_bounds__(896, 1)
stem_conv_tc_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_y,
                    const StemTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t a_base = smem_base;                                   // kTcStagesA x 8192
  const uint32_t b_base = a_base + kTcStagesA * 8192u;                 // 32 rows x 64 B (1024-aligned)
  const uint32_t out_base = b_base + 2048u;                            // 16 warps x 2 buffers x 2048
  const uint32_t x_base = out_base + kTcEpiSets * 4 * 2 * 2048u;       // kTcStagesX x 2432
  const uint32_t bias_smem = x_base + kTcStagesX * kTcPatchStride;     // 32 floats
  const uint32_t bar_base = bias_smem + 128u;
  auto afull = [&](int s) { return bar_base + 8u * s; };
  auto aempty = [&](int s) { return bar_base + 8u * (kTcStagesA + s); };
  auto tfull = [&](int a) { return bar_base + 8u * (2 * kTcStagesA + a); };
  auto tempty = [&](int a) { return bar_base + 8u * (2 * kTcStagesA + kTcAcc + a); };
  auto xfull = [&](int s) { return bar_base + 8u * (2 * kTcStagesA + 2 * kTcAcc + s); };
  auto xempty = [&](int s) { return bar_base + 8u * (2 * kTcStagesA + 2 * kTcAcc + kTcStagesX + s); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * kTcStagesA + 2 * kTcAcc + 2 * kTcStagesX);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // the filter matrix [32 rows (n)][32 k] bf16 as a K-major 64B-swizzled operand (rows >= n_out are zero), and the bias
  if (threadIdx.x < 128) {
    const int n = threadIdx.x >> 2, j = threadIdx.x & 3;
    uint4 q = make_uint4(0, 0, 0, 0);
    if (n < p.n_out) q = __ldg(reinterpret_cast<const uint4*>(p.wp + (size_t)n * 32) + j);
    ptx::st_shared_v4(b_base + (uint32_t)n * 64u + (((uint32_t)j ^ (((uint32_t)n >> 1) & 3u)) << 4), q.x, q.y, q.z, q.w);
  } else if (threadIdx.x < 160) {
    const int n = threadIdx.x - 128;
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(bias_smem + 4u * n), "f"(n < p.n_out ? __ldg(p.bias + n) : 0.f) : "memory");
  }
  if (threadIdx.x == 128) {  // (not a lane of the MMA warp, which allocates TMEM below as a converged warp)
    ptx::prefetch_tmap(&tmap_x);
    ptx::prefetch_tmap(&tmap_y);
    for (int s = 0; s < kTcStagesA; ++s) {
      ptx::mbar_init(afull(s), kTcBuilders);  // every builder thread arrives once per tile
      ptx::mbar_init(aempty(s), 1);           // tcgen05.commit
    }
    for (int a = 0; a < kTcAcc; ++a) {
      ptx::mbar_init(tfull(a), 1);
      ptx::mbar_init(tempty(a), 4);           // the four warps of the tile's epilogue set
    }
    for (int s = 0; s < kTcStagesX; ++s) {
      ptx::mbar_init(xfull(s), 1);            // arrive.expect_tx of the loader + the bytes
      ptx::mbar_init(xempty(s), 4);           // lane 0 of each builder warp, after the warp has read the patch
    }
    ptx::fence_barrier_init();
  }
  ptx::fence_proxy_async();  // the filter rows above were written through the generic proxy; the tensor core reads them
  __syncwarp();
  if (warp == kTcMmaWarp) {
    ptx::tmem_alloc(tmem_slot, 256);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp < kTcLoadWarp) {
    // ===================== builders: patch -> the tile's A operand; group g takes tiles g, g + G, ... =====================
    const int grp = warp >> 2;
    const int pid = threadIdx.x & 127;
    const int pr = pid >> 4, pc = pid & 15;  // this thread's pixel inside the tile
    const uint32_t row_off = (uint32_t)pid * 64u;
    const uint32_t swz = ((uint32_t)pid >> 1) & 3u;
    const uint32_t pix_off = (uint32_t)(pr * kTcPatchW + pc + kTcPatchLeft - 1) * 4u;  // tap (0, 0): one up-left of the pixel
    for (int it = grp, t = blockIdx.x + grp * (int)gridDim.x; t < p.tiles; it += kTcBuildGroups, t += kTcBuildGroups * (int)gridDim.x) {
      const int xs = it % kTcStagesX, as = it % kTcStagesA;
      ptx::mbar_wait(xfull(xs), (uint32_t)((it / kTcStagesX) & 1));
      const uint32_t src = x_base + (uint32_t)xs * kTcPatchStride + pix_off;
      float v[27];  // k = (kh*3+kw)*3 + ci
#pragma unroll
      for (int kh = 0; kh < 3; ++kh)
#pragma unroll
        for (int kw = 0; kw < 3; ++kw)
#pragma unroll
          for (int ci = 0; ci < 3; ++ci)
            v[(kh * 3 + kw) * 3 + ci] = tc_lds(src + (uint32_t)(((ci * kTcPatchH + kh) * kTcPatchW + kw) * 4));
      uint32_t w[14];
#pragma unroll
      for (int j = 0; j < 13; ++j) w[j] = tc_pack(v[2 * j], v[2 * j + 1]);
      w[13] = tc_pack(v[26], 0.f);
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(xempty(xs));  // every lane's loads have returned (their values were just consumed)
      ptx::mbar_wait(aempty(as), (uint32_t)((it / kTcStagesA) & 1) ^ 1u);
      const uint32_t dst = a_base + (uint32_t)as * 8192u + row_off;
      ptx::st_shared_v4(dst + ((0u ^ swz) << 4), w[0], w[1], w[2], w[3]);
      ptx::st_shared_v4(dst + ((1u ^ swz) << 4), w[4], w[5], w[6], w[7]);
      ptx::st_shared_v4(dst + ((2u ^ swz) << 4), w[8], w[9], w[10], w[11]);
      ptx::st_shared_v4(dst + ((3u ^ swz) << 4), w[12], w[13], 0u, 0u);
      ptx::fence_proxy_async();  // generic-proxy stores -> visible to the tensor core's (async-proxy) reads
      ptx::mbar_arrive(afull(as));
    }
  } else if (warp == kTcLoadWarp) {
    // ===================== patch loader =====================
    if (lane == 0) {
      int it = 0;
      TcTileWalk at(blockIdx.x, gridDim.x, p.tiles_h, p.tiles_w);
      for (int t = blockIdx.x; t < p.tiles; t += gridDim.x, ++it, at.next(p.tiles_h, p.tiles_w)) {
        const int xs = it % kTcStagesX;
        ptx::mbar_wait(xempty(xs), (uint32_t)((it / kTcStagesX) & 1) ^ 1u);
        ptx::mbar_arrive_expect_tx(xfull(xs), kTcPatchBytes);
        ptx::tma_load_4d(x_base + (uint32_t)xs * kTcPatchStride, &tmap_x, xfull(xs), at.tw * kTcTileW - kTcPatchLeft,
                         at.th * kTcTileH - 1, 0, at.b);
      }
    }
    __syncwarp();
  } else if (warp == kTcMmaWarp) {
    // ===================== MMA issuer (whole warp converged, one elected lane issues) =====================
    const bool leader = ptx::elect_one();
    const uint32_t idesc = ptx::make_idesc_bf16(128, 32);
    const uint64_t bdesc = ptx::make_smem_desc(b_base, 64);
    int it = 0;
    for (int t = blockIdx.x; t < p.tiles; t += gridDim.x, ++it) {
      const int acc = it % kTcAcc;
      const int stage = it % kTcStagesA;
      ptx::mbar_wait(tempty(acc), (uint32_t)((it / kTcAcc) & 1) ^ 1u);
      ptx::mbar_wait(afull(stage), (uint32_t)((it / kTcStagesA) & 1));
      ptx::tc_fence_after();
      const uint64_t adesc = ptx::make_smem_desc(a_base + (uint32_t)stage * 8192u, 64);
      if (leader) {
        ptx::mma_bf16_ss(tmem_base + acc * 32, adesc, bdesc, idesc, 0u);
        ptx::mma_bf16_ss(tmem_base + acc * 32, adesc + 2u, bdesc + 2u, idesc, 1u);  // K 16..31: +32 bytes inside the swizzle atom
        ptx::mma_commit(aempty(stage));
        ptx::mma_commit(tfull(acc));
      }
      __syncwarp();
    }
  } else {
    // ===================== epilogue: sets of four warps, tiles round robin =====================
    const int ew = warp - kTcEpiWarp0;
    const int set = ew >> 2;
    const int quad = warp & 3;  // the TMEM lane quadrant this warp may read
    const uint32_t sbuf0 = out_base + (uint32_t)ew * 2u * 2048u;
    const uint32_t rowoff = (uint32_t)lane * 64u;
    const uint32_t swz = ((uint32_t)lane >> 1) & 3u;
    uint32_t nst = 0;
    TcTileWalk at(blockIdx.x + set * (int)gridDim.x, kTcEpiSets * (int)gridDim.x, p.tiles_h, p.tiles_w);
    for (int it = set, t = blockIdx.x + set * (int)gridDim.x; t < p.tiles;
         it += kTcEpiSets, t += kTcEpiSets * (int)gridDim.x, at.next(p.tiles_h, p.tiles_w)) {
      const int acc = it % kTcAcc;
      ptx::mbar_wait(tfull(acc), (uint32_t)((it / kTcAcc) & 1));
      ptx::tc_fence_after();
      uint32_t r[32];
      ptx::tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * 32), r);
      ptx::tmem_ld_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(tempty(acc));
      const uint32_t sbuf = sbuf0 + (nst & 1u) * 2048u;
      if (lane == 0) ptx::tma_store_wait_read<1>();  // the store that used this buffer two stores ago has read it
      __syncwarp();
#pragma unroll
      for (int j = 0; j < 4; ++j) {  // eight channels = one 16-byte chunk of the pixel's row, as four packed fp32 pairs
        f32x2 x[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          float b0, b1;
          asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(b0), "=f"(b1) : "r"(bias_smem + 4u * (uint32_t)(8 * j + 2 * e)));
          x[e] = add2(pack2(__uint_as_float(r[8 * j + 2 * e]), __uint_as_float(r[8 * j + 2 * e + 1])), pack2(b0, b1));
        }
        uint32_t o[4];
        if (ACT == YC_MISH) {  // act_mish (yc_common.cuh) on pairs: x + x*r, r = 1/(-q/2), q = n(n+2)+2 on h = n/2
          float h[8];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            float a0, a1;
            unpack2(fma2(x[e], pack2(1.4426950408889634f, 1.4426950408889634f), pack2(-1.f, -1.f)), a0, a1);
            h[2 * e] = ex2_ftz(a0);
            h[2 * e + 1] = ex2_ftz(a1);
          }
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const f32x2 hh = pack2(h[2 * e], h[2 * e + 1]);
            float q0, q1;
            unpack2(fma2(hh, fma2(hh, pack2(-2.f, -2.f), pack2(-2.f, -2.f)), pack2(-1.f, -1.f)), q0, q1);
            h[2 * e] = rcp_ftz(q0);
            h[2 * e + 1] = rcp_ftz(q1);
          }
#pragma unroll
          for (int e = 0; e < 4; ++e) x[e] = fma2(x[e], pack2(h[2 * e], h[2 * e + 1]), x[e]);
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          float y0, y1;
          unpack2(x[e], y0, y1);
          if (ACT == YC_LEAKY) { y0 = fmaxf(y0, 0.1f * y0); y1 = fmaxf(y1, 0.1f * y1); }
          o[e] = tc_pack(y0, y1);
        }
        ptx::st_shared_v4(sbuf + rowoff + (((uint32_t)j ^ swz) << 4), o[0], o[1], o[2], o[3]);
      }
      ptx::fence_proxy_async();
      __syncwarp();
      if (lane == 0) {
        // this warp's 32 rows = tile rows 2*quad, 2*quad+1 (16 pixels each): one {32 ch, 16 w, 2 h, 1} box
        ptx::tma_store_4d(&tmap_y, sbuf, 0, at.tw * kTcTileW, at.th * kTcTileH + 2 * quad, at.b);
        ptx::tma_store_commit();
      }
      ++nst;
    }
    if (lane == 0) ptx::tma_store_wait<0>();
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == kTcMmaWarp) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 256);
  }
}

}  // namespace yc

using namespace yc;

namespace {

typedef CUresult (*StemEncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                      const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                      CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
StemEncodeTiledFn s_encode_tiled = nullptr;

constexpr int kTcSmemBytes = 1024 /*alignment slack*/ + kTcStagesA * 8192 + 2048 + kTcEpiSets * 4 * 2 * 2048 +
                             kTcStagesX * kTcPatchStride + 128 + 8 * (2 * kTcStagesA + 2 * kTcAcc + 2 * kTcStagesX) + 16;

template <int ACT>
int stem_tc_launch(const CUtensorMap& tmx, const CUtensorMap& tm, const StemTcParams& p, int grid, cudaStream_t st) {
  static bool attr_done[8] = {};
  int dev = 0;
  YC_CUDA(cudaGetDevice(&dev));
  if (dev < 8 && !attr_done[dev]) {
    YC_CUDA(cudaFuncSetAttribute(stem_conv_tc_kernel<ACT>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemBytes));
    attr_done[dev] = true;
  } else if (dev >= 8) {
    YC_CUDA(cudaFuncSetAttribute(stem_conv_tc_kernel<ACT>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemBytes));
  }
  stem_conv_tc_kernel<ACT><<<grid, kTcThreads, kTcSmemBytes, st>>>(tmx, tm, p);
  YC_LAUNCH_CHECK();
  return YC_OK;
}

}  // namespace

extern "C" int yc_stem_conv_tc(const float* x_nchw, int batch, int h, int w, const void* w_packed, const float* bias,
                               int activation, const yc_tensor* y, void* stream) {
  YC_REQUIRE(x_nchw && w_packed && bias && y && y->ptr && batch > 0 && h > 0 && w > 0, "yc_stem_conv_tc: bad arguments");
  YC_REQUIRE(y->dtype == YC_BF16 && y->n == batch && y->h == h && y->w == w && y->c % 8 == 0 && y->c >= 8 && y->c <= 32 &&
                 y->ldc >= y->c && y->ldc % 8 == 0 && ((uintptr_t)y->ptr & 15) == 0,
             "yc_stem_conv_tc: the output must be an NHWC bf16 view of the input's size with 8, 16, 24 or 32 channels");
  YC_REQUIRE(activation == YC_MISH || activation == YC_LEAKY || activation == YC_LINEAR,
             "yc_stem_conv_tc: activation %d is not supported (mish / leaky / linear)", activation);
  YC_REQUIRE(((uintptr_t)w_packed & 15) == 0, "yc_stem_conv_tc: the packed filters must be 16-byte aligned");
  if (!s_encode_tiled) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qr;
    YC_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr));
    YC_REQUIRE(fn && qr == cudaDriverEntryPointSuccess, "cuTensorMapEncodeTiled not available");
    s_encode_tiled = (StemEncodeTiledFn)fn;
  }
  // the input map's row pitch must be a multiple of 16 bytes
  if (w % 4 != 0 || ((uintptr_t)x_nchw & 15) != 0) return yc_stem_conv_direct(x_nchw, batch, h, w, w_packed, bias, activation, y, stream);
  StemTcParams p;
  p.wp = (const __nv_bfloat16*)w_packed;
  p.bias = bias;
  p.B = batch; p.H = h; p.W = w; p.n_out = y->c; p.act = activation;
  p.tiles_h = ceil_div(h, kTcTileH);
  p.tiles_w = ceil_div(w, kTcTileW);
  const long long tiles = (long long)batch * p.tiles_h * p.tiles_w;
  YC_REQUIRE(tiles < (1ll << 31), "yc_stem_conv_tc: too many tiles");
  p.tiles = (int)tiles;
  // output map over {c, w, h, n} of the NHWC view; box {32 ch, 16 w, 2 h, 1} = the 32 rows one epilogue warp owns.
  // Channels >= y->c, columns >= W and rows >= H are clipped by the TMA engine.
  CUtensorMap tm;
  const cuuint64_t dims[4] = {(cuuint64_t)y->c, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)batch};
  const cuuint64_t strides[3] = {(cuuint64_t)y->ldc * 2, (cuuint64_t)y->ldc * 2 * w, (cuuint64_t)y->ldc * 2 * w * h};
  const cuuint32_t box[4] = {32, (cuuint32_t)kTcTileW, 2, 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  const CUresult cr = s_encode_tiled(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, y->ptr, dims, strides, box, estr,
                                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B,
                                     CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  YC_REQUIRE(cr == CUDA_SUCCESS, "yc_stem_conv_tc: output tensor map encode failed (CUresult %d, c=%d ldc=%d)", (int)cr, y->c,
             y->ldc);
  // input map over {w, h, c, n} of the fp32 NCHW image; box {24, 10, 3, 1} starting one row above / four columns left of a tile, elements
  // outside the image are zero-filled (= the padding)
  CUtensorMap tmx;
  const cuuint64_t xdims[4] = {(cuuint64_t)w, (cuuint64_t)h, 3, (cuuint64_t)batch};
  const cuuint64_t xstrides[3] = {(cuuint64_t)w * 4, (cuuint64_t)w * h * 4, (cuuint64_t)w * h * 12};
  const cuuint32_t xbox[4] = {(cuuint32_t)kTcPatchW, (cuuint32_t)kTcPatchH, 3, 1};
  const CUresult cx = s_encode_tiled(&tmx, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(x_nchw), xdims, xstrides, xbox,
                                     estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                     CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  YC_REQUIRE(cx == CUDA_SUCCESS, "yc_stem_conv_tc: input tensor map encode failed (CUresult %d, w=%d h=%d)", (int)cx, w, h);
  const int grid = (int)std::min<long long>(tiles, (long long)sm_count());
  cudaStream_t st = (cudaStream_t)stream;
  if (activation == YC_MISH) return stem_tc_launch<YC_MISH>(tmx, tm, p, grid, st);
  if (activation == YC_LEAKY) return stem_tc_launch<YC_LEAKY>(tmx, tm, p, grid, st);
  return stem_tc_launch<YC_LINEAR>(tmx, tm, p, grid, st);
}
