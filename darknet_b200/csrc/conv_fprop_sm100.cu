// Convolution forward as ONE implicit-GEMM kernel for sm_100a.
//
// Replaces ForwardConvolutionalLayerGpu (convolutional_kernels.cu:252-553): cuDNN / per-image
// im2col_gpu_ext + cublasSgemm, then add_bias_gpu / scale_bias_gpu, activate_array_ongpu, and the
// shortcut add that follows in residual blocks (shortcut_layer.c:191-204) — all in one launch, with no
// materialised im2col workspace.
//
// GEMM view:  D[M = B*Ho*Wo pixels][N = out_c] = A[M][K = taps*Cin_p] * W[N][K]^T
//   A rows are gathered by the TMA engine in im2col mode straight from the NHWC activation tensor
//   (one (tap, 64-channel chunk) slab per pipeline stage; zero padding = TMA out-of-bounds fill), or by a
//   plain 2-D tiled tensor map for 1x1/stride-1 layers.  W rows come from the pre-packed K-major bf16
//   weights.  Both land in shared memory in the canonical K-major 128B/64B/32B-swizzled layout that
//   tcgen05.mma reads through shared-memory descriptors.  Accumulators live in TMEM (two buffers, so the
//   epilogue of tile i overlaps the main loop of tile i+1).
//
// CTA = 320 threads, persistent over tiles (grid = min(tiles, #SM)):
//   warp 0 / lane 0 : TMA producer        (full/empty mbarrier ring, `stages` deep)
//   warp 1 / lane 0 : tcgen05.mma issuer  (UMMA 128 x block_n x 16, bf16 -> fp32)
//   warps 2..9      : epilogue            (tcgen05.ld -> *scale +bias -> act -> +residual -> act2 -> store)
#include <cuda.h>
#include <algorithm>

#include "ptx.cuh"
#include "yc_common.cuh"

namespace yc {

constexpr int kBlockM = 128;
constexpr int kMaxStages = 16;
constexpr int kThreads = 64 + 32 * 16;  // producer warp, MMA warp, 16 epilogue warps

struct ConvParams {
  int M, N, n_store;
  int taps, ksize, cchunks, block_k, block_n, stages;
  int m_tiles, n_tiles;
  int a_mode;  // 0: 2-D tiled map (1x1 s1 p0), 1: im2col map, 2: halo patch (3x3 s1 p1, see kHalo*)
  int tiles_h, tiles_w;  // a_mode 2: output tiles of kHaloH x kHaloW pixels per image
  // exact division of a tile index (< 2^20) by tiles_h*tiles_w and by tiles_w: q = (n * magic) >> shift
  uint32_t tpi_magic, tpi_shift, tw_magic, tw_shift;
  int w_resident;        // a_mode 2: all 9*cchunks weight slabs stay in shared memory for the whole (persistent) CTA
  int halo_stages;       // a_mode 2: patch ring depth (2..kHaloMaxStages)
  int halo_pair;         // a_mode 2 + w_resident + one chunk: two tiles' MMAs interleaved into two accumulators
  int tile_pair;         // a_mode 0 / 1, one N tile, block_n <= 128: the same on the tiled / im2col path
  int issue_v2;          // halo mode, resident weights: unrolled tap issue (halo_taps_resident); 0 = the rolled loop (A/B)
  int Ho, Wo, stride, pad, dil;
  int kw, stride_w, pad_w;  // kernel width / stride / left padding along W (== ksize / stride / pad unless rectangular)
  const float* bias;   // padded to n_tiles*block_n
  const float* scale;  // same padding, or nullptr
  int y_f32;
  int chunk;           // output columns per TMA store box: 64 or 32 (bf16), 32 (fp32)
  const __nv_bfloat16* res;
  long long ldres;
  int act, act2;
  uint32_t tmem_cols;
  int epi_sets;        // 1, 2 or 4 groups of epilogue warps working on different tiles at the same time
  int nacc;            // TMEM accumulator buffers in use (>= epi_sets)
  // training-mode batchnorm statistics fused into the epilogue (fast_mean_gpu / fast_variance_gpu,
  // blas_kernels.cu:534-616): per-channel sum and sum of squares of the fp32 results, nullptr = off
  float* stat_sum;
  float* stat_sq;
  int stat_cols;       // n_tiles * block_n: floats per statistic in the CTA's shared-memory accumulator
  // Reproducible statistics (no float atomics anywhere): inside a CTA every (TMEM quadrant, epilogue set) pair of warps
  // accumulates into its OWN shared-memory region (the warps of a pair own disjoint columns), the regions are added in
  // index order into the CTA's partial, partials go to stat_partial[cta][2][stat_cols], and the CTA that draws the last
  // ticket adds the partials in CTA order into stat_sum / stat_sq (overwriting them).
  int stat_regions;    // 4 * epi_sets
  float* stat_partial;
  unsigned* stat_ticket;
  // optional (yc_conv_plan_set_bn_finalize): the CTA that adds the partials also turns the sums into batch mean / unbiased
  // variance and updates the rolling averages — yc_bn_finalize's arithmetic without its launch
  float* bn_mean; float* bn_var; float* bn_rmean; float* bn_rvar;
  float bn_count;
  int out_bufs;        // epilogue staging buffers per warp (1 or 2): with 2 a warp stages block k+1 while the TMA
                       // engine still reads block k
  int split_col;       // > 0: output columns >= split_col go to a second tensor (tmap_y2) at column - split_col: two
                       // 1x1 convolutions that read the same input (the CSP branch pairs) run as ONE GEMM
  int fast_epi;        // the specialised epilogue (epilogue_fast) applies: bf16 output, bias only, mish / leaky / linear
  int mish_nr;         // specialised epilogue: mish with a Newton reciprocal (one MUFU per output instead of two)
  int bias_cols;       // > 0: that many floats of bias (n_tiles * block_n) are staged in shared memory at kernel start
  int cta_pair;        // conv_fprop_pair_kernel: tiles of 256 rows over a cluster of two CTAs (tcgen05 cta_group::2);
                       // m_tiles then counts 256-row tiles
  // YOLO head (yc_conv_plan_set_yolo_decode): the fp32 results are not stored NHWC but decoded (logistic / scale_x_y per
  // entry, yolo_layer.cpp:399-404) and written straight into the [yolo] layer's NCHW fp32 output — ForwardYoloLayerGpu
  // inside the head convolution's epilogue, no intermediate tensor, no decode kernel
  float* dec_out;
  int dec_S, dec_E, dec_gauss;
  float dec_sxy;
  int m_tiles128;      // CTA pairs: the number of real 128-row tiles (an odd count leaves the last pair's second CTA idle)
};

// ---- halo mode (a_mode 2): 3x3 stride-1 pad-1 convolutions whose A operand is staged ONCE per (tile, 64-channel chunk)
// as a (kHaloH+2) x kHaloPitch pixel patch (one tiled TMA box, border zero-filled by the engine) instead of nine im2col
// tap loads.  M tile = kHaloH x kHaloW output pixels, row r = h*8 + w; the tap (kh, kw) operand is the same patch
// read through a descriptor shifted by (kh*kHaloPitch + kw) rows of 128 B: 8-row groups (8 consecutive pixels of one
// patch row) are kHaloPitch*128 B apart (SBO).  Measured on B200 (tools/halo_debug.py): the tensor core applies
// the 128B-swizzle XOR to the ABSOLUTE shared-memory address of each row, like TMA did when writing the patch, so a
// start address shifted by whole rows needs NO base-offset in the descriptor (setting it to (addr>>7)&7 breaks kw != 0).
constexpr int kHaloH = 16, kHaloW = 8, kHaloPitch = 16;  // pitch 10 (exact halo) works but runs 1.5x slower: 8-row groups then straddle 1024-byte swizzle atoms
constexpr int kHaloPatchRows = (kHaloH + 2) * kHaloPitch;  // 288 pixel rows of block_k*2 bytes (128 B, or 64 B when Cin % 64 != 0)
constexpr int kHaloPatchStride = kHaloPatchRows * 128;     // ring slot pitch (36864, 1024-byte aligned), sized for 128-byte rows
constexpr int kHaloMaxStages = 4;
constexpr int kStageOutBytes = 2048;  // one epilogue staging buffer: 32 rows x 64 B (a warp owns out_bufs of them, contiguous)
constexpr int kEpiWarps = 16;         // four per TMEM lane quadrant == four per scheduler: enough to hide ALU/MUFU latency
constexpr int kMaxAcc = 4;            // TMEM accumulator buffers

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

// tile index -> (M tile, N tile); nacc and epi_sets are powers of two.  Runtime integer divisions cost ~25 instructions
// each (two of them on the XU pipe the mish epilogue lives on): ncu on a 1x1 64->64 layer at 304^2 showed 40 % of all
// executed instructions in per-tile index arithmetic before these were removed.
__device__ __forceinline__ void tile_mn(const ConvParams& p, int t, int& mt, int& nt) {
  if (p.n_tiles == 1) {
    mt = t;
    nt = 0;
  } else {
    mt = t / p.n_tiles;
    nt = t - mt * p.n_tiles;
  }
}

// n / d for n < 2^20 through the multiplier the host derived from d (set_fastdiv): one wide multiply and a shift
// instead of the ~20-instruction runtime division (two MUFU among them, on the pipe the mish epilogue lives on).
__device__ __forceinline__ int fastdiv(int n, uint32_t magic, uint32_t shift) {
  return (int)(((unsigned long long)(uint32_t)n * magic) >> shift);
}

template <int ACT>
__device__ __forceinline__ float act_ct(float x) {
  if (ACT == YC_LINEAR) return x;
  if (ACT == YC_LEAKY) return fmaxf(x, 0.1f * x);
  if (ACT == YC_MISH) return act_mish(x);
  if (ACT == YC_LOGISTIC) return act_logistic(x);
  if (ACT == YC_SWISH) return x * act_logistic(x);
  if (ACT == YC_RELU) return x > 0.f ? x : 0.f;
  return x;
}

// act(x) for two values; same formulas as act_ct (mish: yc_common.cuh act_mish)
template <int ACT>
__device__ __forceinline__ void act_ct2(f32x2 x, float& y0, float& y1) {
  float x0, x1;
  unpack2(x, x0, x1);
  if (ACT == YC_MISH) {  // x + x*r, r = 1/(-q/2) (yc_common.cuh act_mish): 4 packed fma + 4 MUFU per pair
    float a0, a1;
    unpack2(fma2(x, pack2(1.4426950408889634f, 1.4426950408889634f), pack2(-1.f, -1.f)), a0, a1);
    const f32x2 h = pack2(ex2_ftz(a0), ex2_ftz(a1));
    float q0, q1;
    unpack2(fma2(h, fma2(h, pack2(-2.f, -2.f), pack2(-2.f, -2.f)), pack2(-1.f, -1.f)), q0, q1);
    unpack2(fma2(x, pack2(rcp_ftz(q0), rcp_ftz(q1)), x), y0, y1);
  } else if (ACT == YC_LEAKY) {
    float m0, m1;
    unpack2(mul2(x, pack2(0.1f, 0.1f)), m0, m1);
    y0 = fmaxf(x0, m0);
    y1 = fmaxf(x1, m1);
  } else {
    y0 = act_ct<ACT>(x0);
    y1 = act_ct<ACT>(x1);
  }
}

// Epilogue of one warp over all tiles of this CTA.  The warp owns TMEM lane quadrant `quad` (rows) and every
// second 32-column group (`half`):
//   tcgen05.ld 32 columns -> *scale + bias -> act -> (+residual -> act2) -> bf16/fp32 -> swizzled smem ->
//   TMA store of a 32-row x 64-byte box (bf16: 32 columns, fp32: 16 columns); M / channel tails are clipped by
//   the tensor map.  No inter-warp synchronisation: each warp has its own two staging buffers.
template <int ACT, bool STATS = false>
__device__ __forceinline__ void epilogue_role(const ConvParams& p, const CUtensorMap* tmap_y, const CUtensorMap* tmap_y2,
                                              uint32_t tmem_base, uint32_t tfull0, uint32_t tempty0, uint32_t sbuf0, int quad,
                                              int slot, int lane, int total_tiles, uint32_t stat_smem) {
  // Narrow tiles (block_n 32 / 64) have fewer 32-column groups than the four warps of a quadrant: the warps then
  // split into `epi_sets` sets that take alternate tiles, so all sixteen stay busy.
  const int wps = 4 / p.epi_sets;        // warps of one quadrant working on the same tile
  const int nacc_shift = 31 - __clz(p.nacc);
  const int set = slot / wps, sub = slot - set * wps;
  const uint32_t rowoff = (uint32_t)lane * 64u;
  const uint32_t swz = ((rowoff >> 7) & 3u) << 4;  // 64B swizzle: 16-byte chunk index ^= (row / 2) % 4
  const bool has_res = p.res != nullptr;
  const bool res_coal = has_res && !p.y_f32;
  // staging buffers: store number `nst` of this warp goes through buffer nst & (out_bufs - 1); before a buffer is
  // written again at most out_bufs - 1 younger stores may still be reading theirs
  const bool two_bufs = p.out_bufs == 2;
  uint32_t nst = 0;
  uint32_t sbuf = sbuf0;
  auto wait_buf = [&]() {
    if (lane == 0) {
      if (two_bufs) ptx::tma_store_wait_read<1>();
      else ptx::tma_store_wait_read<0>();
    }
    __syncwarp();
    sbuf = sbuf0 + (two_bufs ? (nst & 1u) * (uint32_t)kStageOutBytes : 0u);
  };
  // this warp's set takes every epi_sets-th tile of the CTA: it = set, set + epi_sets, ... (it = index of the tile in
  // the CTA's sequence, which is what the MMA warp's accumulator rotation counts)
  const int t_step = (int)gridDim.x * p.epi_sets;
  for (int t = blockIdx.x + set * (int)gridDim.x, it = set; t < total_tiles; t += t_step, it += p.epi_sets) {
    const int acc = it & (p.nacc - 1);
    const uint32_t acc_phase = (it >> nacc_shift) & 1;
    int mt_, nt_;
    tile_mn(p, t, mt_, nt_);
    const int m0 = mt_ * kBlockM;
    const int n0 = nt_ * p.block_n;
    int row0 = m0 + quad * 32;
    long long row = (long long)row0 + lane;
    bool row_ok = row < p.M;
    int hb = 0, hw0 = 0, hh0 = 0;  // halo mode: image, first column / row of this warp's 4 x 8 pixel block
    if (p.a_mode == 2) {
      const int mt = mt_;
      hb = fastdiv(mt, p.tpi_magic, p.tpi_shift);
      const int rem = mt - hb * (p.tiles_h * p.tiles_w);
      const int th = fastdiv(rem, p.tw_magic, p.tw_shift);
      hh0 = th * kHaloH + quad * 4;
      hw0 = (rem - th * p.tiles_w) * kHaloW;
      const int h = hh0 + (lane >> 3), w = hw0 + (lane & 7);
      row_ok = h < p.Ho && w < p.Wo;
      row = ((long long)hb * p.Ho + h) * p.Wo + w;
      row0 = hh0 < p.Ho ? 0 : p.M;  // only used as "this block has something to store"
    }
    const uint32_t t_row = tmem_base + ((uint32_t)(quad * 32) << 16) + acc * p.block_n;
    int n_groups = (min(p.block_n, p.n_store - n0) + 31) >> 5;  // 32-column groups with something to store
    // this warp's groups: sub, sub+wps, ...
    int last_g = -1;
    for (int g = sub; g < n_groups; g += wps) last_g = g;
    // Shortcut operand (bf16 output): this warp's 32 rows x 64 B go straight from global into its staging buffer with
    // 16-byte asynchronous copies, COALESCED (copy j of a lane = chunk (lane & 3) of row (lane >> 2) + 8 j: 8 rows x 64 B
    // per instruction; a row-per-lane fetch costs 32 L1 wavefronts per request), 64B-swizzled like the output staging,
    // no registers held.  The first block of a tile is requested BEFORE waiting for the accumulator.
    auto issue_res = [&](int g) {
      wait_buf();  // the store that used this buffer last has finished reading it
      const int c0 = n0 + g * 32;
      const int rr = lane >> 2, q = lane & 3;
      const bool col_ok = c0 + q * 8 + 8 <= p.n_store;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        long long grow;
        bool ok;
        if (p.a_mode == 2) {
          const int h = hh0 + j, w = hw0 + rr;
          ok = h < p.Ho && w < p.Wo;
          grow = ((long long)hb * p.Ho + h) * p.Wo + w;
        } else {
          grow = (long long)row0 + rr + 8 * j;
          ok = grow < p.M;
        }
        ok = ok && col_ok;
        const uint32_t r = (uint32_t)(rr + 8 * j);
        ptx::cp_async_16(sbuf + r * 64u + (((uint32_t)q ^ ((r >> 1) & 3u)) << 4),
                         ok ? (const void*)(p.res + grow * p.ldres + c0 + q * 8) : (const void*)p.res, ok ? 16u : 0u);
      }
      ptx::cp_async_commit();
    };
    if (res_coal && last_g >= 0) issue_res(sub);
    ptx::mbar_wait(tfull0 + 8u * acc, acc_phase);
    ptx::tc_fence_after();
    if (last_g < 0) {  // nothing to do in this tile: still hand the accumulator back
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(tempty0 + 8u * acc);
      continue;
    }
    for (int g = sub; g < n_groups; g += wps) {
      const int col0 = n0 + g * 32;
      if (res_coal && g != sub) issue_res(g);
      uint32_t r[32];
      ptx::tmem_ld_32x32b_x32(t_row + g * 32, r);
      // residual rows: bf16 output -> already on their way into this warp's staging buffer (issue_res); fp32 output ->
      // fetched row-per-lane while the TMEM load is in flight
      uint4 rq[4];
      if (has_res && !res_coal) {
        const __nv_bfloat16* rp = p.res + row * p.ldres + col0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          rq[j] = make_uint4(0, 0, 0, 0);
          if (row_ok && col0 + j * 8 + 8 <= p.n_store) rq[j] = __ldg(reinterpret_cast<const uint4*>(rp + j * 8));
        }
      }
      ptx::tmem_ld_wait();
      if (g == last_g) {  // accumulator fully read by this warp: hand it back to the MMA warp
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(tempty0 + 8u * acc);
      }
      float v[32];
      if (p.scale) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + j));
          const float4 s4 = __ldg(reinterpret_cast<const float4*>(p.scale + col0 + j));
          act_ct2<ACT>(fma2(pack2(__uint_as_float(r[j + 0]), __uint_as_float(r[j + 1])), pack2(s4.x, s4.y), pack2(b4.x, b4.y)),
                       v[j + 0], v[j + 1]);
          act_ct2<ACT>(fma2(pack2(__uint_as_float(r[j + 2]), __uint_as_float(r[j + 3])), pack2(s4.z, s4.w), pack2(b4.z, b4.w)),
                       v[j + 2], v[j + 3]);
        }
      } else {  // batchnorm folded into the weights (inference): bias only, no unit scale to materialise
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + j));
          act_ct2<ACT>(add2(pack2(__uint_as_float(r[j + 0]), __uint_as_float(r[j + 1])), pack2(b4.x, b4.y)), v[j + 0], v[j + 1]);
          act_ct2<ACT>(add2(pack2(__uint_as_float(r[j + 2]), __uint_as_float(r[j + 3])), pack2(b4.z, b4.w)), v[j + 2], v[j + 3]);
        }
      }
      if (STATS) {
        // column sums over this warp's 32 rows by a transposing butterfly: 31 shuffles per statistic, after
        // which lane j holds column col0+j.  Rows past M contribute nothing.  Accumulated per CTA in shared
        // memory and flushed once at kernel end (grid x N atomics instead of tiles x N).
        float red[2];
#pragma unroll
        for (int pass = 0; pass < 2; ++pass) {  // pass 0: sum, pass 1: sum of squares (sequential: fewer live registers)
          float a[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float x = row_ok ? v[j] : 0.f;
            a[j] = pass ? x * x : x;
          }
#pragma unroll
          for (int off = 16; off >= 1; off >>= 1) {
            const bool upper = (lane & off) != 0;
#pragma unroll
            for (int j = 0; j < off; ++j) {
              const float send = upper ? a[j] : a[j + off], keep = upper ? a[j + off] : a[j];
              a[j] = keep + __shfl_xor_sync(0xffffffffu, send, off);
            }
          }
          red[pass] = a[0];
        }
        const uint32_t so = stat_smem + 4u * (uint32_t)((quad * p.epi_sets + set) * 2 * p.stat_cols + col0 + lane);
        asm volatile("red.shared.add.f32 [%0], %1;" ::"r"(so), "f"(red[0]) : "memory");
        asm volatile("red.shared.add.f32 [%0], %1;" ::"r"(so + 4u * (uint32_t)p.stat_cols), "f"(red[1]) : "memory");
      }
      if (has_res) {
        if (res_coal) {  // the block staged by issue_res has landed: every lane reads its own row
          ptx::cp_async_wait_all();
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 4; ++j) rq[j] = ptx::ld_shared_v4(sbuf + rowoff + (((uint32_t)j * 16u) ^ swz));
          __syncwarp();  // all lanes have read their rows before the results are staged in the same buffer
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t w4[4] = {rq[j].x, rq[j].y, rq[j].z, rq[j].w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {  // bf16 pair -> fp32 pair (shift / mask), one packed add
            const int k = j * 8 + 2 * e;
            unpack2(add2(pack2(v[k], v[k + 1]), pack2(__uint_as_float(w4[e] << 16), __uint_as_float(w4[e] & 0xffff0000u))),
                    v[k], v[k + 1]);
          }
        }
        if (p.act2 != YC_LINEAR) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = apply_act(v[j], p.act2);
        }
      }
      // second output tensor of a fused pair: same rows, columns counted from split_col
      const bool second = p.split_col > 0 && col0 >= p.split_col;
      const CUtensorMap* ty = second ? tmap_y2 : tmap_y;
      const int scol = second ? col0 - p.split_col : col0;
      if (!p.y_f32) {
        if (!res_coal) wait_buf();  // (with a staged shortcut operand the buffer is already this block's)
#pragma unroll
        for (int j = 0; j < 4; ++j)
          ptx::st_shared_v4(sbuf + rowoff + (((uint32_t)j * 16u) ^ swz), pack_bf16x2(v[8 * j], v[8 * j + 1]),
                            pack_bf16x2(v[8 * j + 2], v[8 * j + 3]), pack_bf16x2(v[8 * j + 4], v[8 * j + 5]),
                            pack_bf16x2(v[8 * j + 6], v[8 * j + 7]));
        ptx::fence_proxy_async();  // generic-proxy smem writes -> visible to the TMA engine
        __syncwarp();
        if (lane == 0) {  // always commit (possibly an empty group): the buffer rotation counts groups
          if (row0 < p.M) {
            if (p.a_mode == 2) ptx::tma_store_4d(ty, sbuf, scol, hw0, hh0, hb);
            else ptx::tma_store_2d(ty, sbuf, scol, row0);
          }
          ptx::tma_store_commit();
        }
        ++nst;
      } else if (p.dec_out) {
        // fused YOLO decode: lane = pixel, so for a fixed channel the 32 lanes write 32 consecutive floats of that
        // channel's plane (128-byte requests without any transpose)
        const int img = (int)(row / p.dec_S);
        float* ob = p.dec_out + ((long long)img * p.N) * p.dec_S + (row - (long long)img * p.dec_S);
        const float dbias = -0.5f * (p.dec_sxy - 1.f);
        int e = col0 % p.dec_E;
#pragma unroll 4
        for (int j = 0; j < 32; ++j) {
          const int c = col0 + j;
          bool sig, aff;
          if (p.dec_gauss) { sig = (e <= 3) || e == 5 || e == 7 || e >= 8; aff = (e == 0) || (e == 2); }
          else { sig = (e < 2) || (e >= 4); aff = (e < 2); }
          float yv = v[j];
          if (sig) yv = act_logistic(yv);  // MUFU ex2 + rcp: within 2e-7 of the libm form (with expf + a division this
                                           // epilogue took 546 us on the 76x76 head against 174 + 176 us unfused)
          if (aff) yv = yv * p.dec_sxy + dbias;
          if (row_ok && c < p.N) ob[(long long)c * p.dec_S] = yv;
          if (++e == p.dec_E) e = 0;
        }
      } else {
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {  // fp32: two boxes of 16 columns
          wait_buf();
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int e = hh * 16 + j * 4;
            ptx::st_shared_v4(sbuf + rowoff + (((uint32_t)j * 16u) ^ swz), __float_as_uint(v[e]), __float_as_uint(v[e + 1]),
                              __float_as_uint(v[e + 2]), __float_as_uint(v[e + 3]));
          }
          ptx::fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            if (row0 < p.M && col0 + hh * 16 < p.n_store) {
              if (p.a_mode == 2) ptx::tma_store_4d(ty, sbuf, scol + hh * 16, hw0, hh0, hb);
              else ptx::tma_store_2d(ty, sbuf, scol + hh * 16, row0);
            }
            ptx::tma_store_commit();
          }
          ++nst;
        }
      }
    }
  }
  if (lane == 0) ptx::tma_store_wait<0>();  // smem must stay valid until the last store has read it
}


// ---- specialised epilogue: bf16 output, batchnorm folded (bias only), mish / leaky / linear, optional shortcut ----
// ncu on the 64-channel layers at 304^2 (round 2, profiles/r02_*): the generic epilogue executes ~405 warp instructions
// per 32 x 32 output block of which 129 are the activation math; the rest is per-tile / per-block control (runtime
// switches on the A mode, output type, shortcut, statistics; 64-bit row arithmetic; tile index math) issued as one long
// dependent chain, during which the warp does not feed the XU pipe (MUFU) that bounds these layers: XU 60 % busy.
// Here everything that does not change inside the kernel is a template parameter or hoisted, the bias comes from shared
// memory (one broadcast LDS.128 per four columns instead of an LDG), one warp works through a whole 32-row strip of a
// tile (the per-tile work is paid once per 2-8 blocks, not once per block), and the mish math runs in phases of eight
// independent pairs (16 MUFU.EX2, then 16 MUFU.RCP back to back) so a single warp keeps the XU pipe busy.
constexpr int kMishNR = 117;  // epilogue_fast only: mish with the reciprocal on the FMA pipe (Newton) instead of MUFU.RCP

// PAIR (conv_fprop_pair_kernel, cta_group::2): the tile is 256 GEMM rows over two CTAs, this CTA (cluster rank `rank`)
// owns rows (2 * mt + rank) * 128 ..., tiles are walked per CTA pair (cta_first = pair index, cta_stride = pairs), and the
// accumulator is handed back on the LEADER's barrier (tempty0 is then a shared::cluster address).
template <int ACT, bool HALO, bool RES, bool PAIR = false>
__device__ __forceinline__ void epilogue_fast(const ConvParams& p, const CUtensorMap* tmap_y, const CUtensorMap* tmap_y2,
                                              uint32_t tmem_base, uint32_t tfull0, uint32_t tempty0, uint32_t sbuf0,
                                              uint32_t bias_smem, int quad, int slot, int lane, int total_tiles,
                                              int cta_first, int cta_stride, int rank) {
  const int sets = p.epi_sets, wps = 4 / sets;
  const int set = slot / wps, sub = slot - set * wps;
  const int nacc_mask = p.nacc - 1, nacc_shift = 31 - __clz(p.nacc);
  const uint32_t rowoff = (uint32_t)lane * 64u;
  const uint32_t swz = ((rowoff >> 7) & 3u) << 4;
  const bool two_bufs = p.out_bufs == 2;
  const int M = p.M, n_store = p.n_store, block_n = p.block_n, n_tiles = p.n_tiles, split = p.split_col;
  const int Ho = p.Ho, Wo = p.Wo;
  const int tpi = p.tiles_h * p.tiles_w;
  const __nv_bfloat16* res = p.res;
  const long long ldres = p.ldres;
  // one GENERIC pointer for both cases (a single load path in the block loop): the shared-memory copy or global memory
  const float* bias_ptr = p.bias_cols > 0 ? ptx::smem_generic<float>(bias_smem) : p.bias;
  uint32_t nst = 0, sbuf = sbuf0;
  auto wait_buf = [&]() {
    if (lane == 0) {
      if (two_bufs) ptx::tma_store_wait_read<1>();
      else ptx::tma_store_wait_read<0>();
    }
    __syncwarp();
    sbuf = sbuf0 + (two_bufs ? (nst & 1u) * (uint32_t)kStageOutBytes : 0u);
  };
  const int t_step = cta_stride * sets;
  auto give_back = [&](int acc) {  // one arrival per epilogue warp on the accumulator's "empty" barrier
    if (PAIR) ptx::mbar_arrive_cluster(tempty0 + 8u * acc);
    else ptx::mbar_arrive(tempty0 + 8u * acc);
  };
  for (int t = cta_first + set * cta_stride, it = set; t < total_tiles; t += t_step, it += sets) {
    const int acc = it & nacc_mask;
    const uint32_t acc_phase = (uint32_t)(it >> nacc_shift) & 1u;
    int mt = t, nt = 0;
    if (n_tiles != 1) {
      mt = t / n_tiles;
      nt = t - mt * n_tiles;
    }
    if (PAIR) mt = 2 * mt + rank;
    const int n0 = nt * block_n;
    int row0 = mt * kBlockM + quad * 32;   // first GEMM row of this warp's strip (tiled / im2col)
    int hb = 0, hw0 = 0, hh0 = 0;          // halo: image, first column / row of this warp's 4 x 8 pixel block
    bool any_rows;
    bool real = true;  // PAIR + HALO: an odd number of 128-pixel tiles leaves the last pair's second CTA without a tile
    if (HALO) {
      hb = fastdiv(mt, p.tpi_magic, p.tpi_shift);
      const int rem = mt - hb * tpi;
      const int th = fastdiv(rem, p.tw_magic, p.tw_shift);
      hh0 = th * kHaloH + quad * 4;
      hw0 = (rem - th * p.tiles_w) * kHaloW;
      if (PAIR) real = mt < p.m_tiles128;
      any_rows = real && hh0 < Ho;
    } else {
      any_rows = row0 < M;
    }
    const int n_groups = (min(block_n, n_store - n0) + 31) >> 5;
    const uint32_t t_row = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * block_n);
    // shortcut operand of group g: 32 rows x 64 B straight from global into this warp's staging buffer, coalesced
    // 16-byte cp.async copies (copy j of a lane = chunk (lane & 3) of row (lane >> 2) + 8 j), 64B-swizzled
    auto issue_res = [&](int g) {
      wait_buf();
      const int c0 = n0 + g * 32;
      const int rr = lane >> 2, q = lane & 3;
      const bool col_ok = c0 + q * 8 + 8 <= n_store;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        long long grow;
        bool ok;
        if (HALO) {
          const int h = hh0 + j, w = hw0 + rr;
          ok = real && h < Ho && w < Wo;
          grow = ((long long)hb * Ho + h) * Wo + w;
        } else {
          grow = (long long)row0 + rr + 8 * j;
          ok = grow < M;
        }
        ok = ok && col_ok;
        const uint32_t r = (uint32_t)(rr + 8 * j);
        ptx::cp_async_16(sbuf + r * 64u + (((uint32_t)q ^ ((r >> 1) & 3u)) << 4),
                         ok ? (const void*)(res + grow * ldres + c0 + q * 8) : (const void*)res, ok ? 16u : 0u);
      }
      ptx::cp_async_commit();
    };
    if (RES && sub < n_groups) issue_res(sub);
    ptx::mbar_wait(tfull0 + 8u * acc, acc_phase);
    ptx::tc_fence_after();
    if (sub >= n_groups) {  // nothing of this tile is this warp's: still hand the accumulator back
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) give_back(acc);
      continue;
    }
    for (int g = sub; g < n_groups; g += wps) {
      const int col0 = n0 + g * 32;
      uint32_t r[32];
      ptx::tmem_ld_32x32b_x32(t_row + g * 32, r);
      ptx::tmem_ld_wait();
      if (g + wps >= n_groups) {  // this warp's last read of the accumulator
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) give_back(acc);
      }
      uint4 rq[4];
      if (RES) {  // the block staged by issue_res has landed: every lane reads its own row, then the buffer is reused
        ptx::cp_async_wait_all();
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 4; ++j) rq[j] = ptx::ld_shared_v4(sbuf + rowoff + (((uint32_t)j * 16u) ^ swz));
        __syncwarp();
      } else {
        wait_buf();
      }
      const float* bcol = bias_ptr + col0;
#pragma unroll
      for (int hf = 0; hf < 2; ++hf) {  // two halves of 16 columns = 8 packed pairs each
        f32x2 x[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          // bias: broadcast 16-byte load from the copy staged in shared memory at kernel start, or (where that copy would
          // have cost a ring stage: wide K-heavy tiles, whose epilogue is hidden anyway) from global memory / L1
          const uint4 b4 = *reinterpret_cast<const uint4*>(bcol + hf * 16 + j * 4);
          x[2 * j] = add2(pack2(__uint_as_float(r[hf * 16 + 4 * j]), __uint_as_float(r[hf * 16 + 4 * j + 1])),
                          pack2(__uint_as_float(b4.x), __uint_as_float(b4.y)));
          x[2 * j + 1] = add2(pack2(__uint_as_float(r[hf * 16 + 4 * j + 2]), __uint_as_float(r[hf * 16 + 4 * j + 3])),
                              pack2(__uint_as_float(b4.z), __uint_as_float(b4.w)));
        }
        if (ACT == YC_MISH) {  // x + x*r, r = 1/(-q/2), q = n(n+2)+2 on h = n/2 = 2^(x*log2e - 1)   (yc_common.cuh act_mish)
          float h[16];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            float a0, a1;
            unpack2(fma2(x[j], pack2(1.4426950408889634f, 1.4426950408889634f), pack2(-1.f, -1.f)), a0, a1);
            h[2 * j] = ex2_ftz(a0);
            h[2 * j + 1] = ex2_ftz(a1);
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const f32x2 hh = pack2(h[2 * j], h[2 * j + 1]);
            float q0, q1;
            unpack2(fma2(hh, fma2(hh, pack2(-2.f, -2.f), pack2(-2.f, -2.f)), pack2(-1.f, -1.f)), q0, q1);
            h[2 * j] = rcp_ftz(q0);
            h[2 * j + 1] = rcp_ftz(q1);
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) x[j] = fma2(x[j], pack2(h[2 * j], h[2 * j + 1]), x[j]);
        } else if (ACT == kMishNR) {
          // mish(x) = x * t / (t + 1) with t = n(n+2)/2, n = e^x — the same function without the second MUFU: the
          // reciprocal of d = t + 1 >= 1 comes from the exponent trick (seed, 5 % error) and two Newton steps on the FMA
          // pipe (6.6e-6 relative; as a PRODUCT x*y*t there is no cancellation, the result is accurate to 9e-6 relative
          // everywhere = 1/450 of a bf16 half-ulp).  OPT-IN (YC_PLAN_MISH_NEWTON): measured on yolov4-608 x 64 it is 5-8 %
          // SLOWER than the two-MUFU form on every mish layer (stem 564 vs 524 us, 1x1 64->64 @304 289 vs 269 us): with the XU
          // pipe at 80-85 % the epilogue is as much issue-bound as MUFU-bound, and this form issues 16 instructions per
          // pair instead of 9.
          // h = 2^min(x*log2e - 1, 30): beyond that mish(x) == x in fp32, and d stays finite.
          f32x2 hh[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            float a0, a1;
            unpack2(fma2(x[j], pack2(1.4426950408889634f, 1.4426950408889634f), pack2(-1.f, -1.f)), a0, a1);
            hh[j] = pack2(ex2_ftz(fminf(a0, 30.f)), ex2_ftz(fminf(a1, 30.f)));
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const f32x2 t = mul2(hh[j], fma2(hh[j], pack2(2.f, 2.f), pack2(2.f, 2.f)));  // n(n+2)/2
            const f32x2 nd = fma2(t, pack2(-1.f, -1.f), pack2(-1.f, -1.f));             // -(t + 1)
            float nd0, nd1;
            unpack2(nd, nd0, nd1);
            // 1/d seed: magic - bits(d), with bits(d) = bits(-d) ^ sign folded into the constant
            f32x2 y = pack2(__int_as_float((int)(0xFEF311C7u - (unsigned)__float_as_int(nd0))),
                            __int_as_float((int)(0xFEF311C7u - (unsigned)__float_as_int(nd1))));
            y = mul2(y, fma2(nd, y, pack2(2.f, 2.f)));
            y = mul2(y, fma2(nd, y, pack2(2.f, 2.f)));
            x[j] = mul2(mul2(x[j], y), t);
          }
        } else if (ACT == YC_LEAKY) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            float x0, x1, m0, m1;
            unpack2(x[j], x0, x1);
            unpack2(mul2(x[j], pack2(0.1f, 0.1f)), m0, m1);
            x[j] = pack2(fmaxf(x0, m0), fmaxf(x1, m1));
          }
        }
        if (RES) {
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const uint4 q4 = rq[hf * 2 + j];
            const uint32_t w4[4] = {q4.x, q4.y, q4.z, q4.w};
#pragma unroll
            for (int e = 0; e < 4; ++e)  // bf16 pair -> fp32 pair (shift / mask), one packed add
              x[j * 4 + e] = add2(x[j * 4 + e], pack2(__uint_as_float(w4[e] << 16), __uint_as_float(w4[e] & 0xffff0000u)));
          }
        }
        uint32_t o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float y0, y1;
          unpack2(x[j], y0, y1);
          o[j] = pack_bf16x2(y0, y1);
        }
        ptx::st_shared_v4(sbuf + rowoff + (((uint32_t)(hf * 2) * 16u) ^ swz), o[0], o[1], o[2], o[3]);
        ptx::st_shared_v4(sbuf + rowoff + (((uint32_t)(hf * 2 + 1) * 16u) ^ swz), o[4], o[5], o[6], o[7]);
      }
      ptx::fence_proxy_async();  // generic-proxy smem writes -> visible to the TMA engine
      __syncwarp();
      if (lane == 0) {  // always commit (possibly an empty group): the buffer rotation counts groups
        if (any_rows) {
          const bool second = split > 0 && col0 >= split;
          const CUtensorMap* ty = second ? tmap_y2 : tmap_y;
          const int scol = second ? col0 - split : col0;
          if (HALO) ptx::tma_store_4d(ty, sbuf, scol, hw0, hh0, hb);
          else ptx::tma_store_2d(ty, sbuf, scol, row0);
        }
        ptx::tma_store_commit();
      }
      ++nst;
      if (RES && g + wps < n_groups) issue_res(g + wps);
    }
  }
  if (lane == 0) ptx::tma_store_wait<0>();  // smem must stay valid until the last store has read it
}

template <int ACT>
__device__ __forceinline__ void epilogue_fast_dispatch(const ConvParams& p, const CUtensorMap* tmap_y, const CUtensorMap* tmap_y2,
                                                       uint32_t tmem_base, uint32_t tfull0, uint32_t tempty0, uint32_t sbuf0,
                                                       uint32_t bias_smem, int quad, int slot, int lane, int total_tiles) {
  const bool halo = p.a_mode == 2, has_res = p.res != nullptr;
  const int f = (int)blockIdx.x, st = (int)gridDim.x;
  if (halo) {
    if (has_res) epilogue_fast<ACT, true, true>(p, tmap_y, tmap_y2, tmem_base, tfull0, tempty0, sbuf0, bias_smem, quad, slot, lane, total_tiles, f, st, 0);
    else epilogue_fast<ACT, true, false>(p, tmap_y, tmap_y2, tmem_base, tfull0, tempty0, sbuf0, bias_smem, quad, slot, lane, total_tiles, f, st, 0);
  } else {
    if (has_res) epilogue_fast<ACT, false, true>(p, tmap_y, tmap_y2, tmem_base, tfull0, tempty0, sbuf0, bias_smem, quad, slot, lane, total_tiles, f, st, 0);
    else epilogue_fast<ACT, false, false>(p, tmap_y, tmap_y2, tmem_base, tfull0, tempty0, sbuf0, bias_smem, quad, slot, lane, total_tiles, f, st, 0);
  }
}
// the same for the CTA-pair kernel (tiled / im2col operands only)
template <int ACT>
__device__ __forceinline__ void epilogue_pair_dispatch(const ConvParams& p, const CUtensorMap* tmap_y, const CUtensorMap* tmap_y2,
                                                       uint32_t tmem_base, uint32_t tfull0, uint32_t tempty0_leader, uint32_t sbuf0,
                                                       uint32_t bias_smem, int quad, int slot, int lane, int total_tiles,
                                                       int pair, int npairs, int rank) {
  if (p.a_mode == 2) {
    if (p.res != nullptr) epilogue_fast<ACT, true, true, true>(p, tmap_y, tmap_y2, tmem_base, tfull0, tempty0_leader, sbuf0, bias_smem, quad, slot, lane, total_tiles, pair, npairs, rank);
    else epilogue_fast<ACT, true, false, true>(p, tmap_y, tmap_y2, tmem_base, tfull0, tempty0_leader, sbuf0, bias_smem, quad, slot, lane, total_tiles, pair, npairs, rank);
  } else {
    if (p.res != nullptr) epilogue_fast<ACT, false, true, true>(p, tmap_y, tmap_y2, tmem_base, tfull0, tempty0_leader, sbuf0, bias_smem, quad, slot, lane, total_tiles, pair, npairs, rank);
    else epilogue_fast<ACT, false, false, true>(p, tmap_y, tmap_y2, tmem_base, tfull0, tempty0_leader, sbuf0, bias_smem, quad, slot, lane, total_tiles, pair, npairs, rank);
  }
}

// Halo mode with resident weights: the nine taps of one 64- / 32-channel chunk for a tile (pair), fully unrolled.
// ncu on 3x3 32->64 @304 (the MMA warp never waits, everything else does): the rolled loop spent ~85 instructions per tap —
// tap / 3, the patch shift, three descriptors rebuilt from addresses in vector registers and moved over with R2UR, a
// try_wait on a barrier that completed with the CTA's first tile — i.e. ~900 instructions of ONE warp per tile pair, which
// is what bounded the layer.  Here the descriptors are base + compile-time constant (a descriptor is linear in the
// address: start field = addr >> 4, no carry for shared-memory addresses), and the slab barriers are polled for the first
// tile pair only.
template <int HSTEPS, bool PAIR = false>  // K16 steps per chunk: 4 (64 channels, 128-byte rows) or 2 (32 channels, 64-byte rows);
                                          // PAIR: tcgen05 cta_group::2 (conv_fprop_pair_kernel)
__device__ __forceinline__ void halo_taps_resident(bool leader, bool two, bool first, bool acc_in, uint32_t d0, uint32_t d1,
                                                   uint32_t pa0, uint32_t pa1, uint32_t slab0, uint32_t slab_step,
                                                   uint32_t full0, uint32_t full_step, uint32_t idesc) {
  constexpr uint32_t hrow = 32u * HSTEPS;
  const uint64_t abase0 = ptx::make_smem_desc_shifted(pa0, kHaloPitch * hrow, hrow);
  const uint64_t abase1 = ptx::make_smem_desc_shifted(pa1, kHaloPitch * hrow, hrow);
  const uint64_t bbase = ptx::make_smem_desc(slab0, hrow);
  const uint64_t bstep = (uint64_t)(slab_step >> 4);
#pragma unroll
  for (int tap = 0; tap < 9; ++tap) {
    if (first) {  // the slab of (tap, chunk) arrives with this CTA's first tiles; its barrier is never re-armed
      ptx::mbar_wait(full0 + (uint32_t)tap * full_step, 0);
      ptx::tc_fence_after();
    }
    constexpr uint32_t kp = kHaloPitch;
    const uint64_t sh = (uint64_t)((((uint32_t)(tap / 3) * kp + (uint32_t)(tap % 3)) * hrow) >> 4);
    const uint64_t b = bbase + bstep * (uint64_t)tap;
#pragma unroll
    for (int k = 0; k < HSTEPS; ++k) {
      const uint32_t acc = (tap | k) != 0 ? 1u : (acc_in ? 1u : 0u);
      if (PAIR) {
        if (leader) ptx::mma_bf16_ss_pair(d0, abase0 + sh + 2u * k, b + 2u * k, idesc, acc);
        if (leader && two) ptx::mma_bf16_ss_pair(d1, abase1 + sh + 2u * k, b + 2u * k, idesc, acc);
      } else {
        if (leader) ptx::mma_bf16_ss(d0, abase0 + sh + 2u * k, b + 2u * k, idesc, acc);
        if (leader && two) ptx::mma_bf16_ss(d1, abase1 + sh + 2u * k, b + 2u * k, idesc, acc);
      }
    }
  }
}

// One K block of the tiled / im2col path for a tile (pair): KS steps of 16 channels, unrolled, descriptors advanced by
// constants (ncu on the stride-2 32->64 layer: ~100 instructions of the MMA warp per K block with the rolled loop — three
// descriptors rebuilt from addresses, a runtime trip count, R2UR chains — for eight MMAs).
template <int KS>
__device__ __forceinline__ void mma_kblock(bool two, uint32_t d0, uint32_t d1, uint64_t a0, uint64_t a1, uint64_t b, uint32_t idesc,
                                           bool acc_in) {
#pragma unroll
  for (int k = 0; k < KS; ++k) {
    const uint32_t acc = k != 0 ? 1u : (acc_in ? 1u : 0u);
    // advancing K by 16 bf16 = 32 bytes inside the swizzle atom: +2 in the (addr >> 4) field
    ptx::mma_bf16_ss(d0, a0 + 2u * k, b + 2u * k, idesc, acc);
    if (two) ptx::mma_bf16_ss(d1, a1 + 2u * k, b + 2u * k, idesc, acc);
  }
}

// 96 registers is the ceiling for 576 threads: the register file is allocated to warps in groups of four, so the CTA
// counts as 640 threads (65536 / 640 = 102 -> 96); a build with __maxnreg__(112) (64512 registers by the per-warp
// count) fails to launch on B200 ("too many resources requested").
template <bool STATS>
__global__ void __launch_bounds__(kThreads, 1)
conv_fprop_kernel_t(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                  const __grid_constant__ CUtensorMap tmap_y, const __grid_constant__ CUtensorMap tmap_y2,
                  const ConvParams p) {
  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment: required by the 128B swizzle atom (8 rows x 128 B)
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const bool halo = p.a_mode == 2;
  const uint32_t a_bytes = halo ? 0u : kBlockM * p.block_k * 2;           // halo: the ring holds weight slabs only
  const uint32_t b_bytes = ((uint32_t)(p.block_n * p.block_k * 2) + 1023u) & ~1023u;
  // tiled / im2col path with resident weights (w_resident, one N tile): the taps*cchunks weight slabs sit behind the
  // ring (where halo mode keeps its patches), loaded once per CTA; the ring then carries A slabs only
  const bool wres_t = !halo && p.w_resident;
  // tile pairs on the tiled / im2col path (p.tile_pair): a stage holds the A slabs of TWO tiles and one weight slab
  const uint32_t a_slabs = (!halo && p.tile_pair) ? 2u : 1u;
  const uint32_t stage_bytes = a_slabs * a_bytes + (wres_t ? 0u : b_bytes);
  const uint32_t tx_bytes = a_bytes + (wres_t ? 0u : (uint32_t)(p.block_n * p.block_k * 2));
  const uint32_t patch_base = smem_base + p.stages * stage_bytes;         // 1024-aligned; halo: kHaloStages patches
  const uint32_t out_base = patch_base + (halo ? p.halo_stages * kHaloPatchStride
                                               : (wres_t ? (uint32_t)(p.taps * p.cchunks) * b_bytes : 0u));  // 1024-aligned
  const uint32_t bar_base = out_base + kEpiWarps * kStageOutBytes * p.out_bufs;  // 8-byte aligned
  // barrier map: full[s] @ +8s, empty[s] @ +8(kMaxStages+s), tmem_full[4], tmem_empty[4], tmem ptr
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kMaxStages + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * kMaxStages + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * kMaxStages + kMaxAcc + a); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * kMaxStages + 2 * kMaxAcc);
  auto pfull_bar = [&](int s) { return bar_base + 8u * (2 * kMaxStages + 2 * kMaxAcc + 1 + s); };
  auto pempty_bar = [&](int s) { return bar_base + 8u * (2 * kMaxStages + 2 * kMaxAcc + 1 + kHaloMaxStages + s); };
  const uint32_t stat_smem = bar_base + 512u;  // 2 x stat_cols floats, only when statistics are requested
  if (STATS) {
    for (int i = threadIdx.x; i < 2 * p.stat_cols * p.stat_regions; i += blockDim.x)
      asm volatile("st.shared.f32 [%0], %1;" ::"r"(stat_smem + 4u * i), "f"(0.f) : "memory");
  }
  const uint32_t bias_smem = stat_smem + (STATS ? 8u * (uint32_t)(p.stat_cols * p.stat_regions) : 0u);  // bias_cols floats (fast epilogue)
  if (!STATS && p.fast_epi) {  // (bias_cols == 0: the epilogue reads the bias from global memory)
    for (int i = threadIdx.x; i < p.bias_cols; i += blockDim.x)
      asm volatile("st.shared.f32 [%0], %1;" ::"r"(bias_smem + 4u * i), "f"(__ldg(p.bias + i)) : "memory");
  }

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    ptx::prefetch_tmap(&tmap_a);
    ptx::prefetch_tmap(&tmap_b);
    ptx::prefetch_tmap(&tmap_y);
    if (p.split_col > 0) ptx::prefetch_tmap(&tmap_y2);
    for (int s = 0; s < p.stages; ++s) {
      ptx::mbar_init(full_bar(s), 1);
      ptx::mbar_init(empty_bar(s), 1);
    }
    for (int s2 = 0; s2 < kHaloMaxStages; ++s2) {
      ptx::mbar_init(pfull_bar(s2), 1);
      ptx::mbar_init(pempty_bar(s2), 1);
    }
    for (int a = 0; a < kMaxAcc; ++a) {
      ptx::mbar_init(tfull_bar(a), 1);
      ptx::mbar_init(tempty_bar(a), kEpiWarps / p.epi_sets);  // one arrival per epilogue warp of the tile's set
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
  // Programmatic dependent launch (plans with YC_PLAN_PDL): everything above — barrier init, TMEM allocation, tensor-map
  // prefetch, the bias copy (plan-owned, constant) — ran while the previous layer's last CTAs were still finishing on
  // other SMs; from here on this kernel reads what that layer wrote.  The next layer may be scheduled from now on.
  ptx::pdl_wait();
  ptx::pdl_launch_dependents();

  const int total_tiles = p.m_tiles * p.n_tiles;
  const int k_blocks = p.taps * p.cchunks;

  if (warp == 0 && halo) {
    // ===================== TMA producer, halo mode =====================
    // (whole warp converged, one elected lane issues the TMA instructions: coordinates stay in uniform registers)
    const bool leader = ptx::elect_one();
    // Tiles are processed in PAIRS (two consecutive tiles of this CTA, same N tile): per 64-channel chunk the two patches
    // {64 ch, kHaloPitch, kHaloH+2, 1} are fetched (box starts one pixel up-left of the tile; coordinates outside the
    // image are zero-filled = the convolution's padding), then the nine weight slabs ONCE for both tiles — or never
    // again when the weights are resident.
    int stage = 0, ps = 0;
    uint32_t phase = 0, pphase = 0;
    const int tpi = p.tiles_h * p.tiles_w;
    const int G = p.halo_pair ? 2 : 1;
    bool first = true;
    for (int tb = blockIdx.x; tb < total_tiles; tb += G * gridDim.x) {
      const int nt = (G == 2 && tb + (int)gridDim.x < total_tiles) ? 2 : 1;
      int mtb_, ntb_;
      tile_mn(p, tb, mtb_, ntb_);
      const int n0 = ntb_ * p.block_n;  // halo_pair requires n_tiles == 1, so both tiles share it
      for (int cc = 0; cc < p.cchunks; ++cc) {
        for (int j = 0; j < nt; ++j) {
          int mt, ntj_;
          tile_mn(p, tb + j * (int)gridDim.x, mt, ntj_);
          const int img = fastdiv(mt, p.tpi_magic, p.tpi_shift);
          const int rem = mt - img * tpi;
          const int th = fastdiv(rem, p.tw_magic, p.tw_shift);
          const int h0 = th * kHaloH, w0 = (rem - th * p.tiles_w) * kHaloW;
          ptx::mbar_wait(pempty_bar(ps), pphase ^ 1u);
          if (leader) {
            ptx::mbar_arrive_expect_tx(pfull_bar(ps), (uint32_t)(kHaloPatchRows * p.block_k * 2));
            ptx::tma_load_4d(patch_base + ps * kHaloPatchStride, &tmap_a, pfull_bar(ps), cc * p.block_k, w0 - 1, h0 - 1, img);
          }
          if (++ps == p.halo_stages) { ps = 0; pphase ^= 1u; }
        }
        if (p.w_resident) {
          // weights stationary: slab (tap, cc) lives in ring slot tap*cchunks + cc, loaded with this CTA's first tiles
          if (first)
            for (int tap = 0; tap < 9; ++tap) {
              const int slot = tap * p.cchunks + cc;
              if (leader) {
                ptx::mbar_arrive_expect_tx(full_bar(slot), tx_bytes);
                ptx::tma_load_2d(smem_base + slot * stage_bytes, &tmap_b, full_bar(slot), slot * p.block_k, n0);
              }
            }
          continue;
        }
        for (int tap = 0; tap < 9; ++tap) {
          ptx::mbar_wait(empty_bar(stage), phase ^ 1u);
          if (leader) {
            ptx::mbar_arrive_expect_tx(full_bar(stage), tx_bytes);
            ptx::tma_load_2d(smem_base + stage * stage_bytes, &tmap_b, full_bar(stage), (tap * p.cchunks + cc) * p.block_k, n0);
          }
          if (++stage == p.stages) { stage = 0; phase ^= 1u; }
        }
      }
      first = false;
    }
  } else if (warp == 1 && halo) {
    // ===================== MMA issuer, halo mode =====================
    // The WHOLE warp runs this loop converged and one elected lane issues: addresses and descriptors then stay
    // warp-uniform, i.e. in uniform registers where UTCHMMA wants them (a single-lane branch computes them in vector
    // registers and pays an R2UR chain per instruction — the per-MMA issue cost that bounded narrow tiles).
    const bool leader = ptx::elect_one();
    // The two tiles of a pair accumulate into two different TMEM buffers and their MMAs alternate: back-to-back MMAs
    // into ONE accumulator cost ~170 cycles each for tiles this narrow whatever their width (measured), independent
    // accumulators pipeline (64->32 @152: 246 -> 178 us).  Both tiles read the same weight slab.
    const uint32_t idesc = ptx::make_idesc_bf16(kBlockM, p.block_n);
    const uint32_t hrow = (uint32_t)p.block_k * 2u;  // bytes per pixel row of a patch: 128 (64-channel chunks) or 64 (32)
    const int hsteps = p.block_k / 16;
    int stage = 0, ps = 0;
    uint32_t phase = 0, pphase = 0;
    int it = 0;
    const int G = p.halo_pair ? 2 : 1;
    const int nacc_shift = 31 - __clz(p.nacc);
    bool first_pair = true;  // resident weight slabs arrive with the first tiles (halo_taps_resident)
    for (int tb = blockIdx.x; tb < total_tiles; tb += G * gridDim.x) {
      const bool two = G == 2 && tb + (int)gridDim.x < total_tiles;
      const int acc0 = it & (p.nacc - 1), acc1 = (it + 1) & (p.nacc - 1);
      ptx::mbar_wait(tempty_bar(acc0), ((it >> nacc_shift) & 1) ^ 1u);
      if (two) ptx::mbar_wait(tempty_bar(acc1), (((it + 1) >> nacc_shift) & 1) ^ 1u);
      ptx::tc_fence_after();
      const uint32_t d0 = tmem_base + acc0 * p.block_n, d1 = tmem_base + acc1 * p.block_n;
      for (int cc = 0; cc < p.cchunks; ++cc) {
        const int ps0 = ps;
        ptx::mbar_wait(pfull_bar(ps), pphase);
        if (++ps == p.halo_stages) { ps = 0; pphase ^= 1u; }
        const int ps1 = ps;
        if (two) {
          ptx::mbar_wait(pfull_bar(ps), pphase);
          if (++ps == p.halo_stages) { ps = 0; pphase ^= 1u; }
        }
        ptx::tc_fence_after();
        const uint32_t pa0 = patch_base + ps0 * kHaloPatchStride, pa1 = patch_base + ps1 * kHaloPatchStride;
        if (p.w_resident && p.issue_v2) {  // slab (tap, cc) lives in ring slot tap*cchunks + cc
          const uint32_t slab0 = smem_base + cc * stage_bytes, slab_step = p.cchunks * stage_bytes;
          if (hsteps == 4)
            halo_taps_resident<4>(leader, two, first_pair, cc != 0, d0, d1, pa0, pa1, slab0, slab_step, full_bar(cc), 8u * p.cchunks, idesc);
          else
            halo_taps_resident<2>(leader, two, first_pair, cc != 0, d0, d1, pa0, pa1, slab0, slab_step, full_bar(cc), 8u * p.cchunks, idesc);
        } else
        for (int tap = 0; tap < 9; ++tap) {
          const int kh = tap / 3, kw = tap - kh * 3;
          if (p.w_resident) {  // slot of this slab; its barrier completed phase 0 once and is never re-armed
            stage = tap * p.cchunks + cc;
            phase = 0;
          }
          ptx::mbar_wait(full_bar(stage), phase);
          ptx::tc_fence_after();
          const uint32_t sh = (uint32_t)(kh * kHaloPitch + kw) * hrow;
          const uint64_t a0 = ptx::make_smem_desc_shifted(pa0 + sh, kHaloPitch * hrow, hrow);
          const uint64_t a1 = ptx::make_smem_desc_shifted(pa1 + sh, kHaloPitch * hrow, hrow);
          const uint64_t bdesc = ptx::make_smem_desc(smem_base + stage * stage_bytes, hrow);
          if (hsteps == 4) {  // constant trip counts: the unrolled form keeps the descriptors in uniform registers
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              if (leader) ptx::mma_bf16_ss(d0, a0 + 2u * k, bdesc + 2u * k, idesc, (cc | tap | k) != 0);
              if (leader && two) ptx::mma_bf16_ss(d1, a1 + 2u * k, bdesc + 2u * k, idesc, (cc | tap | k) != 0);
            }
          } else {
#pragma unroll
            for (int k = 0; k < 2; ++k) {
              if (leader) ptx::mma_bf16_ss(d0, a0 + 2u * k, bdesc + 2u * k, idesc, (cc | tap | k) != 0);
              if (leader && two) ptx::mma_bf16_ss(d1, a1 + 2u * k, bdesc + 2u * k, idesc, (cc | tap | k) != 0);
            }
          }
          if (!p.w_resident) {
            if (leader) ptx::mma_commit(empty_bar(stage));
            if (++stage == p.stages) { stage = 0; phase ^= 1u; }
          }
        }
        if (leader) ptx::mma_commit(pempty_bar(ps0));  // the patches are free once the nine taps' MMAs have retired
        if (leader && two) ptx::mma_commit(pempty_bar(ps1));
      }
      first_pair = false;
      if (leader) ptx::mma_commit(tfull_bar(acc0));
      if (leader && two) ptx::mma_commit(tfull_bar(acc1));
      __syncwarp();
      it += two ? 2 : 1;
    }
  } else if (warp == 0) {
    // ===================== TMA producer =====================
    const bool leader = ptx::elect_one();  // whole warp converged, one lane issues
    int stage = 0;
    uint32_t phase = 0;
    const int hw = p.Ho * p.Wo;
    if (wres_t && leader) {  // all weight slabs once (n_tiles == 1): one barrier, never re-armed
      ptx::mbar_arrive_expect_tx(pfull_bar(0), (uint32_t)k_blocks * (uint32_t)(p.block_n * p.block_k * 2));
      for (int kb = 0; kb < k_blocks; ++kb)
        ptx::tma_load_2d(patch_base + kb * b_bytes, &tmap_b, pfull_bar(0), kb * p.block_k, 0);
    }
    __syncwarp();
    // Tile pairs (p.tile_pair, one N tile): two consecutive tiles of this CTA go through the ring together — per K block
    // one stage with both A slabs and ONE weight slab — so that the MMA warp can alternate between two accumulators
    // (see the MMA branch).
    const int G = p.tile_pair ? 2 : 1;
    for (int t = blockIdx.x; t < total_tiles; t += G * gridDim.x) {
      const int nt2 = (G == 2 && t + (int)gridDim.x < total_tiles) ? 2 : 1;
      int mt_, nt_;
      tile_mn(p, t, mt_, nt_);
      const int n0 = nt_ * p.block_n;
      int m0[2], img[2] = {0, 0}, base_w[2] = {0, 0}, base_h[2] = {0, 0};
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        int mtj = mt_, ntj;
        if (j == 1) tile_mn(p, t + (int)gridDim.x, mtj, ntj);
        m0[j] = mtj * kBlockM;
        if (p.a_mode == 1) {
          img[j] = m0[j] / hw;
          const int rem = m0[j] - img[j] * hw;
          const int ho = rem / p.Wo;
          base_h[j] = ho * p.stride - p.pad;
          base_w[j] = (rem - ho * p.Wo) * p.stride_w - p.pad_w;
        }
      }
      int kb = 0;
      int r = 0, s = 0;  // filter row / column of the tap (no division per tap: a lone warp feeds the TMA engine)
      for (int tap = 0; tap < p.taps; ++tap, r += (s + 1 == p.kw) ? 1 : 0, s = (s + 1 == p.kw) ? 0 : s + 1) {
        for (int cc = 0; cc < p.cchunks; ++cc, ++kb) {
          ptx::mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t sa = smem_base + stage * stage_bytes;
          const uint32_t sb = sa + a_slabs * a_bytes;
          if (leader) {
            ptx::mbar_arrive_expect_tx(full_bar(stage), tx_bytes + (nt2 == 2 ? a_bytes : 0u));
            for (int j = 0; j < nt2; ++j) {
              if (p.a_mode == 1)
                ptx::tma_load_im2col_4d(sa + j * a_bytes, &tmap_a, full_bar(stage), cc * p.block_k, base_w[j], base_h[j], img[j],
                                        (uint16_t)(s * p.dil), (uint16_t)(r * p.dil));
              else
                ptx::tma_load_2d(sa + j * a_bytes, &tmap_a, full_bar(stage), cc * p.block_k, m0[j]);
            }
            if (!wres_t) ptx::tma_load_2d(sb, &tmap_b, full_bar(stage), kb * p.block_k, n0);
          }
          if (++stage == p.stages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // whole warp converged, one elected lane issues (see the halo branch): descriptors stay in uniform registers
    const bool leader = ptx::elect_one();
    const uint32_t idesc = ptx::make_idesc_bf16(kBlockM, p.block_n);
    const uint32_t row_bytes = p.block_k * 2;
    const int nacc_shift = 31 - __clz(p.nacc);
    const int k_steps = p.block_k / 16;
    const uint64_t adesc_stage0 = ptx::make_smem_desc(smem_base, row_bytes);
    const uint64_t desc_stage_step = (uint64_t)(stage_bytes >> 4), desc_a1_off = (uint64_t)(a_bytes >> 4);
    const uint64_t desc_b_off = (uint64_t)((a_slabs * a_bytes) >> 4);
    const uint64_t bdesc_res0 = ptx::make_smem_desc(patch_base, row_bytes), desc_bres_step = (uint64_t)(b_bytes >> 4);
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    if (wres_t) {
      ptx::mbar_wait(pfull_bar(0), 0);
      ptx::tc_fence_after();
    }
    // Back-to-back MMAs into ONE accumulator are serialised by the accumulator dependency: measured ~175 cycles per MMA
    // whatever its width up to N = 128 (3x3 s2 32->64 and 64->128: tensor pipe 19 % / 36 % busy with the whole tile time
    // = MMAs x 175 cycles).  With tile pairs the MMAs of two tiles alternate between two accumulators and pipeline.
    const int G = p.tile_pair ? 2 : 1;
    for (int t = blockIdx.x; t < total_tiles; t += G * gridDim.x) {
      const bool two = G == 2 && t + (int)gridDim.x < total_tiles;
      const int acc0 = it & (p.nacc - 1), acc1 = (it + 1) & (p.nacc - 1);
      ptx::mbar_wait(tempty_bar(acc0), ((it >> nacc_shift) & 1) ^ 1u);
      if (two) ptx::mbar_wait(tempty_bar(acc1), (((it + 1) >> nacc_shift) & 1) ^ 1u);
      ptx::tc_fence_after();
      const uint32_t d0 = tmem_base + acc0 * p.block_n, d1 = tmem_base + acc1 * p.block_n;
      for (int kb = 0; kb < k_blocks; ++kb) {
        ptx::mbar_wait(full_bar(stage), phase);
        ptx::tc_fence_after();
        // a descriptor is linear in the address (start field = addr >> 4, no carry for shared-memory addresses)
        const uint64_t adesc0 = adesc_stage0 + desc_stage_step * (uint64_t)stage;
        const uint64_t adesc1 = adesc0 + desc_a1_off;
        const uint64_t bdesc = wres_t ? bdesc_res0 + desc_bres_step * (uint64_t)kb : adesc0 + desc_b_off;
        if (leader) {
          if (!p.issue_v2) {  // (A/B) the rolled loop
            for (int k = 0; k < k_steps; ++k) {
              ptx::mma_bf16_ss(d0, adesc0 + 2u * k, bdesc + 2u * k, idesc, (kb | k) != 0);
              if (two) ptx::mma_bf16_ss(d1, adesc1 + 2u * k, bdesc + 2u * k, idesc, (kb | k) != 0);
            }
          } else if (k_steps == 4) mma_kblock<4>(two, d0, d1, adesc0, adesc1, bdesc, idesc, kb != 0);
          else if (k_steps == 2) mma_kblock<2>(two, d0, d1, adesc0, adesc1, bdesc, idesc, kb != 0);
          else mma_kblock<1>(two, d0, d1, adesc0, adesc1, bdesc, idesc, kb != 0);
          ptx::mma_commit(empty_bar(stage));  // frees the smem slot when these MMAs retire
          if (kb == k_blocks - 1) {
            ptx::mma_commit(tfull_bar(acc0));
            if (two) ptx::mma_commit(tfull_bar(acc1));
          }
        }
        __syncwarp();
        if (++stage == p.stages) { stage = 0; phase ^= 1u; }
      }
      it += two ? 2 : 1;
    }
  } else if (warp >= 2) {
    // ===================== epilogue =====================
    const int quad = warp & 3;  // TMEM lane quadrant this warp may access
    const uint32_t stage_base = out_base + (uint32_t)(warp - 2) * kStageOutBytes * p.out_bufs;
    const int half = (warp - 2) >> 2;  // slot 0..3 within the quadrant
    if (STATS) {
      epilogue_role<YC_LINEAR, true>(p, &tmap_y, &tmap_y2, tmem_base, tfull_bar(0), tempty_bar(0), stage_base, quad, half, lane, total_tiles, stat_smem);
    } else if (p.fast_epi) {
      if (p.act == YC_MISH && p.mish_nr) epilogue_fast_dispatch<kMishNR>(p, &tmap_y, &tmap_y2, tmem_base, tfull_bar(0), tempty_bar(0), stage_base, bias_smem, quad, half, lane, total_tiles);
      else if (p.act == YC_MISH) epilogue_fast_dispatch<YC_MISH>(p, &tmap_y, &tmap_y2, tmem_base, tfull_bar(0), tempty_bar(0), stage_base, bias_smem, quad, half, lane, total_tiles);
      else if (p.act == YC_LEAKY) epilogue_fast_dispatch<YC_LEAKY>(p, &tmap_y, &tmap_y2, tmem_base, tfull_bar(0), tempty_bar(0), stage_base, bias_smem, quad, half, lane, total_tiles);
      else epilogue_fast_dispatch<YC_LINEAR>(p, &tmap_y, &tmap_y2, tmem_base, tfull_bar(0), tempty_bar(0), stage_base, bias_smem, quad, half, lane, total_tiles);
    } else
    switch (p.act) {
      case YC_LEAKY: epilogue_role<YC_LEAKY>(p, &tmap_y, &tmap_y2, tmem_base, tfull_bar(0), tempty_bar(0), stage_base, quad, half, lane, total_tiles, stat_smem); break;
      case YC_MISH: epilogue_role<YC_MISH>(p, &tmap_y, &tmap_y2, tmem_base, tfull_bar(0), tempty_bar(0), stage_base, quad, half, lane, total_tiles, stat_smem); break;
      case YC_LINEAR: epilogue_role<YC_LINEAR>(p, &tmap_y, &tmap_y2, tmem_base, tfull_bar(0), tempty_bar(0), stage_base, quad, half, lane, total_tiles, stat_smem); break;
      case YC_LOGISTIC: epilogue_role<YC_LOGISTIC>(p, &tmap_y, &tmap_y2, tmem_base, tfull_bar(0), tempty_bar(0), stage_base, quad, half, lane, total_tiles, stat_smem); break;
      case YC_SWISH: epilogue_role<YC_SWISH>(p, &tmap_y, &tmap_y2, tmem_base, tfull_bar(0), tempty_bar(0), stage_base, quad, half, lane, total_tiles, stat_smem); break;
      default: epilogue_role<YC_RELU>(p, &tmap_y, &tmap_y2, tmem_base, tfull_bar(0), tempty_bar(0), stage_base, quad, half, lane, total_tiles, stat_smem); break;
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (STATS) {
    // CTA partial = regions added in index order
    float* mine = p.stat_partial + (size_t)blockIdx.x * 2 * p.stat_cols;
    for (int i = threadIdx.x; i < 2 * p.stat_cols; i += blockDim.x) {
      float a = 0.f;
      for (int r = 0; r < p.stat_regions; ++r) {
        float v;
        asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(stat_smem + 4u * (uint32_t)(r * 2 * p.stat_cols + i)));
        a += v;
      }
      mine[i] = a;
    }
    __threadfence();
    __syncthreads();
    const uint32_t flag_slot = bar_base + 504u;  // last word pair of the 512-byte barrier block (unused by the barrier map)
    if (threadIdx.x == 0) {
      const unsigned last = atomicAdd(p.stat_ticket, 1u) == gridDim.x - 1 ? 1u : 0u;
      asm volatile("st.shared.u32 [%0], %1;" ::"r"(flag_slot), "r"(last) : "memory");
    }
    __syncthreads();
    unsigned last_flag;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(last_flag) : "r"(flag_slot));
    if (last_flag) {  // every partial is visible: add them in CTA order
      __threadfence();
      for (int i = threadIdx.x; i < p.N; i += blockDim.x) {
        float a = 0.f, q = 0.f;
        for (unsigned b = 0; b < gridDim.x; ++b) {
          a += __ldcg(p.stat_partial + (size_t)b * 2 * p.stat_cols + i);
          q += __ldcg(p.stat_partial + (size_t)b * 2 * p.stat_cols + p.stat_cols + i);
        }
        p.stat_sum[i] = a;
        p.stat_sq[i] = q;
        if (p.bn_mean) {  // forward_batchnorm_layer (batchnorm_layer.cpp:186-205): mean, unbiased variance, rolling averages
          const float m = a / p.bn_count;
          float v = (q - p.bn_count * m * m) / (p.bn_count - 1.f);
          v = fmaxf(v, 0.f);
          p.bn_mean[i] = m;
          p.bn_var[i] = v;
          if (p.bn_rmean) p.bn_rmean[i] = .9f * p.bn_rmean[i] + .1f * m;
          if (p.bn_rvar) p.bn_rvar[i] = .9f * p.bn_rvar[i] + .1f * v;
        }
      }
      if (threadIdx.x == 0) *p.stat_ticket = 0u;  // ready for the next launch
    }
  }
  if (warp == 1) {
    __syncwarp();
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

// ---- CTA pairs: the same GEMM with tcgen05 cta_group::2 ---------------------------------------------------------------
// Why: with one CTA per MMA a K block of a 128 x 256 tile moves 48 KB INTO shared memory (TMA) and the four MMAs read 48 KB
// OUT of it in 512 tensor-pipe cycles: 192 B/clk against the 128 B/clk a B200 SM's shared memory delivers, i.e. the wide
// layers cannot exceed ~2/3 of the tensor peak whatever the pipeline does (measured: 1.0-1.17 PFLOP/s at 1.74 GHz).  A pair
// of CTAs (cluster of two = one TPC) shares every MMA of M = 256: each CTA stages its own 128 rows of A and only HALF of the
// weight rows, 32 KB in + 32 KB out per K block = exactly 128 B/clk, and the L2 -> SM traffic for weights halves as well.
//   * both CTAs run the TMA producer for their own slabs; all bytes are counted on the LEADER's (rank 0) "full" barrier
//     (cp.async.bulk.tensor ... .cta_group::2 with the leader's barrier address);
//   * the leader's MMA warp issues tcgen05.mma.cta_group::2 (A / B descriptors name the same offsets in both CTAs; each
//     CTA's TMEM receives its own 128 rows) and commits with .multicast::cluster to the "empty" / "accumulator full"
//     barriers of both CTAs;
//   * both CTAs run the specialised epilogue on their 128 rows; every epilogue warp hands the accumulator back on the
//     leader's barrier (remote mbarrier.arrive for the peer).
// Tiled (1x1) and im2col operands, specialised epilogue only (bf16 output, folded batchnorm); everything else stays on
// conv_fprop_kernel_t.
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
conv_fprop_pair_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                       const __grid_constant__ CUtensorMap tmap_y, const __grid_constant__ CUtensorMap tmap_y2,
                       const ConvParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const bool halo = p.a_mode == 2;
  const uint32_t a_bytes = halo ? 0u : kBlockM * p.block_k * 2;           // halo: the ring holds weight half-slabs only
  const uint32_t half_n = (uint32_t)p.block_n >> 1;                       // weight rows this CTA stages
  const uint32_t bh_tx = half_n * (uint32_t)p.block_k * 2u;
  const uint32_t bh_bytes = (bh_tx + 1023u) & ~1023u;
  const uint32_t stage_bytes = a_bytes + bh_bytes;
  const uint32_t patch_base = smem_base + p.stages * stage_bytes;         // 1024-aligned; halo: halo_stages patches
  const uint32_t out_base = patch_base + (halo ? p.halo_stages * kHaloPatchStride : 0u);
  const uint32_t bar_base = out_base + kEpiWarps * kStageOutBytes * p.out_bufs;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kMaxStages + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * kMaxStages + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * kMaxStages + kMaxAcc + a); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * kMaxStages + 2 * kMaxAcc);
  auto pfull_bar = [&](int s) { return bar_base + 8u * (2 * kMaxStages + 2 * kMaxAcc + 1 + s); };
  auto pempty_bar = [&](int s) { return bar_base + 8u * (2 * kMaxStages + 2 * kMaxAcc + 1 + kHaloMaxStages + s); };
  const uint32_t bias_smem = bar_base + 512u;
  for (int i = threadIdx.x; i < p.bias_cols; i += blockDim.x)
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(bias_smem + 4u * i), "f"(__ldg(p.bias + i)) : "memory");

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const int pair = (int)(blockIdx.x >> 1), npairs = (int)(gridDim.x >> 1);

  if (threadIdx.x == 0) {
    ptx::prefetch_tmap(&tmap_a);
    ptx::prefetch_tmap(&tmap_b);
    ptx::prefetch_tmap(&tmap_y);
    if (p.split_col > 0) ptx::prefetch_tmap(&tmap_y2);
    for (int s = 0; s < p.stages; ++s) {
      ptx::mbar_init(full_bar(s), 1);    // the leader's: one arrive.expect_tx for the bytes of both CTAs
      ptx::mbar_init(empty_bar(s), 1);   // each CTA's own, arrived by the leader's multicast commit
    }
    for (int s2 = 0; s2 < kHaloMaxStages; ++s2) {
      ptx::mbar_init(pfull_bar(s2), 1);
      ptx::mbar_init(pempty_bar(s2), 1);
    }
    for (int a = 0; a < kMaxAcc; ++a) {
      ptx::mbar_init(tfull_bar(a), 1);
      ptx::mbar_init(tempty_bar(a), 2 * kEpiWarps / p.epi_sets);  // the leader's: the tile's set of warps of BOTH CTAs
    }
    ptx::fence_barrier_init();
  }
  if (warp == 1) {  // the same warp of both CTAs
    ptx::tmem_alloc_pair(tmem_slot, p.tmem_cols);
    ptx::tmem_relinquish_pair();
  }
  ptx::tc_fence_before();
  ptx::cluster_sync();  // barriers of BOTH CTAs are initialised before any remote arrival / TMA completion
  ptx::tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
  ptx::pdl_wait();               // (see conv_fprop_kernel_t)
  ptx::pdl_launch_dependents();

  const int total_tiles = p.m_tiles * p.n_tiles;  // 256-row tiles x N tiles
  const int k_blocks = p.taps * p.cchunks;

  if (warp == 0 && halo) {
    // ===================== TMA producer, halo mode (both CTAs): own patches, own half of the weight rows =====================
    // As in conv_fprop_kernel_t's halo branch, tiles go in PAIRS (halo_pair): two 256-pixel tiles (one 16 x 8 block per
    // CTA each) accumulate into two TMEM buffers and share every weight slab, so that the leader's MMAs alternate between
    // two independent accumulators.
    const bool leader = ptx::elect_one();
    const uint32_t full0_leader = ptx::mapa(full_bar(0), 0), pfull0_leader = ptx::mapa(pfull_bar(0), 0);
    int stage = 0, ps = 0;
    uint32_t phase = 0, pphase = 0;
    const int tpi = p.tiles_h * p.tiles_w;
    const int G = p.halo_pair ? 2 : 1;
    bool first = true;
    for (int tb = pair; tb < total_tiles; tb += G * npairs) {
      const int nt = (G == 2 && tb + npairs < total_tiles) ? 2 : 1;
      int mtb_, ntb_;
      tile_mn(p, tb, mtb_, ntb_);
      const int n0 = ntb_ * p.block_n + (int)(rank * half_n);
      for (int cc = 0; cc < p.cchunks; ++cc) {
        for (int j = 0; j < nt; ++j) {
          int mt, ntj_;
          tile_mn(p, tb + j * npairs, mt, ntj_);
          mt = 2 * mt + (int)rank;  // (== m_tiles128 for the idle half of an odd last pair: image index past the batch, zero fill)
          const int img = fastdiv(mt, p.tpi_magic, p.tpi_shift);
          const int rem = mt - img * tpi;
          const int th = fastdiv(rem, p.tw_magic, p.tw_shift);
          const int h0 = th * kHaloH, w0 = (rem - th * p.tiles_w) * kHaloW;
          ptx::mbar_wait(pempty_bar(ps), pphase ^ 1u);
          if (leader) {
            if (rank == 0) ptx::mbar_arrive_expect_tx(pfull_bar(ps), 2u * (uint32_t)(kHaloPatchRows * p.block_k * 2));
            ptx::tma_load_4d_pair(patch_base + ps * kHaloPatchStride, &tmap_a, pfull0_leader + 8u * ps, cc * p.block_k, w0 - 1,
                                  h0 - 1, img);
          }
          if (++ps == p.halo_stages) { ps = 0; pphase ^= 1u; }
        }
        if (p.w_resident) {
          if (first)
            for (int tap = 0; tap < 9; ++tap) {
              const int slot = tap * p.cchunks + cc;
              if (leader) {
                if (rank == 0) ptx::mbar_arrive_expect_tx(full_bar(slot), 2u * bh_tx);
                ptx::tma_load_2d_pair(smem_base + slot * stage_bytes, &tmap_b, full0_leader + 8u * slot, slot * p.block_k, n0);
              }
            }
          continue;
        }
        for (int tap = 0; tap < 9; ++tap) {
          ptx::mbar_wait(empty_bar(stage), phase ^ 1u);
          if (leader) {
            if (rank == 0) ptx::mbar_arrive_expect_tx(full_bar(stage), 2u * bh_tx);
            ptx::tma_load_2d_pair(smem_base + stage * stage_bytes, &tmap_b, full0_leader + 8u * stage,
                                  (tap * p.cchunks + cc) * p.block_k, n0);
          }
          if (++stage == p.stages) { stage = 0; phase ^= 1u; }
        }
      }
      first = false;
    }
  } else if (warp == 1 && halo) {
    if (rank == 0) {
      // ===================== MMA issuer, halo mode (leader CTA only) =====================
      const bool leader = ptx::elect_one();
      const uint32_t idesc = ptx::make_idesc_bf16(2 * kBlockM, p.block_n);
      const uint32_t hrow = (uint32_t)p.block_k * 2u;
      const int hsteps = p.block_k / 16;
      int stage = 0, ps = 0;
      uint32_t phase = 0, pphase = 0;
      int it = 0;
      const int G = p.halo_pair ? 2 : 1;
      const int nacc_shift = 31 - __clz(p.nacc);
      bool first_pair = true;
      for (int tb = pair; tb < total_tiles; tb += G * npairs) {
        const bool two = G == 2 && tb + npairs < total_tiles;
        const int acc0 = it & (p.nacc - 1), acc1 = (it + 1) & (p.nacc - 1);
        ptx::mbar_wait(tempty_bar(acc0), ((it >> nacc_shift) & 1) ^ 1u);
        if (two) ptx::mbar_wait(tempty_bar(acc1), (((it + 1) >> nacc_shift) & 1) ^ 1u);
        ptx::tc_fence_after();
        const uint32_t d0 = tmem_base + acc0 * p.block_n, d1 = tmem_base + acc1 * p.block_n;
        for (int cc = 0; cc < p.cchunks; ++cc) {
          const int ps0 = ps;
          ptx::mbar_wait(pfull_bar(ps), pphase);
          if (++ps == p.halo_stages) { ps = 0; pphase ^= 1u; }
          const int ps1 = ps;
          if (two) {
            ptx::mbar_wait(pfull_bar(ps), pphase);
            if (++ps == p.halo_stages) { ps = 0; pphase ^= 1u; }
          }
          ptx::tc_fence_after();
          const uint32_t pa0 = patch_base + ps0 * kHaloPatchStride, pa1 = patch_base + ps1 * kHaloPatchStride;
          if (p.w_resident && p.issue_v2) {  // unrolled taps, constant descriptor offsets (halo_taps_resident)
            const uint32_t slab0 = smem_base + cc * stage_bytes, slab_step = p.cchunks * stage_bytes;
            if (hsteps == 4)
              halo_taps_resident<4, true>(leader, two, first_pair, cc != 0, d0, d1, pa0, pa1, slab0, slab_step, full_bar(cc), 8u * p.cchunks, idesc);
            else
              halo_taps_resident<2, true>(leader, two, first_pair, cc != 0, d0, d1, pa0, pa1, slab0, slab_step, full_bar(cc), 8u * p.cchunks, idesc);
          } else
          for (int tap = 0; tap < 9; ++tap) {
            const int kh = tap / 3, kw = tap - kh * 3;
            if (p.w_resident) {
              stage = tap * p.cchunks + cc;
              phase = 0;
            }
            ptx::mbar_wait(full_bar(stage), phase);
            ptx::tc_fence_after();
            const uint32_t sh = (uint32_t)(kh * kHaloPitch + kw) * hrow;
            const uint64_t a0 = ptx::make_smem_desc_shifted(pa0 + sh, kHaloPitch * hrow, hrow);
            const uint64_t a1 = ptx::make_smem_desc_shifted(pa1 + sh, kHaloPitch * hrow, hrow);
            const uint64_t bdesc = ptx::make_smem_desc(smem_base + stage * stage_bytes, hrow);
            if (hsteps == 4) {
#pragma unroll
              for (int k2 = 0; k2 < 4; ++k2) {
                if (leader) ptx::mma_bf16_ss_pair(d0, a0 + 2u * k2, bdesc + 2u * k2, idesc, (cc | tap | k2) != 0);
                if (leader && two) ptx::mma_bf16_ss_pair(d1, a1 + 2u * k2, bdesc + 2u * k2, idesc, (cc | tap | k2) != 0);
              }
            } else {
#pragma unroll
              for (int k2 = 0; k2 < 2; ++k2) {
                if (leader) ptx::mma_bf16_ss_pair(d0, a0 + 2u * k2, bdesc + 2u * k2, idesc, (cc | tap | k2) != 0);
                if (leader && two) ptx::mma_bf16_ss_pair(d1, a1 + 2u * k2, bdesc + 2u * k2, idesc, (cc | tap | k2) != 0);
              }
            }
            if (!p.w_resident) {
              if (leader) ptx::mma_commit_pair(empty_bar(stage), 3);
              if (++stage == p.stages) { stage = 0; phase ^= 1u; }
            }
          }
          if (leader) ptx::mma_commit_pair(pempty_bar(ps0), 3);
          if (leader && two) ptx::mma_commit_pair(pempty_bar(ps1), 3);
        }
        first_pair = false;
        if (leader) ptx::mma_commit_pair(tfull_bar(acc0), 3);
        if (leader && two) ptx::mma_commit_pair(tfull_bar(acc1), 3);
        __syncwarp();
        it += two ? 2 : 1;
      }
    }
  } else if (warp == 0) {
    // ===================== TMA producer (both CTAs): own 128 rows of A, own half of the weight rows =====================
    const bool leader = ptx::elect_one();
    const uint32_t full0_leader = ptx::mapa(full_bar(0), 0);
    int stage = 0;
    uint32_t phase = 0;
    const int hw = p.Ho * p.Wo;
    for (int t = pair; t < total_tiles; t += npairs) {
      int mt, nt;
      tile_mn(p, t, mt, nt);
      const int m0 = (2 * mt + (int)rank) * kBlockM;
      const int n0 = nt * p.block_n + (int)(rank * half_n);
      int img = 0, base_w = 0, base_h = 0;
      if (p.a_mode == 1) {
        img = m0 / hw;
        const int rem = m0 - img * hw;
        const int ho = rem / p.Wo;
        base_h = ho * p.stride - p.pad;
        base_w = (rem - ho * p.Wo) * p.stride_w - p.pad_w;
      }
      int kb = 0;
      int r = 0, s = 0;  // filter row / column of the tap (no division per tap: a lone warp feeds the TMA engine)
      for (int tap = 0; tap < p.taps; ++tap, r += (s + 1 == p.kw) ? 1 : 0, s = (s + 1 == p.kw) ? 0 : s + 1) {
        for (int cc = 0; cc < p.cchunks; ++cc, ++kb) {
          ptx::mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t sa = smem_base + stage * stage_bytes;
          if (leader) {
            if (rank == 0) ptx::mbar_arrive_expect_tx(full_bar(stage), 2u * (a_bytes + bh_tx));
            const uint32_t fb = full0_leader + 8u * stage;
            if (p.a_mode == 1)
              ptx::tma_load_im2col_4d_pair(sa, &tmap_a, fb, cc * p.block_k, base_w, base_h, img, (uint16_t)(s * p.dil),
                                           (uint16_t)(r * p.dil));
            else
              ptx::tma_load_2d_pair(sa, &tmap_a, fb, cc * p.block_k, m0);
            ptx::tma_load_2d_pair(sa + a_bytes, &tmap_b, fb, kb * p.block_k, n0);
          }
          if (++stage == p.stages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    if (rank == 0) {
      // ===================== MMA issuer (leader CTA only), one elected lane =====================
      const bool leader = ptx::elect_one();
      const uint32_t idesc = ptx::make_idesc_bf16(2 * kBlockM, p.block_n);
      const uint32_t row_bytes = p.block_k * 2;
      const int nacc_shift = 31 - __clz(p.nacc);
      const int k_steps = p.block_k / 16;
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int t = pair; t < total_tiles; t += npairs, ++it) {
        const int acc = it & (p.nacc - 1);
        ptx::mbar_wait(tempty_bar(acc), ((it >> nacc_shift) & 1) ^ 1u);
        ptx::tc_fence_after();
        const uint32_t d = tmem_base + acc * p.block_n;
        for (int kb = 0; kb < k_blocks; ++kb) {
          ptx::mbar_wait(full_bar(stage), phase);
          ptx::tc_fence_after();
          const uint32_t sa = smem_base + stage * stage_bytes;
          const uint64_t adesc = ptx::make_smem_desc(sa, row_bytes);
          const uint64_t bdesc = ptx::make_smem_desc(sa + a_bytes, row_bytes);
          if (leader) {
            for (int k = 0; k < k_steps; ++k)
              ptx::mma_bf16_ss_pair(d, adesc + 2u * k, bdesc + 2u * k, idesc, (kb | k) != 0);
            ptx::mma_commit_pair(empty_bar(stage), 3);  // frees the slot in both CTAs when these MMAs retire
            if (kb == k_blocks - 1) ptx::mma_commit_pair(tfull_bar(acc), 3);
          }
          __syncwarp();
          if (++stage == p.stages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else {
    // ===================== epilogue (both CTAs, own 128 rows) =====================
    const int quad = warp & 3;
    const uint32_t stage_base = out_base + (uint32_t)(warp - 2) * kStageOutBytes * p.out_bufs;
    const int slot = (warp - 2) >> 2;
    const uint32_t tempty0_leader = ptx::mapa(tempty_bar(0), 0);
    const int r = (int)rank;
    if (p.act == YC_MISH) epilogue_pair_dispatch<YC_MISH>(p, &tmap_y, &tmap_y2, tmem_base, tfull_bar(0), tempty0_leader, stage_base, bias_smem, quad, slot, lane, total_tiles, pair, npairs, r);
    else if (p.act == YC_LEAKY) epilogue_pair_dispatch<YC_LEAKY>(p, &tmap_y, &tmap_y2, tmem_base, tfull_bar(0), tempty0_leader, stage_base, bias_smem, quad, slot, lane, total_tiles, pair, npairs, r);
    else epilogue_pair_dispatch<YC_LINEAR>(p, &tmap_y, &tmap_y2, tmem_base, tfull_bar(0), tempty0_leader, stage_base, bias_smem, quad, slot, lane, total_tiles, pair, npairs, r);
  }

  ptx::tc_fence_before();
  ptx::cluster_sync();  // neither CTA's shared memory / TMEM goes away while the other may still use it
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc_pair(tmem_base, p.tmem_cols);
  }
}

// ---- weight packing: OIHW fp32 -> [out_c][tap][in_c_p] bf16 ------------------------------------
__global__ void pack_weights_kernel(const float* __restrict__ w, int out_c, int in_c, int taps, int in_c_p,
                                    __nv_bfloat16* __restrict__ out) {
  const long long total = (long long)out_c * taps * in_c_p;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % in_c_p);
    const int tap = (int)((i / in_c_p) % taps);
    const int n = (int)(i / ((long long)in_c_p * taps));
    float v = 0.f;
    if (c < in_c) v = w[((long long)n * in_c + c) * taps + tap];
    out[i] = __float2bfloat16_rn(v);
  }
}

// ---- dgrad weight packing: OIHW fp32 -> [in_c][flipped tap][out_c_p] bf16 ------------------------
// dX = correlation of dY with the spatially flipped, channel-transposed filter (the col2im(W^T * dY) of
// convolutional_layer.cpp:1360-1376 / convolutional_kernels.cu:765-802 written as a forward conv), so the
// data-gradient runs through conv_fprop_kernel itself.
__global__ void pack_weights_dgrad_kernel(const float* __restrict__ w, int out_c, int in_c, int ksize, int out_c_p,
                                          __nv_bfloat16* __restrict__ out) {
  const int taps = ksize * ksize;
  const long long total = (long long)in_c * taps * out_c_p;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(i % out_c_p);
    const int tap = (int)((i / out_c_p) % taps);
    const int c = (int)(i / ((long long)out_c_p * taps));
    float v = 0.f;
    if (n < out_c) v = w[((long long)n * in_c + c) * taps + (taps - 1 - tap)];
    out[i] = __float2bfloat16_rn(v);
  }
}

// zero-insertion for strided layers: up[b][s*i][s*j][:] = dy[b][i][j][:], zeros elsewhere
__global__ void dgrad_dilate_kernel(const __nv_bfloat16* __restrict__ dy, int B, int Ho, int Wo, int c8, long long ld_dy,
                                    int stride, int Hu, int Wu, long long ld_up, __nv_bfloat16* __restrict__ up) {
  const long long total = (long long)B * Hu * Wu * c8;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i % c8);
    const long long pix = i / c8;
    const int wu = (int)(pix % Wu);
    const int hu = (int)((pix / Wu) % Hu);
    const long long b = pix / ((long long)Wu * Hu);
    uint4 q = make_uint4(0, 0, 0, 0);
    if (hu % stride == 0 && wu % stride == 0) {
      const int ho = hu / stride, wo = wu / stride;
      if (ho < Ho && wo < Wo) q = *reinterpret_cast<const uint4*>(dy + ((b * Ho + ho) * Wo + wo) * ld_dy + g * 8);
    }
    *reinterpret_cast<uint4*>(up + pix * ld_up + g * 8) = q;
  }
}

}  // namespace yc

// ================================== host side ==================================================
using namespace yc;

struct yc_conv_plan {
  CUtensorMap tmap_a, tmap_b, tmap_y, tmap_y2;
  bool rect = false;
  ConvParams p;
  float* bias_pad = nullptr;
  float* scale_pad = nullptr;
  float* stat_partial = nullptr;   // [grid][2][stat_cols] (training statistics)
  unsigned* stat_ticket = nullptr;
  int grid = 0;
  size_t smem = 0;
  bool pdl = false;                // launch with programmatic stream serialization (YC_PLAN_PDL)
  const void* w_packed = nullptr;  // for re-encoding the weight map when a CTA-pair plan falls back (yc_conv_plan_set_stats)
  int w_kdim = 0;                  // taps * padded in_c
};

namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode_tiled = nullptr;
int g_issue_v2 = 1;  // yc_conv_fprop_set_issue_v2 (A/B): plans created from now on use the unrolled halo tap issue
EncodeIm2colFn g_encode_im2col = nullptr;

int load_driver_entry_points() {
  if (g_encode_tiled && g_encode_im2col) return YC_OK;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qr;
  YC_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr));
  YC_REQUIRE(fn && qr == cudaDriverEntryPointSuccess, "cuTensorMapEncodeTiled not available");
  g_encode_tiled = (EncodeTiledFn)fn;
  YC_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &fn, cudaEnableDefault, &qr));
  YC_REQUIRE(fn && qr == cudaDriverEntryPointSuccess, "cuTensorMapEncodeIm2col not available");
  g_encode_im2col = (EncodeIm2colFn)fn;
  return YC_OK;
}

int pick_block_k(int in_c) {
  if (in_c % 64 == 0) return 64;
  if (in_c % 32 == 0) return 32;
  return 16;
}

CUtensorMapSwizzle swizzle_for(int block_k) {
  return block_k == 64 ? CU_TENSOR_MAP_SWIZZLE_128B
                       : (block_k == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
}

}  // namespace

extern "C" int yc_conv_padded_in_c(int in_c) { return round_up(in_c, pick_block_k(in_c)); }

extern "C" size_t yc_conv_packed_bytes(int out_c, int in_c, int ksize) {
  return (size_t)out_c * ksize * ksize * yc_conv_padded_in_c(in_c) * 2;
}

extern "C" int yc_conv_pack(const float* w, int out_c, int in_c, int ksize, void* out, void* stream) {
  YC_REQUIRE(w && out && out_c > 0 && in_c > 0 && ksize > 0, "yc_conv_pack: bad arguments");
  const int in_c_p = yc_conv_padded_in_c(in_c);
  const long long total = (long long)out_c * ksize * ksize * in_c_p;
  const int blocks = (int)std::min<long long>(ceil_div_ll(total, 256), 148 * 8);
  pack_weights_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(w, out_c, in_c, ksize * ksize, in_c_p,
                                                                 (__nv_bfloat16*)out);
  YC_LAUNCH_CHECK();
  return YC_OK;
}

// kw / stride_w / pad_l / pad_r: kernel width, stride and left / right padding along W when they differ from the
// square geometry of the descriptor (used by the network for the pixel-pair view of a 3x3 stride-2 layer, see
// network.cu); the filter is packed [n][ksize*kw][c].  y2 (optional): second output tensor of a fused pair of
// convolutions reading the same input — GEMM columns [0, y->c) go to y, [y->c, y->c + y2->c) to y2.
static int plan_create_impl(yc_conv_plan** out_plan, const yc_conv_desc* d, int kw, int stride_w, int pad_l, int pad_r,
                            const yc_tensor* x, const void* w_packed, const float* bias, const float* scale,
                            const yc_tensor* residual, const yc_tensor* y, const yc_tensor* y2, unsigned flags);
extern "C" int yc_conv_plan_create(yc_conv_plan** out_plan, const yc_conv_desc* d, const yc_tensor* x,
                                   const void* w_packed, const float* bias, const float* scale,
                                   const yc_tensor* residual, const yc_tensor* y) {
  YC_REQUIRE(d, "yc_conv_plan_create: null argument");
  return plan_create_impl(out_plan, d, d->ksize, d->stride, d->pad, d->pad, x, w_packed, bias, scale, residual, y, nullptr, 0);
}
extern "C" int yc_conv_plan_create_ex(yc_conv_plan** out_plan, const yc_conv_desc* d, const yc_tensor* x,
                                      const void* w_packed, const float* bias, const float* scale,
                                      const yc_tensor* residual, const yc_tensor* y, const yc_tensor* y2,
                                      unsigned plan_flags) {
  YC_REQUIRE(d, "yc_conv_plan_create_ex: null argument");
  return plan_create_impl(out_plan, d, d->ksize, d->stride, d->pad, d->pad, x, w_packed, bias, scale, residual, y, y2,
                          plan_flags);
}
extern "C" int yc_conv_plan_create_rect(yc_conv_plan** out_plan, const yc_conv_desc* d, int kw, int stride_w, int pad_l,
                                        int pad_r, const yc_tensor* x, const void* w_packed, const float* bias,
                                        const float* scale, const yc_tensor* residual, const yc_tensor* y,
                                        unsigned plan_flags) {
  YC_REQUIRE(d && kw >= 1 && stride_w >= 1 && pad_l >= 0 && pad_r >= 0, "yc_conv_plan_create_rect: bad arguments");
  return plan_create_impl(out_plan, d, kw, stride_w, pad_l, pad_r, x, w_packed, bias, scale, residual, y, nullptr,
                          plan_flags);
}
static int plan_create_impl(yc_conv_plan** out_plan, const yc_conv_desc* d, int kw, int stride_w, int pad_l, int pad_r,
                            const yc_tensor* x, const void* w_packed, const float* bias, const float* scale,
                            const yc_tensor* residual, const yc_tensor* y, const yc_tensor* y2, unsigned flags) {
  YC_REQUIRE(out_plan && d && x && w_packed && y, "yc_conv_plan_create: null argument");  // bias == NULL: zeros
  if (y2) {
    YC_REQUIRE(!residual && y2->dtype == y->dtype && y2->n == y->n && y2->h == y->h && y2->w == y->w && y->c % 32 == 0 &&
                   y2->ldc % (y->dtype == YC_F32 ? 4 : 8) == 0 && ((uintptr_t)y2->ptr & 15) == 0,
               "fused pair: the second output must have the first one's shape and type, the first a multiple of 32 channels");
  }
  const bool rect = kw != d->ksize || stride_w != d->stride || pad_l != d->pad || pad_r != d->pad;
  YC_REQUIRE(x->dtype == YC_BF16, "conv input must be bf16 NHWC");
  YC_REQUIRE(x->ldc % 8 == 0 && ((uintptr_t)x->ptr & 15) == 0, "conv input view must be 16-byte aligned (ldc%%8)");
  YC_REQUIRE(d->ksize >= 1 && d->stride >= 1 && d->dilation >= 1 && d->pad >= 0, "bad conv geometry");
  const int eff = d->dilation * (d->ksize - 1) + 1;
  const int Ho = (x->h + 2 * d->pad - eff) / d->stride + 1;
  const int Wo = (x->w + pad_l + pad_r - (d->dilation * (kw - 1) + 1)) / stride_w + 1;
  YC_REQUIRE(y->n == x->n && y->h == Ho && y->w == Wo, "conv output shape mismatch: want %dx%d got %dx%d", Ho, Wo,
             y->h, y->w);
  const int elt = y->dtype == YC_F32 ? 4 : 8;  // store granularity in elements (16 bytes)
  YC_REQUIRE(y->ldc % elt == 0 && ((uintptr_t)y->ptr & 15) == 0, "conv output view must be 16-byte aligned");
  if (residual) {
    YC_REQUIRE(residual->dtype == YC_BF16 && residual->n == y->n && residual->h == y->h && residual->w == y->w &&
                   residual->c == y->c && residual->ldc % 8 == 0 && ((uintptr_t)residual->ptr & 15) == 0,
               "residual must be a bf16 view of the output's shape");
  }
  YC_REQUIRE((long long)x->n * Ho * Wo < (1LL << 31), "M too large");
  int rc = load_driver_entry_points();
  if (rc) return rc;

  yc_conv_plan* pl = new yc_conv_plan();
  ConvParams& p = pl->p;
  const int N = y->c + (y2 ? y2->c : 0);
  p.M = x->n * Ho * Wo;
  p.N = N;
  p.split_col = y2 ? y->c : 0;
  // columns we may write: up to the padded pitch when the tensor owns its pitch, else exactly c
  const int n_store2 = y2 ? std::min(round_up(y2->c, elt), (int)y2->ldc) : 0;
  p.n_store = y2 ? y->c + n_store2 : std::min(round_up(N, elt), (int)y->ldc);
  YC_REQUIRE(p.n_store % elt == 0, "output channel count must be a multiple of %d for this view", elt);
  p.taps = d->ksize * kw;
  p.ksize = d->ksize;
  p.kw = kw; p.stride_w = stride_w; p.pad_w = pad_l;
  p.block_k = pick_block_k(x->c);
  const int in_c_p = round_up(x->c, p.block_k);
  p.cchunks = in_c_p / p.block_k;
  p.a_mode = (!rect && d->ksize == 1 && d->stride == 1 && d->pad == 0) ? 0 : 1;
  p.Ho = Ho; p.Wo = Wo; p.stride = d->stride; p.pad = d->pad; p.dil = d->dilation;
  p.m_tiles = ceil_div(p.M, kBlockM);
  p.tiles_h = p.tiles_w = 0;
  {
    // halo mode: 3x3 / stride 1 / pad 1 with whole 64- (or 32-) channel chunks.  It pays where the nine im2col tap loads bound the
    // layer (few output channels per fetched row) and the 16x8 tiling wastes little: YC_PLAN_NO_HALO never,
    // YC_PLAN_FORCE_HALO always (when legal), neither: cost rule.
    const int mode = (flags & YC_PLAN_NO_HALO) ? 0 : ((flags & YC_PLAN_FORCE_HALO) ? 1 : -1);
    const bool legal = !y2 && !rect && d->ksize == 3 && d->stride == 1 && d->pad == 1 && d->dilation == 1 && x->c % 32 == 0;
    const int th = ceil_div(Ho, kHaloH), tw = ceil_div(Wo, kHaloW);
    const double waste = (double)x->n * th * tw * kBlockM / (double)p.M;
    const bool want = mode == 1 || (mode < 0 && N <= 128 && waste <= 1.12);
    if (legal && want && mode != 0 && (long long)x->n * th * tw < (1ll << 20)) {
      p.a_mode = 2;
      p.tiles_h = th;
      p.tiles_w = tw;
      p.m_tiles = x->n * th * tw;
      // q = (n * magic) >> shift == n / d for every n < 2^20: shift = 20 + ceil(log2 d), magic = ceil(2^shift / d)
      auto set_fastdiv = [](uint32_t dd, uint32_t& magic, uint32_t& shift) {
        uint32_t l = 0;
        while ((1u << l) < dd) ++l;
        shift = 20 + l;
        magic = (uint32_t)(((1ull << shift) + dd - 1) / dd);
      };
      set_fastdiv((uint32_t)(th * tw), p.tpi_magic, p.tpi_shift);
      set_fastdiv((uint32_t)tw, p.tw_magic, p.tw_shift);
    }
  }

  // ---- tile width: minimise (waves x per-tile cost); per-tile cost ~ block_n + fixed overhead -----
  const int sms = sm_count();
  int best_bn = 32;
  double best_cost = 1e30;
  for (int bn = 32; bn <= 256; bn += 32) {
    if (bn > round_up(N, 32)) break;
    const int nt = ceil_div(N, bn);
    const long long tiles = (long long)p.m_tiles * nt;
    const double waves = (double)ceil_div_ll(tiles, sms);
    // Per-tile cost in units of one K step.  Measured on B200 (profiles/r01_layer_times.txt): a K=16 step of a 128 x 256
    // tile takes only 1.14x the time of a 128 x 128 one (both are bound by the shared-memory operand reads of the MMA,
    // 4 KB of A either way), and a 128 x 64 one is no faster than 128 x 128.  COST_V1 is the round-1 model.
    const double per_tile = (flags & YC_PLAN_COST_V1) ? std::max(bn, 96) + 24.0 : std::max(bn, 192) + 32.0;
    const double cost = waves * per_tile;
    if (cost < best_cost - 1e-9) { best_cost = cost; best_bn = bn; }
  }
  p.block_n = best_bn;
  p.n_tiles = ceil_div(N, p.block_n);
  const uint32_t a_bytes = p.a_mode == 2 ? 0u : kBlockM * p.block_k * 2;
  const uint32_t b_bytes = (uint32_t)round_up(p.block_n * p.block_k * 2, 1024);
  const uint32_t stage_bytes = a_bytes + b_bytes;
  p.act = d->activation;
  p.act2 = d->residual_activation;
  p.y_f32 = (y->dtype == YC_F32);
  // the specialised epilogue covers what inference runs: bf16 output, batchnorm folded into the weights, mish / leaky /
  // linear, optional fused shortcut with a linear second activation (YC_PLAN_GENERIC_EPI keeps the generic one)
  p.fast_epi = !(flags & YC_PLAN_GENERIC_EPI) && !p.y_f32 && !scale && p.act2 == YC_LINEAR &&
               (p.act == YC_MISH || p.act == YC_LEAKY || p.act == YC_LINEAR);
  p.mish_nr = (flags & YC_PLAN_MISH_NEWTON) ? 1 : 0;
  uint32_t bias_bytes = 0;  // set per layout attempt below
  const int k_blocks = p.taps * p.cchunks;
  const int wres_env = (flags & YC_PLAN_NO_W_RES) ? 0 : ((flags & YC_PLAN_W_RES_1X1) ? 2 : 1);  // 0 never, 1 multi-tap layers, 2 also 1x1
  // Shared-memory layout for a given number of epilogue staging buffers per warp:
  //  * halo mode: tiles go in pairs (two accumulators, one weight slab for both) when the N tile is shared (n_tiles == 1)
  //    and four patches fit next to at least three weight slabs (YC_PLAN_NO_HALO_PAIR switches the pairing off); when
  //    all 9 * cchunks weight slabs fit they stay resident;
  //  * tiled / im2col path, one N tile, many tiles per CTA, small filter: the weight slabs stay resident behind the ring
  //    (the per-tile reload is a third of the L2 -> SM traffic of the 64-channel layers; YC_PLAN_NO_W_RES switches it off);
  //  * otherwise the ring is as deep as shared memory allows: the producer runs ahead across tiles, so the ring depth is
  //    the amount of HBM traffic in flight per SM (what the memory-bound 1x1 layers live on).
  // tile pairs on the tiled / im2col path: worth it where a tile issues enough dependent MMAs for their ~175-cycle
  // accumulator latency to show and the CTA has at least two tiles.  Measured (profiles/r02_ab_tile_pair.txt): 3x3 s2
  // 32->64 (18 MMAs per tile) 863 -> 509 us, 3x3 s2 64->128 (36) 365 -> 238 us; the 1x1 layers with 8 MMAs per tile lose
  // 5-10 % (they are bound by the activation stream, and a pair halves the number of ring stages) -> threshold 12
  p.tile_pair = (p.a_mode != 2 && p.n_tiles == 1 && p.block_n <= 128 && !(flags & YC_PLAN_NO_TILE_PAIR) &&
                 k_blocks * (p.block_k / 16) >= 12 && p.m_tiles >= 2 * sms) ? 1 : 0;
  // CTA pairs (conv_fprop_pair_kernel, tcgen05 cta_group::2): 256-row tiles over a cluster of two CTAs, each staging its own
  // A rows and half of the weight rows.  Default: the 256-wide tiles of the tiled / im2col path with the specialised
  // epilogue (they are shared-memory-bandwidth bound with one CTA per MMA); YC_PLAN_CTA_PAIR_ALL: wherever legal (tests),
  // YC_PLAN_NO_CTA_PAIR: never.
  // Measured (yolov4-608 x 64, same box, tools/ab_layers.py pair / nopair): 3x3 512->1024 @19 214 -> 153 us (1.42 PFLOP/s),
  // 256->256 @38 120 -> 91 us, 512->512 @19 137 -> 100 us, 1x1 1024->512 @19 41 -> 34 us; the 1x1 layers with K <= 512 (8 K
  // blocks or fewer) gain nothing or lose up to 10 % (their tiles are short: the pair's start-up and the remote hand-backs
  // show) -> from 12 K blocks on (first A/B; lowered to 8 by the second one, below).
  // Halo mode: the 128-wide tiles with at least 128 input channels (3x3 128->128 @76: one MMA of a 128 x 128 tile is 64
  // tensor-pipe cycles of work against an issue interval of ~175 for MMAs on one accumulator; a pair of CTAs doubles the
  // work per MMA, two tiles in flight double it again).
  // Second same-box A/B (default / pairs everywhere / default / pairs everywhere, profiles/r02_ab_cta_pair_all.txt): halo
  // 3x3 64->64 @152 233 -> 185 us (-18 %), 1x1 with 8 K blocks (512->256 @38, 512->512 @19 / @38) -7..-15 %; 1x1 with 4 K blocks
  // +-5 % (noise), everything at 304^2 and the stride-2 layers +13..+60 % -> halo from 64 channels on, tiled / im2col from
  // 8 K blocks on.
  const bool pair_rule = p.a_mode == 2 ? (p.n_tiles == 1 && ((p.block_n == 128 && x->c >= 128) || (p.block_n == 64 && x->c == 64)))
                                       : (p.block_n == 256 && k_blocks >= 8);
  p.cta_pair = (p.fast_epi && !(flags & YC_PLAN_NO_CTA_PAIR) && sms >= 2 &&
                (pair_rule || (flags & YC_PLAN_CTA_PAIR_ALL))) ? 1 : 0;
  if (p.cta_pair) p.tile_pair = 0;
  const uint32_t bh_bytes = (uint32_t)round_up((p.block_n / 2) * p.block_k * 2, 1024);
  const uint32_t stage_bytes_g = p.cta_pair ? a_bytes + bh_bytes : (p.tile_pair ? 2u : 1u) * a_bytes + b_bytes;  // tiled / im2col ring stage
  auto layout = [&](int out_bufs) -> bool {
    p.out_bufs = out_bufs;
    p.halo_stages = 2;
    p.halo_pair = 0;
    const uint32_t stage_bytes = p.a_mode == 2 ? (p.cta_pair ? bh_bytes : b_bytes) : stage_bytes_g;
    const uint32_t a_ring = (p.tile_pair ? 2u : 1u) * a_bytes;
    const uint32_t out_bytes = kEpiWarps * kStageOutBytes * out_bufs;
    const uint32_t fixed = 1024 /*align slack*/ + 512 /*barriers*/ + bias_bytes;
    if (p.a_mode == 2 && p.n_tiles == 1 && !(flags & YC_PLAN_NO_HALO_PAIR)) {
      const uint32_t need = out_bytes + 4 * kHaloPatchStride + 3 * stage_bytes + fixed;
      if (need <= 227 * 1024) {
        p.halo_stages = 4;
        p.halo_pair = 1;
      }
    }
    const uint32_t out_stage = out_bytes + (p.a_mode == 2 ? p.halo_stages * kHaloPatchStride : 0);
    if (out_stage + fixed + 2 * stage_bytes > 227 * 1024) return false;
    const uint32_t smem_budget = 227 * 1024 - fixed - out_stage;
    p.stages = std::min<int>(kMaxStages, smem_budget / stage_bytes);
    p.w_resident = 0;
    if (p.a_mode == 2 && p.n_tiles == 1 && 9 * p.cchunks <= p.stages) {
      p.w_resident = 1;  // every weight slab fits: load them once per CTA instead of once per tile
      p.stages = 9 * p.cchunks;
    }
    pl->smem = (size_t)p.stages * stage_bytes + out_stage + fixed;
    if (p.a_mode != 2 && !p.cta_pair && p.n_tiles == 1 && (wres_env == 2 || (wres_env == 1 && p.taps > 1))) {
      const uint32_t wres_bytes = (uint32_t)k_blocks * b_bytes;
      const long long tiles_per_cta = (long long)p.m_tiles / std::max(1, sms);
      if (wres_bytes <= 64 * 1024 && tiles_per_cta >= 8 && smem_budget >= wres_bytes + 4 * a_ring) {
        p.w_resident = 1;
        p.stages = std::min<int>(kMaxStages, (smem_budget - wres_bytes) / a_ring);
        pl->smem = (size_t)p.stages * a_ring + wres_bytes + out_stage + fixed;
      }
    }
    return p.stages >= 2;
  };
  // Staging buffers per epilogue warp: two would let a warp stage block k+1 while the TMA engine still reads block k, at
  // 32 KB of ring.  Measured with the specialised epilogue (tools/ab_layers.py, yolov4-608 x 64, round 2): one buffer is
  // as fast or faster on every layer (11.32 vs 11.40 ms of conv per step; halo mode loses its tile pairs / resident
  // weights with two: 3x3 64->64 @152 209 -> 323 us).  YC_PLAN_TWO_OUT_BUFS forces two where they fit.
  // The bias copy in shared memory is taken only where it does not cost a ring stage (a 128 x 256 tile has four 48 KB
  // stages in 227 KB with 2.5 KB to spare: 512+ columns of bias would make that three; measured -5..-14 % on those layers).
  auto layout_bias = [&](int out_bufs) -> bool {
    bias_bytes = 0;
    p.bias_cols = 0;
    if (!layout(out_bufs)) return false;
    if (p.fast_epi) {
      const int stages0 = p.stages, pair0 = p.halo_pair, wres0 = p.w_resident, hst0 = p.halo_stages;
      bias_bytes = 4u * (uint32_t)(p.n_tiles * p.block_n);
      if (layout(out_bufs) && p.stages == stages0 && p.halo_pair == pair0 && p.w_resident == wres0 && p.halo_stages == hst0) {
        p.bias_cols = p.n_tiles * p.block_n;
      } else {
        bias_bytes = 0;
        layout(out_bufs);
      }
    }
    return true;
  };
  bool ok_layout;
  if (flags & YC_PLAN_TWO_OUT_BUFS) ok_layout = layout_bias(2) || layout_bias(1);
  else ok_layout = layout_bias(1);
  if (!ok_layout) {
    set_error("yc_conv_plan_create: tile does not fit shared memory");
    yc_conv_plan_destroy(pl);
    return YC_ERR_ARG;
  }
  // Epilogue warp sets: set s takes tiles s, s + sets, ... of the CTA, each with its own TMEM accumulator.  The fast
  // epilogue gives every warp a whole 32-row strip of its tile (sets = accumulators that fit: 4 up to 128 columns, else 2),
  // so the per-tile work is paid once per strip; the generic one keeps four warps per quadrant on a wide tile.
  const int groups = p.block_n / 32;
  if (p.fast_epi && !(flags & YC_PLAN_EPI_SETS_V1)) {
    p.epi_sets = p.block_n <= 128 ? 4 : 2;
    p.nacc = p.epi_sets;
  } else {
    p.epi_sets = groups >= 4 ? 1 : (groups >= 2 ? 2 : 4);
    p.nacc = std::max(2, p.epi_sets);
    if (p.block_n <= 128) p.nacc = 4;  // 4 x 128 columns fit: deeper MMA/epilogue decoupling for free
  }
  p.issue_v2 = g_issue_v2;
  uint32_t cols = 32;
  while (cols < (uint32_t)(p.nacc * p.block_n)) cols <<= 1;
  p.tmem_cols = cols;
  p.chunk = p.y_f32 ? 16 : 32;  // columns per TMA store box: 64-byte rows
  p.res = residual ? (const __nv_bfloat16*)residual->ptr : nullptr;
  p.ldres = residual ? residual->ldc : 0;

  // ---- padded bias / scale ------------------------------------------------------------------------
  const int npad = p.n_tiles * p.block_n;
  cudaError_t e = cudaMalloc(&pl->bias_pad, sizeof(float) * npad);
  if (e == cudaSuccess) e = cudaMemset(pl->bias_pad, 0, sizeof(float) * npad);
  if (e == cudaSuccess && bias) e = cudaMemcpy(pl->bias_pad, bias, sizeof(float) * N, cudaMemcpyDeviceToDevice);
  if (e == cudaSuccess && scale) {
    e = cudaMalloc(&pl->scale_pad, sizeof(float) * npad);
    if (e == cudaSuccess) e = cudaMemset(pl->scale_pad, 0, sizeof(float) * npad);
    if (e == cudaSuccess) e = cudaMemcpy(pl->scale_pad, scale, sizeof(float) * N, cudaMemcpyDeviceToDevice);
  }
  // the memsets / copies above run on the legacy default stream, asynchronously to the host; launches go to the
  // caller's (possibly non-blocking) stream, so finish them here (setup time)
  if (e == cudaSuccess) e = cudaStreamSynchronize(cudaStreamLegacy);
  if (e != cudaSuccess) {
    set_error("yc_conv_plan_create: bias staging failed: %s", cudaGetErrorString(e));
    yc_conv_plan_destroy(pl);
    return YC_ERR_CUDA;
  }
  p.bias = pl->bias_pad;
  p.scale = pl->scale_pad;

  // ---- tensor maps ---------------------------------------------------------------------------------
  const CUtensorMapSwizzle swz = swizzle_for(p.block_k);
  CUresult cr;
  if (p.a_mode == 2) {
    // the (kHaloH+2) x kHaloPitch pixel patch around a tile: plain tiled box over {c, w, h, n}, zero fill outside
    cuuint64_t dims[4] = {(cuuint64_t)x->c, (cuuint64_t)x->w, (cuuint64_t)x->h, (cuuint64_t)x->n};
    cuuint64_t strides[3] = {(cuuint64_t)x->ldc * 2, (cuuint64_t)x->ldc * 2 * x->w, (cuuint64_t)x->ldc * 2 * x->w * x->h};
    cuuint32_t box[4] = {(cuuint32_t)p.block_k, (cuuint32_t)kHaloPitch, (cuuint32_t)(kHaloH + 2), 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    cr = g_encode_tiled(&pl->tmap_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, x->ptr, dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  } else if (p.a_mode == 0) {
    cuuint64_t dims[2] = {(cuuint64_t)x->c, (cuuint64_t)p.M};
    cuuint64_t strides[1] = {(cuuint64_t)x->ldc * 2};
    cuuint32_t box[2] = {(cuuint32_t)p.block_k, (cuuint32_t)kBlockM};
    cuuint32_t estr[2] = {1, 1};
    cr = g_encode_tiled(&pl->tmap_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, x->ptr, dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  } else {
    cuuint64_t dims[4] = {(cuuint64_t)x->c, (cuuint64_t)x->w, (cuuint64_t)x->h, (cuuint64_t)x->n};
    cuuint64_t strides[3] = {(cuuint64_t)x->ldc * 2, (cuuint64_t)x->ldc * 2 * x->w,
                             (cuuint64_t)x->ldc * 2 * x->w * x->h};
    // base-pixel bounding box (fprop): lower = -pad, upper = pad - (k-1)*dilation  (relative to W-1/H-1)
    int lower[2] = {-pad_l, -d->pad};
    int upper[2] = {pad_r - (kw - 1) * d->dilation, d->pad - (d->ksize - 1) * d->dilation};
    cuuint32_t estr[4] = {1, (cuuint32_t)stride_w, (cuuint32_t)d->stride, 1};
    cr = g_encode_im2col(&pl->tmap_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, x->ptr, dims, strides, lower, upper,
                         (cuuint32_t)p.block_k, (cuuint32_t)kBlockM, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  if (cr != CUDA_SUCCESS) {
    set_error("yc_conv_plan_create: activation tensor map encode failed (CUresult %d, mode %d, c=%d ldc=%d)", (int)cr,
              p.a_mode, x->c, x->ldc);
    yc_conv_plan_destroy(pl);
    return YC_ERR_CUDA;
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)p.taps * in_c_p, (cuuint64_t)N};
    cuuint64_t strides[1] = {(cuuint64_t)p.taps * in_c_p * 2};
    cuuint32_t box[2] = {(cuuint32_t)p.block_k, (cuuint32_t)(p.cta_pair ? p.block_n / 2 : p.block_n)};
    cuuint32_t estr[2] = {1, 1};
    pl->w_packed = w_packed;
    pl->w_kdim = p.taps * in_c_p;
    cr = g_encode_tiled(&pl->tmap_b, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(w_packed), dims, strides,
                        box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  if (cr != CUDA_SUCCESS) {
    set_error("yc_conv_plan_create: weight tensor map encode failed (CUresult %d)", (int)cr);
    yc_conv_plan_destroy(pl);
    return YC_ERR_CUDA;
  }

  {
    // output map: [M rows][n_store columns] with the view's pixel pitch; box = 32 rows x chunk columns, swizzled
    // like the staging tile the epilogue writes.  Rows >= M and columns >= n_store are clipped by the TMA engine.
    const int elt_b = p.y_f32 ? 4 : 2;
    const CUtensorMapSwizzle oswz = CU_TENSOR_MAP_SWIZZLE_64B;
    const CUtensorMapDataType odt = p.y_f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    auto encode_out = [&](CUtensorMap* tm, const yc_tensor* t, int cols) -> CUresult {
      if (p.a_mode == 2) {
        // halo mode: a warp's 32 rows are a 4 x 8 pixel block of the image -> box {chunk, 8, 4, 1} over {c, w, h, n};
        // the staging tile has the same row order (w fastest), clipping at the right / bottom edge is the engine's job
        cuuint64_t dims[4] = {(cuuint64_t)cols, (cuuint64_t)Wo, (cuuint64_t)Ho, (cuuint64_t)x->n};
        cuuint64_t strides[3] = {(cuuint64_t)t->ldc * elt_b, (cuuint64_t)t->ldc * elt_b * Wo, (cuuint64_t)t->ldc * elt_b * Wo * Ho};
        cuuint32_t box[4] = {(cuuint32_t)p.chunk, (cuuint32_t)kHaloW, 4, 1};
        cuuint32_t estr[4] = {1, 1, 1, 1};
        return g_encode_tiled(tm, odt, 4, t->ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, oswz,
                              CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      }
      cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)p.M};
      cuuint64_t strides[1] = {(cuuint64_t)t->ldc * elt_b};
      cuuint32_t box[2] = {(cuuint32_t)p.chunk, 32};
      cuuint32_t estr[2] = {1, 1};
      return g_encode_tiled(tm, odt, 2, t->ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, oswz,
                            CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    };
    cr = encode_out(&pl->tmap_y, y, y2 ? y->c : p.n_store);
    if (cr == CUDA_SUCCESS) cr = y2 ? encode_out(&pl->tmap_y2, y2, n_store2) : (pl->tmap_y2 = pl->tmap_y, CUDA_SUCCESS);
    if (cr != CUDA_SUCCESS) {
      set_error("yc_conv_plan_create: output tensor map encode failed (CUresult %d, c=%d ldc=%d)", (int)cr, p.n_store, y->ldc);
      yc_conv_plan_destroy(pl);
      return YC_ERR_CUDA;
    }
  }
  pl->rect = rect;
  pl->pdl = (flags & YC_PLAN_PDL) != 0;

  p.m_tiles128 = p.m_tiles;
  if (p.cta_pair) {  // 256-row tiles, one cluster of two CTAs per tile at a time
    p.m_tiles = ceil_div(p.m_tiles, 2);
    pl->grid = 2 * (int)std::min<long long>((long long)p.m_tiles * p.n_tiles, sms / 2);
  } else {
    pl->grid = (int)std::min<long long>((long long)p.m_tiles * p.n_tiles, sms);
  }
  static bool attr_set = false;
  if (!attr_set) {
    e = cudaFuncSetAttribute(conv_fprop_kernel_t<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(conv_fprop_kernel_t<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(conv_fprop_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(max dynamic smem) failed: %s", cudaGetErrorString(e));
      yc_conv_plan_destroy(pl);
      return YC_ERR_CUDA;
    }
    attr_set = true;
  }
  *out_plan = pl;
  return YC_OK;
}

// ---- data gradient through the forward kernel -------------------------------------------------------------
extern "C" int yc_conv_dgrad_channels(int out_c) {  // channel count (and minimum pitch) of the dY view the GEMM reads
  return out_c <= 64 ? round_up(out_c, 16) : round_up(out_c, 64);
}
extern "C" size_t yc_conv_dgrad_packed_bytes(int out_c, int in_c, int ksize) {
  return (size_t)in_c * ksize * ksize * yc_conv_padded_in_c(yc_conv_dgrad_channels(out_c)) * 2;
}
extern "C" int yc_conv_pack_dgrad(const float* w, int out_c, int in_c, int ksize, void* out, void* stream) {
  YC_REQUIRE(w && out && out_c > 0 && in_c > 0 && ksize > 0, "yc_conv_pack_dgrad: bad arguments");
  const int out_c_p = yc_conv_padded_in_c(yc_conv_dgrad_channels(out_c));
  const long long total = (long long)in_c * ksize * ksize * out_c_p;
  const int blocks = (int)std::min<long long>(ceil_div_ll(total, 256), 148 * 8);
  pack_weights_dgrad_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(w, out_c, in_c, ksize, out_c_p, (__nv_bfloat16*)out);
  YC_LAUNCH_CHECK();
  return YC_OK;
}
// geometry of the zero-inserted dY a strided layer's data gradient reads (stride 1: dY itself)
extern "C" int yc_conv_dgrad_dilated_size(const yc_conv_desc* d, int in_h, int in_w, int* hu, int* wu) {
  YC_REQUIRE(d && hu && wu, "yc_conv_dgrad_dilated_size: null argument");
  const int eff = d->dilation * (d->ksize - 1) + 1;
  *hu = in_h + 2 * d->pad - eff + 1;
  *wu = in_w + 2 * d->pad - eff + 1;
  return YC_OK;
}
extern "C" int yc_conv_dgrad_dilate(const yc_tensor* dy, int stride, const yc_tensor* up, void* stream) {
  YC_REQUIRE(dy && up && dy->dtype == YC_BF16 && up->dtype == YC_BF16 && stride >= 1, "yc_conv_dgrad_dilate: bad arguments");
  YC_REQUIRE(up->n == dy->n && up->c == dy->c && up->ldc % 8 == 0 && dy->ldc % 8 == 0 && (dy->h - 1) * stride < up->h &&
                 (dy->w - 1) * stride < up->w && ((uintptr_t)dy->ptr & 15) == 0 && ((uintptr_t)up->ptr & 15) == 0,
             "yc_conv_dgrad_dilate: shape mismatch");
  const int c8 = std::min(dy->ldc, up->ldc) / 8;
  YC_REQUIRE(c8 * 8 >= dy->c, "yc_conv_dgrad_dilate: pitch too small");
  const long long total = (long long)up->n * up->h * up->w * c8;
  const int blocks = (int)std::max<long long>(1, std::min<long long>(ceil_div_ll(total, 256), (long long)sm_count() * 16));
  dgrad_dilate_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)dy->ptr, dy->n, dy->h, dy->w, c8, dy->ldc,
                                                                 stride, up->h, up->w, up->ldc, (__nv_bfloat16*)up->ptr);
  YC_LAUNCH_CHECK();
  return YC_OK;
}
// dX = dgrad(dY): `d` is the FORWARD layer's geometry; dy_s1 is dY for stride-1 layers or its zero-inserted
// version (yc_conv_dgrad_dilate) for strided ones; dx is the forward input's shape.  accumulate != 0 adds to dx
// (cuDNN beta=1 semantics) instead of overwriting it (col2im semantics).
extern "C" int yc_conv_dgrad_plan_create(yc_conv_plan** plan, const yc_conv_desc* d, const yc_tensor* dy_s1,
                                         const void* w_packed_dgrad, const yc_tensor* dx, int accumulate) {
  YC_REQUIRE(plan && d && dy_s1 && w_packed_dgrad && dx, "yc_conv_dgrad_plan_create: null argument");
  yc_conv_desc g;
  g.ksize = d->ksize;
  g.stride = 1;
  g.dilation = d->dilation;
  g.pad = (d->ksize - 1) * d->dilation - d->pad;
  g.activation = YC_LINEAR;
  g.residual_activation = YC_LINEAR;
  YC_REQUIRE(g.pad >= 0, "yc_conv_dgrad: pad larger than the filter reach is not supported");
  return yc_conv_plan_create(plan, &g, dy_s1, w_packed_dgrad, nullptr, nullptr, accumulate ? dx : nullptr, dx);
}

// Training-mode forward: also accumulate per-channel sum / sum of squares of the (pre-batchnorm) results.
extern "C" int yc_conv_plan_set_stats(yc_conv_plan* pl, float* sum, float* sumsq) {
  YC_REQUIRE(pl && ((sum && sumsq) || (!sum && !sumsq)), "yc_conv_plan_set_stats: bad arguments");
  ConvParams& p = pl->p;
  YC_REQUIRE(!sum || p.act == YC_LINEAR, "yc_conv_plan_set_stats: statistics are taken on the linear (pre-batchnorm) output");
  pl->pdl = false;
  p.fast_epi = 0;                       // statistics are reduced by the generic epilogue
  p.bias_cols = 0;
  p.tile_pair = 0;
  if (p.cta_pair) {  // back to one CTA per tile: 128-row tiles, the whole N tile's weight rows per stage
    p.cta_pair = 0;
    p.m_tiles = p.m_tiles128;
    pl->grid = (int)std::min<long long>((long long)p.m_tiles * p.n_tiles, sm_count());
    cuuint64_t dims[2] = {(cuuint64_t)pl->w_kdim, (cuuint64_t)p.N};
    cuuint64_t strides[1] = {(cuuint64_t)pl->w_kdim * 2};
    cuuint32_t box[2] = {(cuuint32_t)p.block_k, (cuuint32_t)p.block_n};
    cuuint32_t estr[2] = {1, 1};
    CUresult cr = g_encode_tiled(&pl->tmap_b, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(pl->w_packed), dims,
                                 strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for(p.block_k),
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    YC_REQUIRE(cr == CUDA_SUCCESS, "yc_conv_plan_set_stats: weight tensor map encode failed (CUresult %d)", (int)cr);
  }
  if (p.a_mode != 2) p.w_resident = 0;  // the ring is re-planned below with weights in every stage
  // the generic epilogue's warp-to-tile assignment: four warps per quadrant on a wide tile (fewer statistics regions)
  const int groups = p.block_n / 32;
  p.epi_sets = groups >= 4 ? 1 : (groups >= 2 ? 2 : 4);
  const uint32_t a_bytes = p.a_mode == 2 ? 0u : kBlockM * p.block_k * 2;
  const uint32_t b_bytes = (uint32_t)round_up(p.block_n * p.block_k * 2, 1024);
  const uint32_t stage_bytes = a_bytes + b_bytes;
  const uint32_t out_stage = kEpiWarps * kStageOutBytes * p.out_bufs + (p.a_mode == 2 ? p.halo_stages * kHaloPatchStride : 0);
  p.stat_cols = sum ? p.n_tiles * p.block_n : 0;
  p.stat_regions = 4 * p.epi_sets;
  const uint32_t stat_bytes = 2u * 4u * (uint32_t)(p.stat_cols * p.stat_regions);
  YC_REQUIRE(1024 + 512 + out_stage + stat_bytes + 2 * stage_bytes <= 227 * 1024, "tile does not fit shared memory");
  const uint32_t smem_budget = 227 * 1024 - 1024 - 512 - out_stage - stat_bytes;
  p.stages = std::min<int>(kMaxStages, smem_budget / stage_bytes);
  if (p.w_resident) {
    if (9 * p.cchunks <= p.stages) p.stages = 9 * p.cchunks;
    else p.w_resident = 0;
  }
  pl->smem = (size_t)p.stages * stage_bytes + out_stage + 1024 + 512 + stat_bytes;
  if (pl->stat_partial) { cudaFree(pl->stat_partial); pl->stat_partial = nullptr; }
  if (pl->stat_ticket) { cudaFree(pl->stat_ticket); pl->stat_ticket = nullptr; }
  if (sum) {
    cudaError_t e = cudaMalloc(&pl->stat_partial, (size_t)pl->grid * 2 * p.stat_cols * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&pl->stat_ticket, sizeof(unsigned));
    if (e == cudaSuccess) e = cudaMemset(pl->stat_ticket, 0, sizeof(unsigned));
    if (e == cudaSuccess) e = cudaStreamSynchronize(cudaStreamLegacy);
    if (e != cudaSuccess) {
      set_error("yc_conv_plan_set_stats: %s", cudaGetErrorString(e));
      return YC_ERR_ALLOC;
    }
  }
  p.stat_partial = pl->stat_partial;
  p.stat_ticket = pl->stat_ticket;
  p.stat_sum = sum;
  p.stat_sq = sumsq;
  return YC_OK;
}

// Fold yc_bn_finalize into a statistics plan's kernel: its last CTA writes mean / var (and updates the rolling averages,
// either may be NULL) from the sums it has just formed.  count = pixels per channel (batch * out_h * out_w).
extern "C" int yc_conv_plan_set_bn_finalize(yc_conv_plan* pl, long long count, float* mean, float* var, float* rolling_mean,
                                            float* rolling_var) {
  YC_REQUIRE(pl && pl->p.stat_sum && mean && var && count > 1,
             "yc_conv_plan_set_bn_finalize: needs a plan with statistics (yc_conv_plan_set_stats) and mean / var buffers");
  pl->p.bn_mean = mean; pl->p.bn_var = var; pl->p.bn_rmean = rolling_mean; pl->p.bn_rvar = rolling_var;
  pl->p.bn_count = (float)count;
  return YC_OK;
}

// ForwardYoloLayerGpu folded into the head convolution: the plan's fp32 results go, decoded, straight into the [yolo]
// layer's NCHW fp32 output (out_nchw: [batch][channels][h*w], channels = anchors * entries); the plan's own output tensor
// is not written any more.  entries = 5 + classes (9 + classes for Gaussian heads).
extern "C" int yc_conv_plan_set_yolo_decode(yc_conv_plan* pl, float* out_nchw, int entries, float scale_x_y, int gaussian) {
  YC_REQUIRE(pl && out_nchw && entries > 0, "yc_conv_plan_set_yolo_decode: bad arguments");
  ConvParams& p = pl->p;
  YC_REQUIRE(p.y_f32 && p.a_mode != 2 && !p.fast_epi && !p.cta_pair && !p.res && !p.stat_sum && p.act == YC_LINEAR &&
                 !p.split_col && p.N % entries == 0,
             "yc_conv_plan_set_yolo_decode: needs a plain fp32-output linear convolution whose channels are whole anchors");
  p.dec_out = out_nchw;
  p.dec_S = p.Ho * p.Wo;
  p.dec_E = entries;
  p.dec_gauss = gaussian ? 1 : 0;
  p.dec_sxy = scale_x_y;
  return YC_OK;
}

// The plan holds a padded COPY of the bias taken when it was created: after the bias changes (an SGD step on a layer
// without batchnorm, i.e. the YOLO heads) the copy has to follow.  Asynchronous on `stream`, N floats device to device.
extern "C" int yc_conv_plan_set_bias(yc_conv_plan* pl, const float* bias, void* stream) {
  YC_REQUIRE(pl && bias && pl->bias_pad, "yc_conv_plan_set_bias: bad arguments");
  YC_CUDA(cudaMemcpyAsync(pl->bias_pad, bias, sizeof(float) * pl->p.N, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return YC_OK;
}

extern "C" int yc_conv_plan_destroy(yc_conv_plan* pl) {
  if (!pl) return YC_OK;
  if (pl->bias_pad) cudaFree(pl->bias_pad);
  if (pl->scale_pad) cudaFree(pl->scale_pad);
  if (pl->stat_partial) cudaFree(pl->stat_partial);
  if (pl->stat_ticket) cudaFree(pl->stat_ticket);
  delete pl;
  return YC_OK;
}

extern "C" int yc_conv_fprop(const yc_conv_plan* pl, void* stream) {
  YC_REQUIRE(pl, "yc_conv_fprop: null plan");
  if (pl->p.stat_sum)
    conv_fprop_kernel_t<true><<<pl->grid, kThreads, pl->smem, (cudaStream_t)stream>>>(pl->tmap_a, pl->tmap_b, pl->tmap_y, pl->tmap_y2, pl->p);
  else if (pl->pdl) {
    // programmatic dependent launch: this kernel's set-up overlaps the tail of the kernel before it in the stream
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)pl->grid);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = pl->smem;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    cudaError_t e = pl->p.cta_pair
        ? cudaLaunchKernelEx(&cfg, conv_fprop_pair_kernel, pl->tmap_a, pl->tmap_b, pl->tmap_y, pl->tmap_y2, pl->p)
        : cudaLaunchKernelEx(&cfg, conv_fprop_kernel_t<false>, pl->tmap_a, pl->tmap_b, pl->tmap_y, pl->tmap_y2, pl->p);
    if (e != cudaSuccess) {
      set_error("yc_conv_fprop: cudaLaunchKernelEx -> %s", cudaGetErrorString(e));
      return YC_ERR_CUDA;
    }
  } else if (pl->p.cta_pair)  // cluster of two CTAs per 256-row tile (static __cluster_dims__)
    conv_fprop_pair_kernel<<<pl->grid, kThreads, pl->smem, (cudaStream_t)stream>>>(pl->tmap_a, pl->tmap_b, pl->tmap_y, pl->tmap_y2, pl->p);
  else
    conv_fprop_kernel_t<false><<<pl->grid, kThreads, pl->smem, (cudaStream_t)stream>>>(pl->tmap_a, pl->tmap_b, pl->tmap_y, pl->tmap_y2, pl->p);
  YC_LAUNCH_CHECK();
  return YC_OK;
}

// (A/B) plans created from now on issue the halo taps unrolled (1, default) or with the rolled loop (0)
extern "C" int yc_conv_fprop_set_issue_v2(int on) {
  g_issue_v2 = on ? 1 : 0;
  return YC_OK;
}

extern "C" int yc_conv_plan_info(const yc_conv_plan* pl, int what) {
  if (!pl) return -1;
  switch (what) {
    case 0: return pl->p.block_n;
    case 1: return pl->p.block_k;
    case 2: return pl->p.stages;
    case 3: return pl->p.m_tiles * pl->p.n_tiles;
    case 4: return pl->grid;
    case 5: return pl->p.a_mode;
    case 6: return (int)pl->smem;
    case 7:  // kernel variant: [0,2) a_mode, 4 halo tile pairs, 8 resident weights, 16 two outputs, 32 two staging buffers, 64 rectangular, 128 fast epilogue, 256 tile pairs, 512 CTA pairs, 1024 programmatic dependent launch
      return pl->p.a_mode | (pl->p.halo_pair ? 4 : 0) | (pl->p.w_resident ? 8 : 0) | (pl->p.split_col ? 16 : 0) |
             (pl->p.out_bufs == 2 ? 32 : 0) | (pl->rect ? 64 : 0) | (pl->p.fast_epi ? 128 : 0) | (pl->p.tile_pair ? 256 : 0) |
             (pl->p.cta_pair ? 512 : 0) | (pl->pdl ? 1024 : 0);
    case 8: return pl->p.epi_sets;
  }
  return -1;
}
