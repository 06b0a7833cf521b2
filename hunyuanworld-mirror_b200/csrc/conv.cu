// Implicit-GEMM convolution on tcgen05 for the DPT prediction heads (SURVEY.md section 8f row 1;
// reference: src/models/heads/dense_head.py:164-344).  The reference runs its heads with autocast disabled
// (worldmirror.py:146), i.e. in fp32; bf16 activations (8-bit mantissa) measured 1.5e-2 relative error on pts3d (the
// inv_log output amplifies), above the stated 1e-2.  A 10-bit mantissa passes (1.2e-3): the first version of this file ran
// kind::tf32 on fp32 maps rounded to TF32; the same 10 bits fit fp16, which the tensor cores read at TWICE the tf32 rate
// and which halves every map in HBM.  So activations and weights are fp16 in memory (round to nearest, saturating: the
// maps of these heads are O(1), fp16 reaches 65504), the MMA is kind::f16 with fp32 accumulation, and every epilogue does
// its bias / residual / activation arithmetic in fp32 before rounding once.  Activations are NHWC, so a 3x3 convolution is a GEMM
//     acc[pixel, co] (fp32, TMEM) = sum over (tap, ci) of  in[pixel + tap offset, ci] * w[co, tap, ci]
// whose A operand needs no im2col buffer: the producer issues one 4D TMA tile load per (tap, 32-channel block) with
// the tile origin shifted by the tap offset, and TMA's out-of-bounds zero fill IS the convolution's zero padding.
// A tile = 8 x 16 pixels x 64 channels (fp16) = 128 rows of 128 B in the SWIZZLE_128B layout the UMMA descriptors expect.
//
// One CTA per SM, persistent over (image, tile row, tile column, Cout tile), 256 threads:
//   warp 0 = TMA producer, warp 1 = tcgen05.mma issuer (one elected lane), warp 2 = TMEM allocator, warp 3 idle,
//   warps 4-7 = epilogue (thread == pixel == accumulator row == TMEM lane).
// 4-stage smem ring; two TMEM accumulators of NT columns so the epilogue of tile i overlaps the main loop of i+1.
//
// Epilogues (the reference's module boundaries they fuse are cited at the launchers):
//   NT = 128 / 256:  out fp16 = [relu]( acc + bias + res1 + res2 ), written as NHWC through a 4D TMA store (pixels of a
//                    border tile that fall outside the image are clipped by the store);
//   NT = 32 (final): relu(acc + bias) -> 1x1 conv (32 -> od) -> activate_head -> fp32 prediction + confidence;
//   NT = 256 (project): relu(acc + bias) -> 1x1 conv (256 -> od <= 16) + bias, raw fp32 [pixels, od]: the Gaussian
//                    parameter head `gs_renderer.gs_head` (models/rasterization.py:149-153).  The hidden map never leaves
//                    tensor memory: the epilogue rewrites it as fp16 in place and a second MMA (A from TMEM, W2 from
//                    shared memory) projects it.
#include <cuda_fp16.h>

#include "common.cuh"
#include "ptx.cuh"

namespace wm {

// Pixels x 32-channel groups of an fp16 output map whose fp32 value left the fp16 range (or was not finite) since the
// last reset: the DPT heads keep their maps in fp16, and a checkpoint or input that overflows them must be reported.
__device__ unsigned int g_fp16_saturated = 0;
// max(m, |a|, |b|) per fp16 lane of packed pairs, and "a lane reached the clamp value 65504 (0x7bff) or is not finite"
__device__ __forceinline__ uint32_t habs_max2(uint32_t m, uint32_t a, uint32_t b) {
  uint32_t d;
  asm("{\n.reg .b32 t;\nmax.NaN.f16x2 t, %1, %2;\nmax.NaN.f16x2 %0, t, %3;\n}" : "=r"(d) : "r"(a & 0x7fff7fffu), "r"(b & 0x7fff7fffu), "r"(m));
  return d;
}
__device__ __forceinline__ bool fp16_pair_saturated(uint32_t m) { return (m & 0xffffu) >= 0x7bffu || (m >> 16) >= 0x7bffu; }


constexpr int CV_TW = 16, CV_TH = 8;   // tile = 8 rows x 16 columns of pixels
constexpr int CV_STAGES = 4;
constexpr int CV_KE = 64;              // fp16 elements per k-block (one 128-byte swizzle row)
constexpr int CV_A_BYTES = 128 * CV_KE * 2;
constexpr int CV_THREADS = 256;
constexpr int CV_STG_BYTES = 128 * 128;
// "Tall" mode (3x3 convolutions with Cout tiles of 32 / 128): the narrow Cout tiles are bound by the nine-fold re-read of
// their input tile from L2, so the producer loads ONE box of (8 + 2) x 16 pixels per (dx, channel block) and the three dy taps
// read it through descriptors that start 0 / 1 / 2 pixel rows (2048 B, a multiple of the 1024-byte swizzle atom) into it:
// three times fewer input bytes.  A stage then holds that box and the three weight tiles of the taps (dy, dx).
constexpr int CV_TALL_A_BYTES = (CV_TH + 2) * CV_TW * 128;
template <int NT, bool TALL>
constexpr int cv_stages() {
  return TALL && NT >= 128 ? 3 : CV_STAGES;
}
template <int NT, bool TALL>
constexpr int cv_stage_bytes() {
  return TALL ? CV_TALL_A_BYTES + 3 * NT * 128 : CV_A_BYTES + NT * 128;
}
template <int NT, bool TALL = false>
constexpr int cv_smem() {
  return cv_stages<NT, TALL>() * cv_stage_bytes<NT, TALL>() + CV_STG_BYTES + 256 + 1024;
}

enum : int { ACT_EXP = 0, ACT_INV_LOG = 1, ACT_NORM = 2 };
enum : int { EPI_STORE = 0, EPI_FINAL = 1, EPI_PROJECT = 2 };
constexpr int CV_PROJ_MAX = 16;  // output channels of the projecting epilogue (its weights live in the staging tile)

struct ConvArgs {
  int n, H, W, Cin, Cout, taps;  // taps: 1 (1x1), 9 (3x3, stride 1, pad 1) or 7 (7x1 vertical, stride 1, pad 3)
  int tiles_x, tiles_y, n_tiles, total_tiles, kblocks;  // kblocks = taps * Cin / 64
  const float* bias;            // [Cout] or null
  const void* res1;             // fp16 [n, H, W, Cout], or fp32 [H, W, Cout] when res1_bcast, or null
  const __half* res2;           // fp16 [n, H, W, Cout] or null
  int res1_bcast;               // res1 is the fp32 positional table, shared by all images
  int relu;                     // 1: on the whole sum; 2: on acc + bias only, before the residuals are added
  // final / projecting epilogue
  const float *w2, *b2;         // [od, NT], [od]
  int od, act;
  float *pred, *conf;           // final: [n, H, W, od - 1], [n, H, W]; project: pred = [n, H, W, od]
};

template <int NT, int EPI, bool TALL>
__global__ void __launch_bounds__(CV_THREADS, 1)
conv_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const __grid_constant__ CUtensorMap tmC, const ConvArgs g) {
  constexpr int B_BYTES = NT * 128;
  constexpr bool FINAL = EPI == EPI_FINAL;
  constexpr bool STORE = EPI == EPI_STORE;
  static_assert(!FINAL || NT == 32, "final epilogue: 32 hidden channels");
  static_assert(EPI != EPI_PROJECT || CV_PROJ_MAX * NT * 2 <= CV_STG_BYTES, "projection weights must fit the staging tile");
  static_assert(EPI != EPI_PROJECT || (NT % 64 == 0 && NT / 2 + CV_PROJ_MAX <= NT), "projecting epilogue reuses the accumulator columns");
  constexpr uint32_t TMEM_COLS = 2 * NT < 32 ? 32 : 2 * NT;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  constexpr int STAGES = TALL && NT >= 128 ? 3 : CV_STAGES;  // = cv_stages<NT, TALL>()
  constexpr int A_STAGE = TALL ? CV_TALL_A_BYTES : CV_A_BYTES;
  constexpr int B_STAGE = TALL ? 3 * B_BYTES : B_BYTES;
  uint8_t* sA = smem;
  uint8_t* sB = sA + STAGES * A_STAGE;
  uint8_t* sC = sB + STAGES * B_STAGE;
  uint64_t* full = reinterpret_cast<uint64_t*>(sC + CV_STG_BYTES);
  uint64_t* empty = full + CV_STAGES;
  uint64_t* tfull = empty + CV_STAGES;
  uint64_t* tempty = tfull + 2;
  uint64_t* hfull = tempty + 2;   // projecting epilogue: the fp16 hidden map of accumulator i is in TMEM
  uint64_t* d2full = hfull + 2;   // projecting epilogue: the projected columns of accumulator i are in TMEM
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(d2full + 2);

  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5;
  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (STORE) tma_prefetch_desc(&tmC);
  }
  if (warp == 1 && elect_one()) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], 4);  // one arrival per epilogue warp
      mbar_init(&hfull[i], 4);
      mbar_init(&d2full[i], 1);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_ptr, TMEM_COLS);
    tmem_relinquish();
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  pdl_wait();

  // tile -> (Cout tile, tile column, tile row, image); the Cout tile varies fastest so CTAs that run together share A
  auto decode = [&](int tile, int& n0, int& x0, int& y0, int& img) {
    n0 = (tile % g.n_tiles) * NT;
    int r = tile / g.n_tiles;
    x0 = (r % g.tiles_x) * CV_TW;
    r /= g.tiles_x;
    y0 = (r % g.tiles_y) * CV_TH;
    img = r / g.tiles_y;
  };
  const int cblocks = g.Cin / CV_KE;

  if (warp == 0) {
    if (elect_one()) {
      uint32_t it = 0;
      for (int tile = blockIdx.x; tile < g.total_tiles; tile += gridDim.x) {
        int n0, x0, y0, img;
        decode(tile, n0, x0, y0, img);
        for (int kb = 0; kb < g.kblocks; ++kb, ++it) {
          const uint32_t s = it % STAGES, ph = (it / STAGES) & 1;
          const int tap = kb / cblocks, cb = kb - tap * cblocks;
          mbar_wait(&empty[s], ph ^ 1);
          mbar_arrive_expect_tx(&full[s], A_STAGE + B_STAGE);
          if constexpr (TALL) {  // tap = dx index: one box of 10 pixel rows, the weight tiles of (dy = -1, 0, 1; dx)
            tma_load_4d(sA + s * A_STAGE, &tmA, &full[s], cb * CV_KE, x0 + tap - 1, y0 - 1, img, kEvictNormal);
#pragma unroll
            for (int dyi = 0; dyi < 3; ++dyi)
              tma_load_2d(sB + s * B_STAGE + dyi * B_BYTES, &tmB, &full[s], ((dyi * 3 + tap) * cblocks + cb) * CV_KE, n0, kEvictLast);
          } else {
            const int dy = g.taps == 9 ? tap / 3 - 1 : (g.taps == 7 ? tap - 3 : 0), dx = g.taps == 9 ? tap % 3 - 1 : 0;
            tma_load_4d(sA + s * A_STAGE, &tmA, &full[s], cb * CV_KE, x0 + dx, y0 + dy, img, kEvictNormal);
            tma_load_2d(sB + s * B_STAGE, &tmB, &full[s], kb * CV_KE, n0, kEvictLast);
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_f16(128, NT);
      uint32_t it = 0, tcount = 0;
      // projecting epilogue: out[128 x 16] = relu(hidden)[128 x NT, fp16 in TMEM] x W2^T[NT x 16, fp16 in smem], issued for
      // the PREVIOUS tile in the middle of the current tile's main loop (its hidden map has been converted by then)
      auto project = [&](uint32_t pc) {
        const uint32_t pacc = pc & 1, pph = (pc >> 1) & 1;
        mbar_wait(&hfull[pacc], pph);
        tcgen05_fence_after();
        constexpr uint32_t idesc2 = umma_idesc_f16(128, CV_PROJ_MAX);
        const uint32_t base = tmem_base + pacc * NT;
#pragma unroll
        for (int k = 0; k < NT / 16; ++k) {  // 16 hidden channels = 8 TMEM columns of packed fp16 per step
          const uint64_t wdesc = umma_desc_sw128(smem_u32(sC) + (k >> 2) * (CV_PROJ_MAX * 128), 16, 1024) + 2 * (k & 3);
          umma_ts(base + NT / 2, base + k * 8, wdesc, idesc2, k != 0);
        }
        umma_commit(&d2full[pacc]);
      };
      for (int tile = blockIdx.x; tile < g.total_tiles; tile += gridDim.x, ++tcount) {
        const uint32_t acc = tcount & 1, aph = (tcount >> 1) & 1;
        mbar_wait(&tempty[acc], aph ^ 1);
        tcgen05_fence_after();
        const uint32_t d_tmem = tmem_base + acc * NT;
        for (int kb = 0; kb < g.kblocks; ++kb, ++it) {
          if constexpr (EPI == EPI_PROJECT) {
            if (kb == g.kblocks / 2 && tcount > 0) project(tcount - 1);
          }
          const uint32_t s = it % STAGES, ph = (it / STAGES) & 1;
          mbar_wait(&full[s], ph);
          tcgen05_fence_after();
#pragma unroll
          for (int dyi = 0; dyi < (TALL ? 3 : 1); ++dyi) {
            // tall mode: the tile of tap dy starts dyi pixel rows (16 pixels x 128 B) into the box
            const uint64_t adesc = umma_desc_sw128(smem_u32(sA + s * A_STAGE + dyi * (CV_TW * 128)), 16, 1024);
            const uint64_t bdesc = umma_desc_sw128(smem_u32(sB + s * B_STAGE + dyi * B_BYTES), 16, 1024);
#pragma unroll
            for (int k = 0; k < 4; ++k)  // K = 16 fp16 = 32 bytes (>> 4 = 2) per instruction inside the swizzle row
              umma_ss(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | dyi | k) != 0);
          }
          umma_commit(&empty[s]);
        }
        umma_commit(&tfull[acc]);
      }
      if constexpr (EPI == EPI_PROJECT) {
        if (tcount > 0) project(tcount - 1);
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    const int et = threadIdx.x - 128;  // accumulator row == pixel of the tile
    const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const uint32_t stg = smem_u32(sC);
    uint32_t tcount = 0;
    if constexpr (EPI == EPI_PROJECT) {
      // w2 [od, NT] fp32 -> fp16 B operand of the projecting MMA: NT / 64 k-blocks of 16 rows (outputs, zero beyond od)
      // x 128 B in the SWIZZLE_128B K-major layout (the layout TMA would have written)
      for (int i = et; i < (NT / 64) * CV_PROJ_MAX * 8; i += 128) {
        const int ch = i & 7, n = (i >> 3) % CV_PROJ_MAX, kb = i / (8 * CV_PROJ_MAX);
        uint32_t w[4] = {0u, 0u, 0u, 0u};
        if (n < g.od) {
          const float* src = g.w2 + (size_t)n * NT + kb * 64 + ch * 8;
#pragma unroll
          for (int q = 0; q < 4; ++q) w[q] = pack_half2_sat(__ldg(src + 2 * q), __ldg(src + 2 * q + 1));
        }
        st_shared_v4(stg + kb * (CV_PROJ_MAX * 128) + sw128_offset(n, ch), w[0], w[1], w[2], w[3]);
      }
      fence_proxy_async_smem();
      named_bar_sync(1, 128);
    }
    for (int tile = blockIdx.x; tile < g.total_tiles; tile += gridDim.x, ++tcount) {
      int n0, x0, y0, img;
      decode(tile, n0, x0, y0, img);
      const int y = y0 + et / CV_TW, x = x0 + et % CV_TW;
      const bool valid = y < g.H && x < g.W;
      const size_t pix = ((size_t)img * g.H + y) * g.W + x;
      const uint32_t acc = tcount & 1, aph = (tcount >> 1) & 1;
      // Residual maps (the skip inputs of a ResidualConvUnit / fusion block) are read by this thread for its own pixel, 64 bytes
      // per map and 32-channel group.  The loads of group g + 1 are issued before group g is processed, and those of group 0
      // before the accumulator is even complete, so their DRAM / L2 latency hides under the main loop and the previous
      // group's arithmetic instead of stalling every group (with two residual maps the 148 x 148 convolution took 228 us
      // against 160 us without: profiles/r02_heads_launches.csv).
      uint4 rbuf[2][2][4];
      auto issue_residuals = [&](int grp, uint4 (&dst)[2][4]) {
        if constexpr (STORE) {
#pragma unroll
          for (int rsel = 0; rsel < 2; ++rsel) {
            const __half* rp = rsel == 0 ? (g.res1_bcast ? nullptr : static_cast<const __half*>(g.res1)) : g.res2;
            if (rp == nullptr || !valid) continue;
            const uint4* r4 = reinterpret_cast<const uint4*>(rp + pix * g.Cout + n0 + grp * 32);
#pragma unroll
            for (int i = 0; i < 4; ++i) dst[rsel][i] = __ldg(r4 + i);
          }
        }
      };
      issue_residuals(0, rbuf[0]);
      mbar_wait(&tfull[acc], aph);
      tcgen05_fence_after();
      const uint32_t taddr = tmem_base + acc * NT + lane_base;
      if constexpr (FINAL) {
        // output_conv2 (dense_head.py:85-89): relu(conv3x3 + b) -> conv1x1(32 -> od) -> activate_head (:297-344)
        uint32_t v[32];
        tmem_ld_x32(taddr, v);
        tmem_wait_ld();
        tcgen05_fence_before();
        __syncwarp();
        if ((threadIdx.x & 31) == 0) mbar_arrive(&tempty[acc]);
        float o[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (j < g.od) o[j] = __ldg(g.b2 + j);
#pragma unroll
        for (int c = 0; c < 32; ++c) {
          const float h = fmaxf(__uint_as_float(v[c]) + __ldg(g.bias + c), 0.f);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (j < g.od) o[j] = fmaf(__ldg(g.w2 + j * 32 + c), h, o[j]);
        }
        if (valid) {
          const int na = g.od - 1;  // 1 (depth) or 3 (points, normals); the last channel is the confidence logit
          const float cl = na == 1 ? o[1] : o[3];
          g.conf[pix] = 1.0f + expf(cl);
          float* pr = g.pred + pix * na;
          float nn = 0.f;
#pragma unroll
          for (int j = 0; j < 3; ++j)
            if (j < na) nn = fmaf(o[j], o[j], nn);
          const float inv = 1.0f / sqrtf(nn);
#pragma unroll
          for (int j = 0; j < 3; ++j) {
            if (j < na) {
              float r;
              if (g.act == ACT_EXP)
                r = expf(o[j]);
              else if (g.act == ACT_INV_LOG)
                r = copysignf(expm1f(fabsf(o[j])), o[j]);
              else
                r = o[j] * inv;
              pr[j] = r;
            }
          }
        }
      } else if constexpr (EPI == EPI_PROJECT) {
        // relu(acc [+ b1]) -> fp16, written back IN PLACE into the first NT / 2 columns of this accumulator (the columns a
        // group is written to have already been read); the MMA warp multiplies it with W2 into columns [NT/2, NT/2 + 16).
        // (The first version did the 256 x od multiply-adds per pixel on the CUDA cores with the weights broadcast from
        // shared memory: 16 useful bytes per shared-memory wavefront, in competition with the MMA's own operand reads -
        // tensor pipe 50 % active.)
#pragma unroll 1
        for (int grp = 0; grp < NT / 32; ++grp) {
          uint32_t v[32];
          tmem_ld_x32(taddr + grp * 32, v);
          tmem_wait_ld();
          uint32_t pk[16];
          uint32_t mx2 = 0u;
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float a = __uint_as_float(v[2 * i]), b = __uint_as_float(v[2 * i + 1]);
            if (g.bias != nullptr) a += __ldg(g.bias + grp * 32 + 2 * i), b += __ldg(g.bias + grp * 32 + 2 * i + 1);
            pk[i] = pack_half2_sat(fmaxf(a, 0.f), fmaxf(b, 0.f));
          }
#pragma unroll
          for (int i = 0; i < 16; i += 2) mx2 = habs_max2(mx2, pk[i], pk[i + 1]);
          if (valid && fp16_pair_saturated(mx2)) atomicAdd(&g_fp16_saturated, 1u);
          tmem_st_x16(taddr + grp * 16, pk);
        }
        tmem_wait_st();
        tcgen05_fence_before();
        __syncwarp();
        if ((threadIdx.x & 31) == 0) mbar_arrive(&hfull[acc]);
        mbar_wait(&d2full[acc], aph);
        tcgen05_fence_after();
        uint32_t d2[16];
        tmem_ld_x16(taddr + NT / 2, d2);
        tmem_wait_ld();
        tcgen05_fence_before();
        __syncwarp();
        if ((threadIdx.x & 31) == 0) mbar_arrive(&tempty[acc]);
        if (valid) {
          float* pr = g.pred + pix * g.od;
#pragma unroll
          for (int j = 0; j < CV_PROJ_MAX; ++j)
            if (j < g.od) pr[j] = __uint_as_float(d2[j]) + __ldg(g.b2 + j);
        }
      } else {
#pragma unroll 2
        for (int grp = 0; grp < NT / 32; ++grp) {
          if (grp + 1 < NT / 32) issue_residuals(grp + 1, rbuf[(grp + 1) & 1]);
          float xv[32];
          {
            uint32_t v0[32];
            tmem_ld_x32(taddr + grp * 32, v0);
            tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) xv[i] = __uint_as_float(v0[i]);
          }
          if (grp == NT / 32 - 1) {
            tcgen05_fence_before();
            __syncwarp();
            if ((threadIdx.x & 31) == 0) mbar_arrive(&tempty[acc]);
          }
          const int col0 = n0 + grp * 32;
          if (g.bias != nullptr) {
            const float4* b4 = reinterpret_cast<const float4*>(g.bias + col0);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float4 b = __ldg(b4 + i);
              xv[4 * i] += b.x, xv[4 * i + 1] += b.y, xv[4 * i + 2] += b.z, xv[4 * i + 3] += b.w;
            }
          }
          if (g.relu == 2) {
#pragma unroll
            for (int i = 0; i < 32; ++i) xv[i] = fmaxf(xv[i], 0.f);
          }
          if (valid) {
            if (g.res1 != nullptr && g.res1_bcast) {  // fp32 positional table [H, W, Cout]
              const float4* r4 = reinterpret_cast<const float4*>(static_cast<const float*>(g.res1) + ((size_t)y * g.W + x) * g.Cout + col0);
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                const float4 r = __ldg(r4 + i);
                xv[4 * i] += r.x, xv[4 * i + 1] += r.y, xv[4 * i + 2] += r.z, xv[4 * i + 3] += r.w;
              }
            }
#pragma unroll
            for (int rsel = 0; rsel < 2; ++rsel) {
              const __half* rp = rsel == 0 ? (g.res1_bcast ? nullptr : static_cast<const __half*>(g.res1)) : g.res2;
              if (rp == nullptr) continue;
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const uint4 r = rbuf[grp & 1][rsel][i];  // loaded one group ahead (issue_residuals)
                const uint32_t u[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&u[k]));
                  xv[8 * i + 2 * k] += f.x, xv[8 * i + 2 * k + 1] += f.y;
                }
              }
            }
          }
          if (g.relu == 1) {
#pragma unroll
            for (int i = 0; i < 32; ++i) xv[i] = fmaxf(xv[i], 0.f);
          }
          // two 32-column groups fill one 128-byte staging row (64 fp16 channels) per pixel
          const int half = grp & 1;
          if (half == 0) {
            if (et == 0) tma_store_wait_read<0>();  // the previous store has drained the staging tile
            named_bar_sync(1, 128);
          }
          // fp16 maps: a value beyond the fp16 range is clamped to +-65504 by cvt.satfinite -- never silently: the packed
          // results are max-reduced (one HMNMX2 per two outputs) and a group that reached the clamp value is counted
          uint32_t mx2 = 0u;
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const uint32_t p0 = pack_half2_sat(xv[8 * c], xv[8 * c + 1]), p1 = pack_half2_sat(xv[8 * c + 2], xv[8 * c + 3]);
            const uint32_t p2 = pack_half2_sat(xv[8 * c + 4], xv[8 * c + 5]), p3 = pack_half2_sat(xv[8 * c + 6], xv[8 * c + 7]);
            mx2 = habs_max2(habs_max2(mx2, p0, p1), p2, p3);
            st_shared_v4(stg + sw128_offset(et, half * 4 + c), p0, p1, p2, p3);
          }
          if (valid && fp16_pair_saturated(mx2)) atomicAdd(&g_fp16_saturated, 1u);
          if (half == 1) {
            fence_proxy_async_smem();
            named_bar_sync(1, 128);
            if (et == 0) {
              tma_store_4d(&tmC, sC, col0 - 32, x0, y0, img);
              tma_store_commit();
            }
          }
        }
      }
    }
    if (STORE && et == 0) tma_store_wait<0>();
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 2) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------ host
template <int NT, int EPI, bool TALL = false>
static wm_status launch_conv(cudaStream_t st, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, ConvArgs g) {
  int sms = 0;
  WM_TRY(require_sm100(&sms));
  static bool configured = false;
  if (!configured) {
    WM_CUDA(cudaFuncSetAttribute(conv_kernel<NT, EPI, TALL>, cudaFuncAttributeMaxDynamicSharedMemorySize, cv_smem<NT, TALL>()));
    configured = true;
  }
  g.tiles_x = (g.W + CV_TW - 1) / CV_TW;
  g.tiles_y = (g.H + CV_TH - 1) / CV_TH;
  g.n_tiles = g.Cout / NT;
  g.total_tiles = g.n * g.tiles_x * g.tiles_y * g.n_tiles;
  g.kblocks = (TALL ? 3 : g.taps) * (g.Cin / CV_KE);
  const int grid = g.total_tiles < sms ? g.total_tiles : sms;
  return launch_pdl(conv_kernel<NT, EPI, TALL>, dim3(grid), dim3(CV_THREADS), cv_smem<NT, TALL>(), st, ta, tb, tc, g);
}

static wm_status conv_check(const void* in, const void* w, int n, int H, int W, int Cin, int Cout, int taps) {
  WM_CHECK(in && w, "conv2d: null operand");
  WM_CHECK(n > 0 && H > 0 && W > 0 && Cin > 0 && Cin % CV_KE == 0 && Cout > 0, "conv2d: bad geometry n=%d H=%d W=%d Cin=%d Cout=%d", n,
           H, W, Cin, Cout);
  WM_CHECK(taps == 1 || taps == 9 || taps == 7, "conv2d: taps must be 1 (1x1), 9 (3x3, pad 1) or 7 (7x1 vertical, pad 3)");
  WM_CHECK((reinterpret_cast<uintptr_t>(in) & 15) == 0 && (reinterpret_cast<uintptr_t>(w) & 15) == 0, "conv2d: operands must be 16-byte aligned");
  return WM_OK;
}

// nn.Conv2d(Cin, Cout, 1 | 3, stride 1, padding 0 | 1) on NHWC fp16 with the surrounding elementwise work fused:
// out = [relu](conv(in) + bias + res1 + res2).  weight: fp16 [Cout, taps * Cin], tap-major (ky, kx, ci).
wm_status conv2d_nhwc(cudaStream_t st, const void* in, int n, int H, int W, int Cin, const void* weight, int taps, int Cout,
                      const float* bias, const void* res1, int res1_bcast, const void* res2, int relu, void* out) {
  WM_TRY(conv_check(in, weight, n, H, W, Cin, Cout, taps));
  WM_CHECK(out && (Cout % 256 == 0 || Cout == 128), "conv2d: Cout must be 128 or a multiple of 256 (got %d)", Cout);
  const CUtensorMapDataType f16 = CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  const int NT = Cout == 128 ? 128 : 256;
  CUtensorMap ta, tb, tc;
  const bool tall = NT == 128 && taps == 9;
  WM_TRY(tmap_nhwc(&ta, f16, 2, in, n, H, W, Cin, CV_KE, CV_TW, tall ? CV_TH + 2 : CV_TH));
  WM_TRY(tmap_2d(&tb, f16, 2, weight, Cout, (uint64_t)taps * Cin, (uint64_t)taps * Cin, CV_KE, NT));
  WM_TRY(tmap_nhwc(&tc, f16, 2, out, n, H, W, Cout, 64, CV_TW, CV_TH));
  ConvArgs g{};
  g.n = n, g.H = H, g.W = W, g.Cin = Cin, g.Cout = Cout, g.taps = taps;
  g.bias = bias, g.res1 = res1, g.res1_bcast = res1_bcast;
  g.res2 = static_cast<const __half*>(res2), g.relu = relu;
  if (tall) return launch_conv<128, EPI_STORE, true>(st, ta, tb, tc, g);
  return NT == 128 ? launch_conv<128, EPI_STORE>(st, ta, tb, tc, g) : launch_conv<256, EPI_STORE>(st, ta, tb, tc, g);
}

// scratch.output_conv2 + activate_head (dense_head.py:85-89, 297-344): conv3x3(Cin -> 32) + ReLU + conv1x1(32 -> od)
// + attribute / confidence activations, fp32 out.  in, w1: fp16 (w1 [32, 9 * Cin] tap-major); w2: fp32 [od, 32].
wm_status conv2d_final(cudaStream_t st, const void* in, int n, int H, int W, int Cin, const void* w1, const float* b1,
                       const float* w2, const float* b2, int od, int act, float* pred, float* conf) {
  WM_TRY(conv_check(in, w1, n, H, W, Cin, 32, 9));
  WM_CHECK(b1 && w2 && b2 && pred && conf && (od == 2 || od == 4) && act >= ACT_EXP && act <= ACT_NORM, "conv2d_final: bad argument");
  const CUtensorMapDataType f16 = CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  CUtensorMap ta, tb;
  WM_TRY(tmap_nhwc(&ta, f16, 2, in, n, H, W, Cin, CV_KE, CV_TW, CV_TH + 2));
  WM_TRY(tmap_2d(&tb, f16, 2, w1, 32, (uint64_t)9 * Cin, (uint64_t)9 * Cin, CV_KE, 32));
  ConvArgs g{};
  g.n = n, g.H = H, g.W = W, g.Cin = Cin, g.Cout = 32, g.taps = 9;
  g.bias = b1, g.w2 = w2, g.b2 = b2, g.od = od, g.act = act, g.pred = pred, g.conf = conf;
  return launch_conv<32, EPI_FINAL, true>(st, ta, tb, ta, g);
}

// gs_renderer.gs_head (models/rasterization.py:149-153): conv3x3(Cin -> 256, bias optional) + ReLU + conv1x1(256 -> od) + b2,
// raw fp32 [n, H, W, od]; the 256-channel hidden map stays in TMEM.  in, w1: fp16 (w1 [256, 9 * Cin] tap-major); w2: fp32 [od, 256].
wm_status conv2d_project(cudaStream_t st, const void* in, int n, int H, int W, int Cin, const void* w1, const float* b1,
                         const float* w2, const float* b2, int od, float* out) {
  WM_TRY(conv_check(in, w1, n, H, W, Cin, 256, 9));
  WM_CHECK(w2 && b2 && out && od > 0 && od <= CV_PROJ_MAX, "conv2d_project: bad argument (od=%d, at most %d)", od, CV_PROJ_MAX);
  const CUtensorMapDataType f16 = CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  CUtensorMap ta, tb;
  WM_TRY(tmap_nhwc(&ta, f16, 2, in, n, H, W, Cin, CV_KE, CV_TW, CV_TH));
  WM_TRY(tmap_2d(&tb, f16, 2, w1, 256, (uint64_t)9 * Cin, (uint64_t)9 * Cin, CV_KE, 256));
  ConvArgs g{};
  g.n = n, g.H = H, g.W = W, g.Cin = Cin, g.Cout = 256, g.taps = 9;
  g.bias = b1, g.w2 = w2, g.b2 = b2, g.od = od, g.pred = out;
  return launch_conv<256, EPI_PROJECT>(st, ta, tb, ta, g);
}

wm_status fp16_saturation_read(cudaStream_t st, unsigned* host_pinned, int reset) {
  WM_CHECK(host_pinned != nullptr, "fp16_saturation_read: null destination");
  WM_CUDA(cudaMemcpyFromSymbolAsync(host_pinned, g_fp16_saturated, sizeof(unsigned), 0, cudaMemcpyDeviceToHost, st));
  if (reset) {
    void* p = nullptr;
    WM_CUDA(cudaGetSymbolAddress(&p, g_fp16_saturated));
    WM_CUDA(cudaMemsetAsync(p, 0, sizeof(unsigned), st));
  }
  return WM_OK;
}

}  // namespace wm
