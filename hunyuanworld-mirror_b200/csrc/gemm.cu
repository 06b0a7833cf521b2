// Persistent, warp-specialised tcgen05 GEMM for the aggregator's linear layers:
//     acc[M,N] (fp32, TMEM) = A[M,K] (bf16, row-major) x W[N,K]^T (bf16, nn.Linear layout)
// with the epilogue fused per reference call site:
//   EPI_BIAS   out bf16 = acc + b                                   generic nn.Linear
//   EPI_GELU   out bf16 = gelu_erf(acc + b)                         Mlp.fc1 + nn.GELU   (mlp.py:30-31)
//   EPI_QKV    per 64-wide head: + b, QK-LayerNorm(64), 2D RoPE     attention.py:50-56, rope.py:148-181
//              q -> out0[M,1024], k|v -> out1[M,2048]  (both bf16)
//   EPI_RESID  x fp32 (in place) = x + gamma * (acc + b), optional second strided copy of the result
//              (attn.proj / Mlp.fc2 + LayerScale + residual: block.py:91-92, layer_scale.py:17, and the
//              `cat([frame, global], -1)` intermediates of visual_transformer.py:337-339)
//   EPI_TABLE  out fp32 = acc + table[row % period][col]            DINOv2 patch embed + pos_embed / cls / reg
//
// One CTA per SM, 256 threads: warp 0 = TMA producer, warp 1 = MMA issuer (one elected lane),
// warp 2 = TMEM allocator, warps 4-7 = epilogue (thread == accumulator row; TMEM lane quarter = warp % 4).
// Tile 128 x 256 x 64, 4-stage smem ring (SWIZZLE_128B, K-major for both operands), two 256-column TMEM
// accumulators so the epilogue of tile i overlaps the main loop of tile i+1.  Outputs leave through a
// swizzled staging tile and TMA stores (row clipping for the M tail comes for free); the fp32 residual is
// TMA-loaded into that same staging tile, updated in place and stored back.
#include "common.cuh"
#include "ptx.cuh"

namespace wm {

enum : int { EPI_BIAS = 0, EPI_GELU = 1, EPI_QKV = 2, EPI_RESID = 3, EPI_TABLE = 4 };

constexpr int BM = 128, BN = 256, BK = 64, STAGES = 4;
constexpr int A_BYTES = BM * BK * 2;        // 16 KB
constexpr int B_BYTES = BN * BK * 2;        // 32 KB
constexpr int STG_BYTES = 128 * 128;        // 128 rows x 128 B
constexpr int GEMM_THREADS = 256;
constexpr int GEMM_SMEM = STAGES * (A_BYTES + B_BYTES) + 2 * STG_BYTES + 256 + 1024;

struct GemmArgs {
  int M, N, K;
  int num_m_tiles, num_n_tiles, num_k_blocks;
  const float* bias;
  // EPI_QKV
  const float *qn_w, *qn_b, *kn_w, *kn_b;  // [64] each, or null (DINOv2: no QK-norm)
  const float *rope_cos, *rope_sin;        // [max_pos + 1][16], or null (DINOv2: no RoPE)
  int tokens_per_view, patch_start, grid_w;
  float qk_eps;
  // EPI_RESID
  const float* gamma;  // null: out = acc + b (no residual read)
  int has_out2, out2_col;
  // EPI_TABLE
  const float* table;
  int table_period;
};

__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f)); }

// LayerNorm over the 64 values of one head held by one thread (attention.py:41-42,52; eps 1e-5, affine).
__device__ __forceinline__ void head_layernorm(float (&x)[64], const float* __restrict__ w,
                                               const float* __restrict__ b, float eps) {
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 64; ++i) s += x[i];
  const float mu = s * (1.0f / 64.0f);
  float v = 0.f;
#pragma unroll
  for (int i = 0; i < 64; ++i) {
    const float d = x[i] - mu;
    v += d * d;
  }
  const float rstd = rsqrtf(v * (1.0f / 64.0f) + eps);
#pragma unroll
  for (int i = 0; i < 64; ++i) x[i] = (x[i] - mu) * rstd * __ldg(w + i) + __ldg(b + i);
}

// 2D RoPE (rope.py:148-181): dims [0,32) rotate by y, [32,64) by x; inside each half the pair is (i, i+16).
__device__ __forceinline__ void head_rope(float (&x)[64], const float* __restrict__ cs, const float* __restrict__ sn,
                                          int py, int px) {
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    const int p = half == 0 ? py : px;
    const float* c = cs + p * 16;
    const float* s = sn + p * 16;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const float co = __ldg(c + i), si = __ldg(s + i);
      const float a = x[half * 32 + i], bb = x[half * 32 + 16 + i];
      x[half * 32 + i] = a * co - bb * si;
      x[half * 32 + 16 + i] = bb * co + a * si;
    }
  }
}

template <int EPI>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const __grid_constant__ CUtensorMap tmC0, const __grid_constant__ CUtensorMap tmC1, const GemmArgs g) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sA = smem;
  uint8_t* sB = sA + STAGES * A_BYTES;
  uint8_t* sC = sB + STAGES * B_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(sC + 2 * STG_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* tfull = empty + STAGES;
  uint64_t* tempty = tfull + 2;
  uint64_t* ldbar = tempty + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(ldbar + 2);

  const int warp = threadIdx.x >> 5;
  const int total_tiles = g.num_m_tiles * g.num_n_tiles;

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC0);
    if (EPI == EPI_QKV || EPI == EPI_RESID) tma_prefetch_desc(&tmC1);
  }
  if (warp == 1 && elect_one()) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], 128);
      mbar_init(&ldbar[i], 1);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_ptr, 512);
    tmem_relinquish();
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (elect_one()) {
      uint32_t it = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int m0 = (tile / g.num_n_tiles) * BM;
        const int n0 = (tile % g.num_n_tiles) * BN;
        for (int kb = 0; kb < g.num_k_blocks; ++kb, ++it) {
          const uint32_t s = it % STAGES, ph = (it / STAGES) & 1;
          mbar_wait(&empty[s], ph ^ 1);
          mbar_arrive_expect_tx(&full[s], A_BYTES + B_BYTES);
          tma_load_2d(sA + s * A_BYTES, &tmA, &full[s], kb * BK, m0, kEvictNormal);
          tma_load_2d(sB + s * B_BYTES, &tmB, &full[s], kb * BK, n0, kEvictLast);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(BM, BN, false, false);
      uint32_t it = 0, tcount = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++tcount) {
        const uint32_t acc = tcount & 1, aph = (tcount >> 1) & 1;
        mbar_wait(&tempty[acc], aph ^ 1);
        tcgen05_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < g.num_k_blocks; ++kb, ++it) {
          const uint32_t s = it % STAGES, ph = (it / STAGES) & 1;
          mbar_wait(&full[s], ph);
          tcgen05_fence_after();
          const uint64_t adesc = umma_desc_sw128(smem_u32(sA + s * A_BYTES), 16, 1024);
          const uint64_t bdesc = umma_desc_sw128(smem_u32(sB + s * B_BYTES), 16, 1024);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)  // +32 bytes (>>4 = 2) per 16-element K step inside the 128B swizzle row
            umma_ss(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
          umma_commit(&empty[s]);
        }
        umma_commit(&tfull[acc]);
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ------------------------------------------------------------ epilogue (128 threads, thread == row)
    const int et = threadIdx.x - 128;
    const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const uint32_t stg0 = smem_u32(sC);
    uint32_t tcount = 0, ccount = 0;

    if (EPI == EPI_RESID) {
      // prologue: residual chunk 0 of this CTA's first tile -> staging buffer 0
      if (et == 0 && g.gamma != nullptr && (int)blockIdx.x < total_tiles) {
        const int m0 = (blockIdx.x / g.num_n_tiles) * BM, n0 = (blockIdx.x % g.num_n_tiles) * BN;
        mbar_arrive_expect_tx(&ldbar[0], STG_BYTES);
        tma_load_2d(sC, &tmC0, &ldbar[0], n0, m0, kEvictNormal);
      }
    }

    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++tcount) {
      const int m0 = (tile / g.num_n_tiles) * BM;
      const int n0 = (tile % g.num_n_tiles) * BN;
      const uint32_t acc = tcount & 1, aph = (tcount >> 1) & 1;
      mbar_wait(&tfull[acc], aph);
      tcgen05_fence_after();
      const uint32_t taddr = tmem_base + acc * BN + lane_base;
      const int row = m0 + et;

      if (EPI == EPI_BIAS || EPI == EPI_GELU || EPI == EPI_QKV) {
        // ---- bf16 outputs: four 64-column groups (one attention head each for EPI_QKV)
        int py = 0, px = 0;
        if (EPI == EPI_QKV && g.rope_cos != nullptr) {
          const int p = row % g.tokens_per_view;
          if (p >= g.patch_start) {
            const int idx = p - g.patch_start;
            py = idx / g.grid_w + 1;
            px = idx % g.grid_w + 1;
          }
        }
#pragma unroll 1
        for (int grp = 0; grp < BN / 64; ++grp, ++ccount) {
          const uint32_t stg = stg0 + (ccount & 1) * STG_BYTES;
          if (et == 0) tma_store_wait_read<1>();  // the store that last used this buffer has drained it
          named_bar_sync(1, 128);
          float x[64];
          {
            uint32_t v0[32], v1[32];
            tmem_ld_x32(taddr + grp * 64, v0);
            tmem_ld_x32(taddr + grp * 64 + 32, v1);
            tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              x[i] = __uint_as_float(v0[i]);
              x[32 + i] = __uint_as_float(v1[i]);
            }
          }
          if (grp == BN / 64 - 1) {  // accumulator fully read: hand the TMEM buffer back to the MMA warp
            tcgen05_fence_before();
            mbar_arrive(&tempty[acc]);
          }
          const int col0 = n0 + grp * 64;
          const float* bias = g.bias + col0;
#pragma unroll
          for (int i = 0; i < 64; ++i) x[i] += __ldg(bias + i);
          if (EPI == EPI_GELU) {
#pragma unroll
            for (int i = 0; i < 64; ++i) x[i] = gelu_erf(x[i]);
          }
          if (EPI == EPI_QKV) {
            const int which = col0 >> 10;  // 0 = q, 1 = k, 2 = v  (attention.py:50: qkv rows are [q | k | v])
            if (which < 2) {
              if (g.qn_w != nullptr)
                head_layernorm(x, which == 0 ? g.qn_w : g.kn_w, which == 0 ? g.qn_b : g.kn_b, g.qk_eps);
              if (g.rope_cos != nullptr) head_rope(x, g.rope_cos, g.rope_sin, py, px);
            }
          }
#pragma unroll
          for (int c = 0; c < 8; ++c)
            st_shared_v4(stg + sw128_offset(et, c), pack_bf16(x[8 * c], x[8 * c + 1]),
                         pack_bf16(x[8 * c + 2], x[8 * c + 3]), pack_bf16(x[8 * c + 4], x[8 * c + 5]),
                         pack_bf16(x[8 * c + 6], x[8 * c + 7]));
          fence_proxy_async_smem();
          named_bar_sync(1, 128);
          if (et == 0) {
            const void* src = sC + (ccount & 1) * STG_BYTES;
            if (EPI == EPI_QKV && col0 >= kEmbed)
              tma_store_2d(&tmC1, src, col0 - kEmbed, m0);
            else
              tma_store_2d(&tmC0, src, col0, m0);
            tma_store_commit();
          }
        }
      } else {
        // ---- fp32 outputs: eight 32-column chunks
#pragma unroll 1
        for (int ch = 0; ch < BN / 32; ++ch, ++ccount) {
          const uint32_t buf = ccount & 1;
          const uint32_t stg = stg0 + buf * STG_BYTES;
          const int col0 = n0 + ch * 32;
          const bool resid = (EPI == EPI_RESID) && g.gamma != nullptr;
          if (et == 0) {
            if (resid) {
              // prefetch the next residual chunk (possibly of this CTA's next tile) into the other buffer
              int nt = tile, nch = ch + 1;
              if (nch == BN / 32) {
                nch = 0;
                nt = tile + gridDim.x;
              }
              tma_store_wait_read<0>();
              if (nt < total_tiles) {
                mbar_arrive_expect_tx(&ldbar[buf ^ 1], STG_BYTES);
                tma_load_2d(sC + (buf ^ 1) * STG_BYTES, &tmC0, &ldbar[buf ^ 1], (nt % g.num_n_tiles) * BN + nch * 32,
                            (nt / g.num_n_tiles) * BM, kEvictNormal);
              }
            } else {
              tma_store_wait_read<1>();
            }
          }
          if (resid)
            mbar_wait(&ldbar[buf], (ccount >> 1) & 1);
          else
            named_bar_sync(1, 128);
          uint32_t v[32];
          tmem_ld_x32(taddr + ch * 32, v);
          tmem_wait_ld();
          if (ch == BN / 32 - 1) {
            tcgen05_fence_before();
            mbar_arrive(&tempty[acc]);
          }
          const float* bias = g.bias + col0;
          const float* tab = nullptr;
          if (EPI == EPI_TABLE) tab = g.table + (size_t)(row % g.table_period) * g.N + col0;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            float o[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) o[i] = __uint_as_float(v[4 * c + i]);
            const uint32_t addr = stg + sw128_offset(et, c);
            if (EPI == EPI_TABLE) {
              const float4 t = __ldg(reinterpret_cast<const float4*>(tab + 4 * c));
              o[0] += t.x, o[1] += t.y, o[2] += t.z, o[3] += t.w;
            } else {
#pragma unroll
              for (int i = 0; i < 4; ++i) o[i] += __ldg(bias + 4 * c + i);
              if (resid) {
                uint32_t r0, r1, r2, r3;
                ld_shared_v4(addr, r0, r1, r2, r3);
                const float* gm = g.gamma + col0 + 4 * c;
                o[0] = __uint_as_float(r0) + __ldg(gm + 0) * o[0];
                o[1] = __uint_as_float(r1) + __ldg(gm + 1) * o[1];
                o[2] = __uint_as_float(r2) + __ldg(gm + 2) * o[2];
                o[3] = __uint_as_float(r3) + __ldg(gm + 3) * o[3];
              }
            }
            st_shared_v4(addr, __float_as_uint(o[0]), __float_as_uint(o[1]), __float_as_uint(o[2]),
                         __float_as_uint(o[3]));
          }
          fence_proxy_async_smem();
          named_bar_sync(1, 128);
          if (et == 0) {
            const void* src = sC + buf * STG_BYTES;
            tma_store_2d(&tmC0, src, col0, m0);
            if (EPI == EPI_RESID && g.has_out2) tma_store_2d(&tmC1, src, g.out2_col + col0, m0);
            tma_store_commit();
          }
        }
      }
    }
    if (et == 0) tma_store_wait<0>();
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 2) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------------------ host
static int g_sm_count = 0;

template <int EPI>
static wm_status launch(cudaStream_t st, const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& c0,
                        const CUtensorMap& c1, GemmArgs g) {
  if (g_sm_count == 0) WM_TRY(require_sm100(&g_sm_count));
  static bool configured = false;
  if (!configured) {
    WM_CUDA(cudaFuncSetAttribute(gemm_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM));
    configured = true;
  }
  g.num_m_tiles = (g.M + BM - 1) / BM;
  g.num_n_tiles = g.N / BN;
  g.num_k_blocks = (g.K + BK - 1) / BK;
  const int tiles = g.num_m_tiles * g.num_n_tiles;
  const int grid = tiles < g_sm_count ? tiles : g_sm_count;
  gemm_kernel<EPI><<<grid, GEMM_THREADS, GEMM_SMEM, st>>>(a, b, c0, c1, g);
  WM_LAUNCH_CHECK();
  return WM_OK;
}

static wm_status check_common(const void* A, const void* W, int M, int N, int K, int lda, int ldw) {
  WM_CHECK(A && W, "gemm: null operand");
  WM_CHECK(M > 0 && K > 0 && N > 0 && N % BN == 0, "gemm: need M,K > 0 and N a positive multiple of %d (M=%d N=%d K=%d)",
           BN, M, N, K);
  WM_CHECK(lda % 8 == 0 && ldw % 8 == 0 && lda >= K && ldw >= K, "gemm: row pitches must be >= K and multiples of 8 elements");
  WM_CHECK((reinterpret_cast<uintptr_t>(A) & 15) == 0 && (reinterpret_cast<uintptr_t>(W) & 15) == 0,
           "gemm: operands must be 16-byte aligned");
  return WM_OK;
}

static wm_status operand_maps(CUtensorMap* ta, CUtensorMap* tb, const void* A, int lda, const void* W, int ldw, int M,
                              int N, int K) {
  WM_TRY(tmap_2d(ta, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, A, M, K, lda, BK, BM));
  WM_TRY(tmap_2d(tb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, W, N, K, ldw, BK, BN));
  return WM_OK;
}

wm_status gemm_bias_act(cudaStream_t st, const void* A, int lda, const void* W, int ldw, const float* bias, int M,
                        int N, int K, void* out, int ldo, int gelu) {
  WM_TRY(check_common(A, W, M, N, K, lda, ldw));
  WM_CHECK(out && bias && ldo % 8 == 0 && ldo >= N, "gemm_bias_act: bad output");
  CUtensorMap ta, tb, tc;
  WM_TRY(operand_maps(&ta, &tb, A, lda, W, ldw, M, N, K));
  WM_TRY(tmap_2d(&tc, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, out, M, N, ldo, 64, 128));
  GemmArgs g{};
  g.M = M, g.N = N, g.K = K, g.bias = bias;
  return gelu ? launch<EPI_GELU>(st, ta, tb, tc, tc, g) : launch<EPI_BIAS>(st, ta, tb, tc, tc, g);
}

wm_status gemm_qkv(cudaStream_t st, const void* A, int lda, const void* W, const float* bias, const float* qn_w,
                   const float* qn_b, const float* kn_w, const float* kn_b, float qk_eps, const float* rope_cos,
                   const float* rope_sin, int tokens_per_view, int patch_start, int grid_w, int M, void* q_out,
                   void* kv_out) {
  const int N = 3 * kEmbed, K = kEmbed;
  WM_TRY(check_common(A, W, M, N, K, lda, K));
  WM_CHECK(q_out && kv_out && bias, "gemm_qkv: null output or bias");
  WM_CHECK((qn_w != nullptr) == (kn_w != nullptr) && (qn_w != nullptr) == (qn_b != nullptr) &&
               (qn_w != nullptr) == (kn_b != nullptr),
           "gemm_qkv: q/k norm parameters must be all present or all null");
  WM_CHECK((rope_cos != nullptr) == (rope_sin != nullptr), "gemm_qkv: rope tables must both be present or both null");
  WM_CHECK(rope_cos == nullptr || (tokens_per_view > 0 && grid_w > 0 && patch_start >= 0), "gemm_qkv: bad token geometry");
  CUtensorMap ta, tb, tq, tkv;
  WM_TRY(operand_maps(&ta, &tb, A, lda, W, K, M, N, K));
  WM_TRY(tmap_2d(&tq, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, q_out, M, kEmbed, kEmbed, 64, 128));
  WM_TRY(tmap_2d(&tkv, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, kv_out, M, 2 * kEmbed, 2 * kEmbed, 64, 128));
  GemmArgs g{};
  g.M = M, g.N = N, g.K = K, g.bias = bias;
  g.qn_w = qn_w, g.qn_b = qn_b, g.kn_w = kn_w, g.kn_b = kn_b, g.qk_eps = qk_eps;
  g.rope_cos = rope_cos, g.rope_sin = rope_sin;
  g.tokens_per_view = tokens_per_view > 0 ? tokens_per_view : 1, g.patch_start = patch_start,
  g.grid_w = grid_w > 0 ? grid_w : 1;
  return launch<EPI_QKV>(st, ta, tb, tq, tkv, g);
}

wm_status gemm_ls_residual(cudaStream_t st, const void* A, int lda, const void* W, int ldw, const float* bias,
                           const float* gamma, int M, int N, int K, float* x, int ldx, float* out2, int ld2,
                           int out2_col) {
  WM_TRY(check_common(A, W, M, N, K, lda, ldw));
  WM_CHECK(x && bias && ldx % 4 == 0 && ldx >= N, "gemm_ls_residual: bad x");
  CUtensorMap ta, tb, tx, t2;
  WM_TRY(operand_maps(&ta, &tb, A, lda, W, ldw, M, N, K));
  WM_TRY(tmap_2d(&tx, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, x, M, N, ldx, 32, 128));
  t2 = tx;
  if (out2) {
    WM_CHECK(ld2 % 4 == 0 && out2_col >= 0 && out2_col + N <= ld2, "gemm_ls_residual: bad out2 geometry");
    WM_TRY(tmap_2d(&t2, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, out2, M, ld2, ld2, 32, 128));
  }
  GemmArgs g{};
  g.M = M, g.N = N, g.K = K, g.bias = bias, g.gamma = gamma, g.has_out2 = out2 != nullptr, g.out2_col = out2_col;
  return launch<EPI_RESID>(st, ta, tb, tx, t2, g);
}

wm_status gemm_table(cudaStream_t st, const void* A, int lda, const void* W, int ldw, const float* table,
                     int table_period, int M, int N, int K, float* out, int ldo) {
  WM_TRY(check_common(A, W, M, N, K, lda, ldw));
  WM_CHECK(out && table && table_period > 0 && ldo % 4 == 0 && ldo >= N, "gemm_table: bad output/table");
  CUtensorMap ta, tb, tc;
  WM_TRY(operand_maps(&ta, &tb, A, lda, W, ldw, M, N, K));
  WM_TRY(tmap_2d(&tc, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, out, M, N, ldo, 32, 128));
  GemmArgs g{};
  g.M = M, g.N = N, g.K = K, g.bias = table, g.table = table, g.table_period = table_period;
  return launch<EPI_TABLE>(st, ta, tb, tc, tc, g);
}

}  // namespace wm
