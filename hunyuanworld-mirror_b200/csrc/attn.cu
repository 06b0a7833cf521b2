// Flash attention for head_dim 64 on tcgen05/TMEM, fed by TMA (F.scaled_dot_product_attention at
// attention.py:59; no mask, no dropout).  One kernel serves both call shapes of the aggregator:
//   frame  attention: batch = views, Lq = Lkv = tokens per view       (visual_transformer.py:313-323)
//   global attention: batch = 1,     Lq = local tokens, Lkv = all tokens (visual_transformer.py:324-334;
//                     under sequence parallelism Lkv spans the all-gathered K|V of every rank)
// q, k, v, o are bf16 with heads side by side in the row (q:[.,1024], k|v:[.,2048], o:[.,1024]) which is
// exactly what the QKV GEMM writes and the proj GEMM reads, so no permute/transposed copy exists.
//
// Persistent CTAs (one per SM, 384 threads).  A work item is (head, batch, 256 query rows):
//   warp 0      TMA producer: Q pair once per item, K/V tiles through a KV_STAGES ring
//   warp 1      MMA issuer (one lane): S_w = Q_w K^T (128x128x64), O_w += P_w V (128x64x128)
//   warp 2      TMEM allocator
//   warps 4-7   softmax / correction / epilogue for query tile 0 (thread == query row)
//   warps 8-11  the same for query tile 1 (ping-pong: tensor core works on one tile while the other is in softmax)
// TMEM: S0 [0,128) S1 [128,256) O0 [256,320) O1 [320,384); P (bf16) overwrites the first 64 columns of its S.
// Online softmax keeps a running max that is only advanced (and O rescaled) when it grew by more than 2^8,
// so the O correction is rare; the row sum and the final 1/l normalisation stay in fp32.
#include "common.cuh"
#include "ptx.cuh"

namespace wm {

constexpr int ATT_THREADS = 384;
constexpr int ATT_TILE = 128 * 64 * 2;  // one 128 x 64 bf16 tile, 16 KB

struct AttnArgs {
  int Lq, Lkv, batch;
  int n_qpairs, n_kv, total_items;
  float scale_log2;
  uint32_t v_lbo, v_sbo;  // V (MN-major B operand) descriptor strides in bytes
};

template <bool P_TMEM>
struct AttnCfg {
  static constexpr int KV_STAGES = P_TMEM ? 4 : 2;
  static constexpr int P_TILES = P_TMEM ? 0 : 4;  // two 64-wide K halves per query tile
  static constexpr int SMEM = (2 + 2 * KV_STAGES + 2 + P_TILES) * ATT_TILE + 512 + 1024;
};

__host__ __device__ constexpr uint32_t s_col(int w) { return 128u * w; }        // S0 / S1
__host__ __device__ constexpr uint32_t o_col(int w) { return 256u + 64u * w; }   // O0 / O1

template <bool P_TMEM>
__global__ void __launch_bounds__(ATT_THREADS, 1)
attn_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
            const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO, const AttnArgs a) {
  using Cfg = AttnCfg<P_TMEM>;
  constexpr int ST = Cfg::KV_STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sQ = smem;                       // 2 tiles
  uint8_t* sK = sQ + 2 * ATT_TILE;          // ST tiles
  uint8_t* sV = sK + ST * ATT_TILE;         // ST tiles
  uint8_t* sO = sV + ST * ATT_TILE;         // 2 tiles (epilogue staging, one per query tile)
  uint8_t* sP = sO + 2 * ATT_TILE;          // 4 tiles when P goes through shared memory
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + Cfg::P_TILES * ATT_TILE);
  uint64_t* q_full = bars;            // 1
  uint64_t* q_empty = bars + 1;       // 1
  uint64_t* k_full = bars + 2;        // ST
  uint64_t* k_empty = k_full + ST;    // ST
  uint64_t* v_full = k_empty + ST;    // ST
  uint64_t* v_empty = v_full + ST;    // ST
  uint64_t* s_full = v_empty + ST;    // 2
  uint64_t* p_full = s_full + 2;      // 2
  uint64_t* o_done = p_full + 2;      // 2
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(o_done + 2);

  const int warp = threadIdx.x >> 5;

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmO);
  }
  if (warp == 1 && elect_one()) {
    mbar_init(q_full, 1);
    mbar_init(q_empty, 1);
    for (int i = 0; i < ST; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&v_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], 128);
      mbar_init(&o_done[i], 1);
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

  if (warp < 4) {
    setmaxnreg_dec<56>();
    if (warp == 0) {
      // -------------------------------------------------------------- TMA producer
      if (elect_one()) {
        uint32_t kvit = 0, icount = 0;
        for (int item = blockIdx.x; item < a.total_items; item += gridDim.x, ++icount) {
          const int per_head = a.batch * a.n_qpairs;
          const int h = item / per_head, rem = item % per_head;
          const int b = rem / a.n_qpairs, q0 = (rem % a.n_qpairs) * 256;
          mbar_wait(q_empty, (icount & 1) ^ 1);
          mbar_arrive_expect_tx(q_full, 2 * ATT_TILE);
          tma_load_3d(sQ, &tmQ, q_full, h * 64, q0, b, kEvictFirst);
          tma_load_3d(sQ + ATT_TILE, &tmQ, q_full, h * 64, q0 + 128, b, kEvictFirst);
          for (int j = 0; j < a.n_kv; ++j, ++kvit) {
            const uint32_t s = kvit % ST, ph = (kvit / ST) & 1;
            mbar_wait(&k_empty[s], ph ^ 1);
            mbar_arrive_expect_tx(&k_full[s], ATT_TILE);
            tma_load_3d(sK + s * ATT_TILE, &tmK, &k_full[s], h * 64, j * 128, b, kEvictLast);
            mbar_wait(&v_empty[s], ph ^ 1);
            mbar_arrive_expect_tx(&v_full[s], ATT_TILE);
            tma_load_3d(sV + s * ATT_TILE, &tmV, &v_full[s], h * 64, j * 128, b, kEvictLast);
          }
        }
      }
      __syncwarp();
    } else if (warp == 1) {
      // -------------------------------------------------------------- MMA issuer
      if (elect_one()) {
        constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128, false, false);
        constexpr uint32_t idesc_pv = umma_idesc_bf16(128, 64, false, true);  // B = V is MN-major (dims contiguous)
        uint32_t kvit = 0, icount = 0;
        uint32_t pcount[2] = {0, 0};
        for (int item = blockIdx.x; item < a.total_items; item += gridDim.x, ++icount) {
          const int q0 = (item % a.n_qpairs) * 256;
          const int nw = (q0 + 128 < a.Lq) ? 2 : 1;  // second query tile entirely out of range: skip it
          mbar_wait(q_full, icount & 1);
          tcgen05_fence_after();
          auto issue_s = [&](int w, uint32_t stage) {
            const uint64_t qd = umma_desc_sw128(smem_u32(sQ + w * ATT_TILE), 16, 1024);
            const uint64_t kd = umma_desc_sw128(smem_u32(sK + stage * ATT_TILE), 16, 1024);
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_ss(tmem_base + s_col(w), qd + 2 * k, kd + 2 * k, idesc_s, k != 0);
            umma_commit(&s_full[w]);
          };
          {
            const uint32_t s0 = kvit % ST, ph0 = (kvit / ST) & 1;
            mbar_wait(&k_full[s0], ph0);
            tcgen05_fence_after();
            for (int w = 0; w < nw; ++w) issue_s(w, s0);
            umma_commit(&k_empty[s0]);
          }
          for (int j = 0; j < a.n_kv; ++j, ++kvit) {
            const uint32_t sj = kvit % ST, phj = (kvit / ST) & 1;
            const uint32_t sn = (kvit + 1) % ST, phn = ((kvit + 1) / ST) & 1;
            const bool more = j + 1 < a.n_kv;
            mbar_wait(&v_full[sj], phj);
            tcgen05_fence_after();
            for (int w = 0; w < nw; ++w) {
              mbar_wait(&p_full[w], pcount[w] & 1);
              ++pcount[w];
              tcgen05_fence_after();
              const uint32_t vbase = smem_u32(sV + sj * ATT_TILE);
#pragma unroll
              for (int k = 0; k < 8; ++k) {  // 16 kv rows (2048 B of the V tile) per step
                const uint64_t vd = umma_desc_sw128(vbase + k * 2048, a.v_lbo, a.v_sbo);
                if (P_TMEM) {
                  umma_ts(tmem_base + o_col(w), tmem_base + s_col(w) + k * 8, vd, idesc_pv, (j | k) != 0);
                } else {
                  const uint64_t pd =
                      umma_desc_sw128(smem_u32(sP + (2 * w + (k >> 2)) * ATT_TILE), 16, 1024) + 2 * (k & 3);
                  umma_ss(tmem_base + o_col(w), pd, vd, idesc_pv, (j | k) != 0);
                }
              }
              umma_commit(&o_done[w]);
              if (more) {
                if (w == 0) {
                  mbar_wait(&k_full[sn], phn);
                  tcgen05_fence_after();
                }
                issue_s(w, sn);
              }
            }
            umma_commit(&v_empty[sj]);
            if (more) umma_commit(&k_empty[sn]);
          }
          umma_commit(q_empty);
        }
      }
      __syncwarp();
    }
  } else {
    // ---------------------------------------------------------------- softmax / correction / epilogue
    setmaxnreg_inc<208>();
    const int w = (warp - 4) >> 2;              // query tile of this warpgroup
    const int wt = threadIdx.x - 128 - w * 128;  // row inside the tile
    const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const uint32_t s_addr = tmem_base + s_col(w) + lane_base;
    const uint32_t o_addr = tmem_base + o_col(w) + lane_base;
    const uint32_t stg = smem_u32(sO + w * ATT_TILE);
    const bool ragged = (a.Lkv & 127) != 0;
    uint32_t scount = 0, ocount = 0;  // completed phases of s_full[w] / o_done[w] consumed so far

    for (int item = blockIdx.x; item < a.total_items; item += gridDim.x) {
      const int per_head = a.batch * a.n_qpairs;
      const int h = item / per_head, rem = item % per_head;
      const int b = rem / a.n_qpairs, q0 = (rem % a.n_qpairs) * 256;
      if (w == 1 && q0 + 128 >= a.Lq) continue;

      float m = -INFINITY, l = 0.f;
      for (int j = 0; j < a.n_kv; ++j, ++scount) {
        mbar_wait(&s_full[w], scount & 1);
        tcgen05_fence_after();
        uint32_t v[4][32];
#pragma unroll
        for (int i = 0; i < 4; ++i) tmem_ld_x32(s_addr + 32 * i, v[i]);
        tmem_wait_ld();
        if (ragged && j == a.n_kv - 1) {
          const int valid = a.Lkv - j * 128;
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int c = 0; c < 32; ++c)
              if (i * 32 + c >= valid) v[i][c] = 0xff800000u;  // -inf
        }
        float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int c = 0; c < 32; ++c) mx[c & 3] = fmaxf(mx[c & 3], __uint_as_float(v[i][c]));
        const float m_new = fmaxf(fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])), m);
        if (j == 0) {
          m = m_new;
        } else {
          const bool need = (m_new - m) * a.scale_log2 > 8.0f;
          if (__any_sync(0xffffffffu, need)) {
            // O correction: wait for PV(j-1), then scale this row of O and l by 2^((m_old - m_new) * scale)
            mbar_wait(&o_done[w], (ocount + j - 1) & 1);
            tcgen05_fence_after();
            const float alpha = need ? ex2_approx((m - m_new) * a.scale_log2) : 1.0f;
            if (need) m = m_new;
            l *= alpha;
            uint32_t o[2][32];
            tmem_ld_x32(o_addr, o[0]);
            tmem_ld_x32(o_addr + 32, o[1]);
            tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < 2; ++i)
#pragma unroll
              for (int c = 0; c < 32; ++c) o[i][c] = __float_as_uint(__uint_as_float(o[i][c]) * alpha);
            tmem_st_x32(o_addr, o[0]);
            tmem_st_x32(o_addr + 32, o[1]);
            tmem_wait_st();
          }
        }
        const float ms = m * a.scale_log2;
        float sum[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint32_t pk[16];
#pragma unroll
          for (int c = 0; c < 32; c += 2) {
            const float p0 = ex2_approx(fmaf(__uint_as_float(v[i][c]), a.scale_log2, -ms));
            const float p1 = ex2_approx(fmaf(__uint_as_float(v[i][c + 1]), a.scale_log2, -ms));
            sum[(c >> 1) & 3] += p0 + p1;
            pk[c >> 1] = pack_bf16(p0, p1);
          }
          if (P_TMEM) {
            tmem_st_x16(s_addr + 16 * i, pk);
          } else {
            // K-major SW128 tile pair: columns [64*half, 64*half+64) of row wt; chunk = 8 bf16 = 16 bytes
            const uint32_t pbase = smem_u32(sP + (2 * w + (i >> 1)) * ATT_TILE);
#pragma unroll
            for (int c = 0; c < 4; ++c)
              st_shared_v4(pbase + sw128_offset(wt, (i & 1) * 4 + c), pk[4 * c], pk[4 * c + 1], pk[4 * c + 2],
                           pk[4 * c + 3]);
          }
        }
        l += (sum[0] + sum[1]) + (sum[2] + sum[3]);
        if (P_TMEM) {
          tmem_wait_st();
          tcgen05_fence_before();
        } else {
          fence_proxy_async_smem();
          tcgen05_fence_before();
        }
        mbar_arrive(&p_full[w]);
      }
      // ---- epilogue: O / l -> bf16 -> swizzled staging tile -> TMA store
      ocount += a.n_kv;
      mbar_wait(&o_done[w], (ocount - 1) & 1);
      tcgen05_fence_after();
      uint32_t o[2][32];
      tmem_ld_x32(o_addr, o[0]);
      tmem_ld_x32(o_addr + 32, o[1]);
      tmem_wait_ld();
      tcgen05_fence_before();
      if (wt == 0) tma_store_wait_read<0>();
      named_bar_sync(1 + w, 128);
      const float inv = 1.0f / l;
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        uint32_t pk[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int e = 8 * c + 2 * i;
          pk[i] = pack_bf16(__uint_as_float(o[e >> 5][e & 31]) * inv, __uint_as_float(o[e >> 5][(e & 31) + 1]) * inv);
        }
        st_shared_v4(stg + sw128_offset(wt, c), pk[0], pk[1], pk[2], pk[3]);
      }
      fence_proxy_async_smem();
      named_bar_sync(1 + w, 128);
      if (wt == 0) {
        tma_store_3d(&tmO, sO + w * ATT_TILE, h * 64, q0 + w * 128, b);
        tma_store_commit();
      }
    }
    if (wt == 0) tma_store_wait<0>();
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 2) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------------------
// Single-tile probe of the two operand forms the attention kernel relies on:
//   D[128,64] (fp32) = P[128,128] (bf16, K-major) x V[128,64] (bf16, rows = K, i.e. MN-major B operand)
// with P either staged in shared memory (mode 0) or written to TMEM by tcgen05.st (mode 1).
// Test infrastructure for descriptor conventions; not used by the forward pass.
__global__ void __launch_bounds__(128, 1)
umma_probe_kernel(const __grid_constant__ CUtensorMap tmP, const __grid_constant__ CUtensorMap tmV,
                  const __nv_bfloat16* __restrict__ p_global, float* __restrict__ d_out, int mode, uint32_t v_lbo,
                  uint32_t v_sbo) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sP = smem;                  // 2 tiles (K halves)
  uint8_t* sV = smem + 2 * ATT_TILE;   // 1 tile
  uint64_t* bar = reinterpret_cast<uint64_t*>(sV + ATT_TILE);
  uint64_t* done = bar + 1;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    mbar_init(done, 1);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(tmem_ptr, 256);
    tmem_relinquish();
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const uint32_t lane_base = static_cast<uint32_t>(warp * 32) << 16;
  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(bar, 3 * ATT_TILE);
    tma_load_2d(sP, &tmP, bar, 0, 0, kEvictNormal);
    tma_load_2d(sP + ATT_TILE, &tmP, bar, 64, 0, kEvictNormal);
    tma_load_2d(sV, &tmV, bar, 0, 0, kEvictNormal);
  }
  if (mode == 1) {
    // thread == row: pack the row of P (128 bf16 = 64 words) into TMEM columns [128, 192)
    const uint32_t* prow = reinterpret_cast<const uint32_t*>(p_global + (size_t)threadIdx.x * 128);
    uint32_t r0[32], r1[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      r0[i] = prow[i];
      r1[i] = prow[32 + i];
    }
    tmem_st_x32(tmem_base + 128 + lane_base, r0);
    tmem_st_x32(tmem_base + 160 + lane_base, r1);
    tmem_wait_st();
  }
  mbar_wait(bar, 0);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  if (threadIdx.x == 0) {
    constexpr uint32_t idesc = umma_idesc_bf16(128, 64, false, true);
    const uint32_t vbase = smem_u32(sV);
    for (int k = 0; k < 8; ++k) {
      const uint64_t vd = umma_desc_sw128(vbase + k * 2048, v_lbo, v_sbo);
      if (mode == 1) {
        umma_ts(tmem_base, tmem_base + 128 + k * 8, vd, idesc, k != 0);
      } else {
        const uint64_t pd = umma_desc_sw128(smem_u32(sP + (k >> 2) * ATT_TILE), 16, 1024) + 2 * (k & 3);
        umma_ss(tmem_base, pd, vd, idesc, k != 0);
      }
    }
    umma_commit(done);
  }
  mbar_wait(done, 0);
  tcgen05_fence_after();
  uint32_t o[2][32];
  tmem_ld_x32(tmem_base + lane_base, o[0]);
  tmem_ld_x32(tmem_base + lane_base + 32, o[1]);
  tmem_wait_ld();
#pragma unroll
  for (int i = 0; i < 64; ++i) d_out[threadIdx.x * 64 + i] = __uint_as_float(o[i >> 5][i & 31]);
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, 256);
  }
}

// ------------------------------------------------------------------------------------------------ host
wm_status flash_attention(cudaStream_t st, const void* q, int ldq, const void* k, const void* v, int ldkv, void* o,
                          int ldo, int batch, int Lq, int Lkv, int heads, float scale, int variant, uint32_t v_lbo,
                          uint32_t v_sbo) {
  static int sm_count = 0;
  if (sm_count == 0) WM_TRY(require_sm100(&sm_count));
  WM_CHECK(q && k && v && o, "flash_attention: null pointer");
  WM_CHECK(batch > 0 && Lq > 0 && Lkv > 0 && heads > 0, "flash_attention: batch, Lq, Lkv, heads must be positive");
  WM_CHECK(ldq >= heads * 64 && ldkv >= heads * 64 && ldo >= heads * 64 && ldq % 8 == 0 && ldkv % 8 == 0 && ldo % 8 == 0,
           "flash_attention: row pitches must be >= heads*64 and multiples of 8 elements");
  CUtensorMap tq, tk, tv, to;
  const CUtensorMapDataType bf = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  WM_TRY(tmap_3d(&tq, bf, 2, q, batch, Lq, heads * 64, ldq, (uint64_t)Lq * ldq, 64, 128));
  WM_TRY(tmap_3d(&tk, bf, 2, k, batch, Lkv, heads * 64, ldkv, (uint64_t)Lkv * ldkv, 64, 128));
  WM_TRY(tmap_3d(&tv, bf, 2, v, batch, Lkv, heads * 64, ldkv, (uint64_t)Lkv * ldkv, 64, 128));
  WM_TRY(tmap_3d(&to, bf, 2, o, batch, Lq, heads * 64, ldo, (uint64_t)Lq * ldo, 64, 128));
  AttnArgs a{};
  a.Lq = Lq, a.Lkv = Lkv, a.batch = batch;
  a.n_qpairs = (Lq + 255) / 256;
  a.n_kv = (Lkv + 127) / 128;
  a.total_items = heads * batch * a.n_qpairs;
  a.scale_log2 = scale * 1.4426950408889634f;
  a.v_lbo = v_lbo ? v_lbo : 1024;
  a.v_sbo = v_sbo ? v_sbo : 1024;
  const int grid = a.total_items < sm_count ? a.total_items : sm_count;
  if (variant == 0) {
    static bool cfg = false;
    if (!cfg) {
      WM_CUDA(cudaFuncSetAttribute(attn_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, AttnCfg<true>::SMEM));
      cfg = true;
    }
    attn_kernel<true><<<grid, ATT_THREADS, AttnCfg<true>::SMEM, st>>>(tq, tk, tv, to, a);
  } else {
    static bool cfg = false;
    if (!cfg) {
      WM_CUDA(cudaFuncSetAttribute(attn_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, AttnCfg<false>::SMEM));
      cfg = true;
    }
    attn_kernel<false><<<grid, ATT_THREADS, AttnCfg<false>::SMEM, st>>>(tq, tk, tv, to, a);
  }
  WM_LAUNCH_CHECK();
  return WM_OK;
}

wm_status umma_probe(cudaStream_t st, const void* p, const void* v, float* d, int mode, uint32_t v_lbo, uint32_t v_sbo) {
  int sm = 0;
  WM_TRY(require_sm100(&sm));
  CUtensorMap tp, tv;
  WM_TRY(tmap_2d(&tp, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, p, 128, 128, 128, 64, 128));
  WM_TRY(tmap_2d(&tv, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, v, 128, 64, 64, 64, 128));
  const int smem = 3 * ATT_TILE + 64 + 1024;
  WM_CUDA(cudaFuncSetAttribute(umma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  umma_probe_kernel<<<1, 128, smem, st>>>(tp, tv, static_cast<const __nv_bfloat16*>(p), d, mode, v_lbo, v_sbo);
  WM_LAUNCH_CHECK();
  return WM_OK;
}

}  // namespace wm
