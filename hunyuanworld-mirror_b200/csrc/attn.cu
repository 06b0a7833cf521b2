// Flash attention for head_dim 64 on tcgen05/TMEM, fed by TMA (F.scaled_dot_product_attention at
// attention.py:59; no mask, no dropout).  One kernel serves both call shapes of the aggregator:
//   frame  attention: batch = views, Lq = Lkv = tokens per view       (visual_transformer.py:313-323)
//   global attention: batch = 1,     Lq = local tokens, Lkv = all tokens (visual_transformer.py:324-334;
//                     under sequence parallelism Lkv spans the all-gathered K|V of every rank)
// q, k, v, o are bf16 with heads side by side in the row (q:[.,1024], k|v:[.,2048], o:[.,1024]) which is
// exactly what the QKV GEMM writes and the proj GEMM reads, so no permute/transposed copy exists.
//
// Persistent CTAs (one per SM, 384 threads).  A work item is (head, batch, 256 query rows):
//   warp 0      TMA producer: Q pair (double-buffered across items), K/V tiles through a 3-stage ring
//   warp 1      MMA issuer (one lane): S_w = Q_w K^T (128x128x64), O_w += P_w V (128x64x128, P from TMEM)
//   warp 2      TMEM allocator
//   warps 4-7   softmax / correction / epilogue for query tile 0 (thread == query row)
//   warps 8-11  the same for query tile 1
// TMEM (512 columns): S0 [0,128) S1 [128,256) O0 [256,320) O1 [320,384) P0 [384,448) P1 [448,512).
// P has its own columns, so the softmax warps release S as soon as it is in registers (s_free) and the MMA
// warp issues S(j+1) while softmax(j) is still running: a warpgroup goes from one tile's softmax straight
// into the next, and the next work item's first S overlaps the current item's epilogue.
// Online softmax keeps a running max that is only advanced (and O rescaled) when it grew by more than 2^8,
// so the O correction is rare; row sum and the final 1/l normalisation stay in fp32.  exp2 runs on the MUFU
// pipe for most elements and, for NPOLY of every 32, as a Cody-Waite split + degree-3 polynomial on the FMA
// pipe (max relative error 7.6e-5, far below the bf16 rounding of P): at head_dim 64 the kernel is bound by
// exponentials, not by the tensor core (SURVEY.md section 7 "hard parts").
#include "attn_common.cuh"

namespace wm {

template <int NPOLY>
__global__ void __launch_bounds__(ATT_THREADS, 1)
attn_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
            const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO, const AttnArgs a) {
  constexpr int ST = KV_ST;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sQ = smem;                       // 2 buffers x 2 tiles
  uint8_t* sK = sQ + 4 * ATT_TILE;          // ST tiles
  uint8_t* sV = sK + ST * ATT_TILE;         // ST tiles
  uint8_t* sO = sV + ST * ATT_TILE;         // 2 tiles (epilogue staging, one per query tile)
  uint64_t* bars = reinterpret_cast<uint64_t*>(sO + 2 * ATT_TILE);
  uint64_t* q_full = bars;             // 2
  uint64_t* q_empty = q_full + 2;      // 2
  uint64_t* k_full = q_empty + 2;      // ST
  uint64_t* k_empty = k_full + ST;     // ST
  uint64_t* v_full = k_empty + ST;     // ST
  uint64_t* v_empty = v_full + ST;     // ST
  uint64_t* s_full = v_empty + ST;     // 2  MMA -> softmax: S_w written
  uint64_t* s_free = s_full + 2;       // 2  softmax -> MMA: S_w is in registers
  uint64_t* p_full = s_free + 2;       // 2  softmax -> MMA: P_w written
  uint64_t* pv_done = p_full + 2;      // 2  MMA -> softmax: O_w += P_w V complete (P_w reusable, O_w readable)
  uint64_t* o_free = pv_done + 2;      // 2  epilogue -> MMA: O_w has been read out
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(o_free + 2);

  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5;

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmO);
  }
  if (warp == 1 && elect_one()) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&q_full[i], 1);
      mbar_init(&q_empty[i], 1);
      mbar_init(&s_full[i], 1);
      mbar_init(&s_free[i], 128);
      mbar_init(&p_full[i], 128);
      mbar_init(&pv_done[i], 1);
      mbar_init(&o_free[i], 128);
    }
    for (int i = 0; i < ST; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&v_empty[i], 1);
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
  pdl_wait();

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
          const uint32_t qb = icount & 1, qph = (icount >> 1) & 1;
          mbar_wait(&q_empty[qb], qph ^ 1);
          mbar_arrive_expect_tx(&q_full[qb], 2 * ATT_TILE);
          tma_load_3d(sQ + (2 * qb) * ATT_TILE, &tmQ, &q_full[qb], h * 64, q0, b, kEvictFirst);
          tma_load_3d(sQ + (2 * qb + 1) * ATT_TILE, &tmQ, &q_full[qb], h * 64, q0 + 128, b, kEvictFirst);
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
        uint32_t tiles0 = 0, tiles1 = 0;  // S tiles issued per warpgroup (s_free phase bookkeeping)
        uint32_t pvs0 = 0, pvs1 = 0;      // PVs issued per warpgroup (p_full phases consumed)
        uint32_t items0 = 0, items1 = 0;  // work items started per warpgroup (o_free phases)
        for (int item = blockIdx.x; item < a.total_items; item += gridDim.x, ++icount) {
          const int q0 = (item % a.n_qpairs) * 256;
          const int nw = (q0 + 128 < a.Lq) ? 2 : 1;  // second query tile entirely out of range: skip it
          const uint32_t qb = icount & 1, qph = (icount >> 1) & 1;
          mbar_wait(&q_full[qb], qph);
          tcgen05_fence_after();
          auto issue_s = [&](int w, uint32_t stage) {
            uint32_t& tiles = w == 0 ? tiles0 : tiles1;
            mbar_wait(&s_free[w], (tiles & 1) ^ 1);  // softmax has read the previous S_w into registers
            ++tiles;
            tcgen05_fence_after();
            const uint64_t qd = umma_desc_sw128(smem_u32(sQ + (2 * qb + w) * ATT_TILE), 16, 1024);
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
            if (j + 1 < a.n_kv) {
              const uint32_t sn = (kvit + 1) % ST, phn = ((kvit + 1) / ST) & 1;
              mbar_wait(&k_full[sn], phn);
              tcgen05_fence_after();
              for (int w = 0; w < nw; ++w) issue_s(w, sn);
              umma_commit(&k_empty[sn]);
            }
            mbar_wait(&v_full[sj], phj);
            tcgen05_fence_after();
            for (int w = 0; w < nw; ++w) {
              uint32_t& pvs = w == 0 ? pvs0 : pvs1;
              mbar_wait(&p_full[w], pvs & 1);
              ++pvs;
              if (j == 0) mbar_wait(&o_free[w], ((w == 0 ? items0 : items1) & 1) ^ 1);  // previous item's O_w read out
              tcgen05_fence_after();
              const uint32_t vbase = smem_u32(sV + sj * ATT_TILE);
#pragma unroll
              for (int k = 0; k < 8; ++k) {  // 16 kv rows (2048 B of the V tile) per step; P: 8 TMEM columns per step
                const uint64_t vd = umma_desc_sw128(vbase + k * 2048, 1024, 1024);
                umma_ts(tmem_base + o_col(w), tmem_base + p_col(w) + k * 8, vd, idesc_pv, (j | k) != 0);
              }
              umma_commit(&pv_done[w]);
            }
            umma_commit(&v_empty[sj]);
          }
          umma_commit(&q_empty[qb]);
          ++items0;
          if (nw == 2) ++items1;
        }
      }
      __syncwarp();
    }
  } else {
    // ---------------------------------------------------------------- softmax / correction / epilogue
    setmaxnreg_inc<208>();
    const int w = (warp - 4) >> 2;               // query tile of this warpgroup
    const int wt = (threadIdx.x - 128) & 127;    // row inside the tile
    const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const uint32_t s_addr = tmem_base + s_col(w) + lane_base;
    const uint32_t o_addr = tmem_base + o_col(w) + lane_base;
    const uint32_t p_addr = tmem_base + p_col(w) + lane_base;
    const uint32_t stg = smem_u32(sO + w * ATT_TILE);
    const bool ragged = (a.Lkv & 127) != 0;
    const float2 scale2 = make_float2(a.scale_log2, a.scale_log2);
    uint32_t scount = 0, pvbase = 0;  // s_full[w] phases consumed / pv_done[w] phases of finished items

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
        tcgen05_fence_before();
        mbar_arrive(&s_free[w]);  // the MMA warp may overwrite S_w with the next tile's scores
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
        bool pv_waited = false;
        if (j == 0) {
          m = m_new;
        } else {
          const bool need = (m_new - m) * a.scale_log2 > 8.0f;
          if (__any_sync(0xffffffffu, need)) {
            // O correction: wait for PV(j-1), then scale this row of O and l by 2^((m_old - m_new) * scale)
            mbar_wait(&pv_done[w], (pvbase + j - 1) & 1);
            pv_waited = true;
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
        const float2 nms2 = make_float2(-ms, -ms);
        float2 sum2[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
        uint32_t pk[2][32];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
#pragma unroll
          for (int c = 0; c < 32; c += 2) {
            const float2 xs = ffma2(make_float2(__uint_as_float(v[i][c]), __uint_as_float(v[i][c + 1])), scale2, nms2);
            float2 p;
            if (c >= 32 - NPOLY) {
              p = exp2_poly2(xs);
            } else {
              p.x = ex2_approx(xs.x);
              p.y = ex2_approx(xs.y);
            }
            sum2[(c >> 1) & 1] = fadd2(sum2[(c >> 1) & 1], p);
            pk[i >> 1][(i & 1) * 16 + (c >> 1)] = pack_bf16(p.x, p.y);
          }
        }
        l += (sum2[0].x + sum2[0].y) + (sum2[1].x + sum2[1].y);
        if (j > 0 && !pv_waited) {  // P_w is still being read by PV(j-1) until this completes (normally long done)
          mbar_wait(&pv_done[w], (pvbase + j - 1) & 1);
          tcgen05_fence_after();
        }
        tmem_st_x32(p_addr, pk[0]);
        tmem_st_x32(p_addr + 32, pk[1]);
        tmem_wait_st();
        tcgen05_fence_before();
        mbar_arrive(&p_full[w]);
      }
      // ---- epilogue: O / l -> bf16 -> swizzled staging tile -> TMA store
      pvbase += a.n_kv;
      mbar_wait(&pv_done[w], (pvbase - 1) & 1);
      tcgen05_fence_after();
      uint32_t o[2][32];
      tmem_ld_x32(o_addr, o[0]);
      tmem_ld_x32(o_addr + 32, o[1]);
      tmem_wait_ld();
      tcgen05_fence_before();
      mbar_arrive(&o_free[w]);  // the next item's first PV may overwrite O_w
      if (wt == 0) tma_store_wait_read<0>();
      named_bar_sync(1 + w, 128);
      const float inv = 1.0f / l;
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        uint32_t pk4[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int e = 8 * c + 2 * i;
          pk4[i] = pack_bf16(__uint_as_float(o[e >> 5][e & 31]) * inv, __uint_as_float(o[e >> 5][(e & 31) + 1]) * inv);
        }
        st_shared_v4(stg + sw128_offset(wt, c), pk4[0], pk4[1], pk4[2], pk4[3]);
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
wm_status flash_attention_split(cudaStream_t st, int sm_count, const CUtensorMap& tq, const CUtensorMap& tk,
                                const CUtensorMap& tv, const CUtensorMap& to, AttnArgs a, int variant, void* out, int ldo,
                                float* partial_ws, size_t partial_ws_bytes, const AttnGather* gather,
                                const AttnGuard* guard);  // attn_split.cu

template <int NPOLY>
static wm_status launch_attn(cudaStream_t st, int grid, const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv,
                             const CUtensorMap& to, const AttnArgs& a) {
  static bool cfg = false;
  if (!cfg) {
    WM_CUDA(cudaFuncSetAttribute(attn_kernel<NPOLY>, cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_SMEM));
    cfg = true;
  }
  return launch_pdl(attn_kernel<NPOLY>, dim3(grid), dim3(ATT_THREADS), ATT_SMEM, st, tq, tk, tv, to, a);
}

// variant: how many of every 32 exponentials run as a polynomial on the FMA pipe (0 -> 0, 1 -> 8, 2 -> 12, 3 -> 16);
// kDefaultVariant is what the forward pass uses.
constexpr int kDefaultVariant = 2;

wm_status flash_attention(cudaStream_t st, const void* q, int ldq, const void* k, const void* v, int ldkv, void* o,
                          int ldo, int batch, int Lq, int Lkv, int heads, float scale, int variant, float* partial_ws,
                          size_t partial_ws_bytes, const AttnGather* gather, const AttnGuard* guard) {
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
  const int grid = a.total_items < sm_count ? a.total_items : sm_count;
  if (variant < 0) variant = kDefaultVariant;
  if (variant >= 4)
    return flash_attention_split(st, sm_count, tq, tk, tv, to, a, variant, o, ldo, partial_ws, partial_ws_bytes, gather, guard);
  WM_CHECK(gather == nullptr && guard == nullptr, "flash_attention: an in-flight gather or a guard needs a split-row variant (4..12)");
  switch (variant) {
    case 0: return launch_attn<0>(st, grid, tq, tk, tv, to, a);
    case 1: return launch_attn<8>(st, grid, tq, tk, tv, to, a);
    case 2: return launch_attn<12>(st, grid, tq, tk, tv, to, a);
    case 3: return launch_attn<16>(st, grid, tq, tk, tv, to, a);
    default: return fail(WM_ERR_INVALID, "flash_attention: variant must be -1 (default) or 0..12");
  }
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
