// Flash attention, head_dim 64, split-row softmax (thread == query row x key half) -- second generation of
// attn_split.cu.  Same tile pipeline (TMA-fed K/V ring, tcgen05.mma with S, O and P in TMEM, two 128-row query tiles
// per CTA, 16 softmax warps, work scheduler with split tails, sequence-parallel arrival words); what changed is WHO
// limits it.  ncu of attn_split_kernel<12, true> (profiles/r02_attn_split_v10_ncu.txt): the softmax warps sleep on
// s_full while the tensor pipe is 47 % busy and MUFU 59 % -- because the ONE thread that issues the MMAs executes ~500
// instructions per key tile (descriptor arithmetic in vector registers, an R2UR per operand, rolled loops over the
// query tiles, the 9-instruction watchdog loop around every mbarrier wait), shares its scheduler with four softmax
// warps and, having the LOWEST warp id of the five, loses the arbitration to them: 75 % of its samples are plain
// instruction issue, not waiting.  The pipeline ran at the speed of that instruction stream (~2100 clk per tile).
//
//   * control warps LAST (warps 16-19; softmax warps 0-15): the issue arbiter prefers the highest warp id, so the MMA
//     warp wins its scheduler whenever it is ready;
//   * TWO issuing threads in two warps: one issues QK^T (waits: K stage, S buffers free), the other P V (waits: V stage,
//     P written, O free).  An mbarrier try_wait costs ~90 clk even when the phase has long completed, and the six of
//     them per tile were serialised in one thread; tcgen05.mma from different threads needs no mutual order here
//     (different TMEM regions, every dependency is an mbarrier);
//   * a leaner instruction stream: descriptors are base + stage * 1024 (one add), the loops over query tiles are
//     unrolled, waits take a one-instruction fast path before the watchdog loop;
//   * HALVES = true additionally issues every 128-key tile as two 64-key halves with their own barriers (QK^T N = 64,
//     PV K = 64): a softmax thread loads its 32 scores of a half with ONE tcgen05.ld, frees the half at once and
//     hands 16 packed columns of P to PV before it touches the other half -- at most 32 scores + 16 packed
//     probabilities live, and QK^T of the next tile is issued a whole tile ahead.
//
// BOUNDED (logits bounded by the QK-LayerNorm weights, q pre-multiplied by scale * log2 e, P = 2^s with no running
// maximum) and the general running-max softmax are both here; NPOLY of every 32 exponentials run as the degree-3
// polynomial on the FMA pipe.
#include "attn_common.cuh"

namespace wm {

constexpr int ATP_THREADS = 640;
constexpr int ATP_XCH = 2 * 2 * 2 * 128 * 4 + 2 * 2 * 128 * 4;  // half-row maxima (double-buffered) + partial row sums
constexpr int ATP_SMEM = (4 + 2 * KV_ST + 2) * ATT_TILE + 512 + ATP_XCH + 1024;

// mbarrier wait whose common case (the phase has already completed) is one try_wait and one branch
__device__ __forceinline__ void mbar_wait_fast(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  mbar_wait(bar, parity);
}

// shared-memory descriptor of a SWIZZLE_128B operand whose address field is filled in later (addr >> 4, one add)
__host__ __device__ constexpr uint64_t umma_desc_base(uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFF) << 16) | (static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFF) << 32) |
         (1ull << 46) | (2ull << 61);
}

template <int NPOLY, bool BOUNDED, bool HALVES>
__global__ void __launch_bounds__(ATP_THREADS, 1)
attn_pipe_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                 const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO, const AttnArgs a) {
  static_assert(!HALVES || BOUNDED, "the half-tile pipeline is built for the bounded-logit softmax");
  constexpr int ST = KV_ST;
  constexpr int NSUB = HALVES ? 2 : 1;  // independent sub-tiles per 128-key tile
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
  uint64_t* s_full = v_empty + ST;     // [2 w + s]  MMA -> softmax: (half s of) S_w written
  uint64_t* s_free = s_full + 4;       // [2 w + s]  softmax -> MMA: it is in registers (8 warp arrivals)
  uint64_t* p_full = s_free + 4;       // [2 w + s]  softmax -> MMA: (half s of) P_w written (8 warp arrivals)
  uint64_t* pv_done = p_full + 4;      // [2 w + s]  MMA -> softmax: O_w += P V complete
  uint64_t* o_free = pv_done + 4;      // 2          epilogue -> MMA: O_w has been read out (8 warp arrivals)
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(o_free + 2);
  float* xmax = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 512);  // [qt][parity][half][128]
  float* xsum = xmax + 2 * 2 * 2 * 128;                                            // [qt][half][128]

  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 16 && elect_one()) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmO);
  }
  if (warp == 17 && elect_one()) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&q_full[i], 1);
      mbar_init(&q_empty[i], 1);
      mbar_init(&o_free[i], 8);
    }
    for (int i = 0; i < 4; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&s_free[i], 8);  // one arrival per softmax warp (lane 0 after __syncwarp)
      mbar_init(&p_full[i], 8);
      mbar_init(&pv_done[i], 1);
    }
    for (int i = 0; i < ST; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&v_empty[i], 1);
    }
    fence_barrier_init();
  }
  if (warp == 19) {
    tmem_alloc(tmem_ptr, 512);
    tmem_relinquish();
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  pdl_wait();  // q, k, v of the preceding GEMM (or all-gather) are complete and visible from here on

  if (warp >= 16) {
    setmaxnreg_dec<64>();
    if (warp == 16) {
      // -------------------------------------------------------------- TMA producer
      if (elect_one()) {
        uint32_t kvit = 0, icount = 0;
        unsigned have = a.shard_rows > 0 ? (1u << a.self_rank) : 0xffffffffu;  // shards known to be present
        AttnSched sched(a);
        for (int item, kv0, kv1, slot; sched.next(a, item, kv0, kv1, slot); ++icount) {
          const int per_head = a.batch * a.n_qpairs;
          const int h = item / per_head, rem = item % per_head;
          const int b = rem / a.n_qpairs, q0 = (rem % a.n_qpairs) * 256;
          const uint32_t qb = icount & 1, qph = (icount >> 1) & 1;
          mbar_wait(&q_empty[qb], qph ^ 1);
          mbar_arrive_expect_tx(&q_full[qb], 2 * ATT_TILE);
          tma_load_3d(sQ + (2 * qb) * ATT_TILE, &tmQ, &q_full[qb], h * 64, q0, b, kEvictFirst);
          tma_load_3d(sQ + (2 * qb + 1) * ATT_TILE, &tmQ, &q_full[qb], h * 64, q0 + 128, b, kEvictFirst);
          for (int j = kv0; j < kv1; ++j, ++kvit) {
            const uint32_t s = kvit % ST, ph = (kvit / ST) & 1;
            const int jp = attn_phys_tile(a, j);
            if (have != 0xffffffffu) attn_wait_shards(a, jp, have);
            mbar_wait(&k_empty[s], ph ^ 1);
            mbar_arrive_expect_tx(&k_full[s], ATT_TILE);
            tma_load_3d(sK + s * ATT_TILE, &tmK, &k_full[s], h * 64, jp * 128, b, kEvictLast);
            mbar_wait(&v_empty[s], ph ^ 1);
            mbar_arrive_expect_tx(&v_full[s], ATT_TILE);
            tma_load_3d(sV + s * ATT_TILE, &tmV, &v_full[s], h * 64, jp * 128, b, kEvictLast);
          }
        }
      }
      __syncwarp();
    } else if (warp == 17) {
      // -------------------------------------------------------------- QK^T issuer (one elected thread)
      if (elect_one()) {
        constexpr uint32_t idesc_s = umma_idesc_bf16(128, HALVES ? 64 : 128, false, false);
        constexpr uint64_t qk_base = umma_desc_base(16, 1024);
        constexpr uint32_t TILE16 = ATT_TILE >> 4;  // one 16 KB tile in descriptor address units
        const uint32_t q_lo = (smem_u32(sQ) & 0x3FFFF) >> 4, k_lo = (smem_u32(sK) & 0x3FFFF) >> 4;
        uint32_t kvit = 0, icount = 0;
        uint32_t tiles0 = 0, tiles1 = 0;  // score tiles issued per query tile: s_free phases consumed
        AttnSched sched(a);
        for (int item, kv0, kv1, slot; sched.next(a, item, kv0, kv1, slot); ++icount) {
          const int q0 = (item % a.n_qpairs) * 256;
          const bool two = q0 + 128 < a.Lq;  // second query tile entirely out of range: skip it
          const uint32_t qb = icount & 1, qph = (icount >> 1) & 1;
          mbar_wait_fast(&q_full[qb], qph);
          const uint64_t qd0 = qk_base | (q_lo + (2 * qb) * TILE16), qd1 = qd0 + TILE16;
          for (int j = kv0; j < kv1; ++j, ++kvit) {
            const uint32_t st = kvit % ST, ph = (kvit / ST) & 1;
            mbar_wait_fast(&k_full[st], ph);
            const uint64_t kd = qk_base | (k_lo + st * TILE16);
#pragma unroll
            for (int s = 0; s < NSUB; ++s) {
              // both row halves hold the previous scores of this (half) tile in registers
              mbar_wait_fast(&s_free[s], (tiles0 & 1) ^ 1);
              tcgen05_fence_after();
#pragma unroll
              for (int k = 0; k < 4; ++k)
                umma_ss(tmem_base + s_col(0) + 64 * s, qd0 + 2 * k, kd + s * (TILE16 / 2) + 2 * k, idesc_s, k != 0);
              umma_commit(&s_full[s]);
              if (two) {
                mbar_wait_fast(&s_free[2 + s], (tiles1 & 1) ^ 1);
                tcgen05_fence_after();
#pragma unroll
                for (int k = 0; k < 4; ++k)
                  umma_ss(tmem_base + s_col(1) + 64 * s, qd1 + 2 * k, kd + s * (TILE16 / 2) + 2 * k, idesc_s, k != 0);
                umma_commit(&s_full[2 + s]);
              }
            }
            ++tiles0;
            if (two) ++tiles1;
            umma_commit(&k_empty[st]);
          }
          umma_commit(&q_empty[qb]);
        }
      }
      __syncwarp();
    } else if (warp == 18) {
      // -------------------------------------------------------------- PV issuer (one elected thread): its own warp, so
      // that the two chains of mbarrier waits (~90 clk each even when complete) do not add up in one thread
      if (elect_one()) {
        constexpr uint32_t idesc_pv = umma_idesc_bf16(128, 64, false, true);  // B = V is MN-major (dims contiguous)
        constexpr uint64_t v_base = umma_desc_base(1024, 1024);
        constexpr uint32_t TILE16 = ATT_TILE >> 4;
        const uint32_t v_lo = (smem_u32(sV) & 0x3FFFF) >> 4;
        uint32_t kvit = 0;
        uint32_t pvs0 = 0, pvs1 = 0;      // P tiles consumed per query tile: p_full phases consumed
        uint32_t items0 = 0, items1 = 0;  // work items started per query tile: o_free phases
        AttnSched sched(a);
        for (int item, kv0, kv1, slot; sched.next(a, item, kv0, kv1, slot);) {
          const int q0 = (item % a.n_qpairs) * 256;
          const bool two = q0 + 128 < a.Lq;
          for (int j = kv0; j < kv1; ++j, ++kvit) {
            const uint32_t st = kvit % ST, ph = (kvit / ST) & 1;
            const bool first = j == kv0;
            mbar_wait_fast(&v_full[st], ph);
            const uint64_t vd = v_base | (v_lo + st * TILE16);
#pragma unroll
            for (int s = 0; s < NSUB; ++s) {
              mbar_wait_fast(&p_full[s], pvs0 & 1);
              if (first && s == 0) mbar_wait_fast(&o_free[0], (items0 & 1) ^ 1);  // the previous segment's O_0 has been read out
              tcgen05_fence_after();
              umma_ts(tmem_base + o_col(0), tmem_base + p_col(0) + 32 * s, vd + s * (TILE16 / 2), idesc_pv, (first && s == 0) ? 0u : 1u);
#pragma unroll
              for (int k = 1; k < 8 / NSUB; ++k)  // 16 keys (2048 B of the V tile, 8 packed columns of P) per step
                umma_ts(tmem_base + o_col(0), tmem_base + p_col(0) + 32 * s + 8 * k, vd + s * (TILE16 / 2) + 128 * k, idesc_pv, 1);
              umma_commit(&pv_done[s]);
              if (two) {
                mbar_wait_fast(&p_full[2 + s], pvs1 & 1);
                if (first && s == 0) mbar_wait_fast(&o_free[1], (items1 & 1) ^ 1);
                tcgen05_fence_after();
                umma_ts(tmem_base + o_col(1), tmem_base + p_col(1) + 32 * s, vd + s * (TILE16 / 2), idesc_pv, (first && s == 0) ? 0u : 1u);
#pragma unroll
                for (int k = 1; k < 8 / NSUB; ++k)
                  umma_ts(tmem_base + o_col(1), tmem_base + p_col(1) + 32 * s + 8 * k, vd + s * (TILE16 / 2) + 128 * k, idesc_pv, 1);
                umma_commit(&pv_done[2 + s]);
              }
            }
            ++pvs0;
            if (two) ++pvs1;
            umma_commit(&v_empty[st]);
          }
          ++items0;
          if (two) ++items1;
        }
      }
      __syncwarp();
    }
  } else {
    // ---------------------------------------------------------------- softmax / correction / epilogue (warps 0-15)
    setmaxnreg_inc<104>();
    const int grp = warp >> 2;
    const int qt = grp >> 1;                     // query tile of this warpgroup
    const int hf = grp & 1;                      // tile mode: keys [64 hf, +64) of a tile; HALVES: keys [64 s + 32 hf, +32)
    const int row = (warp & 3) * 32 + lane;
    const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const uint32_t s_addr = tmem_base + s_col(qt) + (HALVES ? 32u : 64u) * hf + lane_base;  // HALVES: + 64 s
    const uint32_t o_addr = tmem_base + o_col(qt) + 32u * hf + lane_base;  // this thread rescales / reads 32 of the 64 O columns
    const uint32_t p_addr = tmem_base + p_col(qt) + (HALVES ? 16u : 32u) * hf + lane_base;  // HALVES: + 32 s
    const uint32_t stg = smem_u32(sO + qt * ATT_TILE);
    const bool leader = hf == 0 && row == 0;
    const bool ragged = (a.Lkv & 127) != 0;
    const float2 scale2 = make_float2(a.scale_log2, a.scale_log2);
    float* my_max = xmax + (qt * 4 + hf) * 128 + row;          // + parity * 256
    const float* peer_max = xmax + (qt * 4 + (hf ^ 1)) * 128 + row;
    uint32_t scount = 0, pvbase = 0;  // s_full phases consumed / pv_done phases of finished items

    AttnSched sched(a);
    for (int item, kv0, kv1, slot; sched.next(a, item, kv0, kv1, slot);) {
      const int per_head = a.batch * a.n_qpairs;
      const int h = item / per_head, rem = item % per_head;
      const int b = rem / a.n_qpairs, q0 = (rem % a.n_qpairs) * 256;
      if (qt == 1 && q0 + 128 >= a.Lq) continue;

      float m = BOUNDED ? 0.f : -INFINITY, l = 0.f;
      for (int j = kv0; j < kv1; ++j, ++scount) {
        const bool last_ragged = ragged && attn_phys_tile(a, j) == a.n_kv - 1;
        if constexpr (HALVES) {
#pragma unroll
          for (int s = 0; s < 2; ++s) {
            mbar_wait(&s_full[2 * qt + s], scount & 1);
            tcgen05_fence_after();
            uint32_t v[32];
            tmem_ld_x32(s_addr + 64 * s, v);
            tmem_wait_ld();
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&s_free[2 * qt + s]);  // the next tile's QK_s may overwrite this half
            if (last_ragged) {
              const int valid = a.Lkv - (a.n_kv - 1) * 128 - 64 * s - 32 * hf;
#pragma unroll
              for (int c = 0; c < 32; ++c)
                if (c >= valid) v[c] = 0xc2f00000u;  // -120: P = 2^-120, >= 2^40 below any real term (logit bound <= 80)
            }
            uint32_t pk[16];
            float2 sum2[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
#pragma unroll
            for (int c = 0; c < 32; c += 2) {
              const float2 xs = make_float2(__uint_as_float(v[c]), __uint_as_float(v[c + 1]));
              float2 p;
              if (c >= 32 - NPOLY) {
                p = exp2_poly2_bounded(xs);
              } else {
                p.x = ex2_approx(xs.x);
                p.y = ex2_approx(xs.y);
              }
              sum2[(c >> 1) & 1] = fadd2(sum2[(c >> 1) & 1], p);
              pk[c >> 1] = pack_bf16(p.x, p.y);
            }
            l += (sum2[0].x + sum2[0].y) + (sum2[1].x + sum2[1].y);
            if (j > kv0) {  // PV_s of the previous tile reads this half of P_qt until it completes (normally long done)
              mbar_wait(&pv_done[2 * qt + s], (pvbase + j - kv0 - 1) & 1);
              tcgen05_fence_after();
            }
            tmem_st_x16(p_addr + 32 * s, pk);
            tmem_wait_st();
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&p_full[2 * qt + s]);
          }
        } else {
          mbar_wait(&s_full[2 * qt], scount & 1);
          tcgen05_fence_after();
          uint32_t pk[32];
          bool pv_waited = false;
          if constexpr (BOUNDED) {
            // two 32-column chunks, one after the other: at most 32 scores + 16 packed probabilities are live
            float2 sum2[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
#pragma unroll
            for (int i = 0; i < 2; ++i) {
              uint32_t v[32];
              tmem_ld_x32(s_addr + 32 * i, v);
              tmem_wait_ld();
              if (i == 1) {
                tcgen05_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&s_free[2 * qt]);  // the MMA warp may overwrite S_qt once both halves have arrived
              }
              if (last_ragged) {
                const int valid = a.Lkv - (a.n_kv - 1) * 128 - hf * 64 - i * 32;
#pragma unroll
                for (int c = 0; c < 32; ++c)
                  if (c >= valid) v[c] = 0xc2f00000u;  // -120
              }
#pragma unroll
              for (int c = 0; c < 32; c += 2) {
                const float2 xs = make_float2(__uint_as_float(v[c]), __uint_as_float(v[c + 1]));
                float2 p;
                if (c >= 32 - NPOLY) {
                  p = exp2_poly2_bounded(xs);
                } else {
                  p.x = ex2_approx(xs.x);
                  p.y = ex2_approx(xs.y);
                }
                sum2[(c >> 1) & 1] = fadd2(sum2[(c >> 1) & 1], p);
                pk[i * 16 + (c >> 1)] = pack_bf16(p.x, p.y);
              }
            }
            l += (sum2[0].x + sum2[0].y) + (sum2[1].x + sum2[1].y);
          } else {
            // pass 1: the maximum of this thread's 64 scores, 32 at a time (nothing is kept: TMEM is read again below,
            // which costs two more tcgen05.ld per tile and keeps the live set at 32 scores + 32 packed probabilities)
            const int valid = last_ragged ? a.Lkv - (a.n_kv - 1) * 128 - hf * 64 : 64;
            float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
            for (int i = 0; i < 2; ++i) {
              uint32_t v[32];
              tmem_ld_x32(s_addr + 32 * i, v);
              tmem_wait_ld();
#pragma unroll
              for (int c = 0; c < 32; ++c)
                mx[c & 3] = fmaxf(mx[c & 3], (last_ragged && i * 32 + c >= valid) ? -INFINITY : __uint_as_float(v[c]));
            }
            const float mt = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
            my_max[(j & 1) * 256] = mt;
            named_bar_sync(1 + qt, 256);
            const float m_new = fmaxf(fmaxf(mt, peer_max[(j & 1) * 256]), m);
            if (j == kv0) {
              m = m_new;
            } else {
              const bool need = (m_new - m) * a.scale_log2 > 8.0f;
              if (__any_sync(0xffffffffu, need)) {
                // O correction: wait for PV(j-1), then scale this thread's 32 columns of O and its partial l
                mbar_wait(&pv_done[2 * qt], (pvbase + j - kv0 - 1) & 1);
                pv_waited = true;
                tcgen05_fence_after();
                const float alpha = need ? ex2_approx((m - m_new) * a.scale_log2) : 1.0f;
                if (need) m = m_new;
                l *= alpha;
                uint32_t o[32];
                tmem_ld_x32(o_addr, o);
                tmem_wait_ld();
#pragma unroll
                for (int c = 0; c < 32; ++c) o[c] = __float_as_uint(__uint_as_float(o[c]) * alpha);
                tmem_st_x32(o_addr, o);
                tmem_wait_st();
              }
            }
            // pass 2: P = 2^(s * scale - m * scale)
            const float ms = m * a.scale_log2;
            const float2 nms2 = make_float2(-ms, -ms);
            float2 sum2[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
#pragma unroll
            for (int i = 0; i < 2; ++i) {
              uint32_t v[32];
              tmem_ld_x32(s_addr + 32 * i, v);
              tmem_wait_ld();
              if (i == 1) {
                tcgen05_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&s_free[2 * qt]);  // the MMA warp may overwrite S_qt once both halves have arrived
              }
              if (last_ragged) {
#pragma unroll
                for (int c = 0; c < 32; ++c)
                  if (i * 32 + c >= valid) v[c] = 0xff800000u;  // -inf
              }
#pragma unroll
              for (int c = 0; c < 32; c += 2) {
                const float2 xs = ffma2(make_float2(__uint_as_float(v[c]), __uint_as_float(v[c + 1])), scale2, nms2);
                float2 p;
                if (c >= 32 - NPOLY) {
                  p = exp2_poly2(xs);
                } else {
                  p.x = ex2_approx(xs.x);
                  p.y = ex2_approx(xs.y);
                }
                sum2[(c >> 1) & 1] = fadd2(sum2[(c >> 1) & 1], p);
                pk[i * 16 + (c >> 1)] = pack_bf16(p.x, p.y);
              }
            }
            l += (sum2[0].x + sum2[0].y) + (sum2[1].x + sum2[1].y);
          }
          if (j > kv0 && !pv_waited) {  // P_qt is still being read by PV(j-1) until this completes (normally long done)
            mbar_wait(&pv_done[2 * qt], (pvbase + j - kv0 - 1) & 1);
            tcgen05_fence_after();
          }
          tmem_st_x32(p_addr, pk);
          tmem_wait_st();
          tcgen05_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&p_full[2 * qt]);
        }
      }
      // ---- epilogue: merge the partial row sums, O / l -> bf16 -> swizzled staging tile -> TMA store
      pvbase += kv1 - kv0;
      mbar_wait(&pv_done[2 * qt + NSUB - 1], (pvbase - 1) & 1);  // the item's last MMA: every earlier one has completed too
      tcgen05_fence_after();
      uint32_t o[32];
      tmem_ld_x32(o_addr, o);
      tmem_wait_ld();
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&o_free[qt]);  // the next item's first PV may overwrite O_qt
      xsum[(qt * 2 + hf) * 128 + row] = l;
      if (leader) tma_store_wait_read<0>();  // the previous store has drained the staging tile
      named_bar_sync(1 + qt, 256);
      const float l_row = l + xsum[(qt * 2 + (hf ^ 1)) * 128 + row];
      if (slot >= 0) {
        // this CTA holds only part of the item's keys: leave the unnormalised share for attn_merge_kernel
        float4* po = reinterpret_cast<float4*>(a.part_o + ((size_t)slot * 256 + qt * 128 + row) * 64 + 32 * hf);
#pragma unroll
        for (int c = 0; c < 8; ++c)
          po[c] = make_float4(__uint_as_float(o[4 * c]), __uint_as_float(o[4 * c + 1]), __uint_as_float(o[4 * c + 2]),
                              __uint_as_float(o[4 * c + 3]));
        if (hf == 0) {
          a.part_l[(size_t)slot * 256 + qt * 128 + row] = l_row;
          a.part_m[(size_t)slot * 256 + qt * 128 + row] = m;
        }
        named_bar_sync(1 + qt, 256);  // xsum is rewritten by the next segment
        continue;
      }
      const float inv = 1.0f / l_row;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t pk4[4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
          pk4[i] = pack_bf16(__uint_as_float(o[8 * c + 2 * i]) * inv, __uint_as_float(o[8 * c + 2 * i + 1]) * inv);
        st_shared_v4(stg + sw128_offset(row, 4 * hf + c), pk4[0], pk4[1], pk4[2], pk4[3]);
      }
      fence_proxy_async_smem();
      named_bar_sync(1 + qt, 256);
      if (leader) {
        tma_store_3d(&tmO, sO + qt * ATT_TILE, h * 64, q0 + qt * 128, b);
        tma_store_commit();
      }
    }
    if (leader) tma_store_wait<0>();
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 19) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------------------ host
template <int NPOLY, bool BOUNDED, bool HALVES>
static wm_status launch_pipe(cudaStream_t st, int grid, const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv,
                             const CUtensorMap& to, const AttnArgs& a) {
  static int cfg_dev[16] = {0};  // the shared-memory attribute is per device
  int dev = 0;
  WM_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 16 || !cfg_dev[dev]) {
    WM_CUDA(cudaFuncSetAttribute(attn_pipe_kernel<NPOLY, BOUNDED, HALVES>, cudaFuncAttributeMaxDynamicSharedMemorySize, ATP_SMEM));
    if (dev >= 0 && dev < 16) cfg_dev[dev] = 1;
  }
  return launch_pdl(attn_pipe_kernel<NPOLY, BOUNDED, HALVES>, dim3(grid), dim3(ATP_THREADS), ATP_SMEM, st, tq, tk, tv, to, a);
}

// variants 13..19: bounded-logit mode (q pre-scaled); variants 20..23: general (running-max) softmax
wm_status flash_attention_pipe(cudaStream_t st, int grid, const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv,
                               const CUtensorMap& to, const AttnArgs& a, int variant) {
  switch (variant) {
    case 13: return launch_pipe<12, true, false>(st, grid, tq, tk, tv, to, a);
    case 14: return launch_pipe<14, true, false>(st, grid, tq, tk, tv, to, a);
    case 15: return launch_pipe<16, true, false>(st, grid, tq, tk, tv, to, a);
    case 16: return launch_pipe<10, true, false>(st, grid, tq, tk, tv, to, a);
    case 17: return launch_pipe<12, true, true>(st, grid, tq, tk, tv, to, a);
    case 18: return launch_pipe<8, true, true>(st, grid, tq, tk, tv, to, a);
    case 19: return launch_pipe<16, true, true>(st, grid, tq, tk, tv, to, a);
    case 20: return launch_pipe<0, false, false>(st, grid, tq, tk, tv, to, a);
    case 21: return launch_pipe<8, false, false>(st, grid, tq, tk, tv, to, a);
    case 22: return launch_pipe<12, false, false>(st, grid, tq, tk, tv, to, a);
    case 23: return launch_pipe<16, false, false>(st, grid, tq, tk, tv, to, a);
    default: return fail(WM_ERR_INVALID, "flash_attention: unknown pipelined variant %d", variant);
  }
}

}  // namespace wm
