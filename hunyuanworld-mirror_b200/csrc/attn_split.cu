// Flash attention, head_dim 64, "split-row" variant: the same tile pipeline as attn.cu (TMA-fed K/V ring,
// tcgen05.mma with S, O and P in TMEM, two 128-row query tiles per CTA) but every 128 x 128 score tile is
// shared by TWO softmax warpgroups -- thread == (query row, 64-key half) -- so 16 softmax warps (4 per SM
// sub-partition) are in flight instead of 8.  ncu on attn.cu showed no saturated pipe (tensor 35 %, MUFU 43 %,
// FMA 43 %, issue 52 %): two warps per scheduler cannot hide MUFU / FMA / TMEM latencies; four can.
//
// 640 threads: warp 0 TMA producer, warp 1 MMA issuer, warp 2 TMEM allocator, warp 3 idle (setmaxnreg acts on
// whole warpgroups, so the control roles fill one), warps 4-19 softmax:
//   group g = (warp - 4) / 4, query tile qt = g / 2, key half hf = g % 2, TMEM lane quarter = warp % 4.
// Per tile the two halves of a row exchange their half-row maxima through shared memory (one named barrier of
// 256 threads per query tile) so both use the same reference maximum; row sums stay partial per half and are
// merged once per work item in the epilogue.  Everything else (lazy rescale, polynomial exp2 share, ragged
// last tile) follows attn.cu.
#include "attn_common.cuh"

namespace wm {

constexpr int ATS_THREADS = 640;
constexpr int ATS_XCH = 2 * 2 * 2 * 128 * 4 + 2 * 2 * 128 * 4;  // half-row maxima (double-buffered) + partial row sums
constexpr int ATS_SMEM = (4 + 2 * KV_ST + 2) * ATT_TILE + 512 + ATS_XCH + 1024;

// The softmax / correction / epilogue role of attn_split_kernel (warps 4..19: two threads per query row, see the kernel).
template <int NPOLY, bool BOUNDED, int STREAM>
__device__ __forceinline__ void attn_split_softmax(const AttnArgs& a, const CUtensorMap* tmO_p, uint8_t* sO, uint64_t* bars,
                                                   float* xmax, float* xsum, uint32_t tmem_base, int warp) {
  constexpr int ST = KV_ST;
  const int grp = (warp - 4) >> 2;
  const int qt = grp >> 1;                     // query tile of this warpgroup
  const int hf = grp & 1;                      // key half [64 hf, 64 hf + 64) of every score tile
  const int lane = threadIdx.x & 31;
  const int row = (warp & 3) * 32 + lane;
  const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
  const uint32_t s_addr = tmem_base + s_col(qt) + 64u * hf + lane_base;
  const uint32_t o_addr = tmem_base + o_col(qt) + 32u * hf + lane_base;  // this thread rescales / reads 32 of the 64 O columns
  const uint32_t p_addr = tmem_base + p_col(qt) + 32u * hf + lane_base;
  const uint32_t stg = smem_u32(sO + qt * ATT_TILE);
  const bool leader = hf == 0 && row == 0;
  const bool ragged = (a.Lkv & 127) != 0;
  const float2 scale2 = make_float2(a.scale_log2, a.scale_log2);
  float* my_max = xmax + (qt * 4 + hf) * 128 + row;          // + parity * 256
  const float* peer_max = xmax + (qt * 4 + (hf ^ 1)) * 128 + row;
  uint32_t scount = 0, pvbase = 0;  // s_full[qt] phases consumed / pv_done[qt] phases of finished items
  // this warp's five barriers as ONE shared-window address + constant offsets, and the logical index of the ragged key
  // tile, both computed once: the softmax loop spent ~30 of its ~300 instructions per tile rebuilding them
  uint32_t bq = smem_u32(bars) + 8u * qt;
  asm volatile("" : "+r"(bq));  // opaque: kept in a register, not rematerialised from the generic pointer
  constexpr uint32_t B_SFULL = (4 + 4 * ST) * 8, B_SFREE = B_SFULL + 16, B_PFULL = B_SFULL + 32, B_PVDONE = B_SFULL + 48,
                     B_OFREE = B_SFULL + 64;
  const int j_rag = ragged ? (2 * a.n_kv - 1 - a.kv_rot) % a.n_kv : -1;  // attn_phys_tile(a, j_rag) == n_kv - 1

  AttnSched sched(a);
  for (int item, kv0, kv1, slot; sched.next(a, item, kv0, kv1, slot);) {
    const int per_head = a.batch * a.n_qpairs;
    const int h = item / per_head, rem = item % per_head;
    const int b = rem / a.n_qpairs, q0 = (rem % a.n_qpairs) * 256;
    if (qt == 1 && q0 + 128 >= a.Lq) continue;

    float m = BOUNDED ? 0.f : -INFINITY, l = 0.f;
    for (int j = kv0; j < kv1; ++j, ++scount) {
      mbar_wait(bq + B_SFULL, scount & 1);
      tcgen05_fence_after();
      uint32_t pk[32];
      bool pv_waited = false;
      if constexpr (BOUNDED && STREAM != 0) {
        // (Measured, profiles/r02_s10_attn_stream_v1.log, r02_s11_attn_stream_v2.log: with the PV(j-1) wait at the top
        // of the step 764 TFLOP/s; before the first half-store 893; STREAM 2 = unrolled, prefetching 16-column loads and
        // one store at the end 966; the plain form below 986-994.  Not selected by default.)
        // Streaming form: the 64 scores of this thread go through in four chunks of 16 inside a ROLLED loop, every
        // chunk's TMEM load issued one chunk ahead (the load latency hides under the previous chunk's exponentials)
        // and its 8 packed probabilities stored to P straight away.  ptxas cannot pull the MUFU.EX2 of a whole tile
        // into one burst any more (at most 16 * (1 - NPOLY/32) per chunk), and at most 32 scores + 8 packed values
        // are live.  P_qt must be free before the first store: the PV(j-1) wait sits before the first (16-column) store,
        // after half of the tile's exponentials (at the top of the step it was the hottest PC: 764 TFLOP/s).
        pv_waited = true;
        float2 sum2[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
        int valid = 64;
        if (j == j_rag) valid = a.Lkv - (a.n_kv - 1) * 128 - hf * 64;
        uint32_t va[16], vb[16];
        tmem_ld_x16(s_addr, va);
#pragma unroll(STREAM == 2 ? 2 : 1)
        for (int i = 0; i < 2; ++i) {
          tmem_wait_ld();
          tmem_ld_x16(s_addr + 32 * i + 16, vb);
          uint32_t pk16s[16];
          uint32_t* pk16 = STREAM == 2 ? pk + 16 * i : pk16s;  // STREAM 2: all 32 packed values kept, one store at the end
          {
            uint32_t* pk8 = pk16;
            const int lim = valid - 32 * i;
            if (lim < 16) {  // only in the last, ragged key tile: P = 2^-120 for the keys beyond Lkv
              __syncwarp();  // (warp-uniform; keeps this a real branch instead of 30 predicated selects per chunk)
#pragma unroll
              for (int c = 0; c < 16; ++c)
                if (c >= lim) va[c] = 0xc2f00000u;
            }
#pragma unroll
            for (int c = 0; c < 16; c += 2) {
              const float2 xs = make_float2(__uint_as_float(va[c]), __uint_as_float(va[c + 1]));
              float2 p;
              if (c >= 16 - NPOLY / 2) {
                p = exp2_poly2_bounded(xs);
              } else {
                p.x = ex2_approx(xs.x);
                p.y = ex2_approx(xs.y);
              }
              sum2[(c >> 1) & 1] = fadd2(sum2[(c >> 1) & 1], p);
              pk8[c >> 1] = pack_bf16(p.x, p.y);
            }
          }
          tmem_wait_ld();
          if (i == 0) {
            tmem_ld_x16(s_addr + 32, va);
          } else {
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bq + B_SFREE);  // all 64 scores of this thread have left S_qt
          }
          {
            uint32_t* pk8 = pk16 + 8;
            const int lim = valid - 32 * i - 16;
            if (lim < 16) {
              __syncwarp();
#pragma unroll
              for (int c = 0; c < 16; ++c)
                if (c >= lim) vb[c] = 0xc2f00000u;
            }
#pragma unroll
            for (int c = 0; c < 16; c += 2) {
              const float2 xs = make_float2(__uint_as_float(vb[c]), __uint_as_float(vb[c + 1]));
              float2 p;
              if (c >= 16 - NPOLY / 2) {
                p = exp2_poly2_bounded(xs);
              } else {
                p.x = ex2_approx(xs.x);
                p.y = ex2_approx(xs.y);
              }
              sum2[(c >> 1) & 1] = fadd2(sum2[(c >> 1) & 1], p);
              pk8[c >> 1] = pack_bf16(p.x, p.y);
            }
          }
          if constexpr (STREAM == 1) {
            if (i == 0 && j > kv0) {
              mbar_wait(bq + B_PVDONE, (pvbase + j - kv0 - 1) & 1);
              tcgen05_fence_after();
            }
            tmem_st_x16(p_addr + 16 * i, pk16s);
          }
        }
        if constexpr (STREAM == 2) {
          if (j > kv0) {
            mbar_wait(bq + B_PVDONE, (pvbase + j - kv0 - 1) & 1);
            tcgen05_fence_after();
          }
          tmem_st_x32(p_addr, pk);
        }
        l += (sum2[0].x + sum2[0].y) + (sum2[1].x + sum2[1].y);
        tmem_wait_st();
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bq + B_PFULL);
        continue;
      } else if constexpr (BOUNDED) {
        // two 32-column chunks, one after the other: at most 32 scores + 16 packed probabilities are live
        // (measured: reading all 64 scores first and freeing S early is 7 % SLOWER, 910 vs 977 TFLOP/s)
        float2 sum2[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          uint32_t v[32];
          tmem_ld_x32(s_addr + 32 * i, v);
          tmem_wait_ld();
          if (i == 1) {
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bq + B_SFREE);  // the MMA warp may overwrite S_qt once both halves have arrived
          }
          if (j == j_rag) {
            const int valid = a.Lkv - (a.n_kv - 1) * 128 - hf * 64 - i * 32;
#pragma unroll
            for (int c = 0; c < 32; ++c)
              if (c >= valid) v[c] = 0xc2f00000u;  // -120: P = 2^-120, >= 2^40 below any real term (logit bound <= 80)
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
        uint32_t v[2][32];
        tmem_ld_x32(s_addr, v[0]);
        tmem_ld_x32(s_addr + 32, v[1]);
        tmem_wait_ld();
        tcgen05_fence_before();
        __syncwarp();
            if (lane == 0) mbar_arrive(bq + B_SFREE);  // the MMA warp may overwrite S_qt once both halves have arrived
        if (j == j_rag) {
          const int valid = a.Lkv - (a.n_kv - 1) * 128 - hf * 64;
#pragma unroll
          for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int c = 0; c < 32; ++c)
              if (i * 32 + c >= valid) v[i][c] = 0xff800000u;  // -inf
        }
        float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
          for (int c = 0; c < 32; ++c) mx[c & 3] = fmaxf(mx[c & 3], __uint_as_float(v[i][c]));
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
            mbar_wait(bq + B_PVDONE, (pvbase + j - kv0 - 1) & 1);
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
        const float ms = m * a.scale_log2;
        const float2 nms2 = make_float2(-ms, -ms);
        float2 sum2[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
#pragma unroll
        for (int i = 0; i < 2; ++i) {
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
            pk[i * 16 + (c >> 1)] = pack_bf16(p.x, p.y);
          }
        }
        l += (sum2[0].x + sum2[0].y) + (sum2[1].x + sum2[1].y);
      }
      if (j > kv0 && !pv_waited) {  // P_qt is still being read by PV(j-1) until this completes (normally long done)
        mbar_wait(bq + B_PVDONE, (pvbase + j - kv0 - 1) & 1);
        tcgen05_fence_after();
      }
      tmem_st_x32(p_addr, pk);
      tmem_wait_st();
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bq + B_PFULL);
    }
    // ---- epilogue: merge the partial row sums, O / l -> bf16 -> swizzled staging tile -> TMA store
    pvbase += kv1 - kv0;
    mbar_wait(bq + B_PVDONE, (pvbase - 1) & 1);
    tcgen05_fence_after();
    uint32_t o[32];
    tmem_ld_x32(o_addr, o);
    tmem_wait_ld();
    tcgen05_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(bq + B_OFREE);  // the next item's first PV may overwrite O_qt
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
      tma_store_3d(tmO_p, sO + qt * ATT_TILE, h * 64, q0 + qt * 128, b);
      tma_store_commit();
    }
  }
  if (leader) tma_store_wait<0>();
}

// BOUNDED: the caller guarantees |q . k| <= ~96 with q pre-multiplied by scale * log2(e) (QK-LayerNorm bounds the
// logits; the bound is computed from the norm weights at load time, capi.cu).  softmax is shift-invariant, so the
// reference maximum is fixed at 0: P = 2^s needs no running max, no exchange between the halves, no rescale
// and no clamp -- exact up to the same bf16 rounding of P, because bf16 keeps 8 exponent bits.
// NPOLY_G >= 0: the kernel carries the bounded form (NPOLY) AND the running-max form (NPOLY_G) of the softmax role and
// picks one from the guard words at run time (AttnGuard mode 3).
template <int NPOLY, bool BOUNDED, int STREAM, int NPOLY_G = -1>
__global__ void __launch_bounds__(ATS_THREADS, 1)
attn_split_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
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
  uint64_t* s_full = v_empty + ST;     // 2  MMA -> softmax: S_qt written
  uint64_t* s_free = s_full + 2;       // 2  softmax -> MMA: S_qt is in registers (8 warp arrivals)
  uint64_t* p_full = s_free + 2;       // 2  softmax -> MMA: P_qt written (8 warp arrivals)
  uint64_t* pv_done = p_full + 2;      // 2  MMA -> softmax: O_qt += P_qt V complete
  uint64_t* o_free = pv_done + 2;      // 2  epilogue -> MMA: O_qt has been read out (8 warp arrivals)
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(o_free + 2);
  float* xmax = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 512);  // [qt][parity][half][128]
  float* xsum = xmax + 2 * 2 * 2 * 128;                                            // [qt][half][128]

  pdl_launch_dependents();
  if (NPOLY_G < 0 && a.guard_mode != 0) {  // one of a (bounded, running-max) pair of launches: the data decides which one works
    pdl_wait();
    if (attn_guard_skip(a)) return;
  }
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
      mbar_init(&s_free[i], 8);  // one arrival per softmax warp (lane 0 after __syncwarp): every arrival wakes the
      mbar_init(&p_full[i], 8);  // CTA's sleeping try_wait-ers, so 256 per-thread arrivals per tile cost issue slots
      mbar_init(&pv_done[i], 1);
      mbar_init(&o_free[i], 8);
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
  pdl_wait();  // q, k, v of the preceding GEMM (or all-gather) are complete and visible from here on

  if (warp < 4) {
    setmaxnreg_dec<56>();
    if (warp == 0) {
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
    } else if (warp == 1) {
      // -------------------------------------------------------------- MMA issuer
      if (elect_one()) {
        constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128, false, false);
        constexpr uint32_t idesc_pv = umma_idesc_bf16(128, 64, false, true);  // B = V is MN-major (dims contiguous)
        uint32_t kvit = 0, icount = 0;
        uint32_t tiles0 = 0, tiles1 = 0;  // S tiles issued per query tile (s_free phase bookkeeping)
        uint32_t pvs0 = 0, pvs1 = 0;      // PVs issued per query tile (p_full phases consumed)
        uint32_t items0 = 0, items1 = 0;  // work items started per query tile (o_free phases)
        AttnSched sched(a);
        for (int item, kv0, kv1, slot; sched.next(a, item, kv0, kv1, slot); ++icount) {
          const int q0 = (item % a.n_qpairs) * 256;
          const int nw = (q0 + 128 < a.Lq) ? 2 : 1;  // second query tile entirely out of range: skip it
          const uint32_t qb = icount & 1, qph = (icount >> 1) & 1;
          mbar_wait(&q_full[qb], qph);
          tcgen05_fence_after();
          auto issue_s = [&](int w, uint32_t stage) {
            uint32_t& tiles = w == 0 ? tiles0 : tiles1;
            mbar_wait(&s_free[w], (tiles & 1) ^ 1);  // both halves have read the previous S_w into registers
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
          for (int j = kv0; j < kv1; ++j, ++kvit) {
            const uint32_t sj = kvit % ST, phj = (kvit / ST) & 1;
            if (j + 1 < kv1) {
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
              if (j == kv0) mbar_wait(&o_free[w], ((w == 0 ? items0 : items1) & 1) ^ 1);  // previous segment's O_w read out
              tcgen05_fence_after();
              const uint32_t vbase = smem_u32(sV + sj * ATT_TILE);
#pragma unroll
              for (int k = 0; k < 8; ++k) {  // 16 kv rows (2048 B of the V tile) per step; P: 8 TMEM columns per step
                const uint64_t vd = umma_desc_sw128(vbase + k * 2048, 1024, 1024);
                umma_ts(tmem_base + o_col(w), tmem_base + p_col(w) + k * 8, vd, idesc_pv, ((j - kv0) | k) != 0);
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
    setmaxnreg_inc<104>();
    if constexpr (NPOLY_G >= 0) {
      // one launch, the data decides (AttnGuard mode 3): both forms of the role are in the kernel, every thread takes the
      // same one; producer and MMA issuer do not depend on the choice
      if (attn_guard_fits(a))
        attn_split_softmax<NPOLY, true, 0>(a, &tmO, sO, bars, xmax, xsum, tmem_base, warp);
      else
        attn_split_softmax<NPOLY_G, false, 0>(a, &tmO, sO, bars, xmax, xsum, tmem_base, warp);
    } else {
      attn_split_softmax<NPOLY, BOUNDED, STREAM>(a, &tmO, sO, bars, xmax, xsum, tmem_base, warp);
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 2) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------------------ merge
// Adds up the partials of every work item whose keys were split between CTAs (at most a handful per launch) and
// writes the normalised bf16 rows.  One block per tail item; items that were computed whole return at once.
// Partials are visited in CTA order, so the result does not depend on timing.
__global__ void __launch_bounds__(256)
attn_merge_kernel(const AttnArgs a, __nv_bfloat16* __restrict__ out, int ldo) {
  pdl_launch_dependents();
  pdl_wait();
  if (attn_guard_skip(a)) return;
  const int ti = blockIdx.x >> 4;                         // tail item; 16 blocks of 16 rows each
  const int item = a.full_waves * a.merge_grid + ti;
  const int s0 = ti * a.n_kv, s1 = s0 + a.n_kv;           // its steps inside the tail
  const int c0 = s0 / a.per_cta, c1 = (s1 - 1) / a.per_cta;
  if (c0 == c1) return;
  const int per_head = a.batch * a.n_qpairs;
  const int h = item / per_head, rem = item % per_head;
  const int b = rem / a.n_qpairs, q0 = (rem % a.n_qpairs) * 256;
  const int c4 = threadIdx.x & 15;
  const int r = (blockIdx.x & 15) * 16 + (threadIdx.x >> 4);
  if (q0 + r >= a.Lq) return;
  const int first_slot = 2 * c0 + (s0 != c0 * a.per_cta ? 1 : 0);  // later CTAs hold the item in their first segment
  float M = a.part_m[(size_t)first_slot * 256 + r];
#pragma unroll 4
  for (int c = c0 + 1; c <= c1; ++c) M = fmaxf(M, a.part_m[(size_t)(2 * c) * 256 + r]);
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  float L = 0.f;
#pragma unroll 4
  for (int c = c0; c <= c1; ++c) {
    const int slot = c == c0 ? first_slot : 2 * c;
    const float w = ex2_approx((a.part_m[(size_t)slot * 256 + r] - M) * a.scale_log2);
    const float4 o = reinterpret_cast<const float4*>(a.part_o + ((size_t)slot * 256 + r) * 64)[c4];
    acc.x = fmaf(w, o.x, acc.x), acc.y = fmaf(w, o.y, acc.y), acc.z = fmaf(w, o.z, acc.z), acc.w = fmaf(w, o.w, acc.w);
    L = fmaf(w, a.part_l[(size_t)slot * 256 + r], L);
  }
  const float inv = 1.0f / L;
  uint2 pk = make_uint2(pack_bf16(acc.x * inv, acc.y * inv), pack_bf16(acc.z * inv, acc.w * inv));
  *reinterpret_cast<uint2*>(out + ((size_t)b * a.Lq + q0 + r) * ldo + h * 64 + c4 * 4) = pk;
}

// ------------------------------------------------------------------------------------------------ host
template <int NPOLY, bool BOUNDED, int STREAM = 0, int NPOLY_G = -1>
static wm_status launch_split(cudaStream_t st, int grid, const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv,
                              const CUtensorMap& to, const AttnArgs& a) {
  static int cfg_dev[16] = {0};  // the shared-memory attribute is per device, not per process
  int dev = 0;
  WM_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 16 || !cfg_dev[dev]) {
    WM_CUDA(cudaFuncSetAttribute(attn_split_kernel<NPOLY, BOUNDED, STREAM, NPOLY_G>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 ATS_SMEM));
    if (dev >= 0 && dev < 16) cfg_dev[dev] = 1;
  }
  return launch_pdl(attn_split_kernel<NPOLY, BOUNDED, STREAM, NPOLY_G>, dim3(grid), dim3(ATS_THREADS), ATS_SMEM, st, tq, tk, tv, to, a);
}

wm_status flash_attention_pipe(cudaStream_t st, int grid, const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv,
                               const CUtensorMap& to, const AttnArgs& a, int variant);  // attn_pipe.cu

size_t attn_partial_bytes(int sm_count) { return (size_t)2 * sm_count * kAttnSlotFloats * sizeof(float); }

// variants 4..7 of flash_attention(): running-max softmax, polynomial share 0, 8, 10, 12 of every 32 exponentials;
// variants 8..12: bounded-logit mode (q pre-scaled, |logit| bounded), polynomial share 0, 8, 12, 14, 16;
// variants 13..19: bounded-logit mode, half-tile pipelined (attn_pipe.cu);
// variants 24..27: bounded-logit mode with the streaming (rolled, prefetching) softmax loop, polynomial share 12, 16 (P stored in halves) and 12, 16 (one P store at the end).
// With a partial workspace (attn_partial_bytes) the (item, kv tile) steps are dealt out evenly and split items are
// merged by attn_merge_kernel; without one every CTA gets whole items.
wm_status flash_attention_split(cudaStream_t st, int sm_count, const CUtensorMap& tq, const CUtensorMap& tk,
                                const CUtensorMap& tv, const CUtensorMap& to, AttnArgs a, int variant, void* out, int ldo,
                                float* partial_ws, size_t partial_ws_bytes, const AttnGather* gather, const AttnGuard* guard) {
  a.guard_norm2 = nullptr, a.guard_mode = 0, a.guard_heads = 0, a.guard_limit2 = 0.f;
  if (guard) {
    WM_CHECK(guard->norm2 && guard->mode >= 1 && guard->mode <= 3 && guard->limit > 0.f && gather == nullptr,
             "flash_attention: bad guard description (or combined with an in-flight gather)");
    WM_CHECK((variant >= 4 && variant <= 12) || (variant >= 24 && variant <= 27),
             "flash_attention: the device-side guard needs a split-row variant (4..12, 24..27)");
    const bool bounded_variant = variant >= 8;
    WM_CHECK(bounded_variant == (guard->mode != 2), "flash_attention: guard modes 1 and 3 go with a bounded variant, mode 2 with a general one");
    WM_CHECK(guard->mode != 3 || variant == 10, "flash_attention: the one-launch guard (mode 3) is built for variant 10 (+ 4)");
    a.guard_norm2 = guard->norm2, a.guard_mode = guard->mode, a.guard_limit2 = guard->limit * guard->limit;
    a.guard_heads = a.total_items / (a.batch * a.n_qpairs);
  }
  a.kv_rot = 0, a.shard_rows = 0, a.self_rank = 0, a.epoch = 0, a.flags = nullptr, a.err = nullptr;
  if (gather) {
    WM_CHECK(gather->shard_rows > 0 && gather->flags && gather->err && gather->self_rank >= 0 && gather->self_rank < 32 &&
                 (long long)gather->shard_rows * (gather->self_rank + 1) <= a.Lkv && a.batch == 1,
             "flash_attention: bad gather description");
    a.shard_rows = gather->shard_rows, a.self_rank = gather->self_rank, a.epoch = gather->epoch;
    a.flags = gather->flags, a.err = gather->err;
    a.kv_rot = (int)(((long long)gather->self_rank * gather->shard_rows) / 128);  // own shard first (floor: a tile that
  }                                                                               // straddles two shards waits for both)
  const bool balance = partial_ws != nullptr && partial_ws_bytes >= attn_partial_bytes(sm_count) && a.total_items > 0;
  const int grid = a.total_items < sm_count ? a.total_items : sm_count;
  a.full_waves = a.total_items / grid;
  const int tail_items = a.total_items - a.full_waves * grid;
  a.tail_steps = tail_items * a.n_kv;
  a.merge_grid = grid;
  // splitting the tail pays a partial store, a merge launch and a pipeline refill: about ten tile times (measured
  // at 1376 keys, where whole items are faster); take it only when more than that is saved
  const int even_share = (a.tail_steps + grid - 1) / grid;
  if (balance && tail_items > 0 && even_share + 10 <= a.n_kv) {
    a.per_cta = even_share;
    a.part_o = partial_ws;
    a.part_l = partial_ws + (size_t)2 * sm_count * 256 * 64;
    a.part_m = a.part_l + (size_t)2 * sm_count * 256;
  } else {
    a.per_cta = a.n_kv;  // whole items only: no partials; the tail items go to the last CTAs (AttnSched)
    a.part_o = a.part_l = a.part_m = nullptr;
  }
  const int last_rows = a.Lq - (a.n_qpairs - 1) * 256;
  a.n_cheap = (a.part_o == nullptr && last_rows <= 128 && a.total_items > grid) ? a.total_items / a.n_qpairs : 0;
  wm_status r;
  switch (variant) {
    case 4: r = launch_split<0, false>(st, grid, tq, tk, tv, to, a); break;
    case 5: r = launch_split<8, false>(st, grid, tq, tk, tv, to, a); break;
    case 6: r = launch_split<10, false>(st, grid, tq, tk, tv, to, a); break;
    case 7: r = launch_split<12, false>(st, grid, tq, tk, tv, to, a); break;
    case 8: r = launch_split<0, true>(st, grid, tq, tk, tv, to, a); break;
    case 9: r = launch_split<8, true>(st, grid, tq, tk, tv, to, a); break;
    case 10:
      r = a.guard_mode == 3 ? launch_split<12, true, 0, 0>(st, grid, tq, tk, tv, to, a) : launch_split<12, true>(st, grid, tq, tk, tv, to, a);
      break;
    case 11: r = launch_split<14, true>(st, grid, tq, tk, tv, to, a); break;
    case 12: r = launch_split<16, true>(st, grid, tq, tk, tv, to, a); break;
    case 24: r = launch_split<12, true, 1>(st, grid, tq, tk, tv, to, a); break;  // streaming softmax loop, P stored in halves
    case 25: r = launch_split<16, true, 1>(st, grid, tq, tk, tv, to, a); break;
    case 26: r = launch_split<12, true, 2>(st, grid, tq, tk, tv, to, a); break;  // streaming loop, one P store at the end
    case 27: r = launch_split<16, true, 2>(st, grid, tq, tk, tv, to, a); break;
    default:
      if (variant >= 13 && variant <= 23) {  // second-generation kernels (attn_pipe.cu): 13-19 bounded, 20-23 general
        r = flash_attention_pipe(st, grid, tq, tk, tv, to, a, variant);
        break;
      }
      return fail(WM_ERR_INVALID, "flash_attention: unknown split variant %d", variant);
  }
  WM_TRY(r);
  if (a.part_o != nullptr && a.per_cta % a.n_kv != 0)
    return launch_pdl(attn_merge_kernel, dim3(16 * tail_items), dim3(256), 0, st, a, static_cast<__nv_bfloat16*>(out), ldo);
  return WM_OK;
}

}  // namespace wm
