// Persistent, warp-specialised tcgen05 GEMM for the aggregator's linear layers:
//     acc[M,N] (fp32, TMEM) = A[M,K] (bf16, row-major) x W[N,K]^T (bf16, nn.Linear layout)
// with the epilogue fused per reference call site:
//   EPI_BIAS   out bf16 = acc + b                                   generic nn.Linear
//   EPI_GELU   out bf16 = gelu_erf(acc + b)                         Mlp.fc1 + nn.GELU   (mlp.py:30-31)
//   EPI_QKV    per 64-wide head: + b, QK-LayerNorm(64), 2D RoPE     attention.py:50-56, rope.py:148-181
//              q -> out0[M,1024], k|v -> out1[M,2048]  (both bf16)
//   EPI_RESID  x fp32 += gamma * (acc + b) as a TMA reduce-add (the residual is never read by the SM);
//              gamma null: x = acc + b   (attn.proj / Mlp.fc2 + LayerScale + residual: block.py:91-92)
//   EPI_TABLE  out fp32 = acc + table[row % period][col]            DINOv2 patch embed + pos_embed / cls / reg
//
// CTA pairs (clusters of 2, tcgen05 cta_group::2): a pair computes a 256 x 256 tile, CTA r owning rows
// [128 r, 128 r + 128) of A and of the accumulator and rows [128 r, 128 r + 128) of the 256-row W tile.  One
// M = 256 MMA issued by the leader CTA reads both halves, so every SM streams 32 KB instead of 48 KB per k-block
// through its shared memory: the single-CTA 128 x 256 version measured 70-73 % tensor-pipe utilisation on every
// shape, the ratio of the 128 B/clk shared-memory port to its 192 B/clk of TMA writes + MMA operand reads.
// Per CTA, 384 threads: warp 0 = TMA producer, warp 1 = MMA issuer (leader CTA only, one elected lane),
// warp 2 = TMEM allocator, warps 4-11 = epilogue in two teams of four warps (team t owns accumulator columns
// [128 t, 128 t + 128); thread == accumulator row; TMEM lane quarter = warp % 4).
// 5-stage smem ring (SWIZZLE_128B, K-major for both operands), two 256-column TMEM accumulators so the epilogue
// of tile i overlaps the main loop of tile i+1.  Outputs leave through one swizzled staging tile per team and
// TMA stores / reduce-adds (row clipping for the M tail comes for free).
//
// Tail split (EPI_RESID, N = 1024: proj / fc2).  172 pair tiles on 74 CTA pairs are 2.32 waves and used to run as 3.
// With a word of counters per (tail tile, rank, team) the tiles beyond the last FULL wave are cut along K into `split`
// slices, one per CTA pair, and run FIRST.  Every slice adds its share gamma * (acc_s [+ b for s = 0]) to x with the
// same TMA reduce-add as a whole tile, but IN SLICE ORDER: the epilogue team of slice s waits for the counter to reach
// s (slice s-1 has seen its reduce-adds complete, cp.async.bulk.wait_group 0, then released the counter), the last
// slice writes 0 back.  x = ((x + g (p0 + b)) + g p1) + g p2: one fixed order of fp32 additions, independent of
// timing, so the forward stays bit-deterministic; no partial-sum workspace, no fix-up pass.  The slices are the first
// unit of distinct pairs of one co-resident persistent grid (no deadlock), and while a team waits the MMA warp runs
// on with the pair's next tile in the other TMEM accumulator.  2 + 1/3 waves instead of 3.
// (A first version parked fp32 partials in a workspace and let the owner add them at the end: row-strided 16-byte
// accesses and a serial fix-up made it SLOWER than 3 waves, 48 vs 25 us for proj; profiles/r02_s2_selftest.log.)
#include <stdlib.h>

#include "common.cuh"
#include "ops.cuh"
#include "ptx.cuh"

namespace wm {

enum : int { EPI_BIAS = 0, EPI_GELU = 1, EPI_QKV = 2, EPI_RESID = 3, EPI_TABLE = 4 };

constexpr int BM = 128, BN = 256, BK = 64, STAGES = 5;  // BM: rows per CTA (the pair tile has 2 BM)
constexpr int A_BYTES = BM * BK * 2;        // 16 KB
constexpr int B_BYTES = (BN / 2) * BK * 2;  // 16 KB: this CTA's half of the W tile
constexpr int STG_BYTES = 128 * 128;        // 128 rows x 128 B, one per epilogue team
constexpr int GEMM_THREADS = 384;           // 4 control warps + 2 epilogue teams of 4 warps
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
  float q_scale;  // q is multiplied by this after norm + rope (softmax scale * log2 e folded in for bounded-logit attention)
  unsigned* norm2;  // AttnGuard words or null: [head] max |q row|^2 (after q_scale), [16 + head] max |k row|^2, float bits
  // EPI_RESID
  const float* gamma;  // null: out = acc + b (plain store); else out += gamma * (acc + b) (TMA reduce-add)
  // EPI_TABLE
  const float* table;
  int table_period;
  // tail split (EPI_RESID only; split == 1: every pair walks whole tiles round-robin, as the other epilogues do)
  int split, full_tiles, tail_tiles;  // tiles [0, full_tiles) whole; tiles [full_tiles, full_tiles + tail_tiles) in `split` K slices
  // stream-K (EPI_RESID only): the (tile, k block) units of the whole problem are cut into one contiguous range per CTA
  // pair, so every pair does the same amount of MMA work; a tile that straddles two ranges is added to x in two pieces, in a
  // fixed order (see GemmSched).  split == 2 then only describes the flag protocol.
  int streamk;
  unsigned* flags;  // [tail_tiles][2 ranks][2 teams] slice counters; zero before the launch, left zero by it
};

// The work units of one CTA pair, in the order every role of the pair walks them.
struct GemmUnit {
  int tile, kb0, kb1, slice;  // slice < 0: a whole tile with the normal epilogue
  int flag_idx;               // which group of slice counters orders the adds of a split tile
};
struct GemmSched {
  int pair_id, num_pairs, i;
  int u0, u1;  // stream-K: the pair's range of (tile, k block) units, u = tile * num_k_blocks + kb
  __device__ GemmSched(int pair, int pairs) : pair_id(pair), num_pairs(pairs), i(-1), u0(0), u1(0) {}
  __device__ bool next(const GemmArgs& g, GemmUnit& u) {
    if (g.streamk) {
      // Pair p owns units [total p / P, total (p + 1) / P): at least one whole tile's worth (the host takes this mode only
      // for tiles >= pairs), so a tile is cut at most once.  The piece that ENDS a tile (kb0 > 0) is the first unit of its
      // pair and adds first, with the bias (slice 0); the piece that BEGINS it (kb1 < num_k_blocks) is the LAST unit of the
      // previous pair and adds second (slice 1).  The two are a whole kernel apart in time, so the second never waits,
      // and the order of the fp32 additions is fixed: bit-deterministic.  The cut at the start of pair p uses counters p.
      if (i < 0) {
        const long long total = (long long)g.num_m_tiles * g.num_n_tiles * g.num_k_blocks;
        u0 = (int)(total * pair_id / num_pairs);
        u1 = (int)(total * (pair_id + 1) / num_pairs);
        i = 0;
      }
      if (u0 >= u1) return false;
      const int nkb = g.num_k_blocks;
      u.tile = u0 / nkb;
      u.kb0 = u0 - u.tile * nkb;
      const int end = min(u1, (u.tile + 1) * nkb);
      u.kb1 = end - u.tile * nkb;
      if (u.kb0 > 0)
        u.slice = 0, u.flag_idx = pair_id;
      else if (u.kb1 < nkb)
        u.slice = 1, u.flag_idx = pair_id + 1;
      else
        u.slice = -1, u.flag_idx = 0;
      u0 = end;
      return true;
    }
    if (g.split > 1) {
      if (i < 0) {  // the K slice of a tail tile goes first: the ordered adds finish under the whole tiles
        i = 0;
        if (pair_id < g.tail_tiles * g.split) {
          const int t = pair_id / g.split;
          u.slice = pair_id - t * g.split;
          u.tile = g.full_tiles + t;
          u.flag_idx = t;
          u.kb0 = (int)((long long)g.num_k_blocks * u.slice / g.split);
          u.kb1 = (int)((long long)g.num_k_blocks * (u.slice + 1) / g.split);
          return true;
        }
      }
      u.tile = pair_id + i * num_pairs;
      if (u.tile >= g.full_tiles) return false;
    } else {
      if (i < 0) i = 0;
      u.tile = pair_id + i * num_pairs;
      if (u.tile >= g.num_m_tiles * g.num_n_tiles) return false;
    }
    ++i;
    u.kb0 = 0, u.kb1 = g.num_k_blocks, u.slice = -1, u.flag_idx = 0;
    return true;
  }
};

// Spin until *flag >= want (acquire, gpu scope), bounded like every other wait of the library.
__device__ __forceinline__ void wait_flag_ge(const unsigned* flag, unsigned want) {
  uint32_t spins = 0, t0 = 0;
  for (;;) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
    if (v >= want) return;
    __nanosleep(32);
    if ((++spins & 0xfffu) == 0) {
      uint32_t tt;
      asm volatile("mov.u32 %0, %%globaltimer_hi;" : "=r"(tt));
      if (t0 == 0)
        t0 = tt | 0x80000000u;
      else if (((tt | 0x80000000u) - t0) >= WM_MBAR_TIMEOUT_TICKS)
        __trap();
    }
  }
}

// Exact-erf GELU (nn.GELU() default, mlp.py:16,25) with erf from Abramowitz-Stegun 7.1.26 (|err| <= 1.5e-7,
// far below the bf16 rounding of the output): Phi(x) = 1 - h for x >= 0, h for x < 0,
// h = 0.5 * (a1 t + ... + a5 t^5) * exp(-x^2 / 2), t = 1 / (1 + p |x| / sqrt(2)).
__device__ __forceinline__ float gelu_erf(float x) {
  const float z = fabsf(x) * 0.70710678118654752f;
  const float t = rcp_approx(fmaf(z, 0.3275911f, 1.0f));
  float poly = fmaf(t, 1.061405429f, -1.453152027f);
  poly = fmaf(poly, t, 1.421413741f);
  poly = fmaf(poly, t, -0.284496736f);
  poly = fmaf(poly, t, 0.254829592f);
  const float h = 0.5f * poly * t * ex2_approx(z * z * -1.4426950408889634f);
  return x * (x >= 0.f ? 1.0f - h : h);
}

__device__ __forceinline__ void add_vec64(float (&x)[64], const float* __restrict__ b) {
  const float4* b4 = reinterpret_cast<const float4*>(b);
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const float4 v = __ldg(b4 + i);
    x[4 * i] += v.x, x[4 * i + 1] += v.y, x[4 * i + 2] += v.z, x[4 * i + 3] += v.w;
  }
}

// LayerNorm over the 64 values of one head held by one thread (attention.py:41-42,52; eps 1e-5, affine).
__device__ __forceinline__ void head_layernorm(float (&x)[64], const float* __restrict__ w,
                                               const float* __restrict__ b, float eps) {
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
  for (int i = 0; i < 64; i += 4) s0 += x[i], s1 += x[i + 1], s2 += x[i + 2], s3 += x[i + 3];
  const float mu = ((s0 + s1) + (s2 + s3)) * (1.0f / 64.0f);
  float v0 = 0.f, v1 = 0.f, v2 = 0.f, v3 = 0.f;
#pragma unroll
  for (int i = 0; i < 64; i += 4) {
    const float d0 = x[i] - mu, d1 = x[i + 1] - mu, d2 = x[i + 2] - mu, d3 = x[i + 3] - mu;
    v0 = fmaf(d0, d0, v0), v1 = fmaf(d1, d1, v1), v2 = fmaf(d2, d2, v2), v3 = fmaf(d3, d3, v3);
  }
  const float rstd = rsqrtf(((v0 + v1) + (v2 + v3)) * (1.0f / 64.0f) + eps);
  const float4* w4 = reinterpret_cast<const float4*>(w);
  const float4* b4 = reinterpret_cast<const float4*>(b);
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const float4 ww = __ldg(w4 + i), bb = __ldg(b4 + i);
    x[4 * i] = (x[4 * i] - mu) * rstd * ww.x + bb.x;
    x[4 * i + 1] = (x[4 * i + 1] - mu) * rstd * ww.y + bb.y;
    x[4 * i + 2] = (x[4 * i + 2] - mu) * rstd * ww.z + bb.z;
    x[4 * i + 3] = (x[4 * i + 3] - mu) * rstd * ww.w + bb.w;
  }
}

// 2D RoPE (rope.py:148-181): dims [0,32) rotate by y, [32,64) by x; inside each half the pair is (i, i+16).
__device__ __forceinline__ void head_rope(float (&x)[64], const float* __restrict__ cs, const float* __restrict__ sn,
                                          int py, int px) {
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    const int p = half == 0 ? py : px;
    const float4* c4 = reinterpret_cast<const float4*>(cs + p * 16);
    const float4* s4 = reinterpret_cast<const float4*>(sn + p * 16);
#pragma unroll
    for (int i4 = 0; i4 < 4; ++i4) {
      const float4 co = __ldg(c4 + i4), si = __ldg(s4 + i4);
      const float cc[4] = {co.x, co.y, co.z, co.w}, ss[4] = {si.x, si.y, si.z, si.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int i = 4 * i4 + k;
        const float a = x[half * 32 + i], bb = x[half * 32 + 16 + i];
        x[half * 32 + i] = a * cc[k] - bb * ss[k];
        x[half * 32 + 16 + i] = bb * cc[k] + a * ss[k];
      }
    }
  }
}

template <int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
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
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tempty + 2);

  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5;
  const uint32_t rank = cluster_ctarank();
  const int pair_id = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC0);
    if (EPI == EPI_QKV) tma_prefetch_desc(&tmC1);
  }
  if (warp == 1 && elect_one()) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 2);   // leader only: one arrive per producer of the pair + 64 KB of TMA bytes
      mbar_init(&empty[i], 1);  // per CTA: the leader's MMA commit is multicast to both
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);     // per CTA: multicast commit of the tile's last MMA
      mbar_init(&tempty[i], 16);   // leader only: one arrival per epilogue warp of both CTAs
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc_pair(tmem_ptr, 512);
    tmem_relinquish_pair();
  }
  tcgen05_fence_before();
  __syncthreads();
  cluster_sync_all();  // the peer's barriers are initialised before anything arrives on them remotely
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  pdl_wait();  // the predecessor's outputs (A, the residual stream) are complete and visible from here on

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (elect_one()) {
      uint32_t it = 0;
      GemmSched sched(pair_id, num_pairs);
      for (GemmUnit u; sched.next(g, u);) {
        const int m0 = (u.tile / g.num_n_tiles) * (2 * BM) + rank * BM;
        const int n0 = (u.tile % g.num_n_tiles) * BN + rank * (BN / 2);
        for (int kb = u.kb0; kb < u.kb1; ++kb, ++it) {
          const uint32_t s = it % STAGES, ph = (it / STAGES) & 1;
          mbar_wait(&empty[s], ph ^ 1);
          const uint32_t lead_full = mapa_shared(smem_u32(&full[s]), 0);
          if (rank == 0)
            mbar_arrive_expect_tx(&full[s], 2 * (A_BYTES + B_BYTES));
          else
            mbar_arrive_cluster(lead_full);
          tma_load_2d_pair(sA + s * A_BYTES, &tmA, lead_full, kb * BK, m0, kEvictNormal);
          tma_load_2d_pair(sB + s * B_BYTES, &tmB, lead_full, kb * BK, n0, kEvictLast);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (leader CTA of the pair)
    if (rank == 0 && elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(2 * BM, BN, false, false);
      uint32_t it = 0, tcount = 0;
      GemmSched sched(pair_id, num_pairs);
      for (GemmUnit u; sched.next(g, u); ++tcount) {
        const uint32_t acc = tcount & 1, aph = (tcount >> 1) & 1;
        mbar_wait(&tempty[acc], aph ^ 1);
        tcgen05_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = u.kb0; kb < u.kb1; ++kb, ++it) {
          const uint32_t s = it % STAGES, ph = (it / STAGES) & 1;
          mbar_wait(&full[s], ph);
          tcgen05_fence_after();
          const uint64_t adesc = umma_desc_sw128(smem_u32(sA + s * A_BYTES), 16, 1024);
          const uint64_t bdesc = umma_desc_sw128(smem_u32(sB + s * B_BYTES), 16, 1024);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)  // +32 bytes (>>4 = 2) per 16-element K step inside the 128B swizzle row
            umma_ss_pair(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, ((kb - u.kb0) | k) != 0);
          umma_commit_pair(&empty[s], 3);
        }
        umma_commit_pair(&tfull[acc], 3);
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ------------------------------------------------------------ epilogue: two teams of 128 threads, team t owns
    // accumulator columns [128 t, 128 t + 128); thread == accumulator row; one staging tile per team
    const int team = (warp - 4) >> 2;
    const int et = (threadIdx.x - 128) & 127;
    const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
    uint8_t* stg_ptr = sC + team * STG_BYTES;
    const uint32_t stg = smem_u32(stg_ptr);
    const uint32_t bar_id = 1 + team;
    uint32_t tcount = 0;
    const uint32_t lead_tempty[2] = {mapa_shared(smem_u32(&tempty[0]), 0), mapa_shared(smem_u32(&tempty[1]), 0)};

    GemmSched sched(pair_id, num_pairs);
    for (GemmUnit u; sched.next(g, u); ++tcount) {
      const int tile = u.tile;
      const int m0 = (tile / g.num_n_tiles) * (2 * BM) + rank * BM;
      const int n0 = (tile % g.num_n_tiles) * BN + team * 128;
      const uint32_t acc = tcount & 1, aph = (tcount >> 1) & 1;
      mbar_wait(&tfull[acc], aph);
      tcgen05_fence_after();
      const uint32_t taddr = tmem_base + acc * BN + team * 128 + lane_base;
      const int row = m0 + et;

      if (EPI == EPI_BIAS || EPI == EPI_GELU || EPI == EPI_QKV) {
        // ---- bf16 outputs: two 64-column groups per team (one attention head each for EPI_QKV)
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
        for (int grp = 0; grp < 2; ++grp) {
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
          if (grp == 1) {  // this thread's share of the accumulator is in registers: hand TMEM back to the MMA warp
            tcgen05_fence_before();
            __syncwarp();
            if ((threadIdx.x & 31) == 0) mbar_arrive_cluster(lead_tempty[acc]);
          }
          const int col0 = n0 + grp * 64;
          add_vec64(x, g.bias + col0);
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
              if (which == 0 && g.q_scale != 1.0f) {
#pragma unroll
                for (int i = 0; i < 64; ++i) x[i] *= g.q_scale;
              }
              if (g.norm2 != nullptr) {
                // the squared length of this thread's q / k head row; the largest of the launch decides on the device
                // which attention kernel runs (AttnGuard).  Bit patterns of non-negative floats order like the floats,
                // a NaN (either sign) sorts above +inf and survives the maximum.
                float n2 = 0.f;
#pragma unroll
                for (int i = 0; i < 64; ++i) n2 = fmaf(x[i], x[i], n2);
                unsigned nb = row < g.M ? __float_as_uint(n2) : 0u;
                nb = __reduce_max_sync(0xffffffffu, nb);
                unsigned* word = g.norm2 + which * 16 + ((col0 & (kEmbed - 1)) >> 6);
                if ((threadIdx.x & 31) == 0 && nb > __ldcg(word)) atomicMax(word, nb);
              }
            }
          }
          if (et == 0) tma_store_wait_read<0>();  // the team's previous store has drained the staging tile
          named_bar_sync(bar_id, 128);
#pragma unroll
          for (int c = 0; c < 8; ++c)
            st_shared_v4(stg + sw128_offset(et, c), pack_bf16(x[8 * c], x[8 * c + 1]),
                         pack_bf16(x[8 * c + 2], x[8 * c + 3]), pack_bf16(x[8 * c + 4], x[8 * c + 5]),
                         pack_bf16(x[8 * c + 6], x[8 * c + 7]));
          fence_proxy_async_smem();
          named_bar_sync(bar_id, 128);
          if (et == 0) {
            if (EPI == EPI_QKV && col0 >= kEmbed)
              tma_store_2d(&tmC1, stg_ptr, col0 - kEmbed, m0);
            else
              tma_store_2d(&tmC0, stg_ptr, col0, m0);
            tma_store_commit();
          }
        }
      } else {
        // ---- fp32 outputs: four 32-column chunks per team.  A K slice of a tail tile (u.slice >= 0) adds its share of the
        // product to x with the same TMA reduce-add, in SLICE ORDER: slice s waits until slice s-1 has published that its
        // reduce-adds are complete (x = ((x + g (p0 + b)) + g p1) + g p2, one fixed order: bit-deterministic), and only
        // slice 0 carries the bias.  The slices are the first unit of distinct, co-resident pairs, and the MMA warp runs
        // on with the pair's next tile in the other accumulator while the epilogue team waits.
        const bool sliced = EPI == EPI_RESID && u.slice >= 0;
        const bool later = sliced && u.slice > 0;
        unsigned* flag = sliced ? g.flags + (u.flag_idx * 2 + rank) * 2 + team : nullptr;
#pragma unroll 1
        for (int ch = 0; ch < 4; ++ch) {
          const int col0 = n0 + ch * 32;
          uint32_t v[32];
          tmem_ld_x32(taddr + ch * 32, v);
          tmem_wait_ld();
          if (ch == 3) {
            tcgen05_fence_before();
            __syncwarp();
            if ((threadIdx.x & 31) == 0) mbar_arrive_cluster(lead_tempty[acc]);
          }
          float o[32];
          if (EPI == EPI_TABLE) {
            const float4* t4 = reinterpret_cast<const float4*>(g.table + (size_t)(row % g.table_period) * g.N + col0);
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              const float4 t = __ldg(t4 + c);
              o[4 * c] = __uint_as_float(v[4 * c]) + t.x, o[4 * c + 1] = __uint_as_float(v[4 * c + 1]) + t.y;
              o[4 * c + 2] = __uint_as_float(v[4 * c + 2]) + t.z, o[4 * c + 3] = __uint_as_float(v[4 * c + 3]) + t.w;
            }
          } else {
            const float4* b4 = reinterpret_cast<const float4*>(g.bias + col0);
            const float4* g4 = reinterpret_cast<const float4*>(g.gamma + col0);
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
              if (!later) b = __ldg(b4 + c);
              float4 gm = make_float4(1.f, 1.f, 1.f, 1.f);
              if (g.gamma != nullptr) gm = __ldg(g4 + c);
              o[4 * c] = gm.x * (__uint_as_float(v[4 * c]) + b.x), o[4 * c + 1] = gm.y * (__uint_as_float(v[4 * c + 1]) + b.y);
              o[4 * c + 2] = gm.z * (__uint_as_float(v[4 * c + 2]) + b.z), o[4 * c + 3] = gm.w * (__uint_as_float(v[4 * c + 3]) + b.w);
            }
          }
          if (et == 0) tma_store_wait_read<0>();
          named_bar_sync(bar_id, 128);
#pragma unroll
          for (int c = 0; c < 8; ++c)
            st_shared_v4(stg + sw128_offset(et, c), __float_as_uint(o[4 * c]), __float_as_uint(o[4 * c + 1]),
                         __float_as_uint(o[4 * c + 2]), __float_as_uint(o[4 * c + 3]));
          fence_proxy_async_smem();
          named_bar_sync(bar_id, 128);
          if (et == 0) {
            if (later && ch == 0) wait_flag_ge(flag, (unsigned)u.slice);
            if (EPI == EPI_RESID && (g.gamma != nullptr || later))
              tma_reduce_add_2d(&tmC0, stg_ptr, col0, m0);  // x += gamma * (acc + b), added at L2
            else
              tma_store_2d(&tmC0, stg_ptr, col0, m0);
            tma_store_commit();
          }
        }
        if (sliced && et == 0) {
          tma_store_wait<0>();  // the reduce-adds of this slice are complete (not only read from the staging tile)
          __threadfence();
          const unsigned next = (u.slice + 1 == g.split) ? 0u : (unsigned)(u.slice + 1);  // the last slice re-arms the word
          asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(flag), "r"(next) : "memory");
        }
      }
    }
    if (et == 0) tma_store_wait<0>();
  }

  tcgen05_fence_before();
  __syncthreads();
  cluster_sync_all();  // no CTA leaves (or frees TMEM) while its peer may still signal or write into it
  if (warp == 2) {
    tcgen05_fence_after();
    tmem_dealloc_pair(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------------------ host
static int g_sm_count = 0;
constexpr int kGemmMaxSplit = 4;  // more slices than this buy little (the tail is already < 1/4 wave) and lengthen the chain
// The ordered chain costs one reduce-add round trip per slice; it hides under the pair's next whole tile only when
// that tile is long enough.  Measured on B200 (profiles/r02_s3_selftest.log): K = 4096 (fc2) 76 -> 70 us, K = 1024
// (proj) 25 -> 38 us.  Tiles shorter than this many 64-wide K blocks are not split.
constexpr int kGemmSplitMinKBlocks = 32;
static int env_int(const char* name, int dflt) {
  const char* e = getenv(name);
  return e ? atoi(e) : dflt;
}

template <int EPI>
static wm_status launch(cudaStream_t st, const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& c0,
                        const CUtensorMap& c1, GemmArgs g) {
  if (g_sm_count == 0) WM_TRY(require_sm100(&g_sm_count));
  static int cfg_dev[16] = {0};  // the shared-memory attribute is per device, not per process
  int dev = 0;
  WM_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 16 || !cfg_dev[dev]) {
    WM_CUDA(cudaFuncSetAttribute(gemm_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM));
    if (dev >= 0 && dev < 16) cfg_dev[dev] = 1;
  }
  g.num_m_tiles = (g.M + 2 * BM - 1) / (2 * BM);  // pair tiles: 256 rows
  g.num_n_tiles = g.N / BN;
  g.num_k_blocks = (g.K + BK - 1) / BK;
  const int tiles = g.num_m_tiles * g.num_n_tiles;
  const int pairs = g_sm_count / 2;
  int grid = 2 * (tiles < pairs ? tiles : pairs);
  g.split = 1, g.full_tiles = tiles, g.tail_tiles = 0, g.streamk = 0;
  if (EPI == EPI_RESID && g.flags != nullptr) {
    // Measured (profiles/r02_s30_streamk.log): correct and bit-deterministic, but SLOWER at the bench shape: proj 25 -> 30 us,
    // fc2 67 (ordered tail split) -> 76 us; in the step 2.55 -> 3.0 and 5.24 -> 6.22 ms.  Every cut tile pays a second
    // epilogue pass (TMEM -> registers -> staging -> reduce-add), and with 2.3 tiles per pair that is 4 epilogues per pair
    // instead of 3 at most; only M = 5504, K = 4096 gains (43 -> 39 us).  Off by default.
    static const int streamk_on = env_int("WM_GEMM_STREAMK", 0);                      // experiment switches
    static const int max_split = env_int("WM_GEMM_MAX_SPLIT", kGemmMaxSplit);
    static const int min_kb = env_int("WM_GEMM_SPLIT_MIN_KB", kGemmSplitMinKBlocks);
    const int waves = (tiles + pairs - 1) / pairs;
    // stream-K when the last wave of whole tiles would be less than ~85 % full: every pair gets tiles * k blocks / pairs
    // units.  (172 tiles on 74 pairs: 2.32 tile times instead of 3.)  Needs tiles >= pairs (a tile is cut at most once)
    // and (pairs + 1) * 4 slice counters.
    if (streamk_on && tiles >= pairs && (long long)tiles * 100 < (long long)waves * pairs * 85 && g.num_k_blocks >= 4 &&
        (size_t)(pairs + 2) * 4 * sizeof(unsigned) <= kGemmFlagBytes) {
      g.streamk = 1, g.split = 2;
      grid = 2 * pairs;
    } else {
      // the tiles beyond the last full wave, cut along K so that they fill (most of) one more wave of pairs
      const int full = (tiles / pairs) * pairs, tail = tiles - full;
      int sp = tail > 0 ? pairs / tail : 1;
      if (sp > max_split) sp = max_split;
      if (sp > kGemmMaxSplit) sp = kGemmMaxSplit;
      if (sp > g.num_k_blocks) sp = g.num_k_blocks;
      if (g.num_k_blocks < min_kb) sp = 1;
      if (sp >= 2) {
        g.split = sp, g.full_tiles = full, g.tail_tiles = tail;
        grid = 2 * (full > 0 ? pairs : tail * sp);
      }
    }
  }
  return launch_pdl(gemm_kernel<EPI>, dim3(grid), dim3(GEMM_THREADS), GEMM_SMEM, st, a, b, c0, c1, g);
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
  WM_TRY(tmap_2d(tb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, W, N, K, ldw, BK, BN / 2));
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
                   const float* rope_sin, int tokens_per_view, int patch_start, int grid_w, float q_scale, int M,
                   void* q_out, void* kv_out, unsigned* norm2) {
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
  g.qn_w = qn_w, g.qn_b = qn_b, g.kn_w = kn_w, g.kn_b = kn_b, g.qk_eps = qk_eps, g.q_scale = q_scale;
  g.rope_cos = rope_cos, g.rope_sin = rope_sin;
  g.norm2 = norm2;
  g.tokens_per_view = tokens_per_view > 0 ? tokens_per_view : 1, g.patch_start = patch_start,
  g.grid_w = grid_w > 0 ? grid_w : 1;
  return launch<EPI_QKV>(st, ta, tb, tq, tkv, g);
}

wm_status gemm_ls_residual(cudaStream_t st, const void* A, int lda, const void* W, int ldw, const float* bias,
                           const float* gamma, int M, int N, int K, float* x, int ldx, unsigned* flags) {
  WM_TRY(check_common(A, W, M, N, K, lda, ldw));
  WM_CHECK(x && bias && ldx % 4 == 0 && ldx >= N, "gemm_ls_residual: bad x");
  if (g_sm_count == 0) WM_TRY(require_sm100(&g_sm_count));
  CUtensorMap ta, tb, tx;
  WM_TRY(operand_maps(&ta, &tb, A, lda, W, ldw, M, N, K));
  WM_TRY(tmap_2d(&tx, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, x, M, N, ldx, 32, 128));
  GemmArgs g{};
  g.M = M, g.N = N, g.K = K, g.bias = bias, g.gamma = gamma;
  g.flags = flags;
  return launch<EPI_RESID>(st, ta, tb, tx, tx, g);
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
