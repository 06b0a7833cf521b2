// Shared pieces of the two flash-attention kernels (attn.cu: thread == query row, 8 softmax warps;
// attn_split.cu: thread == half a query row, 16 softmax warps): tile constants, TMEM column map, launch
// arguments and the FMA-pipe exp2.
#pragma once
#include "common.cuh"
#include "ptx.cuh"

namespace wm {

constexpr int ATT_THREADS = 384;
constexpr int ATT_TILE = 128 * 64 * 2;  // one 128 x 64 bf16 tile, 16 KB
constexpr int KV_ST = 3;
constexpr int ATT_SMEM = (4 + 2 * KV_ST + 2) * ATT_TILE + 512 + 1024;

struct AttnArgs {
  int Lq, Lkv, batch;
  int n_qpairs, n_kv, total_items;
  int n_cheap;  // > 0 (whole-item schedules only): items whose query pair holds ONE 128-row tile, dealt out last (attn_item)
  float scale_log2;
  // attn_split.cu only.  Work items are dealt out in `full_waves` strided waves (item = wave * grid + cta: CTAs that
  // run at the same time share a head's K|V in L2), then the remaining items' (item, kv tile) steps are split into
  // equal contiguous runs of `per_cta` steps, so a tail item may be shared by several CTAs; a CTA's share of a split
  // item leaves as an unnormalised fp32 partial (O [256][64], row sum l [256], reference max m [256]) in slot
  // 2 * cta + {0: first, 1: last tail segment of the CTA} and attn_merge_kernel adds the partials up in CTA order.
  int full_waves, per_cta, tail_steps, merge_grid;
  float *part_o, *part_l, *part_m;
  // attn_split.cu only: sequence-parallel global attention over a K|V buffer that peer GPUs are still filling
  // (capi.cu, wm_p2p_*).  Logical key tile j reads physical tile (j + kv_rot) % n_kv, so every rank starts with the
  // keys it produced itself; shard r (rows [r * shard_rows, (r+1) * shard_rows)) may be read once
  // (int)(flags[r] - epoch) >= 0.  shard_rows == 0: every key is present at launch (flags unused).
  int kv_rot, shard_rows, self_rank;
  unsigned epoch;
  const unsigned* flags;
  unsigned* err;  // set to 1 when a shard did not arrive within kGatherTimeoutNs (the launch then finishes on stale keys)
  // device-side kernel choice (AttnGuard, common.cuh); guard_mode == 0: none
  const unsigned* guard_norm2;
  int guard_mode, guard_heads;
  float guard_limit2;  // limit^2, compared with max_h |q|^2 |k|^2
};
constexpr unsigned long long kGatherTimeoutNs = 4000000000ull;

__device__ __forceinline__ int attn_phys_tile(const AttnArgs& a, int j) {
  const int jp = j + a.kv_rot;
  return jp >= a.n_kv ? jp - a.n_kv : jp;
}

// Producer side: block until every shard that key tile `jp` touches has arrived.  `have` caches the answer per shard.
__device__ __forceinline__ void attn_wait_shards(const AttnArgs& a, int jp, unsigned& have) {
  const int row0 = jp * 128, row1 = min(row0 + 127, a.Lkv - 1);
  for (int r = row0 / a.shard_rows; r <= row1 / a.shard_rows; ++r) {
    if ((have >> r) & 1u) continue;
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
      unsigned v;
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(a.flags + r) : "memory");
      if (static_cast<int>(v - a.epoch) >= 0) break;
      __nanosleep(256);
      unsigned long long t1;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
      if (t1 - t0 > kGatherTimeoutNs) {
        atomicExch(a.err, 1u);
        break;
      }
    }
    asm volatile("fence.proxy.async;" ::: "memory");  // the keys were written by a peer's copy engine; TMA reads them next
    have |= 1u << r;
  }
}

// True when this launch is the wrong one of the (bounded, running-max) pair for the data at hand: every thread of the
// grid computes the same answer from 2 * heads words and returns before any barrier or TMEM allocation.  Call after
// pdl_wait() (the words are written by the preceding QKV GEMM).  A NaN norm makes the bound fail.
__device__ __forceinline__ bool attn_guard_fits(const AttnArgs& a) {
  // lane h looks at head h: one round trip to L2 for the warp (a loop with an early-out costs one per head, ~12 us for 16
  // heads at the start of every CTA).  Call with all 32 lanes of the warp active; guard_heads <= 32.
  const int lane = threadIdx.x & 31;
  bool ok = true;
  if (lane < a.guard_heads) {
    const float q2 = __uint_as_float(__ldcg(a.guard_norm2 + lane));
    const float k2 = __uint_as_float(__ldcg(a.guard_norm2 + a.guard_heads + lane));
    ok = q2 * k2 <= a.guard_limit2;  // false for NaN
  }
  return __all_sync(0xffffffffu, ok);
}
__device__ __forceinline__ bool attn_guard_skip(const AttnArgs& a) {
  if (a.guard_mode != 1 && a.guard_mode != 2) return false;
  const bool fits = attn_guard_fits(a);
  return a.guard_mode == 1 ? !fits : fits;
}

// Schedule position -> work item.  When the last query pair of every (head, batch entry) holds a single 128-row tile
// (Lq mod 256 in 1..128: 1376 and 1374 rows per view are such lengths) that item costs about half of a full one.  Whole-item
// schedules deal the full items out first and the cheap ones last, and the tail wave goes to the CTAs that drew cheap
// items in the last full wave (AttnSched), so the longest CTA runs 4 full + 2 cheap items instead of 6 full ones at 8
// views x 16 heads on 148 SMs.
__device__ __forceinline__ int attn_item(const AttnArgs& a, int pos) {
  if (a.n_cheap == 0) return pos;
  const int n_full = a.total_items - a.n_cheap, qp_full = a.n_qpairs - 1;
  if (pos < n_full) {
    const int hb = pos / qp_full;
    return hb * a.n_qpairs + (pos - hb * qp_full);
  }
  return (pos - n_full) * a.n_qpairs + qp_full;
}

// The segment sequence of one CTA; producer, MMA issuer and softmax warps walk it in lockstep.
struct AttnSched {
  int w, tstep, tbegin, tend;
  __device__ explicit AttnSched(const AttnArgs& a) : w(0) {
    // whole-item schedules hand the tail items to the LAST CTAs (see attn_item); split schedules keep CTA order, which
    // attn_merge_kernel relies on
    const int c = a.part_o == nullptr ? (int)(gridDim.x - 1 - blockIdx.x) : (int)blockIdx.x;
    tbegin = tstep = c * a.per_cta;
    tend = min(tbegin + a.per_cta, a.tail_steps);
  }
  __device__ bool next(const AttnArgs& a, int& item, int& kv0, int& kv1, int& slot) {
    if (w < a.full_waves) {
      item = attn_item(a, w * gridDim.x + blockIdx.x);
      ++w;
      kv0 = 0, kv1 = a.n_kv, slot = -1;
      return true;
    }
    if (tstep >= tend) return false;
    const int ti = tstep / a.n_kv;
    kv0 = tstep - ti * a.n_kv;
    kv1 = min(a.n_kv, kv0 + (tend - tstep));
    item = attn_item(a, a.full_waves * gridDim.x + ti);
    slot = (kv0 == 0 && kv1 == a.n_kv) ? -1 : 2 * blockIdx.x + (tstep == tbegin ? 0 : 1);
    tstep += kv1 - kv0;
    return true;
  }
};
constexpr size_t kAttnSlotFloats = 256 * 64 + 256 + 256;  // floats per partial slot (O, l, m)

__host__ __device__ constexpr uint32_t s_col(int w) { return 128u * w; }         // S0 / S1
__host__ __device__ constexpr uint32_t o_col(int w) { return 256u + 64u * w; }   // O0 / O1
__host__ __device__ constexpr uint32_t p_col(int w) { return 384u + 64u * w; }   // P0 / P1 (bf16 pairs)

// 2^x for x <= 0 on the FMA pipe, two lanes at a time: x = n + f with n = round(x), f in [-0.5, 0.5];
// 2^f by a degree-3 minimax polynomial, 2^n by adding n to the exponent field.
__device__ __forceinline__ float2 exp2_poly2(float2 x) {
  const float2 magic = make_float2(12582912.0f, 12582912.0f);  // 1.5 * 2^23: the add rounds x to an integer
  x.x = fmaxf(x.x, -126.0f);
  x.y = fmaxf(x.y, -126.0f);
  const float2 xr = fadd2(x, magic);
  const float2 n = fadd2(xr, make_float2(-12582912.0f, -12582912.0f));
  const float2 f = ffma2(n, make_float2(-1.0f, -1.0f), x);
  float2 p = ffma2(make_float2(0.05520551f, 0.05520551f), f, make_float2(0.24261397f, 0.24261397f));
  p = ffma2(p, f, make_float2(0.69325477f, 0.69325477f));
  p = ffma2(p, f, make_float2(0.9999277f, 0.9999277f));
  float2 r;
  r.x = __uint_as_float(__float_as_uint(p.x) + (__float_as_uint(xr.x) << 23));
  r.y = __uint_as_float(__float_as_uint(p.y) + (__float_as_uint(xr.y) << 23));
  return r;
}

// The same without the clamp, for arguments known to lie in [-126, 126] (bounded-logit mode).
__device__ __forceinline__ float2 exp2_poly2_bounded(float2 x) {
  const float2 magic = make_float2(12582912.0f, 12582912.0f);
  const float2 xr = fadd2(x, magic);
  const float2 n = fadd2(xr, make_float2(-12582912.0f, -12582912.0f));
  const float2 f = ffma2(n, make_float2(-1.0f, -1.0f), x);
  float2 p = ffma2(make_float2(0.05520551f, 0.05520551f), f, make_float2(0.24261397f, 0.24261397f));
  p = ffma2(p, f, make_float2(0.69325477f, 0.69325477f));
  p = ffma2(p, f, make_float2(0.9999277f, 0.9999277f));
  float2 r;
  r.x = __uint_as_float(__float_as_uint(p.x) + (__float_as_uint(xr.x) << 23));
  r.y = __uint_as_float(__float_as_uint(p.y) + (__float_as_uint(xr.y) << 23));
  return r;
}

}  // namespace wm
