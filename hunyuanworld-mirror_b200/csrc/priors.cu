// Depth-prior normalisation (reference: src/models/utils/priors.py:107-167 normalize_depth),
// the part of WorldMirror.extract_priors (worldmirror.py:218-251) that touches every pixel.  The reference loops over
// views in Python, calls torch.quantile twice per view and reads a scalar back to the host (`scale.item()`) each time;
// here all views are normalised by ONE launch with no host round trip: one CTA per view finds the two order statistics
// that each percentile interpolates between with an exact 4-pass radix select over the float keys (8 bits per pass,
// all four targets per pass), then rescales the view.  HBM-bound: 6 reads + 1 write of 4 bytes per pixel, L2-resident
// after the first read (a 518 x 518 view is 1 MB).
//
// Semantics restated exactly (fp32 arithmetic in the same order as torch):
//   v = nan_to_num(depth, nan=0, posinf=1e6, neginf=0); use the values v > 0 if there are any, else all values;
//   more than 100 such values: lo = quantile(0.01), hi = quantile(0.99), torch.quantile's linear interpolation
//   (rank = q_f32 * (n - 1) in fp32, lerp between the floor and ceil order statistics); otherwise min / max;
//   hi == lo: hi = lo + 1;  eps = max(1e-6, |hi - lo| * 1e-6) (a Python double, rounded to fp32 when added);
//   out = clamp((v - lo) / (hi - lo + eps), 0, 1).
#include "common.cuh"
#include "ptx.cuh"

namespace wm {

constexpr int ND_THREADS = 1024;
constexpr int ND_COPIES = 8;  // privatised histograms: warps w, w + 8, ... share copy w % 8

__device__ __forceinline__ float nd_sanitize(float v) {
  if (v != v) return 0.f;
  if (v == INFINITY) return 1e6f;
  if (v == -INFINITY) return 0.f;
  return v;
}
// order-preserving map float -> uint32
__device__ __forceinline__ uint32_t nd_key(float v) {
  const uint32_t b = __float_as_uint(v);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float nd_unkey(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}
// torch's lerp for floats (aten/native/Lerp.h): weight < 0.5 ? a + w (b - a) : b - (b - a) (1 - w)
__device__ __forceinline__ float nd_lerp(float a, float b, float w) {
  const float d = __fsub_rn(b, a);  // explicit roundings: no fma contraction, like the reference's scalar CPU path
  return w < 0.5f ? __fadd_rn(a, __fmul_rn(w, d)) : __fsub_rn(b, __fmul_rn(d, __fsub_rn(1.0f, w)));
}

__global__ void __launch_bounds__(ND_THREADS) normalize_depth_kernel(const float* __restrict__ depth, int hw, float q_lo, float q_hi,
                                                                     double eps, float* __restrict__ out) {
  __shared__ unsigned hist[ND_COPIES][4][256];
  __shared__ unsigned s_prefix[4], s_rank[4];
  __shared__ unsigned s_red[32];
  __shared__ float s_lohi[2];
  __shared__ unsigned s_npos;
  const float* src = depth + (size_t)blockIdx.x * hw;
  float* dst = out + (size_t)blockIdx.x * hw;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  // ---- count the positive values
  unsigned cnt = 0;
  for (int i = tid; i < hw; i += ND_THREADS) cnt += nd_sanitize(__ldg(src + i)) > 0.f;
  cnt = __reduce_add_sync(0xffffffffu, cnt);
  if (lane == 0) s_red[warp] = cnt;
  __syncthreads();
  if (tid == 0) {
    unsigned t = 0;
    for (int w = 0; w < ND_THREADS / 32; ++w) t += s_red[w];
    s_npos = t;
  }
  __syncthreads();
  const unsigned npos = s_npos;
  const bool only_pos = npos > 0;
  const unsigned n = only_pos ? npos : (unsigned)hw;

  // ---- the four order statistics (or min / max): ranks as torch.quantile computes them, in fp32
  float w_lo = 0.f, w_hi = 0.f;
  if (n > 100) {
    const float r_lo = q_lo * (float)(n - 1), r_hi = q_hi * (float)(n - 1);
    w_lo = r_lo - floorf(r_lo), w_hi = r_hi - floorf(r_hi);
    if (tid == 0) {
      s_rank[0] = (unsigned)floorf(r_lo), s_rank[1] = (unsigned)ceilf(r_lo);
      s_rank[2] = (unsigned)floorf(r_hi), s_rank[3] = (unsigned)ceilf(r_hi);
    }
  } else if (tid == 0) {
    s_rank[0] = s_rank[1] = 0, s_rank[2] = s_rank[3] = n - 1;  // min and max
  }
  if (tid < 4) s_prefix[tid] = 0;
  __syncthreads();

  for (int pass = 0; pass < 4; ++pass) {
    const int shift = 24 - 8 * pass;
    const uint32_t mask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
    for (int i = tid; i < ND_COPIES * 4 * 256; i += ND_THREADS) (&hist[0][0][0])[i] = 0;
    __syncthreads();
    const uint32_t p0 = s_prefix[0], p1 = s_prefix[1], p2 = s_prefix[2], p3 = s_prefix[3];
    unsigned(*h)[256] = hist[warp % ND_COPIES];
    for (int i = tid; i < hw; i += ND_THREADS) {
      const float v = nd_sanitize(__ldg(src + i));
      if (only_pos && !(v > 0.f)) continue;
      const uint32_t k = nd_key(v), bin = (k >> shift) & 255u;
      if (((k ^ p0) & mask) == 0) atomicAdd(&h[0][bin], 1u);
      if (((k ^ p1) & mask) == 0) atomicAdd(&h[1][bin], 1u);
      if (((k ^ p2) & mask) == 0) atomicAdd(&h[2][bin], 1u);
      if (((k ^ p3) & mask) == 0) atomicAdd(&h[3][bin], 1u);
    }
    __syncthreads();
    if (tid < 4) {  // one thread per target walks its 256 bins
      unsigned rank = s_rank[tid], cum = 0;
      int bin = 0;
      for (; bin < 256; ++bin) {
        unsigned c = 0;
        for (int cpy = 0; cpy < ND_COPIES; ++cpy) c += hist[cpy][tid][bin];
        if (cum + c > rank) break;
        cum += c;
      }
      bin = min(bin, 255);
      s_rank[tid] = rank - cum;
      s_prefix[tid] |= (uint32_t)bin << shift;
    }
    __syncthreads();
  }
  if (tid == 0) {
    float lo = nd_lerp(nd_unkey(s_prefix[0]), nd_unkey(s_prefix[1]), w_lo);
    float hi = nd_lerp(nd_unkey(s_prefix[2]), nd_unkey(s_prefix[3]), w_hi);
    if (hi == lo) hi = lo + 1.0f;
    s_lohi[0] = lo, s_lohi[1] = hi;
  }
  __syncthreads();
  const float lo = s_lohi[0], hi = s_lohi[1];
  // eps_val = max(eps, scale.item() * eps) is a Python double; adding it to the fp32 tensor rounds it to fp32 first
  const double scale = fabs((double)(hi - lo));
  const float eps_f = (float)fmax(eps, scale * eps);
  const float denom = (hi - lo) + eps_f;
  for (int i = tid; i < hw; i += ND_THREADS) {
    const float v = nd_sanitize(__ldg(src + i));
    dst[i] = fminf(fmaxf(__fdiv_rn(v - lo, denom), 0.f), 1.f);
  }
}

wm_status normalize_depth(cudaStream_t st, const float* depth, int n_views, int hw, double eps, int min_percentile,
                          int max_percentile, float* out) {
  WM_CHECK(depth && out && n_views > 0 && hw > 0 && min_percentile >= 0 && max_percentile <= 100 && min_percentile <= max_percentile,
           "normalize_depth: bad argument");
  // min_percentile / 100.0 is a Python double; torch.quantile turns it into an fp32 tensor
  const float q_lo = (float)((double)min_percentile / 100.0), q_hi = (float)((double)max_percentile / 100.0);
  normalize_depth_kernel<<<n_views, ND_THREADS, 0, st>>>(depth, hw, q_lo, q_hi, eps, out);
  WM_LAUNCH_CHECK();
  return WM_OK;
}

}  // namespace wm
