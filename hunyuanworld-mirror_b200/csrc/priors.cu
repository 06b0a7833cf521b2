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

namespace wm {

// ------------------------------------------------------------------------------------------------ pose / ray priors
// The rest of WorldMirror.extract_priors (worldmirror.py:218-251) for one scene in one launch, no host round trip:
//   poses = extrinsics_to_vector(normalize_poses(camera_pose[:, :3]))[:, :7]   (utils/priors.py:4-104,
//            utils/camera_utils.py:25-35, utils/rotation.py:41-97)   -> [S, 7] = [t in the unit cube, quat xyzw]
//   rays  = [fx / W, fy / H, cx / W, cy / H]                                  (worldmirror.py:243-249) -> [S, 4]
// S x 12 numbers: one CTA; the 5 % / 95 % position percentiles (more than 10 views) are exact order statistics found by
// rank counting (O(S^2), S <= 1024) and interpolated as torch.quantile does.
constexpr int PP_MAX_VIEWS = 1024;

__device__ __forceinline__ float pp_sanitize(float v, bool on) {
  if (!on) return v;
  if (v != v) return 0.f;
  if (v == INFINITY) return 1e6f;
  if (v == -INFINITY) return -1e6f;
  return v;
}

__global__ void __launch_bounds__(256) pose_prior_kernel(const float* __restrict__ extr, int extr_stride, const float* __restrict__ intr,
                                                         int S, float Hf, float Wf, float one_minus_2pad,
                                                         float* __restrict__ poses, float* __restrict__ rays) {
  __shared__ float pos[3][PP_MAX_VIEWS];
  __shared__ float stat[3][4];  // per axis: order statistics floor/ceil of the low and the high percentile
  __shared__ int s_bad;
  const int tid = threadIdx.x;
  if (tid == 0) s_bad = 0;
  __syncthreads();
  // normalize_poses :21-27: any NaN / inf anywhere in the scene's extrinsics -> nan_to_num on all of them
  int bad = 0;
  for (int i = tid; i < S * 12; i += blockDim.x) {
    const float v = extr[(size_t)(i / 12) * extr_stride + i % 12];
    bad |= !(fabsf(v) <= 3.402823466e38f);
  }
  if (bad) atomicOr(&s_bad, 1);
  __syncthreads();
  const bool fix = s_bad != 0;
  for (int s = tid; s < S; s += blockDim.x)
#pragma unroll
    for (int a = 0; a < 3; ++a) pos[a][s] = pp_sanitize(extr[(size_t)s * extr_stride + a * 4 + 3], fix);
  __syncthreads();
  // bounds :56-63: 5 % / 95 % percentiles when more than 10 views, else min / max
  int k[4];
  float w_lo = 0.f, w_hi = 0.f;
  if (S > 10) {
    const float r_lo = 0.05f * (float)(S - 1), r_hi = 0.95f * (float)(S - 1);
    w_lo = r_lo - floorf(r_lo), w_hi = r_hi - floorf(r_hi);
    k[0] = (int)floorf(r_lo), k[1] = (int)ceilf(r_lo), k[2] = (int)floorf(r_hi), k[3] = (int)ceilf(r_hi);
  } else {
    k[0] = k[1] = 0, k[2] = k[3] = S - 1;
  }
  for (int s = tid; s < S; s += blockDim.x) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      const float v = pos[a][s];
      int rank = 0;
      for (int j = 0; j < S; ++j) {
        const float u = pos[a][j];
        rank += (u < v) || (u == v && j < s);
      }
#pragma unroll
      for (int t = 0; t < 4; ++t)
        if (rank == k[t]) stat[a][t] = v;
    }
  }
  __syncthreads();
  float mn[3], mx[3];
  float scale = 0.f;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    mn[a] = nd_lerp(stat[a][0], stat[a][1], w_lo);
    mx[a] = nd_lerp(stat[a][2], stat[a][3], w_hi);
    const float eps = fmaxf(1e-6f, __fmul_rn(fabsf(mx[a]), 1e-6f));  // :69-73
    scale = fmaxf(scale, fmaxf(__fsub_rn(mx[a], mn[a]), eps));       // :74-77
  }
  scale = fminf(fmaxf(scale, 1e-6f), 1e6f);                          // :78
  const float actual = __fdiv_rn(scale, one_minus_2pad);             // :84
  for (int s = tid; s < S; s += blockDim.x) {
    const float* e = extr + (size_t)s * extr_stride;
    float m[9];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int c = 0; c < 3; ++c) m[r * 3 + c] = pp_sanitize(e[r * 4 + c], fix);
    float* out = poses + (size_t)s * 7;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      const float center = __fdiv_rn(__fadd_rn(mn[a], mx[a]), 2.0f);                        // :81
      const float p = __fadd_rn(__fdiv_rn(__fsub_rn(pos[a][s], center), actual), 0.5f);    // :85
      out[a] = fminf(fmaxf(p, 0.f), 1.f);                                                   // :88
    }
    // rotmat_to_quat (rotation.py:41-97): candidates for r, i, j, k; the best-conditioned one; xyzw; w >= 0
    const float m00 = m[0], m01 = m[1], m02 = m[2], m10 = m[3], m11 = m[4], m12 = m[5], m20 = m[6], m21 = m[7], m22 = m[8];
    float qa[4] = {__fadd_rn(__fadd_rn(__fadd_rn(1.0f, m00), m11), m22), __fsub_rn(__fsub_rn(__fadd_rn(1.0f, m00), m11), m22),
                   __fsub_rn(__fadd_rn(__fsub_rn(1.0f, m00), m11), m22), __fadd_rn(__fsub_rn(__fsub_rn(1.0f, m00), m11), m22)};
    int best = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) qa[i] = qa[i] > 0.f ? __fsqrt_rn(qa[i]) : 0.f;
#pragma unroll
    for (int i = 1; i < 4; ++i)
      if (qa[i] > qa[best]) best = i;
    const float cand[4][4] = {{__fmul_rn(qa[0], qa[0]), __fsub_rn(m21, m12), __fsub_rn(m02, m20), __fsub_rn(m10, m01)},
                              {__fsub_rn(m21, m12), __fmul_rn(qa[1], qa[1]), __fadd_rn(m10, m01), __fadd_rn(m02, m20)},
                              {__fsub_rn(m02, m20), __fadd_rn(m10, m01), __fmul_rn(qa[2], qa[2]), __fadd_rn(m12, m21)},
                              {__fsub_rn(m10, m01), __fadd_rn(m20, m02), __fadd_rn(m21, m12), __fmul_rn(qa[3], qa[3])}};
    float q[4];  // r, i, j, k
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float sel = 0.f, den = 0.f;
#pragma unroll
      for (int b = 0; b < 4; ++b)
        if (b == best) sel = cand[b][i], den = __fmul_rn(2.0f, fmaxf(qa[b], 0.1f));
      q[i] = __fdiv_rn(sel, den);
    }
    const float sgn = q[0] < 0.f ? -1.f : 1.f;
    out[3] = sgn * q[1], out[4] = sgn * q[2], out[5] = sgn * q[3], out[6] = sgn * q[0];
    if (intr != nullptr && rays != nullptr) {
      const float* K = intr + (size_t)s * 9;
      rays[s * 4 + 0] = __fdiv_rn(K[0], Wf), rays[s * 4 + 1] = __fdiv_rn(K[4], Hf);
      rays[s * 4 + 2] = __fdiv_rn(K[2], Wf), rays[s * 4 + 3] = __fdiv_rn(K[5], Hf);
    }
  }
}

wm_status pose_prior(cudaStream_t st, const float* extr, int extr_stride, const float* intr, int S, int H, int W, double padding,
                     float* poses, float* rays) {
  WM_CHECK(extr && poses && S > 0 && S <= PP_MAX_VIEWS && extr_stride >= 12 && H > 0 && W > 0 && (intr == nullptr) == (rays == nullptr),
           "pose_prior: bad argument (at most %d views)", PP_MAX_VIEWS);
  // `scale_factor / (1 - 2 * padding)`: the divisor is a Python double, rounded to fp32 when it meets the tensor
  pose_prior_kernel<<<1, 256, 0, st>>>(extr, extr_stride, intr, S, (float)H, (float)W, (float)(1.0 - 2.0 * padding), poses, rays);
  WM_LAUNCH_CHECK();
  return WM_OK;
}

}  // namespace wm
