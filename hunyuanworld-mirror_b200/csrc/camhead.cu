// Camera head (reference: src/models/heads/camera_head.py:58-104): 4 refinement steps of
// a 4-block transformer trunk at width 2048 over ONE token per view.  M = number of views (8 .. 128) rows against
// 201 M fp32 trunk parameters: every linear layer is a skinny GEMM whose cost is streaming the weight matrix once
// from HBM, so the kernels below are fp32 SIMT (the reference runs this head in fp32, worldmirror.py:146, and the
// 1e-3 rad rotation tolerance of BASELINE.json leaves no room for reduced precision) and are sized for bandwidth,
// not for the tensor cores.  Algorithmic bytes: 4 * N * K per linear layer call.
#include "common.cuh"
#include "ptx.cuh"

namespace wm {

enum : int { CAM_ACT_NONE = 0, CAM_ACT_GELU = 1 };

// weights are read exactly once per call: bypass L1 allocation
__device__ __forceinline__ float4 ld_stream(const float4* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ float silu(float x) { return x / (1.0f + __expf(-x)); }
__device__ __forceinline__ float gelu_exact(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f)); }

// out[m, n] = epi( sum_k f(A[m, k]) * W[n, k] + bias[n] ),  f = SiLU when silu_in (adapt_norm_gen = Sequential(SiLU, Linear),
// camera_head.py:49), epi = GELU (Mlp.fc1, mlp.py:30-31) or, with resid != null, resid[m, n] + gamma[n] * (.)
// (LayerScale + residual, block.py:91-92; in place when out == resid).
// CTA = 4 warps x 2 output columns = 8 columns of W, 8 rows of A; lane l covers k = 128 i + 4 l .. + 3 (512 B coalesced
// per W row per step); grid = (ceil(M / 8), N / 8).  A is re-read from L1/L2 by the warps (8 x K fp32 <= 256 KB).
// The k loop is unrolled 4x so that every lane has 8 independent 16-byte weight loads in flight (Little: 6.5 TB/s x
// ~1 us needs ~44 KB outstanding per SM; the first version kept 8 KB in flight and measured 1.45 TB/s).
constexpr int SK_ROWS = 8, SK_COLS = 2, SK_WARPS = 4, SK_UNROLL = 4;
__global__ void __launch_bounds__(32 * SK_WARPS) skinny_linear_kernel(const float* __restrict__ A, int lda, const float* __restrict__ W,
                                                                      const float* __restrict__ bias, int M, int N, int K, int silu_in,
                                                                      int act, const float* resid, const float* __restrict__ gamma,
                                                                      float* out, int ldo) {
  pdl_launch_dependents();
  pdl_wait();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // blockIdx.x (fastest in launch order) = row tile: the CTAs that share a set of weight rows run together, so W comes
  // from HBM once and from L2 for the other row tiles (M = 64 views: 8 row tiles)
  const int n0 = (blockIdx.y * SK_WARPS + warp) * SK_COLS;
  const int m0 = blockIdx.x * SK_ROWS;
  if (n0 >= N) return;
  float acc[SK_ROWS][SK_COLS];
#pragma unroll
  for (int r = 0; r < SK_ROWS; ++r)
#pragma unroll
    for (int c = 0; c < SK_COLS; ++c) acc[r][c] = 0.f;
  const float4* w4[SK_COLS];
#pragma unroll
  for (int c = 0; c < SK_COLS; ++c) w4[c] = reinterpret_cast<const float4*>(W + (size_t)min(n0 + c, N - 1) * K);
  for (int kb = lane; kb < K / 4; kb += 32 * SK_UNROLL) {  // K % 128 == 0; chunks past the end are skipped below
    float4 wv[SK_UNROLL][SK_COLS];
#pragma unroll
    for (int u = 0; u < SK_UNROLL; ++u) {
      const int k4 = min(kb + 32 * u, K / 4 - 1);
#pragma unroll
      for (int c = 0; c < SK_COLS; ++c) wv[u][c] = ld_stream(w4[c] + k4);
    }
#pragma unroll
    for (int u = 0; u < SK_UNROLL; ++u) {
      const int k4 = kb + 32 * u;
      if (k4 < K / 4) {
#pragma unroll
        for (int r = 0; r < SK_ROWS; ++r) {
          const int m = min(m0 + r, M - 1);
          float4 a = __ldg(reinterpret_cast<const float4*>(A + (size_t)m * lda) + k4);
          if (silu_in) a = make_float4(silu(a.x), silu(a.y), silu(a.z), silu(a.w));
#pragma unroll
          for (int c = 0; c < SK_COLS; ++c)
            acc[r][c] = fmaf(a.x, wv[u][c].x, fmaf(a.y, wv[u][c].y, fmaf(a.z, wv[u][c].z, fmaf(a.w, wv[u][c].w, acc[r][c]))));
        }
      }
    }
  }
#pragma unroll
  for (int r = 0; r < SK_ROWS; ++r)
#pragma unroll
    for (int c = 0; c < SK_COLS; ++c) {
      float v = acc[r][c];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      acc[r][c] = v;
    }
  // lane r * SK_COLS + c writes element (r, c)
  if (lane < SK_ROWS * SK_COLS) {
    const int r = lane / SK_COLS, c = lane % SK_COLS;
    float v = 0.f;
#pragma unroll
    for (int rr = 0; rr < SK_ROWS; ++rr)
#pragma unroll
      for (int cc = 0; cc < SK_COLS; ++cc)
        if (rr == r && cc == c) v = acc[rr][cc];
    const int m = m0 + r, n = n0 + c;
    if (m < M && n < N) {
      v += bias != nullptr ? __ldg(bias + n) : 0.f;
      if (act == CAM_ACT_GELU) v = gelu_exact(v);
      if (resid != nullptr) v = resid[(size_t)m * ldo + n] + __ldg(gamma + n) * v;
      out[(size_t)m * ldo + n] = v;
    }
  }
}

wm_status skinny_linear(cudaStream_t st, const float* A, int lda, const float* W, const float* bias, int M, int N, int K, int silu_in,
                        int act, const float* resid, const float* gamma, float* out, int ldo) {
  WM_CHECK(A && W && out && M > 0 && N > 0 && K > 0 && K % 128 == 0 && lda % 4 == 0 && lda >= K && ldo >= N,
           "skinny_linear: bad argument (M=%d N=%d K=%d)", M, N, K);
  WM_CHECK((resid == nullptr) == (gamma == nullptr), "skinny_linear: resid and gamma go together");
  const dim3 grid((M + SK_ROWS - 1) / SK_ROWS, (N + SK_WARPS * SK_COLS - 1) / (SK_WARPS * SK_COLS));
  WM_CHECK(grid.y <= 65535, "skinny_linear: N too large");
  return launch_pdl(skinny_linear_kernel, grid, dim3(32 * SK_WARPS), 0, st, A, lda, W, bias, M, N, K, silu_in, act, resid, gamma, out, ldo);
}

// LayerNorm over C = 2048 in fp32, one warp per row; input rows `in_stride` floats apart (the camera token is row 0 of
// every view of token_list[-1], camera_head.py:72-73).  w / b null: no affine (adapt_norm, camera_head.py:52).
// mod != null: adaLN modulation (camera_head.py:81-83): out = gate * (LN(x) * (1 + scale) + shift) + x with
// mod[m] = [shift | scale | gate] (chunk(3), each C wide).
__global__ void __launch_bounds__(256) cam_ln_kernel(const float* __restrict__ x, long long in_stride, const float* __restrict__ w,
                                                     const float* __restrict__ b, float eps, int M, const float* __restrict__ mod,
                                                     float* __restrict__ y) {
  constexpr int C = 2048, V = C / 128;
  pdl_launch_dependents();
  pdl_wait();
  const int m = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (m >= M) return;
  const float4* xr = reinterpret_cast<const float4*>(x + (size_t)m * in_stride);
  float4 v[V];
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < V; ++k) {
    v[k] = __ldg(xr + k * 32 + lane);
    s += (v[k].x + v[k].y) + (v[k].z + v[k].w);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float mu = s * (1.0f / C);
  float q = 0.f;
#pragma unroll
  for (int k = 0; k < V; ++k) {
    const float a = v[k].x - mu, bb = v[k].y - mu, c = v[k].z - mu, d = v[k].w - mu;
    q += (a * a + bb * bb) + (c * c + d * d);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
  const float rstd = rsqrtf(q * (1.0f / C) + eps);
  float4* yr = reinterpret_cast<float4*>(y + (size_t)m * C);
#pragma unroll
  for (int k = 0; k < V; ++k) {
    const int i = k * 32 + lane;
    float4 o = make_float4((v[k].x - mu) * rstd, (v[k].y - mu) * rstd, (v[k].z - mu) * rstd, (v[k].w - mu) * rstd);
    if (w != nullptr) {
      const float4 ww = __ldg(reinterpret_cast<const float4*>(w) + i), bb = __ldg(reinterpret_cast<const float4*>(b) + i);
      o = make_float4(o.x * ww.x + bb.x, o.y * ww.y + bb.y, o.z * ww.z + bb.z, o.w * ww.w + bb.w);
    }
    if (mod != nullptr) {
      const float4* mr = reinterpret_cast<const float4*>(mod + (size_t)m * 3 * C);
      const float4 sh = __ldg(mr + i), sc = __ldg(mr + C / 4 + i), gt = __ldg(mr + 2 * (C / 4) + i);
      o = make_float4(gt.x * (o.x * (1.f + sc.x) + sh.x) + v[k].x, gt.y * (o.y * (1.f + sc.y) + sh.y) + v[k].y,
                      gt.z * (o.z * (1.f + sc.z) + sh.z) + v[k].z, gt.w * (o.w * (1.f + sc.w) + sh.w) + v[k].w);
    }
    yr[i] = o;
  }
}

wm_status cam_layernorm(cudaStream_t st, const float* x, long long in_stride, const float* w, const float* b, float eps, int M,
                        const float* mod, float* y) {
  WM_CHECK(x && y && M > 0 && in_stride >= 2048 && in_stride % 4 == 0 && (w == nullptr) == (b == nullptr), "cam_layernorm: bad argument");
  return launch_pdl(cam_ln_kernel, dim3((M + 7) / 8), dim3(256), 0, st, x, in_stride, w, b, eps, M, mod, y);
}

// Attention over the S camera tokens (attention.py:48-69 without QK-norm / RoPE): qkv [S, 3 * 2048] = [q | k | v], 16 heads
// of 128.  One CTA per (query, head): 4 warps take the keys round-robin (lane = 4 dims), softmax over S scores in shared
// memory, then thread t accumulates output dim t.
__global__ void __launch_bounds__(128) cam_attn_kernel(const float* __restrict__ qkv, int S, float* __restrict__ out) {
  constexpr int C = 2048, D = 128;
  extern __shared__ float sc[];  // S scores + 4 partials
  pdl_launch_dependents();
  pdl_wait();
  const int qi = blockIdx.x, h = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float4 q = __ldg(reinterpret_cast<const float4*>(qkv + (size_t)qi * 3 * C + h * D) + lane);
  for (int j = warp; j < S; j += 4) {
    const float4 k = __ldg(reinterpret_cast<const float4*>(qkv + (size_t)j * 3 * C + C + h * D) + lane);
    float d = q.x * k.x + q.y * k.y + q.z * k.z + q.w * k.w;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
    if (lane == 0) sc[j] = d * 0.08838834764831845f;  // 1 / sqrt(128)
  }
  __syncthreads();
  float mx = -INFINITY;
  for (int j = threadIdx.x; j < S; j += 128) mx = fmaxf(mx, sc[j]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if (lane == 0) sc[S + warp] = mx;
  __syncthreads();
  mx = fmaxf(fmaxf(sc[S], sc[S + 1]), fmaxf(sc[S + 2], sc[S + 3]));
  __syncthreads();
  float sum = 0.f;
  for (int j = threadIdx.x; j < S; j += 128) {
    const float p = expf(sc[j] - mx);
    sc[j] = p;
    sum += p;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if (lane == 0) sc[S + warp] = sum;
  __syncthreads();
  const float inv = 1.0f / ((sc[S] + sc[S + 1]) + (sc[S + 2] + sc[S + 3]));
  float acc = 0.f;
  for (int j = 0; j < S; ++j) acc = fmaf(sc[j], __ldg(qkv + (size_t)j * 3 * C + 2 * C + h * D + threadIdx.x), acc);
  out[(size_t)qi * C + h * D + threadIdx.x] = acc * inv;
}

wm_status cam_attention(cudaStream_t st, const float* qkv, int S, float* out) {
  WM_CHECK(qkv && out && S > 0 && S <= 8192, "cam_attention: bad argument");
  return launch_pdl(cam_attn_kernel, dim3(S, 16), dim3(128), (size_t)(S + 4) * sizeof(float), st, qkv, S, out);
}

// out[m, n] = bias[n] + sum_k W[n, k] * in[m or 0, k], K <= 16 (param_embed = Linear(9, 2048), camera_head.py:47,77-80;
// in_bcast: the learned init_token is shared by all views).
__global__ void __launch_bounds__(256) cam_embed_kernel(const float* __restrict__ in, int in_bcast, int K, const float* __restrict__ W,
                                                        const float* __restrict__ bias, int M, int N, float* __restrict__ out) {
  pdl_launch_dependents();
  pdl_wait();
  const int n = blockIdx.x * 256 + threadIdx.x, m = blockIdx.y;
  if (n >= N || m >= M) return;
  const float* x = in + (in_bcast ? 0 : (size_t)m * K);
  float v = __ldg(bias + n);
  for (int k = 0; k < K; ++k) v = fmaf(__ldg(W + (size_t)n * K + k), __ldg(x + k), v);
  out[(size_t)m * N + n] = v;
}

wm_status cam_embed(cudaStream_t st, const float* in, int in_bcast, int K, const float* W, const float* bias, int M, int N, float* out) {
  WM_CHECK(in && W && bias && out && M > 0 && N > 0 && K > 0 && K <= 16, "cam_embed: bad argument");
  return launch_pdl(cam_embed_kernel, dim3((N + 255) / 256, M), dim3(256), 0, st, in, in_bcast, K, W, bias, M, N, out);
}

// param_predictor.fc2 + the running sum + activations (camera_head.py:88-99, 106-126): one warp per (view, parameter):
// delta = b[n] + <h[m], W[n]>; pred[m, n] = first ? delta : pred[m, n] + delta; act[m, n] = n >= 7 ? relu(pred) : pred.
__global__ void __launch_bounds__(32 * 9) cam_predict_kernel(const float* __restrict__ h, int K, const float* __restrict__ W,
                                                             const float* __restrict__ bias, int M, int first, float* pred,
                                                             float* __restrict__ act) {
  pdl_launch_dependents();
  pdl_wait();
  const int m = blockIdx.x, n = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float v = 0.f;
  for (int k4 = lane; k4 < K / 4; k4 += 32) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(h + (size_t)m * K) + k4);
    const float4 w = __ldg(reinterpret_cast<const float4*>(W + (size_t)n * K) + k4);
    v = fmaf(a.x, w.x, fmaf(a.y, w.y, fmaf(a.z, w.z, fmaf(a.w, w.w, v))));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if (lane == 0) {
    v += __ldg(bias + n);
    if (!first) v += pred[m * 9 + n];
    pred[m * 9 + n] = v;
    act[m * 9 + n] = n >= 7 ? fmaxf(v, 0.f) : v;
  }
}

wm_status cam_predict(cudaStream_t st, const float* h, int K, const float* W, const float* bias, int M, int first, float* pred,
                      float* act) {
  WM_CHECK(h && W && bias && pred && act && M > 0 && K > 0 && K % 128 == 0, "cam_predict: bad argument");
  return launch_pdl(cam_predict_kernel, dim3(M), dim3(32 * 9), 0, st, h, K, W, bias, M, first, pred, act);
}

}  // namespace wm
