// HBM-bound kernels around the DPT heads' convolutions: LayerNorm of the 2048-wide
// patch tokens, bilinear resampling (align_corners=True) of NHWC fp16 maps with the optional sinusoidal UV
// positional embedding, the ConvTranspose2d(k == stride) scatter, and the im2col of the one stride-2 convolution.
// Activations are fp16 with fp32 arithmetic (see conv.cu); every kernel moves 16 bytes (8 channels) per thread per access
// along the contiguous channel axis.
#include <cuda_fp16.h>

#include <algorithm>
#include <cstdlib>

#include "common.cuh"
#include "ptx.cuh"

namespace wm {

// nn.LayerNorm(C) over rows gathered from a strided token tensor: output row r = group g * rows_per_group + i reads
// input row g * in_group_rows + in_offset + i (DPTHead: tokens[:, :, patch_start:], dense_head.py:203-211).
// One warp per row, the row lives in registers (C <= 2048), two-pass variance, fp32 in, fp16 out (conv.cu).
template <int C>
__global__ void __launch_bounds__(256) ln_rows_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                      const float* __restrict__ b, float eps, long long rows,
                                                      int rows_per_group, int in_group_rows, int in_offset,
                                                      __half* __restrict__ y) {
  constexpr int V = C / 128;  // float4 per lane
  pdl_launch_dependents();
  pdl_wait();
  const long long r = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (r >= rows) return;
  const int lane = threadIdx.x & 31;
  const long long grp = r / rows_per_group, i = r - grp * rows_per_group;
  const float4* xr = reinterpret_cast<const float4*>(x + (grp * in_group_rows + in_offset + i) * C);
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
  uint2* yr = reinterpret_cast<uint2*>(y + r * C);
  const float4* w4 = reinterpret_cast<const float4*>(w);
  const float4* b4 = reinterpret_cast<const float4*>(b);
#pragma unroll
  for (int k = 0; k < V; ++k) {
    const float4 ww = __ldg(w4 + k * 32 + lane), bb = __ldg(b4 + k * 32 + lane);
    yr[k * 32 + lane] = make_uint2(pack_half2_sat((v[k].x - mu) * rstd * ww.x + bb.x, (v[k].y - mu) * rstd * ww.y + bb.y),
                                   pack_half2_sat((v[k].z - mu) * rstd * ww.z + bb.z, (v[k].w - mu) * rstd * ww.w + bb.w));
  }
}

wm_status ln_rows(cudaStream_t st, const float* x, const float* w, const float* b, float eps, long long rows,
                  int rows_per_group, int in_group_rows, int in_offset, int C, void* y) {
  WM_CHECK(x && w && b && y && rows > 0 && rows_per_group > 0, "ln_rows: bad argument");
  const unsigned grid = (unsigned)((rows + 7) / 8);
  __half* yo = static_cast<__half*>(y);
  if (C == 2048)
    return launch_pdl(ln_rows_kernel<2048>, dim3(grid), dim3(256), 0, st, x, w, b, eps, rows, rows_per_group, in_group_rows, in_offset, yo);
  if (C == 1024)
    return launch_pdl(ln_rows_kernel<1024>, dim3(grid), dim3(256), 0, st, x, w, b, eps, rows, rows_per_group, in_group_rows, in_offset, yo);
  return fail(WM_ERR_INVALID, "ln_rows: C must be 1024 or 2048 (got %d)", C);
}

// F.interpolate(mode="bilinear", align_corners=True) on NHWC fp16 (fp32 arithmetic) (dense_head.py:216-226, 529 via custom_interpolate
// :539-579): src = dst * (in - 1) / (out - 1).  Optional positional embedding added after resampling
// (_apply_pos_embed, dense_head.py:262-271): channel c < C/2 takes pos_u[x][c], c >= C/2 takes pos_v[y][c - C/2]
// (the UV embedding is separable: utils/grid.py:4-42).  One thread per 4 channels of one output pixel.
__device__ __forceinline__ void unpack8(const uint4& u, float* f) {
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float2 t = __half22float2(*reinterpret_cast<const __half2*>(&w[k]));
    f[2 * k] = t.x, f[2 * k + 1] = t.y;
  }
}

__global__ void __launch_bounds__(256) bilinear_kernel(const __half* __restrict__ in, int n, int Hi, int Wi, int C,
                                                       int Ho, int Wo, float sy, float sx, const float* __restrict__ pos_u,
                                                       const float* __restrict__ pos_v, __half* __restrict__ out) {
  // blockIdx.y = (image, output row): the row interpolation is uniform per block; threads run over (x, 8-channel chunk)
  // with 32-bit index arithmetic only
  pdl_launch_dependents();
  pdl_wait();
  const int c8 = C >> 3;
  const int img = blockIdx.y / Ho, yo = blockIdx.y - img * Ho;
  const float fy = yo * sy;
  const int y0 = min((int)fy, Hi - 1), y1 = min(y0 + 1, Hi - 1);
  const float ly = fy - y0, hy0 = 1.f - ly;
  const uint4* row0 = reinterpret_cast<const uint4*>(in + ((size_t)img * Hi + y0) * Wi * C);
  const uint4* row1 = reinterpret_cast<const uint4*>(in + ((size_t)img * Hi + y1) * Wi * C);
  uint4* orow = reinterpret_cast<uint4*>(out + ((size_t)img * Ho + yo) * Wo * C);
  const int half = C >> 1;
  const int total = Wo * c8;
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += gridDim.x * blockDim.x) {
    const int xo = t / c8, c = t - xo * c8;
    const float fx = xo * sx;
    const int x0 = min((int)fx, Wi - 1), x1 = min(x0 + 1, Wi - 1);
    const float lx = fx - x0, hx0 = 1.f - lx;
    float a[8], b[8], d[8], e[8], r[8];
    unpack8(__ldg(row0 + x0 * c8 + c), a);
    unpack8(__ldg(row0 + x1 * c8 + c), b);
    unpack8(__ldg(row1 + x0 * c8 + c), d);
    unpack8(__ldg(row1 + x1 * c8 + c), e);
    // the order PyTorch's upsample_bilinear2d kernel uses: h0 * (w0 * a + w1 * b) + h1 * (w0 * d + w1 * e)
#pragma unroll
    for (int i = 0; i < 8; ++i) r[i] = hy0 * (hx0 * a[i] + lx * b[i]) + ly * (hx0 * d[i] + lx * e[i]);
    if (pos_u != nullptr) {
      const int ch = c * 8;
      const float* pp = ch < half ? pos_u + xo * half + ch : pos_v + yo * half + (ch - half);
      const float4 p0 = __ldg(reinterpret_cast<const float4*>(pp)), p1 = __ldg(reinterpret_cast<const float4*>(pp) + 1);
      r[0] += p0.x, r[1] += p0.y, r[2] += p0.z, r[3] += p0.w, r[4] += p1.x, r[5] += p1.y, r[6] += p1.z, r[7] += p1.w;
    }
    __stcs(orow + t, make_uint4(pack_half2_sat(r[0], r[1]), pack_half2_sat(r[2], r[3]), pack_half2_sat(r[4], r[5]),
                                pack_half2_sat(r[6], r[7])));
  }
}

// The same for C / 8 a power of two (every map of the DPT heads: C = 128 or 256), written for the instruction count: the
// generic kernel above issues ~150 instructions per 16 output bytes (32 conversions, 50 scalar multiply-adds, an integer
// division) and is bound by issue slots, not by memory (profiles/r02_heads_launches.csv: 3.1 TB/s of algorithmic bytes on
// the 518 x 518 map).  Here the interpolation runs on packed fp32 pairs (FMUL2 / FFMA2, the same IEEE roundings per lane,
// operation order fixed explicitly: fma(lx, b, hx0 * a), fma(ly, t2, hy0 * t1)) and the index split is a shift.
template <int LC8>
__global__ void __launch_bounds__(256) bilinear_p2_kernel(const __half* __restrict__ in, int n, int Hi, int Wi, int Ho, int Wo,
                                                          float sy, float sx, const float* __restrict__ pos_u,
                                                          const float* __restrict__ pos_v, __half* __restrict__ out) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int c8 = 1 << LC8, C = c8 * 8, half = C / 2;
  const int img = blockIdx.y / Ho, yo = blockIdx.y - img * Ho;
  const float fy = yo * sy;
  const int y0 = min((int)fy, Hi - 1), y1 = min(y0 + 1, Hi - 1);
  const float ly = fy - y0, hy0 = 1.f - ly;
  const uint4* row0 = reinterpret_cast<const uint4*>(in + ((size_t)img * Hi + y0) * Wi * C);
  const uint4* row1 = reinterpret_cast<const uint4*>(in + ((size_t)img * Hi + y1) * Wi * C);
  uint4* orow = reinterpret_cast<uint4*>(out + ((size_t)img * Ho + yo) * Wo * C);
  const float2 hy2 = make_float2(hy0, hy0), ly2 = make_float2(ly, ly);
  const int total = Wo << LC8;
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += gridDim.x * blockDim.x) {
    const int xo = t >> LC8, c = t & (c8 - 1);
    const float fx = xo * sx;
    const int x0 = min((int)fx, Wi - 1), x1 = min(x0 + 1, Wi - 1);
    const float lx = fx - x0, hx0 = 1.f - lx;
    const float2 hx2 = make_float2(hx0, hx0), lx2 = make_float2(lx, lx);
    const uint4 ua = __ldg(row0 + (x0 << LC8) + c), ub = __ldg(row0 + (x1 << LC8) + c);
    const uint4 ud = __ldg(row1 + (x0 << LC8) + c), ue = __ldg(row1 + (x1 << LC8) + c);
    const uint32_t wa[4] = {ua.x, ua.y, ua.z, ua.w}, wb[4] = {ub.x, ub.y, ub.z, ub.w};
    const uint32_t wd[4] = {ud.x, ud.y, ud.z, ud.w}, we[4] = {ue.x, ue.y, ue.z, ue.w};
    float2 r[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&wa[k]));
      const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&wb[k]));
      const float2 d = __half22float2(*reinterpret_cast<const __half2*>(&wd[k]));
      const float2 e = __half22float2(*reinterpret_cast<const __half2*>(&we[k]));
      const float2 t1 = ffma2(lx2, b, fmul2(hx2, a));
      const float2 t2 = ffma2(lx2, e, fmul2(hx2, d));
      r[k] = ffma2(ly2, t2, fmul2(hy2, t1));
    }
    if (pos_u != nullptr) {
      const int ch = c * 8;
      const float* pp = ch < half ? pos_u + xo * half + ch : pos_v + yo * half + (ch - half);
      const float4 p0 = __ldg(reinterpret_cast<const float4*>(pp)), p1 = __ldg(reinterpret_cast<const float4*>(pp) + 1);
      r[0] = fadd2(r[0], make_float2(p0.x, p0.y)), r[1] = fadd2(r[1], make_float2(p0.z, p0.w));
      r[2] = fadd2(r[2], make_float2(p1.x, p1.y)), r[3] = fadd2(r[3], make_float2(p1.z, p1.w));
    }
    __stcs(orow + t, make_uint4(pack_half2_sat(r[0].x, r[0].y), pack_half2_sat(r[1].x, r[1].y), pack_half2_sat(r[2].x, r[2].y),
                                pack_half2_sat(r[3].x, r[3].y)));
  }
}

wm_status bilinear_nhwc(cudaStream_t st, const void* in, int n, int Hi, int Wi, int C, int Ho, int Wo, const float* pos_u,
                        const float* pos_v, void* out) {
  WM_CHECK(in && out && n > 0 && Hi > 0 && Wi > 0 && Ho > 0 && Wo > 0 && C > 0 && C % 16 == 0, "bilinear: bad geometry");
  WM_CHECK((pos_u == nullptr) == (pos_v == nullptr), "bilinear: pos_u and pos_v go together");
  const float sy = Ho > 1 ? (float)(Hi - 1) / (float)(Ho - 1) : 0.f, sx = Wo > 1 ? (float)(Wi - 1) / (float)(Wo - 1) : 0.f;
  WM_CHECK((long long)n * Ho <= 65535 && (long long)Wo * C < (1LL << 31), "bilinear: n * Ho must fit a grid dimension");
  const int row_work = Wo * (C / 8);
  const unsigned gx = (unsigned)std::min((row_work + 255) / 256, 8);
  static const bool generic_only = getenv("WM_BILINEAR_GENERIC") != nullptr;  // A/B switch for measurements
  if (!generic_only && C == 128)
    return launch_pdl(bilinear_p2_kernel<4>, dim3(gx, (unsigned)(n * Ho)), dim3(256), 0, st, static_cast<const __half*>(in), n, Hi, Wi,
                      Ho, Wo, sy, sx, pos_u, pos_v, static_cast<__half*>(out));
  if (!generic_only && C == 256)
    return launch_pdl(bilinear_p2_kernel<5>, dim3(gx, (unsigned)(n * Ho)), dim3(256), 0, st, static_cast<const __half*>(in), n, Hi, Wi,
                      Ho, Wo, sy, sx, pos_u, pos_v, static_cast<__half*>(out));
  return launch_pdl(bilinear_kernel, dim3(gx, (unsigned)(n * Ho)), dim3(256), 0, st, static_cast<const __half*>(in), n, Hi, Wi, C,
                    Ho, Wo, sy, sx, pos_u, pos_v, static_cast<__half*>(out));
}

// nn.ConvTranspose2d(C, C, k, stride=k) (dense_head.py:57-64) = GEMM [pixels, Cin] x [Cin, (ky, kx, co)] followed by
// this scatter: in [n*Hi*Wi, k*k*C] -> out [n, Hi*k, Wi*k, C].
__global__ void __launch_bounds__(256) pixel_shuffle_kernel(const uint4* __restrict__ in, int n, int Hi, int Wi, int C, int k,
                                                            uint4* __restrict__ out) {
  pdl_launch_dependents();
  pdl_wait();
  const int c8 = C >> 3;  // 16-byte chunks (8 fp16) per pixel
  const long long total = (long long)n * Hi * k * Wi * k * c8;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(t % c8);
    long long p = t / c8;
    const int xo = (int)(p % (Wi * k));
    p /= Wi * k;
    const int yo = (int)(p % (Hi * k)), img = (int)(p / (Hi * k));
    const int y = yo / k, ky = yo - y * k, x = xo / k, kx = xo - x * k;
    out[t] = __ldg(in + (((size_t)img * Hi + y) * Wi + x) * (k * k * c8) + (size_t)(ky * k + kx) * c8 + c);
  }
}

wm_status pixel_shuffle(cudaStream_t st, const void* in, int n, int Hi, int Wi, int C, int k, void* out) {
  WM_CHECK(in && out && n > 0 && Hi > 0 && Wi > 0 && C > 0 && C % 8 == 0 && k > 0, "pixel_shuffle: bad geometry");
  const long long total = (long long)n * Hi * k * Wi * k * (C / 8);
  const unsigned grid = (unsigned)std::min<long long>((total + 255) / 256, 148LL * 32);
  return launch_pdl(pixel_shuffle_kernel, dim3(grid), dim3(256), 0, st, static_cast<const uint4*>(in), n, Hi, Wi, C, k,
                    static_cast<uint4*>(out));
}

// im2col of nn.Conv2d(C, C, 3, stride=2, padding=1) (dense_head.py:66-68): out [n*Ho*Wo, 9*C], tap-major.
__global__ void __launch_bounds__(256) im2col_s2_kernel(const uint4* __restrict__ in, int n, int Hi, int Wi, int C, int Ho, int Wo,
                                                        uint4* __restrict__ out) {
  pdl_launch_dependents();
  pdl_wait();
  const int c8 = C >> 3;  // 16-byte chunks (8 fp16) per pixel
  const long long total = (long long)n * Ho * Wo * 9 * c8;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(t % c8);
    long long p = t / c8;
    const int tap = (int)(p % 9);
    p /= 9;
    const int xo = (int)(p % Wo);
    p /= Wo;
    const int yo = (int)(p % Ho), img = (int)(p / Ho);
    const int y = 2 * yo + tap / 3 - 1, x = 2 * xo + tap % 3 - 1;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (y >= 0 && y < Hi && x >= 0 && x < Wi) v = __ldg(in + (((size_t)img * Hi + y) * Wi + x) * c8 + c);
    out[t] = v;
  }
}

wm_status im2col_s2(cudaStream_t st, const void* in, int n, int Hi, int Wi, int C, void* out) {
  WM_CHECK(in && out && n > 0 && Hi > 0 && Wi > 0 && C > 0 && C % 8 == 0, "im2col_s2: bad geometry");
  const int Ho = (Hi - 1) / 2 + 1, Wo = (Wi - 1) / 2 + 1;
  const long long total = (long long)n * Ho * Wo * 9 * (C / 8);
  const unsigned grid = (unsigned)std::min<long long>((total + 255) / 256, 148LL * 32);
  return launch_pdl(im2col_s2_kernel, dim3(grid), dim3(256), 0, st, static_cast<const uint4*>(in), n, Hi, Wi, C, Ho, Wo,
                    static_cast<uint4*>(out));
}

}  // namespace wm
