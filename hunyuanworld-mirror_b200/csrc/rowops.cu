// HBM-bound row kernels around the GEMMs: LayerNorm over C = 1024,
// image normalise + 14x14 patch gather, depth pixel-unshuffle, pose / ray prior-token MLPs, special-token
// assembly and fp32 -> bf16 weight packing.  Vectorised 16-byte accesses, one warp per token row.
#include "common.cuh"
#include "ptx.cuh"

namespace wm {

// ------------------------------------------------------------------------------------------------
// nn.LayerNorm(1024) (block.py:44,61 eps 1e-5; vision_transformer.py:89 eps 1e-6): biased variance around
// the mean, fp32 statistics.  Row remap: logical row r -> group = r / rows_per_group, idx = r % rows_per_group;
// input row = group * in_group_rows + in_offset + idx, output row likewise.  This lets the final DINOv2 norm
// read patch rows [5, 5+g) of each view and write them to rows [7, 7+g) of the aggregator's token buffer
// (vision_transformer.py:259-263 -> visual_transformer.py:293-295).
template <typename OutT, bool ACCUM>
__global__ void __launch_bounds__(256)
layernorm_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b, float eps,
                 int rows, int rows_per_group, int in_group_rows, int in_offset, int out_group_rows, int out_offset,
                 OutT* __restrict__ y) {
  pdl_launch_dependents();
  pdl_wait();
  const int r = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (r >= rows) return;
  const int lane = threadIdx.x & 31;
  const int grp = r / rows_per_group, idx = r % rows_per_group;
  const float4* xr = reinterpret_cast<const float4*>(x + ((size_t)grp * in_group_rows + in_offset + idx) * kEmbed);
  float4 v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = xr[i * 32 + lane];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float mu = s * (1.0f / kEmbed);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float a = v[i].x - mu, bb = v[i].y - mu, c = v[i].z - mu, d = v[i].w - mu;
    q += (a * a + bb * bb) + (c * c + d * d);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
  const float rstd = rsqrtf(q * (1.0f / kEmbed) + eps);
  const float4* w4 = reinterpret_cast<const float4*>(w);
  const float4* b4 = reinterpret_cast<const float4*>(b);
  const size_t orow = ((size_t)grp * out_group_rows + out_offset + idx) * kEmbed;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c4 = i * 32 + lane;
    const float4 ww = __ldg(w4 + c4), bb = __ldg(b4 + c4);
    float4 o;
    o.x = (v[i].x - mu) * rstd * ww.x + bb.x;
    o.y = (v[i].y - mu) * rstd * ww.y + bb.y;
    o.z = (v[i].z - mu) * rstd * ww.z + bb.z;
    o.w = (v[i].w - mu) * rstd * ww.w + bb.w;
    if constexpr (sizeof(OutT) == 2) {
      uint2 pk = make_uint2(pack_bf16(o.x, o.y), pack_bf16(o.z, o.w));
      *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(y) + orow + (size_t)c4 * 4) = pk;
    } else {
      float4* dst = reinterpret_cast<float4*>(reinterpret_cast<float*>(y) + orow) + c4;
      if (ACCUM) {
        const float4 old = *dst;
        o.x += old.x, o.y += old.y, o.z += old.z, o.w += old.w;
      }
      *dst = o;
    }
  }
}

wm_status layernorm(cudaStream_t st, const float* x, const float* w, const float* b, float eps, int rows,
                    int rows_per_group, int in_group_rows, int in_offset, int out_group_rows, int out_offset, void* y,
                    int out_f32, int accumulate) {
  WM_CHECK(x && w && b && y && rows > 0 && rows_per_group > 0, "layernorm: bad arguments");
  const int grid = (rows + 7) / 8;
  if (!out_f32) {
    WM_CHECK(!accumulate, "layernorm: accumulate needs an fp32 output");
    return launch_pdl(layernorm_kernel<__nv_bfloat16, false>, dim3(grid), dim3(256), 0, st, x, w, b, eps, rows, rows_per_group,
                      in_group_rows, in_offset, out_group_rows, out_offset, static_cast<__nv_bfloat16*>(y));
  }
  if (accumulate)
    return launch_pdl(layernorm_kernel<float, true>, dim3(grid), dim3(256), 0, st, x, w, b, eps, rows, rows_per_group,
                      in_group_rows, in_offset, out_group_rows, out_offset, static_cast<float*>(y));
  return launch_pdl(layernorm_kernel<float, false>, dim3(grid), dim3(256), 0, st, x, w, b, eps, rows, rows_per_group,
                    in_group_rows, in_offset, out_group_rows, out_offset, static_cast<float*>(y));
}

// ------------------------------------------------------------------------------------------------
// Patch gather.  One thread moves one 14-pixel row segment of one patch:
//   images [N, C, H, W] fp32 -> A [N * rows_per_view, ld] bf16, row = n * rows_per_view + row_offset + patch,
//   column = c * 196 + ky * 14 + kx  (the order of conv weight.flatten(1), patch_embed.py:63-76, and of
//   pixel_unshuffle(14) for the depth prior, patch_embed.py:79-108).
// With `normalize` the ImageNet mean/std of visual_transformer.py:16-17,276-277 is applied first.
// Rows [0,row_offset) of every view and columns [C*196, ld) are zero-filled by the caller (memset).
__global__ void __launch_bounds__(256)
patch_gather_kernel(const float* __restrict__ img, int n_views, int channels, int H, int W, int rows_per_view,
                    int row_offset, int ld, int normalize, __nv_bfloat16* __restrict__ out) {
  const int gh = H / kPatch, gw = W / kPatch;
  const long long total = (long long)n_views * channels * H * gw;
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  const int px = (int)(t % gw);
  long long r = t / gw;
  const int y = (int)(r % H);
  r /= H;
  const int c = (int)(r % channels);
  const int n = (int)(r / channels);
  const int py = y / kPatch, ky = y % kPatch;
  const float* src = img + (((size_t)n * channels + c) * H + y) * W + px * kPatch;
  float mean = 0.f, inv = 1.f;
  if (normalize) {
    mean = c == 0 ? 0.485f : (c == 1 ? 0.456f : 0.406f);
    inv = 1.0f / (c == 0 ? 0.229f : (c == 1 ? 0.224f : 0.225f));
  }
  __nv_bfloat16* dst = out + ((size_t)n * rows_per_view + row_offset + py * gw + px) * ld + c * 196 + ky * kPatch;
  (void)gh;
#pragma unroll
  for (int i = 0; i < kPatch; i += 2) {
    const float2 v = *reinterpret_cast<const float2*>(src + i);
    *reinterpret_cast<uint32_t*>(dst + i) = pack_bf16((v.x - mean) * inv, (v.y - mean) * inv);
  }
}

wm_status patch_gather(cudaStream_t st, const float* img, int n_views, int channels, int H, int W, int rows_per_view,
                       int row_offset, int ld, int normalize, void* out) {
  WM_CHECK(img && out && n_views > 0 && H % kPatch == 0 && W % kPatch == 0 && H > 0 && W > 0,
           "patch_gather: H and W must be positive multiples of 14");
  WM_CHECK(ld >= channels * 196 && ld % 8 == 0 && rows_per_view >= row_offset + (H / kPatch) * (W / kPatch),
           "patch_gather: bad output geometry");
  WM_CHECK((reinterpret_cast<uintptr_t>(img) & 7) == 0, "patch_gather: image pointer must be 8-byte aligned");
  const long long total = (long long)n_views * channels * H * (W / kPatch);
  const int grid = (int)((total + 255) / 256);
  patch_gather_kernel<<<grid, 256, 0, st>>>(img, n_views, channels, H, W, rows_per_view, row_offset, ld, normalize,
                                            static_cast<__nv_bfloat16*>(out));
  WM_LAUNCH_CHECK();
  return WM_OK;
}

// ------------------------------------------------------------------------------------------------
// pose_embed / ray_embed: Linear(d_in -> 1024) -> SiLU -> Linear(1024 -> 1024)  (visual_transformer.py:178-204),
// one token per view, written to row (n * rows_per_view + row) of the fp32 token buffer.  fp32 weights.
__global__ void __launch_bounds__(256)
prior_token_kernel(const float* __restrict__ in, int d_in, const float* __restrict__ w0, const float* __restrict__ b0,
                   const float* __restrict__ w2, const float* __restrict__ b2, int rows_per_view, int row,
                   float* __restrict__ x) {
  __shared__ float h[kEmbed];
  const int n = blockIdx.x;
  for (int o = threadIdx.x; o < kEmbed; o += blockDim.x) {
    float acc = b0[o];
    for (int k = 0; k < d_in; ++k) acc += w0[o * d_in + k] * in[n * d_in + k];
    h[o] = acc / (1.0f + __expf(-acc));  // SiLU
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int o0 = blockIdx.y * 64 + warp * 8;
  for (int o = o0; o < o0 + 8; ++o) {
    const float4* wr = reinterpret_cast<const float4*>(w2 + (size_t)o * kEmbed);
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float4 ww = __ldg(wr + i * 32 + lane);
      const float4 hh = *reinterpret_cast<const float4*>(h + (i * 32 + lane) * 4);
      acc += (ww.x * hh.x + ww.y * hh.y) + (ww.z * hh.z + ww.w * hh.w);
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s);
    if (lane == 0) x[((size_t)n * rows_per_view + row) * kEmbed + o] = acc + b2[o];
  }
}

wm_status prior_token(cudaStream_t st, const float* in, int n_views, int d_in, const float* w0, const float* b0,
                      const float* w2, const float* b2, int rows_per_view, int row, float* x) {
  WM_CHECK(in && w0 && b0 && w2 && b2 && x && n_views > 0 && d_in > 0, "prior_token: bad arguments");
  prior_token_kernel<<<dim3(n_views, kEmbed / 64), 256, 0, st>>>(in, d_in, w0, b0, w2, b2, rows_per_view, row, x);
  WM_LAUNCH_CHECK();
  return WM_OK;
}

// ------------------------------------------------------------------------------------------------
// Special tokens (visual_transformer.py:285-295,397-416): rows 0..4 of each view = cam, reg x4 with variant 0
// for the globally first view and variant 1 for every other; rows 5 (pose) and 6 (ray) are zeroed when their
// prior is off (`zero_pose` / `zero_ray`), otherwise left to prior_token_kernel.
__global__ void __launch_bounds__(256)
special_tokens_kernel(const float* __restrict__ cam, const float* __restrict__ reg, int rows_per_view,
                      int first_view_global, int zero_pose, int zero_ray, float* __restrict__ x) {
  const int n = blockIdx.x, row = blockIdx.y;  // row in [0, 7)
  const int variant = (first_view_global + n) == 0 ? 0 : 1;
  float4* dst = reinterpret_cast<float4*>(x + ((size_t)n * rows_per_view + row) * kEmbed) + threadIdx.x;
  if (row == 0) {
    *dst = reinterpret_cast<const float4*>(cam + (size_t)variant * kEmbed)[threadIdx.x];
  } else if (row < 5) {
    *dst = reinterpret_cast<const float4*>(reg + ((size_t)variant * 4 + (row - 1)) * kEmbed)[threadIdx.x];
  } else if ((row == 5 && zero_pose) || (row == 6 && zero_ray)) {
    *dst = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

wm_status special_tokens(cudaStream_t st, const float* cam, const float* reg, int n_views, int rows_per_view,
                         int first_view_global, int zero_pose, int zero_ray, float* x) {
  WM_CHECK(cam && reg && x && n_views > 0 && rows_per_view >= kPatchStart, "special_tokens: bad arguments");
  special_tokens_kernel<<<dim3(n_views, kPatchStart), 256, 0, st>>>(cam, reg, rows_per_view, first_view_global,
                                                                     zero_pose, zero_ray, x);
  WM_LAUNCH_CHECK();
  return WM_OK;
}

// ------------------------------------------------------------------------------------------------
// Intermediate outputs (visual_transformer.py:337-339): out[:, col0 : col0 + 1024] = x, the two halves of the
// reference's `cat([frame, global], -1)` written in place, one float4 per thread.
__global__ void __launch_bounds__(256)
copy_cols_kernel(const float4* __restrict__ x, long long rows, int ld_out4, int col4, float4* __restrict__ out) {
  pdl_launch_dependents();
  pdl_wait();
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= rows * (kEmbed / 4)) return;
  const long long r = t / (kEmbed / 4);
  const int c = (int)(t % (kEmbed / 4));
  out[r * ld_out4 + col4 + c] = x[t];
}

wm_status copy_cols(cudaStream_t st, const float* x, long long rows, float* out, int ld_out, int col0) {
  WM_CHECK(x && out && rows > 0 && ld_out % 4 == 0 && col0 % 4 == 0 && col0 + kEmbed <= ld_out, "copy_cols: bad arguments");
  const long long total = rows * (kEmbed / 4);
  return launch_pdl(copy_cols_kernel, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, st,
                    reinterpret_cast<const float4*>(x), rows, ld_out / 4, col0 / 4, reinterpret_cast<float4*>(out));
}

// ------------------------------------------------------------------------------------------------
// fp32 -> bf16 packing with optional row padding (weights: [rows, cols] -> [rows, ld], zero tail columns).
__global__ void __launch_bounds__(256)
pack_bf16_kernel(const float* __restrict__ src, long long rows, int cols, int ld, __nv_bfloat16* __restrict__ dst) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= rows * ld) return;
  const long long r = t / ld;
  const int c = (int)(t % ld);
  dst[t] = __float2bfloat16_rn(c < cols ? src[r * cols + c] : 0.0f);
}

wm_status pack_bf16_rows(cudaStream_t st, const float* src, long long rows, int cols, int ld, void* dst) {
  WM_CHECK(src && dst && rows > 0 && cols > 0 && ld >= cols, "pack_bf16_rows: bad arguments");
  const long long total = rows * ld;
  pack_bf16_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(src, rows, cols, ld, static_cast<__nv_bfloat16*>(dst));
  WM_LAUNCH_CHECK();
  return WM_OK;
}

// ------------------------------------------------------------------------------------------------
// RotaryPositionEmbedding2D.forward (rope.py:148-181) as a stand-alone operator (secondary seam of SURVEY.md section 8b;
// inside wm_forward the rotation lives in the QKV epilogue).  tokens [B, Hh, L, 64] fp32, positions [B, L, 2] int64
// (y, x), ANY values in [0, max_pos]: dims [0, 32) rotate with y, [32, 64) with x; within an axis dim j pairs with
// j +- 16 (rotate-half) at frequency j % 16: out[j] = t[j] cos + (j < 16 ? -t[j + 16] : t[j - 16]) sin.
// HBM-bound: 256 B read + 256 B written per row; 16 threads per row, one float4 each (+ the partner float4 from L1).
__global__ void __launch_bounds__(256)
rope2d_kernel(const float* __restrict__ t, const long long* __restrict__ pos, long long rows, int heads, int L, int max_pos,
              const float* __restrict__ cos_t, const float* __restrict__ sin_t, float* __restrict__ out) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < rows * 16; i += (long long)gridDim.x * blockDim.x) {
    const long long row = i >> 4;  // (b * heads + h) * L + l
    const int c = (int)(i & 15), axis = c >> 3, j0 = (c & 7) * 4;  // this thread: dims axis * 32 + j0 .. + 3
    const long long b = row / ((long long)heads * L);
    const int l = (int)(row % L);
    long long p = pos[(b * L + l) * 2 + axis];
    p = p < 0 ? 0 : (p > max_pos ? max_pos : p);
    const float4 x = *reinterpret_cast<const float4*>(t + row * 64 + axis * 32 + j0);
    const float4 y = *reinterpret_cast<const float4*>(t + row * 64 + axis * 32 + (j0 ^ 16));
    const float4 cs = *reinterpret_cast<const float4*>(cos_t + p * 16 + (j0 & 15));
    const float4 sn = *reinterpret_cast<const float4*>(sin_t + p * 16 + (j0 & 15));
    const float sg = j0 < 16 ? -1.f : 1.f;
    float4 o;  // products rounded separately, then added: x * cos + rot * sin as torch evaluates it (rope.py:146)
    o.x = __fadd_rn(__fmul_rn(x.x, cs.x), __fmul_rn(sg * y.x, sn.x));
    o.y = __fadd_rn(__fmul_rn(x.y, cs.y), __fmul_rn(sg * y.y, sn.y));
    o.z = __fadd_rn(__fmul_rn(x.z, cs.z), __fmul_rn(sg * y.z, sn.z));
    o.w = __fadd_rn(__fmul_rn(x.w, cs.w), __fmul_rn(sg * y.w, sn.w));
    *reinterpret_cast<float4*>(out + row * 64 + axis * 32 + j0) = o;
  }
}

wm_status rope2d(cudaStream_t st, const float* tokens, const long long* pos, int batch, int heads, int L, int max_pos,
                 const float* cos_t, const float* sin_t, float* out) {
  WM_CHECK(tokens && pos && cos_t && sin_t && out && batch > 0 && heads > 0 && L > 0 && max_pos >= 0, "rope2d: bad argument");
  WM_CHECK(tokens != out, "rope2d: in place is not supported (a thread reads its partner's input)");
  const long long rows = (long long)batch * heads * L;
  const int grid = (int)std::min<long long>((rows * 16 + 255) / 256, 148 * 16);
  rope2d_kernel<<<grid, 256, 0, st>>>(tokens, pos, rows, heads, L, max_pos, cos_t, sin_t, out);
  WM_LAUNCH_CHECK();
  return WM_OK;
}

}  // namespace wm
