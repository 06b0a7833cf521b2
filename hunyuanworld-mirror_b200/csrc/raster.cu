// Forward Gaussian-splat rasterisation as CUDA compute (last item): what the reference gets
// from the vendored gsplat 1.5.3 through `Rasterizer.rasterize_splats` (src/models/models/rasterization.py:29-66:
// rasterization(..., packed=True, rasterize_mode="classic", render_mode="RGB+ED", sh_degree=0), pinhole cameras).
// Restated from gsplat's CUDA sources (none of it copied; its build is unavailable on this box, glm is missing):
//   projection      csrc/ProjectionEWA3DGSPacked.cu:17-215 + include/Utils.cuh (quat_to_rotmat :142, covariance :191,
//                   persp_proj :498, add_blur :381): world -> camera, EWA 2D covariance + 0.3 px blur, conic,
//                   opacity-aware extent min(3.33, sqrt(2 ln(255 o))), near / far and image culling;
//   tile binning    csrc/IntersectTile.cu:24-110: key = image | tile | depth bits, value = flat (image, gaussian) index,
//                   stable radix sort (CUB, library), per-tile offsets (:209);
//   compositing     csrc/RasterizeToPixels3DGSFwd.cu:18-200: one 16 x 16 block per tile, gaussians staged through shared
//                   memory in batches of 256, front to back, alpha = min(0.999, o exp(-sigma)), skipped below 1/255,
//                   a pixel stops before the gaussian that would take its transmittance to <= 1e-4;
//   colours         spherical harmonics of degree 0: max(C0 sh + 0.5, 0) (gsplat/rendering.py:509-525), expected depth =
//                   sum w z / max(alpha, 1e-10) (:760-766).
// HBM / latency-bound integer and fp32 SIMT work; no tensor cores.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include "common.cuh"
#include "ptx.cuh"

namespace wm {

constexpr int RS_TILE = 16;
constexpr int RS_BLOCK = RS_TILE * RS_TILE;
constexpr float RS_ALPHA_MIN = 1.0f / 255.0f;
constexpr float RS_C0 = 0.28209479177387814f;

struct RasterGeom {
  int C, N, W, H, tw, th;
  float near_plane, far_plane, eps2d;
};

__global__ void __launch_bounds__(256)
raster_project_kernel(const float* __restrict__ means, const float* __restrict__ quats, const float* __restrict__ scales,
                      const float* __restrict__ opacities, const float* __restrict__ viewmats, const float* __restrict__ Ks,
                      RasterGeom g, float* __restrict__ means2d, float* __restrict__ conics, float* __restrict__ depths,
                      int* __restrict__ radii, unsigned* __restrict__ tiles) {
  const size_t total = (size_t)g.C * g.N;
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int cam = (int)(idx / g.N);
    const size_t gi = idx - (size_t)cam * g.N;
    const float* V = viewmats + cam * 16;
    const float* K = Ks + cam * 9;
    const float px = means[gi * 3], py = means[gi * 3 + 1], pz = means[gi * 3 + 2];
    const float mx = V[0] * px + V[1] * py + V[2] * pz + V[3];
    const float my = V[4] * px + V[5] * py + V[6] * pz + V[7];
    const float mz = V[8] * px + V[9] * py + V[10] * pz + V[11];
    int rx = 0, ry = 0;
    unsigned ntiles = 0;
    float m2x = 0.f, m2y = 0.f, ca = 0.f, cb = 0.f, cc = 0.f;
    bool valid = !(mz < g.near_plane || mz > g.far_plane);
    if (valid) {
      // covariance R S S R^T of the (w, x, y, z) quaternion, normalised here as gsplat does
      float w = quats[gi * 4], x = quats[gi * 4 + 1], y = quats[gi * 4 + 2], z = quats[gi * 4 + 3];
      const float inv = rsqrtf(x * x + y * y + z * z + w * w);
      x *= inv, y *= inv, z *= inv, w *= inv;
      const float R[9] = {1.f - 2.f * (y * y + z * z), 2.f * (x * y - w * z),       2.f * (x * z + w * y),
                          2.f * (x * y + w * z),       1.f - 2.f * (x * x + z * z), 2.f * (y * z - w * x),
                          2.f * (x * z - w * y),       2.f * (y * z + w * x),       1.f - 2.f * (x * x + y * y)};
      const float s[3] = {scales[gi * 3], scales[gi * 3 + 1], scales[gi * 3 + 2]};
      float M[9], cov[9], T[9], covc[9];
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) M[r * 3 + c] = R[r * 3 + c] * s[c];
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) cov[r * 3 + c] = M[r * 3] * M[c * 3] + M[r * 3 + 1] * M[c * 3 + 1] + M[r * 3 + 2] * M[c * 3 + 2];
      // camera frame: Rv cov Rv^T
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) T[r * 3 + c] = V[r * 4] * cov[c] + V[r * 4 + 1] * cov[3 + c] + V[r * 4 + 2] * cov[6 + c];
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) covc[r * 3 + c] = T[r * 3] * V[c * 4] + T[r * 3 + 1] * V[c * 4 + 1] + T[r * 3 + 2] * V[c * 4 + 2];
      // perspective projection with the clamped Jacobian (persp_proj)
      const float fx = K[0], fy = K[4], cx = K[2], cy = K[5];
      const float tan_fovx = 0.5f * g.W / fx, tan_fovy = 0.5f * g.H / fy;
      const float lim_xp = (g.W - cx) / fx + 0.3f * tan_fovx, lim_xn = cx / fx + 0.3f * tan_fovx;
      const float lim_yp = (g.H - cy) / fy + 0.3f * tan_fovy, lim_yn = cy / fy + 0.3f * tan_fovy;
      const float rz = 1.f / mz, rz2 = rz * rz;
      const float tx = mz * fminf(lim_xp, fmaxf(-lim_xn, mx * rz)), ty = mz * fminf(lim_yp, fmaxf(-lim_yn, my * rz));
      const float j00 = fx * rz, j02 = -fx * tx * rz2, j11 = fy * rz, j12 = -fy * ty * rz2;
      // cov2d = J covc J^T with J = [[j00, 0, j02], [0, j11, j12]]
      const float a0 = j00 * covc[0] + j02 * covc[6], a1 = j00 * covc[1] + j02 * covc[7], a2 = j00 * covc[2] + j02 * covc[8];
      const float b0 = j11 * covc[3] + j12 * covc[6], b1 = j11 * covc[4] + j12 * covc[7], b2 = j11 * covc[5] + j12 * covc[8];
      float c00 = a0 * j00 + a2 * j02, c01 = a1 * j11 + a2 * j12, c10 = b0 * j00 + b2 * j02, c11 = b1 * j11 + b2 * j12;
      m2x = fx * mx * rz + cx, m2y = fy * my * rz + cy;
      c00 += g.eps2d, c11 += g.eps2d;
      const float det = c00 * c11 - c01 * c10;
      valid = det > 0.f;
      if (valid) {
        const float idet = 1.f / det;
        ca = c11 * idet, cb = -c01 * idet, cc = c00 * idet;  // inverse[0][0], inverse[0][1], inverse[1][1]
        const float o = opacities[gi];
        if (o < RS_ALPHA_MIN) valid = false;
        const float extend = fminf(3.33f, sqrtf(2.0f * __logf(o / RS_ALPHA_MIN)));
        const float frx = ceilf(extend * sqrtf(c00)), fry = ceilf(extend * sqrtf(c11));
        if (frx <= 0.f && fry <= 0.f) valid = false;
        if (m2x + frx <= 0.f || m2x - frx >= g.W || m2y + fry <= 0.f || m2y - fry >= g.H) valid = false;
        if (valid) {
          rx = (int)frx, ry = (int)fry;
          if (rx > 0 && ry > 0) {  // intersect_tile_kernel, first pass
            const float trx = frx / RS_TILE, try_ = fry / RS_TILE, tcx = m2x / RS_TILE, tcy = m2y / RS_TILE;
            const int x0 = min(max(0, (int)floorf(tcx - trx)), g.tw), y0 = min(max(0, (int)floorf(tcy - try_)), g.th);
            const int x1 = min(max(0, (int)ceilf(tcx + trx)), g.tw), y1 = min(max(0, (int)ceilf(tcy + try_)), g.th);
            ntiles = (unsigned)((y1 - y0) * (x1 - x0));
          }
        }
      }
    }
    if (!valid) rx = ry = 0;
    means2d[idx * 2] = m2x, means2d[idx * 2 + 1] = m2y;
    conics[idx * 3] = ca, conics[idx * 3 + 1] = cb, conics[idx * 3 + 2] = cc;
    depths[idx] = mz;
    radii[idx * 2] = rx, radii[idx * 2 + 1] = ry;
    tiles[idx] = ntiles;
  }
}

__global__ void __launch_bounds__(256)
raster_emit_kernel(const float* __restrict__ means2d, const int* __restrict__ radii, const float* __restrict__ depths,
                   const unsigned* __restrict__ cum, RasterGeom g, int tile_bits, long long* __restrict__ keys, unsigned* __restrict__ vals) {
  const size_t total = (size_t)g.C * g.N;
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int rx = radii[idx * 2], ry = radii[idx * 2 + 1];
    if (rx <= 0 || ry <= 0) continue;
    const float trx = (float)rx / RS_TILE, try_ = (float)ry / RS_TILE, tcx = means2d[idx * 2] / RS_TILE, tcy = means2d[idx * 2 + 1] / RS_TILE;
    const int x0 = min(max(0, (int)floorf(tcx - trx)), g.tw), y0 = min(max(0, (int)floorf(tcy - try_)), g.th);
    const int x1 = min(max(0, (int)ceilf(tcx + trx)), g.tw), y1 = min(max(0, (int)ceilf(tcy + try_)), g.th);
    const long long img = (long long)(idx / g.N) << (32 + tile_bits);
    const long long dep = (long long)__float_as_uint(depths[idx]);
    size_t cur = idx == 0 ? 0 : cum[idx - 1];
    for (int i = y0; i < y1; ++i)
      for (int j = x0; j < x1; ++j) {
        keys[cur] = img | ((long long)(i * g.tw + j) << 32) | dep;
        vals[cur] = (unsigned)idx;
        ++cur;
      }
  }
}

// offsets[t] = first sorted intersection whose (image, tile) id is >= t (isect_offset_encode, IntersectTile.cu:209)
__global__ void __launch_bounds__(256)
raster_offsets_kernel(const long long* __restrict__ keys, unsigned n_isects, int tile_bits, int tiles_per_image, unsigned n_tiles,
                      unsigned* __restrict__ offsets) {
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n_isects; i += gridDim.x * blockDim.x) {
    auto tile_of = [&](long long k) {
      const unsigned long long hi = (unsigned long long)k >> 32;
      return (unsigned)((hi >> tile_bits) * tiles_per_image + (hi & ((1ull << tile_bits) - 1)));
    };
    const unsigned cur = tile_of(keys[i]);
    const unsigned first = i == 0 ? 0 : tile_of(keys[i - 1]) + 1;
    for (unsigned t = first; t <= cur; ++t) offsets[t] = i;
    if (i == n_isects - 1)
      for (unsigned t = cur + 1; t <= n_tiles; ++t) offsets[t] = n_isects;
  }
}

__global__ void raster_fill_kernel(unsigned* p, unsigned n, unsigned v) {
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) p[i] = v;
}

__global__ void __launch_bounds__(RS_BLOCK)
raster_pixels_kernel(const float* __restrict__ means2d, const float* __restrict__ conics, const float* __restrict__ depths,
                     const float* __restrict__ opacities, const float* __restrict__ sh, const unsigned* __restrict__ offsets,
                     const unsigned* __restrict__ flat_ids, RasterGeom g, float* __restrict__ out_rgb, float* __restrict__ out_depth,
                     float* __restrict__ out_alpha) {
  __shared__ float s_xy[RS_BLOCK][2], s_conic[RS_BLOCK][3], s_col[RS_BLOCK][4], s_op[RS_BLOCK];
  const int cam = blockIdx.z, ty = blockIdx.y, tx = blockIdx.x;
  const unsigned tile = (unsigned)((cam * g.th + ty) * g.tw + tx);
  const int i = ty * RS_TILE + threadIdx.y, j = tx * RS_TILE + threadIdx.x;
  const float px = (float)j + 0.5f, py = (float)i + 0.5f;
  const bool inside = i < g.H && j < g.W;
  bool done = !inside;
  const unsigned start = offsets[tile], end = offsets[tile + 1];
  const unsigned tr = threadIdx.y * RS_TILE + threadIdx.x;
  float T = 1.0f, acc[4] = {0.f, 0.f, 0.f, 0.f};
  const unsigned nb = (end - start + RS_BLOCK - 1) / RS_BLOCK;
  for (unsigned b = 0; b < nb; ++b) {
    if (__syncthreads_count(done) >= RS_BLOCK) break;
    const unsigned bs = start + RS_BLOCK * b, idx = bs + tr;
    if (idx < end) {
      const unsigned f = flat_ids[idx];
      const size_t gi = f % (unsigned)g.N;
      s_xy[tr][0] = means2d[(size_t)f * 2], s_xy[tr][1] = means2d[(size_t)f * 2 + 1];
      s_conic[tr][0] = conics[(size_t)f * 3], s_conic[tr][1] = conics[(size_t)f * 3 + 1], s_conic[tr][2] = conics[(size_t)f * 3 + 2];
      s_op[tr] = opacities[gi];
#pragma unroll
      for (int k = 0; k < 3; ++k) s_col[tr][k] = fmaxf(RS_C0 * sh[gi * 3 + k] + 0.5f, 0.f);
      s_col[tr][3] = depths[f];
    }
    __syncthreads();
    const unsigned n = min((unsigned)RS_BLOCK, end - bs);
    for (unsigned t = 0; t < n && !done; ++t) {
      const float dx = s_xy[t][0] - px, dy = s_xy[t][1] - py;
      const float sigma = 0.5f * (s_conic[t][0] * dx * dx + s_conic[t][2] * dy * dy) + s_conic[t][1] * dx * dy;
      const float alpha = fminf(0.999f, s_op[t] * __expf(-sigma));
      if (sigma < 0.f || alpha < RS_ALPHA_MIN) continue;
      const float next_T = T * (1.0f - alpha);
      if (next_T <= 1e-4f) {
        done = true;
        break;
      }
      const float vis = alpha * T;
#pragma unroll
      for (int k = 0; k < 4; ++k) acc[k] += s_col[t][k] * vis;
      T = next_T;
    }
  }
  if (inside) {
    const size_t pix = ((size_t)cam * g.H + i) * g.W + j;
    const float a = 1.0f - T;
    out_alpha[pix] = a;
    out_rgb[pix * 3] = acc[0], out_rgb[pix * 3 + 1] = acc[1], out_rgb[pix * 3 + 2] = acc[2];
    out_depth[pix] = acc[3] / fmaxf(a, 1e-10f);
  }
}

static RasterGeom raster_geom(int C, long long N, int W, int H, float near_plane, float far_plane, float eps2d) {
  RasterGeom g{};
  g.C = C, g.N = (int)N, g.W = W, g.H = H, g.tw = (W + RS_TILE - 1) / RS_TILE, g.th = (H + RS_TILE - 1) / RS_TILE;
  g.near_plane = near_plane, g.far_plane = far_plane, g.eps2d = eps2d;
  return g;
}
static int raster_bits(unsigned v) {
  int b = 0;
  while (v) ++b, v >>= 1;
  return b;
}
static size_t rs_align(size_t x) { return (x + 255) & ~size_t(255); }

// Projection + tile counts + their inclusive scan.  cum[C*N - 1] is the number of intersections: the caller reads it,
// sizes the workspace with raster_workspace_bytes and calls raster_render.
wm_status raster_project(cudaStream_t st, const float* means, const float* quats, const float* scales, const float* opacities,
                         const float* viewmats, const float* Ks, int C, long long N, int W, int H, float near_plane, float far_plane,
                         float eps2d, float* means2d, float* conics, float* depths, int* radii, unsigned* tiles, unsigned* cum,
                         void* scan_tmp, size_t scan_tmp_bytes) {
  WM_CHECK(means && quats && scales && opacities && viewmats && Ks && means2d && conics && depths && radii && tiles && cum && scan_tmp,
           "raster_project: null argument");
  WM_CHECK(C > 0 && N > 0 && (long long)C * N < (1ll << 31) && W > 0 && H > 0, "raster_project: bad geometry C=%d N=%lld %dx%d", C, N, W, H);
  const RasterGeom g = raster_geom(C, N, W, H, near_plane, far_plane, eps2d);
  const size_t total = (size_t)C * N;
  const int grid = (int)std::min<size_t>((total + 255) / 256, 148 * 16);
  raster_project_kernel<<<grid, 256, 0, st>>>(means, quats, scales, opacities, viewmats, Ks, g, means2d, conics, depths, radii, tiles);
  WM_LAUNCH_CHECK();
  size_t need = 0;
  WM_CUDA(cub::DeviceScan::InclusiveSum(nullptr, need, (const unsigned*)nullptr, (unsigned*)nullptr, (int)total, st));
  WM_CHECK(scan_tmp_bytes >= need, "raster_project: scan workspace of %zu bytes, %zu needed", scan_tmp_bytes, need);
  WM_CUDA(cub::DeviceScan::InclusiveSum(scan_tmp, need, (const unsigned*)tiles, cum, (int)total, st));
  return WM_OK;
}

wm_status raster_workspace_bytes(long long elems, long long n_isects, size_t* scan_bytes, size_t* render_bytes) {
  WM_CHECK(elems > 0 && elems < (1ll << 31) && n_isects >= 0 && n_isects < (1ll << 31) && scan_bytes && render_bytes,
           "raster_workspace_bytes: bad argument");
  size_t scan = 0, sort = 0;
  WM_CUDA(cub::DeviceScan::InclusiveSum(nullptr, scan, (const unsigned*)nullptr, (unsigned*)nullptr, (int)elems, (cudaStream_t)0));
  const int n = (int)std::max<long long>(n_isects, 1);
  WM_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, sort, (const long long*)nullptr, (long long*)nullptr, (const unsigned*)nullptr,
                                          (unsigned*)nullptr, n, 0, 64, (cudaStream_t)0));
  *scan_bytes = rs_align(scan);
  *render_bytes = 2 * rs_align((size_t)n * 8) + 2 * rs_align((size_t)n * 4) + rs_align(sort);
  return WM_OK;
}

// Binning + compositing.  sh [N, 3] (degree-0 coefficients); offsets [C * th * tw + 1] scratch; out_rgb [C, H, W, 3],
// out_depth / out_alpha [C, H, W].
wm_status raster_render(cudaStream_t st, const float* means2d, const float* conics, const float* depths, const int* radii,
                        const unsigned* cum, const float* opacities, const float* sh, int C, long long N, int W, int H,
                        long long n_isects, void* workspace, size_t workspace_bytes, unsigned* offsets, float* out_rgb,
                        float* out_depth, float* out_alpha) {
  WM_CHECK(means2d && conics && depths && radii && cum && opacities && sh && workspace && offsets && out_rgb && out_depth && out_alpha,
           "raster_render: null argument");
  WM_CHECK(C > 0 && N > 0 && W > 0 && H > 0 && n_isects >= 0 && n_isects < (1ll << 31), "raster_render: bad geometry");
  const RasterGeom g = raster_geom(C, N, W, H, 0.f, 0.f, 0.f);
  const unsigned n_tiles = (unsigned)(C * g.tw * g.th);
  const int tile_bits = raster_bits((unsigned)(g.tw * g.th)), image_bits = raster_bits((unsigned)C);
  WM_CHECK(tile_bits + image_bits <= 32, "raster_render: too many tiles");
  size_t scan_b = 0, need = 0;
  WM_TRY(raster_workspace_bytes((long long)C * N, n_isects, &scan_b, &need));
  WM_CHECK(workspace_bytes >= need, "raster_render: workspace of %zu bytes, %zu needed", workspace_bytes, need);
  const unsigned n = (unsigned)n_isects;
  const size_t nn = std::max<size_t>(n, 1);
  char* p = static_cast<char*>(workspace);
  long long* keys_in = reinterpret_cast<long long*>(p);
  long long* keys_out = reinterpret_cast<long long*>(p + rs_align(nn * 8));
  unsigned* vals_in = reinterpret_cast<unsigned*>(p + 2 * rs_align(nn * 8));
  unsigned* vals_out = reinterpret_cast<unsigned*>(p + 2 * rs_align(nn * 8) + rs_align(nn * 4));
  void* sort_tmp = p + 2 * rs_align(nn * 8) + 2 * rs_align(nn * 4);
  size_t sort_bytes = workspace_bytes - (2 * rs_align(nn * 8) + 2 * rs_align(nn * 4));
  if (n == 0) {
    raster_fill_kernel<<<(n_tiles + 256) / 256, 256, 0, st>>>(offsets, n_tiles + 1, 0u);
  } else {
    const size_t total = (size_t)C * N;
    const int grid = (int)std::min<size_t>((total + 255) / 256, 148 * 16);
    raster_emit_kernel<<<grid, 256, 0, st>>>(means2d, radii, depths, cum, g, tile_bits, keys_in, vals_in);
    WM_LAUNCH_CHECK();
    WM_CUDA(cub::DeviceRadixSort::SortPairs(sort_tmp, sort_bytes, (const long long*)keys_in, keys_out, (const unsigned*)vals_in, vals_out,
                                            (int)n, 0, 32 + tile_bits + image_bits, st));
    const int g2 = (int)std::min<size_t>(((size_t)n + 255) / 256, 148 * 16);
    raster_offsets_kernel<<<g2, 256, 0, st>>>(keys_out, n, tile_bits, g.tw * g.th, n_tiles, offsets);
  }
  WM_LAUNCH_CHECK();
  raster_pixels_kernel<<<dim3(g.tw, g.th, C), dim3(RS_TILE, RS_TILE), 0, st>>>(means2d, conics, depths, opacities, sh, offsets, vals_out, g,
                                                                               out_rgb, out_depth, out_alpha);
  WM_LAUNCH_CHECK();
  return WM_OK;
}

}  // namespace wm
