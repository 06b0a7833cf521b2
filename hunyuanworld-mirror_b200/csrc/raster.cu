// Forward Gaussian-splat rasterisation as CUDA compute (last item): what the reference gets
// from the vendored gsplat 1.5.3 through `Rasterizer.rasterize_splats` (src/models/models/rasterization.py:29-66:
// rasterization(..., packed=True, rasterize_mode="classic", render_mode="RGB+ED", sh_degree=0), pinhole cameras).
// Restated from gsplat's CUDA sources (none of it copied; its build is unavailable on this box, glm is missing):
//   projection      csrc/ProjectionEWA3DGSPacked.cu:17-215 + include/Utils.cuh (quat_to_rotmat :142, covariance :191,
//                   persp_proj :498, add_blur :381): world -> camera, EWA 2D covariance + 0.3 px blur, conic,
//                   opacity-aware extent min(3.33, sqrt(2 ln(255 o))), near / far and image culling;
//   tile binning    csrc/IntersectTile.cu:24-110: key = image | tile | depth bits, value = flat (image, gaussian) index,
//                   stable radix sort (CUB, library), per-tile offsets (:209);
//   compositing     csrc/RasterizeToPixels3DGSFwd.cu:18-200 defines the arithmetic (front to back, alpha = min(0.999,
//                   o exp(-sigma)), skipped below 1/255, a pixel stops before the gaussian that would take its
//                   transmittance to <= 1e-4); the kernel here is organised differently, see raster_stream_kernel:
//                   after the sort every intersection becomes one packed 48-byte record in tile order, and a tile's
//                   block streams its contiguous run of records through a ring of shared-memory stages with bulk
//                   asynchronous copies (one producer thread, mbarrier full / empty pairs), so the compositing warps
//                   never gather, never wait for a block-wide barrier, and the next records land while they work;
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

// ------------------------------------------------------------------------------------------------ streamed compositing
// One record per (tile, gaussian) intersection, written in sorted (image, tile, depth) order: everything a pixel needs.
constexpr int RS_REC = 12;      // floats per record: x, y, conic a b c, opacity, r, g, b, depth, 2 pad (48 bytes, 16-byte aligned)
constexpr int RS_CHUNK = 128;   // records per shared-memory stage (6 KB)
constexpr int RS_STAGES = 4;
constexpr int RS_STREAM_THREADS = RS_BLOCK + 32;  // 8 compositing warps (16 x 2 pixels each) + 1 producer warp

__global__ void __launch_bounds__(256)
raster_pack_kernel(const float* __restrict__ means2d, const float* __restrict__ conics, const float* __restrict__ depths,
                   const float* __restrict__ opacities, const float* __restrict__ sh, const unsigned* __restrict__ flat_ids,
                   unsigned n, unsigned N, float4* __restrict__ rec) {
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const unsigned f = flat_ids[i];
    const size_t gi = f % N;
    const float r = fmaxf(RS_C0 * sh[gi * 3] + 0.5f, 0.f), gg = fmaxf(RS_C0 * sh[gi * 3 + 1] + 0.5f, 0.f),
                b = fmaxf(RS_C0 * sh[gi * 3 + 2] + 0.5f, 0.f);
    rec[(size_t)i * 3] = make_float4(means2d[(size_t)f * 2], means2d[(size_t)f * 2 + 1], conics[(size_t)f * 3], conics[(size_t)f * 3 + 1]);
    rec[(size_t)i * 3 + 1] = make_float4(conics[(size_t)f * 3 + 2], opacities[gi], r, gg);
    rec[(size_t)i * 3 + 2] = make_float4(b, depths[f], 0.f, 0.f);
  }
}

__device__ __forceinline__ void bulk_copy_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
               "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// One block per 16 x 16 pixel tile.  Warp 8 lane 0 is the producer: chunk k of the tile's records goes to stage k % 4 with
// ONE bulk copy that completes on full[stage]; it reuses a stage once all 8 compositing warps have arrived on empty[stage].
// A compositing warp owns 16 x 2 pixels, reads every record of a stage as a shared-memory broadcast, and keeps arriving
// (without computing) once all of its pixels are saturated or outside the image, counting itself in s_done once.  When the
// producer finds all 8 warps counted it publishes the index of the first chunk it will NOT load and completes that
// stage's barrier without data: the warps wake, see the index and leave (at most 3 chunks later than the last useful one).
__global__ void __launch_bounds__(RS_STREAM_THREADS)
raster_stream_kernel(const float* __restrict__ rec, const unsigned* __restrict__ offsets, RasterGeom g, float* __restrict__ out_rgb,
                     float* __restrict__ out_depth, float* __restrict__ out_alpha) {
  __shared__ __align__(128) float s_rec[RS_STAGES][RS_CHUNK * RS_REC];
  __shared__ __align__(8) uint64_t s_full[RS_STAGES], s_empty[RS_STAGES];
  __shared__ unsigned s_done;
  __shared__ int s_stop_at;
  const int cam = blockIdx.z, ty = blockIdx.y, tx = blockIdx.x;
  const unsigned tile = (unsigned)((cam * g.th + ty) * g.tw + tx);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < RS_STAGES; ++s) {
      mbar_init(&s_full[s], 1);
      mbar_init(&s_empty[s], RS_BLOCK / 32);
    }
    s_done = 0;
    s_stop_at = -1;
    fence_barrier_init();
  }
  __syncthreads();
  const unsigned start = offsets[tile], end = offsets[tile + 1];
  const unsigned n = end - start;
  const int nchunks = (int)((n + RS_CHUNK - 1) / RS_CHUNK);

  if (warp == RS_BLOCK / 32) {
    // ---------------------------------------------------------------- producer
    if (lane == 0) {
      for (int k = 0; k < nchunks; ++k) {
        const int s = k % RS_STAGES;
        if (k >= RS_STAGES) {
          mbar_wait(&s_empty[s], ((k / RS_STAGES) - 1) & 1);
          if (*reinterpret_cast<volatile unsigned*>(&s_done) >= RS_BLOCK / 32) {
            *reinterpret_cast<volatile int*>(&s_stop_at) = k;
            mbar_arrive(&s_full[s]);  // release: the index is visible to whoever wakes on this phase
            break;
          }
        }
        const uint32_t cnt = min((unsigned)RS_CHUNK, n - (unsigned)k * RS_CHUNK);
        const uint32_t bytes = cnt * RS_REC * 4;
        mbar_arrive_expect_tx(&s_full[s], bytes);
        bulk_copy_g2s(s_rec[s], rec + ((size_t)start + (size_t)k * RS_CHUNK) * RS_REC, bytes, &s_full[s]);
      }
    }
    return;
  }

  // ------------------------------------------------------------------ compositing warps
  const int i = ty * RS_TILE + warp * 2 + (lane >> 4), j = tx * RS_TILE + (lane & 15);
  const float px = (float)j + 0.5f, py = (float)i + 0.5f;
  const bool inside = i < g.H && j < g.W;
  bool done = !inside;
  bool warp_done = __all_sync(0xffffffffu, done);
  if (warp_done && lane == 0) atomicAdd(&s_done, 1u);
  float T = 1.0f, acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int k = 0; k < nchunks; ++k) {
    const int s = k % RS_STAGES;
    mbar_wait(&s_full[s], (k / RS_STAGES) & 1);
    if (*reinterpret_cast<volatile int*>(&s_stop_at) == k) break;
    if (!warp_done) {
      const unsigned cnt = min((unsigned)RS_CHUNK, n - (unsigned)k * RS_CHUNK);
      const float4* r4 = reinterpret_cast<const float4*>(s_rec[s]);
      for (unsigned t = 0; t < cnt && !done; ++t) {
        const float4 a = r4[3 * t], b = r4[3 * t + 1], c = r4[3 * t + 2];  // x y ca cb | cc op r g | b z
        const float dx = a.x - px, dy = a.y - py;
        const float sigma = 0.5f * (a.z * dx * dx + b.x * dy * dy) + a.w * dx * dy;
        const float alpha = fminf(0.999f, b.y * __expf(-sigma));
        if (sigma < 0.f || alpha < RS_ALPHA_MIN) continue;
        const float next_T = T * (1.0f - alpha);
        if (next_T <= 1e-4f) {
          done = true;
          break;
        }
        const float vis = alpha * T;
        acc[0] += b.z * vis, acc[1] += b.w * vis, acc[2] += c.x * vis, acc[3] += c.y * vis;
        T = next_T;
      }
      warp_done = __all_sync(0xffffffffu, done);
      if (warp_done && lane == 0) atomicAdd(&s_done, 1u);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&s_empty[s]);  // release: orders this warp's reads of the stage (and its s_done count) before the refill
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
  // keys in / out, values in / out, CUB's sort storage, and the packed records of the streamed compositing pass
  *render_bytes = 2 * rs_align((size_t)n * 8) + 2 * rs_align((size_t)n * 4) + rs_align(sort) + rs_align((size_t)n * RS_REC * 4);
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
  const size_t rec_bytes = rs_align(nn * RS_REC * 4);
  size_t sort_bytes = need - (2 * rs_align(nn * 8) + 2 * rs_align(nn * 4)) - rec_bytes;
  float* rec = reinterpret_cast<float*>(p + need - rec_bytes);
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
  if (n > 0) {
    const int g3 = (int)std::min<size_t>(((size_t)n + 255) / 256, 148 * 16);
    raster_pack_kernel<<<g3, 256, 0, st>>>(means2d, conics, depths, opacities, sh, vals_out, n, (unsigned)N, reinterpret_cast<float4*>(rec));
    WM_LAUNCH_CHECK();
  }
  raster_stream_kernel<<<dim3(g.tw, g.th, C), dim3(RS_STREAM_THREADS), 0, st>>>(rec, offsets, g, out_rgb, out_depth, out_alpha);
  WM_LAUNCH_CHECK();
  return WM_OK;
}

}  // namespace wm
