// Gaussian path of WorldMirror.forward (reference: src/models/models/rasterization.py).
// What `GaussianSplatRenderer.render` leaves in the predictions is `splats` only (the rasterised images are dropped,
// rasterization.py:239-246), so the path is: gs DPT head (conv.cu / headops.cu, plus the 7x7 image branch prepared
// here) -> gs_head parameter convolution (conv.cu, projecting epilogue) -> prepare_splats (:389-498) -> prune_gs (:301-387).
// Everything in this file is HBM-bound elementwise / integer work, written for coalesced access, no tensor cores:
//   im2row7_kernel        images NCHW -> NHWC fp16 rows of the 7 horizontal taps (21 of 64 channels), so the 7x7 convolution
//                         of `input_merger` (dense_head.py:91-94) becomes a 7-tap vertical implicit GEMM of conv.cu;
//   splat_camera_kernel   camera_params [S, 9] -> camera-to-world rotation / translation and pinhole intrinsics
//                         (utils/camera_utils.py:46-75, utils/rotation.py:8-38, utils/geometry.py:92-111);
//   splat_prepare_kernel  activations of act_gs.py, RGB2SH, depth un-projection (utils/geometry.py:5-89);
//   prune: voxel index bounds -> flat voxel key -> stable radix sort (CUB, library) -> segment heads -> scan -> one
//                         thread per voxel merges its splats IN ORIGINAL INDEX ORDER, which is the order the reference's
//                         CPU scatter_add_ accumulates in, so the merge is deterministic (the reference on CUDA is not).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include "common.cuh"
#include "ptx.cuh"

namespace wm {

constexpr int SP_THREADS = 256;
constexpr int SP_RAW = 12;  // quats 4 | scales 3 | opacity 1 | residual SH (degree 0) 3 | weight 1  (rasterization.py:131-138)

// ------------------------------------------------------------------------------------------------ 7x7 image branch
__global__ void __launch_bounds__(SP_THREADS) im2row7_kernel(const float* __restrict__ img, int n, int H, int W, uint4* __restrict__ out) {
  const size_t total = (size_t)n * H * W * 8;  // 8 chunks of 8 fp16 channels per pixel
  for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
    const int grp = (int)(t & 7);
    uint4 o = make_uint4(0u, 0u, 0u, 0u);
    if (grp < 3) {  // channels 0..20 carry data
      const size_t pix = t >> 3;
      const int x = (int)(pix % W);
      const int y = (int)((pix / W) % H);
      const size_t im = pix / ((size_t)W * H);
      float v[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int k = grp * 8 + i, kx = k / 3, c = k - kx * 3, xs = x + kx - 3;
        v[i] = (k < 21 && xs >= 0 && xs < W) ? __ldg(img + ((im * 3 + c) * H + y) * W + xs) : 0.f;
      }
      o = make_uint4(pack_half2_sat(v[0], v[1]), pack_half2_sat(v[2], v[3]), pack_half2_sat(v[4], v[5]), pack_half2_sat(v[6], v[7]));
    }
    out[t] = o;
  }
}

wm_status im2row7(cudaStream_t st, const float* img, int n, int H, int W, void* out) {
  WM_CHECK(img && out && n > 0 && H > 0 && W > 0, "im2row7: bad argument");
  const size_t total = (size_t)n * H * W * 8;
  const int grid = (int)std::min<size_t>((total + SP_THREADS - 1) / SP_THREADS, 148 * 32);
  im2row7_kernel<<<grid, SP_THREADS, 0, st>>>(img, n, H, W, static_cast<uint4*>(out));
  WM_LAUNCH_CHECK();
  return WM_OK;
}

// ------------------------------------------------------------------------------------------------ cameras
// cam[v] = { R_c2w (9, row major), t_c2w (3), fx, fy, cx, cy }.  fp32, operations in the reference's order.
__global__ void splat_camera_kernel(const float* __restrict__ params, int n, int H, int W, float* __restrict__ cam) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= n) return;
  const float* p = params + (size_t)v * 9;
  const float t0 = p[0], t1 = p[1], t2 = p[2], i = p[3], j = p[4], k = p[5], r = p[6];
  const float ss = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(i, i), __fmul_rn(j, j)), __fmul_rn(k, k)), __fmul_rn(r, r));
  const float two_s = __fdiv_rn(2.0f, ss);
  auto a = [](float x, float y) { return __fadd_rn(x, y); };
  auto s = [](float x, float y) { return __fsub_rn(x, y); };
  auto m = [](float x, float y) { return __fmul_rn(x, y); };
  float R[9];  // world-to-camera rotation (quat_to_rotmat)
  R[0] = s(1.f, m(two_s, a(m(j, j), m(k, k))));
  R[1] = m(two_s, s(m(i, j), m(k, r)));
  R[2] = m(two_s, a(m(i, k), m(j, r)));
  R[3] = m(two_s, a(m(i, j), m(k, r)));
  R[4] = s(1.f, m(two_s, a(m(i, i), m(k, k))));
  R[5] = m(two_s, s(m(j, k), m(i, r)));
  R[6] = m(two_s, s(m(i, k), m(j, r)));
  R[7] = m(two_s, a(m(j, k), m(i, r)));
  R[8] = s(1.f, m(two_s, a(m(i, i), m(j, j))));
  float* o = cam + (size_t)v * 16;
  // closed_form_inverse_se3: R^T, -R^T t
#pragma unroll
  for (int rr = 0; rr < 3; ++rr) {
#pragma unroll
    for (int cc = 0; cc < 3; ++cc) o[rr * 3 + cc] = R[cc * 3 + rr];
    o[9 + rr] = -a(a(m(R[rr], t0), m(R[3 + rr], t1)), m(R[6 + rr], t2));
  }
  o[12] = __fdiv_rn(m((float)W, 0.5f), tanf(m(p[8], 0.5f)));  // fx from fov_u
  o[13] = __fdiv_rn(m((float)H, 0.5f), tanf(m(p[7], 0.5f)));  // fy from fov_v
  o[14] = m((float)W, 0.5f);
  o[15] = m((float)H, 0.5f);
}

wm_status splat_cameras(cudaStream_t st, const float* params, int n, int H, int W, float* cam) {
  WM_CHECK(params && cam && n > 0 && H > 0 && W > 0, "splat_cameras: bad argument");
  splat_camera_kernel<<<(n + 127) / 128, 128, 0, st>>>(params, n, H, W, cam);
  WM_LAUNCH_CHECK();
  return WM_OK;
}

// ------------------------------------------------------------------------------------------------ prepare_splats
__device__ __forceinline__ float sp_sigmoid(float x) { return __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-x))); }

struct SplatOut {
  float *means, *quats, *scales, *opacities, *sh, *weights;  // [N,3] [N,4] [N,3] [N] [N,3] [N]
};

__global__ void __launch_bounds__(SP_THREADS)
splat_prepare_kernel(const float* __restrict__ raw, const float* __restrict__ img, const float* __restrict__ depth,
                     const float* __restrict__ cam, int n, int H, int W, SplatOut o) {
  const size_t hw = (size_t)H * W, total = (size_t)n * hw;
  for (size_t pix = blockIdx.x * (size_t)blockDim.x + threadIdx.x; pix < total; pix += (size_t)gridDim.x * blockDim.x) {
    const size_t v = pix / hw, rem = pix - v * hw;
    const int y = (int)(rem / W), x = (int)(rem - (size_t)y * W);
    float p[SP_RAW];
    const float4* r4 = reinterpret_cast<const float4*>(raw + pix * SP_RAW);
#pragma unroll
    for (int i = 0; i < SP_RAW / 4; ++i) {
      const float4 t = __ldg(r4 + i);
      p[4 * i] = t.x, p[4 * i + 1] = t.y, p[4 * i + 2] = t.z, p[4 * i + 3] = t.w;
    }
    // reg_dense_rotation: q / (|q| + 1e-8)
    const float qn = __fadd_rn(__fsqrt_rn(__fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(p[0], p[0]), __fmul_rn(p[1], p[1])), __fmul_rn(p[2], p[2])),
                                                    __fmul_rn(p[3], p[3]))), 1e-8f);
#pragma unroll
    for (int i = 0; i < 4; ++i) o.quats[pix * 4 + i] = __fdiv_rn(p[i], qn);
    // reg_dense_scales (exp) clamped at 0.3; sigmoid opacity and weight
#pragma unroll
    for (int i = 0; i < 3; ++i) o.scales[pix * 3 + i] = fminf(expf(p[4 + i]), 0.3f);
    o.opacities[pix] = sp_sigmoid(p[7]);
    o.weights[pix] = sp_sigmoid(p[11]);
    // sh = RGB2SH(image) + residual  (utils/sh_utils.py:112-113, C0 as an fp32 divisor)
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float rgb = __ldg(img + (v * 3 + c) * hw + rem);
      o.sh[pix * 3 + c] = __fadd_rn(__fdiv_rn(__fsub_rn(rgb, 0.5f), 0.28209479177387814f), p[8 + c]);
    }
    // means: depth_to_world_coords_points(gs_depth, camera-to-world, intrinsics)
    const float* cv = cam + v * 16;
    const float z = __ldg(depth + pix);
    const float xc = __fdiv_rn(__fmul_rn(__fsub_rn((float)x, cv[14]), z), cv[12]);
    const float yc = __fdiv_rn(__fmul_rn(__fsub_rn((float)y, cv[15]), z), cv[13]);
#pragma unroll
    for (int jx = 0; jx < 3; ++jx)
      o.means[pix * 3 + jx] = fmaf(cv[jx * 3 + 2], z, fmaf(cv[jx * 3 + 1], yc, cv[jx * 3] * xc)) + cv[9 + jx];
  }
}

wm_status splat_prepare(cudaStream_t st, const float* raw, const float* img, const float* depth, const float* cam, int n, int H,
                        int W, float* means, float* quats, float* scales, float* opacities, float* sh, float* weights) {
  WM_CHECK(raw && img && depth && cam && means && quats && scales && opacities && sh && weights && n > 0 && H > 0 && W > 0,
           "splat_prepare: bad argument");
  const size_t total = (size_t)n * H * W;
  const int grid = (int)std::min<size_t>((total + SP_THREADS - 1) / SP_THREADS, 148 * 16);
  splat_prepare_kernel<<<grid, SP_THREADS, 0, st>>>(raw, img, depth, cam, n, H, W, SplatOut{means, quats, scales, opacities, sh, weights});
  WM_LAUNCH_CHECK();
  return WM_OK;
}

}  // namespace wm

// ------------------------------------------------------------------------------------------------ prune_gs
namespace wm {

struct PruneWs {
  long long* bounds;  // [6] = min xyz, max xyz of the voxel indices
  long long *keys_in, *keys_out;
  unsigned *vals_in, *vals_out, *flags, *segid;
  void* cub_tmp;
  size_t cub_bytes;
};

__device__ __forceinline__ long long sp_voxel(float m, float vs) { return (long long)floorf(__fdiv_rn(m, vs)); }

__global__ void prune_init_kernel(long long* bounds) {
  if (threadIdx.x < 3) bounds[threadIdx.x] = LLONG_MAX;
  else if (threadIdx.x < 6) bounds[threadIdx.x] = LLONG_MIN;
}

__global__ void __launch_bounds__(SP_THREADS) prune_bounds_kernel(const float* __restrict__ means, unsigned N, float vs, long long* bounds) {
  long long lo[3] = {LLONG_MAX, LLONG_MAX, LLONG_MAX}, hi[3] = {LLONG_MIN, LLONG_MIN, LLONG_MIN};
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      const long long v = sp_voxel(__ldg(means + (size_t)i * 3 + d), vs);
      lo[d] = v < lo[d] ? v : lo[d];
      hi[d] = v > hi[d] ? v : hi[d];
    }
  }
#pragma unroll
  for (int d = 0; d < 3; ++d) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      const long long a = __shfl_xor_sync(0xffffffffu, lo[d], off), b = __shfl_xor_sync(0xffffffffu, hi[d], off);
      lo[d] = a < lo[d] ? a : lo[d];
      hi[d] = b > hi[d] ? b : hi[d];
    }
  }
  // warp results -> block result in shared memory -> six global atomics per block
  __shared__ long long s_lo[3][SP_THREADS / 32], s_hi[3][SP_THREADS / 32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
#pragma unroll
    for (int d = 0; d < 3; ++d) s_lo[d][warp] = lo[d], s_hi[d][warp] = hi[d];
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    const int d = threadIdx.x;
    long long a = s_lo[d][0], b2 = s_hi[d][0];
    for (int w = 1; w < SP_THREADS / 32; ++w) {
      a = s_lo[d][w] < a ? s_lo[d][w] : a;
      b2 = s_hi[d][w] > b2 ? s_hi[d][w] : b2;
    }
    atomicMin(bounds + d, a);
    atomicMax(bounds + 3 + d, b2);
  }
}

// flat = (vx - minx) * dimy * dimz + (vy - miny) * dimz + (vz - minz), int64 as torch computes it (rasterization.py:322-331)
__global__ void __launch_bounds__(SP_THREADS)
prune_keys_kernel(const float* __restrict__ means, unsigned N, float vs, const long long* __restrict__ bounds, long long* __restrict__ keys,
                  unsigned* __restrict__ vals) {
  const long long mx = bounds[0], my = bounds[1], mz = bounds[2];
  const long long dy = bounds[4] - my + 1, dz = bounds[5] - mz + 1;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
    const long long vx = sp_voxel(__ldg(means + (size_t)i * 3), vs) - mx;
    const long long vy = sp_voxel(__ldg(means + (size_t)i * 3 + 1), vs) - my;
    const long long vz = sp_voxel(__ldg(means + (size_t)i * 3 + 2), vs) - mz;
    keys[i] = vx * dy * dz + vy * dz + vz;
    vals[i] = i;
  }
}

__global__ void __launch_bounds__(SP_THREADS) prune_flags_kernel(const long long* __restrict__ keys, unsigned N, unsigned* __restrict__ flags) {
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x)
    flags[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1u : 0u;
}

// One thread per voxel (= per segment head of the sorted keys).  The sort is stable, so a segment lists its splats in
// ascending original index: the order torch's CPU scatter_add_ adds them in.  Products are rounded before they are added
// (no fused multiply-add), as `x * weights` followed by scatter_add_ does.
__global__ void __launch_bounds__(SP_THREADS)
prune_merge_kernel(const long long* __restrict__ keys, const unsigned* __restrict__ vals, const unsigned* __restrict__ segid, unsigned N,
                   SplatOut in, SplatOut out, unsigned* __restrict__ count) {
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
    if (i == N - 1) *count = segid[i];
    const long long key = keys[i];
    if (i != 0 && keys[i - 1] == key) continue;
    const unsigned s = segid[i] - 1;
    float wsum = 0.f, op = 0.f, m[3] = {0.f, 0.f, 0.f}, sh[3] = {0.f, 0.f, 0.f}, sc[3] = {0.f, 0.f, 0.f}, q[4] = {0.f, 0.f, 0.f, 0.f};
    unsigned j = i;
    do {
      const size_t idx = vals[j];
      const float w = __ldg(in.weights + idx);
      wsum = __fadd_rn(wsum, w);
      op = __fadd_rn(op, __fmul_rn(w, w));
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        m[d] = __fadd_rn(m[d], __fmul_rn(__ldg(in.means + idx * 3 + d), w));
        sh[d] = __fadd_rn(sh[d], __fmul_rn(__ldg(in.sh + idx * 3 + d), w));
        sc[d] = __fadd_rn(sc[d], __fmul_rn(__ldg(in.scales + idx * 3 + d), w));
      }
#pragma unroll
      for (int d = 0; d < 4; ++d) q[d] = __fadd_rn(q[d], __fmul_rn(__ldg(in.quats + idx * 4 + d), w));
      ++j;
    } while (j < N && keys[j] == key);
    wsum = fmaxf(wsum, 1e-8f);
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      out.means[(size_t)s * 3 + d] = __fdiv_rn(m[d], wsum);
      out.sh[(size_t)s * 3 + d] = __fdiv_rn(sh[d], wsum);
      out.scales[(size_t)s * 3 + d] = __fdiv_rn(sc[d], wsum);
    }
    out.opacities[s] = __fdiv_rn(op, wsum);
    const float qn = fmaxf(__fsqrt_rn(__fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(q[0], q[0]), __fmul_rn(q[1], q[1])), __fmul_rn(q[2], q[2])),
                                                 __fmul_rn(q[3], q[3]))), 1e-8f);
#pragma unroll
    for (int d = 0; d < 4; ++d) out.quats[(size_t)s * 4 + d] = __fdiv_rn(q[d], qn);
  }
}

static size_t sp_align(size_t x) { return (x + 255) & ~size_t(255); }

static wm_status prune_layout(unsigned N, void* base, size_t* total, PruneWs* ws) {
  size_t sort_bytes = 0, scan_bytes = 0;
  WM_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (const long long*)nullptr, (long long*)nullptr, (const unsigned*)nullptr,
                                          (unsigned*)nullptr, (int)N, 0, 64, (cudaStream_t)0));
  WM_CUDA(cub::DeviceScan::InclusiveSum(nullptr, scan_bytes, (const unsigned*)nullptr, (unsigned*)nullptr, (int)N, (cudaStream_t)0));
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* p = base ? static_cast<char*>(base) + off : nullptr;
    off += sp_align(bytes);
    return p;
  };
  PruneWs w{};
  w.bounds = static_cast<long long*>(take(6 * sizeof(long long)));
  w.keys_in = static_cast<long long*>(take((size_t)N * 8));
  w.keys_out = static_cast<long long*>(take((size_t)N * 8));
  w.vals_in = static_cast<unsigned*>(take((size_t)N * 4));
  w.vals_out = static_cast<unsigned*>(take((size_t)N * 4));
  w.flags = static_cast<unsigned*>(take((size_t)N * 4));
  w.segid = static_cast<unsigned*>(take((size_t)N * 4));
  w.cub_bytes = std::max(sort_bytes, scan_bytes);
  w.cub_tmp = take(w.cub_bytes);
  *total = off;
  if (ws) *ws = w;
  return WM_OK;
}

wm_status prune_workspace_bytes(long long N, size_t* bytes) {
  WM_CHECK(N > 0 && N < (1ll << 31) && bytes, "prune_workspace_bytes: bad argument");
  return prune_layout((unsigned)N, nullptr, bytes, nullptr);
}

// prune_gs (rasterization.py:301-387) for one scene: N splats -> *count voxels (device word), outputs sized for N rows.
// Voxel-index bounds of a set of means ([6] int64 on the device: min xyz, max xyz) and the flat keys for GIVEN bounds:
// the two halves of prune_splats' keying, exposed so that ranks holding different views of one scene can agree on the
// bounds (all-reduce min / max) and route every splat to the rank that owns its key range (heads.py).
wm_status voxel_bounds(cudaStream_t st, long long N, const float* means, double voxel_size, long long* bounds) {
  WM_CHECK(N >= 0 && N < (1ll << 31) && bounds && (N == 0 || means) && voxel_size > 0, "voxel_bounds: bad argument");
  prune_init_kernel<<<1, 32, 0, st>>>(bounds);
  if (N > 0) {
    const int grid = (int)std::min<size_t>(((size_t)N + SP_THREADS - 1) / SP_THREADS, 148 * 16);
    prune_bounds_kernel<<<grid, SP_THREADS, 0, st>>>(means, (unsigned)N, (float)voxel_size, bounds);
  }
  WM_LAUNCH_CHECK();
  return WM_OK;
}

__global__ void __launch_bounds__(SP_THREADS)
voxel_keys_kernel(const float* __restrict__ means, unsigned N, float vs, const long long* __restrict__ bounds, long long* __restrict__ keys) {
  const long long mx = bounds[0], my = bounds[1], mz = bounds[2];
  const long long dy = bounds[4] - my + 1, dz = bounds[5] - mz + 1;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
    const long long vx = sp_voxel(__ldg(means + (size_t)i * 3), vs) - mx;
    const long long vy = sp_voxel(__ldg(means + (size_t)i * 3 + 1), vs) - my;
    const long long vz = sp_voxel(__ldg(means + (size_t)i * 3 + 2), vs) - mz;
    keys[i] = vx * dy * dz + vy * dz + vz;
  }
}

wm_status voxel_keys(cudaStream_t st, long long N, const float* means, double voxel_size, const long long* bounds, long long* keys) {
  WM_CHECK(N > 0 && N < (1ll << 31) && means && bounds && keys && voxel_size > 0, "voxel_keys: bad argument");
  const int grid = (int)std::min<size_t>(((size_t)N + SP_THREADS - 1) / SP_THREADS, 148 * 16);
  voxel_keys_kernel<<<grid, SP_THREADS, 0, st>>>(means, (unsigned)N, (float)voxel_size, bounds, keys);
  WM_LAUNCH_CHECK();
  return WM_OK;
}

// ext_bounds (device, [6] int64, or null): voxel-index bounds to key with instead of the bounds of `means` -- the bounds of the
// WHOLE scene when this call merges one key range of it.
wm_status prune_splats(cudaStream_t st, long long N, const float* means, const float* quats, const float* scales,
                       const float* opacities, const float* sh, const float* weights, double voxel_size, void* workspace,
                       size_t workspace_bytes, float* out_means, float* out_quats, float* out_scales, float* out_opacities, float* out_sh,
                       unsigned* count, const long long* ext_bounds) {
  WM_CHECK(N > 0 && N < (1ll << 31) && means && quats && scales && opacities && sh && weights && workspace && out_means && out_quats &&
               out_scales && out_opacities && out_sh && count && voxel_size > 0,
           "prune_splats: bad argument");
  PruneWs w{};
  size_t need = 0;
  WM_TRY(prune_layout((unsigned)N, workspace, &need, &w));
  WM_CHECK(workspace_bytes >= need, "prune_splats: workspace of %zu bytes, %zu needed", workspace_bytes, need);
  const unsigned n = (unsigned)N;
  const float vs = (float)voxel_size;  // the Python float becomes an fp32 divisor
  const int grid = (int)std::min<size_t>(((size_t)n + SP_THREADS - 1) / SP_THREADS, 148 * 16);
  if (ext_bounds != nullptr) {
    WM_CUDA(cudaMemcpyAsync(w.bounds, ext_bounds, 6 * sizeof(long long), cudaMemcpyDeviceToDevice, st));
  } else {
    prune_init_kernel<<<1, 32, 0, st>>>(w.bounds);
    prune_bounds_kernel<<<grid, SP_THREADS, 0, st>>>(means, n, vs, w.bounds);
  }
  prune_keys_kernel<<<grid, SP_THREADS, 0, st>>>(means, n, vs, w.bounds, w.keys_in, w.vals_in);
  WM_LAUNCH_CHECK();
  size_t tmp = w.cub_bytes;
  WM_CUDA(cub::DeviceRadixSort::SortPairs(w.cub_tmp, tmp, (const long long*)w.keys_in, w.keys_out, (const unsigned*)w.vals_in, w.vals_out,
                                          (int)n, 0, 64, st));
  prune_flags_kernel<<<grid, SP_THREADS, 0, st>>>(w.keys_out, n, w.flags);
  WM_LAUNCH_CHECK();
  tmp = w.cub_bytes;
  WM_CUDA(cub::DeviceScan::InclusiveSum(w.cub_tmp, tmp, (const unsigned*)w.flags, w.segid, (int)n, st));
  SplatOut in{const_cast<float*>(means), const_cast<float*>(quats), const_cast<float*>(scales), const_cast<float*>(opacities),
              const_cast<float*>(sh), const_cast<float*>(weights)};
  SplatOut out{out_means, out_quats, out_scales, out_opacities, out_sh, nullptr};
  prune_merge_kernel<<<grid, SP_THREADS, 0, st>>>(w.keys_out, w.vals_out, w.segid, n, in, out, count);
  WM_LAUNCH_CHECK();
  return WM_OK;
}

// ------------------------------------------------------------------------------------------------ confidence filter
// apply_confidence_filter (rasterization.py:248-299): confidences <= 1e-5 become -inf, the K most confident splats are
// kept (torch.topk, sorted=False: its output order and its choice among equal values are unspecified).  Here the choice is
// made deterministic: the K-th largest key is found with an exact 4-pass radix select (8 bits per pass, block-private
// shared-memory histograms, 256 global atomics per block), equal keys are taken in ascending index, and the survivors are
// written in ascending original index, so that the voxel merge behind it keeps accumulating in index order.
// HBM-bound integer work: 4 x 4 B per splat for the select, 16 B for the scan, 64 B read + 64 B written per kept splat.
struct ConfWs {
  unsigned* state;  // [0] key prefix found so far, [1] how many are still to take among the keys matching the prefix
  unsigned* hist;   // [4][256]
  unsigned long long *packed, *scanned;  // low word: keys above the threshold before i; high word: keys equal to it
  void* cub_tmp;
  size_t cub_bytes;
};

// order-preserving key of a masked confidence; NaN (either sign) sorts above +inf, as torch.topk treats it
__device__ __forceinline__ unsigned conf_key(float c) {
  if (c != c) return 0xFFFFFFFFu;
  if (c <= 1e-5f) c = -INFINITY;
  const unsigned u = __float_as_uint(c);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

__global__ void conf_init_kernel(unsigned* state, unsigned* hist, unsigned K) {
  for (int i = threadIdx.x; i < 4 * 256; i += blockDim.x) hist[i] = 0u;
  if (threadIdx.x == 0) state[0] = 0u, state[1] = K;
}

__global__ void __launch_bounds__(SP_THREADS)
conf_hist_kernel(const float* __restrict__ conf, unsigned N, int pass, const unsigned* __restrict__ state, unsigned* __restrict__ hist) {
  __shared__ unsigned s_h[256];
  s_h[threadIdx.x] = 0u;  // SP_THREADS == 256
  __syncthreads();
  const int shift = 24 - 8 * pass;
  const unsigned mask = pass == 0 ? 0u : (0xFFFFFFFFu << (shift + 8)), prefix = state[0];
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
    const unsigned k = conf_key(__ldg(conf + i));
    if ((k & mask) == prefix) atomicAdd(&s_h[(k >> shift) & 255u], 1u);
  }
  __syncthreads();
  if (s_h[threadIdx.x]) atomicAdd(hist + pass * 256 + threadIdx.x, s_h[threadIdx.x]);
}

// the digit of this pass: the largest d with |{digit > d}| < remaining <= |{digit >= d}|
__global__ void conf_pick_kernel(int pass, unsigned* state, const unsigned* __restrict__ hist) {
  __shared__ unsigned s_h[256];
  s_h[threadIdx.x] = hist[pass * 256 + threadIdx.x];  // 256 threads: one coalesced read instead of 256 dependent ones
  __syncthreads();
  if (threadIdx.x != 0) return;
  const unsigned rem = state[1];
  unsigned acc = 0u;
  for (int d = 255; d >= 0; --d) {
    const unsigned h = s_h[d];
    if (acc + h >= rem) {
      state[0] |= (unsigned)d << (24 - 8 * pass);
      state[1] = rem - acc;
      return;
    }
    acc += h;
  }
}

__global__ void __launch_bounds__(SP_THREADS)
conf_mark_kernel(const float* __restrict__ conf, unsigned N, const unsigned* __restrict__ state, unsigned long long* __restrict__ packed) {
  const unsigned T = state[0];
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
    const unsigned k = conf_key(__ldg(conf + i));
    packed[i] = k > T ? 1ull : (k == T ? (1ull << 32) : 0ull);
  }
}

__global__ void __launch_bounds__(SP_THREADS)
conf_gather_kernel(const float* __restrict__ conf, unsigned N, const unsigned* __restrict__ state,
                   const unsigned long long* __restrict__ scanned, SplatOut in, SplatOut out) {
  const unsigned T = state[0], need = state[1];
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
    const unsigned k = conf_key(__ldg(conf + i));
    if (k < T) continue;
    const unsigned long long s = scanned[i];
    const unsigned above = (unsigned)s, ties = (unsigned)(s >> 32);
    if (k == T && ties >= need) continue;
    const size_t o = (size_t)above + (ties < need ? ties : need), src = i;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      out.means[o * 3 + d] = __ldg(in.means + src * 3 + d);
      out.scales[o * 3 + d] = __ldg(in.scales + src * 3 + d);
      out.sh[o * 3 + d] = __ldg(in.sh + src * 3 + d);
    }
#pragma unroll
    for (int d = 0; d < 4; ++d) out.quats[o * 4 + d] = __ldg(in.quats + src * 4 + d);
    out.opacities[o] = __ldg(in.opacities + src);
    out.weights[o] = __ldg(in.weights + src);
  }
}

static wm_status conf_layout(unsigned N, void* base, size_t* total, ConfWs* ws) {
  size_t scan_bytes = 0;
  WM_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, (const unsigned long long*)nullptr, (unsigned long long*)nullptr, (int)N,
                                        (cudaStream_t)0));
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* p = base ? static_cast<char*>(base) + off : nullptr;
    off += sp_align(bytes);
    return p;
  };
  ConfWs w{};
  w.state = static_cast<unsigned*>(take(2 * sizeof(unsigned)));
  w.hist = static_cast<unsigned*>(take(4 * 256 * sizeof(unsigned)));
  w.packed = static_cast<unsigned long long*>(take((size_t)N * 8));
  w.scanned = static_cast<unsigned long long*>(take((size_t)N * 8));
  w.cub_bytes = scan_bytes;
  w.cub_tmp = take(scan_bytes);
  *total = off;
  if (ws) *ws = w;
  return WM_OK;
}

wm_status conf_filter_workspace_bytes(long long N, size_t* bytes) {
  WM_CHECK(N > 0 && N < (1ll << 31) && bytes, "conf_filter_workspace_bytes: bad argument");
  return conf_layout((unsigned)N, nullptr, bytes, nullptr);
}

// N splats -> the K most confident, in ascending original index; outputs sized for K rows.
wm_status conf_filter_splats(cudaStream_t st, long long N, long long K, const float* conf, const float* means, const float* quats,
                             const float* scales, const float* opacities, const float* sh, const float* weights, void* workspace,
                             size_t workspace_bytes, float* out_means, float* out_quats, float* out_scales, float* out_opacities,
                             float* out_sh, float* out_weights) {
  WM_CHECK(N > 0 && N < (1ll << 31) && K > 0 && K <= N && conf && means && quats && scales && opacities && sh && weights && workspace &&
               out_means && out_quats && out_scales && out_opacities && out_sh && out_weights,
           "conf_filter_splats: bad argument");
  ConfWs w{};
  size_t need = 0;
  WM_TRY(conf_layout((unsigned)N, workspace, &need, &w));
  WM_CHECK(workspace_bytes >= need, "conf_filter_splats: workspace of %zu bytes, %zu needed", workspace_bytes, need);
  const unsigned n = (unsigned)N;
  const int grid = (int)std::min<size_t>(((size_t)n + SP_THREADS - 1) / SP_THREADS, 148 * 8);
  conf_init_kernel<<<1, 256, 0, st>>>(w.state, w.hist, (unsigned)K);
  for (int pass = 0; pass < 4; ++pass) {
    conf_hist_kernel<<<grid, SP_THREADS, 0, st>>>(conf, n, pass, w.state, w.hist);
    conf_pick_kernel<<<1, 256, 0, st>>>(pass, w.state, w.hist);
  }
  conf_mark_kernel<<<grid, SP_THREADS, 0, st>>>(conf, n, w.state, w.packed);
  WM_LAUNCH_CHECK();
  size_t tmp = w.cub_bytes;
  WM_CUDA(cub::DeviceScan::ExclusiveSum(w.cub_tmp, tmp, (const unsigned long long*)w.packed, w.scanned, (int)n, st));
  SplatOut in{const_cast<float*>(means), const_cast<float*>(quats), const_cast<float*>(scales), const_cast<float*>(opacities),
              const_cast<float*>(sh), const_cast<float*>(weights)};
  SplatOut out{out_means, out_quats, out_scales, out_opacities, out_sh, out_weights};
  conf_gather_kernel<<<grid, SP_THREADS, 0, st>>>(conf, n, w.state, w.scanned, in, out);
  WM_LAUNCH_CHECK();
  return WM_OK;
}


}  // namespace wm
