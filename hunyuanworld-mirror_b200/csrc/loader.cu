// Image loader on the GPU (second half; reference: src/utils/inference_utils.py:187-248
// prepare_images_to_tensor).  File decoding stays on the host (PIL); everything after it runs here on the decoded RGB
// bytes: Pillow's Image.resize(BICUBIC) (a third-party dependency of the reference, restated from its published
// algorithm, src/libImaging/Resample.c), torchvision's ToTensor (uint8 HWC -> fp32 CHW / 255), the centre crop and the
// constant-1.0 padding, written straight into the view's slot of the [N, 3, H, W] batch.
// Integer / byte work, HBM-bound: the bar is bit-exact.  Pillow resamples in two separable passes with a uint8
// intermediate image: coefficients = bicubic (a = -0.5) over a support stretched by the down-scaling factor, normalised in
// double, rounded to 22-bit fixed point; pixel = clip8((2^21 + sum src * k) >> 22).  The coefficient tables are tiny
// ((new_w + new_h) * ksize ints) and are computed on the host in double exactly as Pillow does, then sent along.
#include <cmath>
#include <vector>

#include "common.cuh"
#include "ptx.cuh"

namespace wm {

constexpr int LD_PRECISION_BITS = 32 - 8 - 2;

static double ld_bicubic(double x) {
  const double a = -0.5;
  if (x < 0.0) x = -x;
  if (x < 1.0) return ((a + 2.0) * x - (a + 3.0)) * x * x + 1;
  if (x < 2.0) return (((x - 5) * x + 8) * x - 4) * a;
  return 0.0;
}

// precompute_coeffs + normalize_coeffs_8bpc for the whole-image box.  table = [out][2 + ksize]: first, count, k[ksize]
static int ld_coefficients(int in_size, int out_size, std::vector<int>& table) {
  double scale = (double)in_size / out_size, filterscale = scale;
  if (filterscale < 1.0) filterscale = 1.0;
  const double support = 2.0 * filterscale;
  const int ksize = (int)std::ceil(support) * 2 + 1;
  table.assign((size_t)out_size * (2 + ksize), 0);
  const double ss = 1.0 / filterscale;
  std::vector<double> w(ksize);
  for (int xx = 0; xx < out_size; ++xx) {
    const double center = (xx + 0.5) * scale;
    int xmin = (int)(center - support + 0.5);
    if (xmin < 0) xmin = 0;
    int xmax = (int)(center + support + 0.5);
    if (xmax > in_size) xmax = in_size;
    xmax -= xmin;
    double ww = 0.0;
    for (int x = 0; x < xmax; ++x) {
      w[x] = ld_bicubic((x + xmin - center + 0.5) * ss);
      ww += w[x];
    }
    int* row = table.data() + (size_t)xx * (2 + ksize);
    row[0] = xmin, row[1] = xmax;
    for (int x = 0; x < xmax; ++x) {
      const double v = ww != 0.0 ? w[x] / ww : w[x];
      row[2 + x] = v < 0 ? (int)(-0.5 + v * (1 << LD_PRECISION_BITS)) : (int)(0.5 + v * (1 << LD_PRECISION_BITS));
    }
  }
  return ksize;
}

__device__ __forceinline__ unsigned ld_clip8(int v) { return (unsigned)min(max(v >> LD_PRECISION_BITS, 0), 255); }

// horizontal pass: src [H, W, 3] uint8 -> tmp [H, new_w, 3] uint8
__global__ void __launch_bounds__(256) loader_horizontal_kernel(const unsigned char* __restrict__ src, int H, int W, int new_w,
                                                                const int* __restrict__ table, int ksize, unsigned char* __restrict__ tmp) {
  const size_t total = (size_t)H * new_w;
  for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
    const int xo = (int)(t % new_w);
    const size_t y = t / new_w;
    const int* row = table + (size_t)xo * (2 + ksize);
    const int first = row[0], count = row[1];
    int a0 = 1 << (LD_PRECISION_BITS - 1), a1 = a0, a2 = a0;
    const unsigned char* p = src + (y * W + first) * 3;
    for (int k = 0; k < count; ++k) {
      const int c = row[2 + k];
      a0 += p[3 * k] * c, a1 += p[3 * k + 1] * c, a2 += p[3 * k + 2] * c;
    }
    unsigned char* o = tmp + t * 3;
    o[0] = (unsigned char)ld_clip8(a0), o[1] = (unsigned char)ld_clip8(a1), o[2] = (unsigned char)ld_clip8(a2);
  }
}

// vertical pass + ToTensor + crop + pad: canvas pixel (yc, xc) <- resized pixel (yc - pad_top + crop_top, xc - pad_left), 1.0 outside
__global__ void __launch_bounds__(256)
loader_vertical_kernel(const unsigned char* __restrict__ tmp, int new_w, int new_h, const int* __restrict__ table, int ksize, int crop_top,
                       int crop_h, int out_H, int out_W, int pad_top, int pad_left, float* __restrict__ out) {
  const size_t plane = (size_t)out_H * out_W;
  for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < plane; t += (size_t)gridDim.x * blockDim.x) {
    const int xc = (int)(t % out_W), yc = (int)(t / out_W);
    const int x = xc - pad_left, yr = yc - pad_top;
    float r = 1.0f, g = 1.0f, b = 1.0f;
    if (x >= 0 && x < new_w && yr >= 0 && yr < crop_h) {
      const int* row = table + (size_t)(yr + crop_top) * (2 + ksize);
      const int first = row[0], count = row[1];
      int a0 = 1 << (LD_PRECISION_BITS - 1), a1 = a0, a2 = a0;
      const unsigned char* p = tmp + ((size_t)first * new_w + x) * 3;
      for (int k = 0; k < count; ++k) {
        const int c = row[2 + k];
        a0 += p[0] * c, a1 += p[1] * c, a2 += p[2] * c;
        p += (size_t)new_w * 3;
      }
      r = __fdiv_rn((float)ld_clip8(a0), 255.0f), g = __fdiv_rn((float)ld_clip8(a1), 255.0f), b = __fdiv_rn((float)ld_clip8(a2), 255.0f);
    }
    out[t] = r, out[plane + t] = g, out[2 * plane + t] = b;
  }
}

static size_t ld_align(size_t x) { return (x + 255) & ~size_t(255); }

wm_status load_image_workspace_bytes(int H, int W, int new_w, int new_h, size_t* bytes) {
  WM_CHECK(H > 0 && W > 0 && new_w > 0 && new_h > 0 && bytes, "load_image_workspace_bytes: bad argument");
  std::vector<int> th, tv;
  ld_coefficients(W, new_w, th);
  ld_coefficients(H, new_h, tv);
  *bytes = ld_align(th.size() * 4) + ld_align(tv.size() * 4) + ld_align((size_t)H * new_w * 3);
  return WM_OK;
}

// rgb: device uint8 [H, W, 3].  out: device fp32 [3, out_H, out_W] = the resized image's rows [crop_top, crop_top + crop_h)
// placed at (pad_top, pad_left), 1.0 elsewhere.
wm_status load_image(cudaStream_t st, const unsigned char* rgb, int H, int W, int new_w, int new_h, int crop_top, int crop_h, int out_H,
                     int out_W, int pad_top, int pad_left, float* out, void* workspace, size_t workspace_bytes) {
  WM_CHECK(rgb && out && workspace && H > 0 && W > 0 && new_w > 0 && new_h > 0, "load_image: bad argument");
  WM_CHECK(crop_top >= 0 && crop_h > 0 && crop_top + crop_h <= new_h && pad_top >= 0 && pad_left >= 0 && pad_top + crop_h <= out_H &&
               pad_left + new_w <= out_W,
           "load_image: crop / pad geometry does not fit (%d+%d of %d rows at %d,%d in %dx%d)", crop_top, crop_h, new_h, pad_top, pad_left,
           out_H, out_W);
  std::vector<int> th, tv;
  const int kh = ld_coefficients(W, new_w, th), kv = ld_coefficients(H, new_h, tv);
  const size_t need = ld_align(th.size() * 4) + ld_align(tv.size() * 4) + ld_align((size_t)H * new_w * 3);
  WM_CHECK(workspace_bytes >= need, "load_image: workspace of %zu bytes, %zu needed", workspace_bytes, need);
  char* p = static_cast<char*>(workspace);
  int* d_th = reinterpret_cast<int*>(p);
  int* d_tv = reinterpret_cast<int*>(p + ld_align(th.size() * 4));
  unsigned char* tmp = reinterpret_cast<unsigned char*>(p + ld_align(th.size() * 4) + ld_align(tv.size() * 4));
  // pageable source: the runtime stages the bytes before returning, so the vectors may go out of scope
  WM_CUDA(cudaMemcpyAsync(d_th, th.data(), th.size() * 4, cudaMemcpyHostToDevice, st));
  WM_CUDA(cudaMemcpyAsync(d_tv, tv.data(), tv.size() * 4, cudaMemcpyHostToDevice, st));
  const size_t n1 = (size_t)H * new_w, n2 = (size_t)out_H * out_W;
  loader_horizontal_kernel<<<(unsigned)std::min<size_t>((n1 + 255) / 256, 148 * 16), 256, 0, st>>>(rgb, H, W, new_w, d_th, kh, tmp);
  WM_LAUNCH_CHECK();
  loader_vertical_kernel<<<(unsigned)std::min<size_t>((n2 + 255) / 256, 148 * 16), 256, 0, st>>>(tmp, new_w, new_h, d_tv, kv, crop_top, crop_h,
                                                                                                 out_H, out_W, pad_top, pad_left, out);
  WM_LAUNCH_CHECK();
  return WM_OK;
}

// ---- no-allocation variant (no counterpart in the reference) -------------
// Callers that must not allocate (CUDA-graph capture around them, tight serving loops) get the same two passes with the
// scratch (both coefficient tables + the uint8 intermediate image) in a file-scope device array instead of a caller
// buffer.  Every GPU that loads the library holds kLoadStaticBytes until exit.  The array is shared by all calls and
// keeps the last call's contents: every byte a call reads it has written itself first (tables by the two copies, the
// intermediate image by the horizontal pass), so nothing has to start at zero; two calls that overlap on the GPU
// (different streams, no ordering between them) race on it.
constexpr size_t kLoadStaticBytes = (size_t)8 << 20;
static __device__ __align__(256) unsigned char g_load_scratch[kLoadStaticBytes];

// In host code `g_load_scratch` is a host-side stand-in, not a device address: ask the runtime once per device, with
// that device current.
static wm_status load_scratch_address(void** out) {
  constexpr int kMaxDev = 64;
  static void* addr[kMaxDev] = {nullptr};
  int dev = -1;
  WM_CUDA(cudaGetDevice(&dev));
  WM_CHECK(dev >= 0 && dev < kMaxDev, "load_image_static: device ordinal %d out of range", dev);
  void* p = __atomic_load_n(&addr[dev], __ATOMIC_ACQUIRE);
  if (p == nullptr) {
    WM_CUDA(cudaGetSymbolAddress(&p, g_load_scratch));
    WM_CHECK(p != nullptr && (reinterpret_cast<uintptr_t>(p) & 255) == 0, "load_image_static: bad scratch address");
    __atomic_store_n(&addr[dev], p, __ATOMIC_RELEASE);  // every thread would store the same value
  }
  *out = p;
  return WM_OK;
}

size_t load_image_static_capacity() { return kLoadStaticBytes; }

wm_status load_image_static(cudaStream_t st, const unsigned char* rgb, int H, int W, int new_w, int new_h, int crop_top, int crop_h,
                            int out_H, int out_W, int pad_top, int pad_left, float* out) {
  size_t need = 0;
  WM_TRY(load_image_workspace_bytes(H, W, new_w, new_h, &need));
  if (need > kLoadStaticBytes)
    return fail(WM_ERR_INVALID,
                "load_image_static: a %dx%d image resized to %dx%d needs %zu bytes of scratch, the static array holds %zu "
                "(use wm_load_image with a caller workspace)",
                W, H, new_w, new_h, need, kLoadStaticBytes);
  void* scratch = nullptr;
  WM_TRY(load_scratch_address(&scratch));
  return load_image(st, rgb, H, W, new_w, new_h, crop_top, crop_h, out_H, out_W, pad_top, pad_left, out, scratch, kLoadStaticBytes);
}

}  // namespace wm
