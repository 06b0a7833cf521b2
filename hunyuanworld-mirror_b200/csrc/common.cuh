// Host-side plumbing shared by every translation unit of libwm_b200: status codes, the thread-local
// error string behind wm_last_error(), CUDA error capture, and CUtensorMap construction through the
// driver entry point (no link-time dependency on libcuda, so the library also loads on a GPU-less box).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/wm_b200.h"

namespace wm {

constexpr int kEmbed = 1024;
constexpr int kHeads = 16;
constexpr int kHeadDim = 64;
constexpr int kMlp = 4096;
constexpr int kPatch = 14;
constexpr int kPatchStart = 7;   // cam + 4 reg + pose + ray (visual_transformer.py:99-101)
constexpr int kDinoSpecial = 5;  // cls + 4 reg (vision_transformer.py:209-221)
constexpr int kNumSms = 148;

char* error_buffer();  // thread-local, 512 bytes (capi.cu)

// What flash_attention() needs to know about a K|V gather that is still in flight when it launches (nullptr: none):
// shard r of the key buffer may be read once (int)(flags[r] - epoch) >= 0 (attn_common.cuh, capi.cu wm_p2p_*).
struct AttnGather {
  int shard_rows, self_rank;
  unsigned epoch;
  const unsigned* flags;
  unsigned* err;
};

// Data-dependent choice between the bounded-logit and the running-max attention kernels, made ON THE DEVICE (no host
// read): the QKV epilogue leaves max |q_row|^2 (q already multiplied by scale * log2 e) and max |k_row|^2 per head in
// norm2 [2 * heads] (float bits, accumulated with atomicMax: zero them before the GEMM), so |log2-logit| <= |q| |k| is
// known when the attention launches.  Mode 3 (variants 10 + 4): ONE launch of a kernel that holds both forms of the softmax
// role and branches on the words.  Modes 1 / 2 (any other pair): both kernels are launched; mode 1 (bounded kernel) returns
// at once unless every head's bound is <= limit, mode 2 (running-max kernel) unless some head's bound is above it or NaN
// (measured: the launch that returns at once still costs ~46 us behind a 200 KB-shared-memory kernel, tools/bench_guard.py).
struct AttnGuard {
  const unsigned* norm2;
  int mode;     // 1: run only if the bound holds; 2: run only if it does not; 3: one launch, the device branches
  float limit;  // on |q| |k| in log2 units (kMaxLogitBound)
};

inline wm_status fail(wm_status code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(error_buffer(), 512, fmt, ap);
  va_end(ap);
  return code;
}

#define WM_CUDA(expr)                                                                                   \
  do {                                                                                                  \
    cudaError_t _e = (expr);                                                                            \
    if (_e != cudaSuccess)                                                                              \
      return ::wm::fail(WM_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                        __LINE__);                                                                      \
  } while (0)

#define WM_CHECK(cond, ...)                                    \
  do {                                                         \
    if (!(cond)) return ::wm::fail(WM_ERR_INVALID, __VA_ARGS__); \
  } while (0)

#define WM_TRY(expr)                 \
  do {                               \
    wm_status _s = (expr);           \
    if (_s != WM_OK) return _s;      \
  } while (0)

// Launch-error check that is legal during stream capture (no sync).
#define WM_LAUNCH_CHECK() WM_CUDA(cudaGetLastError())

// ------------------------------------------------------------------ tensor maps
wm_status tmap_encode(CUtensorMap* out, CUtensorMapDataType dtype, int rank, const void* base, const uint64_t* dims,
                      const uint64_t* strides_bytes /* rank-1 entries */, const uint32_t* box, CUtensorMapSwizzle sw);

// 2D row-major [rows, cols] with a row pitch in elements; box = [box_cols, box_rows]; SWIZZLE_128B
inline wm_status tmap_2d(CUtensorMap* out, CUtensorMapDataType dtype, int elem_bytes, const void* base, uint64_t rows,
                         uint64_t cols, uint64_t pitch_elems, uint32_t box_cols, uint32_t box_rows) {
  uint64_t dims[2] = {cols, rows};
  uint64_t strides[1] = {pitch_elems * (uint64_t)elem_bytes};
  uint32_t box[2] = {box_cols, box_rows};
  return tmap_encode(out, dtype, 2, base, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
}
// 3D [batch, rows, cols]: rows inside one batch entry are `pitch_elems` apart, batch entries `batch_pitch_elems`.
inline wm_status tmap_3d(CUtensorMap* out, CUtensorMapDataType dtype, int elem_bytes, const void* base, uint64_t batch,
                         uint64_t rows, uint64_t cols, uint64_t pitch_elems, uint64_t batch_pitch_elems,
                         uint32_t box_cols, uint32_t box_rows) {
  uint64_t dims[3] = {cols, rows, batch};
  uint64_t strides[2] = {pitch_elems * (uint64_t)elem_bytes, batch_pitch_elems * (uint64_t)elem_bytes};
  uint32_t box[3] = {box_cols, box_rows, 1};
  return tmap_encode(out, dtype, 3, base, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

// 4D NHWC activations [n, H, W, C] (C contiguous); box = [box_c, box_w, box_h, 1]; SWIZZLE_128B (box_c * elem = 128 B)
inline wm_status tmap_nhwc(CUtensorMap* out, CUtensorMapDataType dtype, int elem_bytes, const void* base, uint64_t n,
                           uint64_t H, uint64_t W, uint64_t C, uint32_t box_c, uint32_t box_w, uint32_t box_h) {
  uint64_t dims[4] = {C, W, H, n};
  uint64_t strides[3] = {C * elem_bytes, W * C * elem_bytes, H * W * C * elem_bytes};
  uint32_t box[4] = {box_c, box_w, box_h, 1};
  return tmap_encode(out, dtype, 4, base, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

wm_status require_sm100(int* sm_count);

// Launch with programmatic stream serialization (the kernel must call pdl_wait() before touching global data).
template <typename... KArgs, typename... Args>
inline wm_status launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid, cfg.blockDim = block, cfg.dynamicSmemBytes = smem, cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr, cfg.numAttrs = 1;
  WM_CUDA(cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...));
  return WM_OK;
}

}  // namespace wm
