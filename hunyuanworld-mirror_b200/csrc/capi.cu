// C ABI of libwm_b200 (include/wm_b200.h): handle, weight ingestion under the reference state_dict keys,
// workspace carving, the forward pass of the aggregator as a stream-ordered launch sequence, and the NCCL
// K|V all-gather used by sequence-parallel global attention.  No allocation, sync or host round-trip happens
// inside wm_forward / wm_forward_tokens / wm_embed.
#include <dlfcn.h>
#include <math.h>
#include <stdlib.h>

#include <string>
#include <unordered_map>
#include <vector>

#include "ops.cuh"

namespace wm {

char* error_buffer() {
  static thread_local char buf[512] = {0};
  return buf;
}

wm_status tmap_encode(CUtensorMap* out, CUtensorMapDataType dtype, int rank, const void* base, const uint64_t* dims,
                      const uint64_t* strides_bytes, const uint32_t* box, CUtensorMapSwizzle sw) {
  typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static encode_fn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    WM_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    if (!p || q != cudaDriverEntryPointSuccess) return fail(WM_ERR_CUDA, "cuTensorMapEncodeTiled is not available in this driver");
    fn = reinterpret_cast<encode_fn>(p);
  }
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  cuuint64_t d[5], s[4];
  cuuint32_t b[5];
  for (int i = 0; i < rank; ++i) d[i] = dims[i], b[i] = box[i];
  for (int i = 0; i + 1 < rank; ++i) s[i] = strides_bytes[i];
  const CUresult r = fn(out, dtype, (cuuint32_t)rank, const_cast<void*>(base), d, s, b, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(WM_ERR_CUDA, "cuTensorMapEncodeTiled failed (CUresult %d): rank %d dims %llu,%llu,%llu stride0 %llu box %u,%u",
                (int)r, rank, (unsigned long long)d[0], (unsigned long long)d[1], (unsigned long long)(rank > 2 ? d[2] : 0),
                (unsigned long long)s[0], b[0], b[1]);
  return WM_OK;
}

wm_status require_sm100(int* sm_count) {
  // cudaGetDeviceProperties costs milliseconds: remember the answer per device
  static int cached_dev = -1, cached_sms = 0;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(WM_ERR_ARCH, "no CUDA device: %s (libwm_b200 has no CPU path)", cudaGetErrorString(e));
  if (dev == cached_dev) {
    *sm_count = cached_sms;
    return WM_OK;
  }
  cudaDeviceProp prop;
  e = cudaGetDeviceProperties(&prop, dev);
  if (e != cudaSuccess) return fail(WM_ERR_ARCH, "cudaGetDeviceProperties: %s", cudaGetErrorString(e));
  if (prop.major != 10)
    return fail(WM_ERR_ARCH, "device %d is sm_%d%d; libwm_b200 is built for sm_100a (B200) only", dev, prop.major, prop.minor);
  *sm_count = prop.multiProcessorCount;
  cached_dev = dev, cached_sms = prop.multiProcessorCount;
  return WM_OK;
}

// ------------------------------------------------------------------------------------------------ handle
struct DevTensor {
  float* f32 = nullptr;
  __nv_bfloat16* b16 = nullptr;
  std::vector<int64_t> shape;
  int ld = 0;
};

struct BlockW {
  const float *n1w, *n1b, *n2w, *n2b, *qkv_b, *qn_w, *qn_b, *kn_w, *kn_b, *proj_b, *ls1, *ls2, *fc1_b, *fc2_b;
  const __nv_bfloat16 *qkv_w, *proj_w, *fc1_w, *fc2_w;
  // upper bound of |q . k| * scale * log2(e) implied by the QK-LayerNorm weights (0: no bound, DINOv2 blocks)
  float logit_bound = 0.f;
};

typedef struct ncclComm* nccl_comm_t;
struct nccl_uid {
  char internal[128];
};
struct NcclApi {
  void* lib = nullptr;
  int (*GetUniqueId)(nccl_uid*) = nullptr;
  int (*CommInitRank)(nccl_comm_t*, int, nccl_uid, int) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, nccl_comm_t, cudaStream_t) = nullptr;
  int (*CommDestroy)(nccl_comm_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};

static wm_status nccl_open(const char* path, NcclApi* api) {
  WM_CHECK(path && *path, "nccl: library path is empty");
  api->lib = dlopen(path, RTLD_NOW | RTLD_GLOBAL);
  if (!api->lib) return fail(WM_ERR_NCCL, "dlopen(%s): %s", path, dlerror());
  api->GetUniqueId = reinterpret_cast<int (*)(nccl_uid*)>(dlsym(api->lib, "ncclGetUniqueId"));
  api->CommInitRank = reinterpret_cast<int (*)(nccl_comm_t*, int, nccl_uid, int)>(dlsym(api->lib, "ncclCommInitRank"));
  api->AllGather = reinterpret_cast<int (*)(const void*, void*, size_t, int, nccl_comm_t, cudaStream_t)>(
      dlsym(api->lib, "ncclAllGather"));
  api->CommDestroy = reinterpret_cast<int (*)(nccl_comm_t)>(dlsym(api->lib, "ncclCommDestroy"));
  api->GetErrorString = reinterpret_cast<const char* (*)(int)>(dlsym(api->lib, "ncclGetErrorString"));
  if (!api->GetUniqueId || !api->CommInitRank || !api->AllGather || !api->CommDestroy || !api->GetErrorString)
    return fail(WM_ERR_NCCL, "%s does not export the NCCL entry points", path);
  return WM_OK;
}

// Sequence-parallel K|V exchange over peer memory (NVLink), overlapped with the attention that consumes it.
// Every rank owns one cudaMalloc'ed block [4096 B header | buffer 0 | buffer 1]; header = arrival words flags[r]
// (one per source rank) and an error word at [64].  Buffer b holds the gathered K|V rows of the global block whose
// sequence number (epoch) has parity b: rank r's rows at r * T.  After its QKV GEMM a rank pushes its shard into every
// peer's buffer with the copy engines and then raises flags[self] = epoch there; the attention kernel scans its own
// shard first and waits per shard (attn_common.cuh).  Two buffers make acknowledgements unnecessary: a rank can only
// push epoch e after it attended epoch e-1, which needed every peer's e-1 shard, which each peer sent after it had
// finished attending e-2 -- the last reader of buffer e & 1.
constexpr int kP2pHeader = 4096;
constexpr int kP2pMaxWorld = 32;
struct PeerGather {
  bool on = false;
  int max_rows = 0;
  uint8_t* base = nullptr;
  size_t buf_bytes = 0;
  uint8_t* peer[kP2pMaxWorld] = {};
  cudaStream_t comm = nullptr;
  cudaEvent_t ev_ready = nullptr, ev_done = nullptr;
  unsigned epoch = 0;
  bool used = false;  // pushes were enqueued since the last join with the forward's stream
  // the error word of the gather header is copied to pinned host memory at the end of every forward that used the
  // exchange and looked at (without blocking) when the next one starts: a forward that ran on stale keys because a
  // shard never arrived cannot go unnoticed (wm_p2p_status reads it synchronously)
  unsigned* err_host = nullptr;
  cudaEvent_t ev_err = nullptr;
  bool err_pending = false;
};

__global__ void p2p_flag_kernel(unsigned* flag, unsigned epoch) {
  if (threadIdx.x == 0) {
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(epoch) : "memory");
  }
}

}  // namespace wm

struct wm_handle {
  wm_config cfg{};
  std::unordered_map<std::string, wm::DevTensor> w;
  bool finalized = false;
  std::vector<wm::BlockW> dino, frame, global;
  float *rope_cos = nullptr, *rope_sin = nullptr;
  int rope_max_pos = 0;
  float* dino_table = nullptr;  // [5 + gh*gw, 1024]
  int table_gh = 0, table_gw = 0;
  wm::NcclApi nccl;
  wm::nccl_comm_t comm = nullptr;
  int rank = 0, world = 1;
  wm::PeerGather p2p;
  int sp_mode = 1;  // 1: peer-memory push overlapped with attention when wm_p2p_attach succeeded; 0: ncclAllGather
  // opt-in per-kernel-class timing (CUDA events on the forward's stream) and a launch counter
  unsigned prof_mask = 0;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_ev[WM_PROF_CLASSES];
  size_t prof_used[WM_PROF_CLASSES] = {0};
  long long launches = 0;
  // tuning knobs (wm_set_int): attention kernel variants and the bounded-logit switch
  int attn_general_variant = 4, attn_bounded_variant = 10, attn_allow_bounded = 1, attn_balance = 1;
  int gemm_split = 1;             // K-split of the residual GEMMs' tail tiles (gemm.cu); 0: whole tiles only
  unsigned* gemm_flags = nullptr;  // kGemmFlagBytes of arrival counters, zero between launches
  // device-side choice of the attention kernel for blocks whose weights do not bound the logits (AttnGuard): 32 words
  // per block (DINOv2 blocks first, then frame / global pairs), cleared at the start of run_embed / run_stack
  int attn_guard = 1;
  unsigned* guard_words = nullptr;
};

namespace wm {

static bool ends_with(const std::string& s, const char* suf) {
  const size_t n = strlen(suf);
  return s.size() >= n && s.compare(s.size() - n, n, suf) == 0;
}

// ------------------------------------------------------------------------------------------------ tables
__global__ void dino_table_kernel(const float* __restrict__ cls, const float* __restrict__ reg,
                                  const float* __restrict__ pos, const float* __restrict__ conv_bias, int rows,
                                  float* __restrict__ table) {
  // row 0: cls + pos[0]; rows 1-4: register tokens (no positional term); row 5+p: pos[1+p] + conv bias
  // (vision_transformer.py:209-221; the conv bias is folded here so the GEMM epilogue has one addend).
  const int r = blockIdx.x;
  if (r >= rows) return;
  for (int c = threadIdx.x; c < kEmbed; c += blockDim.x) {
    float v;
    if (r == 0)
      v = cls[c] + pos[c];
    else if (r < kDinoSpecial)
      v = reg[(r - 1) * kEmbed + c];
    else
      v = pos[(size_t)(r - kDinoSpecial + 1) * kEmbed + c] + conv_bias[c];
    table[(size_t)r * kEmbed + c] = v;
  }
}

static wm_status build_dino_table(wm_handle* h, int gh, int gw, const float* pos_dev) {
  const int rows = kDinoSpecial + gh * gw;
  if (h->dino_table) cudaFree(h->dino_table);
  h->dino_table = nullptr;
  WM_CUDA(cudaMalloc(&h->dino_table, (size_t)rows * kEmbed * sizeof(float)));
  dino_table_kernel<<<rows, 256>>>(h->w["patch_embed.cls_token"].f32, h->w["patch_embed.register_tokens"].f32, pos_dev,
                                   h->w["patch_embed.patch_embed.proj.bias"].f32, rows, h->dino_table);
  WM_LAUNCH_CHECK();
  WM_CUDA(cudaDeviceSynchronize());
  h->table_gh = gh, h->table_gw = gw;
  return WM_OK;
}

// ------------------------------------------------------------------------------------------------ workspace
struct Workspace {
  float *xd, *x;
  __nv_bfloat16 *ln, *q, *kv, *att, *hid;
  float* attn_part;  // partials of attention work items split between CTAs (attn_split.cu)
  size_t bytes;
};

static size_t align_up(size_t v) { return (v + 1023) & ~(size_t)1023; }

static Workspace carve(void* base, int n, int gh, int gw, int world) {
  const size_t g = (size_t)gh * gw, Td = n * (kDinoSpecial + g), T = n * (kPatchStart + g);
  uint8_t* p = static_cast<uint8_t*>(base);
  size_t off = 0;
  Workspace ws;
  auto take = [&](size_t bytes) {
    uint8_t* r = p ? p + off : nullptr;
    off += align_up(bytes);
    return r;
  };
  ws.xd = reinterpret_cast<float*>(take(Td * kEmbed * 4));
  ws.x = reinterpret_cast<float*>(take(T * kEmbed * 4));
  ws.ln = reinterpret_cast<__nv_bfloat16*>(take(T * kEmbed * 2));
  ws.q = reinterpret_cast<__nv_bfloat16*>(take(T * kEmbed * 2));
  ws.kv = reinterpret_cast<__nv_bfloat16*>(take(T * (size_t)world * 2 * kEmbed * 2));
  ws.att = reinterpret_cast<__nv_bfloat16*>(take(T * kEmbed * 2));
  ws.hid = reinterpret_cast<__nv_bfloat16*>(take(T * kMlp * 2));
  ws.attn_part = reinterpret_cast<float*>(take(attn_partial_bytes(kNumSms)));
  ws.bytes = off + 1024;  // slack so the caller's base may be aligned up to 1024
  return ws;
}

static void* align_base(void* p) { return reinterpret_cast<void*>((reinterpret_cast<uintptr_t>(p) + 1023) & ~(uintptr_t)1023); }

// ------------------------------------------------------------------------------------------------ blocks
struct BlockShape {
  int T;          // rows of the residual stream on this rank
  int batch;      // attention batch (views for frame / DINOv2, 1 for global)
  int Lq, Lkv;    // per batch entry
  int kv_row0;    // row of this rank's first K|V row inside ws.kv (global blocks under SP)
  int tokens_per_view, patch_start, grid_w;
  bool rope;
  float eps;
  bool global;
};

constexpr float kMaxLogitBound = 80.0f;  // 2^80 * 2^19 keys stays far inside fp32; masked keys use 2^-120

struct ProfScope {
  wm_handle* h;
  cudaStream_t st;
  int cls;
  bool on;
  ProfScope(wm_handle* h_, cudaStream_t st_, int cls_) : h(h_), st(st_), cls(cls_), on((h_->prof_mask >> cls_) & 1u) {
    ++h->launches;
    if (!on) return;
    auto& v = h->prof_ev[cls];
    if (h->prof_used[cls] == v.size()) {
      cudaEvent_t a, b;
      cudaEventCreate(&a);
      cudaEventCreate(&b);
      v.emplace_back(a, b);
    }
    cudaEventRecord(v[h->prof_used[cls]].first, st);
  }
  ~ProfScope() {
    if (on) cudaEventRecord(h->prof_ev[cls][h->prof_used[cls]++].second, st);
  }
};
#define WM_OP(cls, expr)          \
  do {                            \
    ProfScope _ps(h, st, cls);    \
    WM_TRY(expr);                 \
  } while (0)

// Block.forward eval branch (block.py:91-92): x += ls1 * attn(norm1(x)); x += ls2 * mlp(norm2(x))
constexpr int kGuardWords = 2 * kHeads;
constexpr float kLn2 = 0.6931471805599453f;

// Attention whose kernel form is chosen on the device (AttnGuard): variants (10, 4) are one launch of a kernel that holds
// both softmax roles; any other pair is two launches of which one returns at once.
static wm_status guarded_attention(wm_handle* h, cudaStream_t st, int cls, const void* q, int ldq, const void* k, const void* v,
                                   int ldkv, void* o, int ldo, int batch, int Lq, int Lkv, int bounded_variant, int general_variant,
                                   float* part, size_t part_bytes, const unsigned* norm2, float limit) {
  if (bounded_variant == 10 && general_variant == 4) {
    const AttnGuard g3{norm2, 3, limit};
    if (h) {
      WM_OP(cls, flash_attention(st, q, ldq, k, v, ldkv, o, ldo, batch, Lq, Lkv, kHeads, kLn2, 10, part, part_bytes, nullptr, &g3));
      return WM_OK;
    }
    return flash_attention(st, q, ldq, k, v, ldkv, o, ldo, batch, Lq, Lkv, kHeads, kLn2, 10, part, part_bytes, nullptr, &g3);
  }
  const AttnGuard g1{norm2, 1, limit}, g2{norm2, 2, limit};
  if (h) {
    WM_OP(cls, flash_attention(st, q, ldq, k, v, ldkv, o, ldo, batch, Lq, Lkv, kHeads, 1.0f, bounded_variant, part, part_bytes, nullptr, &g1));
    WM_OP(cls, flash_attention(st, q, ldq, k, v, ldkv, o, ldo, batch, Lq, Lkv, kHeads, kLn2, general_variant, part, part_bytes, nullptr, &g2));
    return WM_OK;
  }
  WM_TRY(flash_attention(st, q, ldq, k, v, ldkv, o, ldo, batch, Lq, Lkv, kHeads, 1.0f, bounded_variant, part, part_bytes, nullptr, &g1));
  return flash_attention(st, q, ldq, k, v, ldkv, o, ldo, batch, Lq, Lkv, kHeads, kLn2, general_variant, part, part_bytes, nullptr, &g2);
}

static wm_status run_block(wm_handle* h, cudaStream_t st, const BlockW& w, const BlockShape& s, const Workspace& ws,
                           float* x, float* out2, int out2_col, bool gather, int guard_slot) {
  WM_OP(WM_PROF_LN, layernorm(st, x, w.n1w, w.n1b, s.eps, s.T, s.T, 0, 0, 0, 0, ws.ln, 0, 0));
  const bool p2p = gather && h->p2p.on && h->sp_mode == 1 && s.T <= h->p2p.max_rows;
  __nv_bfloat16* kv_all = ws.kv;
  unsigned epoch = 0;
  if (p2p) {
    epoch = ++h->p2p.epoch;
    kv_all = reinterpret_cast<__nv_bfloat16*>(h->p2p.base + kP2pHeader + (size_t)(epoch & 1u) * h->p2p.buf_bytes);
  }
  __nv_bfloat16* kv_local = kv_all + (size_t)s.kv_row0 * 2 * kEmbed;
  // QK-LayerNorm bounds the logits (|q . k| <= |q| |k|, both norms fixed by the norm weights): when the bound is
  // small the softmax needs no running maximum, so q is pre-multiplied by scale * log2(e) and P = 2^(q . k).
  const bool bounded = h->attn_allow_bounded && w.logit_bound > 0.f && w.logit_bound <= kMaxLogitBound;
  // No bound from the weights (DINOv2: no QK-norm; or norm gains too large): the QKV epilogue measures the q / k row
  // lengths of THIS input and the device picks the kernel (AttnGuard); keys gathered from peers keep the weight rule.
  const bool guarded = !bounded && h->attn_allow_bounded && h->attn_guard && !gather && guard_slot >= 0 && h->guard_words &&
                       h->attn_bounded_variant <= 12 && h->attn_general_variant >= 4 && h->attn_general_variant <= 7;
  unsigned* norm2 = guarded ? h->guard_words + (size_t)guard_slot * kGuardWords : nullptr;
  const float q_scale = (bounded || guarded) ? 0.125f * 1.4426950408889634f : 1.0f;
  WM_OP(WM_PROF_QKV, gemm_qkv(st, ws.ln, kEmbed, w.qkv_w, w.qkv_b, w.qn_w, w.qn_b, w.kn_w, w.kn_b, 1e-5f,
                              s.rope ? h->rope_cos : nullptr, s.rope ? h->rope_sin : nullptr, s.tokens_per_view,
                              s.patch_start, s.grid_w, q_scale, s.T, ws.q, kv_local, norm2));
  AttnGather ag{};
  if (p2p) {
    // push this rank's K|V rows into every peer's buffer (copy engines, side stream) while attention starts on them
    PeerGather& g = h->p2p;
    const size_t bytes = (size_t)s.T * 2 * kEmbed * 2;
    const size_t off = kP2pHeader + (size_t)(epoch & 1u) * g.buf_bytes + (size_t)h->rank * bytes;
    WM_CUDA(cudaEventRecord(g.ev_ready, st));
    WM_CUDA(cudaStreamWaitEvent(g.comm, g.ev_ready, 0));
    {
      ProfScope _ps(h, g.comm, WM_PROF_ALLGATHER);
      for (int k = 1; k < h->world; ++k) {
        const int peer = (h->rank - k + h->world) % h->world;  // the peer that reaches this shard first goes first
        WM_CUDA(cudaMemcpyAsync(g.peer[peer] + off, kv_local, bytes, cudaMemcpyDefault, g.comm));
        p2p_flag_kernel<<<1, 32, 0, g.comm>>>(reinterpret_cast<unsigned*>(g.peer[peer]) + h->rank, epoch);
        WM_LAUNCH_CHECK();
      }
    }
    g.used = true;
    ag.shard_rows = s.T, ag.self_rank = h->rank, ag.epoch = epoch;
    ag.flags = reinterpret_cast<const unsigned*>(g.base);
    ag.err = reinterpret_cast<unsigned*>(g.base) + 64;
  } else if (gather) {
    ProfScope _ps(h, st, WM_PROF_ALLGATHER);
    const size_t bytes = (size_t)s.T * 2 * kEmbed * 2;
    const int r = h->nccl.AllGather(kv_local, ws.kv, bytes, /*ncclChar*/ 0, h->comm, st);
    if (r != 0) return fail(WM_ERR_NCCL, "ncclAllGather: %s", h->nccl.GetErrorString(r));
  }
  const __nv_bfloat16* kbase = gather ? kv_all : kv_local;
  if (guarded) {
    // q carries scale * log2(e) already, so the running-max form gets scale = ln 2 (its scale * log2(e) is then 1)
    WM_TRY(guarded_attention(h, st, s.global ? WM_PROF_ATTN_GLOBAL : WM_PROF_ATTN_LOCAL, ws.q, kEmbed, kbase, kbase + kEmbed,
                             2 * kEmbed, ws.att, kEmbed, s.batch, s.Lq, s.Lkv, h->attn_bounded_variant, h->attn_general_variant,
                             h->attn_balance ? ws.attn_part : nullptr, attn_partial_bytes(kNumSms), norm2, kMaxLogitBound));
  } else {
    WM_OP(s.global ? WM_PROF_ATTN_GLOBAL : WM_PROF_ATTN_LOCAL,
          flash_attention(st, ws.q, kEmbed, kbase, kbase + kEmbed, 2 * kEmbed, ws.att, kEmbed, s.batch, s.Lq, s.Lkv, kHeads,
                          0.125f, bounded ? h->attn_bounded_variant : h->attn_general_variant,
                          h->attn_balance ? ws.attn_part : nullptr, attn_partial_bytes(kNumSms), p2p ? &ag : nullptr));
  }
  unsigned* gflags = h->gemm_split ? h->gemm_flags : nullptr;
  WM_OP(WM_PROF_PROJ, gemm_ls_residual(st, ws.att, kEmbed, w.proj_w, kEmbed, w.proj_b, w.ls1, s.T, kEmbed, kEmbed, x,
                                       kEmbed, gflags));
  WM_OP(WM_PROF_LN, layernorm(st, x, w.n2w, w.n2b, s.eps, s.T, s.T, 0, 0, 0, 0, ws.ln, 0, 0));
  WM_OP(WM_PROF_FC1, gemm_bias_act(st, ws.ln, kEmbed, w.fc1_w, kEmbed, w.fc1_b, s.T, kMlp, kEmbed, ws.hid, kMlp, 1));
  WM_OP(WM_PROF_FC2, gemm_ls_residual(st, ws.hid, kMlp, w.fc2_w, kMlp, w.fc2_b, w.ls2, s.T, kEmbed, kMlp, x, kEmbed, gflags));
  if (out2) WM_OP(WM_PROF_EMBED, copy_cols(st, x, s.T, out2, 2 * kEmbed, out2_col));
  return WM_OK;
}

static wm_status check_ready(const wm_handle* h, const void* workspace, size_t workspace_bytes, int n, int gh, int gw) {
  WM_CHECK(h, "null handle");
  if (!h->finalized) return fail(WM_ERR_STATE, "weights are not finalised (call wm_finalize_weights)");
  WM_CHECK(n > 0 && gh > 0 && gw > 0, "n_views and the patch grid must be positive");
  WM_CHECK(gh + 1 <= h->rope_max_pos && gw + 1 <= h->rope_max_pos, "patch grid %dx%d exceeds the RoPE table (%d)", gh, gw,
           h->rope_max_pos);
  const size_t need = carve(nullptr, n, gh, gw, h->world).bytes;
  if (!workspace || workspace_bytes < need)
    return fail(WM_ERR_STATE, "workspace too small: %zu bytes given, %zu needed", workspace_bytes, need);
  return WM_OK;
}

// A shard that did not arrive within the time limit lets the attention kernel finish on stale keys and sets the error
// word; surface it at the latest when the next forward starts (clearing it, so that a recovered job can go on).
static wm_status p2p_check_previous(wm_handle* h, cudaStream_t st, bool block) {
  PeerGather& g = h->p2p;
  if (!g.err_pending) return WM_OK;
  if (block)
    WM_CUDA(cudaEventSynchronize(g.ev_err));
  else if (cudaEventQuery(g.ev_err) != cudaSuccess) {
    cudaGetLastError();
    return WM_OK;  // still in flight: looked at again by the next call
  }
  g.err_pending = false;
  if (*g.err_host != 0) {
    *g.err_host = 0;
    WM_CUDA(cudaMemsetAsync(g.base + 64 * sizeof(unsigned), 0, sizeof(unsigned), st));
    return fail(WM_ERR_STATE, "peer-memory K|V exchange: a shard did not arrive within 4 s in an earlier forward; its "
                              "outputs were computed on stale keys (a peer is slow, desynchronised or gone)");
  }
  return WM_OK;
}

static wm_status run_stack(wm_handle* h, cudaStream_t st, const Workspace& ws, int n, int gh, int gw, float* const* out) {
  const int P = kPatchStart + gh * gw, T = n * P;
  WM_TRY(p2p_check_previous(h, st, false));
  BlockShape fs{T, n, P, P, 0, P, kPatchStart, gw, true, 1e-5f, false};
  BlockShape gs{T, 1, T, T * h->world, h->rank * T, P, kPatchStart, gw, true, 1e-5f, true};
  int next_out = 0;
  const int slot0 = h->cfg.dino_depth;
  if (h->guard_words && h->attn_guard)
    WM_CUDA(cudaMemsetAsync(h->guard_words + (size_t)slot0 * kGuardWords, 0, (size_t)2 * h->cfg.depth * kGuardWords * 4, st));
  for (int i = 0; i < h->cfg.depth; ++i) {
    float* o = nullptr;
    if (next_out < h->cfg.n_intermediate && h->cfg.intermediate[next_out] == i) o = out[next_out++];
    WM_TRY(run_block(h, st, h->frame[i], fs, ws, ws.x, o, 0, false, slot0 + 2 * i));
    WM_TRY(run_block(h, st, h->global[i], gs, ws, ws.x, o, kEmbed, h->world > 1, slot0 + 2 * i + 1));
  }
  if (h->p2p.used) {  // the caller's stream must also cover the pushes this rank still has in flight
    WM_CUDA(cudaEventRecord(h->p2p.ev_done, h->p2p.comm));
    WM_CUDA(cudaStreamWaitEvent(st, h->p2p.ev_done, 0));
    h->p2p.used = false;
    if (h->p2p.err_host) {
      WM_CUDA(cudaMemcpyAsync(h->p2p.err_host, h->p2p.base + 64 * sizeof(unsigned), sizeof(unsigned), cudaMemcpyDeviceToHost, st));
      WM_CUDA(cudaEventRecord(h->p2p.ev_err, st));
      h->p2p.err_pending = true;
    }
  }
  return WM_OK;
}

static wm_status run_embed(wm_handle* h, cudaStream_t st, const Workspace& ws, const float* images, const float* depth,
                           const float* rays, const float* poses, const int flags[3], int n, int H, int W,
                           int first_view) {
  const int gh = H / kPatch, gw = W / kPatch, g = gh * gw;
  const int Pd = kDinoSpecial + g, P = kPatchStart + g, Td = n * Pd, T = n * P;
  if (h->table_gh != gh || h->table_gw != gw) {
    if (gh == 37 && gw == 37)
      return fail(WM_ERR_STATE, "internal: native DINOv2 table missing");
    return fail(WM_ERR_STATE, "no DINOv2 positional table for a %dx%d grid: call wm_set_dino_pos_table first", gh, gw);
  }
  // a1-a3: normalise + patch gather -> GEMM(588->1024) + pos/cls/reg table
  const int ld_pe = 640;
  __nv_bfloat16* a_pe = ws.hid;  // free until the first fc1
  WM_CUDA(cudaMemsetAsync(a_pe, 0, (size_t)Td * ld_pe * 2, st));
  WM_OP(WM_PROF_EMBED, patch_gather(st, images, n, 3, H, W, Pd, kDinoSpecial, ld_pe, 1, a_pe));
  const DevTensor& pew = h->w["patch_embed.patch_embed.proj.weight"];
  WM_OP(WM_PROF_EMBED, gemm_table(st, a_pe, ld_pe, pew.b16, pew.ld, h->dino_table, Pd, Td, kEmbed, ld_pe, ws.xd, kEmbed));
  // a4: DINOv2 blocks (no rope, no qk-norm, eps 1e-6)
  BlockShape ds{Td, n, Pd, Pd, 0, Pd, 0, gw, false, 1e-6f, false};
  if (h->guard_words && h->attn_guard)
    WM_CUDA(cudaMemsetAsync(h->guard_words, 0, (size_t)h->cfg.dino_depth * kGuardWords * 4, st));
  for (int i = 0; i < h->cfg.dino_depth; ++i) WM_TRY(run_block(h, st, h->dino[i], ds, ws, ws.xd, nullptr, 0, false, i));
  // a6 (depth): pixel-unshuffle -> fc1+GELU -> fc2 written to the patch rows of x, then the final norm accumulates
  const bool pose_on = flags[0] == 1 && poses, depth_on = flags[1] == 1 && depth, ray_on = flags[2] == 1 && rays;
  if (depth_on) {
    const int ld_d = 256;
    __nv_bfloat16* a_d = ws.q;
    WM_CUDA(cudaMemsetAsync(a_d, 0, (size_t)T * ld_d * 2, st));
    WM_OP(WM_PROF_EMBED, patch_gather(st, depth, n, 1, H, W, P, kPatchStart, ld_d, 0, a_d));
    const DevTensor& f1 = h->w["depth_embed.proj.2.fc1.weight"];
    const DevTensor& f2 = h->w["depth_embed.proj.2.fc2.weight"];
    WM_OP(WM_PROF_EMBED, gemm_bias_act(st, a_d, ld_d, f1.b16, f1.ld, h->w["depth_embed.proj.2.fc1.bias"].f32, T, kMlp, ld_d, ws.hid,
                         kMlp, 1));
    WM_OP(WM_PROF_EMBED, gemm_ls_residual(st, ws.hid, kMlp, f2.b16, f2.ld, h->w["depth_embed.proj.2.fc2.bias"].f32, nullptr, T, kEmbed,
                            kMlp, ws.x, kEmbed));
  }
  // final DINOv2 norm of the patch rows -> rows [7, 7+g) of x  (vision_transformer.py:259-263)
  WM_OP(WM_PROF_EMBED, layernorm(st, ws.xd, h->w["patch_embed.norm.weight"].f32, h->w["patch_embed.norm.bias"].f32, 1e-6f, n * g, g, Pd,
                   kDinoSpecial, P, kPatchStart, ws.x, 1, depth_on ? 1 : 0));
  // a5-a7: special + prior tokens
  WM_OP(WM_PROF_EMBED, special_tokens(st, h->w["cam_token"].f32, h->w["reg_token"].f32, n, P, first_view, !pose_on, !ray_on, ws.x));
  if (pose_on)
    WM_OP(WM_PROF_EMBED, prior_token(st, poses, n, 7, h->w["pose_embed.0.weight"].f32, h->w["pose_embed.0.bias"].f32,
                       h->w["pose_embed.2.weight"].f32, h->w["pose_embed.2.bias"].f32, P, 5, ws.x));
  if (ray_on)
    WM_OP(WM_PROF_EMBED, prior_token(st, rays, n, 4, h->w["ray_embed.0.weight"].f32, h->w["ray_embed.0.bias"].f32,
                       h->w["ray_embed.2.weight"].f32, h->w["ray_embed.2.bias"].f32, P, 6, ws.x));
  return WM_OK;
}

static void p2p_release(wm_handle* h) {
  PeerGather& g = h->p2p;
  if (g.comm) cudaStreamSynchronize(g.comm);
  for (int r = 0; r < kP2pMaxWorld; ++r)
    if (g.peer[r] && g.peer[r] != g.base) cudaIpcCloseMemHandle(g.peer[r]);
  if (g.base) cudaFree(g.base);
  if (g.ev_ready) cudaEventDestroy(g.ev_ready);
  if (g.ev_done) cudaEventDestroy(g.ev_done);
  if (g.ev_err) cudaEventDestroy(g.ev_err);
  if (g.err_host) cudaFreeHost(g.err_host);
  if (g.comm) cudaStreamDestroy(g.comm);
  g = PeerGather{};
}

}  // namespace wm

using namespace wm;

// ================================================================================================ C ABI
extern "C" {

int wm_version(void) { return 100; }
const char* wm_last_error(void) { return error_buffer(); }

wm_status wm_create(const wm_config* cfg, wm_handle** out) {
  WM_CHECK(cfg && out, "wm_create: null argument");
  WM_CHECK(cfg->depth >= 0 && cfg->depth <= 64 && cfg->dino_depth >= 0 && cfg->dino_depth <= 64, "wm_create: bad depth");
  WM_CHECK(cfg->n_intermediate >= 0 && cfg->n_intermediate <= 4, "wm_create: n_intermediate must be in [0,4]");
  for (int i = 0; i < cfg->n_intermediate; ++i)
    WM_CHECK(cfg->intermediate[i] >= 0 && cfg->intermediate[i] < cfg->depth &&
                 (i == 0 || cfg->intermediate[i] > cfg->intermediate[i - 1]),
             "wm_create: intermediate layers must be increasing and < depth");
  int sms = 0;
  WM_TRY(require_sm100(&sms));
  wm_handle* h = new (std::nothrow) wm_handle();
  WM_CHECK(h, "wm_create: out of host memory");
  h->cfg = *cfg;
  // experiment switches (the defaults are what the forward was measured with): WM_ATTN_BOUNDED_VARIANT, WM_ATTN_GENERAL_VARIANT
  if (const char* e = getenv("WM_ATTN_BOUNDED_VARIANT")) {
    const int v = atoi(e);
    if ((v >= 8 && v <= 19) || (v >= 24 && v <= 27)) h->attn_bounded_variant = v;
  }
  if (const char* e = getenv("WM_GEMM_SPLIT")) h->gemm_split = atoi(e) != 0;
  if (const char* e = getenv("WM_ATTN_GENERAL_VARIANT")) {
    const int v = atoi(e);
    if ((v >= 0 && v <= 7) || (v >= 20 && v <= 23)) h->attn_general_variant = v;
  }
  h->rope_max_pos = 512;
  std::vector<float> c((h->rope_max_pos + 1) * 16), s((h->rope_max_pos + 1) * 16);
  wm_rope_tables(h->rope_max_pos, c.data(), s.data());
  cudaError_t e = cudaMalloc(&h->rope_cos, c.size() * 4);
  if (e == cudaSuccess) e = cudaMalloc(&h->rope_sin, s.size() * 4);
  if (e == cudaSuccess) e = cudaMemcpy(h->rope_cos, c.data(), c.size() * 4, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(h->rope_sin, s.data(), s.size() * 4, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMalloc(&h->gemm_flags, kGemmFlagBytes);
  if (e == cudaSuccess) e = cudaMemset(h->gemm_flags, 0, kGemmFlagBytes);
  const size_t guard_bytes = (size_t)(h->cfg.dino_depth + 2 * h->cfg.depth) * kGuardWords * 4;
  if (e == cudaSuccess) e = cudaMalloc(&h->guard_words, guard_bytes);
  if (e == cudaSuccess) e = cudaMemset(h->guard_words, 0, guard_bytes);
  if (e != cudaSuccess) {
    wm_destroy(h);
    return fail(WM_ERR_CUDA, "wm_create: %s", cudaGetErrorString(e));
  }
  *out = h;
  return WM_OK;
}

void wm_destroy(wm_handle* h) {
  if (!h) return;
  for (auto& kv : h->w) {
    if (kv.second.f32) cudaFree(kv.second.f32);
    if (kv.second.b16) cudaFree(kv.second.b16);
  }
  if (h->rope_cos) cudaFree(h->rope_cos);
  if (h->rope_sin) cudaFree(h->rope_sin);
  if (h->dino_table) cudaFree(h->dino_table);
  if (h->gemm_flags) cudaFree(h->gemm_flags);
  if (h->guard_words) cudaFree(h->guard_words);
  if (h->comm && h->nccl.CommDestroy) h->nccl.CommDestroy(h->comm);
  p2p_release(h);
  for (auto& v : h->prof_ev)
    for (auto& e : v) {
      cudaEventDestroy(e.first);
      cudaEventDestroy(e.second);
    }
  delete h;
}

wm_status wm_load_weight(wm_handle* h, const char* ref_key, const float* data, const int64_t* shape, int ndim) {
  WM_CHECK(h && ref_key && data && shape && ndim >= 1 && ndim <= 4, "wm_load_weight: bad argument");
  const std::string key(ref_key);
  int64_t numel = 1;
  for (int i = 0; i < ndim; ++i) {
    WM_CHECK(shape[i] > 0, "wm_load_weight(%s): non-positive dimension", ref_key);
    numel *= shape[i];
  }
  DevTensor t;
  t.shape.assign(shape, shape + ndim);
  float* tmp = nullptr;
  WM_CUDA(cudaMalloc(&tmp, (size_t)numel * 4));
  cudaError_t e = cudaMemcpy(tmp, data, (size_t)numel * 4, cudaMemcpyDefault);
  if (e != cudaSuccess) {
    cudaFree(tmp);
    return fail(WM_ERR_CUDA, "wm_load_weight(%s): %s", ref_key, cudaGetErrorString(e));
  }
  const bool gemm_w = ends_with(key, "attn.qkv.weight") || ends_with(key, "attn.proj.weight") ||
                      ends_with(key, "mlp.fc1.weight") || ends_with(key, "mlp.fc2.weight") ||
                      key == "depth_embed.proj.2.fc1.weight" || key == "depth_embed.proj.2.fc2.weight" ||
                      key == "patch_embed.patch_embed.proj.weight";
  if (gemm_w) {
    const int64_t rows = shape[0], cols = numel / rows;
    const int ld = (int)((cols + 63) / 64 * 64);  // K padded to whole 64-element TMA boxes (588 -> 640, 196 -> 256)
    e = cudaMalloc(&t.b16, (size_t)rows * ld * 2);
    if (e != cudaSuccess) {
      cudaFree(tmp);
      return fail(WM_ERR_CUDA, "wm_load_weight(%s): %s", ref_key, cudaGetErrorString(e));
    }
    wm_status s = pack_bf16_rows(0, tmp, rows, (int)cols, ld, t.b16);
    cudaError_t e2 = cudaDeviceSynchronize();
    cudaFree(tmp);
    if (s != WM_OK) return s;
    if (e2 != cudaSuccess) return fail(WM_ERR_CUDA, "wm_load_weight(%s): %s", ref_key, cudaGetErrorString(e2));
    t.ld = ld;
  } else {
    t.f32 = tmp;
    t.ld = (int)shape[ndim - 1];
  }
  auto it = h->w.find(key);
  if (it != h->w.end()) {
    if (it->second.f32) cudaFree(it->second.f32);
    if (it->second.b16) cudaFree(it->second.b16);
  }
  h->w[key] = t;
  h->finalized = false;
  return WM_OK;
}

static wm_status need(wm_handle* h, const std::string& key, int64_t numel, bool bf16, const void** out) {
  auto it = h->w.find(key);
  if (it == h->w.end()) return fail(WM_ERR_STATE, "missing weight: %s", key.c_str());
  int64_t n = 1;
  for (int64_t d : it->second.shape) n *= d;
  if (n != numel) return fail(WM_ERR_STATE, "weight %s has %lld elements, expected %lld", key.c_str(), (long long)n, (long long)numel);
  const void* p = bf16 ? (const void*)it->second.b16 : (const void*)it->second.f32;
  if (!p) return fail(WM_ERR_STATE, "weight %s is stored in the wrong precision", key.c_str());
  *out = p;
  return WM_OK;
}

static wm_status resolve_block(wm_handle* h, const std::string& pre, bool qk_norm, BlockW* b) {
  const int C = kEmbed, M = kMlp;
#define NEEDF(field, name, n) WM_TRY(need(h, pre + name, n, false, reinterpret_cast<const void**>(&b->field)))
#define NEEDB(field, name, n) WM_TRY(need(h, pre + name, n, true, reinterpret_cast<const void**>(&b->field)))
  NEEDF(n1w, "norm1.weight", C);
  NEEDF(n1b, "norm1.bias", C);
  NEEDB(qkv_w, "attn.qkv.weight", 3LL * C * C);
  NEEDF(qkv_b, "attn.qkv.bias", 3 * C);
  b->qn_w = b->qn_b = b->kn_w = b->kn_b = nullptr;
  if (qk_norm) {
    NEEDF(qn_w, "attn.q_norm.weight", kHeadDim);
    NEEDF(qn_b, "attn.q_norm.bias", kHeadDim);
    NEEDF(kn_w, "attn.k_norm.weight", kHeadDim);
    NEEDF(kn_b, "attn.k_norm.bias", kHeadDim);
  }
  NEEDB(proj_w, "attn.proj.weight", (int64_t)C * C);
  NEEDF(proj_b, "attn.proj.bias", C);
  NEEDF(ls1, "ls1.gamma", C);
  NEEDF(n2w, "norm2.weight", C);
  NEEDF(n2b, "norm2.bias", C);
  NEEDB(fc1_w, "mlp.fc1.weight", (int64_t)M * C);
  NEEDF(fc1_b, "mlp.fc1.bias", M);
  NEEDB(fc2_w, "mlp.fc2.weight", (int64_t)C * M);
  NEEDF(fc2_b, "mlp.fc2.bias", C);
  NEEDF(ls2, "ls2.gamma", C);
#undef NEEDF
#undef NEEDB
  b->logit_bound = 0.f;
  if (qk_norm) {
    // LayerNorm output z has |z|_2 <= 8 (64 unit-variance entries), so |z * w + b|_2 <= 8 max|w| + |b|_2; RoPE is a
    // rotation.  1.02 covers the bf16 rounding of q and k.
    float hw[4][kHeadDim];
    const float* src[4] = {b->qn_w, b->qn_b, b->kn_w, b->kn_b};
    for (int i = 0; i < 4; ++i) WM_CUDA(cudaMemcpy(hw[i], src[i], sizeof(hw[i]), cudaMemcpyDeviceToHost));
    auto norm_bound = [&](const float* w, const float* bias) {
      float mw = 0.f, nb = 0.f;
      for (int i = 0; i < kHeadDim; ++i) mw = fmaxf(mw, fabsf(w[i])), nb += bias[i] * bias[i];
      return 8.0f * mw + sqrtf(nb);
    };
    b->logit_bound = 1.02f * norm_bound(hw[0], hw[1]) * norm_bound(hw[2], hw[3]) * 0.125f * 1.4426950408889634f;
  }
  return WM_OK;
}

wm_status wm_finalize_weights(wm_handle* h) {
  WM_CHECK(h, "wm_finalize_weights: null handle");
  h->dino.assign(h->cfg.dino_depth, BlockW{});
  h->frame.assign(h->cfg.depth, BlockW{});
  h->global.assign(h->cfg.depth, BlockW{});
  for (int i = 0; i < h->cfg.dino_depth; ++i)
    WM_TRY(resolve_block(h, "patch_embed.blocks." + std::to_string(i) + ".", false, &h->dino[i]));
  for (int i = 0; i < h->cfg.depth; ++i) {
    WM_TRY(resolve_block(h, "frame_blocks." + std::to_string(i) + ".", true, &h->frame[i]));
    WM_TRY(resolve_block(h, "global_blocks." + std::to_string(i) + ".", true, &h->global[i]));
  }
  const void* p;
  const int C = kEmbed;
  WM_TRY(need(h, "cam_token", 2 * C, false, &p));
  WM_TRY(need(h, "reg_token", 8 * C, false, &p));
  WM_TRY(need(h, "patch_embed.cls_token", C, false, &p));
  WM_TRY(need(h, "patch_embed.pos_embed", 1370LL * C, false, &p));
  WM_TRY(need(h, "patch_embed.register_tokens", 4 * C, false, &p));
  WM_TRY(need(h, "patch_embed.patch_embed.proj.weight", 588LL * C, true, &p));
  WM_TRY(need(h, "patch_embed.patch_embed.proj.bias", C, false, &p));
  WM_TRY(need(h, "patch_embed.norm.weight", C, false, &p));
  WM_TRY(need(h, "patch_embed.norm.bias", C, false, &p));
  WM_TRY(need(h, "pose_embed.0.weight", 7 * C, false, &p));
  WM_TRY(need(h, "pose_embed.0.bias", C, false, &p));
  WM_TRY(need(h, "pose_embed.2.weight", (int64_t)C * C, false, &p));
  WM_TRY(need(h, "pose_embed.2.bias", C, false, &p));
  WM_TRY(need(h, "ray_embed.0.weight", 4 * C, false, &p));
  WM_TRY(need(h, "ray_embed.0.bias", C, false, &p));
  WM_TRY(need(h, "ray_embed.2.weight", (int64_t)C * C, false, &p));
  WM_TRY(need(h, "ray_embed.2.bias", C, false, &p));
  WM_TRY(need(h, "depth_embed.proj.2.fc1.weight", 196LL * kMlp, true, &p));
  WM_TRY(need(h, "depth_embed.proj.2.fc1.bias", kMlp, false, &p));
  WM_TRY(need(h, "depth_embed.proj.2.fc2.weight", (int64_t)C * kMlp, true, &p));
  WM_TRY(need(h, "depth_embed.proj.2.fc2.bias", C, false, &p));
  WM_TRY(build_dino_table(h, 37, 37, h->w["patch_embed.pos_embed"].f32));
  h->finalized = true;
  return WM_OK;
}

wm_status wm_set_dino_pos_table(wm_handle* h, int gh, int gw, const float* pos) {
  WM_CHECK(h && pos && gh > 0 && gw > 0, "wm_set_dino_pos_table: bad argument");
  if (!h->finalized) return fail(WM_ERR_STATE, "wm_set_dino_pos_table: finalise the weights first");
  const size_t bytes = (size_t)(1 + gh * gw) * kEmbed * 4;
  float* tmp = nullptr;
  WM_CUDA(cudaMalloc(&tmp, bytes));
  cudaError_t e = cudaMemcpy(tmp, pos, bytes, cudaMemcpyDefault);
  wm_status s = e == cudaSuccess ? build_dino_table(h, gh, gw, tmp) : fail(WM_ERR_CUDA, "wm_set_dino_pos_table: %s", cudaGetErrorString(e));
  cudaFree(tmp);
  return s;
}

size_t wm_workspace_bytes(const wm_handle* h, int n_views, int H, int W, int world) {
  (void)h;
  if (n_views <= 0 || H <= 0 || W <= 0 || world <= 0 || H % kPatch || W % kPatch) return 0;
  return carve(nullptr, n_views, H / kPatch, W / kPatch, world).bytes;
}

wm_status wm_forward(wm_handle* h, wm_stream stream, const float* images, const float* depth, const float* rays,
                     const float* poses, const int cond_flags[3], int n_views, int H, int W, int first_view,
                     float* const* out, void* workspace, size_t workspace_bytes) {
  WM_CHECK(images && cond_flags && out, "wm_forward: null argument");
  WM_CHECK(H > 0 && W > 0 && H % kPatch == 0 && W % kPatch == 0, "wm_forward: H and W must be positive multiples of 14 (got %dx%d)", H, W);
  const int gh = H / kPatch, gw = W / kPatch;
  WM_TRY(check_ready(h, workspace, workspace_bytes, n_views, gh, gw));
  for (int i = 0; i < h->cfg.n_intermediate; ++i) WM_CHECK(out[i], "wm_forward: out[%d] is null", i);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const Workspace ws = carve(align_base(workspace), n_views, gh, gw, h->world);
  WM_TRY(run_embed(h, st, ws, images, depth, rays, poses, cond_flags, n_views, H, W, first_view));
  return run_stack(h, st, ws, n_views, gh, gw, out);
}

wm_status wm_embed(wm_handle* h, wm_stream stream, const float* images, const float* depth, const float* rays,
                   const float* poses, const int cond_flags[3], int n_views, int H, int W, int first_view, float* x,
                   void* workspace, size_t workspace_bytes) {
  WM_CHECK(images && cond_flags && x, "wm_embed: null argument");
  WM_CHECK(H > 0 && W > 0 && H % kPatch == 0 && W % kPatch == 0, "wm_embed: H and W must be positive multiples of 14 (got %dx%d)", H, W);
  const int gh = H / kPatch, gw = W / kPatch;
  WM_TRY(check_ready(h, workspace, workspace_bytes, n_views, gh, gw));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const Workspace ws = carve(align_base(workspace), n_views, gh, gw, h->world);
  WM_TRY(run_embed(h, st, ws, images, depth, rays, poses, cond_flags, n_views, H, W, first_view));
  WM_CUDA(cudaMemcpyAsync(x, ws.x, (size_t)n_views * (kPatchStart + gh * gw) * kEmbed * 4, cudaMemcpyDeviceToDevice, st));
  return WM_OK;
}

wm_status wm_forward_tokens(wm_handle* h, wm_stream stream, float* x, int n_views, int gh, int gw, float* const* out,
                            void* workspace, size_t workspace_bytes) {
  WM_CHECK(x && out, "wm_forward_tokens: null argument");
  WM_TRY(check_ready(h, workspace, workspace_bytes, n_views, gh, gw));
  for (int i = 0; i < h->cfg.n_intermediate; ++i) WM_CHECK(out[i], "wm_forward_tokens: out[%d] is null", i);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  Workspace ws = carve(align_base(workspace), n_views, gh, gw, h->world);
  ws.x = x;  // run in place on the caller's token buffer
  return run_stack(h, st, ws, n_views, gh, gw, out);
}

// ---------------------------------------------------------------------------------------------- profiling
wm_status wm_profile(wm_handle* h, unsigned class_mask) {
  WM_CHECK(h, "wm_profile: null handle");
  h->prof_mask = class_mask;
  for (int c = 0; c < WM_PROF_CLASSES; ++c) h->prof_used[c] = 0;
  h->launches = 0;
  return WM_OK;
}

wm_status wm_profile_read(wm_handle* h, int cls, float* total_ms, int* count) {
  WM_CHECK(h && total_ms && count && cls >= 0 && cls < WM_PROF_CLASSES, "wm_profile_read: bad argument");
  float sum = 0.f;
  for (size_t i = 0; i < h->prof_used[cls]; ++i) {
    WM_CUDA(cudaEventSynchronize(h->prof_ev[cls][i].second));
    float ms = 0.f;
    WM_CUDA(cudaEventElapsedTime(&ms, h->prof_ev[cls][i].first, h->prof_ev[cls][i].second));
    sum += ms;
  }
  *total_ms = sum;
  *count = (int)h->prof_used[cls];
  return WM_OK;
}

long long wm_launch_count(const wm_handle* h) { return h ? h->launches : 0; }

// ---------------------------------------------------------------------------------------------- NCCL
wm_status wm_nccl_unique_id(const char* nccl_lib, void* unique_id_128) {
  WM_CHECK(unique_id_128, "wm_nccl_unique_id: null output");
  NcclApi api;
  WM_TRY(nccl_open(nccl_lib, &api));
  nccl_uid id;
  const int r = api.GetUniqueId(&id);
  if (r != 0) return fail(WM_ERR_NCCL, "ncclGetUniqueId: %s", api.GetErrorString(r));
  memcpy(unique_id_128, &id, sizeof(id));
  return WM_OK;
}

wm_status wm_nccl_init(wm_handle* h, const char* nccl_lib, const void* unique_id_128, int rank, int world) {
  WM_CHECK(h && unique_id_128 && world >= 1 && rank >= 0 && rank < world, "wm_nccl_init: bad argument");
  WM_TRY(nccl_open(nccl_lib, &h->nccl));
  nccl_uid id;
  memcpy(&id, unique_id_128, sizeof(id));
  const int r = h->nccl.CommInitRank(&h->comm, world, id, rank);
  if (r != 0) return fail(WM_ERR_NCCL, "ncclCommInitRank: %s", h->nccl.GetErrorString(r));
  h->rank = rank, h->world = world;
  return WM_OK;
}

wm_status wm_nccl_allgather(wm_handle* h, wm_stream stream, void* buf, size_t bytes_per_rank) {
  WM_CHECK(h && buf && bytes_per_rank > 0, "wm_nccl_allgather: bad argument");
  if (!h->comm) return fail(WM_ERR_STATE, "wm_nccl_allgather: no communicator (call wm_nccl_init)");
  const int r = h->nccl.AllGather(static_cast<uint8_t*>(buf) + (size_t)h->rank * bytes_per_rank, buf, bytes_per_rank, 0,
                                  h->comm, static_cast<cudaStream_t>(stream));
  if (r != 0) return fail(WM_ERR_NCCL, "ncclAllGather: %s", h->nccl.GetErrorString(r));
  return WM_OK;
}

// ---------------------------------------------------------------------------------------------- peer-memory gather
wm_status wm_p2p_alloc(wm_handle* h, int rank, int world, int max_rows_per_rank, void* ipc_handle_64) {
  WM_CHECK(h && ipc_handle_64 && world >= 2 && world <= kP2pMaxWorld && rank >= 0 && rank < world && max_rows_per_rank > 0,
           "wm_p2p_alloc: bad argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
  int sms = 0;
  WM_TRY(require_sm100(&sms));
  WM_CUDA(cudaDeviceSynchronize());
  p2p_release(h);
  PeerGather& g = h->p2p;
  g.max_rows = max_rows_per_rank;
  g.buf_bytes = (size_t)world * max_rows_per_rank * 2 * kEmbed * 2;
  WM_CUDA(cudaMalloc(&g.base, kP2pHeader + 2 * g.buf_bytes));
  WM_CUDA(cudaMemset(g.base, 0, kP2pHeader));
  cudaIpcMemHandle_t mh;
  WM_CUDA(cudaIpcGetMemHandle(&mh, g.base));
  memcpy(ipc_handle_64, &mh, 64);
  h->rank = rank, h->world = world;
  return WM_OK;
}

wm_status wm_p2p_attach(wm_handle* h, const void* ipc_handles_all) {
  WM_CHECK(h && ipc_handles_all, "wm_p2p_attach: null argument");
  PeerGather& g = h->p2p;
  if (!g.base) return fail(WM_ERR_STATE, "wm_p2p_attach: call wm_p2p_alloc first");
  for (int r = 0; r < h->world; ++r) {
    if (r == h->rank) {
      g.peer[r] = g.base;
      continue;
    }
    cudaIpcMemHandle_t mh;
    memcpy(&mh, static_cast<const uint8_t*>(ipc_handles_all) + 64 * r, 64);
    void* ptr = nullptr;
    const cudaError_t e = cudaIpcOpenMemHandle(&ptr, mh, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      cudaGetLastError();
      return fail(WM_ERR_CUDA, "cudaIpcOpenMemHandle(rank %d): %s (peers must sit on one NVLink domain)", r, cudaGetErrorString(e));
    }
    g.peer[r] = static_cast<uint8_t*>(ptr);
  }
  int lo = 0, hi = 0;
  WM_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
  WM_CUDA(cudaStreamCreateWithPriority(&g.comm, cudaStreamNonBlocking, hi));
  WM_CUDA(cudaEventCreateWithFlags(&g.ev_ready, cudaEventDisableTiming));
  WM_CUDA(cudaEventCreateWithFlags(&g.ev_done, cudaEventDisableTiming));
  WM_CUDA(cudaEventCreateWithFlags(&g.ev_err, cudaEventDisableTiming));
  WM_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&g.err_host), sizeof(unsigned), cudaHostAllocDefault));
  *g.err_host = 0;
  g.err_pending = false;
  g.epoch = 0;
  g.on = true;
  return WM_OK;
}

wm_status wm_p2p_status(wm_handle* h, int* enabled, int* timed_out) {
  WM_CHECK(h && enabled && timed_out, "wm_p2p_status: null argument");
  *enabled = h->p2p.on && h->sp_mode == 1;
  *timed_out = 0;
  if (h->p2p.base) {
    unsigned e = 0;
    WM_CUDA(cudaMemcpy(&e, h->p2p.base + 64 * sizeof(unsigned), sizeof(e), cudaMemcpyDeviceToHost));
    *timed_out = (int)e;
    if (e != 0) {  // reported: clear it so that a job that recovers can go on
      WM_CUDA(cudaMemset(h->p2p.base + 64 * sizeof(unsigned), 0, sizeof(unsigned)));
      if (h->p2p.err_host) *h->p2p.err_host = 0;
      h->p2p.err_pending = false;
    }
  }
  return WM_OK;
}

wm_status wm_flash_attn_gathered(wm_stream stream, const void* q, int ldq, const void* k, const void* v, int ldkv, void* o,
                                 int ldo, int Lq, int Lkv, int heads, float scale, int variant, void* workspace,
                                 size_t workspace_bytes, int shard_rows, int self_rank, const unsigned* flags,
                                 unsigned epoch, unsigned* err) {
  AttnGather ag{shard_rows, self_rank, epoch, flags, err};
  return flash_attention(static_cast<cudaStream_t>(stream), q, ldq, k, v, ldkv, o, ldo, 1, Lq, Lkv, heads, scale, variant,
                         static_cast<float*>(workspace), workspace_bytes, &ag);
}

wm_status wm_p2p_raise_flag(wm_stream stream, unsigned* flag, unsigned epoch) {
  WM_CHECK(flag, "wm_p2p_raise_flag: null flag");
  p2p_flag_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(flag, epoch);
  WM_LAUNCH_CHECK();
  return WM_OK;
}

// ---------------------------------------------------------------------------------------------- micro-ops
wm_status wm_layernorm(wm_stream stream, const float* x, const float* w, const float* b, float eps, int rows, void* y,
                       int out_f32) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return layernorm(static_cast<cudaStream_t>(stream), x, w, b, eps, rows, rows > 0 ? rows : 1, 0, 0, 0, 0, y, out_f32, 0);
}

wm_status wm_gemm_bias_act(wm_stream stream, const void* A, int lda, const void* W, int ldw, const float* bias, int M,
                           int N, int K, void* out, int ldo, int gelu) {
  return gemm_bias_act(static_cast<cudaStream_t>(stream), A, lda, W, ldw, bias, M, N, K, out, ldo, gelu);
}

wm_status wm_gemm_qkv(wm_stream stream, const void* A, const void* W, const float* bias, const float* qn_w,
                      const float* qn_b, const float* kn_w, const float* kn_b, float qk_eps, const float* rope_cos,
                      const float* rope_sin, int tokens_per_view, int patch_start, int grid_w, float q_scale, int M, void* q,
                      void* kv) {
  return gemm_qkv(static_cast<cudaStream_t>(stream), A, kEmbed, W, bias, qn_w, qn_b, kn_w, kn_b, qk_eps, rope_cos,
                  rope_sin, tokens_per_view, patch_start, grid_w, q_scale, M, q, kv);
}

wm_status wm_gemm_qkv_guard(wm_stream stream, const void* A, const void* W, const float* bias, const float* qn_w,
                            const float* qn_b, const float* kn_w, const float* kn_b, float qk_eps, const float* rope_cos,
                            const float* rope_sin, int tokens_per_view, int patch_start, int grid_w, float q_scale, int M, void* q,
                            void* kv, unsigned* norm2) {
  WM_CHECK(norm2 && (reinterpret_cast<uintptr_t>(norm2) & 3) == 0, "wm_gemm_qkv_guard: norm2 must be a 4-byte aligned device pointer");
  return gemm_qkv(static_cast<cudaStream_t>(stream), A, kEmbed, W, bias, qn_w, qn_b, kn_w, kn_b, qk_eps, rope_cos,
                  rope_sin, tokens_per_view, patch_start, grid_w, q_scale, M, q, kv, norm2);
}

wm_status wm_flash_attn_guarded(wm_stream stream, const void* q, int ldq, const void* k, const void* v, int ldkv, void* o,
                                int ldo, int batch, int Lq, int Lkv, int heads, int bounded_variant, int general_variant,
                                void* workspace, size_t workspace_bytes, const unsigned* norm2, float limit) {
  WM_CHECK(heads == kHeads, "wm_flash_attn_guarded: the guard words are laid out for %d heads", kHeads);
  WM_CHECK(norm2 && limit > 0.f && limit <= 96.f, "wm_flash_attn_guarded: norm2 must be set and 0 < limit <= 96");
  return guarded_attention(nullptr, static_cast<cudaStream_t>(stream), 0, q, ldq, k, v, ldkv, o, ldo, batch, Lq, Lkv, bounded_variant,
                           general_variant, static_cast<float*>(workspace), workspace_bytes, norm2, limit);
}

wm_status wm_guard_words(const wm_handle* h, int slot, float* host32) {
  WM_CHECK(h && host32 && h->guard_words, "wm_guard_words: null argument");
  WM_CHECK(slot >= 0 && slot < h->cfg.dino_depth + 2 * h->cfg.depth, "wm_guard_words: slot %d out of range", slot);
  WM_CUDA(cudaDeviceSynchronize());
  WM_CUDA(cudaMemcpy(host32, h->guard_words + (size_t)slot * kGuardWords, kGuardWords * 4, cudaMemcpyDeviceToHost));
  return WM_OK;
}

wm_status wm_set_int(wm_handle* h, const char* name, int value) {
  WM_CHECK(h && name, "wm_set_int: null argument");
  const std::string n(name);
  if (n == "attn_general_variant") {
    WM_CHECK((value >= 0 && value <= 7) || (value >= 20 && value <= 23), "attn_general_variant must be in [0,7] or [20,23]");
    h->attn_general_variant = value;
  } else if (n == "attn_bounded_variant") {
    WM_CHECK((value >= 8 && value <= 19) || (value >= 24 && value <= 27), "attn_bounded_variant must be in [8,19] or [24,27]");
    h->attn_bounded_variant = value;
  } else if (n == "attn_allow_bounded") {
    h->attn_allow_bounded = value != 0;
  } else if (n == "attn_guard") {
    h->attn_guard = value != 0;
  } else if (n == "attn_balance") {
    h->attn_balance = value != 0;
  } else if (n == "gemm_split") {
    h->gemm_split = value != 0;
  } else if (n == "sp_mode") {
    WM_CHECK(value == 0 || value == 1, "sp_mode must be 0 (ncclAllGather) or 1 (peer-memory push)");
    h->sp_mode = value;
  } else {
    return fail(WM_ERR_INVALID, "wm_set_int: unknown option %s", name);
  }
  return WM_OK;
}

int wm_get_int(const wm_handle* h, const char* name) {
  if (!h || !name) return -1;
  const std::string n(name);
  if (n == "attn_general_variant") return h->attn_general_variant;
  if (n == "attn_bounded_variant") return h->attn_bounded_variant;
  if (n == "attn_allow_bounded") return h->attn_allow_bounded;
  if (n == "attn_guard") return h->attn_guard;
  if (n == "attn_balance") return h->attn_balance;
  if (n == "sp_mode") return h->sp_mode;
  if (n == "gemm_split") return h->gemm_split;
  return -1;
}

float wm_logit_bound(const wm_handle* h, int kind, int layer) {
  if (!h || !h->finalized || layer < 0) return -1.f;
  const std::vector<BlockW>& v = kind == 0 ? h->dino : kind == 1 ? h->frame : h->global;
  return layer < (int)v.size() ? v[layer].logit_bound : -1.f;
}

wm_status wm_gemm_ls_residual(wm_stream stream, const void* A, int lda, const void* W, int ldw, const float* bias,
                              const float* gamma, int M, int N, int K, float* x, int ldx) {
  return gemm_ls_residual(static_cast<cudaStream_t>(stream), A, lda, W, ldw, bias, gamma, M, N, K, x, ldx);
}

wm_status wm_gemm_ls_residual_ws(wm_stream stream, const void* A, int lda, const void* W, int ldw, const float* bias,
                                 const float* gamma, int M, int N, int K, float* x, int ldx, void* workspace,
                                 size_t workspace_bytes) {
  // workspace = kGemmFlagBytes of slice counters, zero at first use and left zero by every launch
  WM_CHECK(workspace && workspace_bytes >= wm_gemm_workspace_bytes(), "wm_gemm_ls_residual_ws: workspace too small");
  WM_CHECK((reinterpret_cast<uintptr_t>(workspace) & 3) == 0, "wm_gemm_ls_residual_ws: workspace must be 4-byte aligned");
  return gemm_ls_residual(static_cast<cudaStream_t>(stream), A, lda, W, ldw, bias, gamma, M, N, K, x, ldx,
                          static_cast<unsigned*>(workspace));
}

size_t wm_gemm_workspace_bytes(void) { return kGemmFlagBytes; }

wm_status wm_copy_cols(wm_stream stream, const float* x, long long rows, float* out, int ld_out, int col0) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return copy_cols(static_cast<cudaStream_t>(stream), x, rows, out, ld_out, col0);
}

wm_status wm_gemm_table(wm_stream stream, const void* A, int lda, const void* W, int ldw, const float* table,
                        int period, int M, int N, int K, float* out, int ldo) {
  return gemm_table(static_cast<cudaStream_t>(stream), A, lda, W, ldw, table, period, M, N, K, out, ldo);
}

wm_status wm_flash_attn(wm_stream stream, const void* q, int ldq, const void* k, const void* v, int ldkv, void* o,
                        int ldo, int batch, int Lq, int Lkv, int heads, float scale, int variant) {
  return flash_attention(static_cast<cudaStream_t>(stream), q, ldq, k, v, ldkv, o, ldo, batch, Lq, Lkv, heads, scale,
                         variant, nullptr, 0);
}

wm_status wm_flash_attn_ws(wm_stream stream, const void* q, int ldq, const void* k, const void* v, int ldkv, void* o,
                           int ldo, int batch, int Lq, int Lkv, int heads, float scale, int variant, void* workspace,
                           size_t workspace_bytes) {
  return flash_attention(static_cast<cudaStream_t>(stream), q, ldq, k, v, ldkv, o, ldo, batch, Lq, Lkv, heads, scale,
                         variant, static_cast<float*>(workspace), workspace_bytes);
}

size_t wm_attn_workspace_bytes(void) { return attn_partial_bytes(kNumSms); }

wm_status wm_umma_probe(wm_stream stream, const void* p, const void* v, float* d, int mode, unsigned v_lbo,
                        unsigned v_sbo) {
  return umma_probe(static_cast<cudaStream_t>(stream), p, v, d, mode, v_lbo, v_sbo);
}

wm_status wm_patch_gather(wm_stream stream, const float* img, int n_views, int channels, int H, int W,
                          int rows_per_view, int row_offset, int ld, int normalize, void* out) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return patch_gather(static_cast<cudaStream_t>(stream), img, n_views, channels, H, W, rows_per_view, row_offset, ld,
                      normalize, out);
}

void wm_rope_tables(int max_pos, float* cos_host, float* sin_host) {
  // rope.py:94-111: inv_freq_i = 1 / base^(2i/32) (fp32), angle = pos * inv_freq (fp32), table = cos/sin(angle)
  for (int i = 0; i < 16; ++i) {
    const float expo = (float)(2 * i) / 32.0f;
    const float inv = 1.0f / powf(100.0f, expo);
    for (int p = 0; p <= max_pos; ++p) {
      const float ang = (float)p * inv;
      cos_host[p * 16 + i] = (float)cos((double)ang);
      sin_host[p * 16 + i] = (float)sin((double)ang);
    }
  }
}

// ---------------------------------------------------------------------------------------------- prediction heads
wm_status wm_conv2d_nhwc(wm_stream stream, const void* in, int n, int H, int W, int Cin, const void* weight, int taps, int Cout,
                         const float* bias, const void* res1, int res1_bcast, const void* res2, int relu, void* out) {
  return conv2d_nhwc(static_cast<cudaStream_t>(stream), in, n, H, W, Cin, weight, taps, Cout, bias, res1, res1_bcast, res2, relu,
                     out);
}

wm_status wm_conv2d_final(wm_stream stream, const void* in, int n, int H, int W, int Cin, const void* w1, const float* b1,
                          const float* w2, const float* b2, int od, int act, float* pred, float* conf) {
  return conv2d_final(static_cast<cudaStream_t>(stream), in, n, H, W, Cin, w1, b1, w2, b2, od, act, pred, conf);
}

wm_status wm_fp16_saturation(wm_stream stream, unsigned* host_count, int reset) {
  return fp16_saturation_read(static_cast<cudaStream_t>(stream), host_count, reset);
}

wm_status wm_layernorm_rows(wm_stream stream, const float* x, const float* w, const float* b, float eps, long long rows,
                            int rows_per_group, int in_group_rows, int in_offset, int C, void* y) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return ln_rows(static_cast<cudaStream_t>(stream), x, w, b, eps, rows, rows_per_group, in_group_rows, in_offset, C, y);
}

wm_status wm_bilinear_nhwc(wm_stream stream, const void* in, int n, int Hi, int Wi, int C, int Ho, int Wo, const float* pos_u,
                           const float* pos_v, void* out) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return bilinear_nhwc(static_cast<cudaStream_t>(stream), in, n, Hi, Wi, C, Ho, Wo, pos_u, pos_v, out);
}

wm_status wm_pixel_shuffle(wm_stream stream, const void* in, int n, int Hi, int Wi, int C, int k, void* out) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return pixel_shuffle(static_cast<cudaStream_t>(stream), in, n, Hi, Wi, C, k, out);
}

wm_status wm_im2col_s2(wm_stream stream, const void* in, int n, int Hi, int Wi, int C, void* out) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return im2col_s2(static_cast<cudaStream_t>(stream), in, n, Hi, Wi, C, out);
}

wm_status wm_skinny_linear(wm_stream stream, const float* A, int lda, const float* W, const float* bias, int M, int N, int K,
                           int silu_in, int act, const float* resid, const float* gamma, float* out, int ldo) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return skinny_linear(static_cast<cudaStream_t>(stream), A, lda, W, bias, M, N, K, silu_in, act, resid, gamma, out, ldo);
}

wm_status wm_cam_layernorm(wm_stream stream, const float* x, long long in_stride, const float* w, const float* b, float eps, int M,
                           const float* mod, float* y) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return cam_layernorm(static_cast<cudaStream_t>(stream), x, in_stride, w, b, eps, M, mod, y);
}

wm_status wm_cam_attention(wm_stream stream, const float* qkv, int S, float* out) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return cam_attention(static_cast<cudaStream_t>(stream), qkv, S, out);
}

wm_status wm_cam_embed(wm_stream stream, const float* in, int in_bcast, int K, const float* W, const float* bias, int M, int N,
                       float* out) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return cam_embed(static_cast<cudaStream_t>(stream), in, in_bcast, K, W, bias, M, N, out);
}

wm_status wm_cam_predict(wm_stream stream, const float* h, int K, const float* W, const float* bias, int M, int first, float* pred,
                         float* act) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return cam_predict(static_cast<cudaStream_t>(stream), h, K, W, bias, M, first, pred, act);
}

wm_status wm_normalize_depth(wm_stream stream, const float* depth, int n_views, int hw, double eps, int min_percentile,
                             int max_percentile, float* out) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return normalize_depth(static_cast<cudaStream_t>(stream), depth, n_views, hw, eps, min_percentile, max_percentile, out);
}

wm_status wm_pose_prior(wm_stream stream, const float* extrinsics, int extr_stride, const float* intrinsics, int n_views, int H, int W,
                        double padding, float* poses, float* rays) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return pose_prior(static_cast<cudaStream_t>(stream), extrinsics, extr_stride, intrinsics, n_views, H, W, padding, poses, rays);
}

wm_status wm_im2row7(wm_stream stream, const float* images, int n, int H, int W, void* out) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return im2row7(static_cast<cudaStream_t>(stream), images, n, H, W, out);
}

wm_status wm_conv2d_project(wm_stream stream, const void* in, int n, int H, int W, int Cin, const void* w1, const float* b1,
                            const float* w2, const float* b2, int od, float* out) {
  return conv2d_project(static_cast<cudaStream_t>(stream), in, n, H, W, Cin, w1, b1, w2, b2, od, out);
}

wm_status wm_splat_cameras(wm_stream stream, const float* camera_params, int n, int H, int W, float* cam) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return splat_cameras(static_cast<cudaStream_t>(stream), camera_params, n, H, W, cam);
}

wm_status wm_splat_prepare(wm_stream stream, const float* raw, const float* images, const float* gs_depth, const float* cam, int n,
                           int H, int W, float* means, float* quats, float* scales, float* opacities, float* sh, float* weights) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return splat_prepare(static_cast<cudaStream_t>(stream), raw, images, gs_depth, cam, n, H, W, means, quats, scales, opacities, sh,
                       weights);
}

wm_status wm_prune_workspace_bytes(long long N, size_t* bytes) { return prune_workspace_bytes(N, bytes); }

wm_status wm_prune_splats(wm_stream stream, long long N, const float* means, const float* quats, const float* scales,
                          const float* opacities, const float* sh, const float* weights, double voxel_size, void* workspace,
                          size_t workspace_bytes, float* out_means, float* out_quats, float* out_scales, float* out_opacities,
                          float* out_sh, unsigned* count) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return prune_splats(static_cast<cudaStream_t>(stream), N, means, quats, scales, opacities, sh, weights, voxel_size, workspace,
                      workspace_bytes, out_means, out_quats, out_scales, out_opacities, out_sh, count);
}

wm_status wm_prune_splats_bounds(wm_stream stream, long long N, const float* means, const float* quats, const float* scales,
                                 const float* opacities, const float* sh, const float* weights, double voxel_size, void* workspace,
                                 size_t workspace_bytes, float* out_means, float* out_quats, float* out_scales, float* out_opacities,
                                 float* out_sh, unsigned* count, const long long* bounds) {
  int sms;
  WM_TRY(require_sm100(&sms));
  WM_CHECK(bounds != nullptr, "wm_prune_splats_bounds: null bounds");
  return prune_splats(static_cast<cudaStream_t>(stream), N, means, quats, scales, opacities, sh, weights, voxel_size, workspace,
                      workspace_bytes, out_means, out_quats, out_scales, out_opacities, out_sh, count, bounds);
}

wm_status wm_voxel_bounds(wm_stream stream, long long N, const float* means, double voxel_size, long long* bounds) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return voxel_bounds(static_cast<cudaStream_t>(stream), N, means, voxel_size, bounds);
}

wm_status wm_voxel_keys(wm_stream stream, long long N, const float* means, double voxel_size, const long long* bounds, long long* keys) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return voxel_keys(static_cast<cudaStream_t>(stream), N, means, voxel_size, bounds, keys);
}

wm_status wm_conf_filter_workspace_bytes(long long N, size_t* bytes) { return conf_filter_workspace_bytes(N, bytes); }

wm_status wm_conf_filter_splats(wm_stream stream, long long N, long long K, const float* conf, const float* means, const float* quats,
                                const float* scales, const float* opacities, const float* sh, const float* weights, void* workspace,
                                size_t workspace_bytes, float* out_means, float* out_quats, float* out_scales, float* out_opacities,
                                float* out_sh, float* out_weights) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return conf_filter_splats(static_cast<cudaStream_t>(stream), N, K, conf, means, quats, scales, opacities, sh, weights, workspace,
                            workspace_bytes, out_means, out_quats, out_scales, out_opacities, out_sh, out_weights);
}

wm_status wm_raster_project(wm_stream stream, const float* means, const float* quats, const float* scales, const float* opacities,
                            const float* viewmats, const float* Ks, int C, long long N, int W, int H, float near_plane,
                            float far_plane, float eps2d, float* means2d, float* conics, float* depths, int* radii, unsigned* tiles,
                            unsigned* cum, void* scan_tmp, size_t scan_tmp_bytes) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return raster_project(static_cast<cudaStream_t>(stream), means, quats, scales, opacities, viewmats, Ks, C, N, W, H, near_plane,
                        far_plane, eps2d, means2d, conics, depths, radii, tiles, cum, scan_tmp, scan_tmp_bytes);
}

wm_status wm_raster_workspace_bytes(long long elems, long long n_isects, size_t* scan_bytes, size_t* render_bytes) {
  return raster_workspace_bytes(elems, n_isects, scan_bytes, render_bytes);
}

wm_status wm_raster_render(wm_stream stream, const float* means2d, const float* conics, const float* depths, const int* radii,
                           const unsigned* cum, const float* opacities, const float* sh, int C, long long N, int W, int H,
                           long long n_isects, void* workspace, size_t workspace_bytes, unsigned* offsets, float* out_rgb,
                           float* out_depth, float* out_alpha) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return raster_render(static_cast<cudaStream_t>(stream), means2d, conics, depths, radii, cum, opacities, sh, C, N, W, H, n_isects,
                       workspace, workspace_bytes, offsets, out_rgb, out_depth, out_alpha);
}

wm_status wm_load_image_workspace_bytes(int H, int W, int new_w, int new_h, size_t* bytes) {
  return load_image_workspace_bytes(H, W, new_w, new_h, bytes);
}

wm_status wm_load_image(wm_stream stream, const unsigned char* rgb, int H, int W, int new_w, int new_h, int crop_top, int crop_h,
                        int out_H, int out_W, int pad_top, int pad_left, float* out, void* workspace, size_t workspace_bytes) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return load_image(static_cast<cudaStream_t>(stream), rgb, H, W, new_w, new_h, crop_top, crop_h, out_H, out_W, pad_top, pad_left, out,
                    workspace, workspace_bytes);
}

size_t wm_load_image_static_bytes(void) { return load_image_static_capacity(); }

wm_status wm_load_image_static(wm_stream stream, const unsigned char* rgb, int H, int W, int new_w, int new_h, int crop_top, int crop_h,
                               int out_H, int out_W, int pad_top, int pad_left, float* out) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return load_image_static(static_cast<cudaStream_t>(stream), rgb, H, W, new_w, new_h, crop_top, crop_h, out_H, out_W, pad_top,
                           pad_left, out);
}

wm_status wm_rope2d(wm_stream stream, const float* tokens, const long long* positions, int batch, int heads, int L, int max_pos,
                    const float* rope_cos, const float* rope_sin, float* out) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return rope2d(static_cast<cudaStream_t>(stream), tokens, positions, batch, heads, L, max_pos, rope_cos, rope_sin, out);
}

wm_status wm_pack_bf16(wm_stream stream, const float* src, long long rows, int cols, int ld, void* dst) {
  int sms;
  WM_TRY(require_sm100(&sms));
  return pack_bf16_rows(static_cast<cudaStream_t>(stream), src, rows, cols, ld, dst);
}

}  // extern "C"
