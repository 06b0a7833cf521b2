/* libwm_b200 -- C ABI of the B200-native WorldMirror aggregator hot path.
 *
 * The reference (zubair-irshad/HunyuanWorld-Mirror) has no FFI: the path sits behind plain nn.Module calls.
 * Each entry point below names the reference call it stands in for (paths relative to the reference root);
 * hunyuanworld-mirror_b200/aggregator.py binds them with ctypes and mirrors the module interface
 * (VisualGeometryTransformer.forward, src/models/models/visual_transformer.py:250-341).  INTEGRATION.md shows
 * the stub a reference maintainer would add.
 *
 * Conventions
 *  - Plain pointers and sizes only.  Device pointers unless a parameter says "host".  bf16 buffers are
 *    passed as void* (uint16 storage).  All device buffers must be 16-byte aligned.
 *  - Every call is stream-ordered on `stream` (a cudaStream_t), performs no host synchronisation and no
 *    allocation (wm_load_weight / wm_finalize_weights / wm_set_dino_pos_table excepted), and is CUDA-graph
 *    capturable.  One in-flight forward per handle.
 *  - Return value: WM_OK or an error code; wm_last_error() gives the message of the calling thread's last
 *    failure.  No C++ exception crosses this boundary.  There is no CPU path: on a device that is not
 *    sm_100 every compute entry point returns WM_ERR_ARCH.
 */
#ifndef WM_B200_H
#define WM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int wm_status;
enum {
  WM_OK = 0,
  WM_ERR_INVALID = 1, /* bad shape / pointer / argument */
  WM_ERR_ARCH = 2,    /* device is not sm_100 (B200) or no CUDA device */
  WM_ERR_CUDA = 3,    /* CUDA runtime / driver error */
  WM_ERR_NCCL = 4,    /* NCCL error */
  WM_ERR_STATE = 5    /* weights missing / not finalised / workspace too small */
};

typedef struct wm_handle wm_handle;
typedef void* wm_stream; /* cudaStream_t */

typedef struct wm_config {
  int depth;           /* aggregator depth: frame+global block pairs (reference: 24, visual_transformer.py:52) */
  int dino_depth;      /* DINOv2 ViT-L blocks (reference: 24, vision_transformer.py:364) */
  int intermediate[4]; /* layers whose outputs are returned (reference: 4,11,17,23, visual_transformer.py:69) */
  int n_intermediate;  /* <= 4 */
} wm_config;

int wm_version(void);
const char* wm_last_error(void);

/* ---- lifetime and weights ------------------------------------------------------------------------------ */
wm_status wm_create(const wm_config* cfg, wm_handle** out);
void wm_destroy(wm_handle* h);

/* One tensor of the reference state_dict under its reference key, fp32, host or device
 * memory.  GEMM weights are packed to bf16 on the device; vectors stay fp32. */
wm_status wm_load_weight(wm_handle* h, const char* ref_key, const float* data, const int64_t* shape, int ndim);
/* Checks that every tensor of the schema arrived and resolves the per-block tables. */
wm_status wm_finalize_weights(wm_handle* h);
/* DINOv2 positional table for a (gh, gw) patch grid: pos [1 + gh*gw, 1024] fp32 host or device, i.e. the result
 * of interpolate_pos_encoding (vision_transformer.py:175-207).  Not needed for the native 37 x 37 grid. */
wm_status wm_set_dino_pos_table(wm_handle* h, int gh, int gw, const float* pos);

/* ---- the hot path ---------------------------------------------------------------------------------------- */
size_t wm_workspace_bytes(const wm_handle* h, int n_views, int H, int W, int world);

/* VisualGeometryTransformer.forward (visual_transformer.py:250-341) for this rank's `n_views` views
 * (global view indices first_view .. first_view + n_views - 1 of world * n_views).
 *   images [n_views,3,H,W] fp32 in [0,1]; depth [n_views,H,W] | NULL; rays [n_views,4] | NULL;
 *   poses [n_views,7] | NULL; cond_flags = {pose, depth, rays} as in the reference;
 *   out[i] [n_views, P, 2048] fp32, i < n_intermediate, P = 7 + (H/14)(W/14):
 *   [..., :1024] = frame-block output, [..., 1024:] = global-block output. */
wm_status wm_forward(wm_handle* h, wm_stream stream, const float* images, const float* depth, const float* rays,
                     const float* poses, const int cond_flags[3], int n_views, int H, int W, int first_view,
                     float* const* out, void* workspace, size_t workspace_bytes);

/* The alternating-attention stack alone (visual_transformer.py:309-339) on assembled tokens
 * x [n_views, P, 1024] fp32 (updated in place); gh x gw is the patch grid (P = 7 + gh*gw). */
wm_status wm_forward_tokens(wm_handle* h, wm_stream stream, float* x, int n_views, int gh, int gw, float* const* out,
                            void* workspace, size_t workspace_bytes);

/* DINOv2 front + token assembly only (visual_transformer.py:275-295): x [n_views, P, 1024] fp32 out. */
wm_status wm_embed(wm_handle* h, wm_stream stream, const float* images, const float* depth, const float* rays,
                   const float* poses, const int cond_flags[3], int n_views, int H, int W, int first_view, float* x,
                   void* workspace, size_t workspace_bytes);

/* ---- measurement hooks (bench.py): CUDA-event timing of kernel classes on the forward's own stream -------- */
enum {
  WM_PROF_LN = 0,          /* layernorm_kernel */
  WM_PROF_QKV = 1,         /* gemm_kernel<EPI_QKV> */
  WM_PROF_ATTN_LOCAL = 2,  /* attn_kernel, per-view (frame blocks and DINOv2) */
  WM_PROF_ATTN_GLOBAL = 3, /* attn_kernel over all views' tokens */
  WM_PROF_PROJ = 4,        /* gemm_kernel<EPI_RESID>, K = 1024 */
  WM_PROF_FC1 = 5,         /* gemm_kernel<EPI_GELU> */
  WM_PROF_FC2 = 6,         /* gemm_kernel<EPI_RESID>, K = 4096 */
  WM_PROF_EMBED = 7,       /* patch gather / patch-embed GEMM / prior tokens / assembly */
  WM_PROF_ALLGATHER = 8,   /* ncclAllGather of K|V */
  WM_PROF_CLASSES = 9
};
/* Select the classes to time (bit i = class i; 0 switches timing off) and reset the records and launch counter. */
wm_status wm_profile(wm_handle* h, unsigned class_mask);
/* Synchronises on the recorded events; total_ms = sum over the launches of class `cls` since wm_profile(). */
wm_status wm_profile_read(wm_handle* h, int cls, float* total_ms, int* count);
/* Kernel (and collective) launches issued by forwards since the last wm_profile() call. */
long long wm_launch_count(const wm_handle* h);

/* Tuning knobs: "attn_general_variant" (0-7, 20-23: attn_pipe.cu), "attn_bounded_variant" (8-12: attn_split.cu, 13-19: attn_pipe.cu),
 * "attn_allow_bounded" (0/1), "attn_balance" (0/1: even split of attention work with partial merge), "sp_mode",
 * "gemm_split" (0/1: K-split of the residual GEMMs' tail tiles), "attn_guard" (0/1: blocks whose weights do not bound the
 * logits -- DINOv2, or QK-norm gains past the bound -- let the device choose between the bounded-logit and the running-max
 * attention kernel from the q / k row lengths of the input at hand, wm_gemm_qkv_guard / wm_flash_attn_guarded; 0: such
 * blocks always take the running-max kernel).
 * wm_get_int returns the current value, or -1 for an unknown name. */
wm_status wm_set_int(wm_handle* h, const char* name, int value);
int wm_get_int(const wm_handle* h, const char* name);
/* Bound of |q . k| * scale * log2(e) implied by the QK-LayerNorm weights of block `layer` of kind 0 = DINOv2 (always 0),
 * 1 = frame, 2 = global; the forward uses the bounded-logit attention when 0 < bound <= 80.  -1: invalid argument. */
float wm_logit_bound(const wm_handle* h, int kind, int layer);
/* What the device-side guard of the LAST forward saw for one block (synchronises the device): host32 [32] = per head
 * max |q row|^2 (q scaled by scale * log2 e), then max |k row|^2; all zero for a block that did not use the guard (its
 * weights bound the logits, "attn_guard" 0, or keys gathered from peers).  slot: DINOv2 block i -> i; frame block i ->
 * dino_depth + 2 i; global block i -> dino_depth + 2 i + 1.  The bounded-logit kernel ran iff max_h sqrt(q2_h k2_h) <= 80. */
wm_status wm_guard_words(const wm_handle* h, int slot, float* host32);

/* ---- sequence parallelism: K|V all-gather of the global blocks through NCCL --------- */
/* `nccl_lib` = path of libnccl.so.2 (the one torch loaded).  unique_id: 128 bytes. */
wm_status wm_nccl_unique_id(const char* nccl_lib, void* unique_id_128);
wm_status wm_nccl_init(wm_handle* h, const char* nccl_lib, const void* unique_id_128, int rank, int world);
wm_status wm_nccl_allgather(wm_handle* h, wm_stream stream, void* buf, size_t bytes_per_rank); /* in place, test hook */

/* Peer-memory variant of the same exchange (default once attached; wm_set_int "sp_mode" 0 selects ncclAllGather):
 * after its QKV GEMM every rank pushes its K|V rows into each peer's gather buffer with the copy engines on a side
 * stream and raises an arrival word there; the attention kernel starts on the rank's own keys and waits per shard,
 * so the transfer overlaps the math.  Two buffers alternate per global block (no acknowledgement round trip).
 * wm_p2p_alloc: allocate this rank's block for up to max_rows_per_rank tokens per rank and return its
 * cudaIpcMemHandle_t (64 bytes); exchange the handles (rank order) and hand all of them to wm_p2p_attach.
 * Every rank must run the same sequence of forwards.  A shard that fails to arrive within 4 s sets an error word and
 * lets the launch finish on stale keys; the word is copied to pinned host memory behind every forward and the NEXT
 * wm_forward / wm_forward_tokens returns WM_ERR_STATE once (and clears it) -- nothing blocks.  wm_p2p_status reads the
 * word synchronously: *timed_out != 0 reports (and clears) it. */
wm_status wm_p2p_alloc(wm_handle* h, int rank, int world, int max_rows_per_rank, void* ipc_handle_64);
wm_status wm_p2p_attach(wm_handle* h, const void* ipc_handles_all /* world x 64 bytes */);
wm_status wm_p2p_status(wm_handle* h, int* enabled, int* timed_out);

/* ---- micro-ops (unit parity + microbenchmarks); C = 1024, head_dim = 64 ----------------------------------- */
/* nn.LayerNorm(1024) (block.py:44,61): x [rows,1024] fp32 -> y [rows,1024] bf16 (or fp32 when out_f32). */
wm_status wm_layernorm(wm_stream stream, const float* x, const float* w, const float* b, float eps, int rows, void* y,
                       int out_f32);
/* nn.Linear (+ nn.GELU when gelu != 0; mlp.py:30-31): A [M,K] bf16, W [N,K] bf16, bias [N] fp32 -> out [M,N] bf16. */
wm_status wm_gemm_bias_act(wm_stream stream, const void* A, int lda, const void* W, int ldw, const float* bias, int M,
                           int N, int K, void* out, int ldo, int gelu);
/* Attention.qkv + q_norm/k_norm + rope (attention.py:50-56): A [M,1024] bf16, W [3072,1024] bf16 ->
 * q [M,1024] bf16, kv [M,2048] bf16 (k | v).  Norm / rope pointers may be NULL (DINOv2 blocks).
 * rope_cos/sin: [max_pos+1,16] fp32 from wm_rope_tables.  Token t sits at position
 * (0,0) if t % tokens_per_view < patch_start else (1 + i / grid_w, 1 + i % grid_w), i = t % tokens_per_view - patch_start.
 * q_scale multiplies q after norm and rope (1.0: none; scale * log2(e) for the bounded-logit attention variants). */
wm_status wm_gemm_qkv(wm_stream stream, const void* A, const void* W, const float* bias, const float* qn_w,
                      const float* qn_b, const float* kn_w, const float* kn_b, float qk_eps, const float* rope_cos,
                      const float* rope_sin, int tokens_per_view, int patch_start, int grid_w, float q_scale, int M, void* q,
                      void* kv);
/* wm_gemm_qkv that also records, per head, the largest squared row length of q (after q_scale) and of k:
 * norm2 [32] device words = float bit patterns, [h] for q, [16 + h] for k, accumulated with atomicMax -- ZERO them before
 * the call.  |q . k| <= |q| |k| then bounds this input's logits without a host read (attention.py:50-59). */
wm_status wm_gemm_qkv_guard(wm_stream stream, const void* A, const void* W, const float* bias, const float* qn_w,
                            const float* qn_b, const float* kn_w, const float* kn_b, float qk_eps, const float* rope_cos,
                            const float* rope_sin, int tokens_per_view, int patch_start, int grid_w, float q_scale, int M,
                            void* q, void* kv, unsigned* norm2);
/* Linear + LayerScale + residual (block.py:91-92): x [M,N] fp32 += gamma * (A W^T + bias), the addition done by a
 * TMA reduce-add (each element is added exactly once: deterministic); gamma NULL: x = A W^T + bias. */
wm_status wm_gemm_ls_residual(wm_stream stream, const void* A, int lda, const void* W, int ldw, const float* bias,
                              const float* gamma, int M, int N, int K, float* x, int ldx);
/* The same with wm_gemm_workspace_bytes() (4 KB) of slice counters, ZERO at first use (the kernel leaves them zero):
 * tiles beyond the last full wave of CTA pairs are cut along K and the slices add their shares to x in slice order
 * through the counters (one fixed order of fp32 additions: bit-deterministic; no partial-sum buffer). */
wm_status wm_gemm_ls_residual_ws(wm_stream stream, const void* A, int lda, const void* W, int ldw, const float* bias,
                                 const float* gamma, int M, int N, int K, float* x, int ldx, void* workspace,
                                 size_t workspace_bytes);
size_t wm_gemm_workspace_bytes(void);
/* out[:, col0 : col0+1024] = x [rows,1024] fp32: one half of `cat([frame, global], -1)` (visual_transformer.py:337-339). */
wm_status wm_copy_cols(wm_stream stream, const float* x, long long rows, float* out, int ld_out, int col0);
/* out [M,N] fp32 = A W^T + table[row % period] (DINOv2 patch embed + pos_embed, vision_transformer.py:209-221). */
wm_status wm_gemm_table(wm_stream stream, const void* A, int lda, const void* W, int ldw, const float* table,
                        int period, int M, int N, int K, float* out, int ldo);
/* F.scaled_dot_product_attention (attention.py:59), no mask: q [batch*Lq, ldq], k,v [batch*Lkv, ldkv],
 * o [batch*Lq, ldo] bf16, head h in columns [64h, 64h+64).  variant: -1 default; 0-3 one thread per query row
 * (attn.cu) with 0/8/12/16 of every 32 exponentials on the FMA pipe; 4-7 two threads per query row (attn_split.cu)
 * with 0/8/10/12; 8-12 bounded-logit mode of attn_split.cu (0/8/12/14/16): q must already be multiplied by
 * scale * log2(e) (wm_gemm_qkv q_scale), |q . k| <= 80, `scale` is ignored and P = 2^(q . k) without a running max. */
wm_status wm_flash_attn(wm_stream stream, const void* q, int ldq, const void* k, const void* v, int ldkv, void* o,
                        int ldo, int batch, int Lq, int Lkv, int heads, float scale, int variant);
/* As wm_flash_attn with a scratch buffer of wm_attn_workspace_bytes() (device memory, 16-byte aligned): variants 4-12
 * then deal the (work item, key tile) steps out evenly over the SMs, split items leave fp32 partials in the
 * workspace and a merge kernel adds them in a fixed order (deterministic).  This is what the forward pass uses. */
wm_status wm_flash_attn_ws(wm_stream stream, const void* q, int ldq, const void* k, const void* v, int ldkv, void* o,
                           int ldo, int batch, int Lq, int Lkv, int heads, float scale, int variant, void* workspace,
                           size_t workspace_bytes);
size_t wm_attn_workspace_bytes(void);
/* F.scaled_dot_product_attention (attention.py:59) with the kernel chosen ON THE DEVICE: q must carry scale * log2(e)
 * (wm_gemm_qkv_guard with q_scale), norm2 are that call's words.  Two launches are enqueued: the bounded-logit kernel
 * (bounded_variant 8-12) runs iff max_h |q|_h |k|_h <= limit (80 in the forward), otherwise it returns at once and the
 * running-max kernel (general_variant 4-7) runs instead; a NaN length selects the running-max kernel.  16 heads. */
wm_status wm_flash_attn_guarded(wm_stream stream, const void* q, int ldq, const void* k, const void* v, int ldkv, void* o,
                                int ldo, int batch, int Lq, int Lkv, int heads, int bounded_variant, int general_variant,
                                void* workspace, size_t workspace_bytes, const unsigned* norm2, float limit);
/* wm_flash_attn_ws (batch 1, variants 4-12) over a key buffer of Lkv / shard_rows shards that may still be filling:
 * keys are scanned starting with shard `self_rank`; shard r is read only once (int)(flags[r] - epoch) >= 0
 * (flags: device words written by wm_p2p_raise_flag or a peer's push; *err is set to 1 on a 4 s time-out). */
wm_status wm_flash_attn_gathered(wm_stream stream, const void* q, int ldq, const void* k, const void* v, int ldkv, void* o,
                                 int ldo, int Lq, int Lkv, int heads, float scale, int variant, void* workspace,
                                 size_t workspace_bytes, int shard_rows, int self_rank, const unsigned* flags,
                                 unsigned epoch, unsigned* err);
wm_status wm_p2p_raise_flag(wm_stream stream, unsigned* flag, unsigned epoch);
/* Single-tile probe: d [128,64] fp32 = p [128,128] bf16 x v [128,64] bf16; mode 0 = P in smem, 1 = P in TMEM. */
wm_status wm_umma_probe(wm_stream stream, const void* p, const void* v, float* d, int mode, unsigned v_lbo,
                        unsigned v_sbo);
/* Image normalise + 14x14 patch gather / depth pixel-unshuffle (visual_transformer.py:276-277, patch_embed.py:63-108). */
wm_status wm_patch_gather(wm_stream stream, const float* img, int n_views, int channels, int H, int W,
                          int rows_per_view, int row_offset, int ld, int normalize, void* out);
/* cos / sin of pos * 100^(-2i/32), i < 16 (rope.py:80-111): host arrays [(max_pos+1)*16]. */
void wm_rope_tables(int max_pos, float* cos_host, float* sin_host);
/* RotaryPositionEmbedding2D.forward (rope.py:148-181) as an operator of its own: tokens [batch, heads, L, 64] fp32,
 * positions [batch, L, 2] int64 (y, x) with arbitrary values in [0, max_pos], rope_cos/sin [max_pos+1, 16] from
 * wm_rope_tables -> out (same shape, fp32, not in place).  Inside wm_forward the rotation is fused into wm_gemm_qkv. */
wm_status wm_rope2d(wm_stream stream, const float* tokens, const long long* positions, int batch, int heads, int L, int max_pos,
                    const float* rope_cos, const float* rope_sin, float* out);
/* fp32 [rows, cols] -> bf16 [rows, ld] with zero tail columns. */
wm_status wm_pack_bf16(wm_stream stream, const float* src, long long rows, int cols, int ld, void* dst);

/* ---- prediction heads: DPTHead (src/models/heads/dense_head.py) ------------ */
/* The reference runs its heads in fp32 (autocast disabled, worldmirror.py:146).  Here activation maps are NHWC fp16
 * [n, H, W, C] and convolution weights fp16 [Cout, taps * Cin] with taps ordered (ky, kx, ci): a 10-bit mantissa, like
 * TF32, at twice the tensor rate; accumulation and all epilogue arithmetic are fp32; biases, positional tables and the
 * final predictions are fp32 (bf16 maps measured 1.5e-2 relative error on pts3d, above the stated 1e-2; 10 bits: 1.2e-3).
 * nn.Conv2d(Cin, Cout, 1 | 3, stride 1, padding 0 | 1) as an implicit GEMM on tcgen05 with its neighbours fused:
 * out = [relu](conv(in) + bias + res1 + res2); bias (fp32) / res1 / res2 (fp16 maps) may be NULL; res1_bcast: res1 is the
 * fp32 table [H, W, Cout] (the positional table of _apply_pos_embed, dense_head.py:262-271).  Cin % 64 == 0; Cout 128
 * or a multiple of 256.
 * ConvTranspose2d(k == stride) is this with taps = 1 and Cout = k*k*C followed by wm_pixel_shuffle; the stride-2
 * convolution is wm_im2col_s2 followed by this with taps = 1.
 * Fuses: projects[i] + pos embed (:212-215), scratch.layer*_rn (:278-281), ResidualConvUnit (:436-457, whose in-place
 * ReLU makes the skip relu(x)), FeatureFusionBlock.out_conv and skip_add (:518-534), scratch.output_conv1 (:293). */
wm_status wm_conv2d_nhwc(wm_stream stream, const void* in, int n, int H, int W, int Cin, const void* weight, int taps, int Cout,
                         const float* bias, const void* res1, int res1_bcast, const void* res2, int relu, void* out);
/* scratch.output_conv2 + activate_head (dense_head.py:85-89, 297-344): conv3x3(Cin -> 32) + ReLU + conv1x1(32 -> od) +
 * activations; act 0 = exp (depth), 1 = inv_log (points), 2 = norm (normals); confidence = 1 + exp.
 * in, w1 [32, 9*Cin] fp16; b1 [32], w2 [od, 32], b2 [od] fp32; pred [n, H, W, od-1], conf [n, H, W] fp32; od = 2 or 4. */
wm_status wm_conv2d_final(wm_stream stream, const void* in, int n, int H, int W, int Cin, const void* w1, const float* b1,
                          const float* w2, const float* b2, int od, int act, float* pred, float* conf);
/* The DPT heads keep their maps in fp16: every convolution epilogue counts the (pixel, 32-channel
 * group) entries whose fp32 value left the fp16 range before it was clamped.  Copies the count since the last reset to
 * `host_count` (pinned host memory) on `stream` and, with reset != 0, clears it behind the copy.  A non-zero count means
 * the predictions of that forward are not trustworthy (reference: fp32 / bf16-autocast maps, dense_head.py:164-251). */
wm_status wm_fp16_saturation(wm_stream stream, unsigned* host_count, int reset);
/* nn.LayerNorm(C) of gathered rows (dense_head.py:203-211): output row g * rows_per_group + i = LN(input row
 * g * in_group_rows + in_offset + i), fp32 in, fp16 out; C = 1024 or 2048. */
wm_status wm_layernorm_rows(wm_stream stream, const float* x, const float* w, const float* b, float eps, long long rows,
                            int rows_per_group, int in_group_rows, int in_offset, int C, void* y);
/* F.interpolate(bilinear, align_corners=True) (dense_head.py:216-226, 529) on NHWC fp16 (fp32 arithmetic), optionally followed by the
 * separable UV positional embedding: + pos_u[x][c] for c < C/2, + pos_v[y][c - C/2] otherwise (fp32 tables). */
wm_status wm_bilinear_nhwc(wm_stream stream, const void* in, int n, int Hi, int Wi, int C, int Ho, int Wo, const float* pos_u,
                           const float* pos_v, void* out);
/* Scatter of nn.ConvTranspose2d(C, C, k, stride k) (dense_head.py:57-64): [n*Hi*Wi, k*k*C] -> [n, Hi*k, Wi*k, C]. */
wm_status wm_pixel_shuffle(wm_stream stream, const void* in, int n, int Hi, int Wi, int C, int k, void* out);
/* im2col of nn.Conv2d(C, C, 3, stride 2, padding 1) (dense_head.py:66-68): -> [n*Ho*Wo, 9*C], Ho = (Hi-1)/2 + 1. */
wm_status wm_im2col_s2(wm_stream stream, const void* in, int n, int Hi, int Wi, int C, void* out);

/* ---- camera head: CameraHead (src/models/heads/camera_head.py:58-104), fp32 ---- */
/* nn.Linear on M = views rows (HBM-bound: streams W [N, K] once): out = epi(f(A) W^T + bias); silu_in: f = SiLU
 * (adapt_norm_gen, :49); act 1: exact-erf GELU (Mlp.fc1); resid/gamma non-NULL: out = resid + gamma * (.)
 * (LayerScale + residual of Block, block.py:91-92; out may alias resid).  K % 128 == 0. */
wm_status wm_skinny_linear(wm_stream stream, const float* A, int lda, const float* W, const float* bias, int M, int N, int K,
                           int silu_in, int act, const float* resid, const float* gamma, float* out, int ldo);
/* nn.LayerNorm(2048) fp32 of rows `in_stride` floats apart (the camera token of each view, :72-73); w/b NULL: no affine
 * (adapt_norm, :52); mod [M, 3*2048] = [shift | scale | gate] non-NULL: y = gate * (LN(x) * (1 + scale) + shift) + x (:81-83). */
wm_status wm_cam_layernorm(wm_stream stream, const float* x, long long in_stride, const float* w, const float* b, float eps, int M,
                           const float* mod, float* y);
/* Attention.forward over the S camera tokens, 16 heads of 128, no QK-norm / RoPE: qkv [S, 6144] -> out [S, 2048]. */
wm_status wm_cam_attention(wm_stream stream, const float* qkv, int S, float* out);
/* param_embed = nn.Linear(9, 2048) (:47, 77-80); in_bcast: `in` is the shared init_token [K]. */
wm_status wm_cam_embed(wm_stream stream, const float* in, int in_bcast, int K, const float* W, const float* bias, int M, int N,
                       float* out);
/* param_predictor.fc2 (1024 -> 9) + running sum + activations (:88-99, 106-126): pred [M, 9] is updated in place (set
 * when first != 0), act [M, 9] = [t(3), quat(4), relu(fov)(2)]. */
wm_status wm_cam_predict(wm_stream stream, const float* h, int K, const float* W, const float* bias, int M, int first, float* pred,
                         float* act);

/* ---- prior preparation ------------------------------------------------------- */
/* normalize_depth (src/models/utils/priors.py:107-167) for all views in one launch, no host round trip: per view
 * nan_to_num, percentiles (torch.quantile semantics, exact radix select) over the positive values, rescale, clamp.
 * depth, out: [n_views, hw] fp32, distinct buffers. */
wm_status wm_normalize_depth(wm_stream stream, const float* depth, int n_views, int hw, double eps, int min_percentile,
                             int max_percentile, float* out);
/* The pose and ray priors of one scene (worldmirror.py:228-249): poses [n_views, 7] = extrinsics_to_vector(
 * normalize_poses(extrinsics)) = [t in the unit cube (5 % / 95 % percentile box, padding), quaternion xyzw]
 * (utils/priors.py:4-104, utils/camera_utils.py:25-35, utils/rotation.py:41-97); rays [n_views, 4] =
 * [fx/W, fy/H, cx/W, cy/H].  extrinsics: rows of a [3,4] (or the top of a [4,4]) matrix per view, `extr_stride` floats
 * apart (12 or 16); intrinsics [n_views, 3, 3] and rays may both be NULL.  At most 1024 views. */
wm_status wm_pose_prior(wm_stream stream, const float* extrinsics, int extr_stride, const float* intrinsics, int n_views, int H, int W,
                        double padding, float* poses, float* rays);

/* ---- Gaussian path: what WorldMirror.forward keeps of it is `splats` ------------- */
/* (the rasterised images are discarded, src/models/models/rasterization.py:239-246).  wm_conv2d_nhwc additionally takes
 * taps = 7 (a 7x1 vertical convolution, padding 3) and relu = 2 (ReLU on conv + bias BEFORE res1 / res2 are added): with
 * wm_im2row7 that is DPTHead's `input_merger` = Conv2d(3, 128, 7, 1, 3) + ReLU added to the fused map
 * (heads/dense_head.py:91-94, 239-242). */
/* images NCHW fp32 [n, 3, H, W] -> NHWC fp16 [n, H, W, 64]: channel kx * 3 + c = image[c][y][x + kx - 3] (zero outside),
 * channels 21..63 zero.  The matching weight is fp16 [128, 7 (ky), 64]. */
wm_status wm_im2row7(wm_stream stream, const float* images, int n, int H, int W, void* out);
/* gs_renderer.gs_head (rasterization.py:149-153): conv3x3(Cin -> 256) [+ b1] + ReLU + conv1x1(256 -> od) + b2 in one
 * kernel, the hidden map stays in tensor memory.  in, w1 [256, 9*Cin] (tap-major) fp16; w2 [od, 256], b2 fp32; out [n, H, W, od] fp32; od <= 16. */
wm_status wm_conv2d_project(wm_stream stream, const void* in, int n, int H, int W, int Cin, const void* w1, const float* b1,
                            const float* w2, const float* b2, int od, float* out);
/* camera_params [n, 9] -> cam [n, 16] = camera-to-world R (9, row major), t (3), fx, fy, cx, cy: vector_to_camera_matrices
 * + closed_form_inverse_se3 as prepare_splats uses them (rasterization.py:459-468; utils/camera_utils.py:46-75). */
wm_status wm_splat_cameras(wm_stream stream, const float* camera_params, int n, int H, int W, float* cam);
/* prepare_splats, position_from = "gsdepth+predcamera", sh_degree 0, no offsets (rasterization.py:389-498; utils/act_gs.py):
 * raw [n, H, W, 12] = quats 4 | scales 3 | opacity 1 | residual SH 3 | weight 1; images NCHW [n, 3, H, W]; gs_depth
 * [n, H, W]; cam from wm_splat_cameras.  Outputs over N = n*H*W splats: means [N,3], quats [N,4], scales [N,3] (clamped
 * at 0.3), opacities [N], sh [N,3] (= RGB2SH(image) + residual), weights [N]. */
wm_status wm_splat_prepare(wm_stream stream, const float* raw, const float* images, const float* gs_depth, const float* cam, int n,
                           int H, int W, float* means, float* quats, float* scales, float* opacities, float* sh, float* weights);
/* prune_gs (rasterization.py:301-387) for one scene: splats falling into the same voxel are merged (weighted means,
 * scales, SH; opacity = sum w^2 / sum w; normalised weighted quaternion).  Voxels come out in ascending flat index,
 * as torch.unique orders them; a voxel's splats are accumulated in ascending original index (the CPU scatter_add_ order).
 * Outputs are sized for N rows, the first *count (a device word) are valid.  workspace: wm_prune_workspace_bytes(N). */
wm_status wm_prune_workspace_bytes(long long N, size_t* bytes);
wm_status wm_prune_splats(wm_stream stream, long long N, const float* means, const float* quats, const float* scales,
                          const float* opacities, const float* sh, const float* weights, double voxel_size, void* workspace,
                          size_t workspace_bytes, float* out_means, float* out_quats, float* out_scales, float* out_opacities,
                          float* out_sh, unsigned* count);

/* The same merge for ONE KEY RANGE of a scene whose views are spread over several GPUs (heads.py:
 * B200GaussianSplatRenderer, partitioned merge): every rank computes the voxel-index bounds of its own splats
 * (wm_voxel_bounds: [6] int64 on the device, min xyz then max xyz), the ranks all-reduce them (min / max), key their splats
 * with the scene's bounds (wm_voxel_keys: the flat index of rasterization.py:322-331), exchange the splats so that every
 * rank holds one contiguous key range in (rank, index) order, and merge it with wm_prune_splats_bounds (= wm_prune_splats
 * keyed with the given bounds instead of those of its input).  The ranks' outputs, concatenated in rank order, are the
 * single-GPU result bit for bit. */
wm_status wm_voxel_bounds(wm_stream stream, long long N, const float* means, double voxel_size, long long* bounds);
wm_status wm_voxel_keys(wm_stream stream, long long N, const float* means, double voxel_size, const long long* bounds,
                        long long* keys);
wm_status wm_prune_splats_bounds(wm_stream stream, long long N, const float* means, const float* quats, const float* scales,
                                 const float* opacities, const float* sh, const float* weights, double voxel_size, void* workspace,
                                 size_t workspace_bytes, float* out_means, float* out_quats, float* out_scales,
                                 float* out_opacities, float* out_sh, unsigned* count, const long long* bounds);

/* apply_confidence_filter (rasterization.py:248-299) for one scene: conf [N] (confidences <= 1e-5 count as -inf), keep the K
 * most confident splats, K = max(1, min(max_gaussians, ceil(N * (100 - percent) / 100))) computed by the caller.  The
 * reference takes torch.topk(sorted=False), whose order and tie choice are unspecified; here equal confidences are taken in
 * ascending index and the K survivors come out in ascending original index (exact radix select + scan + gather).  Outputs
 * are sized for K rows.  workspace: wm_conf_filter_workspace_bytes(N). */
wm_status wm_conf_filter_workspace_bytes(long long N, size_t* bytes);
wm_status wm_conf_filter_splats(wm_stream stream, long long N, long long K, const float* conf, const float* means, const float* quats,
                                const float* scales, const float* opacities, const float* sh, const float* weights, void* workspace,
                                size_t workspace_bytes, float* out_means, float* out_quats, float* out_scales, float* out_opacities,
                                float* out_sh, float* out_weights);

/* ---- forward splat rasterisation: gsplat 1.5.3's rasterization() as the reference ---- */
/* calls it (src/models/models/rasterization.py:29-66: pinhole, classic mode, tile 16, eps2d 0.3, SH degree 0,
 * render_mode "RGB+ED"), restated from submodules/gsplat/gsplat/cuda/csrc/{ProjectionEWA3DGSPacked,IntersectTile,
 * RasterizeToPixels3DGSFwd}.cu.  Three calls, one host read between the first two:
 *   wm_raster_project: per (camera, gaussian) 2D mean, conic, depth, integer radii (0 = culled), tile count, and the
 *     inclusive scan `cum` of the tile counts (cum[C*N - 1] = number of intersections); quats are (w, x, y, z);
 *     viewmats [C, 4, 4] world-to-camera row major, Ks [C, 3, 3];
 *   wm_raster_workspace_bytes: scratch sizes for the scan (first call) and for sort + binning + the packed 48-byte
 *     per-intersection records of the compositing pass (second call);
 *   wm_raster_render: emit (image | tile | depth) keys, stable radix sort, tile offsets, one packed record per sorted
 *     intersection, then front-to-back compositing that streams each tile's contiguous run of records through shared
 *     memory with bulk asynchronous copies (gsplat's arithmetic, csrc/RasterizeToPixels3DGSFwd.cu:104-185; its own layout);
 *     sh [N, 3] degree-0 coefficients (colour = max(C0 sh + 0.5, 0)); offsets: scratch of C * tiles + 1 words;
 *     out_rgb [C, H, W, 3], out_depth [C, H, W] = expected depth, out_alpha [C, H, W]. */
wm_status wm_raster_project(wm_stream stream, const float* means, const float* quats, const float* scales, const float* opacities,
                            const float* viewmats, const float* Ks, int C, long long N, int W, int H, float near_plane,
                            float far_plane, float eps2d, float* means2d, float* conics, float* depths, int* radii, unsigned* tiles,
                            unsigned* cum, void* scan_tmp, size_t scan_tmp_bytes);
wm_status wm_raster_workspace_bytes(long long elems, long long n_isects, size_t* scan_bytes, size_t* render_bytes);
wm_status wm_raster_render(wm_stream stream, const float* means2d, const float* conics, const float* depths, const int* radii,
                           const unsigned* cum, const float* opacities, const float* sh, int C, long long N, int W, int H,
                           long long n_isects, void* workspace, size_t workspace_bytes, unsigned* offsets, float* out_rgb,
                           float* out_depth, float* out_alpha);

/* ---- image loader: prepare_images_to_tensor (src/utils/inference_utils.py:187-248) ---- */
/* from the decoded image on: Pillow's Image.resize((new_w, new_h), BICUBIC) (bit-exact: two separable passes, 22-bit
 * fixed-point coefficients, uint8 intermediate), torchvision's ToTensor (/ 255), the centre crop (rows [crop_top,
 * crop_top + crop_h) of the resized image) and the constant-1.0 padding: the result is placed at (pad_top, pad_left) of
 * out [3, out_H, out_W] fp32.  rgb: device uint8 [H, W, 3].  workspace: wm_load_image_workspace_bytes. */
wm_status wm_load_image_workspace_bytes(int H, int W, int new_w, int new_h, size_t* bytes);
wm_status wm_load_image(wm_stream stream, const unsigned char* rgb, int H, int W, int new_w, int new_h, int crop_top, int crop_h,
                        int out_H, int out_W, int pad_top, int pad_left, float* out, void* workspace, size_t workspace_bytes);
/* wm_load_image without a caller workspace and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA-graph capture around them, tight serving
 * loops, harnesses that forbid it).  The scratch (coefficient tables + the uint8 intermediate image) is a file-scope
 * static __device__ array of wm_load_image_static_bytes() = 8 MiB inside the library: every GPU that loads the library
 * holds it until exit.  A geometry whose wm_load_image_workspace_bytes exceeds it is rejected with WM_ERR_INVALID
 * (2048 x 1365 -> 518 wide needs 2.0 MiB; 8 MiB covers H * new_w <= about 2.7 M pixels).  The call itself makes no
 * cudaMalloc / cudaFree or allocating library call.  All calls share the array, which keeps the last call's contents;
 * every call overwrites what it reads, nothing needs to start at zero.  Calls that overlap on the GPU (different
 * streams without an ordering between them, or two host threads) race on it: serialise them or use wm_load_image. */
size_t wm_load_image_static_bytes(void);
wm_status wm_load_image_static(wm_stream stream, const unsigned char* rgb, int H, int W, int new_w, int new_h, int crop_top,
                               int crop_h, int out_H, int out_W, int pad_top, int pad_left, float* out);

#ifdef __cplusplus
}
#endif
#endif /* WM_B200_H */
