// Internal (C++) prototypes of the launchers defined in gemm.cu, attn.cu and rowops.cu.
#pragma once
#include "common.cuh"

namespace wm {

wm_status gemm_bias_act(cudaStream_t st, const void* A, int lda, const void* W, int ldw, const float* bias, int M,
                        int N, int K, void* out, int ldo, int gelu);
wm_status gemm_qkv(cudaStream_t st, const void* A, int lda, const void* W, const float* bias, const float* qn_w,
                   const float* qn_b, const float* kn_w, const float* kn_b, float qk_eps, const float* rope_cos,
                   const float* rope_sin, int tokens_per_view, int patch_start, int grid_w, float q_scale, int M,
                   void* q_out, void* kv_out, unsigned* norm2 = nullptr);  // norm2: AttnGuard words [2 * 16], or null
// flags: kGemmFlagBytes of ZEROED counters (left zero by every launch) that let the kernel cut the tiles beyond the
// last full wave along K and add the slices in fixed order (gemm.cu); null: whole tiles only.
wm_status gemm_ls_residual(cudaStream_t st, const void* A, int lda, const void* W, int ldw, const float* bias,
                           const float* gamma, int M, int N, int K, float* x, int ldx, unsigned* flags = nullptr);
constexpr size_t kGemmFlagBytes = 4096;
wm_status copy_cols(cudaStream_t st, const float* x, long long rows, float* out, int ld_out, int col0);
wm_status gemm_table(cudaStream_t st, const void* A, int lda, const void* W, int ldw, const float* table,
                     int table_period, int M, int N, int K, float* out, int ldo);
wm_status flash_attention(cudaStream_t st, const void* q, int ldq, const void* k, const void* v, int ldkv, void* o,
                          int ldo, int batch, int Lq, int Lkv, int heads, float scale, int variant, float* partial_ws,
                          size_t partial_ws_bytes, const AttnGather* gather = nullptr, const AttnGuard* guard = nullptr);
size_t attn_partial_bytes(int sm_count);  // workspace that lets the split-row kernel balance (item, kv tile) steps
wm_status umma_probe(cudaStream_t st, const void* p, const void* v, float* d, int mode, uint32_t v_lbo, uint32_t v_sbo);
wm_status layernorm(cudaStream_t st, const float* x, const float* w, const float* b, float eps, int rows,
                    int rows_per_group, int in_group_rows, int in_offset, int out_group_rows, int out_offset, void* y,
                    int out_f32, int accumulate);
wm_status patch_gather(cudaStream_t st, const float* img, int n_views, int channels, int H, int W, int rows_per_view,
                       int row_offset, int ld, int normalize, void* out);
wm_status prior_token(cudaStream_t st, const float* in, int n_views, int d_in, const float* w0, const float* b0,
                      const float* w2, const float* b2, int rows_per_view, int row, float* x);
wm_status special_tokens(cudaStream_t st, const float* cam, const float* reg, int n_views, int rows_per_view,
                         int first_view_global, int zero_pose, int zero_ray, float* x);
// prediction heads (conv.cu, headops.cu)
wm_status conv2d_nhwc(cudaStream_t st, const void* in, int n, int H, int W, int Cin, const void* weight, int taps, int Cout,
                      const float* bias, const void* res1, int res1_bcast, const void* res2, int relu, void* out);
// the count of fp16 saturations since the last reset (conv.cu: g_fp16_saturated), copied to pinned host memory on `st`
wm_status fp16_saturation_read(cudaStream_t st, unsigned* host_pinned, int reset);
wm_status conv2d_final(cudaStream_t st, const void* in, int n, int H, int W, int Cin, const void* w1, const float* b1,
                       const float* w2, const float* b2, int od, int act, float* pred, float* conf);
wm_status ln_rows(cudaStream_t st, const float* x, const float* w, const float* b, float eps, long long rows,
                  int rows_per_group, int in_group_rows, int in_offset, int C, void* y);
wm_status bilinear_nhwc(cudaStream_t st, const void* in, int n, int Hi, int Wi, int C, int Ho, int Wo, const float* pos_u,
                        const float* pos_v, void* out);
wm_status pixel_shuffle(cudaStream_t st, const void* in, int n, int Hi, int Wi, int C, int k, void* out);
wm_status im2col_s2(cudaStream_t st, const void* in, int n, int Hi, int Wi, int C, void* out);
// camera head (camhead.cu)
wm_status skinny_linear(cudaStream_t st, const float* A, int lda, const float* W, const float* bias, int M, int N, int K, int silu_in,
                        int act, const float* resid, const float* gamma, float* out, int ldo);
wm_status cam_layernorm(cudaStream_t st, const float* x, long long in_stride, const float* w, const float* b, float eps, int M,
                        const float* mod, float* y);
wm_status cam_attention(cudaStream_t st, const float* qkv, int S, float* out);
wm_status cam_embed(cudaStream_t st, const float* in, int in_bcast, int K, const float* W, const float* bias, int M, int N, float* out);
wm_status cam_predict(cudaStream_t st, const float* h, int K, const float* W, const float* bias, int M, int first, float* pred,
                      float* act);
// prior preparation (priors.cu)
wm_status normalize_depth(cudaStream_t st, const float* depth, int n_views, int hw, double eps, int min_percentile,
                          int max_percentile, float* out);
wm_status pose_prior(cudaStream_t st, const float* extr, int extr_stride, const float* intr, int S, int H, int W, double padding,
                     float* poses, float* rays);
// Gaussian path (conv.cu, splat.cu)
wm_status conv2d_project(cudaStream_t st, const void* in, int n, int H, int W, int Cin, const void* w1, const float* b1,
                         const float* w2, const float* b2, int od, float* out);
wm_status im2row7(cudaStream_t st, const float* img, int n, int H, int W, void* out);
wm_status splat_cameras(cudaStream_t st, const float* params, int n, int H, int W, float* cam);
wm_status splat_prepare(cudaStream_t st, const float* raw, const float* img, const float* depth, const float* cam, int n, int H,
                        int W, float* means, float* quats, float* scales, float* opacities, float* sh, float* weights);
wm_status prune_workspace_bytes(long long N, size_t* bytes);
wm_status prune_splats(cudaStream_t st, long long N, const float* means, const float* quats, const float* scales,
                       const float* opacities, const float* sh, const float* weights, double voxel_size, void* workspace,
                       size_t workspace_bytes, float* out_means, float* out_quats, float* out_scales, float* out_opacities, float* out_sh,
                       unsigned* count, const long long* ext_bounds = nullptr);
wm_status voxel_bounds(cudaStream_t st, long long N, const float* means, double voxel_size, long long* bounds);
wm_status voxel_keys(cudaStream_t st, long long N, const float* means, double voxel_size, const long long* bounds, long long* keys);
wm_status conf_filter_workspace_bytes(long long N, size_t* bytes);
wm_status conf_filter_splats(cudaStream_t st, long long N, long long K, const float* conf, const float* means, const float* quats,
                             const float* scales, const float* opacities, const float* sh, const float* weights, void* workspace,
                             size_t workspace_bytes, float* out_means, float* out_quats, float* out_scales, float* out_opacities,
                             float* out_sh, float* out_weights);
// splat rasterisation (raster.cu)
wm_status raster_project(cudaStream_t st, const float* means, const float* quats, const float* scales, const float* opacities,
                         const float* viewmats, const float* Ks, int C, long long N, int W, int H, float near_plane, float far_plane,
                         float eps2d, float* means2d, float* conics, float* depths, int* radii, unsigned* tiles, unsigned* cum,
                         void* scan_tmp, size_t scan_tmp_bytes);
wm_status raster_workspace_bytes(long long elems, long long n_isects, size_t* scan_bytes, size_t* render_bytes);
wm_status raster_render(cudaStream_t st, const float* means2d, const float* conics, const float* depths, const int* radii,
                        const unsigned* cum, const float* opacities, const float* sh, int C, long long N, int W, int H,
                        long long n_isects, void* workspace, size_t workspace_bytes, unsigned* offsets, float* out_rgb,
                        float* out_depth, float* out_alpha);
// image loader (loader.cu)
wm_status load_image_workspace_bytes(int H, int W, int new_w, int new_h, size_t* bytes);
wm_status load_image(cudaStream_t st, const unsigned char* rgb, int H, int W, int new_w, int new_h, int crop_top, int crop_h, int out_H,
                     int out_W, int pad_top, int pad_left, float* out, void* workspace, size_t workspace_bytes);
size_t load_image_static_capacity();
wm_status load_image_static(cudaStream_t st, const unsigned char* rgb, int H, int W, int new_w, int new_h, int crop_top, int crop_h,
                            int out_H, int out_W, int pad_top, int pad_left, float* out);
wm_status rope2d(cudaStream_t st, const float* tokens, const long long* pos, int batch, int heads, int L, int max_pos,
                 const float* cos_t, const float* sin_t, float* out);
wm_status pack_bf16_rows(cudaStream_t st, const float* src, long long rows, int cols, int ld, void* dst);

}  // namespace wm
