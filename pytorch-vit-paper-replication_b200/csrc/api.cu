// extern "C" surface of libb200vit.so. Declarations (with the reference call sites each entry
// replaces) live in include/b200vit.h; this file only forwards to the kernel launchers.
#include "common.cuh"
#include "launchers.h"

#include "../../include/b200vit.h"

using namespace b200vit;

extern "C" {

int b200vit_gemm(int op, int epilogue, int M, int N, int K, const void* A, long long lda, const void* B,
                 long long ldb, void* C, long long ldc, void* C2, long long ldc2, const float* bias,
                 const void* aux, long long ldaux, float dropout_p, unsigned long long seed,
                 unsigned long long offset, int row_period, int split_k, void* stream) {
    return gemm_launch(op, epilogue, M, N, K, A, lda, B, ldb, C, ldc, C2, ldc2, bias, aux, ldaux, dropout_p, seed,
                       offset, row_period, split_k, (cudaStream_t)stream);
}

void b200vit_tmap_cache_stats(long long* hits, long long* misses) { tmap_cache_stats(hits, misses); }
int b200vit_set_gemm_schedule(int dynamic) { set_gemm_dynamic(dynamic); return 0; }

int b200vit_gemm_ln(int op, int epilogue, int M, int N, int K, const void* A, long long lda, const void* B, long long ldb,
                    void* C, long long ldc, void* C2, long long ldc2, const float* bias, const void* aux, long long ldaux,
                    float dropout_p, unsigned long long seed, unsigned long long offset, int row_period, void* x16,
                    long long ldx16, float* rowstats, const float* ln_stats, const float* ln_s, int ln_dim, float ln_eps,
                    void* stream) {
    GemmLnArgs ln{x16, ldx16, rowstats, ln_stats, ln_s, ln_dim, ln_eps};
    return gemm_launch_ln(op, epilogue, M, N, K, A, lda, B, ldb, C, ldc, C2, ldc2, bias, aux, ldaux, dropout_p, seed, offset,
                          row_period, 1, &ln, (cudaStream_t)stream);
}

int b200vit_fold_ln_weight(const float* w, const float* gamma, const float* beta, const float* bias, void* w16, float* s,
                           float* c, int rows, int cols, void* stream) {
    return fold_ln_weight_launch(w, gamma, beta, bias, (bf16*)w16, s, c, rows, cols, (cudaStream_t)stream);
}

int b200vit_class_token_rows_ln(const float* class_token, const float* pos, float* x, void* x16, float* rowstats, int batch,
                                int tokens, int dim, float dropout_p, unsigned long long seed, unsigned long long offset,
                                void* stream) {
    return cls_rows_launch(class_token, pos, x, batch, tokens, dim, dropout_p, seed, offset, (cudaStream_t)stream, (bf16*)x16,
                           rowstats);
}

int b200vit_cast_f32_to_bf16(const float* src, void* dst, long long n, void* stream) {
    return cast_f32_bf16_launch(src, (bf16*)dst, n, (cudaStream_t)stream);
}

int b200vit_im2col_patches(const float* image, void* patches, int batch, int channels, int height, int width,
                           int patch, void* stream) {
    return im2col_launch(image, (bf16*)patches, batch, channels, height, width, patch, (cudaStream_t)stream);
}
int b200vit_im2col_patches_u8(const unsigned char* image, void* patches, int batch, int channels, int height, int width,
                              int patch, void* stream) {
    return im2col_u8_launch(image, (bf16*)patches, batch, channels, height, width, patch, (cudaStream_t)stream);
}

int b200vit_class_token_rows(const float* class_token, const float* pos, float* x, int batch, int tokens, int dim,
                             float dropout_p, unsigned long long seed, unsigned long long offset, void* stream) {
    return cls_rows_launch(class_token, pos, x, batch, tokens, dim, dropout_p, seed, offset, (cudaStream_t)stream);
}

int b200vit_embed_param_grads(const float* gx, float* g_cls, float* g_pos, float* g_bias, int batch, int tokens,
                              int dim, float dropout_p, unsigned long long seed, unsigned long long offset,
                              void* stream) {
    return embed_grads_launch(gx, g_cls, g_pos, g_bias, batch, tokens, dim, dropout_p, seed, offset,
                              (cudaStream_t)stream);
}

int b200vit_patch_grad_rows(const float* gx, void* gy, int batch, int tokens, int dim, float dropout_p,
                            unsigned long long seed, unsigned long long offset, void* stream) {
    return patch_grad_rows_launch(gx, (bf16*)gy, batch, tokens, dim, dropout_p, seed, offset, (cudaStream_t)stream);
}

int b200vit_layernorm_fwd(const float* x, long long row_stride, const float* gamma, const float* beta, void* y_bf16,
                          float* y_f32, float* mean, float* rstd, long long rows, int dim, float eps, void* stream) {
    return layernorm_fwd_launch(x, row_stride, gamma, beta, (bf16*)y_bf16, y_f32, mean, rstd, rows, dim, eps,
                                (cudaStream_t)stream);
}

int b200vit_layernorm_bwd(const void* dy_bf16, const float* dy_f32, const float* x, long long row_stride,
                          const float* mean, const float* rstd, const float* gamma, const float* gin, float* gout,
                          long long gout_stride, void* gout_bf16, float* dgamma, float* dbeta, float* dbias_next,
                          long long rows, int dim, float dropout_p, unsigned long long seed,
                          unsigned long long offset, void* stream) {
    return layernorm_bwd_launch((const bf16*)dy_bf16, dy_f32, x, row_stride, mean, rstd, gamma, gin, gout, gout_stride,
                                (bf16*)gout_bf16, dgamma, dbeta, dbias_next, rows, dim, dropout_p, seed, offset,
                                (cudaStream_t)stream);
}

int b200vit_colsum_bf16(const void* x, long long ld, float* out, long long rows, int cols, void* stream) {
    return colsum_bf16_launch((const bf16*)x, ld, out, rows, cols, (cudaStream_t)stream);
}

int b200vit_dropout_cast(const float* g, void* out_bf16, float* colsum, long long rows, int cols, float dropout_p,
                         unsigned long long seed, unsigned long long offset, void* stream) {
    return dropout_cast_launch(g, (bf16*)out_bf16, colsum, rows, cols, dropout_p, seed, offset, (cudaStream_t)stream);
}

int b200vit_attention_fwd(const void* qkv, void* out, float* lse, int batch, int tokens, int heads, int head_dim,
                          float scale, void* stream) {
    return attention_fwd_launch((const bf16*)qkv, (bf16*)out, lse, batch, tokens, heads, head_dim, scale,
                                (cudaStream_t)stream);
}

int b200vit_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse, float* delta,
                          void* dqkv, int batch, int tokens, int heads, int head_dim, float scale, void* stream) {
    return attention_bwd_launch((const bf16*)qkv, (const bf16*)out, (const bf16*)dout, lse, delta, (bf16*)dqkv, nullptr,
                                batch, tokens, heads, head_dim, scale, (cudaStream_t)stream);
}
int b200vit_attention_fwd_dropout(const void* qkv, void* out, float* lse, int batch, int tokens, int heads, int head_dim,
                                  float scale, float dropout_p, unsigned long long seed, unsigned long long offset,
                                  void* stream) {
    return attention_fwd_dropout_launch((const bf16*)qkv, (bf16*)out, lse, batch, tokens, heads, head_dim, scale,
                                        dropout_p, seed, offset, (cudaStream_t)stream);
}
int b200vit_attention_bwd_dropout(const void* qkv, const void* out, const void* dout, const float* lse, float* delta,
                                  void* dqkv, float* dbias, int batch, int tokens, int heads, int head_dim, float scale,
                                  float dropout_p, unsigned long long seed, unsigned long long offset, void* stream) {
    return attention_bwd_dropout_launch((const bf16*)qkv, (const bf16*)out, (const bf16*)dout, lse, delta, (bf16*)dqkv,
                                        dbias, batch, tokens, heads, head_dim, scale, dropout_p, seed, offset,
                                        (cudaStream_t)stream);
}
int b200vit_attention_bwd_bias(const void* qkv, const void* out, const void* dout, const float* lse, float* delta,
                               void* dqkv, float* dbias, int batch, int tokens, int heads, int head_dim, float scale,
                               void* stream) {
    return attention_bwd_launch((const bf16*)qkv, (const bf16*)out, (const bf16*)dout, lse, delta, (bf16*)dqkv, dbias,
                                batch, tokens, heads, head_dim, scale, (cudaStream_t)stream);
}

int b200vit_head_fwd(const float* xn, const float* w, const float* b, float* logits, int batch, int dim, int classes,
                     void* stream) {
    return head_fwd_launch(xn, w, b, logits, batch, dim, classes, (cudaStream_t)stream);
}

int b200vit_head_bwd(const float* xn, const float* w, const float* dlogits, float* dxn, float* dw, float* db,
                     int batch, int dim, int classes, void* stream) {
    return head_bwd_launch(xn, w, dlogits, dxn, dw, db, batch, dim, classes, (cudaStream_t)stream);
}

int b200vit_softmax_ce(const float* logits, const long long* labels, float* loss, float* dlogits, int* correct,
                       int batch, int classes, void* stream) {
    return softmax_ce_launch(logits, labels, loss, dlogits, correct, batch, classes, (cudaStream_t)stream);
}

int b200vit_softmax_top1(const float* logits, float* probs, long long* labels, int batch, int classes, void* stream) {
    return softmax_top1_launch(logits, probs, labels, batch, classes, (cudaStream_t)stream);
}

int b200vit_scale_by_scalar(const float* src, const float* scalar_dev, float* dst, long long n, void* stream) {
    return scale_by_scalar_launch(src, scalar_dev, dst, n, (cudaStream_t)stream);
}

int b200vit_l2norm_sq(const void* const* ptrs_dev, const long long* sizes_dev, int num_tensors, long long max_size,
                      float* partials, int num_partials, float* out_sumsq, void* stream) {
    return l2norm_sq_launch((const float* const*)ptrs_dev, sizes_dev, num_tensors, max_size, partials, num_partials,
                            out_sumsq, (cudaStream_t)stream);
}

int b200vit_adam_step(const void* table_dev, int num_tensors, long long max_size, const float* lrs, const float* wds,
                      int num_groups, float beta1, float beta2, float eps, int step, int decoupled,
                      const float* grad_sumsq, float max_norm, float grad_scale, void* stream) {
    return adam_step_launch(table_dev, num_tensors, max_size, lrs, wds, num_groups, beta1, beta2, eps, step, decoupled,
                            grad_sumsq, max_norm, grad_scale, (cudaStream_t)stream);
}

int b200vit_split_bf16x3(const float* src, long long ld_src, long long rows, int cols, void* dst_bf16, long long ld_dst,
                         long long part_stride, int pattern, float dropout_p, unsigned long long seed,
                         unsigned long long offset, int row_period, void* stream) {
    return split_bf16x3_launch(src, ld_src, rows, cols, (bf16*)dst_bf16, ld_dst, part_stride, pattern, dropout_p, seed,
                               offset, row_period, (cudaStream_t)stream);
}

int b200vit_im2col_patches_f32(const float* image, float* patches, int batch, int channels, int height, int width,
                               int patch, void* stream) {
    return im2col_f32_launch(image, patches, batch, channels, height, width, patch, (cudaStream_t)stream);
}

int b200vit_gelu_f32_fwd(const float* a, float* h, long long n, float dropout_p, unsigned long long seed,
                         unsigned long long offset, void* stream) {
    return gelu_f32_fwd_launch(a, h, n, dropout_p, seed, offset, (cudaStream_t)stream);
}

int b200vit_gelu_f32_bwd(const float* a, const float* dh, float* da, long long n, float dropout_p,
                         unsigned long long seed, unsigned long long offset, void* stream) {
    return gelu_f32_bwd_launch(a, dh, da, n, dropout_p, seed, offset, (cudaStream_t)stream);
}

int b200vit_colsum_f32(const float* x, long long ld, long long rows, int cols, float* out, float dropout_p,
                       unsigned long long seed, unsigned long long offset, int row_period, void* stream) {
    return colsum_f32_launch(x, ld, rows, cols, out, dropout_p, seed, offset, row_period, (cudaStream_t)stream);
}

int b200vit_attention_f32_fwd(const float* qkv, float* out, float* lse, int batch, int tokens, int heads, int head_dim,
                              float scale, void* stream) {
    return attention_f32_fwd_launch(qkv, out, lse, batch, tokens, heads, head_dim, scale, (cudaStream_t)stream);
}

int b200vit_attention_f32_bwd(const float* qkv, const float* out, const float* dout, const float* lse, float* delta,
                              float* dqkv, int batch, int tokens, int heads, int head_dim, float scale, void* stream) {
    return attention_f32_bwd_launch(qkv, out, dout, lse, delta, dqkv, batch, tokens, heads, head_dim, scale,
                                    (cudaStream_t)stream);
}

}  // extern "C"
