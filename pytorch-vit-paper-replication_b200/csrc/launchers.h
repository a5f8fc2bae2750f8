// Internal launcher prototypes shared between the kernel translation units and api.cu.
#pragma once
#include "common.cuh"

namespace b200vit {

int gemm_launch(int op, int epi, int M, int N, int K, const void* A, long long lda, const void* B, long long ldb,
                void* C, long long ldc, void* C2, long long ldc2, const float* bias, const void* aux,
                long long ldaux, float drop_p, unsigned long long seed, unsigned long long offset, int row_period,
                int splits, cudaStream_t stream);

void tmap_cache_stats(long long* hits, long long* misses);
void set_gemm_dynamic(int on);

// LayerNorm fold (gemm_sm100.cu): extra operands of epilogues 7-10
struct GemmLnArgs {
    void* x16; long long ldx16; float* rowstats;            // producer side (epilogues 7, 8)
    const float* ln_stats; const float* ln_s; int ln_dim; float ln_eps;   // consumer side (epilogues 9, 10)
};
int gemm_launch_ln(int op, int epi, int M, int N, int K, const void* A, long long lda, const void* B, long long ldb,
                   void* C, long long ldc, void* C2, long long ldc2, const float* bias, const void* aux,
                   long long ldaux, float drop_p, unsigned long long seed, unsigned long long offset, int row_period,
                   int splits, const GemmLnArgs* ln, cudaStream_t stream);
int fold_ln_weight_launch(const float* w, const float* gamma, const float* beta, const float* bias, bf16* w16, float* s,
                          float* c, int rows, int cols, cudaStream_t stream);

int cast_f32_bf16_launch(const float* src, bf16* dst, long long n, cudaStream_t stream);
int im2col_launch(const float* image, bf16* patches, int batch, int channels, int height, int width, int patch,
                  cudaStream_t stream);
int im2col_u8_launch(const uint8_t* image, bf16* patches, int batch, int channels, int height, int width, int patch,
                     cudaStream_t stream);
int cls_rows_launch(const float* cls, const float* pos, float* x, int batch, int tokens, int dim, float drop_p,
                    unsigned long long seed, unsigned long long offset, cudaStream_t stream, bf16* x16 = nullptr,
                    float* rowstats = nullptr);
int embed_grads_launch(const float* gx, float* g_cls, float* g_pos, float* g_bias, int batch, int tokens, int dim,
                       float drop_p, unsigned long long seed, unsigned long long offset, cudaStream_t stream);
int patch_grad_rows_launch(const float* gx, bf16* gy, int batch, int tokens, int dim, float drop_p,
                           unsigned long long seed, unsigned long long offset, cudaStream_t stream);
int colsum_bf16_launch(const bf16* x, long long ld, float* out, long long rows, int cols, cudaStream_t stream);
int dropout_cast_launch(const float* g, bf16* out, float* colsum, long long rows, int cols, float drop_p,
                        unsigned long long seed, unsigned long long offset, cudaStream_t stream);
int scale_by_scalar_launch(const float* src, const float* scalar_dev, float* dst, long long n, cudaStream_t stream);

int layernorm_fwd_launch(const float* x, long long row_stride, const float* gamma, const float* beta, bf16* y_bf16,
                         float* y_f32, float* mean, float* rstd, long long rows, int dim, float eps,
                         cudaStream_t stream);
int layernorm_bwd_launch(const bf16* dy_bf16, const float* dy_f32, const float* x, long long row_stride,
                         const float* mean, const float* rstd, const float* gamma, const float* gin, float* gout,
                         long long gout_stride, bf16* gout_bf16, float* dgamma, float* dbeta, float* dbias_next,
                         long long rows, int dim, float drop_p, unsigned long long seed, unsigned long long offset,
                         cudaStream_t stream);

int attention_fwd_launch(const bf16* qkv, bf16* out, float* lse, int batch, int tokens, int heads, int head_dim,
                         float scale, cudaStream_t stream);
int attention_fwd_tc_launch(const bf16* qkv, bf16* out, float* lse, int batch, int tokens, int heads, float scale,
                            cudaStream_t stream);
int attention_fwd_long_launch(const bf16* qkv, bf16* out, float* lse, int batch, int tokens, int heads, float scale,
                              cudaStream_t stream);
constexpr int AT_LONG_MAX_T = 2048;   // the long backward kernel stages lse/delta of a whole head in shared memory
int attention_bwd_long_launch(const bf16* qkv, const bf16* dout, const float* lse, const float* delta, bf16* dqkv,
                              int batch, int tokens, int heads, float scale, cudaStream_t stream);
int attention_bwd_tc_launch(const bf16* qkv, const bf16* out, const bf16* dout, const float* lse, bf16* dqkv, float* dbias,
                            int batch, int tokens, int heads, float scale, cudaStream_t stream);
int attention_bwd_launch(const bf16* qkv, const bf16* out, const bf16* dout, const float* lse, float* delta,
                         bf16* dqkv, float* dbias, int batch, int tokens, int heads, int head_dim, float scale,
                         cudaStream_t stream);
int attention_fwd_dropout_launch(const bf16* qkv, bf16* out, float* lse, int batch, int tokens, int heads, int head_dim,
                                 float scale, float p, unsigned long long seed, unsigned long long offset,
                                 cudaStream_t stream);
int attention_bwd_dropout_launch(const bf16* qkv, const bf16* out, const bf16* dout, const float* lse, float* delta,
                                 bf16* dqkv, float* dbias, int batch, int tokens, int heads, int head_dim, float scale,
                                 float p, unsigned long long seed, unsigned long long offset, cudaStream_t stream);

int head_fwd_launch(const float* xn, const float* w, const float* b, float* logits, int batch, int dim, int classes,
                    cudaStream_t stream);
int head_bwd_launch(const float* xn, const float* w, const float* dlogits, float* dxn, float* dw, float* db,
                    int batch, int dim, int classes, cudaStream_t stream);
int softmax_ce_launch(const float* logits, const long long* labels, float* loss, float* dlogits, int* correct,
                      int batch, int classes, cudaStream_t stream);

int softmax_top1_launch(const float* logits, float* probs, long long* labels, int batch, int classes, cudaStream_t stream);

int l2norm_sq_launch(const float* const* ptrs_dev, const long long* sizes_dev, int num_tensors, long long max_size,
                     float* partials, int num_partials, float* out_sumsq, cudaStream_t stream);
int adam_step_launch(const void* table_dev, int num_tensors, long long max_size, const float* lrs, const float* wds,
                     int num_groups, float beta1, float beta2, float eps, int step, int decoupled,
                     const float* grad_sumsq, float max_norm, float grad_scale, cudaStream_t stream);

// strict-precision ("fp32 mode") helpers, fp32_mode.cu
int split_bf16x3_launch(const float* src, long long ld_src, long long rows, int cols, bf16* dst, long long ld_dst,
                        long long part_stride, int pattern, float drop_p, unsigned long long seed,
                        unsigned long long offset, int row_period, cudaStream_t stream);
int im2col_f32_launch(const float* image, float* patches, int batch, int channels, int height, int width, int patch,
                      cudaStream_t stream);
int gelu_f32_fwd_launch(const float* a, float* h, long long n, float drop_p, unsigned long long seed,
                        unsigned long long offset, cudaStream_t stream);
int gelu_f32_bwd_launch(const float* a, const float* dh, float* da, long long n, float drop_p, unsigned long long seed,
                        unsigned long long offset, cudaStream_t stream);
int colsum_f32_launch(const float* x, long long ld, long long rows, int cols, float* out, float drop_p,
                      unsigned long long seed, unsigned long long offset, int row_period, cudaStream_t stream);
int attention_f32_fwd_launch(const float* qkv, float* out, float* lse, int batch, int tokens, int heads, int head_dim,
                             float scale, cudaStream_t stream);
int attention_f32_bwd_launch(const float* qkv, const float* out, const float* dout, const float* lse, float* delta,
                             float* dqkv, int batch, int tokens, int heads, int head_dim, float scale,
                             cudaStream_t stream);

}  // namespace b200vit
