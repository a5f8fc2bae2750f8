/* libb200vit.so -- C ABI of the B200-native ViT-B/16 training-step kernels.
 *
 * The reference (AvalonEnjoyer/pytorch-ViT-paper-replication) has no FFI: its hot path is Python calling
 * torch.nn modules. The drop-in boundary is therefore the Python module API (b200vit.models.vit.ViT etc.),
 * and THIS header is what that Python layer binds through ctypes. Every entry point below names the
 * reference call site (file:line under the reference repo) whose arithmetic it replaces.
 *
 * Conventions
 *   - All pointers are DEVICE pointers unless a parameter says "host". Nothing here allocates or frees
 *     memory the caller can see; outputs are written into caller-provided buffers.
 *   - `stream` is a cudaStream_t passed as void*. Calls are asynchronous on that stream.
 *   - Return value: 0 = ok, < 0 = argument / shape / alignment error, > 0 = cudaError_t.
 *     b200vit_last_error() returns a thread-local message for the last non-zero return.
 *   - "bf16" buffers are __nv_bfloat16 (passed as void*); "f32" are float.
 *   - Dropout masks are never stored: a mask is a pure function of (seed, offset, element index), so the
 *     forward and backward entry points of one dropout site must be given the same (p, seed, offset).
 *   - There is no CPU implementation behind any of these symbols.
 */
#ifndef B200VIT_H
#define B200VIT_H

#ifdef __cplusplus
extern "C" {
#endif

/* ABI version of this header (bumped on any signature change). */
int b200vit_abi_version(void);

/* Message describing the last error raised on the calling thread. */
const char* b200vit_last_error(void);

/* Leave `sms` streaming multiprocessors out of the grids of every kernel launched from now on (0 restores the full
 * device). The persistent kernels split their work statically over one CTA per SM, so a kernel of another stream that
 * occupies some SMs (a gradient all-reduce overlapped with backward -- the reference has no counterpart, it is
 * single-device, going_modular/engine.py:180) would stall whole CTAs; with a reserve it finds free SMs instead. Takes
 * effect at launch time of each kernel; process-wide. */
int b200vit_set_sm_reserve(int sms);

/* Sum-all-reduce of one fp32 gradient bucket over peer memory (NVLink / NVSwitch), for data-parallel training of the
 * hot path. No counterpart in the reference, which is single-device (going_modular/engine.py:180); called by
 * b200vit/dp.py in place of ncclAllReduce. bufs[r] / flags[r] (r < world <= 8) are THIS process's mappings of rank r's
 * bucket and of rank r's flag words (B200VIT_PEER_FLAG_WORDS zero-initialised 32-bit words, e.g. a torch
 * symmetric-memory allocation; the library maps and allocates nothing). Every rank must call it once per bucket with the same count (a multiple of 4)
 * and the same seq, and seq must grow by one from call to call on a given set of flag words, starting at 1. The kernel
 * waits (bounded, ~20 s, then a launch error) until all ranks' buckets are complete, sums slice `rank` of all of them
 * in rank order and writes the sum into every rank's bucket, and ends when all slices of the local bucket are final:
 * stream order after it is enough to read the result. `ctas` CTAs of 64 threads without shared memory are launched (more
 * than there are SMs is the intended use: all but the first CTA to land on an SM leave at once, the rest claim the
 * work in chunks), so the kernel shares SMs with the backward kernels it overlaps. Plain 16-byte loads and stores. */
#define B200VIT_PEER_FLAG_WORDS 320
int b200vit_allreduce_peer_f32(void* const* bufs, void* const* flags, int world, int rank, long long count,
                               unsigned seq, int ctas, void* stream);

/* How the persistent 2-CTA tcgen05 GEMM hands out its (tile, split) work items. dynamic != 0 (the default; the
 * environment variable B200VIT_GEMM_SCHED=static turns it off at load): a cluster works on its own index first and then
 * claims further items from a counter in device memory (a file-scope array of the library; nothing is allocated), so
 * a cluster whose SMs are held by another stream's kernel -- the gradient all-reduce that overlaps backward under
 * b200vit.dp.DataParallel; the single-device reference, going_modular/engine.py:180, has no counterpart -- claims less
 * instead of delaying the whole product by its static share. dynamic == 0: item i of cluster c is c + i * clusters.
 * Results do not depend on the setting (each item is computed the same way whoever takes it; split-K sums through
 * atomics in either case). Process-wide, takes effect at the next launch. */
int b200vit_set_gemm_schedule(int dynamic);

/* ------------------------------------------------------------------------------------------------
 * Dense contractions on the tcgen05 tensor cores (bf16 operands, fp32 accumulation in tensor memory).
 *
 *   op 0 (NT): C[M,N] = A[M,K] . B[N,K]^T    forward of nn.Linear / the patch Conv2d-as-GEMM / MHA in- and
 *                                            out-projection  (models/vit.py:28-31, 86-90, 119, 123)
 *   op 1 (NN): C[M,N] = A[M,K] . B[K,N]      input gradients of the same layers (autograd of the above)
 *   op 2 (TN): C[M,N] = A[K,M]^T . B[K,N]    weight gradients (contraction over the B*T token rows)
 *
 * A, B: bf16 row-major with leading dimensions lda, ldb (elements, multiples of 8; 16-byte aligned bases).
 * epilogue:
 *   0 BF16       C bf16 = acc (+ bias[n] if bias != NULL)
 *   1 BIAS_GELU  a = acc + bias ; C2 bf16 = dropout(gelu_erf(a)) ; C bf16 = dropout_mask/(1-p) * gelu_erf'(a), the
 *                operand the backward epilogue 3 multiplies by (skipped if C == NULL)    (models/vit.py:119-122)
 *   2 BIAS_RES   C f32 = dropout(acc + bias) + aux_f32[m,n]  (aux NULL -> no residual)
 *                                                              (models/vit.py:123-125 and the residual adds :167-168)
 *   3 DGELU      C bf16 = acc * aux_bf16[m,n], aux = the C output of epilogue 1      (backward of GELU + Dropout);
 *                if `bias` is non-NULL it is an OUTPUT here: the column sums of C (the bias gradient of the Linear in
 *                front of the GELU, reference autograd of nn.Linear vit.py:119) are ADDED to bias[0..N) with fp32
 *                atomics -- the caller zeroes it first
 *   4 ACCUM      C f32 += acc, atomically; the only epilogue that allows split_k > 1 (weight gradients)
 *   5 F32        C f32 = acc (+ bias)
 *   6 PATCH      C f32[m + m/row_period + 1, n] = dropout(acc + bias[n] + aux_f32[(m % row_period) + 1, n]):
 *                patch rows of the embedded sequence, position embedding added, class-token row skipped
 *                                                              (models/vit.py:53-66)
 * Dropout uses element index (output_row * N + n).
 */
int b200vit_gemm(int op, int epilogue, int M, int N, int K,
                 const void* A, long long lda, const void* B, long long ldb,
                 void* C, long long ldc, void* C2, long long ldc2,
                 const float* bias, const void* aux, long long ldaux,
                 float dropout_p, unsigned long long seed, unsigned long long offset,
                 int row_period, int split_k, void* stream);

/* fp32 -> bf16 copy (bf16 weight shadows of the fp32 master parameters). */
int b200vit_cast_f32_to_bf16(const float* src, void* dst_bf16, long long n, void* stream);

/* Patch extraction for a stride == kernel convolution: image f32 [B,C,H,W] -> patches bf16 [B*(H/P)*(W/P), C*P*P]
 * with column order (c, i, j) = flattened Conv2d weight order.  Replaces the data movement of
 * nn.Conv2d(k=s=P) + Flatten(2,3) + permute(0,2,1)  (models/vit.py:28-34, 53-56). */
int b200vit_im2col_patches(const float* image, void* patches_bf16, int batch, int channels, int height, int width,
                           int patch, void* stream);

/* The same for uint8 images [B, C, H, W], converted to [0, 1] floats on the fly (value / 255): the reference's
 * ToImage + ToDtype(torch.float32, scale=True) transform (08_pytorch_paper_replicating_self.ipynb:253-265) moved to
 * the device, so a batch crosses PCIe as 1 byte per value instead of 4. */
int b200vit_im2col_patches_u8(const unsigned char* image_u8, void* patches_bf16, int batch, int channels, int height,
                              int width, int patch, void* stream);

/* x[b, 0, :] = dropout(class_token + pos[0]) for every b; x is the f32 [B, T, D] embedded sequence whose patch
 * rows the PATCH GEMM epilogue writes.  Replaces expand + torch.cat + add + Dropout for the class row
 * (models/vit.py:50, 59, 63, 66). */
int b200vit_class_token_rows(const float* class_token, const float* pos, float* x, int batch, int tokens, int dim,
                             float dropout_p, unsigned long long seed, unsigned long long offset, void* stream);

/* Gradients of class_token [D], position_embedding [T, D] and the conv bias [D] (g_bias may be NULL) from the
 * gradient gx f32 [B, T, D] of the embedded sequence (autograd of models/vit.py:50-66). Outputs are overwritten. */
int b200vit_embed_param_grads(const float* gx, float* g_cls, float* g_pos, float* g_bias, int batch, int tokens,
                              int dim, float dropout_p, unsigned long long seed, unsigned long long offset,
                              void* stream);

/* gy bf16 [B*(T-1), D] = dropout_mask * gx[:, 1:, :]: the token-gradient operand of the conv weight gradient. */
int b200vit_patch_grad_rows(const float* gx, void* gy_bf16, int batch, int tokens, int dim, float dropout_p,
                            unsigned long long seed, unsigned long long offset, void* stream);

/* nn.LayerNorm(dim) forward (models/vit.py:83, 115, 216; eps 1e-5, biased variance).
 * x: f32 rows of `dim` values, `row_stride` elements apart (row_stride = T*dim normalises only class-token rows).
 * Writes y_bf16 and/or y_f32 (dense [rows, dim]; either may be NULL) and the row statistics mean / rstd. */
int b200vit_layernorm_fwd(const float* x, long long row_stride, const float* gamma, const float* beta,
                          void* y_bf16, float* y_f32, float* mean, float* rstd, long long rows, int dim, float eps,
                          void* stream);

/* nn.LayerNorm backward fused with the residual-gradient add.
 *   dy (exactly one of dy_bf16 / dy_f32, dense [rows, dim]) is the gradient w.r.t. the LayerNorm output.
 *   gout[row*gout_stride + :] = (gin ? gin[same index] : 0) + dLN/dx ; dgamma, dbeta f32 [dim] are overwritten.
 *   gout_bf16 (optional, dense): bf16(dropout_mask * gout); dbias_next (optional) = its column sums.
 * Autograd counterpart of models/vit.py:93, 129, 232 plus the `+ x` of :167-168. dim <= 1024. */
int b200vit_layernorm_bwd(const void* dy_bf16, const float* dy_f32, const float* x, long long row_stride,
                          const float* mean, const float* rstd, const float* gamma, const float* gin, float* gout,
                          long long gout_stride, void* gout_bf16, float* dgamma, float* dbeta, float* dbias_next,
                          long long rows, int dim, float dropout_p, unsigned long long seed,
                          unsigned long long offset, void* stream);

/* out[n] = sum_m x[m, n] for a bf16 matrix: bias gradients of the Linear layers. `out` is overwritten. */
int b200vit_colsum_bf16(const void* x_bf16, long long ld, float* out, long long rows, int cols, void* stream);

/* out_bf16 = bf16(dropout_mask * g), colsum (optional) = column sums of out: turns the f32 residual-stream
 * gradient into the operand of the fc2 backward GEMMs, undoing nn.Dropout of models/vit.py:125. */
int b200vit_dropout_cast(const float* g, void* out_bf16, float* colsum, long long rows, int cols, float dropout_p,
                         unsigned long long seed, unsigned long long offset, void* stream);

/* Fused multi-head self-attention forward (the scaled_dot_product_attention inside nn.MultiheadAttention,
 * models/vit.py:86-97; need_weights=False, no mask).
 *   qkv bf16 [B, T, 3, H, 64] (the packed in-projection output), out bf16 [B, T, H*64], lse f32 [B, H, T]. */
int b200vit_attention_fwd(const void* qkv_bf16, void* out_bf16, float* lse, int batch, int tokens, int heads,
                          int head_dim, float scale, void* stream);

/* Backward of the above. `delta` is SCRATCH f32 [B, H, T] provided by the caller; dqkv bf16 [B, T, 3, H, 64]. */
int b200vit_attention_bwd(const void* qkv_bf16, const void* out_bf16, const void* dout_bf16, const float* lse,
                          float* delta, void* dqkv_bf16, int batch, int tokens, int heads, int head_dim, float scale,
                          void* stream);

/* Same, and additionally dbias f32 [3 * H * 64] = column sums of dqkv over all B*T rows: the gradient of
 * nn.MultiheadAttention.in_proj_bias (models/vit.py:86-90; torch adds the packed bias in F.linear, so its gradient
 * is the row sum of the packed projection's gradient). dbias is overwritten. For tokens <= 256 the sums come out of
 * the backward kernel itself (no second pass over dqkv); longer sequences run a column-sum pass afterwards. */
int b200vit_attention_bwd_bias(const void* qkv_bf16, const void* out_bf16, const void* dout_bf16, const float* lse,
                               float* delta, void* dqkv_bf16, float* dbias, int batch, int tokens, int heads,
                               int head_dim, float scale, void* stream);

/* Attention with dropout on the probabilities: nn.MultiheadAttention(dropout=attn_dropout) in training mode
 * (models/vit.py:86-90; torch drops softmax(S) before the PV product and scales the kept entries by 1/(1-p)).
 * Element (b, h, q, k) of the probability tensor uses index ((b*H + h)*T + q)*T + k of the (seed, offset) mask, so
 * the backward call regenerates the forward's mask from the same triple. dropout_p == 0 forwards to the entry points
 * above; dropout_p > 0 runs the mma.sync kernels (the reference default is 0, the tcgen05 kernels carry no mask code).
 * dbias may be NULL. */
int b200vit_attention_fwd_dropout(const void* qkv_bf16, void* out_bf16, float* lse, int batch, int tokens, int heads,
                                  int head_dim, float scale, float dropout_p, unsigned long long seed,
                                  unsigned long long offset, void* stream);
int b200vit_attention_bwd_dropout(const void* qkv_bf16, const void* out_bf16, const void* dout_bf16, const float* lse,
                                  float* delta, void* dqkv_bf16, float* dbias, int batch, int tokens, int heads,
                                  int head_dim, float scale, float dropout_p, unsigned long long seed,
                                  unsigned long long offset, void* stream);

/* Classifier head on the normalised class-token rows: logits[B, C] = xn[B, D] . w[C, D]^T + b
 * (nn.Linear(D, num_classes), models/vit.py:219-222, 235); fp32 on the master weights. */
int b200vit_head_fwd(const float* xn, const float* w, const float* b, float* logits, int batch, int dim, int classes,
                     void* stream);
/* dxn / dw / db (each optional) from dlogits; outputs are overwritten. */
int b200vit_head_bwd(const float* xn, const float* w, const float* dlogits, float* dxn, float* dw, float* db,
                     int batch, int dim, int classes, void* stream);

/* nn.CrossEntropyLoss() (mean) fused with its gradient and the accuracy counter:
 *   *loss = mean_b(logsumexp(z_b) - z_b[y_b]) ; dlogits (optional) = (softmax(z) - onehot(y)) / B ;
 *   *correct (optional) = #{b : argmax(z_b) == y_b}.
 * A label outside [0, classes) makes *loss and that sample's dlogits row NaN (torch raises a device-side assert there).
 * Replaces loss_fn(y_pred, y) and the argmax(softmax(.)) == y count of going_modular/engine.py:53, 73-74, 121-125. */
int b200vit_softmax_ce(const float* logits, const long long* labels, float* loss, float* dlogits, int* correct,
                       int batch, int classes, void* stream);

/* Prediction post-processing: probs f32 [B, C] = softmax(logits) per row, labels int64 [B] = argmax (first maximal index);
 * either output may be NULL. Replaces torch.softmax(dim=1) + torch.argmax(dim=1) of going_modular/predictions.py:62-66. */
int b200vit_softmax_top1(const float* logits, float* probs, long long* labels, int batch, int classes, void* stream);

/* dst = src * (*scalar_dev): chain rule with the upstream scalar gradient of the loss. */
int b200vit_scale_by_scalar(const float* src, const float* scalar_dev, float* dst, long long n, void* stream);

/* Sum of squares over a list of f32 tensors in one launch (+ a tiny final reduction): the global gradient norm of
 * torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0), going_modular/engine.py:63.
 *   ptrs_dev / sizes_dev: device arrays of num_tensors pointers / element counts; max_size = largest count.
 *   partials: SCRATCH f32 buffer with num_partials >= num_tensors * ceil(max_size / 8192) entries. */
int b200vit_l2norm_sq(const void* const* ptrs_dev, const long long* sizes_dev, int num_tensors, long long max_size,
                      float* partials, int num_partials, float* out_sumsq, void* stream);

/* Same result as b200vit_l2norm_sq, for callers that cannot provide (or allocate) the scratch buffer: a process
 * whose device memory was handed to its own arena up front, so that cudaMalloc fails, or a loop that cannot afford
 * the device-wide synchronisation cudaMalloc / cudaFree imply. The partial sums go to a fixed 1 MiB array that lives
 * inside the library image on every GPU that loads it. This entry point has no counterpart in the reference.
 *   - Limit: num_tensors * ceil(max_size / 8192) <= 262144 partial sums; larger problems return an argument error
 *     (ViT-B/16 needs 43 928, ViT-L/16 151 552).
 *   - It calls no cudaMalloc / cudaFree / allocating library routine.
 *   - The first call on each device looks up the device address of the built-in array (cudaGetSymbolAddress);
 *     make that first call before starting any stream capture.
 *   - All calls on a device share the one array: two calls that overlap on the GPU (different streams) race on it.
 *     Nothing has to be cleared between calls -- every slot that is read is written first in the same call.
 *   - The looked-up address is cached per device ordinal for the life of the process. cudaDeviceReset() destroys the
 *     context the address belongs to; a host that resets a device must not call this entry point on it afterwards. */
int b200vit_l2norm_sq_noalloc(const void* const* ptrs_dev, const long long* sizes_dev, int num_tensors,
                              long long max_size, float* out_sumsq, void* stream);

/* One fused optimizer step over all parameters: torch.optim.Adam as configured by the reference recipe
 * (coupled L2 weight decay per param group; notebook optimizer cell, stepped at going_modular/engine.py:66),
 * or AdamW when decoupled != 0.
 *   table_dev: device array of num_tensors records {f32* p, const f32* g, f32* m, f32* v, bf16* shadow (or NULL),
 *              long long n, int group, int pad}.
 *   lrs / wds: HOST arrays with num_groups (<= 8) entries. step counts from 1.
 *   If grad_sumsq != NULL and max_norm > 0 the clip coefficient min(1, max_norm / (sqrt(*grad_sumsq) * grad_scale
 *   + 1e-6)) is applied on the fly (gradients in memory are not modified); grad_scale multiplies every gradient
 *   first (1/world_size for summed data-parallel gradients). */
int b200vit_adam_step(const void* table_dev, int num_tensors, long long max_size, const float* lrs_host,
                      const float* wds_host, int num_groups, float beta1, float beta2, float eps, int step,
                      int decoupled, const float* grad_sumsq, float max_norm, float grad_scale, void* stream);

/* ------------------------------------------------------------------------------------------------
 * One call per encoder block per direction: the launch sequence of TransformerEncoderBlock.forward
 * (models/vit.py:166-169, with MultiHeadSelfAttentionBlock.forward :92-98 and MLPBlock.forward :128-131 inside) and of
 * what autograd derives from it, issued from C instead of one Python -> C transition per kernel. The same kernels
 * run in the same order as through the single-kernel entry points above; every buffer is provided by the caller
 * (R = batch * tokens token rows, D = dim = heads * 64, M = mlp). Set struct_bytes = sizeof(the struct).
 * Dropout sites: key1 = after the GELU (vit.py:122), key2 = after fc2 (vit.py:125), attn = on the attention
 * probabilities (vit.py:86-90, p = 0 by default); mlp_p / attn_p = 0 outside training.
 * ------------------------------------------------------------------------------------------------ */
typedef struct b200vit_block_fwd_args {
    int struct_bytes;
    int batch, tokens, heads, dim, mlp;
    float eps_msa, eps_mlp;                 /* nn.LayerNorm eps of the two norms */
    float mlp_p, attn_p;
    unsigned long long key1_seed, key1_offset, key2_seed, key2_offset, attn_seed, attn_offset;
    const float* x;                         /* f32 [R, D] block input (residual stream) */
    const float *ln1_w, *ln1_b, *b_in, *b_out, *ln2_w, *ln2_b, *b1, *b2;         /* f32 parameters */
    const void *w_in16, *w_out16, *w1_16, *w2_16;                                /* bf16 copies of [3D,D] [D,D] [M,D] [D,M] */
    /* outputs; everything but x2 is what the backward call needs again */
    void *ln1, *qkv, *o, *ln2, *a, *h;      /* bf16 [R,D] [R,3D] [R,D] [R,D] [R,M] [R,M]; a (mask*gelu') may be NULL when no
                                               backward will follow */
    float *mean1, *rstd1, *lse, *x1, *mean2, *rstd2, *x2;   /* f32 [R] [R] [B,H,T] [R,D] [R] [R] [R,D] */
} b200vit_block_fwd_args;

int b200vit_encoder_block_fwd(const b200vit_block_fwd_args* args, void* stream);

typedef struct b200vit_block_bwd_args {
    int struct_bytes;
    int batch, tokens, heads, dim, mlp;
    float mlp_p, attn_p;
    unsigned long long key1_seed, key1_offset, key2_seed, key2_offset, attn_seed, attn_offset;
    const float* g;                         /* f32 [R, D] gradient w.r.t. the block output */
    int g16_ready;                          /* != 0: g16 already holds bf16(mask2 * g) and db2 its column sums (handed
                                               over by the next block's call, see gx16_out below) */
    int zero_accumulators;                  /* != 0: dw_in, dw_out, dw1, db1, dw2 are zero-filled first (the GEMM epilogues
                                               add into them); 0 accumulates on top of their contents */
    /* saved by the forward call */
    const float *x, *mean1, *rstd1, *lse, *x1, *mean2, *rstd2;
    const void *ln1, *qkv, *o, *ln2, *a, *h;
    const float *ln1_w, *ln2_w;
    const void *w_in16, *w_out16, *w1_16, *w2_16;
    /* scratch, caller-allocated: bf16 g16 [R,D], da [R,M], dout [R,D], dqkv [R,3D], g1_16 [R,D]; f32 g1 [R,D], delta [B,H,T] */
    void *g16, *da, *dout, *dqkv, *g1_16;
    float *g1, *delta;
    /* results */
    float* gx;                              /* f32 [R, D] gradient w.r.t. the block input */
    float *d_ln1_w, *d_ln1_b, *dw_in, *db_in, *dw_out, *db_out, *d_ln2_w, *d_ln2_b, *dw1, *db1, *dw2, *db2;
    /* optional: also emit bf16(consumer mask * gx) and its column sums for the block in front (its g16 / db2) */
    void* gx16_out;
    float* dbn_out;
    float consumer_p;
    unsigned long long consumer_seed, consumer_offset;
} b200vit_block_bwd_args;

int b200vit_encoder_block_bwd(const b200vit_block_bwd_args* args, void* stream);

/* ------------------------------------------------------------------------------------------------
 * LayerNorm folded into the GEMMs around it, for a forward that no backward will follow (inference):
 *     LN(x) W^T + b = rstd (x (gamma o W)^T - mean s) + c,   s[n] = sum_k gamma[k] W[n,k],   c[n] = sum_k beta[k] W[n,k] + b[n]
 * so the projection that consumes a LayerNorm (in_proj models/vit.py:93-97, fc1 :129-130) reads the un-normalised
 * residual stream as bf16 and applies the row statistics in its epilogue, provided the GEMM that produced the stream
 * (patch embedding :53-66, out-projection / fc2 with their residual adds :167-168) also wrote that bf16 copy and each
 * row's (sum, sum of squares). The separate nn.LayerNorm pass (:83, :115) disappears.
 * b200vit_gemm_ln = b200vit_gemm (op 0 only, no split-K) with four more epilogues:
 *   7 RES_STATS    epilogue 2, and also x16 bf16 [M, N] (ld ldx16) = C and rowstats[m][0..1] += (sum_n C, sum_n C^2)
 *   8 PATCH_STATS  epilogue 6 with the same extra outputs (same row mapping as C)
 *   9 LN_BF16      C bf16 = rstd[m] (acc - mean[m] ln_s[n]) + bias[n]   with mean / rstd from ln_stats[m] = (sum, sum of
 *                  squares) over ln_dim values and ln_eps; A = the producer's x16, B = the folded weight, bias = c
 *  10 LN_GELU      a = the same, then epilogue 1
 * rowstats must be zero-filled before the producing call (b200vit_class_token_rows_ln STORES its rows' entries).
 * b200vit_fold_ln_weight prepares B, s and c from the f32 master weight [rows, cols], gamma / beta [cols], bias [rows]. */
int b200vit_gemm_ln(int op, int epilogue, int M, int N, int K, const void* A, long long lda, const void* B, long long ldb,
                    void* C, long long ldc, void* C2, long long ldc2, const float* bias, const void* aux, long long ldaux,
                    float dropout_p, unsigned long long seed, unsigned long long offset, int row_period, void* x16,
                    long long ldx16, float* rowstats, const float* ln_stats, const float* ln_s, int ln_dim, float ln_eps,
                    void* stream);
int b200vit_fold_ln_weight(const float* w, const float* gamma, const float* beta, const float* bias, void* w16_bf16,
                           float* s, float* c, int rows, int cols, void* stream);
int b200vit_class_token_rows_ln(const float* class_token, const float* pos, float* x, void* x16_bf16, float* rowstats,
                                int batch, int tokens, int dim, float dropout_p, unsigned long long seed,
                                unsigned long long offset, void* stream);

/* Tensor-map cache statistics: the GEMM entry points keep the CUtensorMap of every (pointer, shape, leading
 * dimension, box) they have encoded; torch's allocator hands the same addresses out step after step, so a training
 * loop re-encodes next to nothing. Either pointer may be NULL. */
void b200vit_tmap_cache_stats(long long* hits, long long* misses);

/* ------------------------------------------------------------------------------------------------
 * Strict-precision ("fp32 mode") entry points. The reference computes in fp32 end to end (default-dtype nn modules,
 * models/vit.py:28-31, 83, 86-90, 115, 119-124, 216-222, run unchanged by going_modular/engine.py:50-66). In this
 * mode activations stay f32 between kernels and every contraction is still ONE b200vit_gemm launch on the tcgen05
 * tensor cores, fed split operands: x = hi + lo with hi = bf16(x), lo = bf16(x - hi), concatenated along the
 * contraction dimension as A' = [hi | lo | hi], B' = [hi | hi | lo], K' = 3K, so that the fp32 accumulator receives
 * a_hi b_hi + a_lo b_hi + a_hi b_lo (error <= ~2^-16 per product). Logits then agree with the reference to < 1e-3.
 * ------------------------------------------------------------------------------------------------ */

/* Write the three bf16 parts of src f32 [rows, cols] (leading dimension ld_src) to dst + k * part_stride (k = 0..2),
 * each [rows, cols] with leading dimension ld_dst. pattern 0 = (hi, lo, hi) for the A operand, 1 = (hi, hi, lo) for B.
 * Concatenate along columns with part_stride = cols, ld_dst = 3 * cols; along rows with part_stride = rows * ld_dst.
 * dropout_p > 0 multiplies the source by the (seed, offset) mask with element index source_row * cols + col (reading
 * the gradient of an nn.Dropout output in place, models/vit.py:43, 125); row_period > 0 reads source row
 * r + r / row_period + 1 (skips the class-token row of every image, models/vit.py:59). */
int b200vit_split_bf16x3(const float* src, long long ld_src, long long rows, int cols, void* dst_bf16, long long ld_dst,
                         long long part_stride, int pattern, float dropout_p, unsigned long long seed,
                         unsigned long long offset, int row_period, void* stream);

/* b200vit_im2col_patches with f32 output [B*(H/P)*(W/P), C*P*P] (models/vit.py:28-34, 53-56). */
int b200vit_im2col_patches_f32(const float* image, float* patches, int batch, int channels, int height, int width,
                               int patch, void* stream);

/* h = dropout(gelu_erf(a)) on f32 (nn.GELU + nn.Dropout, models/vit.py:121-122; erff, not the polynomial of the bf16
 * epilogue) and its backward da = dh * dropout_mask/(1-p) * gelu_erf'(a). Element index i of the mask. */
int b200vit_gelu_f32_fwd(const float* a, float* h, long long n, float dropout_p, unsigned long long seed,
                         unsigned long long offset, void* stream);
int b200vit_gelu_f32_bwd(const float* a, const float* dh, float* da, long long n, float dropout_p,
                         unsigned long long seed, unsigned long long offset, void* stream);

/* out[n] = sum_m mask[m, n] * x[m, n] for an f32 matrix (bias gradients, autograd of nn.Linear models/vit.py:119, 123
 * and of the MHA projections :86-90); same dropout / row_period conventions as b200vit_split_bf16x3. */
int b200vit_colsum_f32(const float* x, long long ld, long long rows, int cols, float* out, float dropout_p,
                       unsigned long long seed, unsigned long long offset, int row_period, void* stream);

/* f32 self-attention (the scaled_dot_product_attention inside nn.MultiheadAttention, models/vit.py:86-97) on CUDA
 * cores: qkv f32 [B, T, 3, H, 64] -> out f32 [B, T, H*64], lse f32 [B, H, T]; any T, head_dim 64, no dropout.
 * Backward: delta is an OUTPUT here (f32 [B, H, T], rowsum(dout * out)), dqkv f32 [B, T, 3, H, 64] is overwritten. */
int b200vit_attention_f32_fwd(const float* qkv, float* out, float* lse, int batch, int tokens, int heads, int head_dim,
                              float scale, void* stream);
int b200vit_attention_f32_bwd(const float* qkv, const float* out, const float* dout, const float* lse, float* delta,
                              float* dqkv, int batch, int tokens, int heads, int head_dim, float scale, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200VIT_H */
