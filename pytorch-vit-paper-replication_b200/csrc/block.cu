// One C call per encoder block per direction.
//
// A training step of ViT-B/16 is ~340 kernel launches. Issued one by one from Python (ctypes marshalling, torch
// allocations, autograd bookkeeping) they cost the host ~23 ms per step -- hidden behind 35 ms of GPU work at batch 256,
// but the whole step at batch 32 and below. These two entry points run the launch sequence of
// TransformerEncoderBlock.forward (reference models/vit.py:166-169: x1 = msa(x) + x ; x2 = mlp(x1) + x1, with
// MultiHeadSelfAttentionBlock :92-98 and MLPBlock :128-131 inside) and of its backward from C: the caller hands over
// every buffer (nothing is allocated here) in a plain struct, and the same kernels are launched in the same order as
// b200vit/functional.py does when it drives them one by one.
#include "common.cuh"
#include "launchers.h"

#include "../../include/b200vit.h"

namespace b200vit {

enum { OPN_NT = 0, OPN_NN = 1, OPN_TN = 2 };
enum { E_BF16 = 0, E_BIAS_GELU = 1, E_BIAS_RES = 2, E_DGELU = 3, E_ACCUM = 4 };

#define B200_TRY(expr)        \
    do {                      \
        const int _rc = (expr); \
        if (_rc) return _rc;  \
    } while (0)

static int block_fwd(const b200vit_block_fwd_args& a, cudaStream_t st) {
    B200_CHECK_ARG(a.struct_bytes == (int)sizeof(b200vit_block_fwd_args),
                   "encoder_block_fwd: struct size %d does not match this library (%d): header / binding out of date",
                   a.struct_bytes, (int)sizeof(b200vit_block_fwd_args));
    B200_CHECK_ARG(a.batch > 0 && a.tokens > 0 && a.heads > 0 && a.dim == a.heads * 64 && a.mlp > 0,
                   "encoder_block_fwd: bad geometry (batch %d tokens %d heads %d dim %d mlp %d)", a.batch, a.tokens, a.heads,
                   a.dim, a.mlp);
    B200_CHECK_ARG(a.x && a.ln1_w && a.ln1_b && a.w_in16 && a.b_in && a.w_out16 && a.b_out && a.ln2_w && a.ln2_b && a.w1_16 &&
                   a.b1 && a.w2_16 && a.b2, "encoder_block_fwd: null parameter pointer");
    B200_CHECK_ARG(a.ln1 && a.mean1 && a.rstd1 && a.qkv && a.o && a.lse && a.x1 && a.ln2 && a.mean2 && a.rstd2 && a.h && a.x2,
                   "encoder_block_fwd: null activation buffer");
    const long long R = (long long)a.batch * a.tokens;
    B200_CHECK_ARG(R < (1LL << 31), "encoder_block_fwd: too many token rows");
    const int r = (int)R, D = a.dim, M = a.mlp;
    const float scale = 0.125f;     // 1 / sqrt(64)
    // ---- x1 = out_proj(attention(in_proj(LN(x)))) + x ---------------------------------------------------------------
    B200_TRY(layernorm_fwd_launch(a.x, D, a.ln1_w, a.ln1_b, (bf16*)a.ln1, nullptr, a.mean1, a.rstd1, R, D, a.eps_msa, st));
    B200_TRY(gemm_launch(OPN_NT, E_BF16, r, 3 * D, D, a.ln1, D, a.w_in16, D, a.qkv, 3 * D, nullptr, 0, a.b_in, nullptr, 0, 0.f,
                         0, 0, 0, 1, st));
    B200_TRY(attention_fwd_dropout_launch((const bf16*)a.qkv, (bf16*)a.o, a.lse, a.batch, a.tokens, a.heads, 64, scale, a.attn_p,
                                          a.attn_seed, a.attn_offset, st));
    B200_TRY(gemm_launch(OPN_NT, E_BIAS_RES, r, D, D, a.o, D, a.w_out16, D, a.x1, D, nullptr, 0, a.b_out, a.x, D, 0.f, 0, 0, 0, 1,
                         st));
    // ---- x2 = fc2(dropout(gelu(fc1(LN(x1))))) dropped out + x1 --------------------------------------------------------
    B200_TRY(layernorm_fwd_launch(a.x1, D, a.ln2_w, a.ln2_b, (bf16*)a.ln2, nullptr, a.mean2, a.rstd2, R, D, a.eps_mlp, st));
    B200_TRY(gemm_launch(OPN_NT, E_BIAS_GELU, r, M, D, a.ln2, D, a.w1_16, D, a.a /* may be NULL: no backward */, M, a.h, M, a.b1,
                         nullptr, 0, a.mlp_p, a.key1_seed, a.key1_offset, 0, 1, st));
    B200_TRY(gemm_launch(OPN_NT, E_BIAS_RES, r, D, M, a.h, M, a.w2_16, M, a.x2, D, nullptr, 0, a.b2, a.x1, D, a.mlp_p, a.key2_seed,
                         a.key2_offset, 0, 1, st));
    return 0;
}

static int block_bwd(const b200vit_block_bwd_args& a, cudaStream_t st) {
    B200_CHECK_ARG(a.struct_bytes == (int)sizeof(b200vit_block_bwd_args),
                   "encoder_block_bwd: struct size %d does not match this library (%d): header / binding out of date",
                   a.struct_bytes, (int)sizeof(b200vit_block_bwd_args));
    B200_CHECK_ARG(a.batch > 0 && a.tokens > 0 && a.heads > 0 && a.dim == a.heads * 64 && a.mlp > 0,
                   "encoder_block_bwd: bad geometry");
    B200_CHECK_ARG(a.g && a.x && a.mean1 && a.rstd1 && a.ln1 && a.qkv && a.o && a.lse && a.x1 && a.mean2 && a.rstd2 && a.ln2 &&
                   a.a && a.h, "encoder_block_bwd: null saved activation");
    B200_CHECK_ARG(a.ln1_w && a.ln2_w && a.w_in16 && a.w_out16 && a.w1_16 && a.w2_16, "encoder_block_bwd: null weight pointer");
    B200_CHECK_ARG(a.g16 && a.da && a.dout && a.dqkv && a.g1 && a.g1_16 && a.delta && a.gx,
                   "encoder_block_bwd: null scratch / output buffer");
    B200_CHECK_ARG(a.d_ln1_w && a.d_ln1_b && a.dw_in && a.db_in && a.dw_out && a.db_out && a.d_ln2_w && a.d_ln2_b && a.dw1 &&
                   a.db1 && a.dw2 && a.db2, "encoder_block_bwd: null parameter-gradient buffer");
    const long long R = (long long)a.batch * a.tokens;
    const int r = (int)R, D = a.dim, M = a.mlp;
    const float scale = 0.125f;
    if (a.zero_accumulators) {      // the five buffers the GEMM epilogues ADD into (split-K weight gradients, fc1 bias)
        B200_CUDA(cudaMemsetAsync(a.dw2, 0, sizeof(float) * (size_t)D * M, st));
        B200_CUDA(cudaMemsetAsync(a.db1, 0, sizeof(float) * (size_t)M, st));
        B200_CUDA(cudaMemsetAsync(a.dw1, 0, sizeof(float) * (size_t)M * D, st));
        B200_CUDA(cudaMemsetAsync(a.dw_out, 0, sizeof(float) * (size_t)D * D, st));
        B200_CUDA(cudaMemsetAsync(a.dw_in, 0, sizeof(float) * (size_t)3 * D * D, st));
    }
    // ---- MLP half: g = dL/dx2 ------------------------------------------------------------------------------------------
    // g16 = bf16(mask2 * g), db2 = its column sums: handed over by the next block's LayerNorm backward, or made here
    if (!a.g16_ready)
        B200_TRY(dropout_cast_launch(a.g, (bf16*)a.g16, a.db2, R, D, a.mlp_p, a.key2_seed, a.key2_offset, st));
    B200_TRY(gemm_launch(OPN_TN, E_ACCUM, D, M, r, a.g16, D, a.h, M, a.dw2, M, nullptr, 0, nullptr, nullptr, 0, 0.f, 0, 0, 0, 0, st));
    // da = (g16 . W2) * mask1 * gelu'(pre-activation); its column sums accumulate into db1
    B200_TRY(gemm_launch(OPN_NN, E_DGELU, r, M, D, a.g16, D, a.w2_16, M, a.da, M, nullptr, 0, a.db1, a.a, M, 0.f, 0, 0, 0, 1, st));
    B200_TRY(gemm_launch(OPN_TN, E_ACCUM, M, D, r, a.da, M, a.ln2, D, a.dw1, D, nullptr, 0, nullptr, nullptr, 0, 0.f, 0, 0, 0, 0, st));
    bf16* dln = (bf16*)a.g16;      // g16 is dead after the two GEMMs above: reuse it for dL/dLN2
    B200_TRY(gemm_launch(OPN_NN, E_BF16, r, D, M, a.da, M, a.w1_16, D, dln, D, nullptr, 0, nullptr, nullptr, 0, 0.f, 0, 0, 0, 1, st));
    // g1 = g + dLN2/dx1 ; also bf16(g1) (out-proj has no dropout) and its column sums = db_out
    B200_TRY(layernorm_bwd_launch(dln, nullptr, a.x1, D, a.mean2, a.rstd2, a.ln2_w, a.g, a.g1, D, (bf16*)a.g1_16, a.d_ln2_w,
                                  a.d_ln2_b, a.db_out, R, D, 0.f, 0, 0, st));
    // ---- attention half: g1 = dL/dx1 -----------------------------------------------------------------------------------
    B200_TRY(gemm_launch(OPN_TN, E_ACCUM, D, D, r, a.g1_16, D, a.o, D, a.dw_out, D, nullptr, 0, nullptr, nullptr, 0, 0.f, 0, 0, 0, 0,
                         st));
    B200_TRY(gemm_launch(OPN_NN, E_BF16, r, D, D, a.g1_16, D, a.w_out16, D, a.dout, D, nullptr, 0, nullptr, nullptr, 0, 0.f, 0, 0, 0,
                         1, st));
    B200_TRY(attention_bwd_dropout_launch((const bf16*)a.qkv, (const bf16*)a.o, (const bf16*)a.dout, a.lse, a.delta, (bf16*)a.dqkv,
                                          a.db_in, a.batch, a.tokens, a.heads, 64, scale, a.attn_p, a.attn_seed, a.attn_offset,
                                          st));
    B200_TRY(gemm_launch(OPN_TN, E_ACCUM, 3 * D, D, r, a.dqkv, 3 * D, a.ln1, D, a.dw_in, D, nullptr, 0, nullptr, nullptr, 0, 0.f, 0,
                         0, 0, 0, st));
    bf16* dln1 = (bf16*)a.g1_16;    // dead after the two out-proj GEMMs
    B200_TRY(gemm_launch(OPN_NN, E_BF16, r, D, 3 * D, a.dqkv, 3 * D, a.w_in16, D, dln1, D, nullptr, 0, nullptr, nullptr, 0, 0.f, 0, 0,
                         0, 1, st));
    // gx = g1 + dLN1/dx ; optionally also what the PREVIOUS block's MLP backward needs first: bf16(its mask2 * gx) and
    // the column sums (that block's fc2 bias gradient)
    B200_TRY(layernorm_bwd_launch(dln1, nullptr, a.x, D, a.mean1, a.rstd1, a.ln1_w, a.g1, a.gx, D, (bf16*)a.gx16_out, a.d_ln1_w,
                                  a.d_ln1_b, a.gx16_out ? a.dbn_out : nullptr, R, D, a.gx16_out ? a.consumer_p : 0.f,
                                  a.consumer_seed, a.consumer_offset, st));
    return 0;
}

}  // namespace b200vit

extern "C" {

int b200vit_encoder_block_fwd(const b200vit_block_fwd_args* args, void* stream) {
    B200_CHECK_ARG(args != nullptr, "encoder_block_fwd: null argument struct");
    return b200vit::block_fwd(*args, (cudaStream_t)stream);
}

int b200vit_encoder_block_bwd(const b200vit_block_bwd_args* args, void* stream) {
    B200_CHECK_ARG(args != nullptr, "encoder_block_bwd: null argument struct");
    return b200vit::block_bwd(*args, (cudaStream_t)stream);
}

}  // extern "C"
