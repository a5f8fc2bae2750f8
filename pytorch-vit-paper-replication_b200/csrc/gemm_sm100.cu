// tcgen05 / TMEM / TMA bf16 GEMM with fused epilogues for the ViT training step (sm_100a).
//
// One kernel template covers every dense contraction of the path:
//   NT  C[M,N] = A[M,K]  . B[N,K]^T   forward Linear / Conv-as-GEMM  (reference models/vit.py:28,119,123
//                                      and nn.MultiheadAttention in/out projections, vit.py:86)
//   NN  C[M,N] = A[M,K]  . B[K,N]     input gradients  (what autograd derives for those Linears)
//   TN  C[M,N] = A[K,M]^T. B[K,N]     weight gradients (contraction over the token dimension)
// Operands are bf16 in global memory, accumulation is fp32 in tensor memory.
//
// Structure (persistent, warp specialised, one CTA per SM, 384 threads):
//   warp 0      TMA producer: global -> 128B-swizzled smem ring (4 stages of A 128x64 + B 256x64)
//   warp 1      MMA issuer: one thread issues tcgen05.mma 128x256x16, commits stages back to the ring
//   warp 2      TMEM allocator (512 columns = two 128x256 fp32 accumulators, double buffered)
//   warps 4-11  epilogue: tcgen05.ld accumulator -> registers -> fused math -> global
// K-major operands are loaded with one [64 x rows] box; MN-major operands (the "N"/"T" sides above)
// with 64x64 boxes, and the UMMA descriptors carry the major-ness, so no operand is ever transposed
// in memory.
#include "common.cuh"
#include "ptx_sm100.cuh"
#include "launchers.h"

#include <cudaTypedefs.h>
#include <atomic>
#include <mutex>
#include <stdlib.h>

namespace b200vit {

using namespace ptx;

constexpr int BM = 128;
constexpr int BN = 256;
constexpr int BK = 64;
constexpr int STAGES = 4;
constexpr int UMMA_K = 16;
constexpr int A_STAGE_BYTES = BM * BK * 2;   // 16 KB
constexpr int B_STAGE_BYTES = BN * BK * 2;   // 32 KB
constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
constexpr int NUM_EPI_WARPS = 8;
constexpr int GEMM_THREADS = 128 + NUM_EPI_WARPS * 32;
constexpr int TMEM_COLS = 512;
constexpr int EPI_STAGE_BYTES = 4096;   // per epilogue warp: one 32x32 chunk (fp32 or bf16), XOR-swizzled
constexpr int GEMM_SMEM_BYTES = STAGES * STAGE_BYTES + NUM_EPI_WARPS * EPI_STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;

// 2-CTA (cta_group::2) variant: a CTA pair owns a 256x256 output tile. Each CTA stages its own 128 rows of A and
// HALF of the B tile (128 of the 256 N-rows); the pair's tensor cores read both halves, so every CTA moves
// 32 KB per k-block instead of 48 KB (1.5x less L2->smem and smem traffic per FLOP) and the ring gets 6 stages.
constexpr int B2_STAGE_BYTES = (BN / 2) * BK * 2;   // 16 KB
constexpr int STAGE2_BYTES = A_STAGE_BYTES + B2_STAGE_BYTES;


enum GemmOp { OP_NT = 0, OP_NN = 1, OP_TN = 2 };
enum GemmEpi {
    EPI_BF16 = 0,       // C bf16 = acc (+ bias[n])
    EPI_BIAS_GELU = 1,  // a = acc + bias ; C2 bf16 = mask*gelu(a) ; C bf16 = mask*gelu'(a) (for backward; optional)  (fc1, vit.py:119-122)
    EPI_BIAS_RES = 2,   // C f32 = dropout(acc + bias) + aux_f32[m,n]                  (out-proj / fc2 + residual, vit.py:167-168)
    EPI_DGELU = 3,      // C bf16 = acc * aux_bf16[m,n], aux = mask*gelu'(a) saved by EPI_BIAS_GELU   (backward of GELU+Dropout)
    EPI_ACCUM = 4,      // C f32 += acc  (atomic; split-K weight gradients)
    EPI_F32 = 5,        // C f32 = acc (+ bias[n])
    EPI_PATCH = 6,      // C f32[m + m/P + 1, n] = dropout(acc + bias[n] + pos[m%P + 1, n])  (vit.py:53-66)
    // LayerNorm folded into the GEMMs around it (forward without a backward to follow; see "LayerNorm fold" below):
    EPI_RES_STATS = 7,  // EPI_BIAS_RES, and also x16 bf16 = C and rowstats[row] += (sum, sum of squares) of the C row
    EPI_PATCH_STATS = 8,// EPI_PATCH with the same two extra outputs
    EPI_LN_BF16 = 9,    // C bf16 = rstd[m] (acc - mean[m] s[n]) + bias[n]      (acc = bf16(x) . (gamma o W)^T)
    EPI_LN_GELU = 10,   // a = the same ; then as EPI_BIAS_GELU
    EPI_COUNT = 11
};
// LayerNorm fold. LN(x) W^T + b = rstd (x (gamma o W)^T - mean s) + c with s[n] = sum_k gamma[k] W[n,k] and
// c[n] = sum_k beta[k] W[n,k] + b[n]: the GEMM that CONSUMES a LayerNorm output (QKV projection, fc1: models/vit.py:93-97,
// 129-130) can read the un-normalised residual stream (as bf16) and apply the row statistics in its epilogue, if the
// GEMM that PRODUCED the stream (patch embedding, out-projection, fc2) also wrote that bf16 copy and accumulated each
// row's sum and sum of squares (three column tiles per row, fp32 atomics). The stand-alone LayerNorm kernel -- one
// read of the f32 stream and one bf16 write per LayerNorm -- disappears. Only used when no backward will follow
// (the weight gradients want LN(x) itself).
__host__ __device__ constexpr bool is_res(int e) { return e == EPI_BIAS_RES || e == EPI_RES_STATS; }
__host__ __device__ constexpr bool is_patch(int e) { return e == EPI_PATCH || e == EPI_PATCH_STATS; }
__host__ __device__ constexpr bool is_gelu(int e) { return e == EPI_BIAS_GELU || e == EPI_LN_GELU; }
__host__ __device__ constexpr bool is_bf16out(int e) { return e == EPI_BF16 || e == EPI_LN_BF16; }
__host__ __device__ constexpr bool has_stats(int e) { return e == EPI_RES_STATS || e == EPI_PATCH_STATS; }
__host__ __device__ constexpr bool is_lnin(int e) { return e == EPI_LN_BF16 || e == EPI_LN_GELU; }

struct GemmParams {
    int M, N, K;
    int tiles_m, tiles_n, splits, kblocks;
    void* C;  long long ldc;
    void* C2; long long ldc2;
    const float* bias;
    const void* aux; long long ldaux;
    DropoutParams drop;
    int row_period;
    int vec_ok;   // leading dimensions / N allow 16-byte vector access
    int* sched_counter;   // 2-CTA kernel: work items beyond the first are claimed from this counter (NULL: static stride)
    // LayerNorm fold (see the epilogue enum): producer side outputs / consumer side inputs
    bf16* x16; long long ldx16;     // has_stats: bf16 copy of C (same row mapping as C)
    float* rowstats;                // has_stats: [rows of C][2] += (sum, sum of squares); zero-filled by the caller
    const float* ln_stats;          // is_lnin: [M][2] (sum, sum of squares) of the A rows before rounding to bf16
    const float* ln_s;              // is_lnin: s[N]
    float ln_inv_dim, ln_eps;       // is_lnin: 1 / (LayerNorm width), eps
};
// The ALU-heavy epilogue (GELU value + derivative, with dropout) runs 16 epilogue warps (4 per scheduler, 64 columns each)
// and give up one pipeline stage for the extra staging tiles; the others run 8 warps and a 6-deep ring.
template <int EPI> struct Cfg2 {
    static constexpr int EW = is_gelu(EPI) ? 16 : 8;
    static constexpr int STAGES = EW == 16 ? 5 : 6;
    static constexpr int THREADS = 128 + EW * 32;
    static constexpr int SMEM = STAGES * STAGE2_BYTES + EW * EPI_STAGE_BYTES + 1024 + 256;
};

// -------------------------------------------------------------------------------------------------
// Epilogue. After tcgen05.ld a thread holds one ROW (TMEM lane) x 32 consecutive columns of the tile. Accessing
// global memory from that layout makes every warp instruction touch 32 different rows (32 LSU wavefronts per
// instruction) -- measured to dominate the K=768 GEMMs. So every global access of the epilogue is done in a
// row-major "cooperative" layout instead (lane -> (row = it*4 + lane/8, 16-byte chunk = lane%8) for fp32 tiles,
// (it*8 + lane/4, lane%4) for bf16 tiles: 4 resp. 8 full rows per instruction), and a 4 KB XOR-swizzled staging
// tile per epilogue warp in shared memory transposes between the two layouts (bank-conflict free both ways).
// Auxiliary operands (residual / position embedding / pre-activation) are fetched one chunk AHEAD in the
// cooperative layout, so their HBM latency overlaps the TMEM load and the math of the previous chunk.
// -------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t stg_off_f32(int row, int chunk) { return row * 128 + ((chunk ^ (row & 7)) << 4); }
__device__ __forceinline__ uint32_t stg_off_b16(int row, int chunk) { return row * 64 + ((chunk ^ ((row >> 1) & 3)) << 4); }
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
    uint4 r;
    asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr) : "memory");
    return r;
}

// registers of prefetched auxiliary operand per 32x32 chunk, cooperative layout
template <int EPI> struct AuxRegs {
    static constexpr bool F32 = is_res(EPI) || is_patch(EPI);
    static constexpr bool B16 = (EPI == EPI_DGELU);
    static constexpr int N = F32 ? 32 : (B16 ? 16 : 1);
};

template <int EPI>
__device__ __forceinline__ long long epi_out_row(const GemmParams& p, int r) {
    return is_patch(EPI) ? (long long)r + r / p.row_period + 1 : (long long)r;
}

// Coalesced fetch of the aux tile of one chunk into registers (cooperative layout); zero where out of range.
template <int EPI>
__device__ __forceinline__ void aux_prefetch(const GemmParams& p, int lane, int row_base, int col0,
                                             uint32_t (&a)[AuxRegs<EPI>::N]) {
    if (!(AuxRegs<EPI>::F32 || AuxRegs<EPI>::B16)) return;
    if (p.aux == nullptr || col0 >= p.N) return;
    const bool interior = p.vec_ok && col0 + 32 <= p.N && row_base + 32 <= p.M;
    if (interior && is_res(EPI)) {
        const float* ptr = reinterpret_cast<const float*>(p.aux) + (long long)(row_base + (lane >> 3)) * p.ldaux + col0 + (lane & 7) * 4;
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const uint4 q = *reinterpret_cast<const uint4*>(ptr + (long long)(it * 4) * p.ldaux);
            a[(4 * it) % AuxRegs<EPI>::N] = q.x; a[(4 * it + 1) % AuxRegs<EPI>::N] = q.y;
            a[(4 * it + 2) % AuxRegs<EPI>::N] = q.z; a[(4 * it + 3) % AuxRegs<EPI>::N] = q.w;
        }
        return;
    }
    if (interior && EPI == EPI_DGELU) {
        const bf16* ptr = reinterpret_cast<const bf16*>(p.aux) + (long long)(row_base + (lane >> 2)) * p.ldaux + col0 + (lane & 3) * 8;
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const uint4 q = *reinterpret_cast<const uint4*>(ptr + (long long)(it * 8) * p.ldaux);
            a[(4 * it) % AuxRegs<EPI>::N] = q.x; a[(4 * it + 1) % AuxRegs<EPI>::N] = q.y;
            a[(4 * it + 2) % AuxRegs<EPI>::N] = q.z; a[(4 * it + 3) % AuxRegs<EPI>::N] = q.w;
        }
        return;
    }
    if (AuxRegs<EPI>::F32) {
        const float* aux = reinterpret_cast<const float*>(p.aux);
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int r = it * 4 + (lane >> 3), ch = lane & 7;
            const int gr = row_base + r, col = col0 + ch * 4;
            uint4 q = make_uint4(0, 0, 0, 0);
            if (gr < p.M && col < p.N) {
                const long long ar = is_patch(EPI) ? (long long)(gr % p.row_period + 1) : (long long)gr;
                const float* src = aux + ar * p.ldaux + col;
                if (p.vec_ok && col + 4 <= p.N) {
                    q = *reinterpret_cast<const uint4*>(src);
                } else {
                    q.x = __float_as_uint(src[0]);
                    if (col + 1 < p.N) q.y = __float_as_uint(src[1]);
                    if (col + 2 < p.N) q.z = __float_as_uint(src[2]);
                    if (col + 3 < p.N) q.w = __float_as_uint(src[3]);
                }
            }
            a[4 * it] = q.x; a[4 * it + 1] = q.y; a[4 * it + 2] = q.z; a[4 * it + 3] = q.w;
        }
    } else {
        const bf16* aux = reinterpret_cast<const bf16*>(p.aux);
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const int r = it * 8 + (lane >> 2), ch = lane & 3;
            const int gr = row_base + r, col = col0 + ch * 8;
            uint4 q = make_uint4(0, 0, 0, 0);
            if (gr < p.M && col < p.N) {
                const bf16* src = aux + (long long)gr * p.ldaux + col;
                if (p.vec_ok && col + 8 <= p.N) {
                    q = *reinterpret_cast<const uint4*>(src);
                } else {
                    uint32_t w[4] = {0, 0, 0, 0};
                    for (int j = 0; j < 8 && col + j < p.N; ++j)
                        w[j >> 1] |= (uint32_t)__bfloat16_as_ushort(src[j]) << (16 * (j & 1));
                    q = make_uint4(w[0], w[1], w[2], w[3]);
                }
            }
            a[4 * it] = q.x; a[4 * it + 1] = q.y; a[4 * it + 2] = q.z; a[4 * it + 3] = q.w;
        }
    }
}

// Per-row state of the LayerNorm-fold epilogues: mean / rstd of this thread's A row (consumer side), running sum and
// sum of squares of this thread's C row over the warp's chunks (producer side).
struct LnRow {
    float mu, rstd, sum, sq;
};

template <int EPI>
__device__ __forceinline__ void epilogue_chunk(const GemmParams& p, uint32_t stg, int lane, int row_base, int col0,
                                               float (&v)[32], const uint32_t (&auxr)[AuxRegs<EPI>::N], LnRow& ln) {
    if (col0 >= p.N) return;                       // warp-uniform
    const int row = row_base + lane;
    const bool fullw = p.vec_ok && col0 + 32 <= p.N;   // warp-uniform: every 16-byte piece of the chunk is valid

    // ---- cooperative movers (whole warp) -----------------------------------------------------------
    // warp-uniform: the whole 32x32 chunk is inside the matrix and 16-byte accesses are legal -> no per-piece checks
    const bool interior = fullw && row_base + 32 <= p.M;
    auto store_f32_tile = [&](float* base, long long ld) {
        if (interior && !is_patch(EPI)) {
            float* ptr = base + (long long)(row_base + (lane >> 3)) * ld + col0 + (lane & 7) * 4;
#pragma unroll
            for (int it = 0; it < 8; ++it)
                *reinterpret_cast<uint4*>(ptr + (long long)(it * 4) * ld) = lds128(stg + stg_off_f32(it * 4 + (lane >> 3), lane & 7));
            return;
        }
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int r = it * 4 + (lane >> 3), ch = lane & 7;
            const int gr = row_base + r, col = col0 + ch * 4;
            const uint4 q = lds128(stg + stg_off_f32(r, ch));
            if (gr < p.M && col < p.N) {
                float* dst = base + epi_out_row<EPI>(p, gr) * ld + col;
                if (fullw || (p.vec_ok && col + 4 <= p.N)) {
                    *reinterpret_cast<uint4*>(dst) = q;
                } else {
                    dst[0] = __uint_as_float(q.x);
                    if (col + 1 < p.N) dst[1] = __uint_as_float(q.y);
                    if (col + 2 < p.N) dst[2] = __uint_as_float(q.z);
                    if (col + 3 < p.N) dst[3] = __uint_as_float(q.w);
                }
            }
        }
    };
    auto accum_f32_tile = [&](float* base, long long ld) {
        if (interior) {
            float* ptr = base + (long long)(row_base + (lane >> 3)) * ld + col0 + (lane & 7) * 4;
#pragma unroll
            for (int it = 0; it < 8; ++it) {
                const uint4 q = lds128(stg + stg_off_f32(it * 4 + (lane >> 3), lane & 7));
                asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};"
                             ::"l"(ptr + (long long)(it * 4) * ld), "f"(__uint_as_float(q.x)), "f"(__uint_as_float(q.y)),
                               "f"(__uint_as_float(q.z)), "f"(__uint_as_float(q.w)) : "memory");
            }
            return;
        }
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int r = it * 4 + (lane >> 3), ch = lane & 7;
            const int gr = row_base + r, col = col0 + ch * 4;
            const uint4 q = lds128(stg + stg_off_f32(r, ch));
            if (gr < p.M && col < p.N) {
                float* dst = base + (long long)gr * ld + col;
                if (fullw || (p.vec_ok && col + 4 <= p.N)) {
                    asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};"
                                 ::"l"(dst), "f"(__uint_as_float(q.x)), "f"(__uint_as_float(q.y)),
                                   "f"(__uint_as_float(q.z)), "f"(__uint_as_float(q.w)) : "memory");
                } else {
                    atomicAdd(dst, __uint_as_float(q.x));
                    if (col + 1 < p.N) atomicAdd(dst + 1, __uint_as_float(q.y));
                    if (col + 2 < p.N) atomicAdd(dst + 2, __uint_as_float(q.z));
                    if (col + 3 < p.N) atomicAdd(dst + 3, __uint_as_float(q.w));
                }
            }
        }
    };
    auto store_b16_tile = [&](bf16* base, long long ld, uint32_t soff = 0) {
        if (interior && !is_patch(EPI)) {       // (the patch epilogues remap rows: class-token rows are skipped)
            bf16* ptr = base + (long long)(row_base + (lane >> 2)) * ld + col0 + (lane & 3) * 8;
#pragma unroll
            for (int it = 0; it < 4; ++it)
                *reinterpret_cast<uint4*>(ptr + (long long)(it * 8) * ld) = lds128(stg + soff + stg_off_b16(it * 8 + (lane >> 2), lane & 3));
            return;
        }
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const int r = it * 8 + (lane >> 2), ch = lane & 3;
            const int gr = row_base + r, col = col0 + ch * 8;
            const uint4 q = lds128(stg + soff + stg_off_b16(r, ch));
            if (gr < p.M && col < p.N) {
                bf16* dst = base + epi_out_row<EPI>(p, gr) * ld + col;
                if (fullw || (p.vec_ok && col + 8 <= p.N)) {
                    *reinterpret_cast<uint4*>(dst) = q;
                } else {
                    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
                    for (int j = 0; j < 8 && col + j < p.N; ++j)
                        dst[j] = __ushort_as_bfloat16((unsigned short)(w[j >> 1] >> (16 * (j & 1))));
                }
            }
        }
    };
    auto put_row_f32 = [&](const float (&x)[32]) {
#pragma unroll
        for (int c = 0; c < 8; ++c)
            sts128(stg + stg_off_f32(lane, c), __float_as_uint(x[4 * c]), __float_as_uint(x[4 * c + 1]),
                   __float_as_uint(x[4 * c + 2]), __float_as_uint(x[4 * c + 3]));
    };
    auto put_row_b16 = [&](const float (&x)[32]) {
#pragma unroll
        for (int c = 0; c < 4; ++c)
            sts128(stg + stg_off_b16(lane, c), pack_bf16(x[8 * c], x[8 * c + 1]), pack_bf16(x[8 * c + 2], x[8 * c + 3]),
                   pack_bf16(x[8 * c + 4], x[8 * c + 5]), pack_bf16(x[8 * c + 6], x[8 * c + 7]));
    };
    // aux registers (cooperative layout) -> staging -> my row
    auto aux_to_stage = [&]() {
        if (AuxRegs<EPI>::F32) {
#pragma unroll
            for (int it = 0; it < 8; ++it)
                sts128(stg + stg_off_f32(it * 4 + (lane >> 3), lane & 7), auxr[(4 * it) % AuxRegs<EPI>::N],
                       auxr[(4 * it + 1) % AuxRegs<EPI>::N], auxr[(4 * it + 2) % AuxRegs<EPI>::N],
                       auxr[(4 * it + 3) % AuxRegs<EPI>::N]);
        } else if (AuxRegs<EPI>::B16) {
#pragma unroll
            for (int it = 0; it < 4; ++it)
                sts128(stg + stg_off_b16(it * 8 + (lane >> 2), lane & 3), auxr[(4 * it) % AuxRegs<EPI>::N],
                       auxr[(4 * it + 1) % AuxRegs<EPI>::N], auxr[(4 * it + 2) % AuxRegs<EPI>::N],
                       auxr[(4 * it + 3) % AuxRegs<EPI>::N]);
        }
        __syncwarp();
    };
    auto apply_dropout = [&](float (&x)[32]) {
        if (p.drop.p <= 0.f) return;
        const unsigned long long e0 = (unsigned long long)epi_out_row<EPI>(p, row) * (unsigned long long)p.N + (unsigned long long)col0;
        if ((p.N & 7) == 0) {
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                float m[8];
                dropout_mult8(p.drop, (e0 >> 3) + b, m);
#pragma unroll
                for (int i = 0; i < 8; ++i) x[8 * b + i] *= m[i];
            }
        } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) x[i] *= dropout_mult1(p.drop, e0 + i);
        }
    };

    // ---- folded LayerNorm of the A operand: acc -> rstd (acc - mean s[n]); the "+ c[n]" is the bias below --------
    if (is_lnin(EPI)) {
        const float a = ln.rstd, b = -ln.mu * ln.rstd;
        if (fullw) {
            const float4* s4 = reinterpret_cast<const float4*>(p.ln_s + col0);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float4 q = __ldg(s4 + i);
                v[4 * i] = fmaf(b, q.x, a * v[4 * i]); v[4 * i + 1] = fmaf(b, q.y, a * v[4 * i + 1]);
                v[4 * i + 2] = fmaf(b, q.z, a * v[4 * i + 2]); v[4 * i + 3] = fmaf(b, q.w, a * v[4 * i + 3]);
            }
        } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) if (col0 + i < p.N) v[i] = fmaf(b, __ldg(p.ln_s + col0 + i), a * v[i]);
        }
    }

    // ---- bias ---------------------------------------------------------------------------------------
    if (is_bf16out(EPI) || is_gelu(EPI) || is_res(EPI) || EPI == EPI_F32 || is_patch(EPI)) {
        if (p.bias != nullptr) {
            if (fullw) {
                const float4* b4 = reinterpret_cast<const float4*>(p.bias + col0);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const float4 b = __ldg(b4 + i);
                    v[4 * i] += b.x; v[4 * i + 1] += b.y; v[4 * i + 2] += b.z; v[4 * i + 3] += b.w;
                }
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) if (col0 + i < p.N) v[i] += __ldg(p.bias + col0 + i);
            }
        }
    }

    if (is_bf16out(EPI)) {
        put_row_b16(v);
        __syncwarp();
        store_b16_tile(reinterpret_cast<bf16*>(p.C), p.ldc);
        __syncwarp();
    } else if (EPI == EPI_F32) {
        put_row_f32(v);
        __syncwarp();
        store_f32_tile(reinterpret_cast<float*>(p.C), p.ldc);
        __syncwarp();
    } else if (is_gelu(EPI)) {
        // Per group of 8 columns: h = scale*gelu(a) (the hidden activation) and gd = scale*gelu'(a) (saved for the
        // backward epilogue), dropout keep-mask ANDed onto the packed bf16 pairs of both. A bf16 tile needs only
        // 2 KB, so both outputs are staged side by side in the warp's 4 KB tile and stored after one warp sync.
        const bool want_gd = p.C != nullptr;
        const unsigned long long e0 = (unsigned long long)epi_out_row<EPI>(p, row) * (unsigned long long)p.N + (unsigned long long)col0;
        const bool blk8 = (p.N & 7) == 0;
        const GeluConst2 gk = make_gelu_const2(p.drop.scale);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            // two columns per instruction (FFMA2): value and derivative of four column pairs
            uint32_t hp[4], gp[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                f32x2 h2, g2;
                gelu_value_and_grad2(v[8 * c + 2 * j], v[8 * c + 2 * j + 1], gk, h2, g2);
                hp[j] = pack_bf16_2(h2);
                gp[j] = pack_bf16_2(g2);
            }
            uint32_t mk[4] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu};
            if (p.drop.p > 0.f) {
                if (blk8) {
                    uint32_t ws[4];
                    dropout_words4(p.drop, (e0 >> 3) + c, ws);
#pragma unroll
                    for (uint32_t j = 0; j < 4; ++j) mk[j] = dropout_mask2(p.drop, ws[j]);
                } else {
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        mk[j] = (dropout_mult1(p.drop, e0 + 8 * c + 2 * j) != 0.f ? 0x0000FFFFu : 0u) |
                                (dropout_mult1(p.drop, e0 + 8 * c + 2 * j + 1) != 0.f ? 0xFFFF0000u : 0u);
                }
            }
            if (want_gd)
                sts128(stg + stg_off_b16(lane, c), gp[0] & mk[0], gp[1] & mk[1], gp[2] & mk[2], gp[3] & mk[3]);
            sts128(stg + 2048 + stg_off_b16(lane, c), hp[0] & mk[0], hp[1] & mk[1], hp[2] & mk[2], hp[3] & mk[3]);
        }
        __syncwarp();
        if (want_gd) store_b16_tile(reinterpret_cast<bf16*>(p.C), p.ldc);
        store_b16_tile(reinterpret_cast<bf16*>(p.C2), p.ldc2, 2048);
        __syncwarp();
    } else if (is_res(EPI) || is_patch(EPI)) {
        if (is_res(EPI)) apply_dropout(v);
        if (is_patch(EPI) || p.aux != nullptr) {
            aux_to_stage();
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const uint4 q = lds128(stg + stg_off_f32(lane, c));
                v[4 * c] += __uint_as_float(q.x); v[4 * c + 1] += __uint_as_float(q.y);
                v[4 * c + 2] += __uint_as_float(q.z); v[4 * c + 3] += __uint_as_float(q.w);
            }
            __syncwarp();
        }
        if (is_patch(EPI)) apply_dropout(v);   // position embedding is added BEFORE dropout (vit.py:63-66)
        if (has_stats(EPI)) {                  // row statistics of what is stored (fp32, before the bf16 rounding)
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const float x = (fullw || col0 + i < p.N) ? v[i] : 0.f;
                ln.sum += x;
                ln.sq = fmaf(x, x, ln.sq);
            }
        }
        put_row_f32(v);
        __syncwarp();
        store_f32_tile(reinterpret_cast<float*>(p.C), p.ldc);
        __syncwarp();
        if (has_stats(EPI)) {                  // the same tile once more as bf16: the next GEMM's A operand
            put_row_b16(v);
            __syncwarp();
            store_b16_tile(p.x16, p.ldx16);
            __syncwarp();
        }
    } else if (EPI == EPI_DGELU) {
        aux_to_stage();
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const uint4 q = lds128(stg + stg_off_b16(lane, c));
            const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 f = unpack_bf16(w[j]);
                v[8 * c + 2 * j] *= f.x;
                v[8 * c + 2 * j + 1] *= f.y;
            }
        }
        __syncwarp();
        put_row_b16(v);
        __syncwarp();
        store_b16_tile(reinterpret_cast<bf16*>(p.C), p.ldc);
        __syncwarp();
        if (p.bias != nullptr) {
            // optional fused bias gradient of the layer in front of the GELU: column sums of this tile's rows are
            // added to bias[] (an OUTPUT for this epilogue; the caller zeroes it), sparing a separate pass over C
            if (row >= p.M) {
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = 0.f;
            }
            const float cs = warp_colsum32(v, lane);
            if (col0 + lane < p.N)
                asm volatile("red.global.add.f32 [%0], %1;" ::"l"(const_cast<float*>(p.bias) + col0 + lane), "f"(cs) : "memory");
        }
    } else if (EPI == EPI_ACCUM) {
        put_row_f32(v);
        __syncwarp();
        accum_f32_tile(reinterpret_cast<float*>(p.C), p.ldc);
        __syncwarp();
    }
}

// -------------------------------------------------------------------------------------------------
// The kernel
// -------------------------------------------------------------------------------------------------
template <bool A_MN, bool B_MN, int EPI>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const GemmParams p) {
    extern __shared__ uint8_t smem_raw[];
    pdl_launch_dependents();      // the next kernel on the stream may be scheduled as SMs free up
    // 128B swizzle atoms need 1024-byte aligned stage bases.
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t smem_a = smem_base;
    const uint32_t smem_b = smem_base + STAGES * A_STAGE_BYTES;
    const uint32_t smem_stage = smem_base + STAGES * STAGE_BYTES;
    const uint32_t bar_base = smem_stage + NUM_EPI_WARPS * EPI_STAGE_BYTES;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
    auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * STAGES + s); };
    auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * STAGES + 2 + s); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);
    uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA);
        prefetch_tmap(&tmB);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), NUM_EPI_WARPS); }
        fence_barrier_init();
    }
    if (warp == 2) {
        tmem_alloc<1>(tmem_slot, TMEM_COLS);
        tmem_relinquish<1>();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;
    pdl_wait();                   // everything above ran under the previous kernel's tail; global memory from here on

    const int total_work = p.tiles_m * p.tiles_n * p.splits;

    auto decode = [&](int w, int& m0, int& n0, int& kb0, int& kb1) {
        const int tile = w % (p.tiles_m * p.tiles_n);
        const int split = w / (p.tiles_m * p.tiles_n);
        // n fastest: CTAs running together share the same A rows (L2 reuse), weights are small.
        n0 = (tile % p.tiles_n) * BN;
        m0 = (tile / p.tiles_n) * BM;
        kb0 = (int)(((long long)split * p.kblocks) / p.splits);
        kb1 = (int)(((long long)(split + 1) * p.kblocks) / p.splits);
    };

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int w = blockIdx.x; w < total_work; w += gridDim.x) {
                int m0, n0, kb0, kb1;
                decode(w, m0, n0, kb0, kb1);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(empty_bar(stage), phase ^ 1u);
                    mbar_arrive_expect_tx(full_bar(stage), STAGE_BYTES);
                    const uint32_t sa = smem_a + stage * A_STAGE_BYTES;
                    const uint32_t sb = smem_b + stage * B_STAGE_BYTES;
                    const int k0 = kb * BK;
                    if (!A_MN) {
                        tma_load_2d(sa, &tmA, full_bar(stage), k0, m0);
                    } else {
#pragma unroll
                        for (int i = 0; i < BM / 64; ++i)
                            tma_load_2d(sa + i * (BK * 128), &tmA, full_bar(stage), m0 + i * 64, k0);
                    }
                    if (!B_MN) {
                        tma_load_2d(sb, &tmB, full_bar(stage), k0, n0);
                    } else {
#pragma unroll
                        for (int i = 0; i < BN / 64; ++i)
                            tma_load_2d(sb + i * (BK * 128), &tmB, full_bar(stage), n0 + i * 64, k0);
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===================== MMA issuer: all 32 lanes run the loop, one elected lane issues =====================
        {
            constexpr uint32_t idesc = make_idesc_bf16(BM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
            // K-major:  8-row groups 1024 B apart (SBO), one 128B atom along K (LBO unused).
            // MN-major: 8-k-row groups 1024 B apart (SBO), 64-wide MN atoms BK*128 B apart (LBO).
            constexpr uint32_t a_lbo = A_MN ? BK * 128 : 0, b_lbo = B_MN ? BK * 128 : 0;
            constexpr uint32_t a_kstep = A_MN ? UMMA_K * 128 : UMMA_K * 2;
            constexpr uint32_t b_kstep = B_MN ? UMMA_K * 128 : UMMA_K * 2;
            int stage = 0; uint32_t phase = 0;
            int iter = 0;
            for (int w = blockIdx.x; w < total_work; w += gridDim.x, ++iter) {
                int m0, n0, kb0, kb1;
                decode(w, m0, n0, kb0, kb1);
                const int acc = iter & 1;
                const uint32_t acc_phase = (iter >> 1) & 1;
                mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + acc * BN;
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(full_bar(stage), phase);
                    tc_fence_after();
                    const uint32_t sa = smem_a + stage * A_STAGE_BYTES;
                    const uint32_t sb = smem_b + stage * B_STAGE_BYTES;
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k) {
                        const uint64_t da = make_smem_desc_sw128(sa + k * a_kstep, 1024, a_lbo);
                        const uint64_t db = make_smem_desc_sw128(sb + k * b_kstep, 1024, b_lbo);
                        umma_bf16_ss_elect(tmem_d, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
                    }
                    umma_commit_elect(empty_bar(stage));          // frees the smem stage when the MMAs retire
                    if (kb == kb1 - 1) umma_commit_elect(tfull_bar(acc));  // accumulator ready
                    if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                }
            }
        }
        __syncwarp();
    } else if (warp >= 4) {
        // ===================== Epilogue =====================
        const int ew = warp - 4;
        const int quarter = warp & 3;        // TMEM lane quarter this warp may access
        const int half = ew >> 2;            // which 128-column half of the tile
        int iter = 0;
        for (int w = blockIdx.x; w < total_work; w += gridDim.x, ++iter) {
            int m0, n0, kb0, kb1;
            decode(w, m0, n0, kb0, kb1);
            const int acc = iter & 1;
            const uint32_t acc_phase = (iter >> 1) & 1;
            const int row_base = m0 + quarter * 32;
            const uint32_t stg = smem_stage + ew * EPI_STAGE_BYTES;
            uint32_t aux_next[AuxRegs<EPI>::N], aux_cur[AuxRegs<EPI>::N];
            LnRow ln = {0.f, 1.f, 0.f, 0.f};
            if (is_lnin(EPI)) {     // mean / rstd of this thread's row from the producer's (sum, sum of squares)
                const int lr = row_base + lane;
                if (lr < p.M) {
                    const float2 st = __ldg(reinterpret_cast<const float2*>(p.ln_stats) + lr);
                    ln.mu = st.x * p.ln_inv_dim;
                    ln.rstd = rsqrtf(fmaxf(fmaf(-ln.mu, ln.mu, st.y * p.ln_inv_dim), 0.f) + p.ln_eps);
                }
            }
            aux_prefetch<EPI>(p, lane, row_base, n0 + half * 128, aux_next);   // before the accumulator is even ready
            mbar_wait(tfull_bar(acc), acc_phase);
            tc_fence_after();
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
                const int col_in_tile = half * 128 + c * 32;
                const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BN + col_in_tile);
                uint32_t r[32];
                tmem_ld_32x32(taddr, r);
#pragma unroll
                for (int i = 0; i < AuxRegs<EPI>::N; ++i) aux_cur[i] = aux_next[i];
                if (c < 3) aux_prefetch<EPI>(p, lane, row_base, n0 + col_in_tile + 32, aux_next);
                tmem_ld_wait();
                if (c == 3) {
                    // all TMEM reads of this warp are done: hand the accumulator back to the MMA warp
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(tempty_bar(acc));
                }
                float v[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
                epilogue_chunk<EPI>(p, stg, lane, row_base, n0 + col_in_tile, v, aux_cur, ln);
            }
            if (has_stats(EPI)) {   // this warp's share of its rows' sums (other column tiles add theirs)
                const int lr = row_base + lane;
                if (lr < p.M) {
                    float* dst = p.rowstats + 2 * epi_out_row<EPI>(p, lr);
                    atomicAdd(dst, ln.sum);
                    atomicAdd(dst + 1, ln.sq);
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc<1>(tmem_base, TMEM_COLS);
    }
}


// -------------------------------------------------------------------------------------------------
// The 2-CTA kernel (cluster of 2, cta_group::2 MMA, UMMA 256x256x16)
// -------------------------------------------------------------------------------------------------
template <bool A_MN, bool B_MN, int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(Cfg2<EPI>::THREADS, 1)
gemm_tcgen05_2cta_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                         const GemmParams p) {
    constexpr int STAGES2 = Cfg2<EPI>::STAGES;
    constexpr int EW = Cfg2<EPI>::EW;
    constexpr int CHUNKS = 8 / (EW / 4);     // 32-column chunks per epilogue warp
    extern __shared__ uint8_t smem_raw[];
    pdl_launch_dependents();      // the next kernel on the stream may be scheduled as SMs free up
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t smem_a = smem_base;
    const uint32_t smem_b = smem_base + STAGES2 * A_STAGE_BYTES;
    const uint32_t smem_stage = smem_base + STAGES2 * STAGE2_BYTES;
    const uint32_t bar_base = smem_stage + EW * EPI_STAGE_BYTES;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };                       // used in the leader CTA
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES2 + s); };          // one per CTA
    auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * STAGES2 + s); };      // one per CTA
    auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * STAGES2 + 2 + s); }; // used in the leader CTA
    const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES2 + 4);
    uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));
    // dynamic work distribution (see the scheduler warp below): a 2-deep ring of work indices per CTA
    auto sfull_bar = [&](int s) { return bar_base + 8u * (2 * STAGES2 + 5 + s); };      // one per CTA
    auto sempty_bar = [&](int s) { return bar_base + 8u * (2 * STAGES2 + 7 + s); };     // used in the leader CTA
    auto sched_slot = [&](int s) { return bar_base + 8u * (2 * STAGES2 + 9) + 4u * s; };

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    const bool dyn = p.sched_counter != nullptr;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA);
        prefetch_tmap(&tmB);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < STAGES2; ++s) { mbar_init(full_bar(s), 2); mbar_init(empty_bar(s), 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), 2 * EW); }
        // readers of a published work index: the TMA lane and the epilogue warps of both CTAs + the MMA warp
        for (int s = 0; s < 2; ++s) { mbar_init(sfull_bar(s), 1); mbar_init(sempty_bar(s), 2 * (1 + EW) + 1); }
        fence_barrier_init();
    }
    if (warp == 2) {
        tmem_alloc<2>(tmem_slot, TMEM_COLS);
        tmem_relinquish<2>();
    }
    tc_fence_before();
    cluster_sync_all();     // barriers of both CTAs are initialised before any remote arrive / multicast commit
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;
    pdl_wait();                   // everything above ran under the previous kernel's tail; global memory from here on

    const int tiles_m2 = (p.M + 2 * BM - 1) / (2 * BM);
    const int total_work = tiles_m2 * p.tiles_n * p.splits;
    const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;

    auto decode = [&](int w, int& m0, int& n0, int& kb0, int& kb1) {
        const int tile = w % (tiles_m2 * p.tiles_n);
        const int split = w / (tiles_m2 * p.tiles_n);
        n0 = (tile % p.tiles_n) * BN;
        m0 = (tile / p.tiles_n) * (2 * BM) + (int)rank * BM;     // this CTA's 128 rows of the pair tile
        kb0 = (int)(((long long)split * p.kblocks) / p.splits);
        kb1 = (int)(((long long)(split + 1) * p.kblocks) / p.splits);
    };

    // The it-th work item of this cluster. Static: cluster_id + it * num_clusters. Dynamic: whatever the scheduler
    // warp published in ring slot it & 1 (-1 = no more work); every lane of the calling warp reads it, one arrives.
    auto work_item = [&](int it, bool whole_warp) -> int {
        if (!dyn) {
            const int w = cluster_id + it * num_clusters;
            return w < total_work ? w : -1;
        }
        const int s = it & 1;
        mbar_wait_acq_cluster(sfull_bar(s), (uint32_t)(it >> 1) & 1u);
        const int w = (int)ld_shared_u32(sched_slot(s));
        if (whole_warp) __syncwarp();
        // the slot may be overwritten once this arrive lands: make the arrive wait for the load above through a
        // register dependency (w is never INT_MAX) instead of paying for a cluster-scope release in the epilogue warps
        if (!whole_warp || lane == 0) mbar_arrive_cluster(sempty_bar(s) + (w == 0x7FFFFFFF ? 8u : 0u), 0);
        return w;
    };

    if (warp == 0) {
        // ===================== TMA producer (both CTAs) =====================
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int it = 0, w; (w = work_item(it, false)) >= 0; ++it) {
                int m0, n0, kb0, kb1;
                decode(w, m0, n0, kb0, kb1);
                const int nb0 = n0 + (int)rank * (BN / 2);           // this CTA's half of the B tile
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(empty_bar(stage), phase ^ 1u);
                    // all transaction bytes of the pair are counted on the LEADER's barrier
                    const uint32_t fb = full_bar(stage) & 0xFEFFFFFFu;
                    if (leader) mbar_arrive_expect_tx(full_bar(stage), 2 * STAGE2_BYTES);
                    const uint32_t sa = smem_a + stage * A_STAGE_BYTES;
                    const uint32_t sb = smem_b + stage * B2_STAGE_BYTES;
                    const int k0 = kb * BK;
                    if (!A_MN) {
                        tma_load_2d_2sm(sa, &tmA, fb, k0, m0);
                    } else {
#pragma unroll
                        for (int i = 0; i < BM / 64; ++i)
                            tma_load_2d_2sm(sa + i * (BK * 128), &tmA, fb, m0 + i * 64, k0);
                    }
                    if (!B_MN) {
                        tma_load_2d_2sm(sb, &tmB, fb, k0, nb0);
                    } else {
#pragma unroll
                        for (int i = 0; i < (BN / 2) / 64; ++i)
                            tma_load_2d_2sm(sb + i * (BK * 128), &tmB, fb, nb0 + i * 64, k0);
                    }
                    if (!leader) mbar_arrive_cluster(full_bar(stage), 0);
                    if (++stage == STAGES2) { stage = 0; phase ^= 1u; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===================== MMA issuer (leader CTA only): all 32 lanes run the loop, one elected lane issues ==========
        if (leader) {
            constexpr uint32_t idesc = make_idesc_bf16(2 * BM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
            constexpr uint32_t a_lbo = A_MN ? BK * 128 : 0, b_lbo = B_MN ? BK * 128 : 0;
            constexpr uint32_t a_kstep = A_MN ? UMMA_K * 128 : UMMA_K * 2;
            constexpr uint32_t b_kstep = B_MN ? UMMA_K * 128 : UMMA_K * 2;
            int stage = 0; uint32_t phase = 0;
            for (int iter = 0, w; (w = work_item(iter, true)) >= 0; ++iter) {
                int m0, n0, kb0, kb1;
                decode(w, m0, n0, kb0, kb1);
                const int acc = iter & 1;
                const uint32_t acc_phase = (iter >> 1) & 1;
                mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + acc * BN;
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(full_bar(stage), phase);
                    tc_fence_after();
                    const uint32_t sa = smem_a + stage * A_STAGE_BYTES;
                    const uint32_t sb = smem_b + stage * B2_STAGE_BYTES;
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k) {
                        const uint64_t da = make_smem_desc_sw128(sa + k * a_kstep, 1024, a_lbo);
                        const uint64_t db = make_smem_desc_sw128(sb + k * b_kstep, 1024, b_lbo);
                        umma_bf16_ss_2sm_elect(tmem_d, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
                    }
                    umma_commit_2sm_elect(empty_bar(stage), 0x3);                      // frees the stage in BOTH CTAs
                    if (kb == kb1 - 1) umma_commit_2sm_elect(tfull_bar(acc), 0x3);     // accumulator ready in BOTH CTAs
                    if (++stage == STAGES2) { stage = 0; phase ^= 1u; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 3) {
        // ===================== Work scheduler (leader CTA, dynamic distribution only) =====================
        // The first item of a cluster is its own index; every further one is claimed from a global counter, so a
        // cluster that starts late (its SMs were held by another stream's kernel, e.g. a gradient all-reduce) simply
        // claims less instead of holding the whole GEMM back by its share. The index is written into both CTAs' rings
        // and signalled with cluster-scope release arrives. Every cluster makes exactly one claim that fails; the
        // one that sees the last of them puts the counter back to zero for the next launch.
        if (dyn && leader && lane == 0) {
            const int claims_ok = total_work > num_clusters ? total_work - num_clusters : 0;
            auto claim = [&]() -> int {
                const int v = atomicAdd(p.sched_counter, 1);
                if (v == claims_ok + num_clusters - 1) atomicExch(p.sched_counter, 0);
                return num_clusters + v;
            };
            int w = cluster_id;
            if (w >= total_work) claim();            // (keeps the count of failing claims at one per cluster)
            for (int it = 0;; ++it) {
                if (w >= total_work) w = -1;
                const int s = it & 1;
                mbar_wait(sempty_bar(s), ((uint32_t)(it >> 1) & 1u) ^ 1u);
                st_shared_cluster_u32(sched_slot(s), 0, (uint32_t)w);
                st_shared_cluster_u32(sched_slot(s), 1, (uint32_t)w);
                mbar_arrive_cluster_release(sfull_bar(s), 0);
                mbar_arrive_cluster_release(sfull_bar(s), 1);
                if (w < 0) break;
                w = claim();
            }
        }
        __syncwarp();
    } else if (warp >= 4) {
        // ===================== Epilogue (both CTAs, own 128 rows) =====================
        const int ew = warp - 4;
        const int quarter = warp & 3;
        const int colgrp = ew >> 2;          // which (256 / (EW/4))-column slice of the tile
        for (int iter = 0, w; (w = work_item(iter, true)) >= 0; ++iter) {
            int m0, n0, kb0, kb1;
            decode(w, m0, n0, kb0, kb1);
            const int acc = iter & 1;
            const uint32_t acc_phase = (iter >> 1) & 1;
            const int row_base = m0 + quarter * 32;
            const uint32_t stg = smem_stage + ew * EPI_STAGE_BYTES;
            uint32_t aux_next[AuxRegs<EPI>::N], aux_cur[AuxRegs<EPI>::N];
            LnRow ln = {0.f, 1.f, 0.f, 0.f};
            if (is_lnin(EPI)) {     // mean / rstd of this thread's row from the producer's (sum, sum of squares)
                const int lr = row_base + lane;
                if (lr < p.M) {
                    const float2 st = __ldg(reinterpret_cast<const float2*>(p.ln_stats) + lr);
                    ln.mu = st.x * p.ln_inv_dim;
                    ln.rstd = rsqrtf(fmaxf(fmaf(-ln.mu, ln.mu, st.y * p.ln_inv_dim), 0.f) + p.ln_eps);
                }
            }
            aux_prefetch<EPI>(p, lane, row_base, n0 + colgrp * (CHUNKS * 32), aux_next);   // before the accumulator is even ready
            mbar_wait(tfull_bar(acc), acc_phase);
            tc_fence_after();
#pragma unroll 1
            for (int c = 0; c < CHUNKS; ++c) {
                const int col_in_tile = colgrp * (CHUNKS * 32) + c * 32;
                const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BN + col_in_tile);
                uint32_t r[32];
                tmem_ld_32x32(taddr, r);
#pragma unroll
                for (int i = 0; i < AuxRegs<EPI>::N; ++i) aux_cur[i] = aux_next[i];
                if (c < CHUNKS - 1) aux_prefetch<EPI>(p, lane, row_base, n0 + col_in_tile + 32, aux_next);
                tmem_ld_wait();
                if (c == CHUNKS - 1) {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(tempty_bar(acc), 0);    // the MMA issuer lives in CTA 0
                }
                float v[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
                epilogue_chunk<EPI>(p, stg, lane, row_base, n0 + col_in_tile, v, aux_cur, ln);
            }
            if (has_stats(EPI)) {   // this warp's share of its rows' sums (other column tiles add theirs)
                const int lr = row_base + lane;
                if (lr < p.M) {
                    float* dst = p.rowstats + 2 * epi_out_row<EPI>(p, lr);
                    atomicAdd(dst, ln.sum);
                    atomicAdd(dst + 1, ln.sq);
                }
            }
        }
    }

    tc_fence_before();
    cluster_sync_all();     // nobody leaves (or frees TMEM) while the peer may still read its smem / signal its barriers
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc<2>(tmem_base, TMEM_COLS);
    }
}

// -------------------------------------------------------------------------------------------------
// Host side
// -------------------------------------------------------------------------------------------------
static PFN_cuTensorMapEncodeTiled get_encode_fn() {
    static PFN_cuTensorMapEncodeTiled fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled>(ptr);
    });
    return fn;
}

// 2-D bf16 tensor map: inner dimension `inner` elements (contiguous), `outer` rows of `ld` elements.
static int encode_tmap_bf16(CUtensorMap* tm, const void* base, uint64_t inner, uint64_t outer, uint64_t ld,
                            uint32_t box_inner, uint32_t box_outer) {
    PFN_cuTensorMapEncodeTiled enc = get_encode_fn();
    if (!enc) return set_error(-2, "cuTensorMapEncodeTiled not available (no CUDA driver?)");
    cuuint64_t dims[2] = {inner, outer};
    cuuint64_t strides[1] = {ld * 2};
    cuuint32_t box[2] = {box_inner, box_outer};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return set_error(-3, "cuTensorMapEncodeTiled failed (%d): base=%p inner=%llu outer=%llu ld=%llu", (int)r,
                         base, (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)ld);
    return 0;
}

// Cache of encoded tensor maps. A map depends only on (address, extents, pitch, box) -- not on the memory's contents
// -- and torch's caching allocator returns the same addresses step after step, so the ~290 encodes of a training step
// are almost all repeats. Direct-mapped, 1024 entries, one mutex (calls come from the Python main thread and from
// autograd's backward thread).
struct TmapKey {
    const void* base; uint64_t inner, outer, ld; uint32_t box_inner, box_outer;
    bool operator==(const TmapKey& o) const {
        return base == o.base && inner == o.inner && outer == o.outer && ld == o.ld && box_inner == o.box_inner &&
               box_outer == o.box_outer;
    }
};
struct TmapEntry { TmapKey key; CUtensorMap map; bool valid; };
constexpr int TMAP_CACHE_SIZE = 1024;
static TmapEntry g_tmap_cache[TMAP_CACHE_SIZE];
static std::mutex g_tmap_mu;
static long long g_tmap_hits = 0, g_tmap_misses = 0;

void tmap_cache_stats(long long* hits, long long* misses) {
    std::lock_guard<std::mutex> lock(g_tmap_mu);
    if (hits) *hits = g_tmap_hits;
    if (misses) *misses = g_tmap_misses;
}

static int make_tmap_bf16(CUtensorMap* tm, const void* base, uint64_t inner, uint64_t outer, uint64_t ld,
                          uint32_t box_inner, uint32_t box_outer) {
    const TmapKey key{base, inner, outer, ld, box_inner, box_outer};
    uint64_t h = (uint64_t)(uintptr_t)base * 0x9E3779B97F4A7C15ull;
    h ^= (inner * 0xBF58476D1CE4E5B9ull) ^ (outer * 0x94D049BB133111EBull) ^ (ld << 17) ^ ((uint64_t)box_inner << 40) ^
         ((uint64_t)box_outer << 50);
    TmapEntry& e = g_tmap_cache[(h >> 20) & (TMAP_CACHE_SIZE - 1)];
    {
        std::lock_guard<std::mutex> lock(g_tmap_mu);
        if (e.valid && e.key == key) {
            *tm = e.map;
            ++g_tmap_hits;
            return 0;
        }
    }
    const int rc = encode_tmap_bf16(tm, base, inner, outer, ld, box_inner, box_outer);
    if (rc) return rc;
    std::lock_guard<std::mutex> lock(g_tmap_mu);
    e.key = key;
    e.map = *tm;
    e.valid = true;
    ++g_tmap_misses;
    return 0;
}

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) once per (kernel, device) instead of on every launch
static int ensure_smem_attr(const void* kern, int bytes) {
    struct Done { const void* kern; int dev; };
    static Done done[256];
    static int n_done = 0;
    static std::mutex mu;
    int dev = 0;
    B200_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lock(mu);
    for (int i = 0; i < n_done; ++i)
        if (done[i].kern == kern && done[i].dev == dev) return 0;
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    if (n_done < 256) done[n_done++] = Done{kern, dev};
    return 0;
}

typedef void (*gemm_kernel_t)(const CUtensorMap, const CUtensorMap, const GemmParams);

// Work counters of the dynamically scheduled 2-CTA GEMM: zero at load, put back to zero by every launch that used one.
// Launches take them in rotation, so kernels of different streams that overlap do not share one.
constexpr int SCHED_COUNTERS = 64;
static __device__ int g_sched_counters[SCHED_COUNTERS];
static std::atomic<int> g_gemm_dynamic{-1};      // -1: read B200VIT_GEMM_SCHED on first use
static std::atomic<unsigned> g_sched_next{0};

void set_gemm_dynamic(int on) { g_gemm_dynamic.store(on ? 1 : 0); }

static bool gemm_dynamic() {
    int v = g_gemm_dynamic.load();
    if (v < 0) {
        const char* e = getenv("B200VIT_GEMM_SCHED");
        v = (e && (e[0] == 's' || e[0] == '0')) ? 0 : 1;      // "static" / "0" turn it off
        g_gemm_dynamic.store(v);
    }
    return v != 0;
}

static int next_sched_counter(int** out) {
    constexpr int MAX_DEV = 64;
    static int* base[MAX_DEV] = {};
    static std::mutex mu;
    int dev = 0;
    B200_CUDA(cudaGetDevice(&dev));
    B200_CHECK_ARG(dev >= 0 && dev < MAX_DEV, "gemm: device index %d out of range", dev);
    {
        std::lock_guard<std::mutex> lock(mu);
        if (!base[dev]) B200_CUDA(cudaGetSymbolAddress((void**)&base[dev], g_sched_counters));
        *out = base[dev] + (g_sched_next.fetch_add(1) % SCHED_COUNTERS);
    }
    return 0;
}

template <bool A_MN, bool B_MN>
static gemm_kernel_t pick_epi(int epi) {
    switch (epi) {
        case EPI_BF16: return gemm_tcgen05_kernel<A_MN, B_MN, EPI_BF16>;
        case EPI_BIAS_GELU: return gemm_tcgen05_kernel<A_MN, B_MN, EPI_BIAS_GELU>;
        case EPI_BIAS_RES: return gemm_tcgen05_kernel<A_MN, B_MN, EPI_BIAS_RES>;
        case EPI_DGELU: return gemm_tcgen05_kernel<A_MN, B_MN, EPI_DGELU>;
        case EPI_ACCUM: return gemm_tcgen05_kernel<A_MN, B_MN, EPI_ACCUM>;
        case EPI_F32: return gemm_tcgen05_kernel<A_MN, B_MN, EPI_F32>;
        case EPI_PATCH: return gemm_tcgen05_kernel<A_MN, B_MN, EPI_PATCH>;
    }
    if constexpr (!A_MN && !B_MN) {     // the LayerNorm-fold epilogues exist for forward (NT) products only
        switch (epi) {
            case EPI_RES_STATS: return gemm_tcgen05_kernel<false, false, EPI_RES_STATS>;
            case EPI_PATCH_STATS: return gemm_tcgen05_kernel<false, false, EPI_PATCH_STATS>;
            case EPI_LN_BF16: return gemm_tcgen05_kernel<false, false, EPI_LN_BF16>;
            case EPI_LN_GELU: return gemm_tcgen05_kernel<false, false, EPI_LN_GELU>;
        }
    }
    return nullptr;
}

template <bool A_MN, bool B_MN>
static gemm_kernel_t pick_epi_2cta(int epi) {
    switch (epi) {
        case EPI_BF16: return gemm_tcgen05_2cta_kernel<A_MN, B_MN, EPI_BF16>;
        case EPI_BIAS_GELU: return gemm_tcgen05_2cta_kernel<A_MN, B_MN, EPI_BIAS_GELU>;
        case EPI_BIAS_RES: return gemm_tcgen05_2cta_kernel<A_MN, B_MN, EPI_BIAS_RES>;
        case EPI_DGELU: return gemm_tcgen05_2cta_kernel<A_MN, B_MN, EPI_DGELU>;
        case EPI_ACCUM: return gemm_tcgen05_2cta_kernel<A_MN, B_MN, EPI_ACCUM>;
        case EPI_F32: return gemm_tcgen05_2cta_kernel<A_MN, B_MN, EPI_F32>;
        case EPI_PATCH: return gemm_tcgen05_2cta_kernel<A_MN, B_MN, EPI_PATCH>;
    }
    if constexpr (!A_MN && !B_MN) {
        switch (epi) {
            case EPI_RES_STATS: return gemm_tcgen05_2cta_kernel<false, false, EPI_RES_STATS>;
            case EPI_PATCH_STATS: return gemm_tcgen05_2cta_kernel<false, false, EPI_PATCH_STATS>;
            case EPI_LN_BF16: return gemm_tcgen05_2cta_kernel<false, false, EPI_LN_BF16>;
            case EPI_LN_GELU: return gemm_tcgen05_2cta_kernel<false, false, EPI_LN_GELU>;
        }
    }
    return nullptr;
}

static int gemm_mode() {   // B200VIT_GEMM_MODE=1 forces the 1-CTA kernel (debugging / A-B measurements)
    static int mode = -1;
    if (mode < 0) {
        const char* e = getenv("B200VIT_GEMM_MODE");
        mode = e ? atoi(e) : 0;
    }
    return mode;
}

int gemm_launch(int op, int epi, int M, int N, int K, const void* A, long long lda, const void* B, long long ldb,
                void* C, long long ldc, void* C2, long long ldc2, const float* bias, const void* aux,
                long long ldaux, float drop_p, unsigned long long seed, unsigned long long offset, int row_period,
                int splits, cudaStream_t stream) {
    return gemm_launch_ln(op, epi, M, N, K, A, lda, B, ldb, C, ldc, C2, ldc2, bias, aux, ldaux, drop_p, seed, offset,
                          row_period, splits, nullptr, stream);
}

int gemm_launch_ln(int op, int epi, int M, int N, int K, const void* A, long long lda, const void* B, long long ldb,
                   void* C, long long ldc, void* C2, long long ldc2, const float* bias, const void* aux,
                   long long ldaux, float drop_p, unsigned long long seed, unsigned long long offset, int row_period,
                   int splits, const GemmLnArgs* ln, cudaStream_t stream) {
    B200_CHECK_ARG(op >= 0 && op <= 2, "gemm: bad op %d", op);
    B200_CHECK_ARG(epi >= 0 && epi < EPI_COUNT, "gemm: bad epilogue %d", epi);
    B200_CHECK_ARG(M > 0 && N > 0 && K > 0, "gemm: empty problem M=%d N=%d K=%d", M, N, K);
    // C may be NULL only for the GELU epilogue (forward without a backward to follow: mask*gelu' is not needed)
    B200_CHECK_ARG(A && B && (C || is_gelu(epi)), "gemm: null operand");
    B200_CHECK_ARG(epi < EPI_RES_STATS || (ln != nullptr && op == OP_NT), "gemm: LayerNorm-fold epilogue %d needs its arguments and op NT", epi);
    if (has_stats(epi))
        B200_CHECK_ARG(ln->x16 && ln->rowstats && (ln->ldx16 % 8) == 0 && ((uintptr_t)ln->x16 % 16) == 0,
                       "gemm: statistics epilogue needs x16 (16-byte aligned, ld %% 8 == 0) and rowstats");
    if (is_lnin(epi))
        B200_CHECK_ARG(ln->ln_stats && ln->ln_s && ln->ln_dim > 0 && ((uintptr_t)ln->ln_s % 16) == 0 &&
                       ((uintptr_t)ln->ln_stats % 8) == 0, "gemm: folded-LayerNorm epilogue needs row statistics, s[] and the width");
    B200_CHECK_ARG((lda % 8) == 0 && (ldb % 8) == 0, "gemm: lda/ldb must be multiples of 8 elements (TMA 16B stride), got %lld %lld", lda, ldb);
    B200_CHECK_ARG(((uintptr_t)A % 16) == 0 && ((uintptr_t)B % 16) == 0, "gemm: A/B must be 16-byte aligned");
    B200_CHECK_ARG(!is_gelu(epi) || C2 != nullptr, "gemm: GELU epilogue needs the second output");
    B200_CHECK_ARG(epi != EPI_DGELU || aux != nullptr, "gemm: dGELU epilogue needs the saved mask*gelu' operand");
    B200_CHECK_ARG(!is_patch(epi) || (aux != nullptr && row_period > 0), "gemm: patch epilogue needs pos embedding and period");
    B200_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f, "gemm: dropout p must be in [0,1)");

    GemmParams p;
    p.M = M; p.N = N; p.K = K;
    p.tiles_m = (M + BM - 1) / BM;
    p.tiles_n = (N + BN - 1) / BN;
    p.kblocks = (K + BK - 1) / BK;
    const bool two_cta = (M > BM) && gemm_mode() != 1;
    if (splits <= 0) {
        // auto: enough (tile, split) work items to fill every SM (or SM pair) about twice
        splits = 1;
        if (epi == EPI_ACCUM) {
            const int units = two_cta ? sm_count() / 2 : sm_count();
            const int tiles = (two_cta ? (M + 2 * BM - 1) / (2 * BM) : p.tiles_m) * p.tiles_n;
            splits = (2 * units) / tiles;
            if (splits > p.kblocks / 4) splits = p.kblocks / 4;
            if (splits < 1) splits = 1;
        }
    }
    if (splits > p.kblocks) splits = p.kblocks;
    B200_CHECK_ARG(splits == 1 || epi == EPI_ACCUM, "gemm: split-K needs the accumulate epilogue");
    p.splits = splits;
    p.C = C; p.ldc = ldc; p.C2 = C2; p.ldc2 = ldc2;
    p.bias = bias; p.aux = aux; p.ldaux = ldaux;
    p.drop = make_dropout(drop_p, seed, offset);
    p.row_period = row_period;
    const bool out_f32 = (is_res(epi) || epi == EPI_ACCUM || epi == EPI_F32 || is_patch(epi));
    const int ovec = out_f32 ? 4 : 8;
    bool vec = (ldc % ovec) == 0 && ((uintptr_t)C % 16) == 0;
    if (C2) vec = vec && (ldc2 % 8) == 0 && ((uintptr_t)C2 % 16) == 0;
    if (bias) vec = vec && ((uintptr_t)bias % 16) == 0;
    if (aux) {
        const int avec = (epi == EPI_DGELU) ? 8 : 4;
        vec = vec && (ldaux % avec) == 0 && ((uintptr_t)aux % 16) == 0;
    }
    p.vec_ok = vec ? 1 : 0;
    p.sched_counter = nullptr;
    p.x16 = nullptr; p.ldx16 = 0; p.rowstats = nullptr; p.ln_stats = nullptr; p.ln_s = nullptr;
    p.ln_inv_dim = 0.f; p.ln_eps = 0.f;
    if (ln != nullptr) {
        p.x16 = (bf16*)ln->x16; p.ldx16 = ln->ldx16; p.rowstats = ln->rowstats;
        p.ln_stats = ln->ln_stats; p.ln_s = ln->ln_s;
        p.ln_inv_dim = ln->ln_dim > 0 ? 1.0f / (float)ln->ln_dim : 0.f;
        p.ln_eps = ln->ln_eps;
    }

    CUtensorMap tmA, tmB;
    int rc;
    const bool a_mn = (op == OP_TN), b_mn = (op != OP_NT);
    if (!a_mn) rc = make_tmap_bf16(&tmA, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, BK, BM);
    else       rc = make_tmap_bf16(&tmA, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, 64, BK);
    if (rc) return rc;
    // K-major B box: 256 rows for the 1-CTA kernel, 128 (this CTA's half) for the 2-CTA kernel
    if (!b_mn) rc = make_tmap_bf16(&tmB, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, BK, two_cta ? BN / 2 : BN);
    else       rc = make_tmap_bf16(&tmB, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, 64, BK);
    if (rc) return rc;

    gemm_kernel_t kern = nullptr;
    if (two_cta) {
        if (op == OP_NT) kern = pick_epi_2cta<false, false>(epi);
        else if (op == OP_NN) kern = pick_epi_2cta<false, true>(epi);
        else kern = pick_epi_2cta<true, true>(epi);
    } else {
        if (op == OP_NT) kern = pick_epi<false, false>(epi);
        else if (op == OP_NN) kern = pick_epi<false, true>(epi);
        else kern = pick_epi<true, true>(epi);
    }
    B200_CHECK_ARG(kern != nullptr, "gemm: no kernel for op %d epi %d", op, epi);

    if (two_cta) {
        const bool heavy = is_gelu(epi);
        const int smem2 = heavy ? Cfg2<EPI_BIAS_GELU>::SMEM : Cfg2<EPI_BF16>::SMEM;
        const int threads2 = heavy ? Cfg2<EPI_BIAS_GELU>::THREADS : Cfg2<EPI_BF16>::THREADS;
        { const int rc2 = ensure_smem_attr((const void*)kern, smem2); if (rc2) return rc2; }
        const int tiles_m2 = (M + 2 * BM - 1) / (2 * BM);
        const int total = tiles_m2 * p.tiles_n * p.splits;
        const int pairs = sm_count() / 2;
        const int grid = 2 * (total < pairs ? total : pairs);
        p.sched_counter = nullptr;
        if (gemm_dynamic() && total > pairs) { const int rc3 = next_sched_counter(&p.sched_counter); if (rc3) return rc3; }
        B200_CUDA(launch_kernel_pdl(kern, dim3(grid), dim3(threads2), (size_t)smem2, stream, tmA, tmB, p));   // cluster dims are compiled in
    } else {
        { const int rc2 = ensure_smem_attr((const void*)kern, GEMM_SMEM_BYTES); if (rc2) return rc2; }
        const int total = p.tiles_m * p.tiles_n * p.splits;
        const int grid = total < sm_count() ? total : sm_count();
        B200_CUDA(launch_kernel_pdl(kern, dim3(grid), dim3(GEMM_THREADS), (size_t)GEMM_SMEM_BYTES, stream, tmA, tmB, p));
    }
    B200_LAUNCH_CHECK();
    return 0;
}

}  // namespace b200vit
