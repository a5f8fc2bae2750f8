// Fused multi-head self-attention, forward and backward, for head_dim 64: the dispatch between the kernel families
// and the mma.sync family itself (FA2-style; used for attn_dropout > 0 and for sequences beyond 2048 tokens -- the
// 197- and 577-token cases of the benchmark configs run the tcgen05 kernels of attention_tc.cu). Replaces the scaled_dot_product_attention call inside nn.MultiheadAttention
// (reference models/vit.py:86-97, need_weights=False, no mask, softmax scale 1/sqrt(head_dim)) and the
// q/k/v head split / merge copies around it: the kernels read Q, K, V straight out of the packed
// in-projection output [B, T, 3, H, 64] and write the merged-head output [B, T, H*64].
//
// Flash-style: scores never reach HBM; the forward keeps only the per-row log-sum-exp, the backward
// recomputes P. Two backward passes without atomics: (dK, dV) per key block and dQ per query block.
// Tensor-core path: mma.sync m16n8k16 bf16 with fp32 accumulation, ldmatrix from XOR-swizzled smem.
#include "common.cuh"
#include "launchers.h"

#include <stdlib.h>

namespace b200vit {

constexpr int AT_BLOCK = 64;     // rows per tile (queries or keys)
constexpr int AT_HD = 64;        // head dim
constexpr int AT_THREADS = 128;  // 4 warps x 16 rows
constexpr int AT_TILE_BYTES = AT_BLOCK * AT_HD * 2;

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// byte offset of (row, 16-byte chunk) inside a 64x64 bf16 tile, XOR-swizzled against ldmatrix conflicts
__device__ __forceinline__ uint32_t tile_off(int row, int chunk) { return row * 128 + ((chunk ^ (row & 7)) << 4); }

// global [rows, 64] slice (row stride in elements) -> smem tile; rows >= T are zero-filled
__device__ __forceinline__ void load_tile(uint8_t* s, const bf16* g, long long row_stride, int row0, int T) {
    const uint32_t sbase = smem_addr(s);
    for (int c = threadIdx.x; c < AT_BLOCK * 8; c += AT_THREADS) {
        const int r = c >> 3, ch = c & 7;
        const uint32_t dst = sbase + tile_off(r, ch);
        const int t = row0 + r;
        if (t < T) {
            const bf16* src = g + (long long)t * row_stride + ch * 8;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
        } else {
            asm volatile("st.shared.v4.b32 [%0], {%1,%1,%1,%1};" ::"r"(dst), "r"(0u) : "memory");
        }
    }
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}

// A fragments (16 rows starting at r0, all four 16-wide k-steps of the 64-wide tile)
__device__ __forceinline__ void load_a_frags(const uint8_t* s, int r0, int lane, uint32_t (&a)[4][4]) {
    const uint32_t sbase = smem_addr(s);
    const int row = r0 + (lane & 7) + 8 * ((lane >> 3) & 1);
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
        const int chunk = kk * 2 + (lane >> 4);
        ldsm_x4(sbase + tile_off(row, chunk), a[kk][0], a[kk][1], a[kk][2], a[kk][3]);
    }
}

// C[16 x 64] += A[16 x 64(k)] * Bsmem^T where Bsmem is [64 n-rows][64 k] (non-transposed ldmatrix)
__device__ __forceinline__ void mma_a_bT(float (&c)[8][4], const uint32_t (&a)[4][4], const uint8_t* sB, int lane,
                                         int npairs = 4) {
    const uint32_t sbase = smem_addr(sB);
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
        for (int np = 0; np < 4; ++np) {
            if (np >= npairs) continue;
            const int row = np * 16 + (lane & 7) + 8 * (lane >> 4);
            const int chunk = kk * 2 + ((lane >> 3) & 1);
            uint32_t b0, b1, b2, b3;
            ldsm_x4(sbase + tile_off(row, chunk), b0, b1, b2, b3);
            mma_bf16(c[2 * np], a[kk], b0, b1);
            mma_bf16(c[2 * np + 1], a[kk], b2, b3);
        }
    }
}

// C[16 x 64(n)] += P[16 x 64(k)] * Bsmem where Bsmem is [64 k-rows][64 n] (transposed ldmatrix); P comes from
// an fp32 accumulator tile (same thread layout) converted to bf16 A fragments on the fly.
__device__ __forceinline__ void mma_p_b(float (&c)[8][4], const float (&p)[8][4], const uint8_t* sB, int lane,
                                        int npairs = 4) {
    const uint32_t sbase = smem_addr(sB);
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
        if (kk >= npairs) continue;
        uint32_t a[4];
        a[0] = pack_bf16(p[2 * kk][0], p[2 * kk][1]);
        a[1] = pack_bf16(p[2 * kk][2], p[2 * kk][3]);
        a[2] = pack_bf16(p[2 * kk + 1][0], p[2 * kk + 1][1]);
        a[3] = pack_bf16(p[2 * kk + 1][2], p[2 * kk + 1][3]);
#pragma unroll
        for (int np = 0; np < 4; ++np) {
            const int row = kk * 16 + (lane & 7) + 8 * ((lane >> 3) & 1);
            const int chunk = np * 2 + (lane >> 4);
            uint32_t b0, b1, b2, b3;
            ldsm_x4_trans(sbase + tile_off(row, chunk), b0, b1, b2, b3);
            mma_bf16(c[2 * np], a, b0, b1);
            mma_bf16(c[2 * np + 1], a, b2, b3);
        }
    }
}

__device__ __forceinline__ void zero_acc(float (&c)[8][4]) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) c[i][j] = 0.f;
}

// write a 16 x 64 accumulator tile as bf16 rows (row stride in elements); rows >= T are skipped
__device__ __forceinline__ void store_acc_bf16(const float (&c)[8][4], bf16* g, long long row_stride, int row0, int T,
                                               int lane, float mul0, float mul1) {
    const int gq = lane >> 2, tq = lane & 3;
    const int r_lo = row0 + gq, r_hi = row0 + gq + 8;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
        const int col = nt * 8 + 2 * tq;
        if (r_lo < T) *reinterpret_cast<uint32_t*>(g + (long long)r_lo * row_stride + col) = pack_bf16(c[nt][0] * mul0, c[nt][1] * mul0);
        if (r_hi < T) *reinterpret_cast<uint32_t*>(g + (long long)r_hi * row_stride + col) = pack_bf16(c[nt][2] * mul1, c[nt][3] * mul1);
    }
}

// ---------------------------------------------------------------------------------------------
// forward
// ---------------------------------------------------------------------------------------------
// DROP: dropout on the attention probabilities (nn.MultiheadAttention(dropout=p), models/vit.py:86-90, applied by
// torch to softmax(S) before the PV product). Element (b, h, q, k) of the probability tensor has index
// ((b*H + h)*T + q)*T + k in the counter-hash mask of common.cuh, so forward and backward regenerate the same mask.
template <bool DROP>
__global__ void __launch_bounds__(AT_THREADS)
attention_fwd_kernel(const bf16* __restrict__ qkv, bf16* __restrict__ out, float* __restrict__ lse, int T, int H,
                     float scale_log2, float scale, DropoutParams drop) {
    __shared__ __align__(128) uint8_t sQ[AT_TILE_BYTES];
    __shared__ __align__(128) uint8_t sK[AT_TILE_BYTES];
    __shared__ __align__(128) uint8_t sV[AT_TILE_BYTES];
    const int bh = blockIdx.x, b = bh / H, h = bh % H;
    const int q0 = blockIdx.y * AT_BLOCK;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long rs = 3LL * H * AT_HD;
    const bf16* gQ = qkv + (long long)b * T * rs + (long long)h * AT_HD;
    const bf16* gK = gQ + (long long)H * AT_HD;
    const bf16* gV = gQ + 2LL * H * AT_HD;

    load_tile(sQ, gQ, rs, q0, T);
    cp_async_wait_all();
    __syncthreads();
    uint32_t qf[4][4];
    load_a_frags(sQ, warp * 16, lane, qf);

    float o[8][4];
    zero_acc(o);
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
    const int tq = lane & 3;

    for (int k0 = 0; k0 < T; k0 += AT_BLOCK) {
        __syncthreads();  // previous iteration finished reading sK / sV
        load_tile(sK, gK, rs, k0, T);
        load_tile(sV, gV, rs, k0, T);
        cp_async_wait_all();
        __syncthreads();

        float s[8][4];
        zero_acc(s);
        mma_a_bT(s, qf, sK, lane);

        float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
            const int key = k0 + nt * 8 + 2 * tq;
            if (key >= T) { s[nt][0] = -INFINITY; s[nt][2] = -INFINITY; }
            if (key + 1 >= T) { s[nt][1] = -INFINITY; s[nt][3] = -INFINITY; }
            mx0 = fmaxf(mx0, fmaxf(s[nt][0], s[nt][1]));
            mx1 = fmaxf(mx1, fmaxf(s[nt][2], s[nt][3]));
        }
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);   // finite: every block holds >= 1 valid key
        const float c0 = exp2f((m0 - mn0) * scale_log2), c1 = exp2f((m1 - mn1) * scale_log2);
        m0 = mn0; m1 = mn1;
        float rs0 = 0.f, rs1 = 0.f;
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
            s[nt][0] = exp2f((s[nt][0] - mn0) * scale_log2);
            s[nt][1] = exp2f((s[nt][1] - mn0) * scale_log2);
            s[nt][2] = exp2f((s[nt][2] - mn1) * scale_log2);
            s[nt][3] = exp2f((s[nt][3] - mn1) * scale_log2);
            rs0 += s[nt][0] + s[nt][1];
            rs1 += s[nt][2] + s[nt][3];
        }
        l0 = l0 * c0 + rs0;
        l1 = l1 * c1 + rs1;
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) { o[nt][0] *= c0; o[nt][1] *= c0; o[nt][2] *= c1; o[nt][3] *= c1; }
        if (DROP) {     // the row sums above use the un-dropped probabilities; only the PV product sees the mask
            const unsigned long long e_lo = ((unsigned long long)bh * T + (q0 + warp * 16 + (lane >> 2))) * T;
            const unsigned long long e_hi = e_lo + 8ull * T;
#pragma unroll
            for (int nt = 0; nt < 8; ++nt) {
                const int key = k0 + nt * 8 + 2 * tq;
                s[nt][0] *= dropout_mult1(drop, e_lo + key);
                s[nt][1] *= dropout_mult1(drop, e_lo + key + 1);
                s[nt][2] *= dropout_mult1(drop, e_hi + key);
                s[nt][3] *= dropout_mult1(drop, e_hi + key + 1);
            }
        }
        mma_p_b(o, s, sV, lane);
    }
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
    l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 2);

    const int row0 = q0 + warp * 16;
    bf16* gO = out + (long long)b * T * H * AT_HD + (long long)h * AT_HD;
    store_acc_bf16(o, gO, (long long)H * AT_HD, row0, T, lane, 1.f / l0, 1.f / l1);
    if (tq == 0) {
        const int gq = lane >> 2;
        float* L = lse + (long long)bh * T;
        if (row0 + gq < T) L[row0 + gq] = m0 * scale + logf(l0);
        if (row0 + gq + 8 < T) L[row0 + gq + 8] = m1 * scale + logf(l1);
    }
}

// ---------------------------------------------------------------------------------------------
// backward prologue: delta[b, h, t] = sum_d dO[b,t,h,d] * O[b,t,h,d]
// ---------------------------------------------------------------------------------------------
__global__ void attention_delta_kernel(const bf16* __restrict__ out, const bf16* __restrict__ dout,
                                       float* __restrict__ delta, int B, int T, int H) {
    // 8 lanes per (batch, token, head) row of 64 elements: one 16-byte load per lane and operand
    const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    const int sub = threadIdx.x & 7;
    const bool ok = w < (long long)B * T * H;
    float s = 0.f;
    if (ok) {
        const uint4 o = __ldg(reinterpret_cast<const uint4*>(out + w * AT_HD) + sub);
        const uint4 d = __ldg(reinterpret_cast<const uint4*>(dout + w * AT_HD) + sub);
        const uint32_t ow[4] = {o.x, o.y, o.z, o.w}, dw[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float2 fo = unpack_bf16(ow[k]), fd = unpack_bf16(dw[k]);
            s = fmaf(fo.x, fd.x, fmaf(fo.y, fd.y, s));
        }
    }
    s += __shfl_xor_sync(0xffffffffu, s, 4);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    if (ok && sub == 0) {
        const int h = (int)(w % H);
        const long long bt = w / H;
        const int t = (int)(bt % T), b = (int)(bt / T);
        delta[((long long)b * H + h) * T + t] = s;
    }
}

// ---------------------------------------------------------------------------------------------
// backward pass 1: dK, dV for one 64-key block; loops over the query blocks
// ---------------------------------------------------------------------------------------------
template <bool DROP>
__global__ void __launch_bounds__(AT_THREADS)
attention_bwd_dkdv_kernel(const bf16* __restrict__ qkv, const bf16* __restrict__ dout, const float* __restrict__ lse,
                          const float* __restrict__ delta, bf16* __restrict__ dqkv, int T, int H, float scale_log2,
                          float scale, DropoutParams drop) {
    __shared__ __align__(128) uint8_t sK[AT_TILE_BYTES];
    __shared__ __align__(128) uint8_t sV[AT_TILE_BYTES];
    __shared__ __align__(128) uint8_t sQ[AT_TILE_BYTES];
    __shared__ __align__(128) uint8_t sdO[AT_TILE_BYTES];
    __shared__ float sLse[AT_BLOCK], sDelta[AT_BLOCK];
    const int bh = blockIdx.x, b = bh / H, h = bh % H;
    const int k0 = blockIdx.y * AT_BLOCK;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tq = lane & 3;
    const long long rs = 3LL * H * AT_HD;
    const long long os = (long long)H * AT_HD;
    const bf16* gQ = qkv + (long long)b * T * rs + (long long)h * AT_HD;
    const bf16* gK = gQ + (long long)H * AT_HD;
    const bf16* gV = gQ + 2LL * H * AT_HD;
    const bf16* gdO = dout + (long long)b * T * os + (long long)h * AT_HD;
    const float* L = lse + (long long)bh * T;
    const float* Dl = delta + (long long)bh * T;

    load_tile(sK, gK, rs, k0, T);
    load_tile(sV, gV, rs, k0, T);
    cp_async_wait_all();
    __syncthreads();
    uint32_t kf[4][4], vf[4][4];
    load_a_frags(sK, warp * 16, lane, kf);
    load_a_frags(sV, warp * 16, lane, vf);

    float dk[8][4], dv[8][4];
    zero_acc(dk);
    zero_acc(dv);
    const float LOG2E = 1.4426950408889634f;

    for (int q0 = 0; q0 < T; q0 += AT_BLOCK) {
        __syncthreads();
        load_tile(sQ, gQ, rs, q0, T);
        load_tile(sdO, gdO, os, q0, T);
        if (threadIdx.x < AT_BLOCK) {
            const int t = q0 + threadIdx.x;
            sLse[threadIdx.x] = t < T ? L[t] * LOG2E : INFINITY;   // +inf -> P = 0 for padded queries
            sDelta[threadIdx.x] = t < T ? Dl[t] : 0.f;
        }
        cp_async_wait_all();
        __syncthreads();

        // S^T = K Q^T : rows = keys of this warp, cols = queries of the block
        float st[8][4];
        zero_acc(st);
        mma_a_bT(st, kf, sQ, lane);
        // dP^T = V dO^T
        float dpt[8][4];
        zero_acc(dpt);
        mma_a_bT(dpt, vf, sdO, lane);
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
            const int q = nt * 8 + 2 * tq;
            const float ls0 = sLse[q], ls1 = sLse[q + 1];
            const float d0 = sDelta[q], d1 = sDelta[q + 1];
            const float p0 = exp2f(st[nt][0] * scale_log2 - ls0);
            const float p1 = exp2f(st[nt][1] * scale_log2 - ls1);
            const float p2 = exp2f(st[nt][2] * scale_log2 - ls0);
            const float p3 = exp2f(st[nt][3] * scale_log2 - ls1);
            float m0 = 1.f, m1 = 1.f, m2 = 1.f, m3 = 1.f;     // dropout multipliers of (query q / q+1, key lo / hi)
            if (DROP) {
                const int key_lo = k0 + warp * 16 + (lane >> 2);
                const unsigned long long e0 = ((unsigned long long)bh * T + (q0 + q)) * T, e1 = e0 + T;
                m0 = dropout_mult1(drop, e0 + key_lo);
                m1 = dropout_mult1(drop, e1 + key_lo);
                m2 = dropout_mult1(drop, e0 + key_lo + 8);
                m3 = dropout_mult1(drop, e1 + key_lo + 8);
            }
            st[nt][0] = p0 * m0; st[nt][1] = p1 * m1; st[nt][2] = p2 * m2; st[nt][3] = p3 * m3;   // dropped P -> dV
            dpt[nt][0] = p0 * (dpt[nt][0] * m0 - d0) * scale;
            dpt[nt][1] = p1 * (dpt[nt][1] * m1 - d1) * scale;
            dpt[nt][2] = p2 * (dpt[nt][2] * m2 - d0) * scale;
            dpt[nt][3] = p3 * (dpt[nt][3] * m3 - d1) * scale;
        }
        mma_p_b(dv, st, sdO, lane);    // dV += P^T dO
        mma_p_b(dk, dpt, sQ, lane);    // dK += dS^T Q
    }
    bf16* gdK = dqkv + (long long)b * T * rs + (long long)H * AT_HD + (long long)h * AT_HD;
    bf16* gdV = gdK + (long long)H * AT_HD;
    store_acc_bf16(dk, gdK, rs, k0 + warp * 16, T, lane, 1.f, 1.f);
    store_acc_bf16(dv, gdV, rs, k0 + warp * 16, T, lane, 1.f, 1.f);
}

// ---------------------------------------------------------------------------------------------
// backward pass 2: dQ for one 64-query block; loops over the key blocks
// ---------------------------------------------------------------------------------------------
template <bool DROP>
__global__ void __launch_bounds__(AT_THREADS)
attention_bwd_dq_kernel(const bf16* __restrict__ qkv, const bf16* __restrict__ dout, const float* __restrict__ lse,
                        const float* __restrict__ delta, bf16* __restrict__ dqkv, int T, int H, float scale_log2,
                        float scale, DropoutParams drop) {
    __shared__ __align__(128) uint8_t sQ[AT_TILE_BYTES];
    __shared__ __align__(128) uint8_t sdO[AT_TILE_BYTES];
    __shared__ __align__(128) uint8_t sK[AT_TILE_BYTES];
    __shared__ __align__(128) uint8_t sV[AT_TILE_BYTES];
    const int bh = blockIdx.x, b = bh / H, h = bh % H;
    const int q0 = blockIdx.y * AT_BLOCK;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tq = lane & 3, gq = lane >> 2;
    const long long rs = 3LL * H * AT_HD;
    const long long os = (long long)H * AT_HD;
    const bf16* gQ = qkv + (long long)b * T * rs + (long long)h * AT_HD;
    const bf16* gK = gQ + (long long)H * AT_HD;
    const bf16* gV = gQ + 2LL * H * AT_HD;
    const bf16* gdO = dout + (long long)b * T * os + (long long)h * AT_HD;
    const float LOG2E = 1.4426950408889634f;

    load_tile(sQ, gQ, rs, q0, T);
    load_tile(sdO, gdO, os, q0, T);
    cp_async_wait_all();
    __syncthreads();
    uint32_t qf[4][4], dof[4][4];
    load_a_frags(sQ, warp * 16, lane, qf);
    load_a_frags(sdO, warp * 16, lane, dof);

    const int r_lo = q0 + warp * 16 + gq, r_hi = r_lo + 8;
    const float* L = lse + (long long)bh * T;
    const float* Dl = delta + (long long)bh * T;
    const float ls0 = r_lo < T ? L[r_lo] * LOG2E : INFINITY, ls1 = r_hi < T ? L[r_hi] * LOG2E : INFINITY;
    const float dl0 = r_lo < T ? Dl[r_lo] : 0.f, dl1 = r_hi < T ? Dl[r_hi] : 0.f;

    float dq[8][4];
    zero_acc(dq);
    for (int k0 = 0; k0 < T; k0 += AT_BLOCK) {
        __syncthreads();
        load_tile(sK, gK, rs, k0, T);
        load_tile(sV, gV, rs, k0, T);
        cp_async_wait_all();
        __syncthreads();
        float s[8][4], dp[8][4];
        zero_acc(s);
        zero_acc(dp);
        mma_a_bT(s, qf, sK, lane);     // S = Q K^T
        mma_a_bT(dp, dof, sV, lane);   // dP = dO V^T
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
            const int key = k0 + nt * 8 + 2 * tq;
            const bool v0 = key < T, v1 = key + 1 < T;
            const float p0 = v0 ? exp2f(s[nt][0] * scale_log2 - ls0) : 0.f;
            const float p1 = v1 ? exp2f(s[nt][1] * scale_log2 - ls0) : 0.f;
            const float p2 = v0 ? exp2f(s[nt][2] * scale_log2 - ls1) : 0.f;
            const float p3 = v1 ? exp2f(s[nt][3] * scale_log2 - ls1) : 0.f;
            if (DROP) {
                const unsigned long long e_lo = ((unsigned long long)bh * T + r_lo) * T, e_hi = e_lo + 8ull * T;
                dp[nt][0] *= dropout_mult1(drop, e_lo + key);
                dp[nt][1] *= dropout_mult1(drop, e_lo + key + 1);
                dp[nt][2] *= dropout_mult1(drop, e_hi + key);
                dp[nt][3] *= dropout_mult1(drop, e_hi + key + 1);
            }
            dp[nt][0] = p0 * (dp[nt][0] - dl0) * scale;
            dp[nt][1] = p1 * (dp[nt][1] - dl0) * scale;
            dp[nt][2] = p2 * (dp[nt][2] - dl1) * scale;
            dp[nt][3] = p3 * (dp[nt][3] - dl1) * scale;
        }
        mma_p_b(dq, dp, sK, lane);     // dQ += dS K
    }
    bf16* gdQ = dqkv + (long long)b * T * rs + (long long)h * AT_HD;
    store_acc_bf16(dq, gdQ, rs, q0 + warp * 16, T, lane, 1.f, 1.f);
}


constexpr int AT_TC_MAX_T = 256;      // the single-head-resident tcgen05 kernels (attention_tc.cu)

// B200VIT_ATTN_MODE=1 forces the mma.sync kernels below for every length (A/B measurements, debugging).
static int attention_mode() {
    static int mode = -1;
    if (mode < 0) {
        const char* e = getenv("B200VIT_ATTN_MODE");
        mode = e ? atoi(e) : 0;
    }
    return mode;
}

// ---------------------------------------------------------------------------------------------
// Dispatch: tokens <= 256 -> tcgen05 kernels with the whole head resident; <= 2048 -> tcgen05 streaming kernels;
// longer (or B200VIT_ATTN_MODE=1) -> the FA2-style mma.sync kernels of this file.
// ---------------------------------------------------------------------------------------------
int attention_fwd_launch(const bf16* qkv, bf16* out, float* lse, int batch, int tokens, int heads, int head_dim,
                         float scale, cudaStream_t stream) {
    B200_CHECK_ARG(qkv && out && lse, "attention_fwd: null pointer");
    B200_CHECK_ARG(head_dim == AT_HD, "attention_fwd: head_dim must be 64 (got %d)", head_dim);
    B200_CHECK_ARG(batch > 0 && tokens > 0 && heads > 0, "attention_fwd: bad sizes");
    B200_CHECK_ARG(((uintptr_t)qkv % 16) == 0 && ((uintptr_t)out % 16) == 0, "attention_fwd: unaligned pointer");
    if (attention_mode() == 0) {
        if (tokens <= AT_TC_MAX_T) return attention_fwd_tc_launch(qkv, out, lse, batch, tokens, heads, scale, stream);
        if (tokens <= AT_LONG_MAX_T) return attention_fwd_long_launch(qkv, out, lse, batch, tokens, heads, scale, stream);
    }
    dim3 grid(batch * heads, (tokens + AT_BLOCK - 1) / AT_BLOCK);
    attention_fwd_kernel<false><<<grid, AT_THREADS, 0, stream>>>(qkv, out, lse, tokens, heads, scale * 1.4426950408889634f,
                                                                scale, DropoutParams{});
    B200_LAUNCH_CHECK();
    return 0;
}

int attention_bwd_launch(const bf16* qkv, const bf16* out, const bf16* dout, const float* lse, float* delta,
                         bf16* dqkv, float* dbias, int batch, int tokens, int heads, int head_dim, float scale,
                         cudaStream_t stream) {
    B200_CHECK_ARG(qkv && out && dout && lse && delta && dqkv, "attention_bwd: null pointer");
    B200_CHECK_ARG(head_dim == AT_HD, "attention_bwd: head_dim must be 64 (got %d)", head_dim);
    B200_CHECK_ARG(batch > 0 && tokens > 0 && heads > 0, "attention_bwd: bad sizes");
    if (tokens <= AT_TC_MAX_T && attention_mode() == 0)
        return attention_bwd_tc_launch(qkv, out, dout, lse, dqkv, dbias, batch, tokens, heads, scale, stream);
    const long long rows = (long long)batch * tokens * heads;
    attention_delta_kernel<<<(unsigned)((rows * 8 + 255) / 256), 256, 0, stream>>>(out, dout, delta, batch, tokens, heads);
    B200_LAUNCH_CHECK();
    if (attention_mode() == 0 && tokens <= AT_LONG_MAX_T) {
        const int rc = attention_bwd_long_launch(qkv, dout, lse, delta, dqkv, batch, tokens, heads, scale, stream);
        if (rc || !dbias) return rc;
        return colsum_bf16_launch(dqkv, 3LL * heads * AT_HD, dbias, (long long)batch * tokens, 3 * heads * AT_HD, stream);
    }
    dim3 grid(batch * heads, (tokens + AT_BLOCK - 1) / AT_BLOCK);
    const float sl2 = scale * 1.4426950408889634f;
    attention_bwd_dkdv_kernel<false><<<grid, AT_THREADS, 0, stream>>>(qkv, dout, lse, delta, dqkv, tokens, heads, sl2, scale, DropoutParams{});
    B200_LAUNCH_CHECK();
    attention_bwd_dq_kernel<false><<<grid, AT_THREADS, 0, stream>>>(qkv, dout, lse, delta, dqkv, tokens, heads, sl2, scale, DropoutParams{});
    B200_LAUNCH_CHECK();
    return dbias ? colsum_bf16_launch(dqkv, 3LL * heads * AT_HD, dbias, (long long)batch * tokens, 3 * heads * AT_HD, stream) : 0;
}

// ---------------------------------------------------------------------------------------------
// attention-probability dropout (attn_dropout > 0; the reference default is 0): the FA2-style kernels with the mask
// ---------------------------------------------------------------------------------------------
int attention_fwd_dropout_launch(const bf16* qkv, bf16* out, float* lse, int batch, int tokens, int heads, int head_dim,
                                 float scale, float p, unsigned long long seed, unsigned long long offset,
                                 cudaStream_t stream) {
    if (!(p > 0.f)) return attention_fwd_launch(qkv, out, lse, batch, tokens, heads, head_dim, scale, stream);
    B200_CHECK_ARG(qkv && out && lse, "attention_fwd: null pointer");
    B200_CHECK_ARG(head_dim == AT_HD, "attention_fwd: head_dim must be 64 (got %d)", head_dim);
    B200_CHECK_ARG(batch > 0 && tokens > 0 && heads > 0 && p < 1.f, "attention_fwd: bad sizes / dropout probability");
    B200_CHECK_ARG(((uintptr_t)qkv % 16) == 0 && ((uintptr_t)out % 16) == 0, "attention_fwd: unaligned pointer");
    dim3 grid(batch * heads, (tokens + AT_BLOCK - 1) / AT_BLOCK);
    attention_fwd_kernel<true><<<grid, AT_THREADS, 0, stream>>>(qkv, out, lse, tokens, heads, scale * 1.4426950408889634f,
                                                               scale, make_dropout(p, seed, offset));
    B200_LAUNCH_CHECK();
    return 0;
}

int attention_bwd_dropout_launch(const bf16* qkv, const bf16* out, const bf16* dout, const float* lse, float* delta,
                                 bf16* dqkv, float* dbias, int batch, int tokens, int heads, int head_dim, float scale,
                                 float p, unsigned long long seed, unsigned long long offset, cudaStream_t stream) {
    if (!(p > 0.f))
        return attention_bwd_launch(qkv, out, dout, lse, delta, dqkv, dbias, batch, tokens, heads, head_dim, scale, stream);
    B200_CHECK_ARG(qkv && out && dout && lse && delta && dqkv, "attention_bwd: null pointer");
    B200_CHECK_ARG(head_dim == AT_HD, "attention_bwd: head_dim must be 64 (got %d)", head_dim);
    B200_CHECK_ARG(batch > 0 && tokens > 0 && heads > 0 && p < 1.f, "attention_bwd: bad sizes / dropout probability");
    const long long rows = (long long)batch * tokens * heads;
    attention_delta_kernel<<<(unsigned)((rows * 8 + 255) / 256), 256, 0, stream>>>(out, dout, delta, batch, tokens, heads);
    B200_LAUNCH_CHECK();
    dim3 grid(batch * heads, (tokens + AT_BLOCK - 1) / AT_BLOCK);
    const float sl2 = scale * 1.4426950408889634f;
    const DropoutParams drop = make_dropout(p, seed, offset);
    attention_bwd_dkdv_kernel<true><<<grid, AT_THREADS, 0, stream>>>(qkv, dout, lse, delta, dqkv, tokens, heads, sl2, scale, drop);
    B200_LAUNCH_CHECK();
    attention_bwd_dq_kernel<true><<<grid, AT_THREADS, 0, stream>>>(qkv, dout, lse, delta, dqkv, tokens, heads, sl2, scale, drop);
    B200_LAUNCH_CHECK();
    return dbias ? colsum_bf16_launch(dqkv, 3LL * heads * AT_HD, dbias, (long long)batch * tokens, 3 * heads * AT_HD, stream) : 0;
}

}  // namespace b200vit
