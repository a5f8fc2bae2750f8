// Strict-precision ("fp32 mode") support kernels.
//
// The reference computes everything in fp32 (models/vit.py builds default-dtype nn modules and
// going_modular/engine.py:50-66 runs them as they are). The throughput path of this library feeds the tensor cores
// bf16 operands, which moves the logits by ~4e-3 relative. For parity runs that must agree with the reference to
// 1e-3 the same tcgen05 GEMM kernel is used with SPLIT operands instead: every fp32 value x is written as
//     x = hi + lo + r,   hi = bf16(x),  lo = bf16(x - hi),  |r| <= 2^-17 |x|
// and a product a.b is evaluated as  a_hi b_hi + a_lo b_hi + a_hi b_lo  (the dropped a_lo b_lo term is below
// 2^-16 |a b|). The three partial products are not three GEMM launches: the operands are concatenated along the
// contraction dimension,
//     A' = [A_hi | A_lo | A_hi]   (pattern 0),      B' = [B_hi | B_hi | B_lo]   (pattern 1),      K' = 3 K,
// so ONE launch of b200vit_gemm accumulates all three in fp32 in tensor memory. Everything between two GEMMs
// (LayerNorm output, GELU, attention, residual stream, gradients) stays fp32 in this mode; the kernels for that
// are below. They are CUDA-core kernels written for accuracy and simplicity, not for speed: this mode exists to
// pin the numerics of the tensor-core path against the reference, the bf16 path is the product.
#include "common.cuh"
#include "launchers.h"

namespace b200vit {

// Source row of logical row r when class-token rows are skipped (row_period = patches per image): same mapping as
// the EPI_PATCH epilogue uses for its output rows.
__device__ __forceinline__ long long src_row_of(long long r, int row_period) {
    return row_period > 0 ? r + r / row_period + 1 : r;
}

// ---------------------------------------------------------------------------------------------
// split: dst part k (k = 0,1,2) at dst + k * part_stride, each [rows, cols] with leading dimension ld_dst.
//   pattern 0: (hi, lo, hi)   pattern 1: (hi, hi, lo)
// Optional dropout mask on the source (element index = source_row * cols + col, the index the forward used) and
// optional class-row skipping let the backward passes read gradients of dropout outputs in place.
// ---------------------------------------------------------------------------------------------
__global__ void split_bf16x3_kernel(const float* __restrict__ src, long long ld_src, long long rows, int cols,
                                    bf16* __restrict__ dst, long long ld_dst, long long part_stride, int pattern,
                                    DropoutParams drop, int row_period) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * (long long)cols) return;
    const long long r = i / cols;
    const int c = (int)(i % cols);
    const long long sr = src_row_of(r, row_period);
    float v = src[sr * ld_src + c];
    if (drop.p > 0.f) v *= dropout_mult1(drop, (unsigned long long)sr * (unsigned long long)cols + c);
    const bf16 hi = __float2bfloat16(v);
    const bf16 lo = __float2bfloat16(v - __bfloat162float(hi));
    bf16* d = dst + r * ld_dst + c;
    d[0] = hi;
    d[part_stride] = pattern == 0 ? lo : hi;
    d[2 * part_stride] = pattern == 0 ? hi : lo;
}

int split_bf16x3_launch(const float* src, long long ld_src, long long rows, int cols, bf16* dst, long long ld_dst,
                        long long part_stride, int pattern, float drop_p, unsigned long long seed,
                        unsigned long long offset, int row_period, cudaStream_t stream) {
    B200_CHECK_ARG(src && dst, "split_bf16x3: null pointer");
    B200_CHECK_ARG(rows > 0 && cols > 0 && ld_src >= cols && ld_dst >= cols, "split_bf16x3: bad sizes");
    B200_CHECK_ARG(pattern == 0 || pattern == 1, "split_bf16x3: pattern must be 0 (hi,lo,hi) or 1 (hi,hi,lo)");
    B200_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f, "split_bf16x3: dropout p must be in [0,1)");
    const long long total = rows * (long long)cols;
    split_bf16x3_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(
        src, ld_src, rows, cols, dst, ld_dst, part_stride, pattern, make_dropout(drop_p, seed, offset), row_period);
    B200_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------------------------
// im2col with fp32 output (same re-index as im2col_kernel in elementwise.cu; models/vit.py:28-34,53-56)
// ---------------------------------------------------------------------------------------------
__global__ void im2col_f32_kernel(const float* __restrict__ img, float* __restrict__ out, int B, int C, int H, int W,
                                  int P, long long total) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= total) return;
    const int Wp = W / P, Hp = H / P;
    const int K = C * P * P;
    const long long row = t / K;
    const int k = (int)(t % K);
    const int c = k / (P * P);
    const int i = (k % (P * P)) / P;
    const int j = k % P;
    const int b = (int)(row / (Hp * Wp));
    const int pp = (int)(row % (Hp * Wp));
    const int py = pp / Wp, px = pp % Wp;
    out[t] = img[(((long long)b * C + c) * H + (py * P + i)) * W + px * P + j];
}

int im2col_f32_launch(const float* image, float* patches, int batch, int channels, int height, int width, int patch,
                      cudaStream_t stream) {
    B200_CHECK_ARG(image && patches, "im2col_f32: null pointer");
    B200_CHECK_ARG(batch > 0 && channels > 0 && patch > 0, "im2col_f32: bad sizes");
    B200_CHECK_ARG(height % patch == 0 && width % patch == 0, "im2col_f32: image %dx%d not divisible by patch %d",
                   height, width, patch);
    const long long total = (long long)batch * (height / patch) * (width / patch) * channels * patch * patch;
    im2col_f32_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(image, patches, batch, channels, height,
                                                                          width, patch, total);
    B200_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------------------------
// GELU (exact erf form, nn.GELU() default, models/vit.py:121) followed by Dropout (vit.py:122), fp32 in and out,
// and its backward: da = dh * mask/(1-p) * gelu'(a).
// ---------------------------------------------------------------------------------------------
__global__ void gelu_f32_fwd_kernel(const float* __restrict__ a, float* __restrict__ h, long long n, DropoutParams drop) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float x = a[i];
    float y = 0.5f * x * (1.0f + erff(x * 0.70710678118654752f));
    if (drop.p > 0.f) y *= dropout_mult1(drop, (unsigned long long)i);
    h[i] = y;
}

__global__ void gelu_f32_bwd_kernel(const float* __restrict__ a, const float* __restrict__ dh, float* __restrict__ da,
                                    long long n, DropoutParams drop) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float x = a[i];
    const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752f));
    const float pdf = 0.3989422804014327f * expf(-0.5f * x * x);
    float g = dh[i] * (cdf + x * pdf);
    if (drop.p > 0.f) g *= dropout_mult1(drop, (unsigned long long)i);
    da[i] = g;
}

int gelu_f32_fwd_launch(const float* a, float* h, long long n, float drop_p, unsigned long long seed,
                        unsigned long long offset, cudaStream_t stream) {
    B200_CHECK_ARG(a && h && n > 0, "gelu_f32_fwd: bad arguments");
    B200_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f, "gelu_f32_fwd: dropout p must be in [0,1)");
    gelu_f32_fwd_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(a, h, n, make_dropout(drop_p, seed, offset));
    B200_LAUNCH_CHECK();
    return 0;
}

int gelu_f32_bwd_launch(const float* a, const float* dh, float* da, long long n, float drop_p, unsigned long long seed,
                        unsigned long long offset, cudaStream_t stream) {
    B200_CHECK_ARG(a && dh && da && n > 0, "gelu_f32_bwd: bad arguments");
    B200_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f, "gelu_f32_bwd: dropout p must be in [0,1)");
    gelu_f32_bwd_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(a, dh, da, n, make_dropout(drop_p, seed, offset));
    B200_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Column sums of an fp32 matrix with the same optional dropout mask / class-row skipping as the split kernel:
// bias gradients db[n] = sum_m mask[m, n] * g[m, n]. out is overwritten.
// ---------------------------------------------------------------------------------------------
constexpr int CSF_ROWS_PER_BLOCK = 256;

__global__ void colsum_f32_kernel(const float* __restrict__ x, long long ld, long long rows, int cols,
                                  float* __restrict__ out, DropoutParams drop, int row_period) {
    __shared__ float red[8][32];
    const int c = blockIdx.x * 32 + threadIdx.x;
    const long long r0 = (long long)blockIdx.y * CSF_ROWS_PER_BLOCK;
    const long long r1 = min(rows, r0 + CSF_ROWS_PER_BLOCK);
    float acc = 0.f;
    if (c < cols) {
        for (long long r = r0 + threadIdx.y; r < r1; r += 8) {
            const long long sr = src_row_of(r, row_period);
            float v = x[sr * ld + c];
            if (drop.p > 0.f) v *= dropout_mult1(drop, (unsigned long long)sr * (unsigned long long)cols + c);
            acc += v;
        }
    }
    red[threadIdx.y][threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.y == 0 && c < cols) {
        float s = 0.f;
#pragma unroll
        for (int y = 0; y < 8; ++y) s += red[y][threadIdx.x];
        atomicAdd(out + c, s);
    }
}

int colsum_f32_launch(const float* x, long long ld, long long rows, int cols, float* out, float drop_p,
                      unsigned long long seed, unsigned long long offset, int row_period, cudaStream_t stream) {
    B200_CHECK_ARG(x && out, "colsum_f32: null pointer");
    B200_CHECK_ARG(rows > 0 && cols > 0 && ld >= cols, "colsum_f32: bad sizes");
    B200_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f, "colsum_f32: dropout p must be in [0,1)");
    B200_CUDA(cudaMemsetAsync(out, 0, sizeof(float) * cols, stream));
    dim3 grid((cols + 31) / 32, (unsigned)((rows + CSF_ROWS_PER_BLOCK - 1) / CSF_ROWS_PER_BLOCK));
    colsum_f32_kernel<<<grid, dim3(32, 8), 0, stream>>>(x, ld, rows, cols, out, make_dropout(drop_p, seed, offset),
                                                       row_period);
    B200_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------------------------
// fp32 multi-head self-attention (the scaled-dot-product attention inside nn.MultiheadAttention,
// models/vit.py:86-97), head_dim 64, any sequence length.
//   qkv  f32 [B, T, 3, H, 64]  (= the [B*T, 3*D] output of the in-projection GEMM, read in place)
//   out  f32 [B, T, H*64]      lse f32 [B, H, T] = log sum_j exp(scale * q_i . k_j)
// A thread PAIR owns one row (query row in fwd / dq, key row in dk/dv); each thread of the pair holds 32 of the 64
// head dimensions, dot products are completed with one shuffle. The other side is streamed through shared memory in
// tiles of 64 rows (every lane reads the same element: broadcast). Online softmax in the forward; the backward
// recomputes probabilities from lse.
// ---------------------------------------------------------------------------------------------
constexpr int AF_ROWS = 64;           // rows per block (128 threads)
constexpr int AF_TILE = 64;           // rows of the streamed operand per shared-memory tile
constexpr int AF_HD = 64;

__device__ __forceinline__ float pair_sum(float v) { return v + __shfl_xor_sync(0xffffffffu, v, 1); }

// cooperative load of up to AF_TILE rows of one of q/k/v (which = 0/1/2) of head (b, h) into smem [AF_TILE][64]
__device__ __forceinline__ void af_load_tile(float (*dst)[AF_HD], const float* __restrict__ qkv, int b, int h, int T,
                                             int H, int which, int row0) {
    const long long ld = 3LL * H * AF_HD;
    for (int idx = threadIdx.x; idx < AF_TILE * (AF_HD / 4); idx += blockDim.x) {
        const int r = idx / (AF_HD / 4), c4 = idx % (AF_HD / 4);
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row0 + r < T)
            v = *reinterpret_cast<const float4*>(qkv + ((long long)b * T + row0 + r) * ld + (long long)which * H * AF_HD +
                                                 h * AF_HD + c4 * 4);
        *reinterpret_cast<float4*>(&dst[r][c4 * 4]) = v;
    }
}
// the same for an [B, T, H*64] tensor (attention output / its gradient)
__device__ __forceinline__ void af_load_tile_o(float (*dst)[AF_HD], const float* __restrict__ o, int b, int h, int T,
                                               int H, int row0) {
    const long long ld = (long long)H * AF_HD;
    for (int idx = threadIdx.x; idx < AF_TILE * (AF_HD / 4); idx += blockDim.x) {
        const int r = idx / (AF_HD / 4), c4 = idx % (AF_HD / 4);
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row0 + r < T)
            v = *reinterpret_cast<const float4*>(o + ((long long)b * T + row0 + r) * ld + h * AF_HD + c4 * 4);
        *reinterpret_cast<float4*>(&dst[r][c4 * 4]) = v;
    }
}

__global__ void __launch_bounds__(2 * AF_ROWS)
attention_f32_fwd_kernel(const float* __restrict__ qkv, float* __restrict__ out, float* __restrict__ lse, int B, int T,
                         int H, float scale) {
    __shared__ __align__(16) float sK[AF_TILE][AF_HD];
    __shared__ __align__(16) float sV[AF_TILE][AF_HD];
    const int b = blockIdx.z, h = blockIdx.y;
    const int row = blockIdx.x * AF_ROWS + (threadIdx.x >> 1);
    const int half = threadIdx.x & 1;
    const long long ld = 3LL * H * AF_HD;
    const bool valid = row < T;
    float q[32], o[32];
#pragma unroll
    for (int d = 0; d < 32; ++d) {
        q[d] = valid ? qkv[((long long)b * T + row) * ld + h * AF_HD + half * 32 + d] : 0.f;
        o[d] = 0.f;
    }
    float m = -INFINITY, l = 0.f;
    for (int j0 = 0; j0 < T; j0 += AF_TILE) {
        __syncthreads();
        af_load_tile(sK, qkv, b, h, T, H, 1, j0);
        af_load_tile(sV, qkv, b, h, T, H, 2, j0);
        __syncthreads();
        const int jn = min(AF_TILE, T - j0);
        for (int j = 0; j < jn; ++j) {
            float s = 0.f;
#pragma unroll
            for (int d = 0; d < 32; ++d) s = fmaf(q[d], sK[j][half * 32 + d], s);
            s = pair_sum(s) * scale;
            if (s > m) {
                const float corr = expf(m - s);     // exp(-inf) = 0 on the first key
                l *= corr;
#pragma unroll
                for (int d = 0; d < 32; ++d) o[d] *= corr;
                m = s;
            }
            const float p = expf(s - m);
            l += p;
#pragma unroll
            for (int d = 0; d < 32; ++d) o[d] = fmaf(p, sV[j][half * 32 + d], o[d]);
        }
    }
    if (valid) {
        const float inv = 1.0f / l;
        float* dst = out + ((long long)b * T + row) * ((long long)H * AF_HD) + h * AF_HD + half * 32;
#pragma unroll
        for (int d = 0; d < 32; ++d) dst[d] = o[d] * inv;
        if (half == 0) lse[((long long)b * H + h) * T + row] = m + logf(l);
    }
}

// dq and delta = rowsum(dO * O), one query row per thread pair
__global__ void __launch_bounds__(2 * AF_ROWS)
attention_f32_bwd_dq_kernel(const float* __restrict__ qkv, const float* __restrict__ out, const float* __restrict__ dout,
                            const float* __restrict__ lse, float* __restrict__ delta, float* __restrict__ dqkv, int B,
                            int T, int H, float scale) {
    __shared__ __align__(16) float sK[AF_TILE][AF_HD];
    __shared__ __align__(16) float sV[AF_TILE][AF_HD];
    const int b = blockIdx.z, h = blockIdx.y;
    const int row = blockIdx.x * AF_ROWS + (threadIdx.x >> 1);
    const int half = threadIdx.x & 1;
    const long long ld = 3LL * H * AF_HD, ldo = (long long)H * AF_HD;
    const bool valid = row < T;
    float q[32], dO[32], dq[32];
    float dl = 0.f;
#pragma unroll
    for (int d = 0; d < 32; ++d) {
        q[d] = valid ? qkv[((long long)b * T + row) * ld + h * AF_HD + half * 32 + d] : 0.f;
        dO[d] = valid ? dout[((long long)b * T + row) * ldo + h * AF_HD + half * 32 + d] : 0.f;
        const float ov = valid ? out[((long long)b * T + row) * ldo + h * AF_HD + half * 32 + d] : 0.f;
        dl = fmaf(dO[d], ov, dl);
        dq[d] = 0.f;
    }
    dl = pair_sum(dl);
    const float L = valid ? lse[((long long)b * H + h) * T + row] : 0.f;
    if (valid && half == 0) delta[((long long)b * H + h) * T + row] = dl;
    for (int j0 = 0; j0 < T; j0 += AF_TILE) {
        __syncthreads();
        af_load_tile(sK, qkv, b, h, T, H, 1, j0);
        af_load_tile(sV, qkv, b, h, T, H, 2, j0);
        __syncthreads();
        const int jn = min(AF_TILE, T - j0);
        for (int j = 0; j < jn; ++j) {
            float s = 0.f, dp = 0.f;
#pragma unroll
            for (int d = 0; d < 32; ++d) {
                s = fmaf(q[d], sK[j][half * 32 + d], s);
                dp = fmaf(dO[d], sV[j][half * 32 + d], dp);
            }
            s = pair_sum(s) * scale;
            dp = pair_sum(dp);
            const float ds = expf(s - L) * (dp - dl) * scale;
#pragma unroll
            for (int d = 0; d < 32; ++d) dq[d] = fmaf(ds, sK[j][half * 32 + d], dq[d]);
        }
    }
    if (valid) {
        float* dst = dqkv + ((long long)b * T + row) * ld + h * AF_HD + half * 32;
#pragma unroll
        for (int d = 0; d < 32; ++d) dst[d] = dq[d];
    }
}

// dk and dv, one key row per thread pair; queries, dO, lse and delta streamed through shared memory
__global__ void __launch_bounds__(2 * AF_ROWS)
attention_f32_bwd_dkv_kernel(const float* __restrict__ qkv, const float* __restrict__ dout, const float* __restrict__ lse,
                             const float* __restrict__ delta, float* __restrict__ dqkv, int B, int T, int H, float scale) {
    __shared__ __align__(16) float sQ[AF_TILE][AF_HD];
    __shared__ __align__(16) float sDO[AF_TILE][AF_HD];
    __shared__ float sL[AF_TILE], sD[AF_TILE];
    const int b = blockIdx.z, h = blockIdx.y;
    const int row = blockIdx.x * AF_ROWS + (threadIdx.x >> 1);
    const int half = threadIdx.x & 1;
    const long long ld = 3LL * H * AF_HD;
    const bool valid = row < T;
    float k[32], v[32], dk[32], dv[32];
#pragma unroll
    for (int d = 0; d < 32; ++d) {
        k[d] = valid ? qkv[((long long)b * T + row) * ld + (long long)H * AF_HD + h * AF_HD + half * 32 + d] : 0.f;
        v[d] = valid ? qkv[((long long)b * T + row) * ld + 2LL * H * AF_HD + h * AF_HD + half * 32 + d] : 0.f;
        dk[d] = 0.f;
        dv[d] = 0.f;
    }
    for (int i0 = 0; i0 < T; i0 += AF_TILE) {
        __syncthreads();
        af_load_tile(sQ, qkv, b, h, T, H, 0, i0);
        af_load_tile_o(sDO, dout, b, h, T, H, i0);
        if (threadIdx.x < AF_TILE) {
            const bool in = i0 + threadIdx.x < T;
            sL[threadIdx.x] = in ? lse[((long long)b * H + h) * T + i0 + threadIdx.x] : 0.f;
            sD[threadIdx.x] = in ? delta[((long long)b * H + h) * T + i0 + threadIdx.x] : 0.f;
        }
        __syncthreads();
        const int in_ = min(AF_TILE, T - i0);
        for (int i = 0; i < in_; ++i) {
            float s = 0.f, dp = 0.f;
#pragma unroll
            for (int d = 0; d < 32; ++d) {
                s = fmaf(sQ[i][half * 32 + d], k[d], s);
                dp = fmaf(sDO[i][half * 32 + d], v[d], dp);
            }
            s = pair_sum(s) * scale;
            dp = pair_sum(dp);
            const float p = expf(s - sL[i]);
            const float ds = p * (dp - sD[i]) * scale;
#pragma unroll
            for (int d = 0; d < 32; ++d) {
                dv[d] = fmaf(p, sDO[i][half * 32 + d], dv[d]);
                dk[d] = fmaf(ds, sQ[i][half * 32 + d], dk[d]);
            }
        }
    }
    if (valid) {
        float* dkp = dqkv + ((long long)b * T + row) * ld + (long long)H * AF_HD + h * AF_HD + half * 32;
        float* dvp = dqkv + ((long long)b * T + row) * ld + 2LL * H * AF_HD + h * AF_HD + half * 32;
#pragma unroll
        for (int d = 0; d < 32; ++d) {
            dkp[d] = dk[d];
            dvp[d] = dv[d];
        }
    }
}

int attention_f32_fwd_launch(const float* qkv, float* out, float* lse, int batch, int tokens, int heads, int head_dim,
                             float scale, cudaStream_t stream) {
    B200_CHECK_ARG(qkv && out && lse, "attention_f32_fwd: null pointer");
    B200_CHECK_ARG(batch > 0 && tokens > 0 && heads > 0, "attention_f32_fwd: bad sizes");
    B200_CHECK_ARG(head_dim == AF_HD, "attention_f32_fwd: head_dim must be 64 (got %d)", head_dim);
    B200_CHECK_ARG(((uintptr_t)qkv % 16) == 0, "attention_f32_fwd: qkv must be 16-byte aligned");
    B200_CHECK_ARG(heads <= 65535 && batch <= 65535, "attention_f32_fwd: batch / heads exceed the grid limits");
    dim3 grid((tokens + AF_ROWS - 1) / AF_ROWS, heads, batch);
    attention_f32_fwd_kernel<<<grid, 2 * AF_ROWS, 0, stream>>>(qkv, out, lse, batch, tokens, heads, scale);
    B200_LAUNCH_CHECK();
    return 0;
}

int attention_f32_bwd_launch(const float* qkv, const float* out, const float* dout, const float* lse, float* delta,
                             float* dqkv, int batch, int tokens, int heads, int head_dim, float scale,
                             cudaStream_t stream) {
    B200_CHECK_ARG(qkv && out && dout && lse && delta && dqkv, "attention_f32_bwd: null pointer");
    B200_CHECK_ARG(batch > 0 && tokens > 0 && heads > 0, "attention_f32_bwd: bad sizes");
    B200_CHECK_ARG(head_dim == AF_HD, "attention_f32_bwd: head_dim must be 64 (got %d)", head_dim);
    B200_CHECK_ARG(((uintptr_t)qkv % 16) == 0 && ((uintptr_t)dout % 16) == 0,
                   "attention_f32_bwd: qkv / dout must be 16-byte aligned");
    B200_CHECK_ARG(heads <= 65535 && batch <= 65535, "attention_f32_bwd: batch / heads exceed the grid limits");
    dim3 grid((tokens + AF_ROWS - 1) / AF_ROWS, heads, batch);
    attention_f32_bwd_dq_kernel<<<grid, 2 * AF_ROWS, 0, stream>>>(qkv, out, dout, lse, delta, dqkv, batch, tokens, heads,
                                                                 scale);
    B200_LAUNCH_CHECK();
    attention_f32_bwd_dkv_kernel<<<grid, 2 * AF_ROWS, 0, stream>>>(qkv, dout, lse, delta, dqkv, batch, tokens, heads,
                                                                  scale);
    B200_LAUNCH_CHECK();
    return 0;
}

// gy[b*(T-1) + t-1, :] = dropout_mask * gx[b, t, :] is not materialised in this mode: the split and column-sum
// kernels above read gx in place (row_period = T-1).

}  // namespace b200vit
