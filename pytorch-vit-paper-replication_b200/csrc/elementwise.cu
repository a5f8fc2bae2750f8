// Memory-bound helper kernels of the ViT step: casts, patch im2col, class-token rows, the gradients of the
// embedding parameters, column sums (bias gradients) and the dropout-masked cast feeding the backward
// GEMMs. All of them are HBM-bound: 16-byte vector accesses, coalesced along the feature dimension.
#include "common.cuh"
#include "launchers.h"

namespace b200vit {

// ---------------------------------------------------------------------------------------------
// fp32 -> bf16 cast (weight shadows, small conversions)
// ---------------------------------------------------------------------------------------------
__global__ void cast_f32_bf16_kernel(const float* __restrict__ src, bf16* __restrict__ dst, long long n) {
    const long long i8 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 8;
    if (i8 + 8 <= n && (((uintptr_t)(src + i8)) & 15) == 0 && (((uintptr_t)(dst + i8)) & 15) == 0) {
        const float4 a = *reinterpret_cast<const float4*>(src + i8);
        const float4 b = *reinterpret_cast<const float4*>(src + i8 + 4);
        uint4 o;
        o.x = pack_bf16(a.x, a.y); o.y = pack_bf16(a.z, a.w);
        o.z = pack_bf16(b.x, b.y); o.w = pack_bf16(b.z, b.w);
        *reinterpret_cast<uint4*>(dst + i8) = o;
    } else {
        for (long long i = i8; i < n && i < i8 + 8; ++i) dst[i] = __float2bfloat16(src[i]);
    }
}

int cast_f32_bf16_launch(const float* src, bf16* dst, long long n, cudaStream_t stream) {
    B200_CHECK_ARG(n >= 0, "cast: negative size");
    if (n == 0) return 0;
    B200_CHECK_ARG(src && dst, "cast: null pointer");
    const long long threads = (n + 7) / 8;
    const int block = 256;
    cast_f32_bf16_kernel<<<(unsigned)((threads + block - 1) / block), block, 0, stream>>>(src, dst, n);
    B200_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------------------------
// im2col for a stride==kernel convolution (reference: nn.Conv2d(k=s=patch) + Flatten + permute,
// models/vit.py:28-34,53-56). Because stride == kernel the "convolution" is a pure re-index:
//   patches[(b*Np + py*Wp + px), c*P*P + i*P + j] = image[b, c, py*P + i, px*P + j]
// which makes the projection the NT GEMM patches[B*Np, C*P*P] x W[D, C*P*P]^T.
// One thread moves 8 consecutive j (two float4 loads -> one 16B bf16 store).
// ---------------------------------------------------------------------------------------------
__global__ void im2col_kernel(const float* __restrict__ img, bf16* __restrict__ out, int B, int C, int H, int W,
                              int P, long long total8) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= total8) return;
    const int Wp = W / P, Hp = H / P;
    const int K = C * P * P;
    const int k8 = K / 8;
    const long long row = t / k8;
    const int k = (int)(t % k8) * 8;
    const int c = k / (P * P);
    const int i = (k % (P * P)) / P;
    const int j = k % P;
    const int b = (int)(row / (Hp * Wp));
    const int pp = (int)(row % (Hp * Wp));
    const int py = pp / Wp, px = pp % Wp;
    const float* src = img + (((long long)b * C + c) * H + (py * P + i)) * W + px * P + j;
    const float4 a = *reinterpret_cast<const float4*>(src);
    const float4 d = *reinterpret_cast<const float4*>(src + 4);
    uint4 o;
    o.x = pack_bf16(a.x, a.y); o.y = pack_bf16(a.z, a.w);
    o.z = pack_bf16(d.x, d.y); o.w = pack_bf16(d.z, d.w);
    *reinterpret_cast<uint4*>(out + row * K + k) = o;
}

// Same re-index for uint8 images (NCHW), scaled to [0, 1] on the fly: the torchvision pipeline of the reference
// (ToImage + ToDtype(float32, scale=True), 08_..._self.ipynb:253-265) applied on the device, so the host->device copy
// moves 1 byte per pixel-channel instead of 4. value = u8 / 255 (exactly what ToDtype(scale=True) computes).
__global__ void im2col_u8_kernel(const uint8_t* __restrict__ img, bf16* __restrict__ out, int B, int C, int H, int W,
                                 int P, long long total8) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= total8) return;
    const int Wp = W / P, Hp = H / P;
    const int K = C * P * P;
    const int k8 = K / 8;
    const long long row = t / k8;
    const int k = (int)(t % k8) * 8;
    const int c = k / (P * P);
    const int i = (k % (P * P)) / P;
    const int j = k % P;
    const int b = (int)(row / (Hp * Wp));
    const int pp = (int)(row % (Hp * Wp));
    const int py = pp / Wp, px = pp % Wp;
    const uint8_t* src = img + (((long long)b * C + c) * H + (py * P + i)) * W + px * P + j;
    const uint2 q = *reinterpret_cast<const uint2*>(src);
    const float s = 1.0f / 255.0f;
    float f[8];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        f[e] = (float)((q.x >> (8 * e)) & 0xFFu) * s;
        f[4 + e] = (float)((q.y >> (8 * e)) & 0xFFu) * s;
    }
    uint4 o;
    o.x = pack_bf16(f[0], f[1]); o.y = pack_bf16(f[2], f[3]);
    o.z = pack_bf16(f[4], f[5]); o.w = pack_bf16(f[6], f[7]);
    *reinterpret_cast<uint4*>(out + row * K + k) = o;
}

int im2col_u8_launch(const uint8_t* image, bf16* patches, int batch, int channels, int height, int width, int patch,
                     cudaStream_t stream) {
    B200_CHECK_ARG(image && patches, "im2col: null pointer");
    B200_CHECK_ARG(batch > 0 && channels > 0 && patch > 0, "im2col: bad sizes");
    B200_CHECK_ARG(height % patch == 0 && width % patch == 0, "im2col: image %dx%d not divisible by patch %d", height,
                   width, patch);
    B200_CHECK_ARG(patch % 8 == 0 && width % 8 == 0, "im2col: patch must be a multiple of 8 (got %d)", patch);
    B200_CHECK_ARG(((uintptr_t)image % 8) == 0 && ((uintptr_t)patches % 16) == 0, "im2col: unaligned pointer");
    const long long rows = (long long)batch * (height / patch) * (width / patch);
    const long long total8 = rows * (channels * patch * patch / 8);
    const int block = 256;
    im2col_u8_kernel<<<(unsigned)((total8 + block - 1) / block), block, 0, stream>>>(image, patches, batch, channels,
                                                                                    height, width, patch, total8);
    B200_LAUNCH_CHECK();
    return 0;
}

int im2col_launch(const float* image, bf16* patches, int batch, int channels, int height, int width, int patch,
                  cudaStream_t stream) {
    B200_CHECK_ARG(image && patches, "im2col: null pointer");
    B200_CHECK_ARG(batch > 0 && channels > 0 && patch > 0, "im2col: bad sizes");
    B200_CHECK_ARG(height % patch == 0 && width % patch == 0, "im2col: image %dx%d not divisible by patch %d", height,
                   width, patch);
    B200_CHECK_ARG(patch % 8 == 0 && width % 4 == 0, "im2col: patch must be a multiple of 8 (got %d)", patch);
    B200_CHECK_ARG(((uintptr_t)image % 16) == 0 && ((uintptr_t)patches % 16) == 0, "im2col: unaligned pointer");
    const long long rows = (long long)batch * (height / patch) * (width / patch);
    const long long total8 = rows * (channels * patch * patch / 8);
    const int block = 256;
    im2col_kernel<<<(unsigned)((total8 + block - 1) / block), block, 0, stream>>>(image, patches, batch, channels,
                                                                                 height, width, patch, total8);
    B200_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Class-token rows: x[b, 0, :] = dropout(class_token + pos[0])  (models/vit.py:50,59,63,66).
// The patch rows are written by the GEMM epilogue (EPI_PATCH); together they replace
// expand + cat + add + dropout with zero extra passes over the token tensor.
// ---------------------------------------------------------------------------------------------
__global__ void cls_rows_kernel(const float* __restrict__ cls, const float* __restrict__ pos, float* __restrict__ x,
                                int B, int T, int D, DropoutParams drop, bf16* __restrict__ x16, float* __restrict__ rowstats) {
    __shared__ float red[2][8];
    const int b = blockIdx.x;
    float sum = 0.f, sq = 0.f;
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
        float v = cls[d] + pos[d];
        const unsigned long long e = (unsigned long long)b * T * D + d;
        if (drop.p > 0.f) v *= dropout_mult1(drop, e);
        x[(long long)b * T * D + d] = v;
        if (x16) x16[(long long)b * T * D + d] = __float2bfloat16(v);
        sum += v;
        sq = fmaf(v, v, sq);
    }
    if (rowstats == nullptr) return;
    // LayerNorm fold: this block owns the whole row, so its (sum, sum of squares) is stored, not accumulated
    sum = warp_sum(sum);
    sq = warp_sum(sq);
    if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = sum; red[1][threadIdx.x >> 5] = sq; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float a = 0.f, c = 0.f;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { a += red[0][w]; c += red[1][w]; }
        rowstats[2LL * b * T] = a;
        rowstats[2LL * b * T + 1] = c;
    }
}

int cls_rows_launch(const float* cls, const float* pos, float* x, int batch, int tokens, int dim, float drop_p,
                    unsigned long long seed, unsigned long long offset, cudaStream_t stream, bf16* x16, float* rowstats) {
    B200_CHECK_ARG(cls && pos && x, "class_token_rows: null pointer");
    B200_CHECK_ARG(batch > 0 && tokens > 0 && dim > 0, "class_token_rows: bad sizes");
    cls_rows_kernel<<<batch, 256, 0, stream>>>(cls, pos, x, batch, tokens, dim, make_dropout(drop_p, seed, offset), x16,
                                              rowstats);
    B200_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------------------------
// LayerNorm fold, weight side (see gemm_sm100.cu): for a Linear that consumes LayerNorm(x) (models/vit.py:93-97 in_proj,
// :129-130 fc1)   w16[n,k] = bf16(gamma[k] w[n,k]),  s[n] = sum_k w16[n,k]  (of the ROUNDED values: the tensor core
// multiplies those),  c[n] = sum_k beta[k] w[n,k] + bias[n].   One warp per output row.
// ---------------------------------------------------------------------------------------------
__global__ void fold_ln_weight_kernel(const float* __restrict__ w, const float* __restrict__ gamma,
                                      const float* __restrict__ beta, const float* __restrict__ bias,
                                      bf16* __restrict__ w16, float* __restrict__ s, float* __restrict__ c, int rows, int cols) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    float ss = 0.f, cc = 0.f;
    for (int k = lane; k < cols; k += 32) {
        const float wv = w[(long long)row * cols + k];
        const bf16 h = __float2bfloat16(wv * gamma[k]);
        w16[(long long)row * cols + k] = h;
        ss += __bfloat162float(h);
        cc = fmaf(wv, beta[k], cc);
    }
    ss = warp_sum(ss);
    cc = warp_sum(cc);
    if (lane == 0) {
        s[row] = ss;
        c[row] = cc + (bias ? bias[row] : 0.f);
    }
}

int fold_ln_weight_launch(const float* w, const float* gamma, const float* beta, const float* bias, bf16* w16, float* s,
                          float* c, int rows, int cols, cudaStream_t stream) {
    B200_CHECK_ARG(w && gamma && beta && w16 && s && c, "fold_ln_weight: null pointer");
    B200_CHECK_ARG(rows > 0 && cols > 0, "fold_ln_weight: bad sizes");
    fold_ln_weight_kernel<<<(rows + 7) / 8, 256, 0, stream>>>(w, gamma, beta, bias, w16, s, c, rows, cols);
    B200_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Gradients of class_token / position_embedding / conv bias from the gradient of the embedded
// sequence gx[B, T, D] (what autograd derives from vit.py:50-66):
//   gd = dropout_mask * gx ; g_pos[t] = sum_b gd[b,t] ; g_cls = g_pos[0] ; g_bias = sum_{t>=1} g_pos[t]
// ---------------------------------------------------------------------------------------------
__global__ void embed_grads_kernel(const float* __restrict__ gx, float* __restrict__ g_cls, float* __restrict__ g_pos,
                                   float* __restrict__ g_bias, int B, int T, int D, DropoutParams drop) {
    const int t = blockIdx.x;
    for (int d = blockIdx.y * blockDim.x + threadIdx.x; d < D; d += gridDim.y * blockDim.x) {
        float acc = 0.f;
#pragma unroll 4
        for (int b = 0; b < B; ++b) {
            const unsigned long long e = ((unsigned long long)b * T + t) * D + d;
            float v = gx[e];
            if (drop.p > 0.f) v *= dropout_mult1(drop, e);
            acc += v;
        }
        g_pos[(long long)t * D + d] = acc;
        if (t == 0) g_cls[d] = acc;
        else if (g_bias) atomicAdd(g_bias + d, acc);
    }
}

int embed_grads_launch(const float* gx, float* g_cls, float* g_pos, float* g_bias, int batch, int tokens, int dim,
                       float drop_p, unsigned long long seed, unsigned long long offset, cudaStream_t stream) {
    B200_CHECK_ARG(gx && g_cls && g_pos, "embed_param_grads: null pointer");
    B200_CHECK_ARG(batch > 0 && tokens > 0 && dim > 0, "embed_param_grads: bad sizes");
    if (g_bias) B200_CUDA(cudaMemsetAsync(g_bias, 0, sizeof(float) * dim, stream));
    embed_grads_kernel<<<dim3(tokens, (dim + 255) / 256), 256, 0, stream>>>(gx, g_cls, g_pos, g_bias, batch, tokens, dim,
                                                   make_dropout(drop_p, seed, offset));
    B200_LAUNCH_CHECK();
    return 0;
}

// gy[b*(T-1) + t-1, :] = bf16(dropout_mask * gx[b, t, :]) for t >= 1: the A operand of the patch wgrad GEMM.
__global__ void patch_grad_rows_kernel(const float* __restrict__ gx, bf16* __restrict__ gy, int B, int T, int D,
                                       DropoutParams drop, long long total) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int d = (int)(i % D);
    const long long r = i / D;
    const int b = (int)(r / (T - 1));
    const int t = (int)(r % (T - 1)) + 1;
    const unsigned long long e = ((unsigned long long)b * T + t) * D + d;
    float v = gx[e];
    if (drop.p > 0.f) v *= dropout_mult1(drop, e);
    gy[i] = __float2bfloat16(v);
}

// 8 elements per thread (D % 8 == 0, 16-byte aligned pointers): two float4 loads, one mask block, one 16-byte store
__global__ void patch_grad_rows_vec8_kernel(const float* __restrict__ gx, bf16* __restrict__ gy, int B, int T, int D,
                                            DropoutParams drop, long long total8) {
    const long long i8 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i8 >= total8) return;
    const int d8 = D >> 3;
    const int d = (int)(i8 % d8) * 8;
    const long long r = i8 / d8;
    const int b = (int)(r / (T - 1));
    const int t = (int)(r % (T - 1)) + 1;
    const unsigned long long e = ((unsigned long long)b * T + t) * D + d;
    const float4 a = *reinterpret_cast<const float4*>(gx + e);
    const float4 c = *reinterpret_cast<const float4*>(gx + e + 4);
    float v[8] = {a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w};
    if (drop.p > 0.f) {
        float m[8];
        dropout_mult8(drop, e >> 3, m);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] *= m[j];
    }
    uint4 o;
    o.x = pack_bf16(v[0], v[1]); o.y = pack_bf16(v[2], v[3]);
    o.z = pack_bf16(v[4], v[5]); o.w = pack_bf16(v[6], v[7]);
    *reinterpret_cast<uint4*>(gy + r * D + d) = o;
}

int patch_grad_rows_launch(const float* gx, bf16* gy, int batch, int tokens, int dim, float drop_p,
                           unsigned long long seed, unsigned long long offset, cudaStream_t stream) {
    B200_CHECK_ARG(gx && gy, "patch_grad_rows: null pointer");
    B200_CHECK_ARG(batch > 0 && tokens > 1 && dim > 0, "patch_grad_rows: bad sizes");
    const long long total = (long long)batch * (tokens - 1) * dim;
    const DropoutParams drop = make_dropout(drop_p, seed, offset);
    if ((dim & 7) == 0 && (((uintptr_t)gx) & 15) == 0 && (((uintptr_t)gy) & 15) == 0) {
        const long long total8 = total / 8;
        patch_grad_rows_vec8_kernel<<<(unsigned)((total8 + 255) / 256), 256, 0, stream>>>(gx, gy, batch, tokens, dim, drop,
                                                                                         total8);
    } else {
        patch_grad_rows_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(gx, gy, batch, tokens, dim, drop, total);
    }
    B200_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Column sums of a bf16 matrix (bias gradients: db[n] = sum_m dY[m, n]).
// Block = 32 x 8 threads; thread (x, y) owns 8 consecutive columns and rows y, y+8, ... of its row slab.
// ---------------------------------------------------------------------------------------------
constexpr int CS_ROWS_PER_BLOCK = 512;   // colsum: 64 rows per thread, loads of 4 rows issued together
constexpr int DC_ROWS_PER_BLOCK = 128;   // dropout_cast: 16 rows per thread

__global__ void colsum_bf16_kernel(const bf16* __restrict__ x, long long ld, float* __restrict__ out, long long rows,
                                   int cols) {
    __shared__ float red[8][32][8 + 1];
    const int c0 = (blockIdx.x * 32 + threadIdx.x) * 8;
    const long long r0 = (long long)blockIdx.y * CS_ROWS_PER_BLOCK;
    const long long r1 = min(rows, r0 + CS_ROWS_PER_BLOCK);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (c0 < cols) {
        const bool vec = (c0 + 8 <= cols) && ((ld & 7) == 0) && ((((uintptr_t)x) & 15) == 0);
        if (vec) {
            for (long long r = r0 + threadIdx.y; r < r1; r += 32) {
                uint4 q[4];
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    q[u] = (r + 8 * u < r1) ? *reinterpret_cast<const uint4*>(x + (r + 8 * u) * ld + c0) : make_uint4(0, 0, 0, 0);
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const uint32_t w[4] = {q[u].x, q[u].y, q[u].z, q[u].w};
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float2 f = unpack_bf16(w[j]);
                        acc[2 * j] += f.x; acc[2 * j + 1] += f.y;
                    }
                }
            }
        } else {
            for (long long r = r0 + threadIdx.y; r < r1; r += 8) {
                const bf16* p = x + r * ld + c0;
                for (int j = 0; j < 8 && c0 + j < cols; ++j) acc[j] += __bfloat162float(p[j]);
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) red[threadIdx.y][threadIdx.x][j] = acc[j];
    __syncthreads();
    if (threadIdx.y == 0 && c0 < cols) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float s = 0.f;
#pragma unroll
            for (int y = 0; y < 8; ++y) s += red[y][threadIdx.x][j];
            if (c0 + j < cols) atomicAdd(out + c0 + j, s);
        }
    }
}

int colsum_bf16_launch(const bf16* x, long long ld, float* out, long long rows, int cols, cudaStream_t stream) {
    B200_CHECK_ARG(x && out, "colsum: null pointer");
    B200_CHECK_ARG(rows > 0 && cols > 0 && ld >= cols, "colsum: bad sizes");
    B200_CUDA(cudaMemsetAsync(out, 0, sizeof(float) * cols, stream));
    dim3 grid((cols + 255) / 256, (unsigned)((rows + CS_ROWS_PER_BLOCK - 1) / CS_ROWS_PER_BLOCK));
    colsum_bf16_kernel<<<grid, dim3(32, 8), 0, stream>>>(x, ld, out, rows, cols);
    B200_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------------------------
// out_bf16 = bf16(dropout_mask * g) and (optionally) colsum[n] = sum_m out[m, n].
// Turns the fp32 residual-stream gradient into the bf16 operand of the fc2 / out-proj backward
// GEMMs (undoing the Dropout of vit.py:125) and yields the bias gradient in the same pass.
// ---------------------------------------------------------------------------------------------
__global__ void dropout_cast_kernel(const float* __restrict__ g, bf16* __restrict__ out, float* __restrict__ colsum,
                                    long long rows, int cols, DropoutParams drop) {
    __shared__ float red[8][32][8 + 1];
    const int c0 = (blockIdx.x * 32 + threadIdx.x) * 8;
    const long long r0 = (long long)blockIdx.y * DC_ROWS_PER_BLOCK;
    const long long r1 = min(rows, r0 + DC_ROWS_PER_BLOCK);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (c0 < cols) {
        const bool vec = (c0 + 8 <= cols) && ((cols & 7) == 0) && ((((uintptr_t)g) & 15) == 0) &&
                         ((((uintptr_t)out) & 15) == 0);
        for (long long rr = r0 + threadIdx.y; rr < r1; rr += 16) {
          float4 la[2], lb[2];
          if (vec) {
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const long long r = rr + 8 * u;
                if (r < r1) {
                    la[u] = *reinterpret_cast<const float4*>(g + (unsigned long long)r * cols + c0);
                    lb[u] = *reinterpret_cast<const float4*>(g + (unsigned long long)r * cols + c0 + 4);
                }
            }
          }
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const long long r = rr + 8 * u;
            if (r >= r1) continue;
            const unsigned long long e0 = (unsigned long long)r * cols + c0;
            float v[8];
            if (vec) {
                const float4 a = la[u];
                const float4 b = lb[u];
                v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
                if (drop.p > 0.f) {
                    float m[8];
                    dropout_mult8(drop, e0 >> 3, m);
#pragma unroll
                    for (int j = 0; j < 8; ++j) v[j] *= m[j];
                }
                uint4 o;
                o.x = pack_bf16(v[0], v[1]); o.y = pack_bf16(v[2], v[3]);
                o.z = pack_bf16(v[4], v[5]); o.w = pack_bf16(v[6], v[7]);
                *reinterpret_cast<uint4*>(out + e0) = o;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    // sum what the GEMM will actually consume (the bf16-rounded value)
                    acc[j] += __bfloat162float(__float2bfloat16(v[j]));
                }
            } else {
                for (int j = 0; j < 8 && c0 + j < cols; ++j) {
                    float t = g[e0 + j];
                    if (drop.p > 0.f) t *= dropout_mult1(drop, e0 + j);
                    const bf16 h = __float2bfloat16(t);
                    out[e0 + j] = h;
                    acc[j] += __bfloat162float(h);
                }
            }
          }
        }
    }
    if (colsum == nullptr) return;
#pragma unroll
    for (int j = 0; j < 8; ++j) red[threadIdx.y][threadIdx.x][j] = acc[j];
    __syncthreads();
    if (threadIdx.y == 0 && c0 < cols) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float s = 0.f;
#pragma unroll
            for (int y = 0; y < 8; ++y) s += red[y][threadIdx.x][j];
            if (c0 + j < cols) atomicAdd(colsum + c0 + j, s);
        }
    }
}

int dropout_cast_launch(const float* g, bf16* out, float* colsum, long long rows, int cols, float drop_p,
                        unsigned long long seed, unsigned long long offset, cudaStream_t stream) {
    B200_CHECK_ARG(g && out, "dropout_cast: null pointer");
    B200_CHECK_ARG(rows > 0 && cols > 0, "dropout_cast: bad sizes");
    if (colsum) B200_CUDA(cudaMemsetAsync(colsum, 0, sizeof(float) * cols, stream));
    dim3 grid((cols + 255) / 256, (unsigned)((rows + DC_ROWS_PER_BLOCK - 1) / DC_ROWS_PER_BLOCK));
    dropout_cast_kernel<<<grid, dim3(32, 8), 0, stream>>>(g, out, colsum, rows, cols,
                                                         make_dropout(drop_p, seed, offset));
    B200_LAUNCH_CHECK();
    return 0;
}

// dst = src * (*scalar_dev): backward of the mean cross-entropy (chain rule with the upstream scalar).
__global__ void scale_by_scalar_kernel(const float* __restrict__ src, const float* __restrict__ s,
                                       float* __restrict__ dst, long long n) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[i] * __ldg(s);
}

int scale_by_scalar_launch(const float* src, const float* scalar_dev, float* dst, long long n, cudaStream_t stream) {
    B200_CHECK_ARG(src && scalar_dev && dst, "scale_by_scalar: null pointer");
    if (n <= 0) return 0;
    scale_by_scalar_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(src, scalar_dev, dst, n);
    B200_LAUNCH_CHECK();
    return 0;
}

}  // namespace b200vit
