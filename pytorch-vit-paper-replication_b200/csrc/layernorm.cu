// LayerNorm forward / backward (nn.LayerNorm(D), eps 1e-5, biased variance: reference models/vit.py:83,115,216).
// HBM-bound: one warp per row, float4 accesses, warp-shuffle row reductions; the backward keeps the
// dgamma / dbeta (and optional next-bias) column partials in per-warp shared-memory slabs across the rows a
// warp visits and finishes with one cross-warp reduction + one atomic per column per block.
#include "common.cuh"
#include "launchers.h"

namespace b200vit {

constexpr int LN_WARPS = 8;

// ---------------------------------------------------------------------------------------------
// forward: y = (x - mean) * rstd * gamma + beta ; writes bf16 (GEMM operand) and/or fp32, plus row stats
// ---------------------------------------------------------------------------------------------
template <int NV>
__global__ void __launch_bounds__(LN_WARPS * 32)
layernorm_fwd_kernel(const float* __restrict__ x, long long row_stride, const float* __restrict__ gamma,
                     const float* __restrict__ beta, bf16* __restrict__ y_bf16, float* __restrict__ y_f32,
                     float* __restrict__ mean_out, float* __restrict__ rstd_out, long long rows, int D, float eps) {
    pdl_launch_dependents();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nv = D >> 2;
    pdl_wait();
    for (long long row = (long long)blockIdx.x * LN_WARPS + warp; row < rows; row += (long long)gridDim.x * LN_WARPS) {
        const float4* xr = reinterpret_cast<const float4*>(x + row * row_stride);
        float4 v[NV];                       // the whole row lives in registers: one global read
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            const int i = lane + 32 * k;
            v[k] = i < nv ? xr[i] : make_float4(0, 0, 0, 0);
            s += (v[k].x + v[k].y) + (v[k].z + v[k].w);
        }
        const float mean = warp_sum(s) / (float)D;
        float q = 0.f;
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            const int i = lane + 32 * k;
            if (i < nv) {
                const float a = v[k].x - mean, b = v[k].y - mean, c = v[k].z - mean, d = v[k].w - mean;
                q += (a * a + b * b) + (c * c + d * d);
            }
        }
        const float var = warp_sum(q) / (float)D;
        const float rstd = rsqrtf(var + eps);
        if (lane == 0) {
            if (mean_out) mean_out[row] = mean;
            if (rstd_out) rstd_out[row] = rstd;
        }
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            const int i = lane + 32 * k;
            if (i < nv) {
                const float4 g = __ldg(reinterpret_cast<const float4*>(gamma) + i);
                const float4 b = __ldg(reinterpret_cast<const float4*>(beta) + i);
                float4 o;
                o.x = (v[k].x - mean) * rstd * g.x + b.x;
                o.y = (v[k].y - mean) * rstd * g.y + b.y;
                o.z = (v[k].z - mean) * rstd * g.z + b.z;
                o.w = (v[k].w - mean) * rstd * g.w + b.w;
                if (y_f32) reinterpret_cast<float4*>(y_f32 + row * (long long)D)[i] = o;
                if (y_bf16) {
                    uint2 p;
                    p.x = pack_bf16(o.x, o.y);
                    p.y = pack_bf16(o.z, o.w);
                    reinterpret_cast<uint2*>(y_bf16 + row * (long long)D)[i] = p;
                }
            }
        }
    }
}

int layernorm_fwd_launch(const float* x, long long row_stride, const float* gamma, const float* beta, bf16* y_bf16,
                         float* y_f32, float* mean, float* rstd, long long rows, int dim, float eps,
                         cudaStream_t stream) {
    B200_CHECK_ARG(x && gamma && beta, "layernorm_fwd: null pointer");
    B200_CHECK_ARG(y_bf16 || y_f32, "layernorm_fwd: no output requested");
    B200_CHECK_ARG(rows > 0 && dim > 0 && (dim % 4) == 0, "layernorm_fwd: dim must be a positive multiple of 4 (got %d)", dim);
    B200_CHECK_ARG((row_stride % 4) == 0 && row_stride >= dim, "layernorm_fwd: row stride %lld must be a multiple of 4 and >= dim", row_stride);
    B200_CHECK_ARG(((uintptr_t)x % 16) == 0 && ((uintptr_t)gamma % 16) == 0 && ((uintptr_t)beta % 16) == 0,
                   "layernorm_fwd: pointers must be 16-byte aligned");
    long long blocks = (rows + LN_WARPS - 1) / LN_WARPS;
    const long long cap = (long long)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    const int nvl = (dim / 4 + 31) / 32;
    B200_CHECK_ARG(nvl <= 16, "layernorm_fwd: dim must be <= 2048 (got %d)", dim);
#define LN_FWD_CASE(NV) B200_CUDA(launch_kernel_pdl(layernorm_fwd_kernel<NV>, dim3((unsigned)blocks), dim3(LN_WARPS * 32), (size_t)0, stream, \
        x, row_stride, gamma, beta, y_bf16, y_f32, mean, rstd, rows, dim, eps))
    if (nvl <= 1) LN_FWD_CASE(1);
    else if (nvl <= 2) LN_FWD_CASE(2);
    else if (nvl <= 4) LN_FWD_CASE(4);
    else if (nvl <= 6) LN_FWD_CASE(6);
    else if (nvl <= 8) LN_FWD_CASE(8);
    else LN_FWD_CASE(16);
#undef LN_FWD_CASE
    B200_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------------------------
// backward. With xhat = (x - mean) * rstd and dyg = dy * gamma:
//   dx = rstd * (dyg - mean_D(dyg) - xhat * mean_D(dyg * xhat))
//   gout = gin + dx                         (gin = gradient arriving through the residual connection)
//   dgamma += sum_rows dy * xhat ; dbeta += sum_rows dy
//   gout_bf16 = bf16(dropout_mask * gout) ; dbias_next += sum_rows gout_bf16   (operand + bias gradient of the
//                                                                              Linear that produced this residual)
// NV = float4 chunks per lane (ceil(D / 128)).
// ---------------------------------------------------------------------------------------------
template <int NV, bool HAS_NEXT>
__global__ void __launch_bounds__(LN_WARPS * 32, (NV <= 6) ? 3 : 2)
layernorm_bwd_kernel(const bf16* __restrict__ dy_bf16, const float* __restrict__ dy_f32, const float* __restrict__ x,
                     long long row_stride, const float* __restrict__ mean_in, const float* __restrict__ rstd_in,
                     const float* __restrict__ gamma, const float* __restrict__ gin, float* __restrict__ gout,
                     long long gout_stride, bf16* __restrict__ gout_bf16, float* __restrict__ dgamma,
                     float* __restrict__ dbeta, float* __restrict__ dbias_next, long long rows, int D,
                     DropoutParams drop) {
    // Column accumulators live in shared memory, one private slab per warp ([acc][warp][D]): keeping them in
    // registers (24 values per lane per accumulator at D=768) starved the kernel of occupancy.
    extern __shared__ float sacc[];
    pdl_launch_dependents();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nv = D >> 2;
    float4* acc_g = reinterpret_cast<float4*>(sacc + (size_t)(0 * LN_WARPS + warp) * D);
    float4* acc_b = reinterpret_cast<float4*>(sacc + (size_t)(1 * LN_WARPS + warp) * D);
    float4* acc_n = reinterpret_cast<float4*>(sacc + (size_t)(2 * LN_WARPS + warp) * D);
    for (int i = lane; i < nv; i += 32) {
        acc_g[i] = make_float4(0, 0, 0, 0);
        acc_b[i] = make_float4(0, 0, 0, 0);
        if (HAS_NEXT) acc_n[i] = make_float4(0, 0, 0, 0);
    }
    __syncwarp();
    pdl_wait();       // accumulators were cleared under the previous kernel's tail; global memory from here on
    const float4* gamma4 = reinterpret_cast<const float4*>(gamma);
    const bool want_cols = dgamma != nullptr || dbeta != nullptr;

    for (long long row = (long long)blockIdx.x * LN_WARPS + warp; row < rows; row += (long long)gridDim.x * LN_WARPS) {
        const float mean = mean_in[row], rstd = rstd_in[row];
        const float4* xr = reinterpret_cast<const float4*>(x + row * row_stride);
        float4 xh[NV], dy[NV];
        // issue the x / dy loads of the row up front (the residual gradient is fetched after the row reductions:
        // keeping it in registers too would cost the third resident block per SM)
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            const int i = lane + 32 * k;
            if (i < nv) {
                xh[k] = xr[i];
                if (dy_bf16) {
                    const uint2 q = reinterpret_cast<const uint2*>(dy_bf16 + row * (long long)D)[i];
                    const float2 a = unpack_bf16(q.x), b = unpack_bf16(q.y);
                    dy[k] = make_float4(a.x, a.y, b.x, b.y);
                } else {
                    dy[k] = reinterpret_cast<const float4*>(dy_f32 + row * (long long)D)[i];
                }
            }
        }
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            const int i = lane + 32 * k;
            if (i < nv) {
                const float4 v = xh[k];
                xh[k] = make_float4((v.x - mean) * rstd, (v.y - mean) * rstd, (v.z - mean) * rstd, (v.w - mean) * rstd);
                const float4 g = __ldg(gamma4 + i);
                const float a0 = dy[k].x * g.x, a1 = dy[k].y * g.y, a2 = dy[k].z * g.z, a3 = dy[k].w * g.w;
                s1 += (a0 + a1) + (a2 + a3);
                s2 += (a0 * xh[k].x + a1 * xh[k].y) + (a2 * xh[k].z + a3 * xh[k].w);
                if (want_cols) {
                    float4 ag = acc_g[i], ab = acc_b[i];
                    ag.x += dy[k].x * xh[k].x; ag.y += dy[k].y * xh[k].y; ag.z += dy[k].z * xh[k].z; ag.w += dy[k].w * xh[k].w;
                    ab.x += dy[k].x; ab.y += dy[k].y; ab.z += dy[k].z; ab.w += dy[k].w;
                    acc_g[i] = ag;
                    acc_b[i] = ab;
                }
            }
        }
        const float c1 = warp_sum(s1) / (float)D;
        const float c2 = warp_sum(s2) / (float)D;
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            const int i = lane + 32 * k;
            if (i < nv) {
                const float4 g = __ldg(gamma4 + i);
                const float4 rs = gin ? reinterpret_cast<const float4*>(gin + row * gout_stride)[i] : make_float4(0, 0, 0, 0);
                float4 o;
                o.x = rstd * (dy[k].x * g.x - c1 - xh[k].x * c2) + rs.x;
                o.y = rstd * (dy[k].y * g.y - c1 - xh[k].y * c2) + rs.y;
                o.z = rstd * (dy[k].z * g.z - c1 - xh[k].z * c2) + rs.z;
                o.w = rstd * (dy[k].w * g.w - c1 - xh[k].w * c2) + rs.w;
                reinterpret_cast<float4*>(gout + row * gout_stride)[i] = o;
                if (HAS_NEXT) {
                    float v[4] = {o.x, o.y, o.z, o.w};
                    if (drop.p > 0.f) {
                        // elements e0 .. e0+3 (e0 % 4 == 0) sit in two consecutive 32-bit words of one mask block:
                        // two hashes per float4 instead of one per element (same decisions as dropout_mult1)
                        const unsigned long long e0 = (unsigned long long)row * D + 4ull * i;
                        const uint32_t j0 = (uint32_t)(e0 & 7) >> 1;
                        const uint32_t w0 = dropout_word(drop, e0 >> 3, j0), w1 = dropout_word(drop, e0 >> 3, j0 + 1);
                        v[0] *= ((w0 & 0x7FFFu) >= drop.thr15) ? drop.scale : 0.f;
                        v[1] *= (((w0 >> 16) & 0x7FFFu) >= drop.thr15) ? drop.scale : 0.f;
                        v[2] *= ((w1 & 0x7FFFu) >= drop.thr15) ? drop.scale : 0.f;
                        v[3] *= (((w1 >> 16) & 0x7FFFu) >= drop.thr15) ? drop.scale : 0.f;
                    }
                    uint2 pk;
                    pk.x = pack_bf16(v[0], v[1]);
                    pk.y = pack_bf16(v[2], v[3]);
                    reinterpret_cast<uint2*>(gout_bf16 + row * (long long)D)[i] = pk;
                    const float2 r0 = unpack_bf16(pk.x), r1 = unpack_bf16(pk.y);
                    float4 an = acc_n[i];
                    an.x += r0.x; an.y += r0.y; an.z += r1.x; an.w += r1.y;
                    acc_n[i] = an;
                }
            }
        }
    }

    __syncthreads();
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
        float sg = 0.f, sb = 0.f, sn = 0.f;
#pragma unroll
        for (int w = 0; w < LN_WARPS; ++w) {
            sg += sacc[(size_t)(0 * LN_WARPS + w) * D + d];
            sb += sacc[(size_t)(1 * LN_WARPS + w) * D + d];
            if (HAS_NEXT) sn += sacc[(size_t)(2 * LN_WARPS + w) * D + d];
        }
        if (dgamma) atomicAdd(dgamma + d, sg);
        if (dbeta) atomicAdd(dbeta + d, sb);
        if (HAS_NEXT && dbias_next) atomicAdd(dbias_next + d, sn);
    }
}

// Wide rows (D > 768, e.g. the 1024 of ViT-L): the one-warp-per-row kernel above needs ~116 registers and 96 KB of
// accumulator slabs per block there, i.e. 2 resident blocks per SM and half the achieved bandwidth. Here TWO warps
// share a row (each owns one half of the columns: half the registers, half the slab) and exchange their partial row
// sums through shared memory with one 64-thread named barrier per row (exchange slots alternate by row parity).
template <int NV, bool HAS_NEXT>
__global__ void __launch_bounds__(LN_WARPS * 32, NV <= 3 ? 4 : 3)
layernorm_bwd_split_kernel(const bf16* __restrict__ dy_bf16, const float* __restrict__ dy_f32, const float* __restrict__ x,
                     long long row_stride, const float* __restrict__ mean_in, const float* __restrict__ rstd_in,
                     const float* __restrict__ gamma, const float* __restrict__ gin, float* __restrict__ gout,
                     long long gout_stride, bf16* __restrict__ gout_bf16, float* __restrict__ dgamma,
                     float* __restrict__ dbeta, float* __restrict__ dbias_next, long long rows, int D,
                     DropoutParams drop) {
    // Column accumulators live in shared memory, one private slab per warp ([acc][warp][D]): keeping them in
    // registers (24 values per lane per accumulator at D=768) starved the kernel of occupancy.
    extern __shared__ float sacc[];
    pdl_launch_dependents();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int pair = warp >> 1, half = warp & 1;
    const int Dh = D >> 1;              // columns owned by this warp: [half * Dh, half * Dh + Dh)
    const int c0 = half * Dh;
    const int nv = Dh >> 2;
    float4* acc_g = reinterpret_cast<float4*>(sacc + (size_t)(0 * LN_WARPS + warp) * Dh);
    float4* acc_b = reinterpret_cast<float4*>(sacc + (size_t)(1 * LN_WARPS + warp) * Dh);
    float4* acc_n = reinterpret_cast<float4*>(sacc + (size_t)(2 * LN_WARPS + warp) * Dh);
    float* xch = sacc + (size_t)(HAS_NEXT ? 3 : 2) * LN_WARPS * Dh;     // [parity][pair][half][s1, s2]
    for (int i = lane; i < nv; i += 32) {
        acc_g[i] = make_float4(0, 0, 0, 0);
        acc_b[i] = make_float4(0, 0, 0, 0);
        if (HAS_NEXT) acc_n[i] = make_float4(0, 0, 0, 0);
    }
    __syncwarp();
    pdl_wait();       // accumulators were cleared under the previous kernel's tail; global memory from here on
    const float4* gamma4 = reinterpret_cast<const float4*>(gamma + c0);
    const bool want_cols = dgamma != nullptr || dbeta != nullptr;

    int par = 0;
    for (long long row = (long long)blockIdx.x * (LN_WARPS / 2) + pair; row < rows;
         row += (long long)gridDim.x * (LN_WARPS / 2), par ^= 1) {
        const float mean = mean_in[row], rstd = rstd_in[row];
        const float4* xr = reinterpret_cast<const float4*>(x + row * row_stride + c0);
        float4 xh[NV], dy[NV];
        // issue the x / dy loads of the row up front (the residual gradient is fetched after the row reductions:
        // keeping it in registers too would cost the third resident block per SM)
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            const int i = lane + 32 * k;
            if (i < nv) {
                xh[k] = xr[i];
                if (dy_bf16) {
                    const uint2 q = reinterpret_cast<const uint2*>(dy_bf16 + row * (long long)D + c0)[i];
                    const float2 a = unpack_bf16(q.x), b = unpack_bf16(q.y);
                    dy[k] = make_float4(a.x, a.y, b.x, b.y);
                } else {
                    dy[k] = reinterpret_cast<const float4*>(dy_f32 + row * (long long)D + c0)[i];
                }
            }
        }
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            const int i = lane + 32 * k;
            if (i < nv) {
                const float4 v = xh[k];
                xh[k] = make_float4((v.x - mean) * rstd, (v.y - mean) * rstd, (v.z - mean) * rstd, (v.w - mean) * rstd);
                const float4 g = __ldg(gamma4 + i);
                const float a0 = dy[k].x * g.x, a1 = dy[k].y * g.y, a2 = dy[k].z * g.z, a3 = dy[k].w * g.w;
                s1 += (a0 + a1) + (a2 + a3);
                s2 += (a0 * xh[k].x + a1 * xh[k].y) + (a2 * xh[k].z + a3 * xh[k].w);
                if (want_cols) {
                    float4 ag = acc_g[i], ab = acc_b[i];
                    ag.x += dy[k].x * xh[k].x; ag.y += dy[k].y * xh[k].y; ag.z += dy[k].z * xh[k].z; ag.w += dy[k].w * xh[k].w;
                    ab.x += dy[k].x; ab.y += dy[k].y; ab.z += dy[k].z; ab.w += dy[k].w;
                    acc_g[i] = ag;
                    acc_b[i] = ab;
                }
            }
        }
        s1 = warp_sum(s1);
        s2 = warp_sum(s2);
        float* mine = xch + ((par * (LN_WARPS / 2) + pair) * 2 + half) * 2;
        const float* other = xch + ((par * (LN_WARPS / 2) + pair) * 2 + (half ^ 1)) * 2;
        if (lane == 0) { mine[0] = s1; mine[1] = s2; }
        asm volatile("bar.sync %0, 64;" ::"r"(1 + pair) : "memory");
        const float c1 = (s1 + other[0]) / (float)D;
        const float c2 = (s2 + other[1]) / (float)D;
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            const int i = lane + 32 * k;
            if (i < nv) {
                const float4 g = __ldg(gamma4 + i);
                const float4 rs = gin ? reinterpret_cast<const float4*>(gin + row * gout_stride + c0)[i] : make_float4(0, 0, 0, 0);
                float4 o;
                o.x = rstd * (dy[k].x * g.x - c1 - xh[k].x * c2) + rs.x;
                o.y = rstd * (dy[k].y * g.y - c1 - xh[k].y * c2) + rs.y;
                o.z = rstd * (dy[k].z * g.z - c1 - xh[k].z * c2) + rs.z;
                o.w = rstd * (dy[k].w * g.w - c1 - xh[k].w * c2) + rs.w;
                reinterpret_cast<float4*>(gout + row * gout_stride + c0)[i] = o;
                if (HAS_NEXT) {
                    float v[4] = {o.x, o.y, o.z, o.w};
                    if (drop.p > 0.f) {
                        // elements e0 .. e0+3 (e0 % 4 == 0) sit in two consecutive 32-bit words of one mask block:
                        // two hashes per float4 instead of one per element (same decisions as dropout_mult1)
                        const unsigned long long e0 = (unsigned long long)row * D + c0 + 4ull * i;
                        const uint32_t j0 = (uint32_t)(e0 & 7) >> 1;
                        const uint32_t w0 = dropout_word(drop, e0 >> 3, j0), w1 = dropout_word(drop, e0 >> 3, j0 + 1);
                        v[0] *= ((w0 & 0x7FFFu) >= drop.thr15) ? drop.scale : 0.f;
                        v[1] *= (((w0 >> 16) & 0x7FFFu) >= drop.thr15) ? drop.scale : 0.f;
                        v[2] *= ((w1 & 0x7FFFu) >= drop.thr15) ? drop.scale : 0.f;
                        v[3] *= (((w1 >> 16) & 0x7FFFu) >= drop.thr15) ? drop.scale : 0.f;
                    }
                    uint2 pk;
                    pk.x = pack_bf16(v[0], v[1]);
                    pk.y = pack_bf16(v[2], v[3]);
                    reinterpret_cast<uint2*>(gout_bf16 + row * (long long)D + c0)[i] = pk;
                    const float2 r0 = unpack_bf16(pk.x), r1 = unpack_bf16(pk.y);
                    float4 an = acc_n[i];
                    an.x += r0.x; an.y += r0.y; an.z += r1.x; an.w += r1.y;
                    acc_n[i] = an;
                }
            }
        }
    }

    __syncthreads();
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
        float sg = 0.f, sb = 0.f, sn = 0.f;
        const int hh = d >= Dh ? 1 : 0, dd = d - hh * Dh;
#pragma unroll
        for (int w = hh; w < LN_WARPS; w += 2) {        // the warps that own this column's half
            sg += sacc[(size_t)(0 * LN_WARPS + w) * Dh + dd];
            sb += sacc[(size_t)(1 * LN_WARPS + w) * Dh + dd];
            if (HAS_NEXT) sn += sacc[(size_t)(2 * LN_WARPS + w) * Dh + dd];
        }
        if (dgamma) atomicAdd(dgamma + d, sg);
        if (dbeta) atomicAdd(dbeta + d, sb);
        if (HAS_NEXT && dbias_next) atomicAdd(dbias_next + d, sn);
    }
}

static const bool B200VIT_LN_SPLIT = [] { const char* e = getenv("B200VIT_LN_SPLIT"); return e ? atoi(e) != 0 : true; }();

int layernorm_bwd_launch(const bf16* dy_bf16, const float* dy_f32, const float* x, long long row_stride,
                         const float* mean, const float* rstd, const float* gamma, const float* gin, float* gout,
                         long long gout_stride, bf16* gout_bf16, float* dgamma, float* dbeta, float* dbias_next,
                         long long rows, int dim, float drop_p, unsigned long long seed, unsigned long long offset,
                         cudaStream_t stream) {
    B200_CHECK_ARG((dy_bf16 != nullptr) != (dy_f32 != nullptr), "layernorm_bwd: pass exactly one of dy_bf16 / dy_f32");
    B200_CHECK_ARG(x && mean && rstd && gamma && gout, "layernorm_bwd: null pointer");
    B200_CHECK_ARG(rows > 0 && dim > 0 && (dim % 4) == 0 && dim <= 1024,
                   "layernorm_bwd: dim must be a multiple of 4 and <= 1024 (got %d)", dim);
    B200_CHECK_ARG((row_stride % 4) == 0 && (gout_stride % 4) == 0, "layernorm_bwd: strides must be multiples of 4");
    // dgamma / dbeta / dbias_next are accumulated with atomics: start from zero (outputs are overwritten).
    if (dgamma) B200_CUDA(cudaMemsetAsync(dgamma, 0, sizeof(float) * dim, stream));
    if (dbeta) B200_CUDA(cudaMemsetAsync(dbeta, 0, sizeof(float) * dim, stream));
    if (dbias_next) B200_CUDA(cudaMemsetAsync(dbias_next, 0, sizeof(float) * dim, stream));
    const int nvl = (dim / 4 + 31) / 32;
    long long blocks = (rows + LN_WARPS - 1) / LN_WARPS;
    const long long cap = (long long)sm_count() * ((dim % 8) == 0 && nvl >= 4 && nvl <= 6 && B200VIT_LN_SPLIT ? 4 : 3);   // persistent: 3 (4) resident blocks per SM, rows strided
    if (blocks > cap) blocks = cap;
    const DropoutParams drop = make_dropout(drop_p, seed, offset);
    const size_t smem = sizeof(float) * (gout_bf16 ? 3 : 2) * LN_WARPS * dim;
#define LN_BWD_CASE(NV)                                                                                         \
    do {                                                                                                        \
        B200_CUDA(cudaFuncSetAttribute(layernorm_bwd_kernel<NV, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * 4 * LN_WARPS * 1024)); \
        B200_CUDA(cudaFuncSetAttribute(layernorm_bwd_kernel<NV, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 4 * LN_WARPS * 1024)); \
        if (gout_bf16)                                                                                          \
            B200_CUDA(launch_kernel_pdl(layernorm_bwd_kernel<NV, true>, dim3((unsigned)blocks), dim3(LN_WARPS * 32), smem, stream, \
                dy_bf16, dy_f32, x, row_stride, mean, rstd, gamma, gin, gout, gout_stride, gout_bf16, dgamma,   \
                dbeta, dbias_next, rows, dim, drop));                                                           \
        else                                                                                                    \
            B200_CUDA(launch_kernel_pdl(layernorm_bwd_kernel<NV, false>, dim3((unsigned)blocks), dim3(LN_WARPS * 32), smem, stream, \
                dy_bf16, dy_f32, x, row_stride, mean, rstd, gamma, gin, gout, gout_stride, gout_bf16, dgamma,   \
                dbeta, dbias_next, rows, dim, drop));                                                           \
    } while (0)
#define LN_BWD_SPLIT_CASE(NVH)                                                                                  \
    do {                                                                                                        \
        long long sblocks = (rows + LN_WARPS / 2 - 1) / (LN_WARPS / 2);                                         \
        if (sblocks > cap) sblocks = cap;                                                                       \
        const size_t ssmem = sizeof(float) * ((gout_bf16 ? 3 : 2) * LN_WARPS * (dim / 2) + 2 * (LN_WARPS / 2) * 2 * 2); \
        B200_CUDA(cudaFuncSetAttribute(layernorm_bwd_split_kernel<NVH, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * 4 * LN_WARPS * 512 + 256)); \
        B200_CUDA(cudaFuncSetAttribute(layernorm_bwd_split_kernel<NVH, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 4 * LN_WARPS * 512 + 256)); \
        if (gout_bf16)                                                                                          \
            B200_CUDA(launch_kernel_pdl(layernorm_bwd_split_kernel<NVH, true>, dim3((unsigned)sblocks), dim3(LN_WARPS * 32), ssmem, stream, \
                dy_bf16, dy_f32, x, row_stride, mean, rstd, gamma, gin, gout, gout_stride, gout_bf16, dgamma,   \
                dbeta, dbias_next, rows, dim, drop));                                                           \
        else                                                                                                    \
            B200_CUDA(launch_kernel_pdl(layernorm_bwd_split_kernel<NVH, false>, dim3((unsigned)sblocks), dim3(LN_WARPS * 32), ssmem, stream, \
                dy_bf16, dy_f32, x, row_stride, mean, rstd, gamma, gin, gout, gout_stride, gout_bf16, dgamma,   \
                dbeta, dbias_next, rows, dim, drop));                                                           \
    } while (0)
    // rows of >= 512 columns: two warps per row (layernorm_bwd_split_kernel) -- fewer registers and a smaller
    // accumulator slab per warp, more bytes in flight per SM (D = 768: 4.1 -> 4.6 TB/s; D = 1024: 2.3 -> 4.6 TB/s)
    const bool split = (dim % 8) == 0 && nvl >= 4 && B200VIT_LN_SPLIT;
    if (split && nvl <= 4) LN_BWD_SPLIT_CASE(2);
    else if (split && nvl <= 6) LN_BWD_SPLIT_CASE(3);
    else if (split) LN_BWD_SPLIT_CASE(4);
    else if (nvl <= 1) LN_BWD_CASE(1);
    else if (nvl <= 2) LN_BWD_CASE(2);
    else if (nvl <= 4) LN_BWD_CASE(4);
    else if (nvl <= 6) LN_BWD_CASE(6);
    else LN_BWD_CASE(8);
#undef LN_BWD_SPLIT_CASE
#undef LN_BWD_CASE
    B200_LAUNCH_CHECK();
    return 0;
}

}  // namespace b200vit
