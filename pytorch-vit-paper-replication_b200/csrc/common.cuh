// Shared device/host helpers for libb200vit (sm_100a only).
//
// Everything in csrc/ is new code: the reference (models/vit.py, going_modular/engine.py)
// has no native component; the arithmetic it delegates to torch.nn is re-implemented here.
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace b200vit {

typedef __nv_bfloat16 bf16;

// ---------------------------------------------------------------------------------------------
// Error convention of the C ABI: every entry returns int (0 ok, <0 argument error, >0 cudaError_t)
// and leaves a message in a thread-local buffer read by b200vit_last_error().
// ---------------------------------------------------------------------------------------------
char* last_error_buf();
int set_error(int code, const char* fmt, ...);

#define B200_CHECK_ARG(cond, ...)                          \
    do {                                                   \
        if (!(cond)) return b200vit::set_error(-1, __VA_ARGS__); \
    } while (0)

#define B200_CUDA(expr)                                                                     \
    do {                                                                                    \
        cudaError_t _e = (expr);                                                            \
        if (_e != cudaSuccess)                                                              \
            return b200vit::set_error((int)_e, "%s failed: %s (%s:%d)", #expr,              \
                                      cudaGetErrorString(_e), __FILE__, __LINE__);          \
    } while (0)

#define B200_LAUNCH_CHECK() B200_CUDA(cudaGetLastError())

int sm_count();  // SMs of the current device (cached per device)

// ---------------------------------------------------------------------------------------------
// Small device helpers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float2 unpack_bf16(uint32_t v) {
    __nv_bfloat162 t = *reinterpret_cast<__nv_bfloat162*>(&v);
    return __bfloat1622float2(t);
}

// Exact (erf) GELU and its derivative: nn.GELU() default, reference models/vit.py:121.
__device__ __forceinline__ float gelu_exact(float x) {
    return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f));
}
__device__ __forceinline__ float gelu_exact_grad(float x) {
    const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752f));
    const float pdf = 0.3989422804014327f * __expf(-0.5f * x * x);
    return cdf + x * pdf;
}

// ---------------------------------------------------------------------------------------------
// Counter-based dropout mask (Philox4x32-10). One Philox block yields 4x32 random bits = eight
// 16-bit lanes; element e of a tensor uses block e/8, lane e%8, so any kernel (forward epilogue,
// backward epilogue, standalone) regenerates the same mask from (seed, offset) alone and no
// mask tensor is ever stored. keep  <=>  lane16 >= thr16 where thr16 = round(p * 65536).
// Semantics mirror nn.Dropout (models/vit.py:43,122,125): kept values are scaled by 1/(1-p).
// ---------------------------------------------------------------------------------------------
struct DropoutParams {
    float p;            // drop probability; 0 disables
    float scale;        // 1/(1-p)
    uint32_t thr16;     // drop if lane16 < thr16
    uint64_t seed;
    uint64_t offset;    // distinguishes dropout sites / steps
};

__device__ __forceinline__ uint4 philox4x32_10(uint64_t seed, uint64_t offset, uint64_t ctr) {
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
    uint32_t c0 = (uint32_t)ctr, c1 = (uint32_t)(ctr >> 32);
    uint32_t c2 = (uint32_t)offset, c3 = (uint32_t)(offset >> 32);
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
}

// Multipliers (0 or scale) for the 8 consecutive elements [8*blk, 8*blk+8).
__device__ __forceinline__ void dropout_mult8(const DropoutParams& d, uint64_t blk, float (&m)[8]) {
    const uint4 r = philox4x32_10(d.seed, d.offset, blk);
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        m[2 * i]     = ((w[i] & 0xFFFFu) >= d.thr16) ? d.scale : 0.0f;
        m[2 * i + 1] = ((w[i] >> 16)     >= d.thr16) ? d.scale : 0.0f;
    }
}
// Multiplier for a single element (slow path, used by small kernels).
__device__ __forceinline__ float dropout_mult1(const DropoutParams& d, uint64_t elem) {
    const uint4 r = philox4x32_10(d.seed, d.offset, elem >> 3);
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
    const uint32_t lane = (uint32_t)(elem & 7);
    const uint32_t v = (lane & 1) ? (w[lane >> 1] >> 16) : (w[lane >> 1] & 0xFFFFu);
    return (v >= d.thr16) ? d.scale : 0.0f;
}

inline DropoutParams make_dropout(float p, uint64_t seed, uint64_t offset) {
    DropoutParams d;
    d.p = p;
    d.scale = (p > 0.f && p < 1.f) ? 1.0f / (1.0f - p) : (p >= 1.f ? 0.f : 1.0f);
    double t = (double)p * 65536.0 + 0.5;
    d.thr16 = p <= 0.f ? 0u : (t >= 65536.0 ? 65536u : (uint32_t)t);
    d.seed = seed;
    d.offset = offset;
    return d;
}

}  // namespace b200vit
