// Shared device/host helpers for libb200vit (sm_100a only).
//
// Everything in csrc/ is new code: the reference (models/vit.py, going_modular/engine.py)
// has no native component; the arithmetic it delegates to torch.nn is re-implemented here.
#pragma once
#include <stdlib.h>

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

// erf-form GELU and its derivative (nn.GELU() default, reference models/vit.py:121). erf is evaluated with an
// Abramowitz-Stegun rational form: one MUFU.RCP + one MUFU.EX2 + a short Horner polynomial, no branches. With
// u = |x|/sqrt(2),  E = exp(-u^2) = exp(-x^2/2),  erf(u) = 1 - poly(t) E,  t = 1/(1 + p u)
// and the same E gives the Gaussian density needed by the derivative.
__device__ __forceinline__ float fast_rcp(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float fast_ex2(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
// erf by Abramowitz & Stegun 7.1.25 (three coefficients, |error| <= 2.5e-5): every consumer rounds the result to bf16
// (half an ulp is 2e-3 relative), so the five-coefficient 7.1.26 (1.5e-7) bought nothing for two more FMAs and a
// multiply per element in the most instruction-bound epilogue of the step. u = |x|/sqrt2 is folded into the constants.
__device__ __forceinline__ void gelu_terms(float x, float& one_minus_pe, float& e) {
    const float t = fast_rcp(fmaf(0.47047f * 0.70710678118654752f, fabsf(x), 1.0f));
    float poly = fmaf(t, 0.7478556f, -0.0958798f);
    poly = fmaf(poly, t, 0.3480242f);
    poly *= t;
    e = fast_ex2(x * x * (-0.5f * 1.4426950408889634f));     // exp(-x^2 / 2)
    one_minus_pe = fmaf(-poly, e, 1.0f);          // erf(|x|/sqrt2)
}
// h = s * gelu(x), g = s * gelu'(x) from one set of shared terms (s = dropout scale 1/(1-p), folded into constants).
__device__ __forceinline__ void gelu_value_and_grad(float x, float s, float& h, float& g) {
    float erf_abs, e;
    gelu_terms(x, erf_abs, e);
    const float hs = 0.5f * s;
    h = hs * fmaf(fabsf(x), erf_abs, x);
    g = fmaf(x * (0.3989422804014327f * s), e, fmaf(copysignf(hs, x), erf_abs, hs));
}
__device__ __forceinline__ float gelu_exact(float x) {
    float erf_abs, e;
    gelu_terms(x, erf_abs, e);
    return 0.5f * fmaf(fabsf(x), erf_abs, x);      // 0.5 x (1 + erf(x/sqrt2))
}
__device__ __forceinline__ float gelu_exact_grad(float x) {
    float erf_abs, e;
    gelu_terms(x, erf_abs, e);
    const float cdf = fmaf(copysignf(0.5f, x), erf_abs, 0.5f);
    return fmaf(x * 0.3989422804014327f, e, cdf);  // cdf + x pdf
}

// ---------------------------------------------------------------------------------------------
// Packed fp32 arithmetic (sm_100: FFMA2 / FMUL2 / FADD2 operate on an aligned register pair per issue slot). The GELU
// epilogue of the fc1 GEMM is bound by instruction issue, not by the tensor pipe: evaluating the erf polynomial,
// the value and the derivative for TWO adjacent columns per instruction halves its floating-point issue count.
// ---------------------------------------------------------------------------------------------
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
    f32x2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
    f32x2 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
// Constants of gelu_value_and_grad2, pre-multiplied by the dropout scale s and broadcast to both halves.
struct GeluConst2 {
    f32x2 c_p, one, na, nb, nc, k_exp, hs, c2s;
    uint32_t hs_bits;
};
__device__ __forceinline__ GeluConst2 make_gelu_const2(float s) {
    GeluConst2 g;
    const float hs = 0.5f * s;
    g.c_p = pack2(0.47047f * 0.70710678118654752f, 0.47047f * 0.70710678118654752f);
    g.one = pack2(1.0f, 1.0f);
    g.na = pack2(-0.7478556f, -0.7478556f);       // the polynomial is built NEGATED: erf = 1 + (-poly) e
    g.nb = pack2(0.0958798f, 0.0958798f);
    g.nc = pack2(-0.3480242f, -0.3480242f);
    g.k_exp = pack2(-0.5f * 1.4426950408889634f, -0.5f * 1.4426950408889634f);
    g.hs = pack2(hs, hs);
    g.c2s = pack2(0.3989422804014327f * s, 0.3989422804014327f * s);
    g.hs_bits = __float_as_uint(hs);
    return g;
}
// (h, g) = s * (gelu(x), gelu'(x)) for two values at once; same formulas as gelu_value_and_grad.
__device__ __forceinline__ void gelu_value_and_grad2(float x0, float x1, const GeluConst2& k, f32x2& h, f32x2& g) {
    const f32x2 x = pack2(x0, x1);
    const f32x2 ax = pack2(fabsf(x0), fabsf(x1));
    const f32x2 d = fma2(k.c_p, ax, k.one);
    float d0, d1;
    unpack2(d, d0, d1);
    const f32x2 t = pack2(fast_rcp(d0), fast_rcp(d1));
    f32x2 np = fma2(t, k.na, k.nb);
    np = fma2(np, t, k.nc);
    np = mul2(np, t);                                   // -poly(t)
    float q0, q1;
    unpack2(mul2(mul2(x, x), k.k_exp), q0, q1);
    const f32x2 e = pack2(fast_ex2(q0), fast_ex2(q1));  // exp(-x^2 / 2)
    const f32x2 erf_abs = fma2(np, e, k.one);           // erf(|x| / sqrt2)
    h = mul2(fma2(ax, erf_abs, x), k.hs);
    const f32x2 shs = pack2(__uint_as_float((__float_as_uint(x0) & 0x80000000u) | k.hs_bits),
                            __uint_as_float((__float_as_uint(x1) & 0x80000000u) | k.hs_bits));   // copysign(hs, x)
    g = fma2(mul2(x, k.c2s), e, fma2(shs, erf_abs, k.hs));
}
__device__ __forceinline__ uint32_t pack_bf16_2(f32x2 v) {
    float lo, hi;
    unpack2(v, lo, hi);
    return pack_bf16(lo, hi);
}

// Column sums of a 32 x 32 block held one ROW per lane (v[j] = element (row = lane, column = j)): returns, in lane
// j, the sum over the 32 rows of column j. Butterfly transpose-reduce: at every stage a lane keeps one half of its
// columns and sends the other half to its partner (31 shuffles in total); v is destroyed.
__device__ __forceinline__ float warp_colsum32(float (&v)[32], int lane) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
        const bool upper = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < off; ++i) {
            const float send = upper ? v[i] : v[i + off];
            const float keep = upper ? v[i + off] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
    }
    return v[0];
}

// ---------------------------------------------------------------------------------------------
// Programmatic dependent launch. A step is ~340 back-to-back kernels on one stream, and between two of them the SMs
// idle for ~2 us (grid launch latency) plus the next kernel's prologue (barrier init, TMEM allocation, descriptor
// prefetch). Kernels launched through launch_kernel_pdl() may be scheduled while their predecessor is still running:
// every CTA announces "dependents may launch" at its very start (pdl_launch_dependents) and waits for the
// predecessor grid to complete and flush (pdl_wait) after its own prologue, before it touches global memory.
// Without the launch attribute (B200VIT_PDL=0, or a predecessor that is not ours) both instructions are no-ops.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

inline bool pdl_enabled() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("B200VIT_PDL");
        v = e ? atoi(e) : 1;
    }
    return v != 0;
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_kernel_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                     Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ---------------------------------------------------------------------------------------------
// Counter-based dropout mask. Element e of a tensor belongs to block e/8, lane e%8; a block draws 4x32 random
// bits = eight 16-bit lanes from a 32-bit avalanche hash of (counter ^ key), where key is derived on the host
// from (seed, offset). Any kernel (forward epilogue, backward epilogue, standalone) regenerates the same mask
// from (seed, offset) alone, so no mask tensor is ever stored. keep <=> (lane16 & 0x7fff) >= thr15,
// thr15 = round(p*32768): 15-bit lanes leave a guard bit per half-word, so a keep-mask for a packed bf16 pair
// can be produced with five integer instructions (dropout_mask2) instead of per-element compare/select/multiply.
// Semantics mirror nn.Dropout (models/vit.py:43,122,125): kept values are scaled by 1/(1-p).
// (A 10-round Philox here cost more instructions than the GELU it sits next to in the fc1 epilogue.)
// ---------------------------------------------------------------------------------------------
struct DropoutParams {
    float p;            // drop probability; 0 disables
    float scale;        // 1/(1-p)
    uint32_t thr15;     // drop if (lane16 & 0x7fff) < thr15
    uint32_t key_lo;    // hash key derived from (seed, offset)
    uint32_t key_hi;
};

__device__ __forceinline__ uint32_t mix32(uint32_t x) {     // "lowbias32" integer hash
    x ^= x >> 16; x *= 0x7feb352du;
    x ^= x >> 15; x *= 0x846ca68bu;
    x ^= x >> 16;
    return x;
}
// Random words of a block of 8 elements. One multiply-xorshift hash of (block counter, key) gives the block's base
// word (word 0); words 1-3 are cheap bijections of it (xor a constant, multiply, fold the high half down). A kernel
// that needs all four pays 6 + 3 x 4 integer instructions per 8 elements instead of 4 full hashes (the fc1 epilogue is
// bound by instruction issue); one that needs a single word pays 6 + 4.
__device__ __forceinline__ uint32_t dropout_base(const DropoutParams& d, uint64_t blk) {
    const uint32_t hi = (uint32_t)(blk >> 32);              // zero for tensors below 2^35 elements
    uint32_t x = (uint32_t)blk * 0x9E3779B1u + (d.key_lo ^ (hi * 0x85EBCA6Bu + (hi ? d.key_hi : 0u)));
    x ^= x >> 15; x *= 0x846ca68bu;
    x ^= x >> 16;
    return x;
}
__device__ __forceinline__ uint32_t dropout_derive(uint32_t base, uint32_t j) {
    if (j == 0) return base;
    const uint32_t c = j == 1 ? 0x7feb352du : (j == 2 ? 0xC2B2AE35u : 0x27D4EB2Fu);
    uint32_t x = (base ^ c) * 0x9E3779B1u;
    return x ^ (x >> 15);
}
__device__ __forceinline__ uint32_t dropout_word(const DropoutParams& d, uint64_t blk, uint32_t j) {
    return dropout_derive(dropout_base(d, blk), j);
}
__device__ __forceinline__ void dropout_words4(const DropoutParams& d, uint64_t blk, uint32_t (&w)[4]) {
    const uint32_t b = dropout_base(d, blk);
    w[0] = b;
    w[1] = dropout_derive(b, 1);
    w[2] = dropout_derive(b, 2);
    w[3] = dropout_derive(b, 3);
}

// Multipliers (0 or scale) for the 8 consecutive elements [8*blk, 8*blk+8).
__device__ __forceinline__ void dropout_mult8(const DropoutParams& d, uint64_t blk, float (&m)[8]) {
    uint32_t ws[4];
    dropout_words4(d, blk, ws);
#pragma unroll
    for (uint32_t i = 0; i < 4; ++i) {
        const uint32_t w = ws[i];
        m[2 * i]     = ((w & 0x7FFFu) >= d.thr15) ? d.scale : 0.0f;
        m[2 * i + 1] = (((w >> 16) & 0x7FFFu) >= d.thr15) ? d.scale : 0.0f;
    }
}
// Multiplier for a single element (used by small kernels).
__device__ __forceinline__ float dropout_mult1(const DropoutParams& d, uint64_t elem) {
    const uint32_t lane = (uint32_t)(elem & 7);
    const uint32_t w = dropout_word(d, elem >> 3, lane >> 1);
    const uint32_t v = ((lane & 1) ? (w >> 16) : w) & 0x7FFFu;
    return (v >= d.thr15) ? d.scale : 0.0f;
}
// 0xFFFF / 0x0000 half-word masks for the element pair (2i, 2i+1) drawn from random word w: AND them onto a packed
// bf16 pair. Same keep decision as dropout_mult8 / dropout_mult1 (the scale must be applied separately).
__device__ __forceinline__ uint32_t dropout_mask2(const DropoutParams& d, uint32_t w) {
    const uint32_t t = ((w & 0x7FFF7FFFu) | 0x80008000u) - (d.thr15 * 0x00010001u);
    uint32_t m;     // bits 15 / 31 (no borrow = keep) replicated over their half-words: PRMT with sign-replicating selectors
    asm("prmt.b32 %0, %1, %1, 0xBB99;" : "=r"(m) : "r"(t));
    return m;
}

inline DropoutParams make_dropout(float p, uint64_t seed, uint64_t offset) {
    DropoutParams d;
    d.p = p;
    d.scale = (p > 0.f && p < 1.f) ? 1.0f / (1.0f - p) : (p >= 1.f ? 0.f : 1.0f);
    double t = (double)p * 32768.0 + 0.5;
    d.thr15 = p <= 0.f ? 0u : (t >= 32768.0 ? 32768u : (uint32_t)t);
    // splitmix64 of (seed, offset) -> 64-bit key
    uint64_t z = seed * 0x9E3779B97F4A7C15ull + offset * 0xD1B54A32D192ED03ull + 0x2545F4914F6CDD1Dull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    d.key_lo = (uint32_t)z;
    d.key_hi = (uint32_t)(z >> 32);
    return d;
}

}  // namespace b200vit
