// Multi-tensor global gradient norm and fused Adam for the training step.
//   * l2norm_sq: one launch over all gradient tensors -> sum of squares (what
//     torch.nn.utils.clip_grad_norm_(params, 1.0) computes with ~153 kernels, going_modular/engine.py:63).
//   * adam_step: one launch over all parameters; torch.optim.Adam semantics of the reference recipe
//     (coupled L2 weight decay per param group, bias correction, eps outside the corrected sqrt:
//     notebook cell "optimizer = torch.optim.Adam(...)", stepped at engine.py:66). The clip coefficient
//     min(1, max_norm / (norm + 1e-6)) is folded in, so gradients are never rewritten, and the bf16
//     shadow copy the GEMMs consume is refreshed in the same pass.
// HBM-bound: 16-byte vector accesses; a 2-D grid (chunk, tensor) walks a device-resident tensor table.
#include "common.cuh"
#include "launchers.h"

#include "../../include/b200vit.h"

#include <mutex>

namespace b200vit {

constexpr int MT_BLOCK = 256;
constexpr int MT_CHUNK = MT_BLOCK * 4 * 8;  // 8192 elements per block

// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(MT_BLOCK)
l2norm_partial_kernel(const float* const* __restrict__ ptrs, const long long* __restrict__ sizes,
                      float* __restrict__ partials) {
    __shared__ float s_red[MT_BLOCK / 32];
    const int t = blockIdx.y;
    const long long n = sizes[t];
    const long long start = (long long)blockIdx.x * MT_CHUNK;
    float acc = 0.f;
    if (start < n) {
        const float* g = ptrs[t];
        const long long end = min(n, start + MT_CHUNK);
        if ((((uintptr_t)g) & 15) == 0) {
            const long long vend = start + ((end - start) / 4) * 4;
            for (long long i = start + (long long)threadIdx.x * 4; i < vend; i += MT_BLOCK * 4) {
                const float4 v = *reinterpret_cast<const float4*>(g + i);
                acc += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
            }
            for (long long i = vend + threadIdx.x; i < end; i += MT_BLOCK) acc += g[i] * g[i];
        } else {
            for (long long i = start + threadIdx.x; i < end; i += MT_BLOCK) acc += g[i] * g[i];
        }
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        float v = threadIdx.x < MT_BLOCK / 32 ? s_red[threadIdx.x] : 0.f;
        v = warp_sum(v);
        if (threadIdx.x == 0) partials[(long long)blockIdx.y * gridDim.x + blockIdx.x] = v;
    }
}

__global__ void l2norm_final_kernel(const float* __restrict__ partials, int n, float* __restrict__ out) {
    __shared__ double s_red[32];
    double acc = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) acc += (double)partials[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        double v = threadIdx.x < (blockDim.x >> 5) ? s_red[threadIdx.x] : 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (threadIdx.x == 0) *out = (float)v;
    }
}

int l2norm_sq_launch(const float* const* ptrs_dev, const long long* sizes_dev, int num_tensors, long long max_size,
                     float* partials, int num_partials, float* out_sumsq, cudaStream_t stream) {
    B200_CHECK_ARG(ptrs_dev && sizes_dev && partials && out_sumsq, "l2norm_sq: null pointer");
    B200_CHECK_ARG(num_tensors > 0 && num_tensors <= 65535 && max_size > 0, "l2norm_sq: bad sizes");
    const long long chunks = (max_size + MT_CHUNK - 1) / MT_CHUNK;
    B200_CHECK_ARG(chunks * num_tensors <= (long long)num_partials,
                   "l2norm_sq: scratch too small: need %lld partial sums, got %d", chunks * num_tensors, num_partials);
    dim3 grid((unsigned)chunks, (unsigned)num_tensors);
    l2norm_partial_kernel<<<grid, MT_BLOCK, 0, stream>>>(ptrs_dev, sizes_dev, partials);
    B200_LAUNCH_CHECK();
    l2norm_final_kernel<<<1, 1024, 0, stream>>>(partials, (int)(chunks * num_tensors), out_sumsq);
    B200_LAUNCH_CHECK();
    return 0;
}


// ---------------------------------------------------------------------------------------------
// Allocation-free gradient norm.
//
// b200vit_l2norm_sq() above needs a scratch buffer for its per-block partial sums, and somebody has to allocate
// it. Some hosts cannot: a serving process that gave all of HBM to its own arena at start-up (a later cudaMalloc
// simply fails), or a training loop that must not stall -- cudaMalloc / cudaFree synchronise the whole device,
// which would serialise the compute and communication streams of an overlapped data-parallel step.
// For those callers the partial sums live in a fixed array inside the library image instead. The price: every
// GPU that loads libb200vit.so carries L2NORM_STATIC_PARTIALS * 4 bytes (1 MiB) until the process exits, and
// problems that need more partial sums than that are refused.
//
// NB: in host code `g_l2norm_static_scratch` is only a handle for the symbol, not a device address; the address
// a kernel can use comes from cudaGetSymbolAddress(), queried once per device with that device current.
// ---------------------------------------------------------------------------------------------
constexpr int L2NORM_STATIC_PARTIALS = 262144;
static __device__ __align__(256) float g_l2norm_static_scratch[L2NORM_STATIC_PARTIALS];

static int l2norm_static_scratch(float** out) {
    static std::mutex mu;
    static float* per_device[64] = {nullptr};
    int dev = -1;
    B200_CUDA(cudaGetDevice(&dev));
    B200_CHECK_ARG(dev >= 0 && dev < 64, "l2norm_sq_noalloc: unsupported device ordinal %d", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (per_device[dev] == nullptr) {
        void* p = nullptr;
        B200_CUDA(cudaGetSymbolAddress(&p, g_l2norm_static_scratch));
        per_device[dev] = static_cast<float*>(p);
    }
    *out = per_device[dev];
    return 0;
}

}  // namespace b200vit

extern "C" int b200vit_l2norm_sq_noalloc(const void* const* ptrs_dev, const long long* sizes_dev, int num_tensors,
                                         long long max_size, float* out_sumsq, void* stream) {
    using namespace b200vit;
    B200_CHECK_ARG(ptrs_dev && sizes_dev && out_sumsq, "l2norm_sq_noalloc: null pointer");
    B200_CHECK_ARG(num_tensors > 0 && num_tensors <= 65535 && max_size > 0, "l2norm_sq_noalloc: bad sizes");
    const long long chunks = (max_size + MT_CHUNK - 1) / MT_CHUNK;
    const long long need = chunks * num_tensors;
    B200_CHECK_ARG(need <= (long long)L2NORM_STATIC_PARTIALS,
                   "l2norm_sq_noalloc: %d tensors of up to %lld elements need %lld partial sums but the built-in "
                   "scratch holds %d; use b200vit_l2norm_sq with a caller-provided buffer",
                   num_tensors, max_size, need, L2NORM_STATIC_PARTIALS);
    float* scratch = nullptr;
    int rc = l2norm_static_scratch(&scratch);
    if (rc) return rc;
    // Every (chunk, tensor) block overwrites its own slot (blocks past the end of a tensor write 0), so nothing
    // left over from the previous call is ever read: no clearing pass is needed.
    dim3 grid((unsigned)chunks, (unsigned)num_tensors);
    l2norm_partial_kernel<<<grid, MT_BLOCK, 0, (cudaStream_t)stream>>>((const float* const*)ptrs_dev, sizes_dev, scratch);
    B200_LAUNCH_CHECK();
    l2norm_final_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(scratch, (int)need, out_sumsq);
    B200_LAUNCH_CHECK();
    return 0;
}

namespace b200vit {

// ---------------------------------------------------------------------------------------------
// Adam
// ---------------------------------------------------------------------------------------------
struct AdamTensor {        // must match the ctypes Structure in b200vit/optim.py
    float* p;
    const float* g;
    float* m;
    float* v;
    bf16* shadow;          // nullable
    long long n;
    int group;
    int pad;
};

constexpr int MAX_GROUPS = 8;
struct AdamHyper {
    float lr[MAX_GROUPS];
    float wd[MAX_GROUPS];
    float beta1, beta2, eps;
    float bc1, bc2_sqrt;   // 1 - beta1^t, sqrt(1 - beta2^t)
    float max_norm, grad_scale;
    int decoupled;
};

__global__ void __launch_bounds__(MT_BLOCK)
adam_kernel(const AdamTensor* __restrict__ table, AdamHyper h, const float* __restrict__ grad_sumsq) {
    const AdamTensor t = table[blockIdx.y];
    const long long start = (long long)blockIdx.x * MT_CHUNK;
    if (start >= t.n) return;
    const long long end = min(t.n, start + MT_CHUNK);

    float gscale = h.grad_scale;
    if (grad_sumsq != nullptr && h.max_norm > 0.f) {
        const float norm = sqrtf(__ldg(grad_sumsq)) * h.grad_scale;
        const float coef = fminf(1.0f, h.max_norm / (norm + 1e-6f));   // torch clip_grad_norm_ semantics
        gscale *= coef;
    }
    const float lr = h.lr[t.group], wd = h.wd[t.group];
    const float step_size = lr / h.bc1;
    const float decay_mult = h.decoupled ? (1.0f - lr * wd) : 1.0f;

    auto update = [&](float& p, float g, float& m, float& v) {
        g *= gscale;
        if (h.decoupled) p *= decay_mult;
        else g = fmaf(wd, p, g);
        m = m + (g - m) * (1.0f - h.beta1);
        v = h.beta2 * v + (1.0f - h.beta2) * g * g;
        const float denom = sqrtf(v) / h.bc2_sqrt + h.eps;
        p -= step_size * (m / denom);
    };

    const bool vec = ((((uintptr_t)t.p) | ((uintptr_t)t.g) | ((uintptr_t)t.m) | ((uintptr_t)t.v)) & 15) == 0 &&
                     (t.shadow == nullptr || (((uintptr_t)t.shadow) & 7) == 0);
    if (vec) {
        const long long vend = start + ((end - start) / 4) * 4;
        for (long long i = start + (long long)threadIdx.x * 4; i < vend; i += MT_BLOCK * 4) {
            float4 p = *reinterpret_cast<float4*>(t.p + i);
            const float4 g = *reinterpret_cast<const float4*>(t.g + i);
            float4 m = *reinterpret_cast<float4*>(t.m + i);
            float4 v = *reinterpret_cast<float4*>(t.v + i);
            update(p.x, g.x, m.x, v.x);
            update(p.y, g.y, m.y, v.y);
            update(p.z, g.z, m.z, v.z);
            update(p.w, g.w, m.w, v.w);
            *reinterpret_cast<float4*>(t.p + i) = p;
            *reinterpret_cast<float4*>(t.m + i) = m;
            *reinterpret_cast<float4*>(t.v + i) = v;
            if (t.shadow) {
                uint2 s;
                s.x = pack_bf16(p.x, p.y);
                s.y = pack_bf16(p.z, p.w);
                *reinterpret_cast<uint2*>(t.shadow + i) = s;
            }
        }
        for (long long i = vend + threadIdx.x; i < end; i += MT_BLOCK) {
            float p = t.p[i], m = t.m[i], v = t.v[i];
            update(p, t.g[i], m, v);
            t.p[i] = p; t.m[i] = m; t.v[i] = v;
            if (t.shadow) t.shadow[i] = __float2bfloat16(p);
        }
    } else {
        for (long long i = start + threadIdx.x; i < end; i += MT_BLOCK) {
            float p = t.p[i], m = t.m[i], v = t.v[i];
            update(p, t.g[i], m, v);
            t.p[i] = p; t.m[i] = m; t.v[i] = v;
            if (t.shadow) t.shadow[i] = __float2bfloat16(p);
        }
    }
}

int adam_step_launch(const void* table_dev, int num_tensors, long long max_size, const float* lrs, const float* wds,
                     int num_groups, float beta1, float beta2, float eps, int step, int decoupled,
                     const float* grad_sumsq, float max_norm, float grad_scale, cudaStream_t stream) {
    B200_CHECK_ARG(table_dev && lrs && wds, "adam_step: null pointer");
    B200_CHECK_ARG(num_tensors > 0 && num_tensors <= 65535 && max_size > 0, "adam_step: bad sizes");
    B200_CHECK_ARG(num_groups > 0 && num_groups <= MAX_GROUPS, "adam_step: at most %d param groups (got %d)", MAX_GROUPS, num_groups);
    B200_CHECK_ARG(step >= 1, "adam_step: step counts from 1");
    AdamHyper h;
    for (int i = 0; i < MAX_GROUPS; ++i) {
        h.lr[i] = i < num_groups ? lrs[i] : 0.f;
        h.wd[i] = i < num_groups ? wds[i] : 0.f;
    }
    h.beta1 = beta1; h.beta2 = beta2; h.eps = eps;
    h.bc1 = (float)(1.0 - pow((double)beta1, (double)step));
    h.bc2_sqrt = (float)sqrt(1.0 - pow((double)beta2, (double)step));
    h.max_norm = max_norm; h.grad_scale = grad_scale; h.decoupled = decoupled;
    const long long chunks = (max_size + MT_CHUNK - 1) / MT_CHUNK;
    dim3 grid((unsigned)chunks, (unsigned)num_tensors);
    adam_kernel<<<grid, MT_BLOCK, 0, stream>>>(reinterpret_cast<const AdamTensor*>(table_dev), h, grad_sumsq);
    B200_LAUNCH_CHECK();
    return 0;
}

}  // namespace b200vit
