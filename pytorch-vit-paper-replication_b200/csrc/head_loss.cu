// Classifier head (nn.Linear(D, C) on the class-token row, reference models/vit.py:219-222,235) and the
// fused softmax cross-entropy (nn.CrossEntropyLoss(), mean reduction, notebook cell `loss_fn`;
// called at going_modular/engine.py:53) with the argmax accuracy counter of engine.py:73-74.
// Both are tiny (B x C outputs), so they run as fp32 SIMT kernels on the fp32 master weights.
#include "common.cuh"
#include "launchers.h"

namespace b200vit {

// logits[b, c] = xn[b, :] . w[c, :] + bias[c] ; one warp per output.
__global__ void head_fwd_kernel(const float* __restrict__ xn, const float* __restrict__ w, const float* __restrict__ bias,
                                float* __restrict__ logits, int B, int D, int C) {
    const long long o = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (o >= (long long)B * C) return;
    const int b = (int)(o / C), c = (int)(o % C);
    const float* xr = xn + (long long)b * D;
    const float* wr = w + (long long)c * D;
    float s = 0.f;
    for (int d = lane; d < D; d += 32) s += xr[d] * __ldg(wr + d);
    s = warp_sum(s);
    if (lane == 0) logits[o] = s + (bias ? bias[c] : 0.f);
}

int head_fwd_launch(const float* xn, const float* w, const float* b, float* logits, int batch, int dim, int classes,
                    cudaStream_t stream) {
    B200_CHECK_ARG(xn && w && logits, "head_fwd: null pointer");
    B200_CHECK_ARG(batch > 0 && dim > 0 && classes > 0, "head_fwd: bad sizes");
    const long long threads = (long long)batch * classes * 32;
    head_fwd_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, stream>>>(xn, w, b, logits, batch, dim, classes);
    B200_LAUNCH_CHECK();
    return 0;
}

// dxn[b, d] = sum_c dlogits[b, c] * w[c, d]
__global__ void head_dx_kernel(const float* __restrict__ w, const float* __restrict__ dl, float* __restrict__ dxn,
                               int B, int D, int C) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)B * D) return;
    const int b = (int)(i / D), d = (int)(i % D);
    float s = 0.f;
    for (int c = 0; c < C; ++c) s += dl[(long long)b * C + c] * __ldg(w + (long long)c * D + d);
    dxn[i] = s;
}
// dw[c, d] = sum_b dlogits[b, c] * xn[b, d] ; db[c] = sum_b dlogits[b, c]
__global__ void head_dw_kernel(const float* __restrict__ xn, const float* __restrict__ dl, float* __restrict__ dw,
                               float* __restrict__ db, int B, int D, int C) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)C * D) return;
    const int c = (int)(i / D), d = (int)(i % D);
    float s = 0.f, sb = 0.f;
    for (int b = 0; b < B; ++b) {
        const float g = dl[(long long)b * C + c];
        s += g * xn[(long long)b * D + d];
        sb += g;
    }
    dw[i] = s;
    if (d == 0 && db) db[c] = sb;
}

int head_bwd_launch(const float* xn, const float* w, const float* dlogits, float* dxn, float* dw, float* db,
                    int batch, int dim, int classes, cudaStream_t stream) {
    B200_CHECK_ARG(xn && w && dlogits, "head_bwd: null pointer");
    B200_CHECK_ARG(batch > 0 && dim > 0 && classes > 0, "head_bwd: bad sizes");
    if (dxn) {
        const long long n = (long long)batch * dim;
        head_dx_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(w, dlogits, dxn, batch, dim, classes);
        B200_LAUNCH_CHECK();
    }
    if (dw) {
        const long long n = (long long)classes * dim;
        head_dw_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(xn, dlogits, dw, db, batch, dim, classes);
        B200_LAUNCH_CHECK();
    }
    return 0;
}

// One block per sample: loss += (logsumexp(z) - z[y]) / B ; dlogits = (softmax(z) - onehot(y)) / B ;
// correct += (argmax(z) == y)   (first maximal index wins, like torch.argmax).
__global__ void softmax_ce_kernel(const float* __restrict__ logits, const long long* __restrict__ labels,
                                  float* __restrict__ loss, float* __restrict__ dlogits, int* __restrict__ correct,
                                  int B, int C) {
    __shared__ float s_val[32];
    __shared__ int s_idx[32];
    const int b = blockIdx.x;
    const float* z = logits + (long long)b * C;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;

    float mx = -INFINITY;
    int arg = 0x7fffffff;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        const float v = z[c];
        if (v > mx || (v == mx && c < arg)) { mx = v; arg = c; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, mx, o);
        const int oi = __shfl_xor_sync(0xffffffffu, arg, o);
        if (ov > mx || (ov == mx && oi < arg)) { mx = ov; arg = oi; }
    }
    if (lane == 0) { s_val[warp] = mx; s_idx[warp] = arg; }
    __syncthreads();
    if (warp == 0) {
        mx = lane < nwarps ? s_val[lane] : -INFINITY;
        arg = lane < nwarps ? s_idx[lane] : 0x7fffffff;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, mx, o);
            const int oi = __shfl_xor_sync(0xffffffffu, arg, o);
            if (ov > mx || (ov == mx && oi < arg)) { mx = ov; arg = oi; }
        }
        if (lane == 0) { s_val[0] = mx; s_idx[0] = arg; }
    }
    __syncthreads();
    mx = s_val[0];
    arg = s_idx[0];
    __syncthreads();

    float se = 0.f;
    for (int c = threadIdx.x; c < C; c += blockDim.x) se += expf(z[c] - mx);
    se = warp_sum(se);
    if (lane == 0) s_val[warp] = se;
    __syncthreads();
    if (warp == 0) {
        se = lane < nwarps ? s_val[lane] : 0.f;
        se = warp_sum(se);
        if (lane == 0) s_val[0] = se;
    }
    __syncthreads();
    se = s_val[0];

    const long long y = labels[b];
    const bool y_ok = y >= 0 && y < C;
    const float invB = 1.0f / (float)B;
    if (dlogits) {
        for (int c = threadIdx.x; c < C; c += blockDim.x) {
            const float p = expf(z[c] - mx) / se;
            // a label outside [0, C) poisons its row (and the loss, below) with NaN: torch raises for it
            // (device-side assert); a kernel cannot raise without a host sync, but it must not pass silently
            dlogits[(long long)b * C + c] = y_ok ? (p - (c == y ? 1.f : 0.f)) * invB : __int_as_float(0x7fc00000);
        }
    }
    if (threadIdx.x == 0) {
        atomicAdd(loss, y_ok ? (logf(se) + mx - z[y]) * invB : __int_as_float(0x7fc00000));
        if (correct && y_ok && arg == (int)y) atomicAdd(correct, 1);
    }
}

int softmax_ce_launch(const float* logits, const long long* labels, float* loss, float* dlogits, int* correct,
                      int batch, int classes, cudaStream_t stream) {
    B200_CHECK_ARG(logits && labels && loss, "softmax_ce: null pointer");
    B200_CHECK_ARG(batch > 0 && classes > 0, "softmax_ce: bad sizes");
    B200_CUDA(cudaMemsetAsync(loss, 0, sizeof(float), stream));
    if (correct) B200_CUDA(cudaMemsetAsync(correct, 0, sizeof(int), stream));
    const int block = classes >= 512 ? 256 : (classes >= 64 ? 128 : 32);
    softmax_ce_kernel<<<batch, block, 0, stream>>>(logits, labels, loss, dlogits, correct, batch, classes);
    B200_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------------------------
// probs = softmax(logits) per row and labels = argmax (first maximal index): the post-processing of a prediction
// (going_modular/predictions.py:62-66: torch.softmax(dim=1) then torch.argmax(dim=1)). One block per sample.
// ---------------------------------------------------------------------------------------------
__global__ void softmax_top1_kernel(const float* __restrict__ logits, float* __restrict__ probs,
                                    long long* __restrict__ labels, int B, int C) {
    __shared__ float s_val[32];
    __shared__ int s_idx[32];
    const int b = blockIdx.x;
    const float* z = logits + (long long)b * C;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    float mx = -INFINITY;
    int arg = 0x7fffffff;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        const float v = z[c];
        if (v > mx || (v == mx && c < arg)) { mx = v; arg = c; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, mx, o);
        const int oi = __shfl_xor_sync(0xffffffffu, arg, o);
        if (ov > mx || (ov == mx && oi < arg)) { mx = ov; arg = oi; }
    }
    if (lane == 0) { s_val[warp] = mx; s_idx[warp] = arg; }
    __syncthreads();
    if (warp == 0) {
        mx = lane < nwarps ? s_val[lane] : -INFINITY;
        arg = lane < nwarps ? s_idx[lane] : 0x7fffffff;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, mx, o);
            const int oi = __shfl_xor_sync(0xffffffffu, arg, o);
            if (ov > mx || (ov == mx && oi < arg)) { mx = ov; arg = oi; }
        }
        if (lane == 0) { s_val[0] = mx; s_idx[0] = arg; }
    }
    __syncthreads();
    mx = s_val[0];
    arg = s_idx[0];
    __syncthreads();
    float se = 0.f;
    for (int c = threadIdx.x; c < C; c += blockDim.x) se += expf(z[c] - mx);
    se = warp_sum(se);
    if (lane == 0) s_val[warp] = se;
    __syncthreads();
    if (warp == 0) {
        se = lane < nwarps ? s_val[lane] : 0.f;
        se = warp_sum(se);
        if (lane == 0) s_val[0] = se;
    }
    __syncthreads();
    const float inv = 1.0f / s_val[0];
    if (probs)
        for (int c = threadIdx.x; c < C; c += blockDim.x) probs[(long long)b * C + c] = expf(z[c] - mx) * inv;
    if (labels && threadIdx.x == 0) labels[b] = arg;
}

int softmax_top1_launch(const float* logits, float* probs, long long* labels, int batch, int classes, cudaStream_t stream) {
    B200_CHECK_ARG(logits && (probs || labels), "softmax_top1: null pointer");
    B200_CHECK_ARG(batch > 0 && classes > 0, "softmax_top1: bad sizes");
    const int block = classes >= 512 ? 256 : (classes >= 64 ? 128 : 32);
    softmax_top1_kernel<<<batch, block, 0, stream>>>(logits, probs, labels, batch, classes);
    B200_LAUNCH_CHECK();
    return 0;
}

}  // namespace b200vit
