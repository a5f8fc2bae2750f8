// Gradient all-reduce of the data-parallel step over peer memory (NVLink 5 / NVSwitch), as ONE small-footprint kernel
// per bucket that can sit on the same SMs as the backward kernels it overlaps.
//
// The reference is single-device (going_modular/engine.py:180 `model.to(device)`); this is the one exchange step
// that data-parallel training of its hot path needs (b200vit/dp.py). NCCL's all-reduce kernels take 32 CTAs of 544
// threads, 96 registers and 101 KB of shared memory each: they cannot share an SM with the persistent tcgen05 GEMM /
// attention CTAs (225 KB of shared memory), so every bucket held those SMs -- and, through the cluster pairs, up to 64
// of them -- away from the weight-gradient GEMM that follows it for ~100 us (measured: 137 -> 237 us per bucket,
// 1.4 ms of a 34.4 ms step on 2 GPUs, 3.8 ms on 8). This kernel needs no shared memory, 64 threads and <= 64 registers
// per CTA, which fits beside one GEMM or attention CTA on an SM.
//
// Every rank's bucket lives at the same offset of a symmetric allocation (torch.distributed._symmetric_memory: each
// rank maps its peers' buffers; plumbing, not the product). Rank r owns slice r of the bucket:
//   A  tells every peer "my gradients of call `seq` are complete" (one flag per source rank in the peer's flag words)
//      and waits for the same from all of them;
//   B  reads its slice from all `world` buffers (plain 16-byte loads over NVLink, no multicast), sums them in rank
//      order -- so every rank ends with bit-identical sums -- and writes the result into all `world` buffers;
//   C  the last CTA to finish tells every peer "slice r of call `seq` is final" and waits for the same from all of
//      them; when the kernel ends, the whole local bucket holds the sum.
// Flags only ever grow (seq increases by one per call, compared as a signed difference), spins are bounded.
#include "common.cuh"
#include "launchers.h"

#include "../../include/b200vit.h"

#include <mutex>

namespace b200vit {

constexpr int PEER_MAX_WORLD = 8;
constexpr int PEER_THREADS = 64;
constexpr int PEER_FLAG_READY = 0;      // [PEER_MAX_WORLD]  written by the peers
constexpr int PEER_FLAG_DONE = 16;      // [PEER_MAX_WORLD]  written by the peers
constexpr int PEER_FLAG_COUNT = 32;     // CTAs of the local kernel that are through (wraps to 0 with the last one)
constexpr int PEER_FLAG_ENTER = 33;     // CTAs that have started (the first one signals the peers)
constexpr int PEER_FLAG_CHUNK = 34;     // next unclaimed chunk of the slice
constexpr int PEER_FLAG_SM = 64;        // [PEER_SM_SLOTS] CTAs of this launch that landed on each SM
constexpr int PEER_SM_SLOTS = 256;
constexpr int PEER_FLAG_WORDS = PEER_FLAG_SM + PEER_SM_SLOTS;      // = B200VIT_PEER_FLAG_WORDS (include/b200vit.h)
constexpr int PEER_ROUNDS = 8;          // rounds per claimed chunk

struct PeerArgs {
    float* buf[PEER_MAX_WORLD];
    unsigned* flags[PEER_MAX_WORLD];
    long long count;                    // fp32 elements, multiple of 4
    unsigned seq;
    int rank;
};

__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 ld_stream_f4(const float* p) {
    float4 v;
    asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_stream_f4(float* p, const float4& v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};"
                 ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ unsigned sm_id() {
    unsigned v;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(v));
    return v;
}
// Bounded: a peer that never arrives (a crashed rank) becomes a launch error here instead of a GPU that spins forever.
__device__ __forceinline__ void wait_flag(const unsigned* p, unsigned seq) {
    const long long t0 = clock64();
    while ((int)(ld_acquire_sys(p) - seq) < 0) {
        __nanosleep(200);
        if (clock64() - t0 > 40000000000LL) __trap();      // ~20 s
    }
}

// One CTA of this kernel fits beside a 225 KB GEMM / attention CTA on an SM, two do not (each resident CTA costs 1 KB
// of the SM's shared memory), and where CTAs land is the hardware's choice. So the grid is larger than the SM count,
// every CTA but the first to land on its SM leaves at once, and the ones that stay claim the work in chunks from a
// counter instead of owning a fixed share.
template <int W>
__global__ void __launch_bounds__(PEER_THREADS, 16)
allreduce_peer_kernel(const PeerArgs a) {
    constexpr int U = (8 / W) > 0 ? (8 / W) : 1;       // 16-byte vectors per lane per round: U * W = 8 loads in flight
    const int tid = threadIdx.x, lane = tid & 31;
    unsigned* my_flags = a.flags[a.rank];

    int stay = 0, first = 0;
    if (tid == 0) {
        first = atomicAdd(my_flags + PEER_FLAG_ENTER, 1u) == 0u;
        stay = atomicAdd(my_flags + PEER_FLAG_SM + (sm_id() & (PEER_SM_SLOTS - 1)), 1u) == 0u;
    }
    first = __syncthreads_or(first);
    stay = __syncthreads_or(stay);

    // ---- A: my gradients are complete (stream order put every producer kernel before this one); are everyone's? ----
    if (first && tid < W && tid != a.rank) {
        __threadfence_system();
        st_release_sys(a.flags[tid] + PEER_FLAG_READY + a.rank, a.seq);
    }
    if (stay) {
        if (tid < W && tid != a.rank) wait_flag(my_flags + PEER_FLAG_READY + tid, a.seq);
        __syncthreads();

        // ---- B: slice `rank` of the bucket, one warp-sized chunk (PEER_ROUNDS rounds of 32 * U vectors) at a time ----
        const long long nvec = a.count >> 2;
        const long long per = (nvec + W - 1) / W;
        const long long lo = per * a.rank;
        const long long hi = lo + per < nvec ? lo + per : nvec;
        constexpr long long CHUNK = 32LL * U * PEER_ROUNDS;
        for (;;) {
            unsigned c = 0;
            if (lane == 0) c = atomicAdd(my_flags + PEER_FLAG_CHUNK, 1u);
            c = __shfl_sync(0xffffffffu, c, 0);
            const long long c0 = lo + (long long)c * CHUNK;
            if (c0 >= hi) break;
#pragma unroll 1
            for (int round = 0; round < PEER_ROUNDS; ++round) {
                const long long i0 = c0 + (long long)round * (32 * U) + lane;
                float4 v[U][W];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const long long i = i0 + u * 32;
                    if (i < hi) {
#pragma unroll
                        for (int r = 0; r < W; ++r) v[u][r] = ld_stream_f4(a.buf[r] + 4 * i);
                    }
                }
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const long long i = i0 + u * 32;
                    if (i < hi) {
                        float4 s = v[u][0];
#pragma unroll
                        for (int r = 1; r < W; ++r) { s.x += v[u][r].x; s.y += v[u][r].y; s.z += v[u][r].z; s.w += v[u][r].w; }
#pragma unroll
                        for (int r = 0; r < W; ++r) st_stream_f4(a.buf[r] + 4 * i, s);
                    }
                }
            }
        }
    }

    // ---- C: slice final everywhere -> tell the peers; the kernel ends when all slices of the local bucket are final ----
    __syncthreads();
    int last = 0;
    if (tid == 0) {
        __threadfence_system();
        const unsigned prev = atomicInc(my_flags + PEER_FLAG_COUNT, gridDim.x - 1);      // back to 0 after the last CTA
        last = (prev == gridDim.x - 1);
    }
    if (!__syncthreads_or(last)) return;
    // every CTA of this launch has passed its counters: clear them for the next launch
    for (int i = tid; i < PEER_SM_SLOTS; i += PEER_THREADS) my_flags[PEER_FLAG_SM + i] = 0u;
    if (tid == 0) { my_flags[PEER_FLAG_ENTER] = 0u; my_flags[PEER_FLAG_CHUNK] = 0u; }
    if (tid < W && tid != a.rank) {
        __threadfence_system();
        st_release_sys(a.flags[tid] + PEER_FLAG_DONE + a.rank, a.seq);
        wait_flag(my_flags + PEER_FLAG_DONE + tid, a.seq);
    }
}

int allreduce_peer_launch(void* const* bufs, void* const* flags, int world, int rank, long long count, unsigned seq,
                          int ctas, cudaStream_t stream) {
    B200_CHECK_ARG(bufs && flags, "allreduce_peer: null pointer table");
    B200_CHECK_ARG(world >= 2 && world <= PEER_MAX_WORLD, "allreduce_peer: world %d outside [2, %d]", world, PEER_MAX_WORLD);
    B200_CHECK_ARG(rank >= 0 && rank < world, "allreduce_peer: rank %d outside the world of %d", rank, world);
    B200_CHECK_ARG(count > 0 && (count & 3) == 0, "allreduce_peer: count %lld must be a positive multiple of 4", count);
    B200_CHECK_ARG(ctas >= 1 && ctas <= 4096, "allreduce_peer: %d CTAs outside [1, 4096]", ctas);
    PeerArgs a;
    for (int r = 0; r < PEER_MAX_WORLD; ++r) { a.buf[r] = nullptr; a.flags[r] = nullptr; }
    for (int r = 0; r < world; ++r) {
        B200_CHECK_ARG(bufs[r] && flags[r], "allreduce_peer: null buffer / flag pointer of rank %d", r);
        B200_CHECK_ARG(((uintptr_t)bufs[r] & 15) == 0 && ((uintptr_t)flags[r] & 3) == 0, "allreduce_peer: rank %d's buffer must be 16-byte aligned", r);
        a.buf[r] = (float*)bufs[r];
        a.flags[r] = (unsigned*)flags[r];
    }
    a.count = count; a.seq = seq; a.rank = rank;
    const dim3 grid((unsigned)ctas), block(PEER_THREADS);
    void (*kern)(const PeerArgs) = nullptr;
    switch (world) {
        case 2: kern = allreduce_peer_kernel<2>; break;
        case 3: kern = allreduce_peer_kernel<3>; break;
        case 4: kern = allreduce_peer_kernel<4>; break;
        case 5: kern = allreduce_peer_kernel<5>; break;
        case 6: kern = allreduce_peer_kernel<6>; break;
        case 7: kern = allreduce_peer_kernel<7>; break;
        default: kern = allreduce_peer_kernel<8>; break;
    }
    // The kernel uses no shared memory, but it must run on SMs configured for the LARGEST shared-memory carveout:
    // an SM's L1 / shared split only changes while the SM is idle, so a CTA of this kernel sitting on an SM in the
    // default (L1-heavy) split would keep the 225 KB GEMM / attention CTAs off that SM until it leaves.
    {
        static std::mutex mu;
        static bool done[PEER_MAX_WORLD + 1][64] = {};
        int dev = 0;
        B200_CUDA(cudaGetDevice(&dev));
        std::lock_guard<std::mutex> lock(mu);
        if (dev >= 0 && dev < 64 && !done[world][dev]) {
            B200_CUDA(cudaFuncSetAttribute((const void*)kern, cudaFuncAttributePreferredSharedMemoryCarveout,
                                           (int)cudaSharedmemCarveoutMaxShared));
            done[world][dev] = true;
        }
    }
    kern<<<grid, block, 0, stream>>>(a);
    B200_LAUNCH_CHECK();
    return 0;
}

}  // namespace b200vit

extern "C" int b200vit_allreduce_peer_f32(void* const* bufs, void* const* flags, int world, int rank, long long count,
                                          unsigned seq, int ctas, void* stream) {
    return b200vit::allreduce_peer_launch(bufs, flags, world, rank, count, seq, ctas, (cudaStream_t)stream);
}
