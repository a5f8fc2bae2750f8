// tcgen05 / TMEM fused attention forward for short sequences (T <= 256, head_dim 64): the 197-token ViT case.
// One CTA per (batch, head, 128-query tile):
//   warp 0      issues the TMA loads (Q tile, all K, all V of the head) and both UMMAs
//   warp 1      TMEM allocator (256 columns -> two CTAs co-reside per SM and overlap each other's phases)
//   warps 2-5   softmax + epilogue: thread <-> query row (TMEM lane)
// Dataflow:  S = Q K^T (UMMA 128 x Nk x 64, smem x smem -> TMEM fp32)
//            P = exp2((S - rowmax) * scale*log2e) as bf16 pairs written IN PLACE over S with tcgen05.st
//            O = P V   (UMMA 128 x 64 x Nk, A from TMEM, V as MN-major smem operand) ; out = O / rowsum
// The whole key range of a head (Nk = T rounded to 16 <= 256) fits one accumulator, so the softmax is a plain
// two-pass row softmax: no online rescaling. Scores never leave the SM; only O and the log-sum-exp do.
// (Reference semantics: scaled_dot_product_attention inside nn.MultiheadAttention, models/vit.py:86-97.)
#include "common.cuh"
#include "launchers.h"
#include "ptx_sm100.cuh"

#include <cudaTypedefs.h>
#include <mutex>
#include <stdlib.h>

namespace b200vit {

using namespace ptx;

constexpr int ATC_BM = 128;
constexpr int ATC_SM_WARPS = 16;    // softmax warps: 2 query tiles x 4 TMEM lane quarters x 2 column halves
constexpr int ATC_THREADS = 64 + ATC_SM_WARPS * 32;   // warp 0: TMA producer, warp 1: TMEM alloc + MMA issue
constexpr int ATC_TMEM_COLS = 512;  // two 256-column slots, one per 128-query tile of the head in flight
constexpr int ATC_O_COL = 192;      // O accumulator columns [192, 256) of a slot: clear of both packed-P ranges
constexpr int ATC_Q_ROWS = 256;     // Q stage holds up to two 128-row tiles (rows >= T are zero-filled by TMA)

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// Persistent: each CTA (one per SM) walks over (batch, head) work items. Per head the producer stages Q (all
// rows), K and V once (2-deep smem ring); the MMA thread issues S = Q_t K^T for both query tiles into the two TMEM
// slots; softmax group g (8 warps: thread <-> query row, two warps per row splitting the key columns) owns tile g:
// P = exp2(...) written in place as bf16 (each column half packs into the front of its own range), then
// O = P V (A operand from TMEM), then the epilogue. The two groups, the MMA thread and the producer run decoupled
// through mbarriers, so loads, tensor work and softmax of neighbouring tiles / heads overlap.
__global__ void __launch_bounds__(ATC_THREADS, 1)
attention_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                        bf16* __restrict__ out, float* __restrict__ lse, int num_heads_total, int T, int Nk, int H,
                        float scale_log2, float scale) {
    extern __shared__ uint8_t smem_raw[];
    pdl_launch_dependents();
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t kv_bytes = (uint32_t)Nk * 128u;
    const uint32_t kv_stride = (kv_bytes + 1023u) & ~1023u;
    const uint32_t stage_bytes = ATC_Q_ROWS * 128 + 2 * kv_stride;
    const uint32_t xchg_base = smem_base + 2 * stage_bytes;         // float [2 groups][2 halves][128]: max, then sum
    const uint32_t bar_base = xchg_base + 2u * 2u * 2u * 128u * 4u;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 16u + 8u * s; };
    auto s_ready = [&](int g) { return bar_base + 32u + 8u * g; };
    auto p_ready = [&](int g) { return bar_base + 48u + 8u * g; };
    auto o_ready = [&](int g) { return bar_base + 64u + 8u * g; };
    auto slot_free = [&](int g) { return bar_base + 80u + 8u * g; };
    const uint32_t tmem_slot = bar_base + 96u;
    uint8_t* gen = smem_raw - smem_u32(smem_raw);
    uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(gen + tmem_slot);
    float* sMax = reinterpret_cast<float*>(gen + xchg_base);        // [g][half][128]
    float* sSum = sMax + 2 * 2 * 128;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nt = (T + ATC_BM - 1) / ATC_BM;          // query tiles per head: 1 or 2
    const int split = ((Nk / 2) + 15) & ~15;           // key columns [0, split) -> half 0, [split, Nk) -> half 1
    auto packed_col = [&](int src_col) { return src_col < split ? (src_col >> 1) : split + ((src_col - split) >> 1); };

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmQ);
        prefetch_tmap(&tmKV);
        for (int i = 0; i < 2; ++i) {
            mbar_init(full_bar(i), 1);
            mbar_init(empty_bar(i), 1);
            mbar_init(s_ready(i), 1);
            mbar_init(p_ready(i), 8);
            mbar_init(o_ready(i), 1);
            mbar_init(slot_free(i), 8);
        }
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc<1>(tmem_slot, ATC_TMEM_COLS);
        tmem_relinquish<1>();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot_ptr;
    pdl_wait();       // the prologue above overlapped the previous kernel; global memory is touched from here on

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int it = 0;
            for (int w = blockIdx.x; w < num_heads_total; w += gridDim.x, ++it) {
                const int s = it & 1;
                const uint32_t ph = (it >> 1) & 1;
                const int b = w / H, h = w % H;
                mbar_wait(empty_bar(s), ph ^ 1u);
                const uint32_t sQ = smem_base + s * stage_bytes, sK = sQ + ATC_Q_ROWS * 128, sV = sK + kv_stride;
                mbar_arrive_expect_tx(full_bar(s), (uint32_t)nt * ATC_BM * 128 + 2 * kv_bytes);
                tma_load_4d(sQ, &tmQ, full_bar(s), 0, 0, h, b);
                tma_load_4d(sK, &tmKV, full_bar(s), 0, 0, H + h, b);
                tma_load_4d(sV, &tmKV, full_bar(s), 0, 0, 2 * H + h, b);
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        {       // all 32 lanes run the loop (uniform operands); one elected lane issues each tcgen05 instruction
            const uint32_t idesc_s = make_idesc_bf16(ATC_BM, Nk, 0, 0);
            const uint32_t idesc_o = make_idesc_bf16(ATC_BM, 64, 0, 1);
            int it = 0;
            for (int w = blockIdx.x; w < num_heads_total; w += gridDim.x, ++it) {
                const int s = it & 1;
                const uint32_t ph = (it >> 1) & 1, hp = it & 1;
                const uint32_t sQ = smem_base + s * stage_bytes, sK = sQ + ATC_Q_ROWS * 128, sV = sK + kv_stride;
                mbar_wait(full_bar(s), ph);
                tc_fence_after();
                for (int g = 0; g < nt; ++g) {
                    mbar_wait(slot_free(g), hp ^ 1u);
                    tc_fence_after();
                    // S = Q_g K^T : A = Q rows [128 g, 128 g + 128) K-major, B = K [Nk x 64] K-major
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_bf16_ss_elect(tmem + g * 256, make_smem_desc_sw128(sQ + g * (ATC_BM * 128) + k * 32, 1024, 0),
                                     make_smem_desc_sw128(sK + k * 32, 1024, 0), idesc_s, k > 0 ? 1u : 0u);
                    umma_commit_elect(s_ready(g));
                }
                for (int g = 0; g < nt; ++g) {
                    mbar_wait(p_ready(g), hp);
                    tc_fence_after();
                    // O = P V : A = P in TMEM (bf16 pairs: 8 columns per 16 keys), B = V [Nk keys x 64] MN-major
                    for (int k = 0; k < Nk / 16; ++k)
                        umma_bf16_ts_elect(tmem + g * 256 + ATC_O_COL, tmem + g * 256 + packed_col(16 * k),
                                     make_smem_desc_sw128(sV + k * 2048, 1024, 0), idesc_o, k > 0 ? 1u : 0u);
                    umma_commit_elect(o_ready(g));
                }
                umma_commit_elect(empty_bar(s));      // every MMA that reads this smem stage has retired
            }
        }
        __syncwarp();
    } else {
        // ===================== softmax + epilogue groups =====================
        const int sw = warp - 2;
        const int g = sw >> 3;                  // query tile / TMEM slot
        const int half = (sw >> 2) & 1;         // key-column half
        const int quarter = warp & 3;           // TMEM lane quarter this warp may access
        if (g < nt) {
            const uint32_t tbase = tmem + g * 256 + ((uint32_t)(quarter * 32) << 16);
            const int r_in_tile = quarter * 32 + lane;
            const int r_in_head = g * ATC_BM + r_in_tile;       // query index inside the head
            const int c_begin = half == 0 ? 0 : split, c_end = half == 0 ? split : Nk;
            float* myMax = sMax + (g * 2 + half) * 128, *otherMax = sMax + (g * 2 + (half ^ 1)) * 128;
            float* mySum = sSum + (g * 2 + half) * 128, *otherSum = sSum + (g * 2 + (half ^ 1)) * 128;
            int it = 0;
            for (int w = blockIdx.x; w < num_heads_total; w += gridDim.x, ++it) {
                const uint32_t hp = it & 1;
                const int b = w / H, h = w % H;
                mbar_wait(s_ready(g), hp);
                tc_fence_after();
                // pass 1: max over this half's valid keys
                float mx = -INFINITY;
                for (int c = c_begin; c < c_end; c += 16) {
                    uint32_t r[16];
                    tmem_ld_32x16(tbase + c, r);
                    tmem_ld_wait();
                    if (c + 16 <= T) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) mx = fmaxf(mx, __uint_as_float(r[i]));
                    } else {
#pragma unroll
                        for (int i = 0; i < 16; ++i) if (c + i < T) mx = fmaxf(mx, __uint_as_float(r[i]));
                    }
                }
                myMax[r_in_tile] = mx;
                named_bar_sync(1 + g, 256);
                mx = fmaxf(mx, otherMax[r_in_tile]);
                // pass 2: P = exp2((S - max) * scale*log2e) as bf16 pairs, packed into the front of this half's range
                const float mneg = -mx * scale_log2;
                float sum = 0.f;
                for (int c = c_begin; c < c_end; c += 16) {
                    uint32_t r[16];
                    tmem_ld_32x16(tbase + c, r);
                    tmem_ld_wait();
                    float p[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) p[i] = fast_ex2(fmaf(__uint_as_float(r[i]), scale_log2, mneg));
                    if (c + 16 > T) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) if (c + i >= T) p[i] = 0.f;
                    }
                    uint32_t pk[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        pk[i] = pack_bf16(p[2 * i], p[2 * i + 1]);
                        sum += p[2 * i] + p[2 * i + 1];
                    }
                    tmem_st_32x8(tbase + c_begin + ((c - c_begin) >> 1), pk);
                }
                mySum[r_in_tile] = sum;
                tmem_st_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(p_ready(g));
                named_bar_sync(1 + g, 256);           // both halves' partial sums are visible (and max slots reusable)
                sum += otherSum[r_in_tile];
                const float inv = 1.0f / sum;
                const bool row_ok = r_in_head < T;
                if (row_ok && half == 0) lse[(long long)w * T + r_in_head] = mx * scale + logf(sum);
                // epilogue: this half converts 32 of the 64 output columns: O / sum -> bf16 -> out[b, row, h, :]
                mbar_wait(o_ready(g), hp);
                tc_fence_after();
                uint32_t r[32];
                tmem_ld_32x32(tbase + ATC_O_COL + half * 32, r);
                tmem_ld_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(slot_free(g));     // all TMEM reads of this warp for this tile are done
                if (row_ok) {
                    bf16* dst = out + (((long long)b * T + r_in_head) * H + h) * 64 + half * 32;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        uint4 o;
                        o.x = pack_bf16(__uint_as_float(r[8 * i]) * inv, __uint_as_float(r[8 * i + 1]) * inv);
                        o.y = pack_bf16(__uint_as_float(r[8 * i + 2]) * inv, __uint_as_float(r[8 * i + 3]) * inv);
                        o.z = pack_bf16(__uint_as_float(r[8 * i + 4]) * inv, __uint_as_float(r[8 * i + 5]) * inv);
                        o.w = pack_bf16(__uint_as_float(r[8 * i + 6]) * inv, __uint_as_float(r[8 * i + 7]) * inv);
                        reinterpret_cast<uint4*>(dst)[i] = o;
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<1>(tmem, ATC_TMEM_COLS);
    }
}

constexpr int ATB_TAIL_PAD = 8192;    // a 128-row X tile starting at row 128 over-reads a 208-row stage by 6 KB

// =============================================================================================
// Backward (T <= 256), single pass. Earlier versions built the probabilities twice per head (once with keys on
// the TMEM lanes for dK/dV, once with queries on the lanes for dQ). Here they are built ONCE, keys on the
// lanes: per 128-key tile t and 64-query chunk c
//     S^T = K_t Q_c^T, dP^T = V_t dO_c^T  ->  P^T, dS^T (bf16, packed in place)  ->  dV_t += P^T dO_c, dK_t += dS^T Q_c
// and dS^T is ALSO written to shared memory as bf16, one 128-byte row of 64 queries per key: that is exactly the
// 128B-swizzled MN-major layout of an A operand with M = queries, so after two chunks (128 queries)
//     dQ_m += dS[128 queries x keys of tile t] K_t
// is one more UMMA chain, A from shared memory, accumulating over the key tiles in TMEM. All exponentials, the
// dS math and the S / dP MMAs are done once instead of twice.
// TMEM: [0,128) chunk accumulators of group 0 (S^T | dP^T), [128,256) of group 1, [256,320) dV_t, [320,384) dK_t,
//       [384,448) dQ rows 0-127, [448,512) dQ rows 128-255.
// Warps: 0 = TMA + statistics, 1 = MMA issue, 2-17 = two math groups (8 warps: thread <-> key row, two warps per
// row splitting a chunk's 64 query columns); the groups alternate chunks. The Q/K/V/dO tiles of a head are staged once
// (single stage: 4 tiles + 2 x 32 KB of dS leave no room for a second), statistics are double-buffered.
// =============================================================================================
constexpr int AT3_W = 64;
constexpr int AT3_THREADS = ATC_THREADS + 64;     // + two statistics warps (registers are allotted per 4 warps anyway)

__global__ void __launch_bounds__(AT3_THREADS, 1)
attention_bwd_tc3_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO,
                         const bf16* __restrict__ out, const bf16* __restrict__ dout, const float* __restrict__ lse,
                         bf16* __restrict__ dqkv, float* __restrict__ dbias, int num_heads_total, int T, int Nk, int H,
                         float scale_log2, float scale) {
    extern __shared__ uint8_t smem_raw[];
    pdl_launch_dependents();
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t tile_bytes = (uint32_t)Nk * 128u;
    const uint32_t tile_stride = (tile_bytes + 1023u) & ~1023u;
    const uint32_t sQ = smem_base, sK = sQ + tile_stride, sV = sK + tile_stride, sdO = sV + tile_stride;
    const uint32_t ds_base = smem_base + 4 * tile_stride + ATB_TAIL_PAD;     // 2 buffers x [2 query blocks][128 keys][128 B]
    const uint32_t stat_base = ds_base + 2u * 32768u;                        // 2 stages x (lse2 [Nk], delta*scale [Nk])
    const uint32_t bar_base = stat_base + 2u * 2u * Nk * 4u;
    const uint32_t full_bar = bar_base, head_empty = bar_base + 8u;
    auto stat_bar = [&](int s) { return bar_base + 16u + 8u * s; };
    auto s_ready = [&](int g) { return bar_base + 32u + 8u * g; };
    auto p_ready = [&](int g) { return bar_base + 48u + 8u * g; };
    auto ds_free = [&](int m) { return bar_base + 64u + 8u * m; };
    const uint32_t kv_ready = bar_base + 80u, kv_free = bar_base + 88u, dq_ready = bar_base + 96u, dq_free = bar_base + 104u;
    auto stat_free = [&](int s) { return bar_base + 112u + 8u * s; };
    const uint32_t tmem_slot = bar_base + 128u;
    const uint32_t kv_started = bar_base + 144u;
    auto ds_half = [&](int m) { return bar_base + 152u + 8u * m; };
    // optional fused in-projection bias gradient: per-CTA column sums of dQ | dK | dV, [3][H][64] fp32, flushed at the end
    float* sBias = reinterpret_cast<float*>(smem_raw - smem_u32(smem_raw) + bar_base + 256u);
    if (dbias != nullptr)
        for (int k = threadIdx.x; k < 3 * H * 64; k += blockDim.x) sBias[k] = 0.f;
    uint8_t* gen = smem_raw - smem_u32(smem_raw);
    uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(gen + tmem_slot);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nt = (T + ATC_BM - 1) / ATC_BM;           // key tiles
    const int nc = (Nk + AT3_W - 1) / AT3_W;            // query chunks
    const int nsteps = nt * nc;
    const int nm = (nc + 1) / 2;                        // 128-query dQ tiles

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmQKV);
        prefetch_tmap(&tmDO);
        mbar_init(full_bar, 1);
        mbar_init(head_empty, 2);       // "2": one commit per MMA-issuing warp
        for (int i = 0; i < 2; ++i) {
            mbar_init(stat_bar(i), 2);
            mbar_init(stat_free(i), 2);
            mbar_init(s_ready(i), 1);
            mbar_init(p_ready(i), 8);
            mbar_init(ds_free(i), 1);
            mbar_init(ds_half(i), 8);
        }
        mbar_init(kv_ready, 2);
        mbar_init(kv_free, 16);
        mbar_init(kv_started, 1);
        mbar_init(dq_ready, 2);
        mbar_init(dq_free, 16);
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc<1>(tmem_slot, 512);
        tmem_relinquish<1>();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot_ptr;
    pdl_wait();       // the prologue above overlapped the previous kernel; global memory is touched from here on

    if (warp < 2) {
        // ===================== two MMA issuers: warp 1 serves math group 0, warp 0 serves group 1 and issues the TMA loads ======
        // One thread can only issue ~1 small UMMA per 100 cycles, and a head is ~140 of them: with a single issuer that is
        // half of the head's time. All 32 lanes run the loop (uniform operands); one elected lane issues each instruction.
        // Cross-issuer ordering: tcgen05.commit only tracks the MMAs of its own thread, so barriers that mean "all MMAs of
        // the tile / head are done" (kv_ready, dq_ready, head_empty, stat_free) take one commit from EACH issuer; the
        // issuer of a tile's first chunk (accumulate = 0 into dV/dK) signals kv_started, the other one waits for it;
        // the dQ MMA waits until the other group's half of the dS tile is in shared memory (ds_half).
        const int my_g = warp == 1 ? 0 : 1;
        const uint32_t idesc2 = make_idesc_bf16(ATC_BM, 64, 0, 1);      // A from TMEM, B MN-major
        const uint32_t idesc3 = make_idesc_bf16(ATC_BM, 64, 1, 1);      // A (dS) and B (K) both MN-major
        const uint32_t acc = tmem + my_g * 128;
        uint32_t n_p = 0;
        uint32_t S0 = 0;                // global step counter at the start of the head
        int it = 0;
        for (int w = blockIdx.x; w < num_heads_total; w += gridDim.x, ++it, S0 += nsteps) {
            if (warp == 0) {
                const int b = w / H, h = w % H;
                mbar_wait(head_empty, (it & 1) ^ 1u);
                if (lane == 0) {
                    mbar_arrive_expect_tx(full_bar, 4 * tile_bytes);
                    tma_load_4d(sQ, &tmQKV, full_bar, 0, 0, h, b);
                    tma_load_4d(sK, &tmQKV, full_bar, 0, 0, H + h, b);
                    tma_load_4d(sV, &tmQKV, full_bar, 0, 0, 2 * H + h, b);
                    tma_load_4d(sdO, &tmDO, full_bar, 0, 0, h, b);
                }
                __syncwarp();
            }
            mbar_wait(full_bar, it & 1);
            tc_fence_after();
            auto width = [&](int c) { return min(AT3_W, Nk - c * AT3_W); };
            // (t, c) of step s is carried along instead of divided out: a runtime integer division is ~25 instructions with
            // two dependent MUFU/I2F round trips, on the critical issue path of every step
            auto issue_1 = [&](int t, int c) {  // S^T and dP^T of (my) step (t, c) into my group's chunk accumulators
                const uint32_t xo = (uint32_t)t * (ATC_BM * 128), yo = (uint32_t)c * (AT3_W * 128);
                const uint32_t idesc1 = make_idesc_bf16(ATC_BM, width(c), 0, 0);
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    umma_bf16_ss_elect(acc, make_smem_desc_sw128(sK + xo + k * 32, 1024, 0),
                                       make_smem_desc_sw128(sQ + yo + k * 32, 1024, 0), idesc1, k > 0 ? 1u : 0u);
                    umma_bf16_ss_elect(acc + 64, make_smem_desc_sw128(sV + xo + k * 32, 1024, 0),
                                       make_smem_desc_sw128(sdO + yo + k * 32, 1024, 0), idesc1, k > 0 ? 1u : 0u);
                }
                umma_commit_elect(s_ready(my_g));
            };
            const int s_first = ((S0 & 1u) == (uint32_t)my_g) ? 0 : 1;
            if (s_first < nsteps) issue_1(nc == 1 ? s_first : 0, nc == 1 ? 0 : s_first);
            bool dq_waited = false;
            for (int t = 0; t < nt; ++t) {
                const uint32_t tile_idx = (uint32_t)it * nt + t;        // key tiles so far, all heads
                bool first_mine = true;
                for (int c = 0; c < nc; ++c) {
                    const int s = t * nc + c;
                    if (((S0 + s) & 1u) != (uint32_t)my_g) continue;
                    mbar_wait(p_ready(my_g), n_p & 1u);
                    ++n_p;
                    if (c == 0) mbar_wait(kv_free, (tile_idx & 1u) ^ 1u);       // the previous tile's dV / dK have been read out
                    else if (first_mine) mbar_wait(kv_started, tile_idx & 1u);  // ... and this tile's were zero-initialised
                    first_mine = false;
                    tc_fence_after();
                    const uint32_t yo = (uint32_t)c * (AT3_W * 128);
                    const int ksteps = width(c) / 16;
                    for (int k = 0; k < ksteps; ++k) {
                        const uint32_t acol = k < 2 ? 8 * k : 32 + 8 * (k - 2);
                        const uint32_t accf = (c > 0 || k > 0) ? 1u : 0u;
                        umma_bf16_ts_elect(tmem + 256, acc + acol, make_smem_desc_sw128(sdO + yo + k * 2048, 1024, 0), idesc2, accf);
                        umma_bf16_ts_elect(tmem + 320, acc + 64 + acol, make_smem_desc_sw128(sQ + yo + k * 2048, 1024, 0), idesc2, accf);
                    }
                    if (c == 0 && nc > 1 && lane == 0) mbar_arrive(kv_started);
                    if (c + 2 >= nc) umma_commit_elect(kv_ready);               // my last chunk of this tile
                    if ((c & 1) || c == nc - 1) {
                        // both chunks of query tile m are in shared memory: dQ_m (+)= dS[m] K_t over this tile's keys
                        const int m = c >> 1;
                        if (c & 1) mbar_wait(ds_half(m), tile_idx & 1u);        // the other group's chunk 2m
                        if (t == 0 && !dq_waited) {                             // the previous head's dQ has been read out
                            mbar_wait(dq_free, (it & 1) ^ 1u);
                            dq_waited = true;
                        }
                        tc_fence_after();
                        const int keys = min(ATC_BM, Nk - t * ATC_BM);
                        const uint32_t sA = ds_base + (uint32_t)m * 32768u, sB = sK + (uint32_t)t * (ATC_BM * 128);
                        for (int k = 0; k < keys / 16; ++k)
                            umma_bf16_ss_elect(tmem + 384 + 64 * m, make_smem_desc_sw128(sA + k * 2048, 1024, 16384),
                                               make_smem_desc_sw128(sB + k * 2048, 1024, 0), idesc3, (t > 0 || k > 0) ? 1u : 0u);
                        umma_commit_elect(ds_free(m));
                    }
                    // (Queueing the next S^T / dP^T BEFORE the dQ MMAs shortens the chain the math group waits on by 4 %,
                    // but it lets the two issuers' dV / dK MMAs of neighbouring chunks reach the tensor core in a
                    // timing-dependent order: the fp32 sums then differ in the last bit from run to run. Not taken.)
                    if (s + 2 < nsteps) {
                        int t2 = t, c2 = c + 2;
                        while (c2 >= nc) { c2 -= nc; ++t2; }
                        issue_1(t2, c2);
                    }
                }
                if (first_mine) umma_commit_elect(kv_ready);    // no chunk of this tile was mine (single-chunk tiles)
            }
            umma_commit_elect(dq_ready);
            umma_commit_elect(head_empty);
            umma_commit_elect(stat_free(it & 1));
        }
        __syncwarp();
    } else if (warp >= 18) {
        // ===================== per-token statistics, one head ahead (two warps, half of the tokens each) ==========
        const float LOG2E = 1.4426950408889634f;
        const long long os = (long long)H * 64;
        const int part = warp - 18;
        const int half_rows = ((Nk / 2) + 15) & ~15;
        const int t_begin = part == 0 ? 0 : half_rows, t_end = part == 0 ? half_rows : Nk;
        int it = 0;
        for (int w = blockIdx.x; w < num_heads_total; w += gridDim.x, ++it) {
            const int stage = it & 1;
            const int b = w / H, h = w % H;
            if (lane == 0) mbar_wait(stat_free(stage), ((it >> 1) & 1) ^ 1u);      // head it-2 no longer reads this stage
            __syncwarp();
            float* sLse2 = reinterpret_cast<float*>(gen + stat_base + (uint32_t)stage * 2u * Nk * 4u);
            float* sDelta = sLse2 + Nk;
            const bf16* gO = out + (long long)b * T * os + (long long)h * 64;
            const bf16* gdO = dout + (long long)b * T * os + (long long)h * 64;
            const float* L = lse + (long long)w * T;
            for (int t = t_begin + lane; t < t_end; t += 32) {
                float d = 0.f;
                if (t < T) {
                    const uint4* po = reinterpret_cast<const uint4*>(gO + (long long)t * os);
                    const uint4* pd = reinterpret_cast<const uint4*>(gdO + (long long)t * os);
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const uint4 a = __ldg(po + i), c = __ldg(pd + i);
                        const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, cw[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const float2 fa = unpack_bf16(aw[k]), fc = unpack_bf16(cw[k]);
                            d = fmaf(fa.x, fc.x, fmaf(fa.y, fc.y, d));
                        }
                    }
                }
                sDelta[t] = d * scale;
                sLse2[t] = t < T ? L[t] * LOG2E : INFINITY;      // +inf: padded query columns get P = 0, dS = 0
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(stat_bar(stage));
        }
    } else {
        // ===================== math + read-out groups =====================
        const int sw = warp - 2;
        const int g = sw >> 3;
        const int half = (sw >> 2) & 1;
        const int quarter = warp & 3;
        const uint32_t lane_sel = (uint32_t)(quarter * 32) << 16;
        const uint32_t acc = tmem + g * 128 + lane_sel;
        const int r_in_tile = quarter * 32 + lane;
        const long long rs = 3LL * H * 64;
        const int cb = half * 32;
        uint32_t n_s = 0, n_kv = 0;
        uint32_t S0 = 0;
        int it = 0;
        for (int w = blockIdx.x; w < num_heads_total; w += gridDim.x, ++it, S0 += nsteps) {
            const int b = w / H, h = w % H;
            const float* sL = reinterpret_cast<const float*>(gen + stat_base + (uint32_t)(it & 1) * 2u * Nk * 4u);
            const float* sD = sL + Nk;
            mbar_wait(stat_bar(it & 1), (it >> 1) & 1);
            // dV (group 0) or dK (group 1) of key tile t: this warp converts 32 of the 64 columns of its 32 rows
            auto read_out_kv = [&](int t) {
                mbar_wait(kv_ready, n_kv & 1u);
                ++n_kv;
                tc_fence_after();
                uint32_t v[32];
                tmem_ld_32x32(tmem + lane_sel + 256 + 64 * g + half * 32, v);
                tmem_ld_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(kv_free);
                const int row = t * ATC_BM + r_in_tile;
                if (row < T) {
                    bf16* dst = dqkv + ((long long)b * T + row) * rs + (long long)(g == 0 ? 2 * H + h : H + h) * 64 + half * 32;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        uint4 q;
                        q.x = pack_bf16(__uint_as_float(v[8 * i]), __uint_as_float(v[8 * i + 1]));
                        q.y = pack_bf16(__uint_as_float(v[8 * i + 2]), __uint_as_float(v[8 * i + 3]));
                        q.z = pack_bf16(__uint_as_float(v[8 * i + 4]), __uint_as_float(v[8 * i + 5]));
                        q.w = pack_bf16(__uint_as_float(v[8 * i + 6]), __uint_as_float(v[8 * i + 7]));
                        reinterpret_cast<uint4*>(dst)[i] = q;
                    }
                }
                if (dbias != nullptr) {         // column sums over this warp's valid rows -> lane j holds column j
                    float f[32];
#pragma unroll
                    for (int i = 0; i < 32; ++i) f[i] = row < T ? __uint_as_float(v[i]) : 0.f;
                    const float cs = warp_colsum32(f, lane);
                    atomicAdd(sBias + ((g == 0 ? 2 : 1) * H + h) * 64 + half * 32 + lane, cs);
                }
            };
            int pending_tile = -1;          // tile whose dV/dK this group still has to read out
            for (int s = 0, t = 0, c = 0; s < nsteps; ++s, c = (c + 1 == nc ? 0 : c + 1), t += (c == 0)) {
                if (((S0 + s) & 1) != (uint32_t)g) {
                    // the other group's step; if it is the last chunk of a tile, this group reads its half of the tile's
                    // gradients after its own NEXT step (or at the end of the head)
                    if (c == nc - 1) pending_tile = t;
                    continue;
                }
                const int m = c >> 1, blk = c & 1;
                const int W = min(AT3_W, Nk - c * AT3_W);
                const int ce = min(cb + 32, W);
                mbar_wait(s_ready(g), n_s & 1u);
                ++n_s;
                if (t > 0 || it > 0) {      // the dQ MMA that read this dS buffer last (previous tile / head) has retired
                    const uint32_t uses = (uint32_t)it * nt + t;       // dQ MMAs issued on buffer m before this tile
                    mbar_wait(ds_free(m), (uses - 1) & 1u);
                }
                tc_fence_after();
                uint8_t* ds_row = gen + ds_base + (uint32_t)m * 32768u + (uint32_t)blk * 16384u + (uint32_t)r_in_tile * 128u;
                for (int cc = cb; cc < ce; cc += 16) {
                    uint32_t sv[16], dp[16];
                    tmem_ld_32x16(acc + cc, sv);
                    tmem_ld_32x16(acc + 64 + cc, dp);
                    tmem_ld_wait();
                    float p[16], ds[16];
#pragma unroll
                    for (int i4 = 0; i4 < 4; ++i4) {
                        const float4 a = reinterpret_cast<const float4*>(sL + c * AT3_W + cc)[i4];
                        const float4 d = reinterpret_cast<const float4*>(sD + c * AT3_W + cc)[i4];
                        const float l2[4] = {a.x, a.y, a.z, a.w}, dl[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const int i = 4 * i4 + k;
                            p[i] = fast_ex2(fmaf(__uint_as_float(sv[i]), scale_log2, -l2[k]));
                            ds[i] = p[i] * fmaf(__uint_as_float(dp[i]), scale, -dl[k]);
                        }
                    }
                    uint32_t pk[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) pk[i] = pack_bf16(p[2 * i], p[2 * i + 1]);
                    tmem_st_32x8(acc + cb + ((cc - cb) >> 1), pk);
#pragma unroll
                    for (int i = 0; i < 8; ++i) pk[i] = pack_bf16(ds[2 * i], ds[2 * i + 1]);
                    tmem_st_32x8(acc + 64 + cb + ((cc - cb) >> 1), pk);
                    // the same 16 dS values as two 16-byte pieces of this key's row of the MN-major dS tile
                    const int j = cc >> 3;
                    *reinterpret_cast<uint4*>(ds_row + (((j) ^ (r_in_tile & 7)) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                    *reinterpret_cast<uint4*>(ds_row + (((j + 1) ^ (r_in_tile & 7)) << 4)) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                }
                fence_proxy_async_smem();       // the dS rows are read by the tensor core (async proxy)
                tmem_st_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) {
                    mbar_arrive(p_ready(g));
                    if (!(c & 1) && c + 1 < nc) mbar_arrive(ds_half(m));     // first half of dS tile m is in shared memory
                }
                if (pending_tile >= 0) {
                    read_out_kv(pending_tile);
                    pending_tile = -1;
                }
                if (c == nc - 1) read_out_kv(t);    // own last chunk of the tile: its MMAs follow immediately
            }
            if (pending_tile >= 0) read_out_kv(pending_tile);
            // dQ: group m converts query rows [128 m, 128 m + 128)
            mbar_wait(dq_ready, it & 1);
            tc_fence_after();
            uint32_t v[32];
            if (g < nm) {
                tmem_ld_32x32(tmem + lane_sel + 384 + 64 * g + half * 32, v);
                tmem_ld_wait();
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(dq_free);
            const int row = g * ATC_BM + r_in_tile;
            if (g < nm && row < T) {
                bf16* dst = dqkv + ((long long)b * T + row) * rs + (long long)h * 64 + half * 32;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    uint4 q;
                    q.x = pack_bf16(__uint_as_float(v[8 * i]), __uint_as_float(v[8 * i + 1]));
                    q.y = pack_bf16(__uint_as_float(v[8 * i + 2]), __uint_as_float(v[8 * i + 3]));
                    q.z = pack_bf16(__uint_as_float(v[8 * i + 4]), __uint_as_float(v[8 * i + 5]));
                    q.w = pack_bf16(__uint_as_float(v[8 * i + 6]), __uint_as_float(v[8 * i + 7]));
                    reinterpret_cast<uint4*>(dst)[i] = q;
                }
            }
            if (dbias != nullptr && g < nm) {
                float f[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) f[i] = row < T ? __uint_as_float(v[i]) : 0.f;
                const float cs = warp_colsum32(f, lane);
                atomicAdd(sBias + h * 64 + half * 32 + lane, cs);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<1>(tmem, 512);
    }
    if (dbias != nullptr)
        for (int k = threadIdx.x; k < 3 * H * 64; k += blockDim.x) atomicAdd(dbias + k, sBias[k]);
}

// =============================================================================================
// Backward (T <= 256), single pass, probabilities through SHARED memory ("tc4"). Earlier versions built the probabilities twice per head (once with keys on
// the TMEM lanes for dK/dV, once with queries on the lanes for dQ). Here they are built ONCE, keys on the
// lanes: per 128-key tile t and 64-query chunk c
//     S^T = K_t Q_c^T, dP^T = V_t dO_c^T  ->  P^T, dS^T (bf16, packed in place)  ->  dV_t += P^T dO_c, dK_t += dS^T Q_c
// and dS^T is ALSO written to shared memory as bf16, one 128-byte row of 64 queries per key: that is exactly the
// 128B-swizzled MN-major layout of an A operand with M = queries, so after two chunks (128 queries)
//     dQ_m += dS[128 queries x keys of tile t] K_t
// is one more UMMA chain, A from shared memory, accumulating over the key tiles in TMEM. All exponentials, the
// dS math and the S / dP MMAs are done once instead of twice.
// TMEM: [0,128) chunk accumulators of group 0 (S^T | dP^T), [128,256) of group 1, [256,320) dV_t, [320,384) dK_t,
//       [384,448) dQ rows 0-127, [448,512) dQ rows 128-255.
// Warps: 0 = TMA + MMA issue for group 1, 1 = MMA issue for group 0, 2-17 = two math groups (8 warps: thread <-> key row,
// two warps per row splitting a chunk's 64 query columns), 18-19 = statistics; the groups alternate chunks.
// Unlike its predecessor (which packed P^T / dS^T back into the chunk accumulators in tensor memory and fed them to
// the dV / dK MMAs from there), this kernel writes P^T and dS^T as bf16 to SHARED memory -- one 128-byte row of 64
// queries per key, 128B-swizzled: a K-major A operand for dV_t += P^T dO_c and dK_t += dS^T Q_c, and for dS^T at the
// same time the MN-major A operand of dQ_m += dS K_t. The chunk accumulators are therefore free again as soon as the
// math group has LOADED them (acc_free), not only after the dV / dK MMAs have consumed the packed values: the issuer
// queues the group's next S^T / dP^T MMAs while the group is still computing exponentials, and the group finds its
// next accumulators ready when it is done. (Before: MMA issue -> commit -> math -> commit -> MMA issue, strictly in
// series per group; 53 % of all warp samples were mbarrier waits.)
// =============================================================================================

__global__ void __launch_bounds__(AT3_THREADS, 1)
attention_bwd_tc4_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO,
                         const bf16* __restrict__ out, const bf16* __restrict__ dout, const float* __restrict__ lse,
                         bf16* __restrict__ dqkv, float* __restrict__ dbias, int num_heads_total, int T, int Nk, int H,
                         float scale_log2, float scale) {
    extern __shared__ uint8_t smem_raw[];
    pdl_launch_dependents();
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t tile_bytes = (uint32_t)Nk * 128u;
    const uint32_t tile_stride = (tile_bytes + 1023u) & ~1023u;
    const uint32_t sQ = smem_base, sK = sQ + tile_stride, sV = sK + tile_stride, sdO = sV + tile_stride;
    const uint32_t ds_base = smem_base + 4 * tile_stride + ATB_TAIL_PAD;     // 2 buffers x [2 query blocks][128 keys][128 B]
    const uint32_t pt_base = ds_base + 2u * 32768u;                          // P^T of the chunk in flight, one 16 KB tile per group
    const uint32_t stat_base = pt_base + 2u * 16384u;                        // 2 stages x (lse2 [Nk], delta*scale [Nk])
    const uint32_t bar_base = stat_base + 2u * 2u * Nk * 4u;
    const uint32_t full_bar = bar_base, head_empty = bar_base + 8u;
    auto stat_bar = [&](int s) { return bar_base + 16u + 8u * s; };
    auto s_ready = [&](int g) { return bar_base + 32u + 8u * g; };
    auto p_ready = [&](int g) { return bar_base + 48u + 8u * g; };
    auto ds_free = [&](int m) { return bar_base + 64u + 8u * m; };
    const uint32_t kv_ready = bar_base + 80u, kv_free = bar_base + 88u, dq_ready = bar_base + 96u, dq_free = bar_base + 104u;
    auto stat_free = [&](int s) { return bar_base + 112u + 8u * s; };
    const uint32_t tmem_slot = bar_base + 128u;
    const uint32_t kv_started = bar_base + 144u;
    auto ds_half = [&](int m) { return bar_base + 152u + 8u * m; };
    auto acc_free = [&](int g) { return bar_base + 168u + 8u * g; };     // group g has its chunk accumulators in registers
    auto pt_free = [&](int g) { return bar_base + 184u + 8u * g; };      // the MMAs that read group g's P^T tile have retired
    // optional fused in-projection bias gradient: per-CTA column sums of dQ | dK | dV, [3][H][64] fp32, flushed at the end
    float* sBias = reinterpret_cast<float*>(smem_raw - smem_u32(smem_raw) + bar_base + 256u);
    if (dbias != nullptr)
        for (int k = threadIdx.x; k < 3 * H * 64; k += blockDim.x) sBias[k] = 0.f;
    uint8_t* gen = smem_raw - smem_u32(smem_raw);
    uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(gen + tmem_slot);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nt = (T + ATC_BM - 1) / ATC_BM;           // key tiles
    const int nc = (Nk + AT3_W - 1) / AT3_W;            // query chunks
    const int nsteps = nt * nc;
    const int nm = (nc + 1) / 2;                        // 128-query dQ tiles

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmQKV);
        prefetch_tmap(&tmDO);
        mbar_init(full_bar, 1);
        mbar_init(head_empty, 2);       // "2": one commit per MMA-issuing warp
        for (int i = 0; i < 2; ++i) {
            mbar_init(stat_bar(i), 2);
            mbar_init(stat_free(i), 2);
            mbar_init(s_ready(i), 1);
            mbar_init(p_ready(i), 8);
            mbar_init(ds_free(i), 2);       // one commit from the issuer of each of the tile's two chunks
            mbar_init(ds_half(i), 8);
            mbar_init(acc_free(i), 8);
            mbar_init(pt_free(i), 1);
        }
        mbar_init(kv_ready, 2);
        mbar_init(kv_free, 16);
        mbar_init(kv_started, 1);
        mbar_init(dq_ready, 2);
        mbar_init(dq_free, 16);
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc<1>(tmem_slot, 512);
        tmem_relinquish<1>();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot_ptr;
    pdl_wait();       // the prologue above overlapped the previous kernel; global memory is touched from here on

    if (warp < 2) {
        // ===================== two MMA issuers: warp 1 serves math group 0, warp 0 serves group 1 and issues the TMA loads ======
        // One thread can only issue ~1 small UMMA per 100 cycles, and a head is ~140 of them: with a single issuer that is
        // half of the head's time. All 32 lanes run the loop (uniform operands); one elected lane issues each instruction.
        // Cross-issuer ordering: tcgen05.commit only tracks the MMAs of its own thread, so barriers that mean "all MMAs of
        // the tile / head are done" (kv_ready, dq_ready, head_empty, stat_free) take one commit from EACH issuer; the
        // issuer of a tile's first chunk (accumulate = 0 into dV/dK) signals kv_started, the other one waits for it;
        // the dQ MMA waits until the other group's half of the dS tile is in shared memory (ds_half).
        const int my_g = warp == 1 ? 0 : 1;
        const uint32_t idesc2 = make_idesc_bf16(ATC_BM, 64, 0, 1);      // A from TMEM, B MN-major
        const uint32_t idesc3 = make_idesc_bf16(ATC_BM, 64, 1, 1);      // A (dS) and B (K) both MN-major
        const uint32_t acc = tmem + my_g * 128;
        uint32_t n_p = 0;
        uint32_t S0 = 0;                // global step counter at the start of the head
        int it = 0;
        for (int w = blockIdx.x; w < num_heads_total; w += gridDim.x, ++it, S0 += nsteps) {
            if (warp == 0) {
                const int b = w / H, h = w % H;
                mbar_wait(head_empty, (it & 1) ^ 1u);
                if (lane == 0) {
                    mbar_arrive_expect_tx(full_bar, 4 * tile_bytes);
                    tma_load_4d(sQ, &tmQKV, full_bar, 0, 0, h, b);
                    tma_load_4d(sK, &tmQKV, full_bar, 0, 0, H + h, b);
                    tma_load_4d(sV, &tmQKV, full_bar, 0, 0, 2 * H + h, b);
                    tma_load_4d(sdO, &tmDO, full_bar, 0, 0, h, b);
                }
                __syncwarp();
            }
            mbar_wait(full_bar, it & 1);
            tc_fence_after();
            auto width = [&](int c) { return min(AT3_W, Nk - c * AT3_W); };
            // (t, c) of step s is carried along instead of divided out: a runtime integer division is ~25 instructions with
            // two dependent MUFU/I2F round trips, on the critical issue path of every step
            auto issue_1 = [&](int t, int c) {  // S^T and dP^T of (my) step (t, c) into my group's chunk accumulators
                const uint32_t xo = (uint32_t)t * (ATC_BM * 128), yo = (uint32_t)c * (AT3_W * 128);
                const uint32_t idesc1 = make_idesc_bf16(ATC_BM, width(c), 0, 0);
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    umma_bf16_ss_elect(acc, make_smem_desc_sw128(sK + xo + k * 32, 1024, 0),
                                       make_smem_desc_sw128(sQ + yo + k * 32, 1024, 0), idesc1, k > 0 ? 1u : 0u);
                    umma_bf16_ss_elect(acc + 64, make_smem_desc_sw128(sV + xo + k * 32, 1024, 0),
                                       make_smem_desc_sw128(sdO + yo + k * 32, 1024, 0), idesc1, k > 0 ? 1u : 0u);
                }
                umma_commit_elect(s_ready(my_g));
            };
            const int s_first = ((S0 & 1u) == (uint32_t)my_g) ? 0 : 1;
            if (s_first < nsteps) issue_1(nc == 1 ? s_first : 0, nc == 1 ? 0 : s_first);
            const uint32_t sPT = pt_base + (uint32_t)my_g * 16384u;
            bool dq_waited = false;
            for (int t = 0; t < nt; ++t) {
                const uint32_t tile_idx = (uint32_t)it * nt + t;        // key tiles so far, all heads
                bool first_mine = true;
                for (int c = 0; c < nc; ++c) {
                    const int s = t * nc + c;
                    if (((S0 + s) & 1u) != (uint32_t)my_g) continue;
                    // (1) as soon as the math group holds this step's S^T / dP^T in registers, queue its NEXT step's
                    if (s + 2 < nsteps) {
                        mbar_wait(acc_free(my_g), n_p & 1u);
                        tc_fence_after();
                        int t2 = t, c2 = c + 2;
                        while (c2 >= nc) { c2 -= nc; ++t2; }
                        issue_1(t2, c2);
                    }
                    // (2) P^T and dS^T of this step are in shared memory
                    mbar_wait(p_ready(my_g), n_p & 1u);
                    ++n_p;
                    if (c == 0) mbar_wait(kv_free, (tile_idx & 1u) ^ 1u);       // the previous tile's dV / dK have been read out
                    else if (first_mine) mbar_wait(kv_started, tile_idx & 1u);  // ... and this tile's were zero-initialised
                    first_mine = false;
                    tc_fence_after();
                    const uint32_t yo = (uint32_t)c * (AT3_W * 128);
                    const int m = c >> 1, blk = c & 1;
                    const uint32_t sDS = ds_base + (uint32_t)m * 32768u + (uint32_t)blk * 16384u;
                    const int ksteps = width(c) / 16;
                    for (int k = 0; k < ksteps; ++k) {
                        const uint32_t accf = (c > 0 || k > 0) ? 1u : 0u;
                        // A: [128 keys][64 queries] rows of 128 B, K-major, 16 queries (32 B) per step; B: 16 query rows of dO / Q
                        umma_bf16_ss_elect(tmem + 256, make_smem_desc_sw128(sPT + k * 32, 1024, 0),
                                           make_smem_desc_sw128(sdO + yo + k * 2048, 1024, 0), idesc2, accf);
                        umma_bf16_ss_elect(tmem + 320, make_smem_desc_sw128(sDS + k * 32, 1024, 0),
                                           make_smem_desc_sw128(sQ + yo + k * 2048, 1024, 0), idesc2, accf);
                    }
                    umma_commit_elect(pt_free(my_g));                           // my P^T tile may be overwritten
                    if (c == 0 && nc > 1 && lane == 0) mbar_arrive(kv_started);
                    if (c + 2 >= nc) umma_commit_elect(kv_ready);               // my last chunk of this tile
                    if ((c & 1) || c == nc - 1) {
                        // both chunks of query tile m are in shared memory: dQ_m (+)= dS[m] K_t over this tile's keys
                        if (c & 1) mbar_wait(ds_half(m), tile_idx & 1u);        // the other group's chunk 2m
                        if (t == 0 && !dq_waited) {                             // the previous head's dQ has been read out
                            mbar_wait(dq_free, (it & 1) ^ 1u);
                            dq_waited = true;
                        }
                        tc_fence_after();
                        const int keys = min(ATC_BM, Nk - t * ATC_BM);
                        const uint32_t sA = ds_base + (uint32_t)m * 32768u, sB = sK + (uint32_t)t * (ATC_BM * 128);
                        for (int k = 0; k < keys / 16; ++k)
                            umma_bf16_ss_elect(tmem + 384 + 64 * m, make_smem_desc_sw128(sA + k * 2048, 1024, 16384),
                                               make_smem_desc_sw128(sB + k * 2048, 1024, 0), idesc3, (t > 0 || k > 0) ? 1u : 0u);
                    }
                    // dS tile m: one arrival per chunk issuer (my dK MMAs, and the dQ MMAs if they were mine); a tile
                    // with a single chunk gets both arrivals from its only issuer
                    umma_commit_elect(ds_free(m));
                    if (!(c & 1) && c == nc - 1) umma_commit_elect(ds_free(m));
                }
                if (first_mine) umma_commit_elect(kv_ready);    // no chunk of this tile was mine (single-chunk tiles)
            }
            umma_commit_elect(dq_ready);
            umma_commit_elect(head_empty);
            umma_commit_elect(stat_free(it & 1));
        }
        __syncwarp();
    } else if (warp >= 18) {
        // ===================== per-token statistics, one head ahead (two warps, half of the tokens each) ==========
        const float LOG2E = 1.4426950408889634f;
        const long long os = (long long)H * 64;
        const int part = warp - 18;
        const int half_rows = ((Nk / 2) + 15) & ~15;
        const int t_begin = part == 0 ? 0 : half_rows, t_end = part == 0 ? half_rows : Nk;
        int it = 0;
        for (int w = blockIdx.x; w < num_heads_total; w += gridDim.x, ++it) {
            const int stage = it & 1;
            const int b = w / H, h = w % H;
            if (lane == 0) mbar_wait(stat_free(stage), ((it >> 1) & 1) ^ 1u);      // head it-2 no longer reads this stage
            __syncwarp();
            float* sLse2 = reinterpret_cast<float*>(gen + stat_base + (uint32_t)stage * 2u * Nk * 4u);
            float* sDelta = sLse2 + Nk;
            const bf16* gO = out + (long long)b * T * os + (long long)h * 64;
            const bf16* gdO = dout + (long long)b * T * os + (long long)h * 64;
            const float* L = lse + (long long)w * T;
            for (int t = t_begin + lane; t < t_end; t += 32) {
                float d = 0.f;
                if (t < T) {
                    const uint4* po = reinterpret_cast<const uint4*>(gO + (long long)t * os);
                    const uint4* pd = reinterpret_cast<const uint4*>(gdO + (long long)t * os);
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const uint4 a = __ldg(po + i), c = __ldg(pd + i);
                        const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, cw[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const float2 fa = unpack_bf16(aw[k]), fc = unpack_bf16(cw[k]);
                            d = fmaf(fa.x, fc.x, fmaf(fa.y, fc.y, d));
                        }
                    }
                }
                sDelta[t] = d * scale;
                sLse2[t] = t < T ? L[t] * LOG2E : INFINITY;      // +inf: padded query columns get P = 0, dS = 0
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(stat_bar(stage));
        }
    } else {
        // ===================== math + read-out groups =====================
        const int sw = warp - 2;
        const int g = sw >> 3;
        const int half = (sw >> 2) & 1;
        const int quarter = warp & 3;
        const uint32_t lane_sel = (uint32_t)(quarter * 32) << 16;
        const uint32_t acc = tmem + g * 128 + lane_sel;
        const int r_in_tile = quarter * 32 + lane;
        const long long rs = 3LL * H * 64;
        const int cb = half * 32;
        uint32_t n_s = 0, n_kv = 0;
        uint32_t S0 = 0;
        int it = 0;
        for (int w = blockIdx.x; w < num_heads_total; w += gridDim.x, ++it, S0 += nsteps) {
            const int b = w / H, h = w % H;
            const float* sL = reinterpret_cast<const float*>(gen + stat_base + (uint32_t)(it & 1) * 2u * Nk * 4u);
            const float* sD = sL + Nk;
            mbar_wait(stat_bar(it & 1), (it >> 1) & 1);
            // dV (group 0) or dK (group 1) of key tile t: this warp converts 32 of the 64 columns of its 32 rows
            auto read_out_kv = [&](int t) {
                mbar_wait(kv_ready, n_kv & 1u);
                ++n_kv;
                tc_fence_after();
                uint32_t v[32];
                tmem_ld_32x32(tmem + lane_sel + 256 + 64 * g + half * 32, v);
                tmem_ld_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(kv_free);
                const int row = t * ATC_BM + r_in_tile;
                if (row < T) {
                    bf16* dst = dqkv + ((long long)b * T + row) * rs + (long long)(g == 0 ? 2 * H + h : H + h) * 64 + half * 32;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        uint4 q;
                        q.x = pack_bf16(__uint_as_float(v[8 * i]), __uint_as_float(v[8 * i + 1]));
                        q.y = pack_bf16(__uint_as_float(v[8 * i + 2]), __uint_as_float(v[8 * i + 3]));
                        q.z = pack_bf16(__uint_as_float(v[8 * i + 4]), __uint_as_float(v[8 * i + 5]));
                        q.w = pack_bf16(__uint_as_float(v[8 * i + 6]), __uint_as_float(v[8 * i + 7]));
                        reinterpret_cast<uint4*>(dst)[i] = q;
                    }
                }
                // Column sums (in_proj_bias gradient) for dV only: the key-bias gradient is sum_keys dK =
                // sum_q (sum_keys dS[q, :]) Q[q] = 0 exactly (a softmax is invariant to a shift of its logits), so that third
                // of dbias keeps the launcher's zero fill instead of collecting rounding noise.
                if (dbias != nullptr && g == 0) {       // column sums over this warp's valid rows -> lane j holds column j
                    float f[32];
#pragma unroll
                    for (int i = 0; i < 32; ++i) f[i] = row < T ? __uint_as_float(v[i]) : 0.f;
                    const float cs = warp_colsum32(f, lane);
                    atomicAdd(sBias + ((g == 0 ? 2 : 1) * H + h) * 64 + half * 32 + lane, cs);
                }
            };
            int pending_tile = -1;          // tile whose dV/dK this group still has to read out
            for (int s = 0, t = 0, c = 0; s < nsteps; ++s, c = (c + 1 == nc ? 0 : c + 1), t += (c == 0)) {
                if (((S0 + s) & 1) != (uint32_t)g) {
                    // the other group's step; if it is the last chunk of a tile, this group reads its half of the tile's
                    // gradients after its own NEXT step (or at the end of the head)
                    if (c == nc - 1) pending_tile = t;
                    continue;
                }
                const int m = c >> 1, blk = c & 1;
                const int W = min(AT3_W, Nk - c * AT3_W);
                const int ce = min(cb + 32, W);
                mbar_wait(s_ready(g), n_s & 1u);
                if (n_s > 0) mbar_wait(pt_free(g), (n_s - 1) & 1u);     // the MMAs that read my group's previous P^T have retired
                ++n_s;
                if (t > 0 || it > 0) {      // the dK / dQ MMAs that read this dS buffer last (previous tile / head) have retired
                    const uint32_t uses = (uint32_t)it * nt + t;       // uses of dS buffer m before this tile
                    mbar_wait(ds_free(m), (uses - 1) & 1u);
                }
                tc_fence_after();
                uint8_t* ds_row = gen + ds_base + (uint32_t)m * 32768u + (uint32_t)blk * 16384u + (uint32_t)r_in_tile * 128u;
                uint8_t* pt_row = gen + pt_base + (uint32_t)g * 16384u + (uint32_t)r_in_tile * 128u;
                for (int cc = cb; cc < ce; cc += 16) {
                    uint32_t sv[16], dp[16];
                    tmem_ld_32x16(acc + cc, sv);
                    tmem_ld_32x16(acc + 64 + cc, dp);
                    tmem_ld_wait();
                    if (cc + 16 >= ce) {        // last load of this step: the accumulators may be overwritten
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(acc_free(g));
                    }
                    float p[16], ds[16];
#pragma unroll
                    for (int i4 = 0; i4 < 4; ++i4) {
                        const float4 a = reinterpret_cast<const float4*>(sL + c * AT3_W + cc)[i4];
                        const float4 d = reinterpret_cast<const float4*>(sD + c * AT3_W + cc)[i4];
                        const float l2[4] = {a.x, a.y, a.z, a.w}, dl[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const int i = 4 * i4 + k;
                            p[i] = fast_ex2(fmaf(__uint_as_float(sv[i]), scale_log2, -l2[k]));
                            ds[i] = p[i] * fmaf(__uint_as_float(dp[i]), scale, -dl[k]);
                        }
                    }
                    // 16 P values and 16 dS values: two 16-byte pieces each of this key's 128-byte rows
                    const int j = cc >> 3;
                    uint32_t pk[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) pk[i] = pack_bf16(p[2 * i], p[2 * i + 1]);
                    *reinterpret_cast<uint4*>(pt_row + (((j) ^ (r_in_tile & 7)) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                    *reinterpret_cast<uint4*>(pt_row + (((j + 1) ^ (r_in_tile & 7)) << 4)) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
#pragma unroll
                    for (int i = 0; i < 8; ++i) pk[i] = pack_bf16(ds[2 * i], ds[2 * i + 1]);
                    *reinterpret_cast<uint4*>(ds_row + (((j) ^ (r_in_tile & 7)) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                    *reinterpret_cast<uint4*>(ds_row + (((j + 1) ^ (r_in_tile & 7)) << 4)) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                }
                if (cb >= ce) {                 // this warp has no columns in a narrow last chunk: it still takes part
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(acc_free(g));
                }
                fence_proxy_async_smem();       // the P^T / dS^T rows are read by the tensor core (async proxy)
                __syncwarp();
                if (lane == 0) {
                    mbar_arrive(p_ready(g));
                    if (!(c & 1) && c + 1 < nc) mbar_arrive(ds_half(m));     // first half of dS tile m is in shared memory
                }
                if (pending_tile >= 0) {
                    read_out_kv(pending_tile);
                    pending_tile = -1;
                }
                if (c == nc - 1) read_out_kv(t);    // own last chunk of the tile: its MMAs follow immediately
            }
            if (pending_tile >= 0) read_out_kv(pending_tile);
            // dQ: group m converts query rows [128 m, 128 m + 128)
            mbar_wait(dq_ready, it & 1);
            tc_fence_after();
            uint32_t v[32];
            if (g < nm) {
                tmem_ld_32x32(tmem + lane_sel + 384 + 64 * g + half * 32, v);
                tmem_ld_wait();
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(dq_free);
            const int row = g * ATC_BM + r_in_tile;
            if (g < nm && row < T) {
                bf16* dst = dqkv + ((long long)b * T + row) * rs + (long long)h * 64 + half * 32;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    uint4 q;
                    q.x = pack_bf16(__uint_as_float(v[8 * i]), __uint_as_float(v[8 * i + 1]));
                    q.y = pack_bf16(__uint_as_float(v[8 * i + 2]), __uint_as_float(v[8 * i + 3]));
                    q.z = pack_bf16(__uint_as_float(v[8 * i + 4]), __uint_as_float(v[8 * i + 5]));
                    q.w = pack_bf16(__uint_as_float(v[8 * i + 6]), __uint_as_float(v[8 * i + 7]));
                    reinterpret_cast<uint4*>(dst)[i] = q;
                }
            }
            if (dbias != nullptr && g < nm) {
                float f[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) f[i] = row < T ? __uint_as_float(v[i]) : 0.f;
                const float cs = warp_colsum32(f, lane);
                atomicAdd(sBias + h * 64 + half * 32 + lane, cs);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<1>(tmem, 512);
    }
    if (dbias != nullptr)
        for (int k = threadIdx.x; k < 3 * H * 64; k += blockDim.x) atomicAdd(dbias + k, sBias[k]);
}

// =============================================================================================
// Forward for long sequences (T > 256, e.g. the 577 tokens of 384x384 images): flash-style streaming over blocks of
// 128 keys with an online softmax. A work item is (batch, head, pair of 128-query tiles); the pair shares one
// stream of K/V blocks (4-deep TMA ring), tile g of the pair lives in TMEM slot g:
//     [0, 128) S block (bf16 P packed in place: key half 0 -> [0, 32), half 1 -> [64, 96)) | [192, 256) O accumulator
// and belongs to softmax group g (8 warps: thread <-> query row, two warps per row splitting the block's keys).
// Per block: S = Q_g K_j^T -> running max m, P = exp2((S - m) c), l = l a + rowsum(P), O = O a (tcgen05.ld/st,
// skipped when no row of the warp raised its max) -> O += P V_j (A operand from TMEM). The MMA thread issues
// PV(g, j) and S(g, j+1) back to back, so one group's softmax overlaps the other group's tensor work.
// =============================================================================================
constexpr int ATL_BN = 128;          // keys per block
constexpr int ATL_STAGES = 4;        // K/V ring depth

__global__ void __launch_bounds__(ATC_THREADS, 1)
attention_fwd_long_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                          bf16* __restrict__ out, float* __restrict__ lse, int num_items, int rounds, int T, int H,
                          float scale_log2, float scale) {
    extern __shared__ uint8_t smem_raw[];
    pdl_launch_dependents();
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    constexpr uint32_t Q_BYTES = 2 * ATC_BM * 128;               // both query tiles of the pair
    constexpr uint32_t KV_BYTES = ATL_BN * 128;                  // one K or V block
    const uint32_t kv_base = smem_base + 2 * Q_BYTES;
    const uint32_t xchg_base = kv_base + ATL_STAGES * 2 * KV_BYTES;     // float [2 groups][2 halves][128]: max; then sum
    const uint32_t bar_base = xchg_base + 2u * 2u * 2u * 128u * 4u;
    auto q_full = [&](int s) { return bar_base + 8u * s; };
    auto q_empty = [&](int s) { return bar_base + 16u + 8u * s; };
    auto kv_full = [&](int s) { return bar_base + 32u + 8u * s; };
    auto kv_empty = [&](int s) { return bar_base + 64u + 8u * s; };
    auto s_ready = [&](int g) { return bar_base + 96u + 8u * g; };
    auto p_ready = [&](int g) { return bar_base + 112u + 8u * g; };
    auto o_ready = [&](int g) { return bar_base + 128u + 8u * g; };
    auto slot_free = [&](int g) { return bar_base + 144u + 8u * g; };
    const uint32_t tmem_slot = bar_base + 160u;
    uint8_t* gen = smem_raw - smem_u32(smem_raw);
    uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(gen + tmem_slot);
    float* sMax = reinterpret_cast<float*>(gen + xchg_base);
    float* sSum = sMax + 2 * 2 * 128;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nb = (T + ATL_BN - 1) / ATL_BN;                    // key blocks per head

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmQ);
        prefetch_tmap(&tmKV);
        for (int i = 0; i < 2; ++i) {
            mbar_init(q_full(i), 1);
            mbar_init(q_empty(i), 1);
            mbar_init(s_ready(i), 1);
            mbar_init(p_ready(i), 8);
            mbar_init(o_ready(i), 1);
            mbar_init(slot_free(i), 8);
        }
        for (int i = 0; i < ATL_STAGES; ++i) {
            mbar_init(kv_full(i), 1);
            mbar_init(kv_empty(i), 1);
        }
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc<1>(tmem_slot, ATC_TMEM_COLS);
        tmem_relinquish<1>();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot_ptr;
    pdl_wait();       // the prologue above overlapped the previous kernel; global memory is touched from here on

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int it = 0;
            uint32_t kvn = 0;
            for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++it) {
                const int w = item / rounds, r = item % rounds;
                const int b = w / H, h = w % H;
                const int qs = it & 1;
                mbar_wait(q_empty(qs), ((it >> 1) & 1) ^ 1u);
                mbar_arrive_expect_tx(q_full(qs), Q_BYTES);
                tma_load_4d(smem_base + qs * Q_BYTES, &tmQ, q_full(qs), 0, r * 2 * ATC_BM, h, b);
                for (int j = 0; j < nb; ++j, ++kvn) {
                    const int st = kvn % ATL_STAGES;
                    mbar_wait(kv_empty(st), ((kvn / ATL_STAGES) & 1) ^ 1u);
                    const uint32_t sK = kv_base + st * 2 * KV_BYTES;
                    mbar_arrive_expect_tx(kv_full(st), 2 * KV_BYTES);
                    tma_load_4d(sK, &tmKV, kv_full(st), 0, j * ATL_BN, H + h, b);
                    tma_load_4d(sK + KV_BYTES, &tmKV, kv_full(st), 0, j * ATL_BN, 2 * H + h, b);
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        {       // all 32 lanes run the loop (uniform operands); one elected lane issues each tcgen05 instruction
            const uint32_t idesc_o = make_idesc_bf16(ATC_BM, 64, 0, 1);
            uint32_t n_p[2] = {0, 0}, n_free[2] = {0, 0};
            int it = 0;
            uint32_t kvn = 0;
            for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++it) {
                const int r = item % rounds;
                const int nact = T - r * 2 * ATC_BM > ATC_BM ? 2 : 1;       // active query tiles of this pair
                const int qs = it & 1;
                const uint32_t sQ = smem_base + qs * Q_BYTES;
                mbar_wait(q_full(qs), (it >> 1) & 1);
                tc_fence_after();
                auto keys_in = [&](int j) { return min(ATL_BN, (T - j * ATL_BN + 15) & ~15); };
                auto issue_S = [&](int g, int j) {     // S_g = Q_g K_j^T
                    const uint32_t sK = kv_base + ((kvn + j) % ATL_STAGES) * 2 * KV_BYTES;
                    const uint32_t idesc_s = make_idesc_bf16(ATC_BM, keys_in(j), 0, 0);
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_bf16_ss_elect(tmem + g * 256, make_smem_desc_sw128(sQ + g * (ATC_BM * 128) + k * 32, 1024, 0),
                                     make_smem_desc_sw128(sK + k * 32, 1024, 0), idesc_s, k > 0 ? 1u : 0u);
                    umma_commit_elect(s_ready(g));
                };
                mbar_wait(kv_full(kvn % ATL_STAGES), (kvn / ATL_STAGES) & 1);
                tc_fence_after();
                for (int g = 0; g < nact; ++g) {
                    mbar_wait(slot_free(g), (n_free[g] & 1u) ^ 1u);
                    ++n_free[g];
                    tc_fence_after();
                    issue_S(g, 0);
                }
                for (int j = 0; j < nb; ++j) {
                    const uint32_t sV = kv_base + ((kvn + j) % ATL_STAGES) * 2 * KV_BYTES + KV_BYTES;
                    const int ksteps = keys_in(j) / 16;
                    for (int g = 0; g < nact; ++g) {
                        mbar_wait(p_ready(g), n_p[g] & 1u);
                        ++n_p[g];
                        tc_fence_after();
                        // O_g (+)= P V_j : A = P in TMEM (8 packed columns per 16 keys), B = V_j [keys x 64] MN-major
                        for (int k = 0; k < ksteps; ++k)
                            umma_bf16_ts_elect(tmem + g * 256 + ATC_O_COL, tmem + g * 256 + (k < 4 ? 8 * k : 64 + 8 * (k - 4)),
                                         make_smem_desc_sw128(sV + k * 2048, 1024, 0), idesc_o, (j > 0 || k > 0) ? 1u : 0u);
                        umma_commit_elect(o_ready(g));
                        if (g == nact - 1) umma_commit_elect(kv_empty((kvn + j) % ATL_STAGES));   // both tiles are done with block j
                        if (j + 1 < nb) {
                            if (g == 0) {
                                mbar_wait(kv_full((kvn + j + 1) % ATL_STAGES), ((kvn + j + 1) / ATL_STAGES) & 1);
                                tc_fence_after();
                            }
                            issue_S(g, j + 1);
                        }
                    }
                }
                umma_commit_elect(q_empty(qs));
                kvn += nb;
            }
        }
        __syncwarp();
    } else {
        // ===================== softmax + epilogue groups =====================
        const int sw = warp - 2;
        const int g = sw >> 3;
        const int half = (sw >> 2) & 1;
        const int quarter = warp & 3;
        const uint32_t tbase = tmem + g * 256 + ((uint32_t)(quarter * 32) << 16);
        const int r_in_tile = quarter * 32 + lane;
        float* myMax = sMax + (g * 2 + half) * 128, *otherMax = sMax + (g * 2 + (half ^ 1)) * 128;
        float* mySum = sSum + (g * 2 + half) * 128, *otherSum = sSum + (g * 2 + (half ^ 1)) * 128;
        uint32_t n_s = 0, n_o = 0;
        for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
            const int w = item / rounds, r = item % rounds;
            const int b = w / H, h = w % H;
            const int row0 = (r * 2 + g) * ATC_BM;
            if (row0 >= T) continue;                      // this pair has a single tile
            const int r_in_head = row0 + r_in_tile;
            if (row0 + quarter * 32 >= T) {
                // all 32 rows of this warp lie beyond T: no TMEM reads, no math -- only the barrier protocol (TMEM read
                // bandwidth and the schedulers are what bound this kernel). The P rows stay garbage, which only reaches
                // O rows that are never stored. Kept as a separate loop so the hot path's register allocation is untouched.
                for (int j = 0; j < nb; ++j) {
                    mbar_wait(s_ready(g), n_s & 1u);
                    ++n_s;
                    myMax[r_in_tile] = -INFINITY;
                    named_bar_sync(1 + g, 256);
                    if (j > 0) {
                        mbar_wait(o_ready(g), n_o & 1u);
                        ++n_o;
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(p_ready(g));
                }
                mySum[r_in_tile] = 0.f;
                named_bar_sync(1 + g, 256);
                mbar_wait(o_ready(g), n_o & 1u);
                ++n_o;
                __syncwarp();
                if (lane == 0) mbar_arrive(slot_free(g));
                continue;
            }
            float m = -INFINITY, l = 0.f;
            for (int j = 0; j < nb; ++j) {
                const int nvalid = min(ATL_BN, T - j * ATL_BN);
                const int c_begin = half * 64, c_end = min(c_begin + 64, (nvalid + 15) & ~15);
                mbar_wait(s_ready(g), n_s & 1u);
                ++n_s;
                tc_fence_after();
                float mx = -INFINITY;
                for (int c = c_begin; c < c_end; c += 16) {
                    uint32_t v[16];
                    tmem_ld_32x16(tbase + c, v);
                    tmem_ld_wait();
                    if (c + 16 <= nvalid) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) mx = fmaxf(mx, __uint_as_float(v[i]));
                    } else {
#pragma unroll
                        for (int i = 0; i < 16; ++i) if (c + i < nvalid) mx = fmaxf(mx, __uint_as_float(v[i]));
                    }
                }
                myMax[r_in_tile] = mx;
                named_bar_sync(1 + g, 256);
                const float m_new = fmaxf(m, fmaxf(mx, otherMax[r_in_tile]));
                const float alpha = fast_ex2((m - m_new) * scale_log2);
                const float mneg = -m_new * scale_log2;
                float sum = 0.f;
                for (int c = c_begin; c < c_end; c += 16) {
                    uint32_t v[16];
                    tmem_ld_32x16(tbase + c, v);
                    tmem_ld_wait();
                    float p[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) p[i] = fast_ex2(fmaf(__uint_as_float(v[i]), scale_log2, mneg));
                    if (c + 16 > nvalid) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) if (c + i >= nvalid) p[i] = 0.f;
                    }
                    uint32_t pk[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        pk[i] = pack_bf16(p[2 * i], p[2 * i + 1]);
                        sum += p[2 * i] + p[2 * i + 1];
                    }
                    tmem_st_32x8(tbase + c_begin + ((c - c_begin) >> 1), pk);
                }
                l = fmaf(l, alpha, sum);
                m = m_new;
                if (j > 0) {
                    mbar_wait(o_ready(g), n_o & 1u);          // PV of the previous block has retired: O may be rescaled
                    ++n_o;
                    tc_fence_after();
                    if (__any_sync(0xffffffffu, alpha < 1.f)) {
                        uint32_t o[32];
                        tmem_ld_32x32(tbase + ATC_O_COL + half * 32, o);
                        tmem_ld_wait();
#pragma unroll
                        for (int q8 = 0; q8 < 4; ++q8) {
                            uint32_t t8[8];
#pragma unroll
                            for (int i = 0; i < 8; ++i) t8[i] = __float_as_uint(__uint_as_float(o[8 * q8 + i]) * alpha);
                            tmem_st_32x8(tbase + ATC_O_COL + half * 32 + 8 * q8, t8);
                        }
                    }
                }
                tmem_st_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(p_ready(g));
            }
            mySum[r_in_tile] = l;
            named_bar_sync(1 + g, 256);
            l += otherSum[r_in_tile];
            const float inv = 1.0f / l;
            const bool row_ok = r_in_head < T;
            if (row_ok && half == 0) lse[(long long)w * T + r_in_head] = m * scale + logf(l);
            mbar_wait(o_ready(g), n_o & 1u);
            ++n_o;
            tc_fence_after();
            uint32_t o[32];
            tmem_ld_32x32(tbase + ATC_O_COL + half * 32, o);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(slot_free(g));
            if (row_ok) {
                bf16* dst = out + (((long long)b * T + r_in_head) * H + h) * 64 + half * 32;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    uint4 q;
                    q.x = pack_bf16(__uint_as_float(o[8 * i]) * inv, __uint_as_float(o[8 * i + 1]) * inv);
                    q.y = pack_bf16(__uint_as_float(o[8 * i + 2]) * inv, __uint_as_float(o[8 * i + 3]) * inv);
                    q.z = pack_bf16(__uint_as_float(o[8 * i + 4]) * inv, __uint_as_float(o[8 * i + 5]) * inv);
                    q.w = pack_bf16(__uint_as_float(o[8 * i + 6]) * inv, __uint_as_float(o[8 * i + 7]) * inv);
                    reinterpret_cast<uint4*>(dst)[i] = q;
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<1>(tmem, ATC_TMEM_COLS);
    }
}

// =============================================================================================
// Backward for long sequences (T > 256). Same two item types as the short kernels -- "KV" items (rows = 128 keys:
// S^T = K_t Q^T, dP^T = V_t dO^T -> dV_t = P^T dO, dK_t = dS^T Q) and "Q" items (rows = 128 queries: S = Q_t K^T,
// dP = dO_t V^T -> dQ_t = dS K) -- but the other dimension is streamed in chunks of 64 columns through a 4-deep
// TMA ring, and the gradients accumulate across the chunks in TMEM. A work item is (batch, head, type, pair of
// row tiles); the pair shares the chunk stream, tile g of the pair owns TMEM slot g:
//     [0, 64) acc0 chunk (packed P in place) | [64, 128) acc1 chunk (packed dS in place) | [128, 192) out0 (dV) |
//     [192, 256) out1 (dK or dQ)
// and math group g (8 warps: thread <-> row, two warps per row splitting the chunk's columns; each column half packs
// into the front of its own range). The MMA thread issues the gradient MMAs of chunk c and the S/dP MMAs of chunk
// c+1 back to back, so one group's math overlaps the other group's tensor work. delta = rowsum(dO * O) comes from
// a small pre-pass kernel ([B, H, T] fp32, like lse); the producer warp stages each chunk's column statistics.
// =============================================================================================
constexpr int ATLB_W = 64;           // columns per chunk
constexpr int ATLB_STAGES = 4;

__global__ void __launch_bounds__(ATC_THREADS, 1)
attention_bwd_long_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmXdO,
                          const __grid_constant__ CUtensorMap tmY, const __grid_constant__ CUtensorMap tmYdO,
                          const float* __restrict__ lse, const float* __restrict__ delta, bf16* __restrict__ dqkv,
                          int num_items, int rounds, int T, int H, float scale_log2, float scale) {
    extern __shared__ uint8_t smem_raw[];
    pdl_launch_dependents();
    const int nc = (T + ATLB_W - 1) / ATLB_W;                    // column chunks per item
    const uint32_t Tp = (uint32_t)nc * ATLB_W;
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    constexpr uint32_t X_BYTES = 2 * ATC_BM * 128;               // one operand, both row tiles of the pair
    constexpr uint32_t Y_BYTES = ATLB_W * 128;                   // one operand chunk
    const uint32_t ch_base = smem_base + 2 * 2 * X_BYTES;
    const uint32_t stat_base = ch_base + ATLB_STAGES * 2 * Y_BYTES;     // per X stage: lse*log2e [Tp], delta*scale [Tp]
    const uint32_t bar_base = stat_base + 2u * 2u * Tp * 4u;
    auto x_full = [&](int s) { return bar_base + 8u * s; };
    auto x_empty = [&](int s) { return bar_base + 16u + 8u * s; };
    auto ch_full = [&](int s) { return bar_base + 32u + 8u * s; };
    auto ch_empty = [&](int s) { return bar_base + 64u + 8u * s; };
    auto s_ready = [&](int g) { return bar_base + 96u + 8u * g; };
    auto p_ready = [&](int g) { return bar_base + 112u + 8u * g; };
    auto o_ready = [&](int g) { return bar_base + 128u + 8u * g; };
    auto slot_free = [&](int g) { return bar_base + 144u + 8u * g; };
    const uint32_t tmem_slot = bar_base + 160u;
    uint8_t* gen = smem_raw - smem_u32(smem_raw);
    uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(gen + tmem_slot);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int items_per_head = 2 * rounds;
    const float LOG2E = 1.4426950408889634f;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmX);
        prefetch_tmap(&tmXdO);
        prefetch_tmap(&tmY);
        prefetch_tmap(&tmYdO);
        for (int i = 0; i < 2; ++i) {
            mbar_init(x_full(i), 2);          // TMA transaction + the head's statistics
            mbar_init(x_empty(i), 1);
            mbar_init(s_ready(i), 1);
            mbar_init(p_ready(i), 8);
            mbar_init(o_ready(i), 1);
            mbar_init(slot_free(i), 8);
        }
        for (int i = 0; i < ATLB_STAGES; ++i) {
            mbar_init(ch_full(i), 1);
            mbar_init(ch_empty(i), 1);
        }
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc<1>(tmem_slot, 512);
        tmem_relinquish<1>();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot_ptr;
    pdl_wait();       // the prologue above overlapped the previous kernel; global memory is touched from here on

    if (warp == 0) {
        // ===================== TMA producer + column statistics =====================
        int it = 0;
        uint32_t chn = 0;
        for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++it) {
            const int w = item / items_per_head, idx = item % items_per_head;
            const bool kv_item = idx < rounds;
            const int r = kv_item ? idx : idx - rounds;
            const int b = w / H, h = w % H;
            const int xs = it & 1;
            if (lane == 0) {
                mbar_wait(x_empty(xs), ((it >> 1) & 1) ^ 1u);
                const uint32_t sX = smem_base + xs * 2 * X_BYTES;
                mbar_arrive_expect_tx(x_full(xs), 2 * X_BYTES);
                if (kv_item) {
                    tma_load_4d(sX, &tmX, x_full(xs), 0, r * 2 * ATC_BM, H + h, b);                  // K rows
                    tma_load_4d(sX + X_BYTES, &tmX, x_full(xs), 0, r * 2 * ATC_BM, 2 * H + h, b);    // V rows
                } else {
                    tma_load_4d(sX, &tmX, x_full(xs), 0, r * 2 * ATC_BM, h, b);                      // Q rows
                    tma_load_4d(sX + X_BYTES, &tmXdO, x_full(xs), 0, r * 2 * ATC_BM, h, b);          // dO rows
                }
            }
            __syncwarp();
            {   // the head's statistics: lse*log2e (+inf beyond T, so padded columns get P = 0) and delta*scale
                float* sL = reinterpret_cast<float*>(gen + stat_base + (uint32_t)xs * 2u * Tp * 4u);
                float* sD = sL + Tp;
                for (int t = lane; t < (int)Tp; t += 32) {
                    sL[t] = t < T ? __ldg(lse + (long long)w * T + t) * LOG2E : INFINITY;
                    sD[t] = t < T ? __ldg(delta + (long long)w * T + t) * scale : 0.f;
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(x_full(xs));
            }
            for (int c = 0; c < nc; ++c, ++chn) {
                const int st = chn % ATLB_STAGES;
                if (lane == 0) {
                    mbar_wait(ch_empty(st), ((chn / ATLB_STAGES) & 1) ^ 1u);
                    const uint32_t sY = ch_base + st * 2 * Y_BYTES;
                    mbar_arrive_expect_tx(ch_full(st), 2 * Y_BYTES);
                    if (kv_item) {
                        tma_load_4d(sY, &tmY, ch_full(st), 0, c * ATLB_W, h, b);                     // Q chunk
                        tma_load_4d(sY + Y_BYTES, &tmYdO, ch_full(st), 0, c * ATLB_W, h, b);         // dO chunk
                    } else {
                        tma_load_4d(sY, &tmY, ch_full(st), 0, c * ATLB_W, H + h, b);                 // K chunk
                        tma_load_4d(sY + Y_BYTES, &tmY, ch_full(st), 0, c * ATLB_W, 2 * H + h, b);   // V chunk
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        {       // all 32 lanes run the loop (uniform operands); one elected lane issues each tcgen05 instruction
            const uint32_t idesc2 = make_idesc_bf16(ATC_BM, 64, 0, 1);
            uint32_t n_p[2] = {0, 0}, n_free[2] = {0, 0};
            int it = 0;
            uint32_t chn = 0;
            for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++it) {
                const int idx = item % items_per_head;
                const bool kv_item = idx < rounds;
                const int r = kv_item ? idx : idx - rounds;
                const int nact = T - r * 2 * ATC_BM > ATC_BM ? 2 : 1;
                const int xs = it & 1;
                const uint32_t sX = smem_base + xs * 2 * X_BYTES, sX2 = sX + X_BYTES;
                mbar_wait(x_full(xs), (it >> 1) & 1);
                tc_fence_after();
                auto width = [&](int c) { return min(ATLB_W, (T - c * ATLB_W + 15) & ~15); };
                auto issue_1 = [&](int g, int c) {      // acc0 = X_g Y_c^T, acc1 = X2_g Y2_c^T (independent chains, interleaved)
                    const uint32_t sY = ch_base + ((chn + c) % ATLB_STAGES) * 2 * Y_BYTES, sY2 = sY + Y_BYTES;
                    const uint32_t idesc1 = make_idesc_bf16(ATC_BM, width(c), 0, 0);
                    const uint32_t slot = tmem + g * 256, xo = g * (ATC_BM * 128);
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        umma_bf16_ss_elect(slot, make_smem_desc_sw128(sX + xo + k * 32, 1024, 0),
                                     make_smem_desc_sw128(sY + k * 32, 1024, 0), idesc1, k > 0 ? 1u : 0u);
                        umma_bf16_ss_elect(slot + 64, make_smem_desc_sw128(sX2 + xo + k * 32, 1024, 0),
                                     make_smem_desc_sw128(sY2 + k * 32, 1024, 0), idesc1, k > 0 ? 1u : 0u);
                    }
                    umma_commit_elect(s_ready(g));
                };
                mbar_wait(ch_full(chn % ATLB_STAGES), (chn / ATLB_STAGES) & 1);
                tc_fence_after();
                for (int g = 0; g < nact; ++g) issue_1(g, 0);
                for (int c = 0; c < nc; ++c) {
                    const uint32_t sY = ch_base + ((chn + c) % ATLB_STAGES) * 2 * Y_BYTES, sY2 = sY + Y_BYTES;
                    const int ksteps = width(c) / 16;
                    for (int g = 0; g < nact; ++g) {
                        const uint32_t slot = tmem + g * 256;
                        mbar_wait(p_ready(g), n_p[g] & 1u);
                        ++n_p[g];
                        if (c == 0) {       // the previous item's gradients must have been read out of this slot
                            mbar_wait(slot_free(g), (n_free[g] & 1u) ^ 1u);
                            ++n_free[g];
                        }
                        tc_fence_after();
                        for (int k = 0; k < ksteps; ++k) {
                            const uint32_t acol = k < 2 ? 8 * k : 32 + 8 * (k - 2);
                            const uint32_t acc = (c > 0 || k > 0) ? 1u : 0u;
                            if (kv_item)                  // dV_t += P^T dO_c
                                umma_bf16_ts_elect(slot + 128, slot + acol, make_smem_desc_sw128(sY2 + k * 2048, 1024, 0), idesc2, acc);
                            // dK_t += dS^T Q_c   or   dQ_t += dS K_c
                            umma_bf16_ts_elect(slot + 192, slot + 64 + acol, make_smem_desc_sw128(sY + k * 2048, 1024, 0), idesc2, acc);
                        }
                        if (c == nc - 1) umma_commit_elect(o_ready(g));
                        if (g == nact - 1) umma_commit_elect(ch_empty((chn + c) % ATLB_STAGES));
                        if (c + 1 < nc) {
                            if (g == 0) {
                                mbar_wait(ch_full((chn + c + 1) % ATLB_STAGES), ((chn + c + 1) / ATLB_STAGES) & 1);
                                tc_fence_after();
                            }
                            issue_1(g, c + 1);
                        }
                    }
                }
                umma_commit_elect(x_empty(xs));
                chn += nc;
            }
        }
        __syncwarp();
    } else {
        // ===================== math + read-out groups =====================
        const int sw = warp - 2;
        const int g = sw >> 3;
        const int half = (sw >> 2) & 1;
        const int quarter = warp & 3;
        const uint32_t slot = tmem + g * 256 + ((uint32_t)(quarter * 32) << 16);
        const long long rs = 3LL * H * 64;
        uint32_t n_s = 0, n_o = 0;
        int it = 0;
        for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++it) {
            const int w = item / items_per_head, idx = item % items_per_head;
            const bool kv_item = idx < rounds;
            const int r = kv_item ? idx : idx - rounds;
            const int b = w / H, h = w % H;
            const int row0 = (r * 2 + g) * ATC_BM;
            if (row0 >= T) continue;                      // this pair has a single tile
            if (row0 + quarter * 32 >= T) {
                // all 32 rows of this warp lie beyond T: only the barrier protocol (see the forward kernel)
                mbar_wait(x_full(it & 1), (it >> 1) & 1);
                for (int c = 0; c < nc; ++c) {
                    mbar_wait(s_ready(g), n_s & 1u);
                    ++n_s;
                    __syncwarp();
                    if (lane == 0) mbar_arrive(p_ready(g));
                }
                mbar_wait(o_ready(g), n_o & 1u);
                ++n_o;
                __syncwarp();
                if (lane == 0) mbar_arrive(slot_free(g));
                continue;
            }
            const int row = row0 + quarter * 32 + lane;
            const bool row_ok = row < T;
                        const float* sL = reinterpret_cast<const float*>(gen + stat_base + (uint32_t)(it & 1) * 2u * Tp * 4u);
            const float* sD = sL + Tp;
            mbar_wait(x_full(it & 1), (it >> 1) & 1);                 // the head's statistics are staged
            const float row_lse2 = row_ok ? sL[row] : 0.f, row_delta = row_ok ? sD[row] : 0.f;
            const int cb = half * 32;
            for (int c = 0; c < nc; ++c) {
                const int W = min(ATLB_W, (T - c * ATLB_W + 15) & ~15);
                const int ce = min(cb + 32, W);
                mbar_wait(s_ready(g), n_s & 1u);
                ++n_s;
                tc_fence_after();
                for (int cc = cb; cc < ce; cc += 16) {
                    uint32_t sv[16], dp[16];
                    tmem_ld_32x16(slot + cc, sv);
                    tmem_ld_32x16(slot + 64 + cc, dp);
                    tmem_ld_wait();
                    float p[16], ds[16];
#pragma unroll
                    for (int i4 = 0; i4 < 4; ++i4) {
                        float4 a = make_float4(row_lse2, row_lse2, row_lse2, row_lse2);
                        float4 d = make_float4(row_delta, row_delta, row_delta, row_delta);
                        if (kv_item) {
                            a = reinterpret_cast<const float4*>(sL + c * ATLB_W + cc)[i4];
                            d = reinterpret_cast<const float4*>(sD + c * ATLB_W + cc)[i4];
                        }
                        const float l2[4] = {a.x, a.y, a.z, a.w}, dl[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const int i = 4 * i4 + k;
                            p[i] = fast_ex2(fmaf(__uint_as_float(sv[i]), scale_log2, -l2[k]));
                            ds[i] = p[i] * fmaf(__uint_as_float(dp[i]), scale, -dl[k]);
                        }
                    }
                    if (!kv_item && c * ATLB_W + cc + 16 > T) {      // Q item: padded key columns must not contribute to dQ
#pragma unroll
                        for (int i = 0; i < 16; ++i) if (c * ATLB_W + cc + i >= T) ds[i] = 0.f;
                    }
                    uint32_t pk[8];
                    if (kv_item) {
#pragma unroll
                        for (int i = 0; i < 8; ++i) pk[i] = pack_bf16(p[2 * i], p[2 * i + 1]);
                        tmem_st_32x8(slot + cb + ((cc - cb) >> 1), pk);
                    }
#pragma unroll
                    for (int i = 0; i < 8; ++i) pk[i] = pack_bf16(ds[2 * i], ds[2 * i + 1]);
                    tmem_st_32x8(slot + 64 + cb + ((cc - cb) >> 1), pk);
                }
                tmem_st_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(p_ready(g));
            }
            // read-out: this warp converts 32 of the 64 gradient columns of its rows
            mbar_wait(o_ready(g), n_o & 1u);
            ++n_o;
            tc_fence_after();
            uint32_t r1[32], r0[32];
            tmem_ld_32x32(slot + 192 + half * 32, r1);
            if (kv_item) tmem_ld_32x32(slot + 128 + half * 32, r0);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(slot_free(g));
            if (row_ok) {
                bf16* base = dqkv + ((long long)b * T + row) * rs + (long long)h * 64 + half * 32;
                auto emit = [&](const uint32_t (&v)[32], bf16* dst) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        uint4 q;
                        q.x = pack_bf16(__uint_as_float(v[8 * i]), __uint_as_float(v[8 * i + 1]));
                        q.y = pack_bf16(__uint_as_float(v[8 * i + 2]), __uint_as_float(v[8 * i + 3]));
                        q.z = pack_bf16(__uint_as_float(v[8 * i + 4]), __uint_as_float(v[8 * i + 5]));
                        q.w = pack_bf16(__uint_as_float(v[8 * i + 6]), __uint_as_float(v[8 * i + 7]));
                        reinterpret_cast<uint4*>(dst)[i] = q;
                    }
                };
                if (kv_item) {
                    emit(r1, base + (long long)H * 64);        // dK
                    emit(r0, base + 2LL * H * 64);             // dV
                } else {
                    emit(r1, base);                            // dQ
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<1>(tmem, 512);
    }
}

// ---------------------------------------------------------------------------------------------
static PFN_cuTensorMapEncodeTiled atc_encode_fn() {
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

// 4-D view of a packed [B, T, S, 64] bf16 tensor (S slabs of 64 per token: 3*H for qkv, H for out/dout):
// dims (d, t, slab, b); a box is [rows x 64] of one slab; rows beyond T are zero-filled by TMA.
int make_tmap_rows64(CUtensorMap* tm, const void* base, int B, int T, int slabs, int box_rows) {
    PFN_cuTensorMapEncodeTiled enc = atc_encode_fn();
    if (!enc) return set_error(-2, "cuTensorMapEncodeTiled not available (no CUDA driver?)");
    cuuint64_t dims[4] = {64, (cuuint64_t)T, (cuuint64_t)slabs, (cuuint64_t)B};
    cuuint64_t strides[3] = {(cuuint64_t)slabs * 64 * 2, 64 * 2, (cuuint64_t)T * slabs * 64 * 2};
    cuuint32_t box[4] = {64, (cuuint32_t)box_rows, 1, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(-3, "cuTensorMapEncodeTiled (attention) failed (%d)", (int)r);
    return 0;
}

int attention_fwd_tc_launch(const bf16* qkv, bf16* out, float* lse, int batch, int tokens, int heads, float scale,
                            cudaStream_t stream) {
    const int Nk = (tokens + 15) & ~15;
    const int nt = (tokens + ATC_BM - 1) / ATC_BM;
    CUtensorMap tmQ, tmKV;
    int rc = make_tmap_rows64(&tmQ, qkv, batch, tokens, 3 * heads, nt * ATC_BM);
    if (rc) return rc;
    rc = make_tmap_rows64(&tmKV, qkv, batch, tokens, 3 * heads, Nk);
    if (rc) return rc;
    const size_t kv_stride = ((size_t)Nk * 128 + 1023) & ~(size_t)1023;
    const size_t smem = 1024 + 2 * (ATC_Q_ROWS * 128 + 2 * kv_stride) + 2 * 2 * 2 * 128 * 4 + 128;
    B200_CUDA(cudaFuncSetAttribute(attention_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int total = batch * heads;
    const int grid = total < sm_count() ? total : sm_count();
    B200_CUDA(launch_kernel_pdl(attention_fwd_tc_kernel, dim3(grid), dim3(ATC_THREADS), smem, stream, tmQ, tmKV, out, lse, total,
                                tokens, Nk, heads, scale * 1.4426950408889634f, scale));
    B200_LAUNCH_CHECK();
    return 0;
}

int attention_fwd_long_launch(const bf16* qkv, bf16* out, float* lse, int batch, int tokens, int heads, float scale,
                              cudaStream_t stream) {
    const int nt = (tokens + ATC_BM - 1) / ATC_BM;
    const int rounds = (nt + 1) / 2;
    CUtensorMap tmQ, tmKV;
    int rc = make_tmap_rows64(&tmQ, qkv, batch, tokens, 3 * heads, 2 * ATC_BM);
    if (rc) return rc;
    rc = make_tmap_rows64(&tmKV, qkv, batch, tokens, 3 * heads, ATL_BN);
    if (rc) return rc;
    const size_t smem = 1024 + 2 * (2 * ATC_BM * 128) + (size_t)ATL_STAGES * 2 * ATL_BN * 128 + 2 * 2 * 2 * 128 * 4 + 256;
    B200_CUDA(cudaFuncSetAttribute(attention_fwd_long_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long items = (long long)batch * heads * rounds;
    B200_CHECK_ARG(items < (1LL << 31), "attention_fwd: too many work items");
    const int grid = items < sm_count() ? (int)items : sm_count();
    B200_CUDA(launch_kernel_pdl(attention_fwd_long_kernel, dim3(grid), dim3(ATC_THREADS), smem, stream, tmQ, tmKV, out, lse,
                                (int)items, rounds, tokens, heads, scale * 1.4426950408889634f, scale));
    B200_LAUNCH_CHECK();
    return 0;
}

int attention_bwd_long_launch(const bf16* qkv, const bf16* dout, const float* lse, const float* delta, bf16* dqkv,
                              int batch, int tokens, int heads, float scale, cudaStream_t stream) {
    const int nt = (tokens + ATC_BM - 1) / ATC_BM;
    const int rounds = (nt + 1) / 2;
    const int Tp = (tokens + ATLB_W - 1) / ATLB_W * ATLB_W;
    CUtensorMap tx, txd, ty, tyd;
    int rc = make_tmap_rows64(&tx, qkv, batch, tokens, 3 * heads, 2 * ATC_BM);
    if (rc) return rc;
    rc = make_tmap_rows64(&txd, dout, batch, tokens, heads, 2 * ATC_BM);
    if (rc) return rc;
    rc = make_tmap_rows64(&ty, qkv, batch, tokens, 3 * heads, ATLB_W);
    if (rc) return rc;
    rc = make_tmap_rows64(&tyd, dout, batch, tokens, heads, ATLB_W);
    if (rc) return rc;
    const size_t smem = 1024 + 2 * 2 * (size_t)(2 * ATC_BM * 128) + (size_t)ATLB_STAGES * 2 * ATLB_W * 128 +
                        2 * 2 * (size_t)Tp * 4 + 256;
    B200_CHECK_ARG(smem <= 232448, "attention_bwd: %d tokens exceed the staged-statistics limit (%d)", tokens, AT_LONG_MAX_T);
    B200_CUDA(cudaFuncSetAttribute(attention_bwd_long_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long items = (long long)batch * heads * 2 * rounds;
    B200_CHECK_ARG(items < (1LL << 31), "attention_bwd: too many work items");
    const int grid = items < sm_count() ? (int)items : sm_count();
    B200_CUDA(launch_kernel_pdl(attention_bwd_long_kernel, dim3(grid), dim3(ATC_THREADS), smem, stream, tx, txd, ty, tyd, lse,
                                delta, dqkv, (int)items, rounds, tokens, heads, scale * 1.4426950408889634f, scale));
    B200_LAUNCH_CHECK();
    return 0;
}

int attention_bwd_tc_launch(const bf16* qkv, const bf16* out, const bf16* dout, const float* lse, bf16* dqkv, float* dbias,
                            int batch, int tokens, int heads, float scale, cudaStream_t stream) {
    const int Nk = (tokens + 15) & ~15;
    CUtensorMap tq, td;
    int rc = make_tmap_rows64(&tq, qkv, batch, tokens, 3 * heads, Nk);
    if (rc) return rc;
    rc = make_tmap_rows64(&td, dout, batch, tokens, heads, Nk);
    if (rc) return rc;
    const size_t tile_stride = ((size_t)Nk * 128 + 1023) & ~(size_t)1023;
    const int total = batch * heads;
    const int grid = total < sm_count() ? total : sm_count();
    const size_t smem_base = 1024 + 4 * tile_stride + ATB_TAIL_PAD + 2 * 32768 + 2 * 2 * (size_t)Nk * 4 + 256;
    B200_CHECK_ARG(smem_base <= 232448, "attention_bwd: %d tokens need %zu bytes of shared memory", tokens, smem_base);
    // The in_proj_bias gradient is accumulated in shared memory ([3][heads][64] fp32) when that fits beside the tiles;
    // with very many heads it comes from a column-sum pass over dqkv afterwards.
    const size_t smem_bias = (size_t)3 * heads * 64 * 4;
    static int variant = -1;       // B200VIT_ATTN_BWD=3 forces the kernel that keeps P^T / dS^T in tensor memory (A/B runs)
    if (variant < 0) {
        const char* e = getenv("B200VIT_ATTN_BWD");
        variant = e ? atoi(e) : 0;
    }
    // tc4 (P^T / dS^T through shared memory) needs two more 16 KB tiles: up to 208 tokens with the fused bias gradient,
    // 224 without; longer sequences (<= 256) run tc3
    const size_t smem4_base = smem_base + 2 * 16384;
    const bool use4 = variant != 3 && smem4_base <= 232448;
    const size_t sb = use4 ? smem4_base : smem_base;
    const bool fused_bias = dbias != nullptr && sb + smem_bias <= 232448;
    const size_t smem3 = sb + (fused_bias ? smem_bias : 0);
    if (fused_bias) B200_CUDA(cudaMemsetAsync(dbias, 0, sizeof(float) * 3 * heads * 64, stream));
    auto kern = use4 ? attention_bwd_tc4_kernel : attention_bwd_tc3_kernel;
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem3));
    B200_CUDA(launch_kernel_pdl(kern, dim3(grid), dim3(AT3_THREADS), smem3, stream, tq, td, out, dout, lse,
                                dqkv, fused_bias ? dbias : (float*)nullptr, total, tokens, Nk, heads,
                                scale * 1.4426950408889634f, scale));
    B200_LAUNCH_CHECK();
    if (dbias != nullptr && !fused_bias)
        return colsum_bf16_launch(dqkv, 3LL * heads * 64, dbias, (long long)batch * tokens, 3 * heads * 64, stream);
    return 0;
}

}  // namespace b200vit
