// Fused causal attention (flash-style) on tcgen05: forward and backward.
//
// For BASELINE.json config 4 (GPT-2 small, seq 1024, head dim 64).  ref: none in the
// reference; its GPT-2 baseline reaches a library flash-attention kernel through torch SDPA.
// The batched-GEMM attention in ops/functional.py materialises the [T, T] score matrix in HBM
// (it is DRAM-bound on reading it back); this kernel keeps scores on chip:
//
//   one CTA per (sequence, head, 128-query block); for each 128-key block up to the diagonal
//     S = Q K^T          tcgen05.mma 128x128x64 from TMA-loaded, 128B-swizzled tiles -> TMEM
//     softmax            4 warps, one query row per thread, tcgen05.ld S, online max/sum in
//                        the log2 domain, P written as bf16 into swizzled shared memory
//     O_j = P V          tcgen05.mma 128x64x128 (V read MN-major: no transpose) -> TMEM
//     o += O_j           rescaled accumulation in registers
//   K/V tiles are double-buffered; the next S = Q K^T is issued while the softmax warps fold
//   O_j.  Outputs: O [B, T, H*D] bf16 and the log-sum-exp per row (fp32) for the backward.
//
// Q, K, V are read in place from the packed qkv activation [B, T, 3, H, D] through one 4-D
// tensor map (no head split copies).
#include "common.cuh"

#include <cstdlib>
#include <type_traits>

void b200_count_launch(int n);

namespace b200 {

constexpr int FA_BLOCK = 128;   // queries per CTA and keys per step
constexpr int FA_D = 64;        // head dimension
constexpr int FA_THREADS = 160; // warps 0-3 softmax (one row per thread), warp 4 TMA + MMA
constexpr uint32_t FA_TILE_BYTES = FA_BLOCK * FA_D * 2;  // 16 KB
// 7 tiles + barriers = 114,944 B: with the 1 KB per-block reserve, two CTAs fit in 227 KB.
constexpr uint32_t FA_SMEM = FA_TILE_BYTES * (1 + 2 + 2) + 2 * FA_TILE_BYTES /*P*/ + 256;

struct FaParams {
    int B, T, H;
    __nv_bfloat16* out;  // [B, T, H*D]
    float* lse;          // [B, H, T]
    float scale_log2e;   // (1/sqrt(D)) * log2(e)
    int trace;
};

__device__ __forceinline__ float fa_exp2(float x) {   // MUFU ex2; inputs are <= 0 here
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// Tried and dropped (B200, ncu + in-kernel step traces, profiles/flash_attention_r2.md): computing a
// share of the exponentials with a cubic on the FMA pipe (the FA4 recipe) and packing bf16 on the ALU
// pipe instead of F2FP.  Both made the kernels slower: a polynomial exp2 costs 6-7 FMA-pipe
// instructions at 2 cycles each against 8 cycles of MUFU.EX2 per warp, so the softmax is fastest with
// every exponential on MUFU and sits near that pipe's rate (16 per clock per SM).
__device__ __forceinline__ float fa_max3(float a, float b, float c) {
    float r;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}
__device__ __forceinline__ uint32_t fa_idesc(int n, bool b_mn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(b_mn ? 1 : 0) << 16) | ((uint32_t)(n >> 3) << 17) |
           ((uint32_t)(128 >> 4) << 24);
}

// 113 KB of shared memory and 256 TMEM columns per CTA: two CTAs share an SM, so one CTA's
// softmax overlaps the other's MMAs and loads.
__global__ void __launch_bounds__(FA_THREADS, 2)
flash_attn_fwd_kernel(const __grid_constant__ CUtensorMap tma_qkv, const FaParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];   // swizzled tiles need 1 KB alignment
    if (threadIdx.x == 0 && (smem_u32(smem) & 1023u) != 0) __trap();
    uint8_t* s_q = smem;
    uint8_t* s_k = s_q + FA_TILE_BYTES;            // 2 stages
    uint8_t* s_v = s_k + 2 * FA_TILE_BYTES;        // 2 stages
    uint8_t* s_p = s_v + 2 * FA_TILE_BYTES;        // 32 KB: two [128 x 64] swizzled blocks
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_p + 2 * FA_TILE_BYTES);
    uint64_t* q_full = bars;          // 1
    uint64_t* k_full = bars + 1;      // 2
    uint64_t* v_full = bars + 3;      // 2
    uint64_t* kv_empty = bars + 5;    // 2
    uint64_t* s_full = bars + 7;      // 1
    uint64_t* p_full = bars + 8;      // 1 (128 arrivals)
    uint64_t* o_full = bars + 9;      // 1
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nq = p.T / FA_BLOCK;
    const int qi = nq - 1 - (int)blockIdx.x;   // heaviest query blocks first
    const int h = blockIdx.y, b = blockIdx.z;
    const int nblocks = qi + 1;
    const int q0 = qi * FA_BLOCK;

    if (warp == 4 && lane == 0) {
        tma_prefetch_desc(&tma_qkv);
        mbar_init(q_full, 1);
        for (int s = 0; s < 2; ++s) { mbar_init(&k_full[s], 1); mbar_init(&v_full[s], 1); mbar_init(&kv_empty[s], 1); }
        mbar_init(s_full, 1);
        mbar_init(p_full, FA_BLOCK);
        mbar_init(o_full, 1);
        fence_mbar_init();
    }
    if (warp == 4) {
        tmem_alloc(tmem_slot, 256);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tmem_s = tmem_base;          // 128 columns
    const uint32_t tmem_o = tmem_base + 128;    // 64 columns
    asm volatile("griddepcontrol.wait;" ::: "memory");

    if (warp == 4) {
        // The whole warp walks the loop; one elected lane issues TMA / MMA / commits.  Addresses and
        // descriptors stay warp-uniform (uniform registers), so each UTCHMMA is a single instruction
        // instead of an elect/broadcast loop -- the MMA issue sits on the critical path between the
        // softmax phases.  Descriptors differ only in the start-address field (bytes >> 4).
        auto load_kv = [&](int j) {
            const int st = j & 1;
            mbar_arrive_expect_tx(&k_full[st], FA_TILE_BYTES);
            tma_load_4d(s_k + st * FA_TILE_BYTES, &tma_qkv, &k_full[st], 0, 1 * p.H + h, j * FA_BLOCK, b);
            mbar_arrive_expect_tx(&v_full[st], FA_TILE_BYTES);
            tma_load_4d(s_v + st * FA_TILE_BYTES, &tma_qkv, &v_full[st], 0, 2 * p.H + h, j * FA_BLOCK, b);
        };
        if (elect_one()) {
            mbar_arrive_expect_tx(q_full, FA_TILE_BYTES);
            tma_load_4d(s_q, &tma_qkv, q_full, 0, 0 * p.H + h, q0, b);
            load_kv(0);
            if (nblocks > 1) load_kv(1);
        }
        __syncwarp();
        const uint32_t idesc_s = fa_idesc(128, false), idesc_o = fa_idesc(64, true);
        const uint64_t dq_k = umma_smem_desc(smem_u32(s_q), 16, 1024);       // Q, K-major
        const uint64_t dk_k0 = umma_smem_desc(smem_u32(s_k), 16, 1024);      // K stage 0, K-major
        const uint64_t dp_k = umma_smem_desc(smem_u32(s_p), 16, 1024);       // P, K-major
        const uint64_t dv_mn0 = umma_smem_desc(smem_u32(s_v), 8192, 1024);   // V stage 0, MN-major
        mbar_wait(q_full, 0);
        for (int j = 0; j < nblocks; ++j) {
            const int st = j & 1;
            const uint32_t kv_par = (uint32_t)((j >> 1) & 1);
            const uint64_t dk_k = dk_k0 + (uint64_t)(st * (FA_TILE_BYTES >> 4));
            const uint64_t dv_mn = dv_mn0 + (uint64_t)(st * (FA_TILE_BYTES >> 4));
            // ---- S = Q K^T
            mbar_wait(&k_full[st], kv_par);
            tc_fence_after();
            if (elect_one()) {
#pragma unroll
                for (int k = 0; k < FA_D / 16; ++k)
                    umma_bf16(tmem_s, dq_k + (uint64_t)(2 * k), dk_k + (uint64_t)(2 * k), idesc_s, k != 0 ? 1u : 0u);
                umma_commit(s_full);
            }
            __syncwarp();
            // ---- O_j = P V once the softmax warps have written P (and left S)
            mbar_wait(p_full, (uint32_t)(j & 1));
            tc_fence_after();
            mbar_wait(&v_full[st], kv_par);
            tc_fence_after();
            if (elect_one()) {
#pragma unroll
                for (int kk = 0; kk < FA_BLOCK / 16; ++kk)
                    umma_bf16(tmem_o, dp_k + (uint64_t)((kk >> 2) * (FA_TILE_BYTES >> 4) + (kk & 3) * 2),
                              dv_mn + (uint64_t)(kk * 128), idesc_o, kk != 0 ? 1u : 0u);
                umma_commit(o_full);
                umma_commit(&kv_empty[st]);
            }
            __syncwarp();
            if (j + 2 < nblocks) {  // refill this K/V stage for block j+2
                mbar_wait(&kv_empty[st], kv_par);
                if (elect_one()) load_kv(j + 2);
                __syncwarp();
            }
        }
    } else {
        // ===================== softmax warps: thread r owns query row q0 + r =====================
        const int r = threadIdx.x;
        const uint32_t lane_addr = ((uint32_t)(warp * 32)) << 16;
        float m = -INFINITY, l = 0.f;
        float o[FA_D];
#pragma unroll
        for (int d = 0; d < FA_D; ++d) o[d] = 0.f;
        for (int j = 0; j < nblocks; ++j) {
            const bool diag = (j == qi);
            mbar_wait(s_full, (uint32_t)(j & 1));
            tc_fence_after();
            // pass A: row maximum (scores are re-read from TMEM in pass B; tcgen05.ld is cheap)
            float mx = -INFINITY;
#pragma unroll 1
            for (int c = 0; c < FA_BLOCK / 32; ++c) {
                uint32_t v[32];
                tmem_ld_32x32(tmem_s + lane_addr + c * 32, v);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const float s = __uint_as_float(v[i]);
                    if (!diag || c * 32 + i <= r) mx = fmaxf(mx, s);
                }
            }
            const float m_new = fmaxf(m, mx * p.scale_log2e);
            const float alpha = fa_exp2(m - m_new);   // 0 on the first block (m = -inf)
            l *= alpha;
#pragma unroll
            for (int d = 0; d < FA_D; ++d) o[d] *= alpha;
            m = m_new;
            // pass B: P = exp2(s * scale - m), bf16 into swizzled smem ([128 x 64] blocks, 128 B rows)
            float psum = 0.f;
#pragma unroll 1
            for (int c = 0; c < FA_BLOCK / 32; ++c) {
                uint32_t v[32];
                tmem_ld_32x32(tmem_s + lane_addr + c * 32, v);
                tmem_ld_wait();
                float pv[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const float e = fa_exp2(__uint_as_float(v[i]) * p.scale_log2e - m);
                    pv[i] = (!diag || c * 32 + i <= r) ? e : 0.f;
                    psum += pv[i];
                }
                uint8_t* blk = s_p + (c >> 1) * FA_TILE_BYTES + r * 128;
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    const int vec = (c & 1) * 4 + t;   // 16-byte chunk index inside the 128-byte row
                    uint4 w;
                    w.x = pack_bf16(pv[8 * t + 0], pv[8 * t + 1]);
                    w.y = pack_bf16(pv[8 * t + 2], pv[8 * t + 3]);
                    w.z = pack_bf16(pv[8 * t + 4], pv[8 * t + 5]);
                    w.w = pack_bf16(pv[8 * t + 6], pv[8 * t + 7]);
                    *reinterpret_cast<uint4*>(blk + ((vec ^ (r & 7)) << 4)) = w;
                }
            }
            l += psum;
            fence_proxy_async();   // generic-proxy smem writes -> visible to the tensor core
            tc_fence_before();
            mbar_arrive(p_full);
            // fold O_j
            mbar_wait(o_full, (uint32_t)(j & 1));
            tc_fence_after();
#pragma unroll
            for (int c = 0; c < FA_D / 32; ++c) {
                uint32_t v[32];
                tmem_ld_32x32(tmem_o + lane_addr + c * 32, v);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) o[c * 32 + i] += __uint_as_float(v[i]);
            }
            tc_fence_before();
        }
        const float inv = 1.f / l;
        __nv_bfloat16* dst = p.out + ((int64_t)b * p.T + q0 + r) * ((int64_t)p.H * FA_D) + (int64_t)h * FA_D;
#pragma unroll
        for (int d = 0; d < FA_D; d += 8) {
            uint4 w;
            w.x = pack_bf16(o[d] * inv, o[d + 1] * inv);
            w.y = pack_bf16(o[d + 2] * inv, o[d + 3] * inv);
            w.z = pack_bf16(o[d + 4] * inv, o[d + 5] * inv);
            w.w = pack_bf16(o[d + 6] * inv, o[d + 7] * inv);
            *reinterpret_cast<uint4*>(dst + d) = w;
        }
        // natural-log LSE of the scaled scores
        p.lse[((int64_t)b * p.H + h) * p.T + q0 + r] = (m + log2f(l)) * 0.6931471805599453f;
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 4) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 256);
    }
}

// ---------------------------------------------------------------------------------------
// Forward, second version (default).  ncu of the kernel above: tensor pipe 13 % active, a third of
// all stall samples in mbarrier spins — S, softmax, PV and the fold of O form ONE serial chain per
// CTA, and two co-resident CTAs are all the overlap there is.  Here a CTA owns TWO 128-query tiles
// that ping-pong on the tensor core:
//
//   warps 0-3 / 4-7   softmax of tile 0 / tile 1, one query row per thread.  ONE pass over S
//                     (128 fp32 in registers): row max, then P = exp2(s*c - m) as bf16 into the
//                     tile's swizzled P buffer.  O is never folded in registers: it accumulates in
//                     TMEM across KV blocks, and is rescaled there (tcgen05.ld / st) only when a
//                     row's maximum grew by more than 2^8 since the value m the probabilities are
//                     expressed in — after the first blocks that is rare, and until then P <= 256
//                     is harmless in bf16 / fp32 (the final 1/l undoes it).
//   warp 8            MMA issuer: ... PV_0(i), S_0(i+1), PV_1(i), S_1(i+1) ...  — while one tile's
//                     softmax runs, the other tile's products occupy the tensor core.
//   warp 9            TMA producer: both Q tiles once, K/V blocks through a 2-stage ring; every
//                     K/V tile feeds four products (two tiles x {S, PV}).
//
// TMEM: S0 128 + S1 128 + O0 64 + O1 64 = 384 of 512 columns.  Shared memory: Q 2 x 16 KB,
// K/V 2 x 2 x 16 KB, P 2 x 32 KB = 160 KB: one CTA per SM, heaviest tile pairs first.
constexpr int FA2_THREADS = 320;
// step timeline of one softmax thread (block 0,0,0; tile 0) for tuning: B200DDP_FA_TRACE=1
__device__ long long g_fa_trace[16 * 8];
#define FA_TRACE(i, k) do { if (p.trace && threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && (i) < 16) g_fa_trace[(i) * 8 + (k)] = clock64(); } while (0)
constexpr int FA2_STAGES = 3;   // K/V ring: S(i+1) is issued while the softmax of block i runs, so block i+2 must be in flight
constexpr uint32_t FA2_SMEM = FA_TILE_BYTES * (2 + 2 * FA2_STAGES) + 2 * 2 * FA_TILE_BYTES /*P0, P1*/ + 256;

__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
        "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
        "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]),
        "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]),
        "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
        : "memory");
}
__device__ __forceinline__ void tmem_st_32x8(uint32_t taddr, const uint32_t (&v)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr), "r"(v[0]),
                 "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
                 : "memory");
}
__device__ __forceinline__ void tmem_ld_32x8(uint32_t taddr, uint32_t (&v)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__global__ void __launch_bounds__(FA2_THREADS, 1)
flash_attn_fwd2_kernel(const __grid_constant__ CUtensorMap tma_qkv, const FaParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    if (threadIdx.x == 0 && (smem_u32(smem) & 1023u) != 0) __trap();
    uint8_t* s_q = smem;                           // 2 tiles
    uint8_t* s_k = s_q + 2 * FA_TILE_BYTES;              // FA2_STAGES stages
    uint8_t* s_v = s_k + FA2_STAGES * FA_TILE_BYTES;     // FA2_STAGES stages
    uint8_t* s_p = s_v + FA2_STAGES * FA_TILE_BYTES;     // 2 tiles x 32 KB
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_p + 4 * FA_TILE_BYTES);
    uint64_t* q_full = bars;          // 1
    uint64_t* k_full = bars + 1;      // 3
    uint64_t* v_full = bars + 4;      // 3
    uint64_t* kv_empty = bars + 7;    // 3
    uint64_t* s_full = bars + 10;     // 2 (per tile)
    uint64_t* p_full = bars + 12;     // 2 (per tile, 128 arrivals)
    uint64_t* o_full = bars + 14;     // 2 (per tile)
    uint64_t* s_free = bars + 16;     // 2 (per tile, 128 arrivals): the scores are in registers, S may be overwritten
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 18);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nq = p.T / FA_BLOCK;                 // 128-query tiles in the sequence
    const int npairs = (nq + 1) / 2;
    // the pair index is the SLOWEST grid axis: every heaviest pair (4u + 3 tile steps) is scheduled
    // before the first lighter one, so the tail of the launch is made of short CTAs
    const int u = npairs - 1 - (int)blockIdx.z;
    const int h = blockIdx.x, b = blockIdx.y;
    // tile t covers query tile 2u + t and walks KV blocks 0 .. 2u + t; tile 1 may not exist (odd nq)
    const int steps0 = 2 * u + 1;
    const int steps1 = (2 * u + 1 < nq) ? 2 * u + 2 : 0;
    const int nblocks = steps1 ? steps1 : steps0;

    if (warp == 8 && lane == 0) {
        tma_prefetch_desc(&tma_qkv);
        mbar_init(q_full, 1);
        for (int s = 0; s < FA2_STAGES; ++s) { mbar_init(&k_full[s], 1); mbar_init(&v_full[s], 1); mbar_init(&kv_empty[s], 1); }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&s_full[s], 1); mbar_init(&p_full[s], FA_BLOCK); mbar_init(&o_full[s], 1); mbar_init(&s_free[s], FA_BLOCK);
        }
        fence_mbar_init();
    }
    if (warp == 8) {
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    asm volatile("griddepcontrol.wait;" ::: "memory");

    if (warp == 9) {
        // ===================== TMA producer =====================
        if (elect_one()) {
            mbar_arrive_expect_tx(q_full, (steps1 ? 2u : 1u) * FA_TILE_BYTES);
            tma_load_4d(s_q, &tma_qkv, q_full, 0, 0 * p.H + h, (2 * u) * FA_BLOCK, b);
            if (steps1) tma_load_4d(s_q + FA_TILE_BYTES, &tma_qkv, q_full, 0, 0 * p.H + h, (2 * u + 1) * FA_BLOCK, b);
        }
        __syncwarp();
        for (int j = 0; j < nblocks; ++j) {
            const int st = j % FA2_STAGES;
            if (j >= FA2_STAGES) mbar_wait(&kv_empty[st], (uint32_t)((j / FA2_STAGES - 1) & 1));
            if (elect_one()) {
                mbar_arrive_expect_tx(&k_full[st], FA_TILE_BYTES);
                tma_load_4d(s_k + st * FA_TILE_BYTES, &tma_qkv, &k_full[st], 0, 1 * p.H + h, j * FA_BLOCK, b);
                mbar_arrive_expect_tx(&v_full[st], FA_TILE_BYTES);
                tma_load_4d(s_v + st * FA_TILE_BYTES, &tma_qkv, &v_full[st], 0, 2 * p.H + h, j * FA_BLOCK, b);
            }
            __syncwarp();
        }
    } else if (warp == 8) {
        // ===================== MMA issuer =====================
        const uint32_t idesc_s = fa_idesc(128, false), idesc_o = fa_idesc(64, true);
        const uint64_t dq0 = umma_smem_desc(smem_u32(s_q), 16, 1024);
        const uint64_t dk0 = umma_smem_desc(smem_u32(s_k), 16, 1024);
        const uint64_t dp0 = umma_smem_desc(smem_u32(s_p), 16, 1024);
        const uint64_t dv0 = umma_smem_desc(smem_u32(s_v), 8192, 1024);
        auto issue_s = [&](int t, int j) {      // S_t = Q_t K_j^T
            const int st = j % FA2_STAGES;
            mbar_wait(&k_full[st], (uint32_t)((j / FA2_STAGES) & 1));
            tc_fence_after();
            if (elect_one()) {
                const uint64_t dq = dq0 + (uint64_t)(t * (FA_TILE_BYTES >> 4)), dk = dk0 + (uint64_t)(st * (FA_TILE_BYTES >> 4));
#pragma unroll
                for (int k = 0; k < FA_D / 16; ++k)
                    umma_bf16(tmem_base + t * 128, dq + (uint64_t)(2 * k), dk + (uint64_t)(2 * k), idesc_s, k != 0 ? 1u : 0u);
                umma_commit(&s_full[t]);
            }
            __syncwarp();
        };
        auto issue_pv = [&](int t, int i, bool last_user) {   // O_t (+)= P_t V_i
            const int st = i % FA2_STAGES;
            mbar_wait(&p_full[t], (uint32_t)(i & 1));
            tc_fence_after();
            mbar_wait(&v_full[st], (uint32_t)((i / FA2_STAGES) & 1));
            tc_fence_after();
            if (elect_one()) {
                const uint64_t dp = dp0 + (uint64_t)(t * 2 * (FA_TILE_BYTES >> 4)), dv = dv0 + (uint64_t)(st * (FA_TILE_BYTES >> 4));
#pragma unroll
                for (int kk = 0; kk < FA_BLOCK / 16; ++kk)
                    umma_bf16(tmem_base + 256 + t * 64, dp + (uint64_t)((kk >> 2) * (FA_TILE_BYTES >> 4) + (kk & 3) * 2),
                              dv + (uint64_t)(kk * 128), idesc_o, (i > 0 || kk != 0) ? 1u : 0u);
                umma_commit(&o_full[t]);
                if (last_user) umma_commit(&kv_empty[st]);
            }
            __syncwarp();
        };
        mbar_wait(q_full, 0);
        issue_s(0, 0);
        if (steps1) issue_s(1, 0);
        for (int i = 0; i < nblocks; ++i) {
            // next scores first: S_t is free as soon as its softmax warps hold block i in registers
            // (a few hundred cycles into their step), so S_t(i+1) is ready long before they finish
            if (i + 1 < steps0) { mbar_wait(&s_free[0], (uint32_t)(i & 1)); tc_fence_after(); issue_s(0, i + 1); }
            if (i + 1 < steps1) { mbar_wait(&s_free[1], (uint32_t)(i & 1)); tc_fence_after(); issue_s(1, i + 1); }
            // KV block i is last used by tile 1's PV when tile 1 exists (it always walks at least as far)
            if (i < steps0) issue_pv(0, i, steps1 == 0);
            if (i < steps1) issue_pv(1, i, true);
        }
    } else {
        // ===================== softmax warps: tile t = warp / 4, thread r owns one query row =====================
        const int t = warp >> 2;
        const int r = threadIdx.x & 127;
        const int steps = t ? steps1 : steps0;
        const int qi = 2 * u + t;
        const uint32_t lane_addr = ((uint32_t)((warp & 3) * 32)) << 16;
        const uint32_t tmem_s = tmem_base + t * 128 + lane_addr, tmem_o = tmem_base + 256 + t * 64 + lane_addr;
        uint8_t* my_p = s_p + t * 2 * FA_TILE_BYTES;
        float m = -INFINITY, l = 0.f;    // m: the maximum the stored probabilities (and O, l) are relative to
        for (int i = 0; i < steps; ++i) {
            const bool diag = (i == qi);
            FA_TRACE(i, 0);
            mbar_wait(&s_full[t], (uint32_t)(i & 1));
            tc_fence_after();
            FA_TRACE(i, 1);
            uint32_t v0[32], v1[32], v2[32], v3[32];
            tmem_ld_32x32(tmem_s, v0);
            tmem_ld_32x32(tmem_s + 32, v1);
            tmem_ld_32x32(tmem_s + 64, v2);
            tmem_ld_32x32(tmem_s + 96, v3);
            tmem_ld_wait();
            tc_fence_before();
            mbar_arrive(&s_free[t]);
            FA_TRACE(i, 2);
            float mx = -INFINITY;
            if (!diag) {     // 29 of 36 steps: no masking arithmetic at all
                // FMNMX3 (two new values per instruction) on four independent chains: a single chain
                // of 64 dependent maxima was 450 cycles of pure latency
                float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
#pragma unroll
                for (int c = 0; c < 32; c += 2) {
                    m0 = fa_max3(m0, __uint_as_float(v0[c]), __uint_as_float(v0[c + 1]));
                    m1 = fa_max3(m1, __uint_as_float(v1[c]), __uint_as_float(v1[c + 1]));
                    m2 = fa_max3(m2, __uint_as_float(v2[c]), __uint_as_float(v2[c + 1]));
                    m3 = fa_max3(m3, __uint_as_float(v3[c]), __uint_as_float(v3[c + 1]));
                }
                mx = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
            } else {
#pragma unroll
                for (int c = 0; c < 32; ++c) {
                    if (c <= r) mx = fmaxf(mx, __uint_as_float(v0[c]));
                    if (32 + c <= r) mx = fmaxf(mx, __uint_as_float(v1[c]));
                    if (64 + c <= r) mx = fmaxf(mx, __uint_as_float(v2[c]));
                    if (96 + c <= r) mx = fmaxf(mx, __uint_as_float(v3[c]));
                }
            }
            const float m_new = fmaxf(m, mx * p.scale_log2e);
            FA_TRACE(i, 3);
            if (i > 0) {
                // P_t of the previous block has been read (and O_t updated) once PV_t(i-1) retired
                mbar_wait(&o_full[t], (uint32_t)((i - 1) & 1));
                tc_fence_after();
                FA_TRACE(i, 4);
                const bool grow = m_new - m > 8.f;
                if (__any_sync(0xffffffffu, grow)) {      // warp-uniform: TMEM ld / st are warp-collective
                    const float m_next = grow ? m_new : m;
                    const float alpha = fa_exp2(m - m_next);
                    // rare path, eight columns at a time: the 128 scores stay in registers meanwhile
#pragma unroll 1
                    for (int c = 0; c < FA_D / 8; ++c) {
                        uint32_t o[8];
                        tmem_ld_32x8(tmem_o + c * 8, o);
                        tmem_ld_wait();
#pragma unroll
                        for (int e = 0; e < 8; ++e) o[e] = __float_as_uint(__uint_as_float(o[e]) * alpha);
                        tmem_st_32x8(tmem_o + c * 8, o);
                    }
                    tmem_st_wait();
                    l *= alpha;
                    m = m_next;
                }
            } else {
                m = m_new;
            }
            float psum = 0.f;
            // eight probabilities at a time (16 bytes of bf16): the 128 scores already fill the register file
            auto emit = [&](const uint32_t (&v)[32], int c32, auto is_diag) {
                constexpr bool DIAG = decltype(is_diag)::value;
                uint8_t* blk = my_p + (c32 >> 1) * FA_TILE_BYTES + r * 128;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    float pv[8];
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        const float arg = fmaf(__uint_as_float(v[8 * q + e]), p.scale_log2e, -m);
                        const float x = fa_exp2(arg);
                        pv[e] = (!DIAG || c32 * 32 + 8 * q + e <= r) ? x : 0.f;
                    }
                    // pairwise: no 128-long chain of dependent additions
                    psum += ((pv[0] + pv[1]) + (pv[2] + pv[3])) + ((pv[4] + pv[5]) + (pv[6] + pv[7]));
                    const int vec = (c32 & 1) * 4 + q;
                    *reinterpret_cast<uint4*>(blk + ((vec ^ (r & 7)) << 4)) =
                        make_uint4(pack_bf16(pv[0], pv[1]), pack_bf16(pv[2], pv[3]), pack_bf16(pv[4], pv[5]), pack_bf16(pv[6], pv[7]));
                }
            };
            if (diag) {
                emit(v0, 0, std::true_type{}); emit(v1, 1, std::true_type{}); emit(v2, 2, std::true_type{}); emit(v3, 3, std::true_type{});
            } else {
                emit(v0, 0, std::false_type{}); emit(v1, 1, std::false_type{}); emit(v2, 2, std::false_type{}); emit(v3, 3, std::false_type{});
            }
            l += psum;
            FA_TRACE(i, 5);
            fence_proxy_async();
            tc_fence_before();
            mbar_arrive(&p_full[t]);
            FA_TRACE(i, 6);
        }
        if (steps > 0) {
            mbar_wait(&o_full[t], (uint32_t)((steps - 1) & 1));
            tc_fence_after();
            const float inv = 1.f / l;
            __nv_bfloat16* dst = p.out + ((int64_t)b * p.T + qi * FA_BLOCK + r) * ((int64_t)p.H * FA_D) + (int64_t)h * FA_D;
#pragma unroll
            for (int c = 0; c < FA_D / 32; ++c) {
                uint32_t o[32];
                tmem_ld_32x32(tmem_o + c * 32, o);
                tmem_ld_wait();
#pragma unroll
                for (int d = 0; d < 32; d += 8) {
                    uint4 w;
                    w.x = pack_bf16(__uint_as_float(o[d]) * inv, __uint_as_float(o[d + 1]) * inv);
                    w.y = pack_bf16(__uint_as_float(o[d + 2]) * inv, __uint_as_float(o[d + 3]) * inv);
                    w.z = pack_bf16(__uint_as_float(o[d + 4]) * inv, __uint_as_float(o[d + 5]) * inv);
                    w.w = pack_bf16(__uint_as_float(o[d + 6]) * inv, __uint_as_float(o[d + 7]) * inv);
                    *reinterpret_cast<uint4*>(dst + c * 32 + d) = w;
                }
            }
            p.lse[((int64_t)b * p.H + h) * p.T + qi * FA_BLOCK + r] = (m + log2f(l)) * 0.6931471805599453f;
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 8) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ---------------------------------------------------------------------------------------
// Backward.  One CTA per (sequence, head, 128-key block j); it keeps K_j, V_j resident and
// walks the query blocks i >= j (causal).  Five tensor-core products per step, all from the same
// swizzled tiles read through K-major or MN-major descriptors (nothing is transposed):
//     S  = Q_i K_j^T          dP = dO_i V_j^T                      -> TMEM (fresh)
//     P = exp2(S*c - lse),  dS = P * (dP - delta) * scale          4 warps, one query row each,
//                                                                  bf16 P and dS into smem
//     dV_j += P^T dO_i        dK_j += dS^T Q_i                     -> TMEM (accumulate over i)
//     dQ_i partial = dS K_j                                        -> TMEM -> fp32 vector atomics
// TMEM: S 128 + dP 128 + dV 64 + dK 64 + dQ 64 = 448 of 512 columns.
constexpr uint32_t FAB_DQ_BYTES = FA_BLOCK * FA_D * 4;   // fp32 dQ tile staged for the bulk reduce-add
constexpr uint32_t FAB_SMEM = FA_TILE_BYTES * (2 + 4) + 4 * FA_TILE_BYTES /*P, dS*/ + FAB_DQ_BYTES + 256;

struct FaBwdParams {
    int B, T, H;
    const float* lse;     // [B, H, T]
    const float* delta;   // [B, H, T]
    float* dq_acc;        // [B, H, T, D] fp32, zeroed
    __nv_bfloat16* dqkv;  // [B, T, 3, H, D]
    float scale, scale_log2e;
    int trace;
};

__device__ __forceinline__ uint32_t fab_idesc(int n, bool a_mn, bool b_mn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(a_mn ? 1 : 0) << 15) | ((uint32_t)(b_mn ? 1 : 0) << 16) |
           ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}

// 8 softmax warps: TWO threads per query row (64 key columns each; warps w and w + 4 share TMEM
// lanes), warp 8 issues TMA and MMA.  lse and delta come from memory, so the two halves of a row
// never talk to each other; with two warps per scheduler the exp2 / dS arithmetic of one hides
// the latencies of the other (one warp per scheduler was the round-1 diagnosis).
constexpr int FAB_THREADS = 288;
__global__ void __launch_bounds__(FAB_THREADS, 1)
flash_attn_bwd_kernel(const __grid_constant__ CUtensorMap tma_qkv, const __grid_constant__ CUtensorMap tma_do,
                      const FaBwdParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    if (threadIdx.x == 0 && (smem_u32(smem) & 1023u) != 0) __trap();
    uint8_t* s_k = smem;                          // K_j
    uint8_t* s_v = s_k + FA_TILE_BYTES;           // V_j
    uint8_t* s_q = s_v + FA_TILE_BYTES;           // Q_i, 2 stages
    uint8_t* s_do = s_q + 2 * FA_TILE_BYTES;      // dO_i, 2 stages
    uint8_t* s_p = s_do + 2 * FA_TILE_BYTES;      // P   [128 q x 128 keys] as two swizzled blocks
    uint8_t* s_ds = s_p + 2 * FA_TILE_BYTES;      // dS  same layout
    uint8_t* s_dq = s_ds + 2 * FA_TILE_BYTES;     // dQ_i partial, fp32 [128][64], 16-byte chunks XOR-swizzled per row
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_dq + FAB_DQ_BYTES);
    uint64_t* kv_full = bars;         // 1
    uint64_t* qdo_full = bars + 1;    // 2
    uint64_t* qdo_empty = bars + 3;   // 2
    uint64_t* sdp_full = bars + 5;    // 1
    uint64_t* pds_full = bars + 6;    // 1 (256 arrivals)
    uint64_t* dq_full = bars + 7;     // 1
    uint64_t* dkv_full = bars + 8;    // 1
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 9);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nq = p.T / FA_BLOCK;
    const int kj = blockIdx.z;                 // key block on the slowest grid axis: block 0 (most query blocks) first
    const int h = blockIdx.x, b = blockIdx.y;
    const int niter = nq - kj;
    const int k0 = kj * FA_BLOCK;

    if (warp == 8 && lane == 0) {
        tma_prefetch_desc(&tma_qkv);
        tma_prefetch_desc(&tma_do);
        mbar_init(kv_full, 1);
        for (int s = 0; s < 2; ++s) { mbar_init(&qdo_full[s], 1); mbar_init(&qdo_empty[s], 1); }
        mbar_init(sdp_full, 1);
        mbar_init(pds_full, 2 * FA_BLOCK);
        mbar_init(dq_full, 1);
        mbar_init(dkv_full, 1);
        fence_mbar_init();
    }
    if (warp == 8) {
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t t_s = tmem_base, t_dp = tmem_base + 128, t_dv = tmem_base + 256, t_dk = tmem_base + 320,
                   t_dq = tmem_base + 384;

    if (warp == 8) {
        // whole warp walks the loop, one elected lane issues (see the forward kernel)
        auto load_qdo = [&](int it) {
            const int st = it & 1, i = kj + it;
            mbar_arrive_expect_tx(&qdo_full[st], 2 * FA_TILE_BYTES);
            tma_load_4d(s_q + st * FA_TILE_BYTES, &tma_qkv, &qdo_full[st], 0, 0 * p.H + h, i * FA_BLOCK, b);
            tma_load_4d(s_do + st * FA_TILE_BYTES, &tma_do, &qdo_full[st], 0, h, i * FA_BLOCK, b);
        };
        if (elect_one()) {
            mbar_arrive_expect_tx(kv_full, 2 * FA_TILE_BYTES);
            tma_load_4d(s_k, &tma_qkv, kv_full, 0, 1 * p.H + h, k0, b);
            tma_load_4d(s_v, &tma_qkv, kv_full, 0, 2 * p.H + h, k0, b);
            load_qdo(0);
            if (niter > 1) load_qdo(1);
        }
        __syncwarp();
        const uint32_t id_s = fab_idesc(128, false, false);   // S, dP: both operands K-major
        const uint32_t id_kv = fab_idesc(64, true, true);     // dV, dK: both MN-major
        const uint32_t id_dq = fab_idesc(64, false, true);    // dQ: dS K-major, K_j MN-major
        constexpr uint32_t TILE16 = FA_TILE_BYTES >> 4;
        const uint64_t k_k = umma_smem_desc(smem_u32(s_k), 16, 1024);         // K_j K-major (S)
        const uint64_t k_mn = umma_smem_desc(smem_u32(s_k), 8192, 1024);      // K_j MN-major (dQ)
        const uint64_t v_k = umma_smem_desc(smem_u32(s_v), 16, 1024);         // V_j K-major (dP)
        const uint64_t q_k0 = umma_smem_desc(smem_u32(s_q), 16, 1024);        // Q_i K-major (S)
        const uint64_t q_mn0 = umma_smem_desc(smem_u32(s_q), 8192, 1024);     // Q_i MN-major (dK)
        const uint64_t do_k0 = umma_smem_desc(smem_u32(s_do), 16, 1024);      // dO_i K-major (dP)
        const uint64_t do_mn0 = umma_smem_desc(smem_u32(s_do), 8192, 1024);   // dO_i MN-major (dV)
        const uint64_t p_mn = umma_smem_desc(smem_u32(s_p), 16384, 1024);     // P^T view (dV)
        const uint64_t ds_mn = umma_smem_desc(smem_u32(s_ds), 16384, 1024);   // dS^T view (dK)
        const uint64_t ds_k = umma_smem_desc(smem_u32(s_ds), 16, 1024);       // dS K-major (dQ)
        mbar_wait(kv_full, 0);
        auto issue_s_dp = [&](int it) {      // S = Q_i K_j^T and dP = dO_i V_j^T of iteration `it`
            const int st = it & 1;
            const uint64_t so = (uint64_t)(st * TILE16);
            mbar_wait(&qdo_full[st], (uint32_t)((it >> 1) & 1));
            tc_fence_after();
            if (elect_one()) {
#pragma unroll
                for (int k = 0; k < FA_D / 16; ++k)
                    umma_bf16(t_s, q_k0 + so + (uint64_t)(2 * k), k_k + (uint64_t)(2 * k), id_s, k != 0 ? 1u : 0u);
#pragma unroll
                for (int k = 0; k < FA_D / 16; ++k)
                    umma_bf16(t_dp, do_k0 + so + (uint64_t)(2 * k), v_k + (uint64_t)(2 * k), id_s, k != 0 ? 1u : 0u);
                umma_commit(sdp_full);
            }
            __syncwarp();
        };
        issue_s_dp(0);
        for (int it = 0; it < niter; ++it) {
            const int st = it & 1;
            const uint32_t par2 = (uint32_t)((it >> 1) & 1);
            const uint64_t so = (uint64_t)(st * TILE16);
            mbar_wait(pds_full, (uint32_t)(it & 1));   // P and dS are in smem, S/dP consumed
            tc_fence_after();
            // Tensor-pipe order: dQ (the softmax warps wait for it), then the NEXT block's S and dP (they wait
            // for those next), and only then dV / dK, which nobody waits for until the last iteration and
            // which therefore run under the next block's softmax arithmetic.
            if (elect_one()) {
#pragma unroll
                for (int kk = 0; kk < FA_BLOCK / 16; ++kk)     // reduce over the 128 keys
                    umma_bf16(t_dq, ds_k + (uint64_t)((kk >> 2) * TILE16 + (kk & 3) * 2), k_mn + (uint64_t)(kk * 128),
                              id_dq, kk != 0 ? 1u : 0u);
                umma_commit(dq_full);
            }
            __syncwarp();
            if (it + 1 < niter) issue_s_dp(it + 1);
            if (elect_one()) {
#pragma unroll
                for (int kk = 0; kk < FA_BLOCK / 16; ++kk) {   // reduce over the 128 query rows
                    // A^T views of P / dS: [q rows, keys] read MN-major (64-key chunks 16 KB apart)
                    umma_bf16(t_dv, p_mn + (uint64_t)(kk * 128), do_mn0 + so + (uint64_t)(kk * 128), id_kv,
                              (it | kk) != 0 ? 1u : 0u);
                    umma_bf16(t_dk, ds_mn + (uint64_t)(kk * 128), q_mn0 + so + (uint64_t)(kk * 128), id_kv,
                              (it | kk) != 0 ? 1u : 0u);
                }
                umma_commit(&qdo_empty[st]);
                if (it == niter - 1) umma_commit(dkv_full);
            }
            __syncwarp();
            if (it + 2 < niter) {
                mbar_wait(&qdo_empty[st], par2);
                if (elect_one()) load_qdo(it + 2);
                __syncwarp();
            }
        }
    } else {
        const int r = threadIdx.x & 127;          // query row inside the block
        const int half = threadIdx.x >> 7;        // key columns [64 half, 64 half + 64)
        const uint32_t lane_addr = ((uint32_t)((warp & 3) * 32)) << 16;
        const int64_t bh = (int64_t)b * p.H + h;
        for (int it = 0; it < niter; ++it) {
            const int i = kj + it;
            const bool diag = (it == 0);
            const int64_t qrow = bh * p.T + (int64_t)i * FA_BLOCK + r;
            const float lse2 = p.lse[qrow] * 1.4426950408889634f;
            const float dlt = p.delta[qrow];
            FA_TRACE(it, 0);
            mbar_wait(sdp_full, (uint32_t)(it & 1));
            tc_fence_after();
            FA_TRACE(it, 1);
            auto chunk = [&](int c, auto is_diag) {
                constexpr bool DIAG = decltype(is_diag)::value;
                uint32_t vs[32], vd[32];
                tmem_ld_32x32(t_s + lane_addr + c * 32, vs);
                tmem_ld_32x32(t_dp + lane_addr + c * 32, vd);
                tmem_ld_wait();
                uint8_t* bp = s_p + (c >> 1) * FA_TILE_BYTES + r * 128;
                uint8_t* bd = s_ds + (c >> 1) * FA_TILE_BYTES + r * 128;
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    float pv[8], dv[8];
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        const float arg = fmaf(__uint_as_float(vs[8 * t + e]), p.scale_log2e, -lse2);
                        const float pe = fa_exp2(arg);
                        const bool ok = !DIAG || c * 32 + 8 * t + e <= r;   // key column <= query row
                        pv[e] = ok ? pe : 0.f;
                        dv[e] = ok ? pe * (__uint_as_float(vd[8 * t + e]) - dlt) * p.scale : 0.f;
                    }
                    const int off = (((c & 1) * 4 + t) ^ (r & 7)) << 4;
                    *reinterpret_cast<uint4*>(bp + off) =
                        make_uint4(pack_bf16(pv[0], pv[1]), pack_bf16(pv[2], pv[3]), pack_bf16(pv[4], pv[5]), pack_bf16(pv[6], pv[7]));
                    *reinterpret_cast<uint4*>(bd + off) =
                        make_uint4(pack_bf16(dv[0], dv[1]), pack_bf16(dv[2], dv[3]), pack_bf16(dv[4], dv[5]), pack_bf16(dv[6], dv[7]));
                }
            };
            if (diag) { chunk(2 * half, std::true_type{}); chunk(2 * half + 1, std::true_type{}); }
            else { chunk(2 * half, std::false_type{}); chunk(2 * half + 1, std::false_type{}); }
            FA_TRACE(it, 2);
            fence_proxy_async();
            tc_fence_before();
            mbar_arrive(pds_full);
            // dQ_i partial -> fp32 accumulator (other key blocks add to the same rows)
            mbar_wait(dq_full, (uint32_t)(it & 1));
            tc_fence_after();
            FA_TRACE(it, 3);
            // dQ_i partial -> staged in shared memory -> ONE bulk reduce-add (cp.reduce.async.bulk) into
            // the fp32 accumulator: the 128 x 64 block is 32 KB contiguous in dq_acc [B, H, T, D].  Round 1
            // issued 8 vector atomics per thread here: 1900 of a step's 6250 cycles.  The staging rows
            // are XOR-swizzled in 16-byte chunks (conflict-free stores); dq_acc keeps that permutation
            // and fa_dq_convert_kernel undoes it.
            {
                uint32_t v[32];
                tmem_ld_32x32(t_dq + lane_addr + half * 32, v);
                if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // previous tile has left s_dq
                tmem_ld_wait();
                asm volatile("bar.sync 1, 256;" ::: "memory");
                uint8_t* row = s_dq + r * 256 + half * 128;
#pragma unroll
                for (int c = 0; c < 8; ++c)
                    *reinterpret_cast<uint4*>(row + ((c ^ (r & 7)) << 4)) = make_uint4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
                fence_proxy_async();
                asm volatile("bar.sync 1, 256;" ::: "memory");
                if (threadIdx.x == 0) {
                    float* dst = p.dq_acc + (bh * p.T + (int64_t)i * FA_BLOCK) * FA_D;
                    asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(dst),
                                 "r"(smem_u32(s_dq)), "r"(FAB_DQ_BYTES)
                                 : "memory");
                    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                }
            }
            tc_fence_before();
            FA_TRACE(it, 4);
        }
        if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // last dQ tile has landed
        // dV_j, dK_j: thread r owns key row k0 + r
        mbar_wait(dkv_full, 0);
        tc_fence_after();
        const int64_t row3 = ((int64_t)b * p.T + k0 + r) * 3;
        __nv_bfloat16* dk_dst = p.dqkv + (row3 + 1) * ((int64_t)p.H * FA_D) + (int64_t)h * FA_D;
        __nv_bfloat16* dv_dst = p.dqkv + (row3 + 2) * ((int64_t)p.H * FA_D) + (int64_t)h * FA_D;
#pragma unroll
        for (int which = 0; which < 2; ++which) {
            const uint32_t taddr = (which == 0 ? t_dk : t_dv) + lane_addr + half * 32;
            __nv_bfloat16* dst = (which == 0 ? dk_dst : dv_dst) + half * 32;
            uint32_t v[32];
            tmem_ld_32x32(taddr, v);
            tmem_ld_wait();
#pragma unroll
            for (int e = 0; e < 32; e += 8)
                *reinterpret_cast<uint4*>(dst + e) = make_uint4(
                    pack_bf16(__uint_as_float(v[e]), __uint_as_float(v[e + 1])),
                    pack_bf16(__uint_as_float(v[e + 2]), __uint_as_float(v[e + 3])),
                    pack_bf16(__uint_as_float(v[e + 4]), __uint_as_float(v[e + 5])),
                    pack_bf16(__uint_as_float(v[e + 6]), __uint_as_float(v[e + 7])));
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 8) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// delta[b,h,t] = sum_d dO[b,t,h,d] * O[b,t,h,d]   (one thread per (b,t,h) row of 64)
__global__ void __launch_bounds__(256) fa_delta_kernel(const __nv_bfloat16* __restrict__ out,
                                                       const __nv_bfloat16* __restrict__ dout,
                                                       float* __restrict__ delta, int B, int T, int H) {
    const int64_t n = (int64_t)B * T * H;
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (int64_t)gridDim.x * blockDim.x) {
        const int hh = (int)(idx % H);
        const int64_t bt = idx / H;
        const int t = (int)(bt % T);
        const int bb = (int)(bt / T);
        const uint4* po = reinterpret_cast<const uint4*>(out + idx * FA_D);
        const uint4* pd = reinterpret_cast<const uint4*>(dout + idx * FA_D);
        float acc = 0.f;
#pragma unroll
        for (int v = 0; v < FA_D / 8; ++v) {
            const uint4 a = po[v], d = pd[v];
            const uint32_t wa[4] = {a.x, a.y, a.z, a.w}, wd[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float2 fa = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&wa[q]));
                const float2 fd = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&wd[q]));
                acc += fa.x * fd.x + fa.y * fd.y;
            }
        }
        delta[((int64_t)bb * H + hh) * T + t] = acc;
    }
}

// dq_acc [B,H,T,D] fp32 -> dqkv[b,t,0,h,:] bf16
__global__ void __launch_bounds__(256) fa_dq_convert_kernel(const float* __restrict__ dq_acc,
                                                            __nv_bfloat16* __restrict__ dqkv, int B, int T, int H) {
    const int64_t nvec = (int64_t)B * H * T * (FA_D / 8);
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < nvec; idx += (int64_t)gridDim.x * blockDim.x) {
        const int v = (int)(idx % (FA_D / 8));
        int64_t r = idx / (FA_D / 8);
        const int t = (int)(r % T); r /= T;
        const int hh = (int)(r % H);
        const int bb = (int)(r / H);
        // the backward kernel stores each 128-byte half row with its 16-byte chunks XOR-swizzled by (t & 7)
        const int q0 = 2 * v, q1 = 2 * v + 1, sw = t & 7;
        const float4* rowp = reinterpret_cast<const float4*>(dq_acc) + (idx / (FA_D / 8)) * (FA_D / 4);
        const float4 a = rowp[(q0 & 8) | ((q0 & 7) ^ sw)], c = rowp[(q1 & 8) | ((q1 & 7) ^ sw)];
        __nv_bfloat16* dst = dqkv + (((int64_t)bb * T + t) * 3 + 0) * ((int64_t)H * FA_D) + (int64_t)hh * FA_D + v * 8;
        *reinterpret_cast<uint4*>(dst) = make_uint4(pack_bf16(a.x, a.y), pack_bf16(a.z, a.w), pack_bf16(c.x, c.y), pack_bf16(c.z, c.w));
    }
}

typedef CUresult (*FaEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                               const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                               CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// [B, T, 3, H, D] bf16 viewed as dims {D, 3H, T, B}; one box = 128 tokens of one (q|k|v, head).
static int fa_make_map(CUtensorMap* map, const void* base, int B, int T, int H, int which_count) {
    static FaEncodeFn enc = nullptr;
    if (!enc) {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) != cudaSuccess || !ptr) {
            b200_set_error("flash_attention: cuTensorMapEncodeTiled unavailable");
            return B200_ERR_INVALID;
        }
        enc = reinterpret_cast<FaEncodeFn>(ptr);
    }
    const int64_t row = (int64_t)which_count * H * FA_D;  // elements per token
    cuuint64_t dims[4] = {(cuuint64_t)FA_D, (cuuint64_t)(which_count * H), (cuuint64_t)T, (cuuint64_t)B};
    cuuint64_t strides[3] = {(cuuint64_t)FA_D * 2, (cuuint64_t)row * 2, (cuuint64_t)row * T * 2};
    cuuint32_t box[4] = {(cuuint32_t)FA_D, 1, (cuuint32_t)FA_BLOCK, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B200_REQUIRE(r == CUDA_SUCCESS, "flash_attention: cuTensorMapEncodeTiled failed (%d)", (int)r);
    return 0;
}

static int fa_bind_ctx() {
    static thread_local bool bound = false;
    if (!bound) {
        int dev = 0;
        B200_CUDA_OK(cudaGetDevice(&dev));
        B200_CUDA_OK(cudaSetDevice(dev));
        bound = true;
    }
    return 0;
}

}  // namespace b200

int b200_flash_init() {
    B200_CUDA_OK(cudaFuncSetAttribute(b200::flash_attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)b200::FA_SMEM));
    B200_CUDA_OK(cudaFuncSetAttribute(b200::flash_attn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)b200::FAB_SMEM));
    B200_CUDA_OK(cudaFuncSetAttribute(b200::flash_attn_fwd2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)b200::FA2_SMEM));
    return 0;
}

extern "C" int b200_flash_debug_trace(long long* out) {
    B200_CUDA_OK(cudaDeviceSynchronize());
    B200_CUDA_OK(cudaMemcpyFromSymbol(out, b200::g_fa_trace, sizeof(long long) * 128));
    return 0;
}

extern "C" int b200_flash_attn_fwd(const void* qkv, void* out, float* lse, int B, int T, int H, int D,
                                   void* stream) {
    using namespace b200;
    B200_REQUIRE(qkv && out && lse, "flash_attn_fwd: null pointer");
    B200_REQUIRE(D == FA_D, "flash_attn_fwd: head dimension must be %d (got %d)", FA_D, D);
    B200_REQUIRE(T > 0 && T % FA_BLOCK == 0, "flash_attn_fwd: T must be a multiple of %d", FA_BLOCK);
    B200_REQUIRE(B > 0 && H > 0 && H <= 65535 && B <= 65535, "flash_attn_fwd: bad batch/heads");
    if (int rc = fa_bind_ctx()) return rc;
    static int inited = b200_flash_init();
    if (inited) return inited;
    CUtensorMap map;
    if (int rc = fa_make_map(&map, qkv, B, T, H, 3)) return rc;
    static const int trace = getenv("B200DDP_FA_TRACE") != nullptr;
    FaParams p{B, T, H, (__nv_bfloat16*)out, lse, 1.4426950408889634f / sqrtf((float)D), trace};
    static const bool v1 = getenv("B200DDP_FLASH_FWD_V1") != nullptr;    // the one-tile-per-CTA kernel, for comparison
    if (v1) {
        dim3 grid((unsigned)(T / FA_BLOCK), (unsigned)H, (unsigned)B);
        flash_attn_fwd_kernel<<<grid, FA_THREADS, FA_SMEM, (cudaStream_t)stream>>>(map, p);
    } else {
        dim3 grid((unsigned)H, (unsigned)B, (unsigned)((T / FA_BLOCK + 1) / 2));
        flash_attn_fwd2_kernel<<<grid, FA2_THREADS, FA2_SMEM, (cudaStream_t)stream>>>(map, p);
    }
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int b200_flash_attn_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                                   float* dq_acc, float* delta, int B, int T, int H, int D, void* stream) {
    using namespace b200;
    B200_REQUIRE(qkv && out && dout && lse && dqkv && dq_acc && delta, "flash_attn_bwd: null pointer");
    B200_REQUIRE(D == FA_D, "flash_attn_bwd: head dimension must be %d (got %d)", FA_D, D);
    B200_REQUIRE(T > 0 && T % FA_BLOCK == 0, "flash_attn_bwd: T must be a multiple of %d", FA_BLOCK);
    B200_REQUIRE(B > 0 && H > 0 && H <= 65535 && B <= 65535, "flash_attn_bwd: bad batch/heads");
    if (int rc = fa_bind_ctx()) return rc;
    static int inited = b200_flash_init();
    if (inited) return inited;
    CUtensorMap map_qkv, map_do;
    if (int rc = fa_make_map(&map_qkv, qkv, B, T, H, 3)) return rc;
    if (int rc = fa_make_map(&map_do, dout, B, T, H, 1)) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const int64_t rows = (int64_t)B * T * H;
    B200_CUDA_OK(cudaMemsetAsync(dq_acc, 0, (size_t)rows * FA_D * sizeof(float), s));
    int blocks = (int)((rows + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    fa_delta_kernel<<<blocks, 256, 0, s>>>((const __nv_bfloat16*)out, (const __nv_bfloat16*)dout, delta, B, T, H);
    const float scale = 1.f / sqrtf((float)D);
    static const int trace = getenv("B200DDP_FA_TRACE_BWD") != nullptr;
    FaBwdParams p{B, T, H, lse, delta, dq_acc, (__nv_bfloat16*)dqkv, scale, scale * 1.4426950408889634f, trace};
    dim3 grid((unsigned)H, (unsigned)B, (unsigned)(T / FA_BLOCK));
    flash_attn_bwd_kernel<<<grid, FAB_THREADS, FAB_SMEM, s>>>(map_qkv, map_do, p);
    fa_dq_convert_kernel<<<148 * 8, 256, 0, s>>>(dq_acc, (__nv_bfloat16*)dqkv, B, T, H);
    b200_count_launch(3);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}
