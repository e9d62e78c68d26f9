// Fused causal attention (flash-style) on tcgen05: forward.
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
};

__device__ __forceinline__ float fa_exp2(float x) {   // MUFU ex2; inputs are <= 0 here
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
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
        if (lane == 0) {
            auto load_kv = [&](int j) {
                const int st = j & 1;
                mbar_arrive_expect_tx(&k_full[st], FA_TILE_BYTES);
                tma_load_4d(s_k + st * FA_TILE_BYTES, &tma_qkv, &k_full[st], 0, 1 * p.H + h, j * FA_BLOCK, b);
                mbar_arrive_expect_tx(&v_full[st], FA_TILE_BYTES);
                tma_load_4d(s_v + st * FA_TILE_BYTES, &tma_qkv, &v_full[st], 0, 2 * p.H + h, j * FA_BLOCK, b);
            };
            mbar_arrive_expect_tx(q_full, FA_TILE_BYTES);
            tma_load_4d(s_q, &tma_qkv, q_full, 0, 0 * p.H + h, q0, b);
            load_kv(0);
            if (nblocks > 1) load_kv(1);
            const uint32_t idesc_s = fa_idesc(128, false), idesc_o = fa_idesc(64, true);
            const uint32_t aq = smem_u32(s_q), ap = smem_u32(s_p);
            mbar_wait(q_full, 0);
            for (int j = 0; j < nblocks; ++j) {
                const int st = j & 1;
                const uint32_t kv_par = (uint32_t)((j >> 1) & 1);
                // ---- S = Q K^T
                mbar_wait(&k_full[st], kv_par);
                tc_fence_after();
                const uint32_t ak = smem_u32(s_k + st * FA_TILE_BYTES);
#pragma unroll
                for (int k = 0; k < FA_D / 16; ++k)
                    umma_bf16(tmem_s, umma_smem_desc(aq + k * 32, 16, 1024), umma_smem_desc(ak + k * 32, 16, 1024),
                              idesc_s, k != 0 ? 1u : 0u);
                umma_commit(s_full);
                // ---- O_j = P V once the softmax warps have written P (and left S)
                mbar_wait(p_full, (uint32_t)(j & 1));
                tc_fence_after();
                mbar_wait(&v_full[st], kv_par);
                tc_fence_after();
                const uint32_t av = smem_u32(s_v + st * FA_TILE_BYTES);
#pragma unroll
                for (int kk = 0; kk < FA_BLOCK / 16; ++kk)
                    umma_bf16(tmem_o, umma_smem_desc(ap + (kk >> 2) * FA_TILE_BYTES + (kk & 3) * 32, 16, 1024),
                              umma_smem_desc(av + kk * 2048, 8192, 1024), idesc_o, kk != 0 ? 1u : 0u);
                umma_commit(o_full);
                umma_commit(&kv_empty[st]);
                if (j + 2 < nblocks) {  // refill this K/V stage for block j+2
                    mbar_wait(&kv_empty[st], kv_par);
                    load_kv(j + 2);
                }
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
    FaParams p{B, T, H, (__nv_bfloat16*)out, lse, 1.4426950408889634f / sqrtf((float)D)};
    dim3 grid((unsigned)(T / FA_BLOCK), (unsigned)H, (unsigned)B);
    flash_attn_fwd_kernel<<<grid, FA_THREADS, FA_SMEM, (cudaStream_t)stream>>>(map, p);
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}
