// Token + position embedding: gather forward, deterministic scatter-add backward (sm_100a).
//
// BASELINE.json config 4 (GPT-2 small); the reference's loop reaches these through
// `output = model(data)` / `loss.backward()` (ref train.py:136,138), where torch runs
// aten::embedding (index_select) and embedding_dense_backward (sort + segmented reduce on CUB).
//
//   forward    x[b, t, :] = wte[idx[b, t], :] + wpe[t, :]            one warp per token, 16-byte vectors
//   backward   g_wte[v, :] += sum over tokens with idx == v of dx     (on top of the tied LM-head gradient
//                                                                      the wgrad GEMM already left there)
//              g_wpe[t, :]  = sum over b of dx[b, t, :]
//
// The scatter-add is deterministic without atomics: one CTA sorts the (token id, position) pairs
// with an in-shared-memory bitonic network — 8192 tokens x 8 bytes = 64 KB, ~90 compare-exchange
// passes — so every vocabulary row's contributors are contiguous and in position order; then one
// warp per run adds its rows in fp32 in that fixed order and does ONE read-modify-write of the
// gradient row.  Up to 16384 tokens per sort (128 KB of shared memory); longer batches are
// processed in consecutive segments, each added in order.
#include "common.cuh"

#include <mutex>

void b200_count_launch(int n);

namespace b200 {

constexpr int EMB_SORT_MAX = 16384;
constexpr int EMB_SORT_THREADS = 1024;

__device__ __forceinline__ void emb_unpack8(const uint4& raw, float (&f)[8]) {
    const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const float2 t = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[q]));
        f[2 * q] = t.x;
        f[2 * q + 1] = t.y;
    }
}

// one warp per token; C % 8 == 0
__global__ void __launch_bounds__(256) embedding_fwd_kernel(const int64_t* __restrict__ idx,
                                                            const __nv_bfloat16* __restrict__ wte,
                                                            const __nv_bfloat16* __restrict__ wpe,
                                                            __nv_bfloat16* __restrict__ out, int64_t tokens, int T,
                                                            int C, int V) {
    const int lane = threadIdx.x & 31;
    const int64_t tok = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (tok >= tokens) return;
    int64_t v = idx[tok];
    v = v < 0 ? 0 : (v >= V ? V - 1 : v);       // clamp: an out-of-range id must not read outside the table
    const int t = (int)(tok % T);
    const uint4* a = reinterpret_cast<const uint4*>(wte + v * C);
    const uint4* b = reinterpret_cast<const uint4*>(wpe + (int64_t)t * C);
    uint4* o = reinterpret_cast<uint4*>(out + tok * C);
    for (int k = lane; k < (C >> 3); k += 32) {
        float x[8], y[8];
        emb_unpack8(__ldg(a + k), x);
        emb_unpack8(__ldg(b + k), y);
        o[k] = make_uint4(pack_bf16(x[0] + y[0], x[1] + y[1]), pack_bf16(x[2] + y[2], x[3] + y[3]),
                          pack_bf16(x[4] + y[4], x[5] + y[5]), pack_bf16(x[6] + y[6], x[7] + y[7]));
    }
}

// keys[i] = (token id << 32) | position, ascending; P = n rounded up to a power of two (padding sorts last)
__global__ void __launch_bounds__(EMB_SORT_THREADS) embedding_sort_kernel(const int64_t* __restrict__ idx, int n, int P,
                                                                          int V, uint64_t* __restrict__ keys) {
    extern __shared__ uint64_t s_keys[];
    for (int i = threadIdx.x; i < P; i += EMB_SORT_THREADS) {
        uint64_t k = ~0ull;
        if (i < n) {
            int64_t v = idx[i];
            v = v < 0 ? 0 : (v >= V ? V - 1 : v);
            k = ((uint64_t)v << 32) | (uint32_t)i;
        }
        s_keys[i] = k;
    }
    __syncthreads();
    for (int k = 2; k <= P; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < P; i += EMB_SORT_THREADS) {
                const int p = i ^ j;
                if (p > i) {
                    const uint64_t a = s_keys[i], b = s_keys[p];
                    const bool up = (i & k) == 0;
                    if ((a > b) == up) { s_keys[i] = b; s_keys[p] = a; }
                }
            }
            __syncthreads();
        }
    for (int i = threadIdx.x; i < n; i += EMB_SORT_THREADS) keys[i] = s_keys[i];
}

// one warp per sorted position; the warp at the head of a run sums the run's dx rows (fp32, position
// order) and adds them to the gradient row.  `base` = position offset of this segment.
template <bool DST_BF16>
__global__ void __launch_bounds__(256) embedding_scatter_kernel(const uint64_t* __restrict__ keys, int n, int64_t base,
                                                                const __nv_bfloat16* __restrict__ dx,
                                                                void* __restrict__ dst_, int C) {
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= n) return;
    const uint32_t v = (uint32_t)(keys[i] >> 32);
    if (i > 0 && (uint32_t)(keys[i - 1] >> 32) == v) return;   // not the head of its run
    for (int k = lane; k < (C >> 3); k += 32) {
        float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (int j = i; j < n; ++j) {
            const uint64_t key = keys[j];
            if ((uint32_t)(key >> 32) != v) break;
            const int64_t pos = base + (uint32_t)key;
            float x[8];
            emb_unpack8(__ldg(reinterpret_cast<const uint4*>(dx + pos * C) + k), x);
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[e] += x[e];
        }
        if (DST_BF16) {
            uint4* d = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(dst_) + (int64_t)v * C) + k;
            float cur[8];
            emb_unpack8(*d, cur);
            *d = make_uint4(pack_bf16(cur[0] + acc[0], cur[1] + acc[1]), pack_bf16(cur[2] + acc[2], cur[3] + acc[3]),
                            pack_bf16(cur[4] + acc[4], cur[5] + acc[5]), pack_bf16(cur[6] + acc[6], cur[7] + acc[7]));
        } else {
            float4* d = reinterpret_cast<float4*>(reinterpret_cast<float*>(dst_) + (int64_t)v * C) + 2 * k;
            float4 lo = d[0], hi = d[1];
            lo.x += acc[0]; lo.y += acc[1]; lo.z += acc[2]; lo.w += acc[3];
            hi.x += acc[4]; hi.y += acc[5]; hi.z += acc[6]; hi.w += acc[7];
            d[0] = lo;
            d[1] = hi;
        }
    }
}

// g_wpe[t, :] = sum_b dx[b, t, :] for t < T, zero for T <= t < rows; one warp per row
template <bool DST_BF16>
__global__ void __launch_bounds__(256) position_grad_kernel(const __nv_bfloat16* __restrict__ dx, void* __restrict__ dst_,
                                                            int Bsz, int T, int rows, int C) {
    const int lane = threadIdx.x & 31;
    const int t = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (t >= rows) return;
    for (int k = lane; k < (C >> 3); k += 32) {
        float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if (t < T)
            for (int b = 0; b < Bsz; ++b) {
                float x[8];
                emb_unpack8(__ldg(reinterpret_cast<const uint4*>(dx + ((int64_t)b * T + t) * C) + k), x);
#pragma unroll
                for (int e = 0; e < 8; ++e) acc[e] += x[e];
            }
        if (DST_BF16) {
            reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(dst_) + (int64_t)t * C)[k] =
                make_uint4(pack_bf16(acc[0], acc[1]), pack_bf16(acc[2], acc[3]), pack_bf16(acc[4], acc[5]),
                           pack_bf16(acc[6], acc[7]));
        } else {
            float4* d = reinterpret_cast<float4*>(reinterpret_cast<float*>(dst_) + (int64_t)t * C) + 2 * k;
            d[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
            d[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
        }
    }
}

}  // namespace b200

extern "C" int b200_embedding_fwd(const int64_t* idx, const void* wte, const void* wpe, void* out, int64_t tokens,
                                  int T, int C, int V, void* stream) {
    B200_REQUIRE(idx && wte && wpe && out, "embedding_fwd: null pointer");
    B200_REQUIRE(tokens > 0 && T > 0 && V > 0 && C > 0 && (C % 8) == 0, "embedding_fwd: C must be a multiple of 8");
    B200_REQUIRE(((uintptr_t)wte | (uintptr_t)wpe | (uintptr_t)out) % 16 == 0, "embedding_fwd: 16-byte alignment");
    const int64_t blocks = (tokens + 7) / 8;
    b200::embedding_fwd_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        idx, (const __nv_bfloat16*)wte, (const __nv_bfloat16*)wpe, (__nv_bfloat16*)out, tokens, T, C, V);
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int64_t b200_embedding_scratch_bytes(int64_t tokens) {
    const int64_t seg = tokens < b200::EMB_SORT_MAX ? tokens : b200::EMB_SORT_MAX;
    return seg * 8;
}

extern "C" int b200_embedding_bwd(const int64_t* idx, const void* dx, void* g_wte, int g_wte_bf16, void* g_wpe,
                                  int g_wpe_bf16, int Bsz, int T, int wpe_rows, int C, int V, void* scratch,
                                  void* stream) {
    using namespace b200;
    B200_REQUIRE(idx && dx && scratch, "embedding_bwd: null pointer");
    B200_REQUIRE(Bsz > 0 && T > 0 && V > 0 && C > 0 && (C % 8) == 0, "embedding_bwd: C must be a multiple of 8");
    B200_REQUIRE(((uintptr_t)dx | (uintptr_t)g_wte | (uintptr_t)g_wpe) % 16 == 0, "embedding_bwd: 16-byte alignment");
    static std::once_flag once;
    static cudaError_t attr_rc = cudaSuccess;
    std::call_once(once, [] {
        attr_rc = cudaFuncSetAttribute(embedding_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       EMB_SORT_MAX * 8);
    });
    B200_CUDA_OK(attr_rc);
    cudaStream_t s = (cudaStream_t)stream;
    const int64_t tokens = (int64_t)Bsz * T;
    if (g_wte) {
        uint64_t* keys = (uint64_t*)scratch;
        for (int64_t base = 0; base < tokens; base += EMB_SORT_MAX) {   // segments are added in order
            const int n = (int)((tokens - base) < EMB_SORT_MAX ? (tokens - base) : EMB_SORT_MAX);
            int P = 1;
            while (P < n) P <<= 1;
            embedding_sort_kernel<<<1, EMB_SORT_THREADS, (size_t)P * 8, s>>>(idx + base, n, P, V, keys);
            const unsigned blocks = (unsigned)((n + 7) / 8);
            if (g_wte_bf16)
                embedding_scatter_kernel<true><<<blocks, 256, 0, s>>>(keys, n, base, (const __nv_bfloat16*)dx, g_wte, C);
            else
                embedding_scatter_kernel<false><<<blocks, 256, 0, s>>>(keys, n, base, (const __nv_bfloat16*)dx, g_wte, C);
            b200_count_launch(2);
        }
    }
    if (g_wpe) {
        B200_REQUIRE(wpe_rows >= T, "embedding_bwd: position table shorter than the sequence");
        const unsigned blocks = (unsigned)((wpe_rows + 7) / 8);
        if (g_wpe_bf16)
            position_grad_kernel<true><<<blocks, 256, 0, s>>>((const __nv_bfloat16*)dx, g_wpe, Bsz, T, wpe_rows, C);
        else
            position_grad_kernel<false><<<blocks, 256, 0, s>>>((const __nv_bfloat16*)dx, g_wpe, Bsz, T, wpe_rows, C);
        b200_count_launch(1);
    }
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}
