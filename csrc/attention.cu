// Causal softmax forward/backward over materialised score matrices.
//
// For BASELINE.json config 4 (GPT-2 small, seq 1024).  The reference has no attention; its
// GPT-2 baseline reaches a library flash-attention kernel through torch SDPA.  Here the
// attention core is built from our own pieces: batched causal tcgen05 GEMMs (gemm.cu:
// S = QK^T, O = PV, dP = dO V^T, dV = P^T dO, dQ = dS K, dK = dS^T Q, with tiles / K ranges
// outside the causal triangle skipped) around these two row kernels.
//
// Layout: scores [batch*heads, T, T] bf16, row-major; only columns <= row are meaningful.
// Both kernels also write zeros in (row, end of the row's 128-block) because the GEMMs that
// consume P / dS read whole 128-row / 64-column tiles up to that boundary.
#include "common.cuh"

void b200_count_launch(int n);

namespace b200 {

__device__ __forceinline__ void unpack8f(const uint4& raw, float (&f)[8]) {
    const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const float2 t = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[q]));
        f[2 * q] = t.x;
        f[2 * q + 1] = t.y;
    }
}
__device__ __forceinline__ uint4 pack8f(const float (&f)[8]) {
    return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}

// One warp per row.  P = softmax(scale * S) over columns [0, row].
__global__ void __launch_bounds__(256) causal_softmax_fwd_kernel(__nv_bfloat16* __restrict__ s, int T,
                                                                 int64_t rows, float scale_log2e) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int64_t gr = (int64_t)blockIdx.x * 8 + warp; gr < rows; gr += (int64_t)gridDim.x * 8) {
        const int r = (int)(gr % T);
        uint4* row = reinterpret_cast<uint4*>(s + gr * T);
        const int nvalid = r + 1;
        const int nv = (nvalid + 7) >> 3;                      // vectors holding valid columns
        const int nv_end = min(T, ((r >> 7) + 1) << 7) >> 3;   // vectors up to the 128-block end
        float mx = -INFINITY;
        for (int v = lane; v < nv; v += 32) {
            float f[8];
            unpack8f(row[v], f);
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if (v * 8 + k < nvalid) mx = fmaxf(mx, f[k]);
        }
        mx = warp_max(mx) * scale_log2e;
        float sum = 0.f;
        for (int v = lane; v < nv; v += 32) {
            float f[8];
            unpack8f(row[v], f);
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if (v * 8 + k < nvalid) sum += exp2f(f[k] * scale_log2e - mx);
        }
        const float inv = 1.f / warp_sum(sum);
        for (int v = lane; v < nv_end; v += 32) {
            float f[8];
            if (v < nv) unpack8f(row[v], f);
#pragma unroll
            for (int k = 0; k < 8; ++k)
                f[k] = (v * 8 + k < nvalid) ? exp2f(f[k] * scale_log2e - mx) * inv : 0.f;
            row[v] = pack8f(f);
        }
    }
}

// dS = P * (dP - sum_c dP*P) * scale, in place over dP.
__global__ void __launch_bounds__(256) causal_softmax_bwd_kernel(const __nv_bfloat16* __restrict__ p,
                                                                 __nv_bfloat16* __restrict__ dp, int T,
                                                                 int64_t rows, float scale) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int64_t gr = (int64_t)blockIdx.x * 8 + warp; gr < rows; gr += (int64_t)gridDim.x * 8) {
        const int r = (int)(gr % T);
        const uint4* prow = reinterpret_cast<const uint4*>(p + gr * T);
        uint4* drow = reinterpret_cast<uint4*>(dp + gr * T);
        const int nvalid = r + 1;
        const int nv = (nvalid + 7) >> 3;
        const int nv_end = min(T, ((r >> 7) + 1) << 7) >> 3;
        float dot = 0.f;
        for (int v = lane; v < nv; v += 32) {
            float a[8], b[8];
            unpack8f(prow[v], a);
            unpack8f(drow[v], b);
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if (v * 8 + k < nvalid) dot += a[k] * b[k];
        }
        dot = warp_sum(dot);
        for (int v = lane; v < nv_end; v += 32) {
            float a[8], b[8], o[8];
            if (v < nv) { unpack8f(prow[v], a); unpack8f(drow[v], b); }
#pragma unroll
            for (int k = 0; k < 8; ++k)
                o[k] = (v * 8 + k < nvalid) ? a[k] * (b[k] - dot) * scale : 0.f;
            drow[v] = pack8f(o);
        }
    }
}

static int sm_grid(int64_t rows) {
    int64_t b = (rows + 7) / 8;
    if (b > 148 * 8) b = 148 * 8;
    return (int)(b < 1 ? 1 : b);
}

}  // namespace b200

extern "C" int b200_causal_softmax_fwd(void* scores, int64_t batch_heads, int T, float scale, void* stream) {
    using namespace b200;
    B200_REQUIRE(scores && batch_heads > 0 && T > 0 && T % 8 == 0, "causal_softmax_fwd: T must be a multiple of 8");
    const int64_t rows = batch_heads * T;
    causal_softmax_fwd_kernel<<<sm_grid(rows), 256, 0, (cudaStream_t)stream>>>(
        (__nv_bfloat16*)scores, T, rows, scale * 1.4426950408889634f);
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int b200_causal_softmax_bwd(const void* probs, void* dprobs, int64_t batch_heads, int T, float scale,
                                       void* stream) {
    using namespace b200;
    B200_REQUIRE(probs && dprobs && batch_heads > 0 && T > 0 && T % 8 == 0, "causal_softmax_bwd: T must be a multiple of 8");
    const int64_t rows = batch_heads * T;
    causal_softmax_bwd_kernel<<<sm_grid(rows), 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)probs, (__nv_bfloat16*)dprobs, T, rows, scale);
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}
