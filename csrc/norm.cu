// LayerNorm forward/backward and column sums (bias gradients) for the
// transformer configuration (BASELINE.json config 4: GPT-2 small).
//
// The reference has no normalisation layer of its own; its GPT-2 baseline gets
// these from ATen (native_layer_norm / native_layer_norm_backward / sum).  Here:
//   * forward: one warp per row, 128-bit bf16 loads, fp32 statistics saved;
//   * backward: the same warp recomputes x-hat, produces dx, and carries its
//     partial dgamma/dbeta in registers over all rows it owns; partials go to
//     scratch and a second tiny kernel reduces them in a fixed order
//     (deterministic, no float atomics);
//   * colsum: bias gradient of a [rows, cols] bf16 tensor with the same
//     two-stage scheme.
#include "common.cuh"

void b200_count_launch(int n);

namespace b200 {

constexpr int LN_WARPS = 8;

__device__ __forceinline__ void unpack8(const uint4& raw, float (&f)[8]) {
    const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const float2 t = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[q]));
        f[2 * q] = t.x;
        f[2 * q + 1] = t.y;
    }
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
    uint4 o;
    o.x = pack_bf16(f[0], f[1]);
    o.y = pack_bf16(f[2], f[3]);
    o.z = pack_bf16(f[4], f[5]);
    o.w = pack_bf16(f[6], f[7]);
    return o;
}

// C = ITER * 256 columns; lane owns 8-element vectors at column (it*32 + lane)*8.
template <int ITER>
__global__ void __launch_bounds__(LN_WARPS * 32)
layernorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ gamma,
                     const __nv_bfloat16* __restrict__ beta, __nv_bfloat16* __restrict__ y,
                     float* __restrict__ mean_out, float* __restrict__ rstd_out, int64_t rows, float eps) {
    constexpr int C = ITER * 256;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float g[ITER][8], b[ITER][8];
#pragma unroll
    for (int it = 0; it < ITER; ++it) {
        unpack8(reinterpret_cast<const uint4*>(gamma)[it * 32 + lane], g[it]);
        unpack8(reinterpret_cast<const uint4*>(beta)[it * 32 + lane], b[it]);
    }
    for (int64_t row = (int64_t)blockIdx.x * LN_WARPS + warp; row < rows; row += (int64_t)gridDim.x * LN_WARPS) {
        const uint4* xr = reinterpret_cast<const uint4*>(x + row * C);
        float v[ITER][8];
        float s = 0.f;
#pragma unroll
        for (int it = 0; it < ITER; ++it) {
            unpack8(xr[it * 32 + lane], v[it]);
#pragma unroll
            for (int k = 0; k < 8; ++k) s += v[it][k];
        }
        const float mean = warp_sum(s) * (1.f / C);
        float q = 0.f;
#pragma unroll
        for (int it = 0; it < ITER; ++it)
#pragma unroll
            for (int k = 0; k < 8; ++k) { const float d = v[it][k] - mean; q += d * d; }
        const float rstd = rsqrtf(warp_sum(q) * (1.f / C) + eps);
        uint4* yr = reinterpret_cast<uint4*>(y + row * C);
#pragma unroll
        for (int it = 0; it < ITER; ++it) {
            float o[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) o[k] = (v[it][k] - mean) * rstd * g[it][k] + b[it][k];
            yr[it * 32 + lane] = pack8(o);
        }
        if (lane == 0) { mean_out[row] = mean; rstd_out[row] = rstd; }
    }
}

// partials layout: [gridDim.x * LN_WARPS][2][C] fp32 (dgamma then dbeta)
template <int ITER>
__global__ void __launch_bounds__(LN_WARPS * 32)
layernorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                     const __nv_bfloat16* __restrict__ gamma, const float* __restrict__ mean_in,
                     const float* __restrict__ rstd_in, __nv_bfloat16* __restrict__ dx,
                     const __nv_bfloat16* __restrict__ dx_add, float* __restrict__ partials, int64_t rows) {
    constexpr int C = ITER * 256;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float g[ITER][8], dg[ITER][8], db[ITER][8];
#pragma unroll
    for (int it = 0; it < ITER; ++it) {
        unpack8(reinterpret_cast<const uint4*>(gamma)[it * 32 + lane], g[it]);
#pragma unroll
        for (int k = 0; k < 8; ++k) { dg[it][k] = 0.f; db[it][k] = 0.f; }
    }
    for (int64_t row = (int64_t)blockIdx.x * LN_WARPS + warp; row < rows; row += (int64_t)gridDim.x * LN_WARPS) {
        const uint4* xr = reinterpret_cast<const uint4*>(x + row * C);
        const uint4* dyr = reinterpret_cast<const uint4*>(dy + row * C);
        const float mean = mean_in[row], rstd = rstd_in[row];
        float xh[ITER][8], dxh[ITER][8];
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int it = 0; it < ITER; ++it) {
            float xv[8], dv[8];
            unpack8(xr[it * 32 + lane], xv);
            unpack8(dyr[it * 32 + lane], dv);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                xh[it][k] = (xv[k] - mean) * rstd;
                dxh[it][k] = dv[k] * g[it][k];
                s1 += dxh[it][k];
                s2 += dxh[it][k] * xh[it][k];
                dg[it][k] += dv[k] * xh[it][k];
                db[it][k] += dv[k];
            }
        }
        s1 = warp_sum(s1) * (1.f / C);
        s2 = warp_sum(s2) * (1.f / C);
        uint4* dxr = reinterpret_cast<uint4*>(dx + row * C);
#pragma unroll
        for (int it = 0; it < ITER; ++it) {
            float o[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) o[k] = rstd * (dxh[it][k] - s1 - xh[it][k] * s2);
            if (dx_add) {  // residual-stream gradient joins here
                float r[8];
                unpack8(reinterpret_cast<const uint4*>(dx_add + row * C)[it * 32 + lane], r);
#pragma unroll
                for (int k = 0; k < 8; ++k) o[k] += r[k];
            }
            dxr[it * 32 + lane] = pack8(o);
        }
    }
    // block-level reduction over the 8 warps (fixed order), one partial row per block;
    // the [8][C] staging buffer is used twice (dgamma, then dbeta): 8*2048*4 = 64 KB at most
    extern __shared__ float ln_stage[];
    float* p = partials + (int64_t)blockIdx.x * 2 * C;
#pragma unroll
    for (int pass = 0; pass < 2; ++pass) {
#pragma unroll
        for (int it = 0; it < ITER; ++it)
#pragma unroll
            for (int k = 0; k < 8; ++k)
                ln_stage[warp * C + (it * 32 + lane) * 8 + k] = pass == 0 ? dg[it][k] : db[it][k];
        __syncthreads();
        for (int c = threadIdx.x; c < C; c += LN_WARPS * 32) {
            float t = 0.f;
#pragma unroll
            for (int w = 0; w < LN_WARPS; ++w) t += ln_stage[w * C + c];
            p[pass * C + c] = t;
        }
        __syncthreads();
    }
}

// out[c] = sum_p partials[p * pstride + c].  Block = 8 columns x 128 row groups: thread (x, y)
// sums rows y, y+128, ... (32-byte segments), then the 128 groups are added in a fixed order in
// two levels through shared memory -> deterministic.  Output bf16 or fp32.  blockIdx.y selects a
// second (partials, out) pair so dgamma and dbeta share a launch.
// (First version: 32 columns x 8 row groups — with 296 partial rows every thread walked 37
// dependent L2 round trips, ~25 us per call, and LayerNorm backward made two calls.)
constexpr int RP_THREADS = 1024;
__global__ void __launch_bounds__(RP_THREADS) reduce_partials_kernel(const float* __restrict__ partials0,
                                                                      const float* __restrict__ partials1,
                                                                      int64_t nparts, int64_t pstride, int cols,
                                                                      void* __restrict__ out0, void* __restrict__ out1,
                                                                      int out_bf16) {
    __shared__ float sm[128][9];
    __shared__ float sm2[16][9];
    const float* __restrict__ partials = blockIdx.y ? partials1 : partials0;
    void* out = blockIdx.y ? out1 : out0;
    const int x = threadIdx.x & 7, y = threadIdx.x >> 3;
    const int c = blockIdx.x * 8 + x;
    float s = 0.f;
    if (c < cols) {
#pragma unroll 4
        for (int64_t p = y; p < nparts; p += 128) s += partials[p * pstride + c];
    }
    sm[y][x] = s;
    __syncthreads();
    if (y < 16) {
        float t = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) t += sm[y * 8 + k][x];
        sm2[y][x] = t;
    }
    __syncthreads();
    if (y == 0 && c < cols) {
        float t = 0.f;
#pragma unroll
        for (int k = 0; k < 16; ++k) t += sm2[k][x];
        if (out_bf16) reinterpret_cast<__nv_bfloat16*>(out)[c] = __float2bfloat16_rn(t);
        else reinterpret_cast<float*>(out)[c] = t;
    }
}

// Stage 1 of a column sum: block (bx, by) sums rows [by*rows_per, ...) of the 8-column
// vectors bx*256 + tid.  partials [gridDim.y][cols_padded] fp32.
__global__ void __launch_bounds__(256) colsum_stage1_kernel(const __nv_bfloat16* __restrict__ x, int64_t ld,
                                                            int64_t rows, int cols, int64_t rows_per,
                                                            float* __restrict__ partials, int64_t pstride) {
    const int vec = blockIdx.x * 256 + threadIdx.x;
    const int c0 = vec * 8;
    if (c0 >= cols) return;
    const int64_t r0 = (int64_t)blockIdx.y * rows_per;
    const int64_t r1 = r0 + rows_per < rows ? r0 + rows_per : rows;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (c0 + 8 <= cols && (ld & 7) == 0) {
#pragma unroll 4
        for (int64_t r = r0; r < r1; ++r) {
            float v[8];
            unpack8(*reinterpret_cast<const uint4*>(x + r * ld + c0), v);
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] += v[k];
        }
    } else {
        for (int64_t r = r0; r < r1; ++r)
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if (c0 + k < cols) acc[k] += __bfloat162float(x[r * ld + c0 + k]);
    }
    float* p = partials + (int64_t)blockIdx.y * pstride + c0;
#pragma unroll
    for (int k = 0; k < 8; ++k)
        if (c0 + k < cols) p[k] = acc[k];
}

static int ln_grid(int64_t rows) {
    int64_t b = (rows + LN_WARPS - 1) / LN_WARPS;
    if (b > 148 * 2) b = 148 * 2;
    if (b < 1) b = 1;
    return (int)b;
}

}  // namespace b200

extern "C" int64_t b200_layernorm_bwd_scratch_bytes(int64_t rows, int cols) {
    return (int64_t)b200::ln_grid(rows) * 2 * cols * (int64_t)sizeof(float);
}

extern "C" int b200_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y,
                                  float* mean, float* rstd, int64_t rows, int cols, float eps,
                                  void* stream) {
    using namespace b200;
    B200_REQUIRE(x && gamma && beta && y && mean && rstd, "layernorm_fwd: null pointer");
    B200_REQUIRE(cols % 256 == 0 && cols >= 256 && cols <= 2048, "layernorm_fwd: cols must be a multiple of 256 in [256, 2048] (got %d)", cols);
    B200_REQUIRE(rows > 0, "layernorm_fwd: no rows");
    const int grid = ln_grid(rows);
    cudaStream_t s = (cudaStream_t)stream;
#define LN_FWD(I) case I: layernorm_fwd_kernel<I><<<grid, LN_WARPS * 32, 0, s>>>(                       \
        (const __nv_bfloat16*)x, (const __nv_bfloat16*)gamma, (const __nv_bfloat16*)beta,               \
        (__nv_bfloat16*)y, mean, rstd, rows, eps); break;
    switch (cols / 256) { LN_FWD(1) LN_FWD(2) LN_FWD(3) LN_FWD(4) LN_FWD(5) LN_FWD(6) LN_FWD(7) LN_FWD(8) }
#undef LN_FWD
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int b200_layernorm_bwd(const void* dy, const void* x, const void* gamma, const float* mean,
                                  const float* rstd, void* dx, const void* dx_add, void* dgamma,
                                  void* dbeta, int grads_bf16, int64_t rows, int cols, void* scratch,
                                  void* stream) {
    using namespace b200;
    B200_REQUIRE(dy && x && gamma && mean && rstd && dx && dgamma && dbeta && scratch, "layernorm_bwd: null pointer");
    B200_REQUIRE(cols % 256 == 0 && cols >= 256 && cols <= 2048, "layernorm_bwd: cols must be a multiple of 256 in [256, 2048] (got %d)", cols);
    const int grid = ln_grid(rows);
    cudaStream_t s = (cudaStream_t)stream;
    float* partials = (float*)scratch;
    const size_t ln_smem = (size_t)LN_WARPS * cols * sizeof(float);
    static bool ln_attr = false;
    if (!ln_attr) {  // > 48 KB for cols > 1536
        cudaFuncSetAttribute(layernorm_bwd_kernel<7>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
        cudaFuncSetAttribute(layernorm_bwd_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
        ln_attr = true;
    }
#define LN_BWD(I) case I: layernorm_bwd_kernel<I><<<grid, LN_WARPS * 32, ln_smem, s>>>(                       \
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (const __nv_bfloat16*)gamma, mean, rstd,     \
        (__nv_bfloat16*)dx, (const __nv_bfloat16*)dx_add, partials, rows); break;
    switch (cols / 256) { LN_BWD(1) LN_BWD(2) LN_BWD(3) LN_BWD(4) LN_BWD(5) LN_BWD(6) LN_BWD(7) LN_BWD(8) }
#undef LN_BWD
    const int64_t nparts = (int64_t)grid;
    reduce_partials_kernel<<<dim3((cols + 7) / 8, 2), RP_THREADS, 0, s>>>(partials, partials + cols, nparts, 2 * (int64_t)cols,
                                                                        cols, dgamma, dbeta, grads_bf16);
    b200_count_launch(2);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int64_t b200_colsum_scratch_bytes(int64_t rows, int cols) {
    (void)rows;
    return 512 * (int64_t)((cols + 7) / 8 * 8) * (int64_t)sizeof(float);
}

extern "C" int b200_colsum(const void* x, int64_t ld, int64_t rows, int cols, void* out, int out_bf16,
                           void* scratch, void* stream) {
    using namespace b200;
    B200_REQUIRE(x && out && scratch && rows > 0 && cols > 0 && ld >= cols, "colsum: bad arguments");
    const int64_t pstride = (cols + 7) / 8 * 8;
    // enough row chunks to put ~4 blocks on every SM (the stage is latency-bound otherwise)
    const int col_blocks = (int)((pstride / 8 + 255) / 256);
    int chunks = (int)((rows + 31) / 32);
    const int want = (148 * 4 + col_blocks - 1) / col_blocks;
    if (chunks > want) chunks = want;
    if (chunks > 512) chunks = 512;
    if (chunks < 1) chunks = 1;
    const int64_t rows_per = (rows + chunks - 1) / chunks;
    dim3 grid((unsigned)((pstride / 8 + 255) / 256), (unsigned)chunks);
    cudaStream_t s = (cudaStream_t)stream;
    colsum_stage1_kernel<<<grid, 256, 0, s>>>((const __nv_bfloat16*)x, ld, rows, cols, rows_per,
                                             (float*)scratch, pstride);
    reduce_partials_kernel<<<dim3((cols + 7) / 8, 1), RP_THREADS, 0, s>>>((const float*)scratch, nullptr, chunks, pstride, cols,
                                                                        out, nullptr, out_bf16);
    b200_count_launch(2);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}
