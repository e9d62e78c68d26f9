// BatchNorm (training and inference) for NHWC bf16 activations, fused with ReLU and the
// residual add, forward and backward.
//
// For BASELINE.json config 3 (ResNet-50).  The reference has no normalisation; its
// ResNet-50 baseline runs ATen's batch_norm_collect_statistics / transform_input /
// backward_reduce / backward_elemt plus separate ReLU and add kernels (~12 passes over the
// activation per layer).  Here a layer is 2 passes forward (statistics; normalise + affine +
// residual + ReLU) and 2 backward (reductions with the ReLU mask applied on the fly;
// dx and the residual-branch gradient), all 16-byte vectorised along channels.
//
// Activations are [rows = N*H*W, C] row-major, C a power-of-two multiple of 8 up to 2048.
// Statistics are fp32 and reduced deterministically: per-block partials -> fixed-order sum.
#include "common.cuh"

void b200_count_launch(int n);

namespace b200 {

constexpr int BN_THREADS = 256;
constexpr int BN_MAX_BLOCKS = 592;   // 4 resident blocks per SM

__device__ __forceinline__ void bn_unpack8(const uint4& raw, float (&f)[8]) {
    const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const float2 t = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[q]));
        f[2 * q] = t.x;
        f[2 * q + 1] = t.y;
    }
}
__device__ __forceinline__ uint4 bn_pack8(const float (&f)[8]) {
    return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}

// Per-block partial sums of two per-channel quantities.
//   MODE 0 (forward):  (x, x^2)
//   MODE 1 (backward): (g, g * xhat) with g = dy * [y > 0] when relu, xhat = (x - mean) * rstd
// partials layout: [gridDim.x][2][C].
template <int MODE>
__global__ void __launch_bounds__(BN_THREADS) bn_reduce_kernel(
    const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ x,
    const __nv_bfloat16* __restrict__ y, const float* __restrict__ mean, const float* __restrict__ rstd,
    float* __restrict__ partials, int64_t rows, int C, int relu) {
    __shared__ float sm[BN_THREADS][17];
    const int cv = C >> 3;
    const int lanes_r = BN_THREADS / cv;
    const int v = threadIdx.x % cv, rl = threadIdx.x / cv;
    float s0[8] = {0, 0, 0, 0, 0, 0, 0, 0}, s1[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    float mu[8], rs[8];
    if (MODE == 1) {
#pragma unroll
        for (int k = 0; k < 8; ++k) { mu[k] = mean[v * 8 + k]; rs[k] = rstd[v * 8 + k]; }
    }
    if (rl < lanes_r) {
#pragma unroll 4
        for (int64_t row = (int64_t)blockIdx.x * lanes_r + rl; row < rows; row += (int64_t)gridDim.x * lanes_r) {
            const int64_t idx = row * cv + v;
            float f[8];
            bn_unpack8(reinterpret_cast<const uint4*>(a)[idx], f);
            if (MODE == 0) {
#pragma unroll
                for (int k = 0; k < 8; ++k) { s0[k] += f[k]; s1[k] += f[k] * f[k]; }
            } else {
                float xv[8];
                bn_unpack8(reinterpret_cast<const uint4*>(x)[idx], xv);
                if (relu) {
                    float yv[8];
                    bn_unpack8(reinterpret_cast<const uint4*>(y)[idx], yv);
#pragma unroll
                    for (int k = 0; k < 8; ++k) f[k] = yv[k] > 0.f ? f[k] : 0.f;
                }
#pragma unroll
                for (int k = 0; k < 8; ++k) { s0[k] += f[k]; s1[k] += f[k] * (xv[k] - mu[k]) * rs[k]; }
            }
        }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) { sm[threadIdx.x][k] = s0[k]; sm[threadIdx.x][8 + k] = s1[k]; }
    __syncthreads();
    if (rl == 0) {
        float* p = partials + (int64_t)blockIdx.x * 2 * C;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            float t0 = 0.f, t1 = 0.f;
            for (int r = 0; r < lanes_r; ++r) { t0 += sm[r * cv + v][k]; t1 += sm[r * cv + v][8 + k]; }
            p[v * 8 + k] = t0;
            p[C + v * 8 + k] = t1;
        }
    }
}

// Sums the per-block partials of both quantities for 32 channels per block: thread (x, y) adds
// partial rows y, y+32, ... (coalesced; up to 592 rows, so 32 row groups keep each thread's chain
// of dependent L2 round trips short), the groups are combined in a fixed order in smem.
constexpr int BN_FIN_THREADS = 1024;
__device__ __forceinline__ bool bn_sum_partials(const float* __restrict__ partials, int nparts, int C, int c,
                                                float& s, float& q) {
    __shared__ float sm[2][32][33];
    const int x = threadIdx.x & 31, y = threadIdx.x >> 5;
    float a0 = 0.f, a1 = 0.f;
    if (c < C) {
#pragma unroll 4
        for (int p = y; p < nparts; p += 32) {
            a0 += partials[(int64_t)p * 2 * C + c];
            a1 += partials[(int64_t)p * 2 * C + C + c];
        }
    }
    sm[0][y][x] = a0;
    sm[1][y][x] = a1;
    __syncthreads();
    s = 0.f; q = 0.f;
    if (y != 0 || c >= C) return false;
#pragma unroll
    for (int k = 0; k < 32; ++k) { s += sm[0][k][x]; q += sm[1][k][x]; }
    return true;
}

// Forward finalize: mean / rstd from the partial sums, running statistics update.
__global__ void __launch_bounds__(BN_FIN_THREADS) bn_finalize_fwd_kernel(const float* __restrict__ partials, int nparts, int C,
                                                              int64_t rows, float eps, float momentum,
                                                              float* __restrict__ mean, float* __restrict__ rstd,
                                                              float* running_mean, float* running_var) {
    const int c = blockIdx.x * 32 + (threadIdx.x & 31);
    float s, q;
    if (!bn_sum_partials(partials, nparts, C, c, s, q)) return;
    const float m = s / (float)rows;
    const float var = fmaxf(q / (float)rows - m * m, 0.f);
    mean[c] = m;
    rstd[c] = rsqrtf(var + eps);
    if (running_mean) {
        const float unbiased = rows > 1 ? var * (float)rows / (float)(rows - 1) : var;
        running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * m;
        running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
    }
}

// Backward finalize: dbeta = sum g, dgamma = sum g*xhat (bf16 or fp32 out) + fp32 copies for dx.
__global__ void __launch_bounds__(BN_FIN_THREADS) bn_finalize_bwd_kernel(const float* __restrict__ partials, int nparts, int C,
                                                              float* __restrict__ sums, void* dgamma, void* dbeta,
                                                              int grads_bf16) {
    const int c = blockIdx.x * 32 + (threadIdx.x & 31);
    float s, q;
    if (!bn_sum_partials(partials, nparts, C, c, s, q)) return;
    sums[c] = s;
    sums[C + c] = q;
    if (grads_bf16) {
        reinterpret_cast<__nv_bfloat16*>(dbeta)[c] = __float2bfloat16_rn(s);
        reinterpret_cast<__nv_bfloat16*>(dgamma)[c] = __float2bfloat16_rn(q);
    } else {
        reinterpret_cast<float*>(dbeta)[c] = s;
        reinterpret_cast<float*>(dgamma)[c] = q;
    }
}

// y = relu?( (x - mean) * rstd * gamma + beta (+ residual) ).  In inference `rstd` already holds
// rsqrt(running_var + eps) and `mean` the running mean.
// The grid stride is a multiple of C/8 (C is a power of two <= 2048, the stride a multiple of
// 256), so every thread stays on the same 8 channels: their scale/shift live in registers.
__global__ void __launch_bounds__(256) bn_apply_kernel(const __nv_bfloat16* __restrict__ x,
                                                       const __nv_bfloat16* __restrict__ residual,
                                                       __nv_bfloat16* __restrict__ y, const float* __restrict__ mean,
                                                       const float* __restrict__ rstd, const float* __restrict__ gamma,
                                                       const float* __restrict__ beta, int64_t total_vecs, int C, int relu) {
    const int cv = C >> 3;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t first = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int c0 = (int)(first % cv) * 8;
    float sc[8], sh[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        sc[k] = rstd[c0 + k] * gamma[c0 + k];
        sh[k] = beta[c0 + k] - mean[c0 + k] * sc[k];
    }
    for (int64_t i = first; i < total_vecs; i += stride) {
        float f[8];
        bn_unpack8(reinterpret_cast<const uint4*>(x)[i], f);
#pragma unroll
        for (int k = 0; k < 8; ++k) f[k] = fmaf(f[k], sc[k], sh[k]);
        if (residual) {
            float r[8];
            bn_unpack8(reinterpret_cast<const uint4*>(residual)[i], r);
#pragma unroll
            for (int k = 0; k < 8; ++k) f[k] += r[k];
        }
        if (relu) {
#pragma unroll
            for (int k = 0; k < 8; ++k) f[k] = fmaxf(f[k], 0.f);
        }
        reinterpret_cast<uint4*>(y)[i] = bn_pack8(f);
    }
}

// dx = gamma * rstd * (g - sum_g/rows - xhat * sum_gxhat/rows), g = dy * [y > 0] when relu;
// dres (optional) = g: the gradient of the fused residual branch.  Per-channel terms are
// folded into three register constants per channel (same channel-invariance as above):
//   dx = a * g + b2 * x + c,  a = gamma*rstd,  b2 = -a*rstd*q,  c = a*(rstd*mean*q - s)
// with s = sum_g/rows, q = sum_gxhat/rows.
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(
    const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ y,
    const float* __restrict__ mean, const float* __restrict__ rstd, const float* __restrict__ gamma,
    const float* __restrict__ sums, __nv_bfloat16* __restrict__ dx, __nv_bfloat16* __restrict__ dres,
    int64_t total_vecs, int C, float inv_rows, int relu) {
    const int cv = C >> 3;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t first = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int c0 = (int)(first % cv) * 8;
    float ka[8], kb[8], kc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const float rs = rstd[c0 + k], mu = mean[c0 + k];
        const float sg = sums[c0 + k] * inv_rows, sq = sums[C + c0 + k] * inv_rows;
        ka[k] = gamma[c0 + k] * rs;
        kb[k] = -ka[k] * rs * sq;
        kc[k] = ka[k] * (rs * mu * sq - sg);
    }
    for (int64_t i = first; i < total_vecs; i += stride) {
        float g[8], xv[8], o[8];
        bn_unpack8(reinterpret_cast<const uint4*>(dy)[i], g);
        bn_unpack8(reinterpret_cast<const uint4*>(x)[i], xv);
        if (relu) {
            float yv[8];
            bn_unpack8(reinterpret_cast<const uint4*>(y)[i], yv);
#pragma unroll
            for (int k = 0; k < 8; ++k) g[k] = yv[k] > 0.f ? g[k] : 0.f;
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = fmaf(ka[k], g[k], fmaf(kb[k], xv[k], kc[k]));
        reinterpret_cast<uint4*>(dx)[i] = bn_pack8(o);
        if (dres) reinterpret_cast<uint4*>(dres)[i] = bn_pack8(g);
    }
}

static int bn_check(int64_t rows, int C) {
    B200_REQUIRE(rows > 0, "batchnorm: no rows");
    B200_REQUIRE(C >= 8 && C <= 2048 && (C & (C - 1)) == 0, "batchnorm: channels must be a power of two in [8, 2048] (got %d)", C);
    return 0;
}
static int bn_reduce_grid(int64_t rows, int C) {
    const int lanes_r = BN_THREADS / (C >> 3);
    int64_t b = (rows + lanes_r * 8 - 1) / (lanes_r * 8);
    if (b > BN_MAX_BLOCKS) b = BN_MAX_BLOCKS;
    return (int)(b < 1 ? 1 : b);
}
static int bn_ew_grid(int64_t vecs) {
    int64_t b = (vecs + 255) / 256;
    if (b > 148 * 16) b = 148 * 16;
    return (int)(b < 1 ? 1 : b);
}

}  // namespace b200

extern "C" int64_t b200_batchnorm_scratch_bytes(int C) {
    return (int64_t)b200::BN_MAX_BLOCKS * 2 * C * (int64_t)sizeof(float) + 2 * (int64_t)C * sizeof(float);
}

extern "C" int b200_batchnorm_fwd(const void* x, const void* residual, void* y, const float* gamma,
                                  const float* beta, float* mean, float* rstd, float* running_mean,
                                  float* running_var, int64_t rows, int C, float eps, float momentum,
                                  int training, int relu, void* scratch, void* stream) {
    using namespace b200;
    B200_REQUIRE(x && y && gamma && beta && mean && rstd, "batchnorm_fwd: null pointer");
    if (int rc = bn_check(rows, C)) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    if (training) {
        B200_REQUIRE(scratch, "batchnorm_fwd: training needs scratch");
        const int grid = bn_reduce_grid(rows, C);
        bn_reduce_kernel<0><<<grid, BN_THREADS, 0, s>>>((const __nv_bfloat16*)x, nullptr, nullptr, nullptr, nullptr,
                                                        (float*)scratch, rows, C, 0);
        bn_finalize_fwd_kernel<<<(C + 31) / 32, BN_FIN_THREADS, 0, s>>>((const float*)scratch, grid, C, rows, eps, momentum,
                                                               mean, rstd, running_mean, running_var);
        b200_count_launch(2);
    }
    const int64_t vecs = rows * (C >> 3);
    bn_apply_kernel<<<bn_ew_grid(vecs), 256, 0, s>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)residual,
                                                     (__nv_bfloat16*)y, mean, rstd, gamma, beta, vecs, C, relu);
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int b200_batchnorm_bwd(const void* dy, const void* x, const void* y, const float* mean, const float* rstd,
                                  const float* gamma, void* dx, void* dres, void* dgamma, void* dbeta,
                                  int grads_bf16, int64_t rows, int C, int relu, void* scratch, void* stream) {
    using namespace b200;
    B200_REQUIRE(dy && x && mean && rstd && gamma && dx && dgamma && dbeta && scratch, "batchnorm_bwd: null pointer");
    B200_REQUIRE(!relu || y, "batchnorm_bwd: the ReLU mask needs the forward output");
    if (int rc = bn_check(rows, C)) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const int grid = bn_reduce_grid(rows, C);
    float* partials = (float*)scratch;
    float* sums = partials + (int64_t)BN_MAX_BLOCKS * 2 * C;
    bn_reduce_kernel<1><<<grid, BN_THREADS, 0, s>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)x,
                                                    (const __nv_bfloat16*)y, mean, rstd, partials, rows, C, relu);
    bn_finalize_bwd_kernel<<<(C + 31) / 32, BN_FIN_THREADS, 0, s>>>(partials, grid, C, sums, dgamma, dbeta, grads_bf16);
    const int64_t vecs = rows * (C >> 3);
    bn_bwd_apply_kernel<<<bn_ew_grid(vecs), 256, 0, s>>>(
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (const __nv_bfloat16*)y, mean, rstd, gamma, sums,
        (__nv_bfloat16*)dx, (__nv_bfloat16*)dres, vecs, C, 1.f / (float)rows, relu);
    b200_count_launch(3);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}

// Inference helper: rstd[c] = rsqrt(running_var[c] + eps)
__global__ void bn_rstd_kernel(const float* var, float* rstd, int C, float eps) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < C) rstd[c] = rsqrtf(var[c] + eps);
}
extern "C" int b200_batchnorm_eval_rstd(const float* running_var, float* rstd, int C, float eps, void* stream) {
    B200_REQUIRE(running_var && rstd && C > 0, "batchnorm_eval_rstd: bad arguments");
    bn_rstd_kernel<<<(C + 255) / 256, 256, 0, (cudaStream_t)stream>>>(running_var, rstd, C, eps);
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}
