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

__device__ __forceinline__ void bn_load8(const float* __restrict__ p, float (&f)[8]) {
    const float4 a = reinterpret_cast<const float4*>(p)[0], b = reinterpret_cast<const float4*>(p)[1];
    f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w;
    f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}
// g *= [y > 0]: from the one-byte-per-8-channels mask the forward pass wrote, else from y itself.
__device__ __forceinline__ void bn_relu_gate(float (&g)[8], const __nv_bfloat16* __restrict__ y,
                                             const uint8_t* __restrict__ mask, int64_t idx) {
    if (mask) {
        const uint32_t m = mask[idx];
#pragma unroll
        for (int k = 0; k < 8; ++k) g[k] = (m >> k) & 1u ? g[k] : 0.f;
    } else {
        float yv[8];
        bn_unpack8(reinterpret_cast<const uint4*>(y)[idx], yv);
#pragma unroll
        for (int k = 0; k < 8; ++k) g[k] = yv[k] > 0.f ? g[k] : 0.f;
    }
}

// Per-block partial sums of two per-channel quantities.
//   MODE 0 (forward):  (x, x^2)
//   MODE 1 (backward): (g, g * xhat) with g = dy * [y > 0] when relu, xhat = (x - mean) * rstd
// partials layout: [gridDim.x][2][C].
template <int MODE>
__global__ void __launch_bounds__(BN_THREADS) bn_reduce_kernel(
    const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ x,
    const __nv_bfloat16* __restrict__ y, const uint8_t* __restrict__ mask, const float* __restrict__ mean,
    const float* __restrict__ rstd, float* __restrict__ partials, int64_t rows, int C, int relu) {
    __shared__ float sm[BN_THREADS][17];
    const int cv = C >> 3;
    const int lanes_r = BN_THREADS / cv;
    const int v = threadIdx.x % cv, rl = threadIdx.x / cv;
    float s0[8] = {0, 0, 0, 0, 0, 0, 0, 0}, s1[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    float mu[8], rs[8];
    if (MODE == 1) {
        bn_load8(mean + v * 8, mu);
        bn_load8(rstd + v * 8, rs);
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
                if (relu) bn_relu_gate(f, y, mask, idx);
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

// Sums the per-block partials of both quantities for 8 channels per block: thread (x, y) adds
// partial rows y, y+128, ... of channel x (32-byte segments; up to 592 rows, so every thread's
// chain is at most five dependent L2 round trips and C/8 blocks share the work — the first version
// used 32 channels x 32 row groups per block: two blocks for a 64-channel layer, 5.4 us per call,
// 0.56 ms per ResNet-50 step), then the 128 groups are combined in a fixed order in two levels.
constexpr int BN_FIN_THREADS = 1024;
constexpr int BN_FIN_CH = 8;
__device__ __forceinline__ bool bn_sum_partials(const float* __restrict__ partials, int nparts, int C, int c,
                                                float& s, float& q) {
    __shared__ float sm[2][128][BN_FIN_CH + 1];
    __shared__ float sm2[2][16][BN_FIN_CH + 1];
    const int x = threadIdx.x & (BN_FIN_CH - 1), y = threadIdx.x >> 3;
    float a0 = 0.f, a1 = 0.f;
    if (c < C) {
#pragma unroll 5
        for (int p = y; p < nparts; p += 128) {
            a0 += partials[(int64_t)p * 2 * C + c];
            a1 += partials[(int64_t)p * 2 * C + C + c];
        }
    }
    sm[0][y][x] = a0;
    sm[1][y][x] = a1;
    __syncthreads();
    if (y < 16) {
        float t0 = 0.f, t1 = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) { t0 += sm[0][y * 8 + k][x]; t1 += sm[1][y * 8 + k][x]; }
        sm2[0][y][x] = t0;
        sm2[1][y][x] = t1;
    }
    __syncthreads();
    s = 0.f; q = 0.f;
    if (y != 0 || c >= C) return false;
#pragma unroll
    for (int k = 0; k < 16; ++k) { s += sm2[0][k][x]; q += sm2[1][k][x]; }
    return true;
}

// Forward finalize: mean / rstd from the partial sums, running statistics update, and the two
// per-channel constants of the apply pass: coef = [scale = rstd*gamma | shift = beta - mean*scale].
__global__ void __launch_bounds__(BN_FIN_THREADS) bn_finalize_fwd_kernel(const float* __restrict__ partials, int nparts, int C,
                                                              int64_t rows, float eps, float momentum,
                                                              const float* __restrict__ gamma, const float* __restrict__ beta,
                                                              float* __restrict__ mean, float* __restrict__ rstd,
                                                              float* running_mean, float* running_var,
                                                              float* __restrict__ coef) {
    const int c = blockIdx.x * BN_FIN_CH + (threadIdx.x & (BN_FIN_CH - 1));
    float s, q;
    if (!bn_sum_partials(partials, nparts, C, c, s, q)) return;
    const float m = s / (float)rows;
    const float var = fmaxf(q / (float)rows - m * m, 0.f);
    const float r = rsqrtf(var + eps);
    mean[c] = m;
    rstd[c] = r;
    const float sc = r * gamma[c];
    coef[c] = sc;
    coef[C + c] = beta[c] - m * sc;
    if (running_mean) {
        const float unbiased = rows > 1 ? var * (float)rows / (float)(rows - 1) : var;
        running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * m;
        running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
    }
}
// Inference: the same two constants from statistics that already exist.
__global__ void bn_coef_kernel(const float* __restrict__ mean, const float* __restrict__ rstd,
                               const float* __restrict__ gamma, const float* __restrict__ beta,
                               float* __restrict__ coef, int C) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const float sc = rstd[c] * gamma[c];
    coef[c] = sc;
    coef[C + c] = beta[c] - mean[c] * sc;
}

// Backward finalize: dbeta = sum g, dgamma = sum g*xhat (bf16 or fp32 out), and the three
// per-channel constants of the dx pass (see bn_bwd_apply_kernel): coef = [a | b2 | c].
__global__ void __launch_bounds__(BN_FIN_THREADS) bn_finalize_bwd_kernel(const float* __restrict__ partials, int nparts, int C,
                                                              const float* __restrict__ mean, const float* __restrict__ rstd,
                                                              const float* __restrict__ gamma, float inv_rows,
                                                              float* __restrict__ coef, void* dgamma, void* dbeta,
                                                              int grads_bf16) {
    const int c = blockIdx.x * BN_FIN_CH + (threadIdx.x & (BN_FIN_CH - 1));
    float s, q;
    if (!bn_sum_partials(partials, nparts, C, c, s, q)) return;
    const float rs = rstd[c], mu = mean[c];
    const float sg = s * inv_rows, sq = q * inv_rows;
    const float ka = gamma[c] * rs;
    coef[c] = ka;
    coef[C + c] = -ka * rs * sq;
    coef[2 * C + c] = ka * (rs * mu * sq - sg);
    if (grads_bf16) {
        reinterpret_cast<__nv_bfloat16*>(dbeta)[c] = __float2bfloat16_rn(s);
        reinterpret_cast<__nv_bfloat16*>(dgamma)[c] = __float2bfloat16_rn(q);
    } else {
        reinterpret_cast<float*>(dbeta)[c] = s;
        reinterpret_cast<float*>(dgamma)[c] = q;
    }
}

// y = relu?( x * scale + shift (+ residual) ), and with `mask` one byte per 8 channels holding
// [y > 0] for the backward pass (1/16 of the bytes of y).
// The grid stride is a multiple of C/8 (C is a power of two <= 2048, the stride a multiple of
// 256), so every thread stays on the same 8 channels: their scale/shift live in registers.
// Two vectors per thread are in flight per trip.
__global__ void __launch_bounds__(256) bn_apply_kernel(const __nv_bfloat16* __restrict__ x,
                                                       const __nv_bfloat16* __restrict__ residual,
                                                       __nv_bfloat16* __restrict__ y, uint8_t* __restrict__ mask,
                                                       const float* __restrict__ coef, int64_t total_vecs, int C,
                                                       int relu) {
    const int cv = C >> 3;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t first = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int c0 = (int)(first % cv) * 8;
    float sc[8], sh[8];
    bn_load8(coef + c0, sc);
    bn_load8(coef + C + c0, sh);
    for (int64_t i = first; i < total_vecs; i += 2 * stride) {
        const int64_t j = i + stride;
        const bool two = j < total_vecs;
        const uint4 z = make_uint4(0, 0, 0, 0);
        const uint4 xa = __ldcs(reinterpret_cast<const uint4*>(x) + i);
        const uint4 xb = two ? __ldcs(reinterpret_cast<const uint4*>(x) + j) : z;
        uint4 ra = z, rb = z;
        if (residual) {
            ra = __ldcs(reinterpret_cast<const uint4*>(residual) + i);
            if (two) rb = __ldcs(reinterpret_cast<const uint4*>(residual) + j);
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (h == 1 && !two) break;
            float f[8], r[8];
            bn_unpack8(h ? xb : xa, f);
            bn_unpack8(h ? rb : ra, r);
#pragma unroll
            for (int k = 0; k < 8; ++k) f[k] = fmaf(f[k], sc[k], sh[k]) + r[k];
            if (relu) {
                uint32_t m = 0;
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    m |= f[k] > 0.f ? 1u << k : 0u;
                    f[k] = fmaxf(f[k], 0.f);
                }
                if (mask) mask[h ? j : i] = (uint8_t)m;
            }
            reinterpret_cast<uint4*>(y)[h ? j : i] = bn_pack8(f);
        }
    }
}

// dx = gamma * rstd * (g - sum_g/rows - xhat * sum_gxhat/rows), g = dy * [y > 0] when relu;
// dres (optional) = g: the gradient of the fused residual branch.  Per-channel terms are
// folded into three constants per channel by the finalize kernel:
//   dx = a * g + b2 * x + c,  a = gamma*rstd,  b2 = -a*rstd*q,  c = a*(rstd*mean*q - s)
// with s = sum_g/rows, q = sum_gxhat/rows.
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(
    const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ y,
    const uint8_t* __restrict__ mask, const float* __restrict__ coef, __nv_bfloat16* __restrict__ dx,
    __nv_bfloat16* __restrict__ dres, int64_t total_vecs, int C, int relu) {
    const int cv = C >> 3;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t first = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int c0 = (int)(first % cv) * 8;
    float ka[8], kb[8], kc[8];
    bn_load8(coef + c0, ka);
    bn_load8(coef + C + c0, kb);
    bn_load8(coef + 2 * C + c0, kc);
    for (int64_t i = first; i < total_vecs; i += 2 * stride) {
        const int64_t j = i + stride;
        const bool two = j < total_vecs;
        const uint4 z = make_uint4(0, 0, 0, 0);
        const uint4 ga = __ldcs(reinterpret_cast<const uint4*>(dy) + i);
        const uint4 xa = __ldcs(reinterpret_cast<const uint4*>(x) + i);
        const uint4 gb = two ? __ldcs(reinterpret_cast<const uint4*>(dy) + j) : z;
        const uint4 xb = two ? __ldcs(reinterpret_cast<const uint4*>(x) + j) : z;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (h == 1 && !two) break;
            float g[8], xv[8], o[8];
            bn_unpack8(h ? gb : ga, g);
            bn_unpack8(h ? xb : xa, xv);
            if (relu) bn_relu_gate(g, y, mask, h ? j : i);
#pragma unroll
            for (int k = 0; k < 8; ++k) o[k] = fmaf(ka[k], g[k], fmaf(kb[k], xv[k], kc[k]));
            reinterpret_cast<uint4*>(dx)[h ? j : i] = bn_pack8(o);
            if (dres) reinterpret_cast<uint4*>(dres)[h ? j : i] = bn_pack8(g);
        }
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
    return (int64_t)b200::BN_MAX_BLOCKS * 2 * C * (int64_t)sizeof(float) + 3 * (int64_t)C * sizeof(float);
}

extern "C" int b200_batchnorm_fwd(const void* x, const void* residual, void* y, void* relu_mask, const float* gamma,
                                  const float* beta, float* mean, float* rstd, float* running_mean,
                                  float* running_var, int64_t rows, int C, float eps, float momentum,
                                  int training, int relu, void* scratch, void* stream) {
    using namespace b200;
    B200_REQUIRE(x && y && gamma && beta && mean && rstd, "batchnorm_fwd: null pointer");
    if (int rc = bn_check(rows, C)) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    B200_REQUIRE(scratch, "batchnorm_fwd: needs scratch");
    float* coef = (float*)scratch + (int64_t)BN_MAX_BLOCKS * 2 * C;
    if (training) {
        const int grid = bn_reduce_grid(rows, C);
        bn_reduce_kernel<0><<<grid, BN_THREADS, 0, s>>>((const __nv_bfloat16*)x, nullptr, nullptr, nullptr, nullptr,
                                                        nullptr, (float*)scratch, rows, C, 0);
        bn_finalize_fwd_kernel<<<(C + BN_FIN_CH - 1) / BN_FIN_CH, BN_FIN_THREADS, 0, s>>>((const float*)scratch, grid, C, rows, eps, momentum,
                                                               gamma, beta, mean, rstd, running_mean, running_var, coef);
        b200_count_launch(2);
    } else {
        bn_coef_kernel<<<(C + 255) / 256, 256, 0, s>>>(mean, rstd, gamma, beta, coef, C);
        b200_count_launch(1);
    }
    const int64_t vecs = rows * (C >> 3);
    bn_apply_kernel<<<bn_ew_grid(vecs), 256, 0, s>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)residual,
                                                     (__nv_bfloat16*)y, (uint8_t*)relu_mask, coef, vecs, C, relu);
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int b200_batchnorm_bwd(const void* dy, const void* x, const void* y, const void* relu_mask,
                                  const float* mean, const float* rstd,
                                  const float* gamma, void* dx, void* dres, void* dgamma, void* dbeta,
                                  int grads_bf16, int64_t rows, int C, int relu, void* scratch, void* stream) {
    using namespace b200;
    B200_REQUIRE(dy && x && mean && rstd && gamma && dx && dgamma && dbeta && scratch, "batchnorm_bwd: null pointer");
    B200_REQUIRE(!relu || y || relu_mask, "batchnorm_bwd: the ReLU gate needs the forward output or its mask");
    if (int rc = bn_check(rows, C)) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const int grid = bn_reduce_grid(rows, C);
    float* partials = (float*)scratch;
    float* coef = partials + (int64_t)BN_MAX_BLOCKS * 2 * C;
    bn_reduce_kernel<1><<<grid, BN_THREADS, 0, s>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)x,
                                                    (const __nv_bfloat16*)y, (const uint8_t*)relu_mask, mean, rstd,
                                                    partials, rows, C, relu);
    bn_finalize_bwd_kernel<<<(C + BN_FIN_CH - 1) / BN_FIN_CH, BN_FIN_THREADS, 0, s>>>(partials, grid, C, mean, rstd, gamma,
                                                                    1.f / (float)rows, coef, dgamma, dbeta, grads_bf16);
    const int64_t vecs = rows * (C >> 3);
    bn_bwd_apply_kernel<<<bn_ew_grid(vecs), 256, 0, s>>>(
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (const __nv_bfloat16*)y, (const uint8_t*)relu_mask, coef,
        (__nv_bfloat16*)dx, (__nv_bfloat16*)dres, vecs, C, relu);
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
