// Max pooling and global average pooling on channels-last (NHWC) bf16 activations.
//
// For BASELINE.json config 3 (ResNet-50: 3x3 stride-2 max pool after the stem, global average pool
// before the classifier); the reference has neither — its ResNet-50 baseline runs ATen's NHWC
// pooling kernels.  Both directions are gathers with 16-byte vectors along channels: the forward
// records, per channel, which tap won (first maximum in scan order, as torch does); the backward
// visits, for each input pixel, the few windows that contain it and adds the gradients of those
// whose winner is this pixel — no atomics, deterministic.
#include "common.cuh"

void b200_count_launch(int n);

namespace b200 {

struct PoolGeom {
    int N, H, W, C, R, S, P, Q, stride, pad;
};

__device__ __forceinline__ void pool_unpack8(const uint4& raw, float (&f)[8]) {
    const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const float2 t = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[q]));
        f[2 * q] = t.x;
        f[2 * q + 1] = t.y;
    }
}
__device__ __forceinline__ uint4 pool_pack8(const float (&f)[8]) {
    return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}

// y[n,p,q,c] = max over the window; idx[n,p,q,c] = r*S + s of the first maximum.
__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                                          uint8_t* __restrict__ idx, PoolGeom g, int64_t total) {
    const int cv = g.C >> 3;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        int64_t t = i;
        const int c8 = (int)(t % cv); t /= cv;
        const int q = (int)(t % g.Q); t /= g.Q;
        const int p = (int)(t % g.P); t /= g.P;
        const int n = (int)t;
        float best[8];
        uint32_t arg[8];
        bool first = true;
#pragma unroll
        for (int k = 0; k < 8; ++k) { best[k] = -INFINITY; arg[k] = 0; }
        for (int r = 0; r < g.R; ++r) {
            const int h = p * g.stride - g.pad + r;
            if (h < 0 || h >= g.H) continue;
            for (int s = 0; s < g.S; ++s) {
                const int w = q * g.stride - g.pad + s;
                if (w < 0 || w >= g.W) continue;
                float v[8];
                pool_unpack8(reinterpret_cast<const uint4*>(x)[(((int64_t)n * g.H + h) * g.W + w) * cv + c8], v);
                const uint32_t tap = (uint32_t)(r * g.S + s);
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    if (first || v[k] > best[k]) { best[k] = v[k]; arg[k] = tap; }
                first = false;
            }
        }
        reinterpret_cast<uint4*>(y)[i] = pool_pack8(best);
        uint2 packed;
        packed.x = arg[0] | (arg[1] << 8) | (arg[2] << 16) | (arg[3] << 24);
        packed.y = arg[4] | (arg[5] << 8) | (arg[6] << 16) | (arg[7] << 24);
        reinterpret_cast<uint2*>(idx)[i] = packed;
    }
}

// dx[n,h,w,c] = sum of dy over the windows (p,q) that contain (h,w) and whose winner is (h,w).
__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const uint8_t* __restrict__ idx,
                                                          __nv_bfloat16* __restrict__ dx, PoolGeom g, int64_t total) {
    const int cv = g.C >> 3;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        int64_t t = i;
        const int c8 = (int)(t % cv); t /= cv;
        const int w = (int)(t % g.W); t /= g.W;
        const int h = (int)(t % g.H); t /= g.H;
        const int n = (int)t;
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int r = 0; r < g.R; ++r) {
            const int hp = h + g.pad - r;
            if (hp < 0 || hp % g.stride) continue;
            const int p = hp / g.stride;
            if (p >= g.P) continue;
            for (int s = 0; s < g.S; ++s) {
                const int wq = w + g.pad - s;
                if (wq < 0 || wq % g.stride) continue;
                const int q = wq / g.stride;
                if (q >= g.Q) continue;
                const int64_t o = (((int64_t)n * g.P + p) * g.Q + q) * cv + c8;
                const uint2 packed = reinterpret_cast<const uint2*>(idx)[o];
                float v[8];
                pool_unpack8(reinterpret_cast<const uint4*>(dy)[o], v);
                const uint32_t tap = (uint32_t)(r * g.S + s);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const uint32_t a = ((k < 4 ? packed.x : packed.y) >> (8 * (k & 3))) & 0xffu;
                    if (a == tap) acc[k] += v[k];
                }
            }
        }
        reinterpret_cast<uint4*>(dx)[i] = pool_pack8(acc);
    }
}

// y[n,c] = mean over the HW pixels of x[n,:,:,c]
__global__ void __launch_bounds__(256) avgpool_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                                          int N, int HW, int C) {
    const int cv = C >> 3;
    const int64_t total = (int64_t)N * cv;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int c8 = (int)(i % cv);
        const int n = (int)(i / cv);
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int px = 0; px < HW; ++px) {
            float v[8];
            pool_unpack8(reinterpret_cast<const uint4*>(x)[((int64_t)n * HW + px) * cv + c8], v);
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] += v[k];
        }
        const float inv = 1.f / (float)HW;
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] *= inv;
        reinterpret_cast<uint4*>(y)[i] = pool_pack8(acc);
    }
}

// dx[n,h,w,c] = dy[n,c] / HW
__global__ void __launch_bounds__(256) avgpool_bwd_kernel(const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ dx,
                                                          int N, int HW, int C) {
    const int cv = C >> 3;
    const int64_t total = (int64_t)N * HW * cv;
    const float inv = 1.f / (float)HW;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int c8 = (int)(i % cv);
        const int n = (int)(i / ((int64_t)HW * cv));
        float v[8];
        pool_unpack8(reinterpret_cast<const uint4*>(dy)[(int64_t)n * cv + c8], v);
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] *= inv;
        reinterpret_cast<uint4*>(dx)[i] = pool_pack8(v);
    }
}

static int pool_grid(int64_t items) {
    int64_t b = (items + 255) / 256;
    if (b > 148 * 16) b = 148 * 16;
    return (int)(b < 1 ? 1 : b);
}

static int pool_geom(PoolGeom* g, int N, int H, int W, int C, int R, int S, int stride, int pad) {
    B200_REQUIRE(N > 0 && H > 0 && W > 0 && C > 0 && R > 0 && S > 0 && stride > 0 && pad >= 0, "pool: bad geometry");
    B200_REQUIRE(C % 8 == 0, "pool: channels must be a multiple of 8 (got %d)", C);
    B200_REQUIRE(R * S <= 255 && 2 * pad <= R && 2 * pad <= S, "pool: window too large or padding above half the window");
    g->N = N; g->H = H; g->W = W; g->C = C; g->R = R; g->S = S; g->stride = stride; g->pad = pad;
    g->P = (H + 2 * pad - R) / stride + 1;
    g->Q = (W + 2 * pad - S) / stride + 1;
    B200_REQUIRE(g->P > 0 && g->Q > 0, "pool: empty output");
    return 0;
}

}  // namespace b200

extern "C" int b200_maxpool_fwd(const void* x, void* y, void* idx, int N, int H, int W, int C, int R, int S,
                                int stride, int pad, void* stream) {
    using namespace b200;
    B200_REQUIRE(x && y && idx, "maxpool_fwd: null pointer");
    PoolGeom g;
    if (int rc = pool_geom(&g, N, H, W, C, R, S, stride, pad)) return rc;
    const int64_t total = (int64_t)N * g.P * g.Q * (C / 8);
    maxpool_fwd_kernel<<<pool_grid(total), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y,
                                                                         (uint8_t*)idx, g, total);
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int b200_maxpool_bwd(const void* dy, const void* idx, void* dx, int N, int H, int W, int C, int R, int S,
                                int stride, int pad, void* stream) {
    using namespace b200;
    B200_REQUIRE(dy && idx && dx, "maxpool_bwd: null pointer");
    PoolGeom g;
    if (int rc = pool_geom(&g, N, H, W, C, R, S, stride, pad)) return rc;
    const int64_t total = (int64_t)N * H * W * (C / 8);
    maxpool_bwd_kernel<<<pool_grid(total), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)dy, (const uint8_t*)idx,
                                                                         (__nv_bfloat16*)dx, g, total);
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int b200_avgpool_fwd(const void* x, void* y, int N, int HW, int C, void* stream) {
    using namespace b200;
    B200_REQUIRE(x && y && N > 0 && HW > 0 && C > 0 && C % 8 == 0, "avgpool_fwd: bad arguments");
    avgpool_fwd_kernel<<<pool_grid((int64_t)N * (C / 8)), 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)x, (__nv_bfloat16*)y, N, HW, C);
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int b200_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, void* stream) {
    using namespace b200;
    B200_REQUIRE(dy && dx && N > 0 && HW > 0 && C > 0 && C % 8 == 0, "avgpool_bwd: bad arguments");
    avgpool_bwd_kernel<<<pool_grid((int64_t)N * HW * (C / 8)), 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)dy, (__nv_bfloat16*)dx, N, HW, C);
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}
