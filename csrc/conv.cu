// Convolution as GEMM on the tcgen05 kernel: im2col / col2im for NHWC bf16 tensors.
//
// For BASELINE.json config 3 (ResNet-50).  The reference has no convolution; its
// ResNet-50 baseline calls cuDNN through ATen.  Here a convolution is
//
//   forward   Y[(n,p,q), k]       = col(X)[(n,p,q), (r,s,c)] * W[k, (r,s,c)]^T      (GEMM, K-major)
//   dgrad     dcol[(n,p,q),(r,s,c)] = dY[(n,p,q), k] * W[k, (r,s,c)]   (GEMM, B MN-major) -> col2im
//   wgrad     dW[k, (r,s,c)]      = dY^T * col(X)                     (GEMM, both MN-major)
//
// with 1x1 stride-1 convolutions skipping the gather entirely (the NHWC tensor IS the
// GEMM operand).  im2col is a pure gather (16-byte vectors along channels, zero fill at
// the borders); col2im is written as a gather over the taps that touch each input
// pixel, so it needs no atomics and is deterministic.
#include "common.cuh"

void b200_count_launch(int n);

namespace b200 {

struct ConvGeom {
    int N, H, W, C;      // input NHWC
    int R, S;            // filter
    int P, Q;            // output spatial
    int stride, pad;
};

// col[(n,p,q), (r,s,c)].  One thread owns 8 channels (16 B) of one output pixel and walks the
// R*S taps: the pixel decomposition (three integer divisions) is paid once per R*S vectors, and
// the taps' loads are independent, so several are in flight per thread.
__global__ void __launch_bounds__(256) im2col_kernel(const __nv_bfloat16* __restrict__ x,
                                                     __nv_bfloat16* __restrict__ col, ConvGeom g,
                                                     int64_t total_items) {
    const int cv = g.C >> 3;  // vectors per pixel
    const int taps = g.R * g.S;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total_items; i += stride) {
        int64_t t = i;
        const int c8 = (int)(t % cv); t /= cv;
        const int64_t pix = t;
        const int q = (int)(t % g.Q); t /= g.Q;
        const int p = (int)(t % g.P); t /= g.P;
        const int n = (int)t;
        const int h0 = p * g.stride - g.pad, w0 = q * g.stride - g.pad;
        uint4* dst = reinterpret_cast<uint4*>(col) + (pix * taps) * cv + c8;
        const uint4* src = reinterpret_cast<const uint4*>(x) + ((int64_t)n * g.H * g.W) * cv + c8;
        for (int r = 0; r < g.R; ++r) {
            const int h = h0 + r;
            const bool hok = h >= 0 && h < g.H;
#pragma unroll 4
            for (int s2 = 0; s2 < g.S; ++s2) {
                const int w = w0 + s2;
                uint4 v = make_uint4(0, 0, 0, 0);
                if (hok && w >= 0 && w < g.W) v = src[((int64_t)h * g.W + w) * cv];
                dst[(int64_t)(r * g.S + s2) * cv] = v;
            }
        }
    }
}

// Few channels per pixel (the 3 -> 8 channel stem): one thread per 16-byte vector of the column
// matrix, consecutive threads on consecutive vectors, so a warp's store is one contiguous 512 bytes
// (the per-pixel walk above would put 32 lanes on 32 rows R*S*C*2 bytes apart: 16 bytes per
// sector written, 0.66 ms for the 7x7 stem at batch 64).
__global__ void __launch_bounds__(256) im2col_flat_kernel(const __nv_bfloat16* __restrict__ x,
                                                          __nv_bfloat16* __restrict__ col, ConvGeom g,
                                                          int64_t total_vecs) {
    const unsigned cv = (unsigned)g.C >> 3, taps = (unsigned)(g.R * g.S), S = (unsigned)g.S;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total_vecs; i += stride) {
        const int64_t pt = i / cv;                       // (pixel, tap)
        const unsigned c8 = (unsigned)(i - pt * cv);
        const int64_t pix = pt / taps;
        const unsigned tap = (unsigned)(pt - pix * taps);
        const unsigned r = tap / S, s2 = tap - r * S;
        int64_t t = pix;
        const int q = (int)(t % g.Q); t /= g.Q;
        const int p = (int)(t % g.P); t /= g.P;
        const int n = (int)t;
        const int h = p * g.stride - g.pad + (int)r, w = q * g.stride - g.pad + (int)s2;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (h >= 0 && h < g.H && w >= 0 && w < g.W)
            v = __ldg(reinterpret_cast<const uint4*>(x) + (((int64_t)n * g.H + h) * g.W + w) * cv + c8);
        __stcs(reinterpret_cast<uint4*>(col) + i, v);
    }
}

__device__ __forceinline__ void acc8(float (&a)[8], const uint4& raw) {
    const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const float2 t = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[q]));
        a[2 * q] += t.x;
        a[2 * q + 1] += t.y;
    }
}

// dx[n,h,w,c] = sum over taps (r,s) and outputs (p,q) with p*stride + r - pad == h, ... of
// dcol[(n,p,q),(r,s,c)].  One thread produces 8 channels of one input pixel.
__global__ void __launch_bounds__(256) col2im_kernel(const __nv_bfloat16* __restrict__ dcol,
                                                     __nv_bfloat16* __restrict__ dx, ConvGeom g,
                                                     int64_t total_vecs) {
    const int cv = g.C >> 3;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total_vecs; i += stride) {
        int64_t t = i;
        const int c8 = (int)(t % cv); t /= cv;
        const int w = (int)(t % g.W); t /= g.W;
        const int h = (int)(t % g.H); t /= g.H;
        const int n = (int)t;
        float a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
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
                const int64_t row = ((int64_t)n * g.P + p) * g.Q + q;
                acc8(a, reinterpret_cast<const uint4*>(dcol)[(row * g.R * g.S + (int64_t)r * g.S + s) * cv + c8]);
            }
        }
        uint4 o;
        o.x = pack_bf16(a[0], a[1]);
        o.y = pack_bf16(a[2], a[3]);
        o.z = pack_bf16(a[4], a[5]);
        o.w = pack_bf16(a[6], a[7]);
        reinterpret_cast<uint4*>(dx)[i] = o;
    }
}

static int conv_grid(int64_t vecs) {
    int64_t b = (vecs + 255) / 256;
    if (b > 148 * 16) b = 148 * 16;
    if (b < 1) b = 1;
    return (int)b;
}

static int make_geom(ConvGeom* g, int N, int H, int W, int C, int R, int S, int stride, int pad) {
    B200_REQUIRE(N > 0 && H > 0 && W > 0 && C > 0 && R > 0 && S > 0 && stride > 0 && pad >= 0, "conv: bad geometry");
    B200_REQUIRE(C % 8 == 0, "conv: channels must be a multiple of 8 (got %d); pad the input", C);
    g->N = N; g->H = H; g->W = W; g->C = C; g->R = R; g->S = S; g->stride = stride; g->pad = pad;
    g->P = (H + 2 * pad - R) / stride + 1;
    g->Q = (W + 2 * pad - S) / stride + 1;
    B200_REQUIRE(g->P > 0 && g->Q > 0, "conv: empty output");
    return 0;
}

}  // namespace b200

extern "C" int b200_im2col(const void* x, void* col, int N, int H, int W, int C, int R, int S, int stride,
                           int pad, void* stream) {
    using namespace b200;
    B200_REQUIRE(x && col, "im2col: null pointer");
    ConvGeom g;
    if (int rc = make_geom(&g, N, H, W, C, R, S, stride, pad)) return rc;
    const int64_t vecs = (int64_t)N * g.P * g.Q * (C / 8);   // one thread item = 8 channels of a pixel
    if (C < 64) {
        const int64_t all = vecs * R * S;
        im2col_flat_kernel<<<conv_grid(all), 256, 0, (cudaStream_t)stream>>>(
            (const __nv_bfloat16*)x, (__nv_bfloat16*)col, g, all);
    } else {
        im2col_kernel<<<conv_grid(vecs), 256, 0, (cudaStream_t)stream>>>(
            (const __nv_bfloat16*)x, (__nv_bfloat16*)col, g, vecs);
    }
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int b200_col2im(const void* dcol, void* dx, int N, int H, int W, int C, int R, int S, int stride,
                           int pad, void* stream) {
    using namespace b200;
    B200_REQUIRE(dcol && dx, "col2im: null pointer");
    ConvGeom g;
    if (int rc = make_geom(&g, N, H, W, C, R, S, stride, pad)) return rc;
    const int64_t vecs = (int64_t)N * H * W * (C / 8);
    col2im_kernel<<<conv_grid(vecs), 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)dcol, (__nv_bfloat16*)dx, g, vecs);
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}
