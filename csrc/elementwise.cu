// Dtype casts (vectorised, grid-stride).  ref: train.py:133 moves an fp32 batch
// to the device; the bf16 compute path needs it in bf16 once.
#include "common.cuh"

void b200_count_launch(int n);

namespace b200 {

__global__ void __launch_bounds__(256) cast_f32_bf16(const float* __restrict__ src,
                                                     __nv_bfloat16* __restrict__ dst, int64_t n) {
    const int64_t n8 = n >> 3;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (int64_t i = tid; i < n8; i += stride) {
        const float4 a = reinterpret_cast<const float4*>(src)[2 * i];
        const float4 b = reinterpret_cast<const float4*>(src)[2 * i + 1];
        uint4 o;
        o.x = pack_bf16(a.x, a.y);
        o.y = pack_bf16(a.z, a.w);
        o.z = pack_bf16(b.x, b.y);
        o.w = pack_bf16(b.z, b.w);
        reinterpret_cast<uint4*>(dst)[i] = o;
    }
    for (int64_t i = (n8 << 3) + tid; i < n; i += stride) dst[i] = __float2bfloat16_rn(src[i]);
}

__global__ void __launch_bounds__(256) cast_bf16_f32(const __nv_bfloat16* __restrict__ src,
                                                     float* __restrict__ dst, int64_t n) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        dst[i] = __bfloat162float(src[i]);
}

// Batch staging for the captured step program: cast the fp32 features to bf16 and copy
// the int64 labels into the graph's fixed input buffers, in one launch.
// It also zeroes the few fp32 words the step accumulates into with atomics (bias
// gradients, loss) so the captured graph needs no memset nodes.
struct ZeroList {
    float* ptr[B200_STAGE_MAX_ZERO];
    int count[B200_STAGE_MAX_ZERO];
    int n;
};

__global__ void __launch_bounds__(256) stage_batch(const float* __restrict__ x,
                                                   const int64_t* __restrict__ y,
                                                   __nv_bfloat16* __restrict__ xo,
                                                   int64_t* __restrict__ yo, int64_t nx, int64_t ny,
                                                   const ZeroList zl) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (int r = 0; r < zl.n; ++r)
        for (int64_t i = tid; i < zl.count[r]; i += stride) zl.ptr[r][i] = 0.f;
    const int64_t n4 = nx >> 2;
    for (int64_t i = tid; i < n4; i += stride) {
        const float4 a = reinterpret_cast<const float4*>(x)[i];
        uint2 o;
        o.x = pack_bf16(a.x, a.y);
        o.y = pack_bf16(a.z, a.w);
        reinterpret_cast<uint2*>(xo)[i] = o;
    }
    for (int64_t i = (n4 << 2) + tid; i < nx; i += stride) xo[i] = __float2bfloat16_rn(x[i]);
    for (int64_t i = tid; i < ny; i += stride) yo[i] = y[i];
}

static int grid_for(int64_t work) {
    int64_t b = (work + 255) / 256;
    if (b < 1) b = 1;
    if (b > 148 * 8) b = 148 * 8;
    return (int)b;
}

}  // namespace b200

extern "C" int b200_cast_f32_to_bf16(const float* src, void* dst, int64_t n, void* stream) {
    B200_REQUIRE(src && dst && n >= 0, "cast: bad arguments");
    B200_REQUIRE(((uintptr_t)src % 16 == 0) && ((uintptr_t)dst % 16 == 0), "cast: pointers must be 16-byte aligned");
    if (n == 0) return 0;
    b200::cast_f32_bf16<<<b200::grid_for(n / 8), 256, 0, (cudaStream_t)stream>>>(
        src, (__nv_bfloat16*)dst, n);
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int b200_cast_bf16_to_f32(const void* src, float* dst, int64_t n, void* stream) {
    B200_REQUIRE(src && dst && n >= 0, "cast: bad arguments");
    if (n == 0) return 0;
    b200::cast_bf16_f32<<<b200::grid_for(n), 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)src, dst, n);
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int b200_stage_batch(const float* x, const int64_t* y, void* x_bf16, int64_t* y_out,
                                int64_t nx, int64_t ny, float* const* zero_ptrs,
                                const int* zero_counts, int nzero, void* stream) {
    B200_REQUIRE(x && y && x_bf16 && y_out && nx >= 0 && ny >= 0, "stage_batch: bad arguments");
    B200_REQUIRE(nzero >= 0 && nzero <= B200_STAGE_MAX_ZERO, "stage_batch: at most %d zero regions", B200_STAGE_MAX_ZERO);
    b200::ZeroList zl{};
    zl.n = nzero;
    for (int r = 0; r < nzero; ++r) { zl.ptr[r] = zero_ptrs[r]; zl.count[r] = zero_counts[r]; }
    B200_REQUIRE(((uintptr_t)x % 16 == 0) && ((uintptr_t)x_bf16 % 8 == 0), "stage_batch: x must be 16-byte aligned");
    b200::stage_batch<<<b200::grid_for(nx / 4), 256, 0, (cudaStream_t)stream>>>(
        x, y, (__nv_bfloat16*)x_bf16, y_out, nx, ny, zl);
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}
