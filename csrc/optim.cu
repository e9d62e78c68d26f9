// Flat-arena optimizer: Adam / AdamW / SGD-momentum in one pass.
//
// Replaces the ~50 small launches of torch's multi-tensor Adam (ref train.py:139,
// torch:optim/adam.py) for the single-GPU and NCCL paths; the peer-memory path
// inlines the same update (opt_update) after its in-kernel reduction (comm.cu).
// Memory-bound: 16 B read (p,m,v,g fp32) + 12 B written (+2 B bf16 shadow) per
// parameter, 128-bit accesses, grid sized to fill 148 SMs.
#include "common.cuh"
#include "optim_update.cuh"

#include <cstdlib>

void b200_count_launch(int n);

namespace b200 {

template <bool GRAD_BF16>
__global__ void __launch_bounds__(256) optimizer_kernel(
    b200_opt_hparams hp, float* __restrict__ p, const void* __restrict__ g_, float* __restrict__ m,
    float* __restrict__ v, __nv_bfloat16* __restrict__ shadow, int64_t n, int* step_dev,
    int step_host, int advance, const OptScalars* __restrict__ scalars) {
    __shared__ OptScalars sc;
    // Programmatic dependent launch (no-ops on an ordinary launch): let the next kernel's
    // prologue start, then wait for the producer of the gradients.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    // `scalars`: bias corrections etc. of this step, precomputed by the producer kernel
    // (csrc/mlp_fused.cu) — make_opt_scalars runs two double-precision pow(): microseconds of
    // serial latency in front of every block's first load when done here.
    if (threadIdx.x == 0) {
        if (scalars) {
            sc = *scalars;
        } else {
            const int t = step_dev ? (*(volatile int*)step_dev + 1) : step_host;
            sc = make_opt_scalars(hp, t);
        }
    }
    __syncthreads();
    const OptScalars s = sc;
    const int64_t n4 = n >> 2;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 pp = reinterpret_cast<float4*>(p)[i];
        float4 mm = reinterpret_cast<float4*>(m)[i];
        float4 vv = make_float4(0, 0, 0, 0);
        if (hp.kind != 2) vv = reinterpret_cast<float4*>(v)[i];
        float4 gg;
        if (GRAD_BF16) {
            const uint2 raw = reinterpret_cast<const uint2*>(g_)[i];
            const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.x));
            const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.y));
            gg = make_float4(a.x, a.y, b.x, b.y);
        } else {
            gg = reinterpret_cast<const float4*>(g_)[i];
        }
        opt_update(hp.kind, s, pp.x, mm.x, vv.x, gg.x * hp.grad_scale);
        opt_update(hp.kind, s, pp.y, mm.y, vv.y, gg.y * hp.grad_scale);
        opt_update(hp.kind, s, pp.z, mm.z, vv.z, gg.z * hp.grad_scale);
        opt_update(hp.kind, s, pp.w, mm.w, vv.w, gg.w * hp.grad_scale);
        reinterpret_cast<float4*>(p)[i] = pp;
        reinterpret_cast<float4*>(m)[i] = mm;
        if (hp.kind != 2) reinterpret_cast<float4*>(v)[i] = vv;
        if (shadow) {
            uint2 o;
            o.x = pack_bf16(pp.x, pp.y);
            o.y = pack_bf16(pp.z, pp.w);
            reinterpret_cast<uint2*>(shadow)[i] = o;
        }
    }
    // scalar tail (n not a multiple of 4; arenas are padded so this is rare)
    for (int64_t i = (n4 << 2) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        float pp = p[i], mm = m[i], vv = hp.kind != 2 ? v[i] : 0.f;
        float gg = GRAD_BF16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(g_)[i])
                             : reinterpret_cast<const float*>(g_)[i];
        opt_update(hp.kind, s, pp, mm, vv, gg * hp.grad_scale);
        p[i] = pp;
        m[i] = mm;
        if (hp.kind != 2) v[i] = vv;
        if (shadow) shadow[i] = __float2bfloat16_rn(pp);
    }
    if (step_dev && advance) {
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            const int ticket = atomicAdd(step_dev + 1, 1);
            if (ticket == (int)gridDim.x - 1) {
                step_dev[1] = 0;
                step_dev[0] = s.t;
                __threadfence();
            }
        }
    }
}

}  // namespace b200

static int optimizer_step_impl(const b200_opt_hparams* hp, float* master, const void* grad,
                               int grad_is_bf16, float* exp_avg, float* exp_avg_sq,
                               void* shadow_bf16, int64_t n, int* step_dev, int step_host,
                               int advance, void* stream, bool pdl, const void* scalars = nullptr) {
    B200_REQUIRE(hp && master && grad && exp_avg, "optimizer_step: null pointer");
    B200_REQUIRE(hp->kind >= 0 && hp->kind <= 2, "optimizer_step: unknown kind %d", hp->kind);
    B200_REQUIRE(hp->kind == 2 || exp_avg_sq, "optimizer_step: Adam needs exp_avg_sq");
    B200_REQUIRE(n >= 0, "optimizer_step: negative length");
    B200_REQUIRE(((uintptr_t)master | (uintptr_t)grad | (uintptr_t)exp_avg | (uintptr_t)exp_avg_sq |
                  (uintptr_t)shadow_bf16) % 16 == 0,
                 "optimizer_step: arena slices must be 16-byte aligned");
    if (n == 0) return 0;
    int sms = 148;
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int64_t want = (n / 4 + 255) / 256;
    int blocks = (int)(want < 1 ? 1 : (want > (int64_t)sms * 8 ? (int64_t)sms * 8 : want));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)blocks);
    cfg.blockDim = dim3(256);
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    const b200_opt_hparams h = *hp;
    __nv_bfloat16* sh = (__nv_bfloat16*)shadow_bf16;
    if (grad_is_bf16)
        B200_CUDA_OK(cudaLaunchKernelEx(&cfg, b200::optimizer_kernel<true>, h, master, grad, exp_avg, exp_avg_sq, sh, n,
                                        step_dev, step_host, advance, (const b200::OptScalars*)scalars));
    else
        B200_CUDA_OK(cudaLaunchKernelEx(&cfg, b200::optimizer_kernel<false>, h, master, grad, exp_avg, exp_avg_sq, sh, n,
                                        step_dev, step_host, advance, (const b200::OptScalars*)scalars));
    b200_count_launch(1);
    return 0;
}

extern "C" int b200_optimizer_step(const b200_opt_hparams* hp, float* master, const void* grad,
                                   int grad_is_bf16, float* exp_avg, float* exp_avg_sq,
                                   void* shadow_bf16, int64_t n, int* step_dev, int step_host,
                                   int advance, void* stream) {
    return optimizer_step_impl(hp, master, grad, grad_is_bf16, exp_avg, exp_avg_sq, shadow_bf16, n, step_dev,
                               step_host, advance, stream, false);
}

// Launched with programmatic stream serialization: the kernel's blocks become resident while the
// producer of the gradients is still running and start the moment it completes (csrc/mlp_fused.cu).
// `scalars` (device, b200::OptScalars of this step) may be null.
int b200_optimizer_step_pdl(const b200_opt_hparams* hp, float* master, const void* grad, int grad_is_bf16,
                            float* exp_avg, float* exp_avg_sq, void* shadow_bf16, int64_t n, int* step_dev,
                            int step_host, int advance, const void* scalars, void* stream) {
    static const bool pdl = getenv("B200DDP_NO_PDL") == nullptr;
    return optimizer_step_impl(hp, master, grad, grad_is_bf16, exp_avg, exp_avg_sq, shadow_bf16, n, step_dev,
                               step_host, advance, stream, pdl, scalars);
}
