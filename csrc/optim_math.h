// The per-element optimizer update, written once for the device kernels (optim.cu, comm.cu) and
// for the host oracle (oracle/oracle.cpp): arithmetic follows torch/optim/adam.py
// (_single_tensor_adam, non-capturable) and sgd.py term for term, with every rounding step spelled
// out (fmaf / IEEE division / IEEE square root) so that host and device agree bit for bit.
// Host builds must not contract floating-point expressions (-ffp-contract=off).
#pragma once
#include <math.h>

#include "../include/b200ddp.h"

#ifdef __CUDACC__
#define B200_HD __host__ __device__ __forceinline__
#else
#define B200_HD inline
#endif

namespace b200 {

struct OptScalars {
    int t;
    float lr, b1, b2, one_minus_b1, one_minus_b2, eps, wd, mu;
    float step_size;     // lr / (1 - b1^t)
    float inv_bc2_sqrt;  // 1 / sqrt(1 - b2^t)
    float bc2_sqrt;
    float decay;         // AdamW: 1 - lr*wd
};

B200_HD float div_rn(float a, float b) {
#ifdef __CUDA_ARCH__
    return __fdiv_rn(a, b);
#else
    return a / b;
#endif
}
B200_HD float sqrt_rn(float a) {
#ifdef __CUDA_ARCH__
    return __fsqrt_rn(a);
#else
    return sqrtf(a);
#endif
}

B200_HD OptScalars make_opt_scalars(const b200_opt_hparams& hp, int t) {
    OptScalars s;
    s.t = t;
    s.lr = hp.lr;
    s.b1 = hp.beta1;
    s.b2 = hp.beta2;
    s.one_minus_b1 = (float)(1.0 - (double)hp.beta1);
    s.one_minus_b2 = (float)(1.0 - (double)hp.beta2);
    s.eps = hp.eps;
    s.wd = hp.weight_decay;
    s.mu = hp.momentum;
    // Bias corrections in double, like the Python scalars torch computes them with.
    const double bc1 = 1.0 - pow((double)hp.beta1, (double)t);
    const double bc2 = 1.0 - pow((double)hp.beta2, (double)t);
    s.step_size = (float)((double)hp.lr / bc1);
    s.bc2_sqrt = (float)sqrt(bc2);
    s.inv_bc2_sqrt = (float)(1.0 / sqrt(bc2));
    s.decay = (float)(1.0 - (double)hp.lr * (double)hp.weight_decay);
    return s;
}

// g is already scaled by 1/world.
B200_HD void opt_update(int kind, const OptScalars& s, float& p, float& m, float& v, float g) {
    if (kind == 2) {  // SGD with momentum (dampening 0, no nesterov)
        if (s.wd != 0.f) g = fmaf(s.wd, p, g);
        if (s.mu != 0.f) {
            m = (s.t == 1) ? g : fmaf(s.mu, m, g);
            g = m;
        }
        p = fmaf(-s.lr, g, p);
        return;
    }
    if (kind == 1) {
        p *= s.decay;  // decoupled weight decay
    } else if (s.wd != 0.f) {
        g = fmaf(s.wd, p, g);
    }
    m = fmaf(s.one_minus_b1, g - m, m);              // m.lerp_(g, 1-b1)
    v = fmaf(s.one_minus_b2 * g, g, s.b2 * v);       // v.mul_(b2).addcmul_(g, g, 1-b2)
    const float denom = div_rn(sqrt_rn(v), s.bc2_sqrt) + s.eps;
    p = fmaf(-s.step_size, div_rn(m, denom), p);     // p.addcdiv_(m, denom, -step_size)
}

}  // namespace b200
