// Host restatements of the device kernels' arithmetic (plain C++17, no CUDA): the numerics oracle
// for the tests that run without a GPU and the timed CPU baseline of tests/test_oracle.py.
//
// Each function states, in the simplest loops possible, WHAT the corresponding sm_100a kernel
// computes and in which precision: bf16 operands, fp32 accumulation, where the result is rounded.
// The optimizer update is not restated but SHARED with the kernels (csrc/optim_math.h), so the
// bit-exact comparison against torch.optim on the host covers the device code's formula.
//
// Build: g++ -O2 -std=c++17 -ffp-contract=off -shared -fPIC (b200ddp/_native/build.py).
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "../csrc/optim_math.h"

namespace {

inline float bf16_to_f32(uint16_t h) {
    uint32_t u = (uint32_t)h << 16;
    float f;
    std::memcpy(&f, &u, 4);
    return f;
}
inline uint16_t f32_to_bf16(float f) {   // round to nearest even, NaN kept quiet
    uint32_t u;
    std::memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);
    u += 0x7fffu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}
inline float gelu_tanh(float z) {
    const float u = 0.7978845608028654f * (z + 0.044715f * z * z * z);
    return 0.5f * z * (1.f + std::tanh(u));
}
inline float gelu_tanh_grad(float z) {
    const float u = 0.7978845608028654f * (z + 0.044715f * z * z * z);
    const float t = std::tanh(u);
    return 0.5f * (1.f + t) + 0.5f * z * (1.f - t * t) * 0.7978845608028654f * (1.f + 3.f * 0.044715f * z * z);
}

}  // namespace

extern "C" {

// ---- optimizer (csrc/optim.cu, the update stage of csrc/comm.cu).  grad is fp32; kind as in
// b200_opt_hparams; t is the 1-based step number; grad_scale = 1/world.
void oracle_optimizer_step(const b200_opt_hparams* hp, float* p, const float* g, float* m, float* v,
                           uint16_t* shadow_bf16, int64_t n, int t) {
    const b200::OptScalars s = b200::make_opt_scalars(*hp, t);
    float dummy = 0.f;
    for (int64_t i = 0; i < n; ++i) {
        b200::opt_update(hp->kind, s, p[i], m[i], v ? v[i] : dummy, g[i] * hp->grad_scale);
        if (shadow_bf16) shadow_bf16[i] = f32_to_bf16(p[i]);
    }
}

// ---- ordered all-reduce (csrc/comm.cu): ranks summed 0..W-1 in fp32, then scaled.
void oracle_allreduce_sum(const float* const* parts, int world, int64_t n, float scale, float* out) {
    for (int64_t i = 0; i < n; ++i) {
        float acc = parts[0][i];
        for (int r = 1; r < world; ++r) acc += parts[r][i];
        out[i] = acc * scale;
    }
}

// ---- GEMM + epilogue (csrc/gemm.cu): D[M,N] = epi(A[M,K] * B[N,K]^T), bf16 operands (row-major,
// K contiguous), fp32 accumulation in k order, bias fp32, output rounded once to bf16.
// epilogue (B200_EPI_* codes): NONE (+ residual aux when given), RELU_DROPOUT with p = 0,
// GELU (aux_out = pre-activation), GELU_BWD (D = acc * gelu'(aux)).
void oracle_gemm_bf16(const uint16_t* A, const uint16_t* B, uint16_t* D, int M, int N, int K, const float* bias,
                      int epilogue, const uint16_t* aux, uint16_t* aux_out) {
    for (int m = 0; m < M; ++m)
        for (int n = 0; n < N; ++n) {
            float acc = 0.f;
            for (int k = 0; k < K; ++k) acc += bf16_to_f32(A[(int64_t)m * K + k]) * bf16_to_f32(B[(int64_t)n * K + k]);
            if (bias) acc += bias[n];
            const int64_t o = (int64_t)m * N + n;
            switch (epilogue) {
                case B200_EPI_RELU_DROPOUT: acc = acc > 0.f ? acc : 0.f; break;   // dropout p = 0
                case B200_EPI_GELU:
                    if (aux_out) aux_out[o] = f32_to_bf16(acc);
                    acc = gelu_tanh(acc);
                    break;
                case B200_EPI_GELU_BWD: acc *= gelu_tanh_grad(bf16_to_f32(aux[o])); break;
                default:
                    if (aux) acc += bf16_to_f32(aux[o]);
                    break;
            }
            D[o] = f32_to_bf16(acc);
        }
}

// ---- softmax cross-entropy, forward + backward (csrc/loss.cu, the fused classifier epilogue):
// fp32 logits; loss_out[0] = mean loss, loss_out[1] = rows whose argmax is the target;
// dlogits = (softmax - onehot) * grad_scale.
void oracle_cross_entropy(const float* logits, const int64_t* targets, int64_t rows, int cols, float* dlogits,
                          float grad_scale, float* loss_out) {
    double total = 0.0;
    float correct = 0.f;
    for (int64_t r = 0; r < rows; ++r) {
        const float* z = logits + r * cols;
        float mx = z[0];
        int arg = 0;
        for (int c = 1; c < cols; ++c)
            if (z[c] > mx) { mx = z[c]; arg = c; }
        float sum = 0.f;
        for (int c = 0; c < cols; ++c) sum += std::exp(z[c] - mx);
        const int t = (int)targets[r];
        total += (double)(mx + std::log(sum) - z[t]);
        correct += arg == t ? 1.f : 0.f;
        if (dlogits)
            for (int c = 0; c < cols; ++c)
                dlogits[r * cols + c] = (std::exp(z[c] - mx) / sum - (c == t ? 1.f : 0.f)) * grad_scale;
    }
    loss_out[0] = (float)(total / (double)rows);
    loss_out[1] = correct;
}

// ---- LayerNorm (csrc/norm.cu): fp32 statistics over bf16 rows, biased variance, output bf16.
void oracle_layernorm_fwd(const uint16_t* x, const uint16_t* gamma, const uint16_t* beta, uint16_t* y, float* mean,
                          float* rstd, int64_t rows, int C, float eps) {
    for (int64_t r = 0; r < rows; ++r) {
        double s = 0.0, q = 0.0;
        for (int c = 0; c < C; ++c) s += bf16_to_f32(x[r * C + c]);
        const float mu = (float)(s / C);
        for (int c = 0; c < C; ++c) {
            const double d = bf16_to_f32(x[r * C + c]) - mu;
            q += d * d;
        }
        const float rs = 1.f / std::sqrt((float)(q / C) + eps);
        mean[r] = mu;
        rstd[r] = rs;
        for (int c = 0; c < C; ++c)
            y[r * C + c] = f32_to_bf16((bf16_to_f32(x[r * C + c]) - mu) * rs * bf16_to_f32(gamma[c]) + bf16_to_f32(beta[c]));
    }
}

// dx and the fp32 column sums dgamma / dbeta.
void oracle_layernorm_bwd(const uint16_t* dy, const uint16_t* x, const uint16_t* gamma, const float* mean,
                          const float* rstd, uint16_t* dx, float* dgamma, float* dbeta, int64_t rows, int C) {
    std::vector<double> dg(C, 0.0), db(C, 0.0);
    for (int64_t r = 0; r < rows; ++r) {
        double a = 0.0, b = 0.0;   // sum(dyhat), sum(dyhat * xhat)
        for (int c = 0; c < C; ++c) {
            const float g = bf16_to_f32(dy[r * C + c]);
            const float xh = (bf16_to_f32(x[r * C + c]) - mean[r]) * rstd[r];
            const float dyh = g * bf16_to_f32(gamma[c]);
            a += dyh;
            b += (double)dyh * xh;
            dg[c] += (double)g * xh;
            db[c] += g;
        }
        for (int c = 0; c < C; ++c) {
            const float xh = (bf16_to_f32(x[r * C + c]) - mean[r]) * rstd[r];
            const float dyh = bf16_to_f32(dy[r * C + c]) * bf16_to_f32(gamma[c]);
            dx[r * C + c] = f32_to_bf16(rstd[r] * (dyh - (float)(a / C) - xh * (float)(b / C)));
        }
    }
    for (int c = 0; c < C; ++c) { dgamma[c] = (float)dg[c]; dbeta[c] = (float)db[c]; }
}

// ---- BatchNorm forward in training mode, channels-last, fused ReLU / residual (csrc/batchnorm.cu):
// batch statistics in fp32, running statistics updated with the unbiased variance.
void oracle_batchnorm_fwd(const uint16_t* x, const uint16_t* residual, uint16_t* y, const float* gamma,
                          const float* beta, float* running_mean, float* running_var, int64_t rows, int C,
                          float eps, float momentum, int relu) {
    for (int c = 0; c < C; ++c) {
        double s = 0.0, q = 0.0;
        for (int64_t r = 0; r < rows; ++r) {
            const double v = bf16_to_f32(x[r * C + c]);
            s += v;
            q += v * v;
        }
        const float m = (float)(s / rows);
        float var = (float)(q / rows) - m * m;
        if (var < 0.f) var = 0.f;
        const float rs = 1.f / std::sqrt(var + eps);
        if (running_mean) {
            const float unbiased = rows > 1 ? var * (float)rows / (float)(rows - 1) : var;
            running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * m;
            running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
        }
        const float sc = rs * gamma[c], sh = beta[c] - m * sc;
        for (int64_t r = 0; r < rows; ++r) {
            float v = std::fma(bf16_to_f32(x[r * C + c]), sc, sh);
            if (residual) v += bf16_to_f32(residual[r * C + c]);
            if (relu && v < 0.f) v = 0.f;
            y[r * C + c] = f32_to_bf16(v);
        }
    }
}

// ---- im2col / col2im (csrc/conv.cu), NHWC bf16: col[(n,p,q), (r,s,c)], zero fill at the borders;
// col2im sums the taps that touch each input pixel in fp32 and rounds once.
void oracle_im2col(const uint16_t* x, uint16_t* col, int N, int H, int W, int C, int R, int S, int stride, int pad) {
    const int P = (H + 2 * pad - R) / stride + 1, Q = (W + 2 * pad - S) / stride + 1;
    for (int n = 0; n < N; ++n)
        for (int p = 0; p < P; ++p)
            for (int q = 0; q < Q; ++q)
                for (int r = 0; r < R; ++r)
                    for (int s = 0; s < S; ++s) {
                        const int h = p * stride - pad + r, w = q * stride - pad + s;
                        uint16_t* dst = col + ((((int64_t)n * P + p) * Q + q) * R * S + r * S + s) * C;
                        if (h < 0 || h >= H || w < 0 || w >= W) std::memset(dst, 0, (size_t)C * 2);
                        else std::memcpy(dst, x + (((int64_t)n * H + h) * W + w) * C, (size_t)C * 2);
                    }
}
void oracle_col2im(const uint16_t* dcol, uint16_t* dx, int N, int H, int W, int C, int R, int S, int stride, int pad) {
    const int P = (H + 2 * pad - R) / stride + 1, Q = (W + 2 * pad - S) / stride + 1;
    std::vector<float> acc((size_t)N * H * W * C, 0.f);
    for (int n = 0; n < N; ++n)
        for (int p = 0; p < P; ++p)
            for (int q = 0; q < Q; ++q)
                for (int r = 0; r < R; ++r)
                    for (int s = 0; s < S; ++s) {
                        const int h = p * stride - pad + r, w = q * stride - pad + s;
                        if (h < 0 || h >= H || w < 0 || w >= W) continue;
                        const uint16_t* src = dcol + ((((int64_t)n * P + p) * Q + q) * R * S + r * S + s) * C;
                        float* a = acc.data() + (((int64_t)n * H + h) * W + w) * C;
                        for (int c = 0; c < C; ++c) a[c] += bf16_to_f32(src[c]);
                    }
    for (size_t i = 0; i < acc.size(); ++i) dx[i] = f32_to_bf16(acc[i]);
}

// ---- max pooling (csrc/pool.cu), NHWC bf16: first maximum in scan order wins; the backward routes
// each output gradient to its winner, summing in fp32.
void oracle_maxpool_fwd(const uint16_t* x, uint16_t* y, uint8_t* idx, int N, int H, int W, int C, int R, int S,
                        int stride, int pad) {
    const int P = (H + 2 * pad - R) / stride + 1, Q = (W + 2 * pad - S) / stride + 1;
    for (int n = 0; n < N; ++n)
        for (int p = 0; p < P; ++p)
            for (int q = 0; q < Q; ++q)
                for (int c = 0; c < C; ++c) {
                    float best = -INFINITY;
                    int arg = 0;
                    bool first = true;
                    for (int r = 0; r < R; ++r)
                        for (int s = 0; s < S; ++s) {
                            const int h = p * stride - pad + r, w = q * stride - pad + s;
                            if (h < 0 || h >= H || w < 0 || w >= W) continue;
                            const float v = bf16_to_f32(x[(((int64_t)n * H + h) * W + w) * C + c]);
                            if (first || v > best) { best = v; arg = r * S + s; }
                            first = false;
                        }
                    const int64_t o = (((int64_t)n * P + p) * Q + q) * C + c;
                    y[o] = f32_to_bf16(best);
                    idx[o] = (uint8_t)arg;
                }
}
void oracle_maxpool_bwd(const uint16_t* dy, const uint8_t* idx, uint16_t* dx, int N, int H, int W, int C, int R, int S,
                        int stride, int pad) {
    const int P = (H + 2 * pad - R) / stride + 1, Q = (W + 2 * pad - S) / stride + 1;
    std::vector<float> acc((size_t)N * H * W * C, 0.f);
    for (int n = 0; n < N; ++n)
        for (int p = 0; p < P; ++p)
            for (int q = 0; q < Q; ++q)
                for (int c = 0; c < C; ++c) {
                    const int64_t o = (((int64_t)n * P + p) * Q + q) * C + c;
                    const int r = idx[o] / S, s = idx[o] % S;
                    const int h = p * stride - pad + r, w = q * stride - pad + s;
                    acc[(((int64_t)n * H + h) * W + w) * C + c] += bf16_to_f32(dy[o]);
                }
    for (size_t i = 0; i < acc.size(); ++i) dx[i] = f32_to_bf16(acc[i]);
}
// global average pool: x [N, HW, C] -> y [N, C]
void oracle_avgpool_fwd(const uint16_t* x, uint16_t* y, int N, int HW, int C) {
    for (int n = 0; n < N; ++n)
        for (int c = 0; c < C; ++c) {
            float acc = 0.f;
            for (int px = 0; px < HW; ++px) acc += bf16_to_f32(x[((int64_t)n * HW + px) * C + c]);
            y[(int64_t)n * C + c] = f32_to_bf16(acc * (1.f / (float)HW));
        }
}

// ---- causal attention forward (csrc/flash_attention.cu): qkv [B, T, 3, H, D] bf16 -> out [B, T, H*D]
// bf16 and the natural-log LSE of the scaled scores [B, H, T]; probabilities rounded to bf16
// before the PV product, as on the device.
void oracle_causal_attention(const uint16_t* qkv, uint16_t* out, float* lse, int B, int T, int H, int D) {
    const float scale = 1.f / std::sqrt((float)D);
    std::vector<float> s(T), o(D);
    for (int b = 0; b < B; ++b)
        for (int h = 0; h < H; ++h)
            for (int i = 0; i < T; ++i) {
                const uint16_t* q = qkv + (((int64_t)b * T + i) * 3 + 0) * H * D + (int64_t)h * D;
                float mx = -INFINITY;
                for (int j = 0; j <= i; ++j) {
                    const uint16_t* k = qkv + (((int64_t)b * T + j) * 3 + 1) * H * D + (int64_t)h * D;
                    float acc = 0.f;
                    for (int d = 0; d < D; ++d) acc += bf16_to_f32(q[d]) * bf16_to_f32(k[d]);
                    s[j] = acc * scale;
                    mx = std::fmax(mx, s[j]);
                }
                float l = 0.f;
                std::fill(o.begin(), o.end(), 0.f);
                for (int j = 0; j <= i; ++j) {
                    const float pj = std::exp(s[j] - mx);
                    l += pj;
                    const float pr = bf16_to_f32(f32_to_bf16(pj));
                    const uint16_t* v = qkv + (((int64_t)b * T + j) * 3 + 2) * H * D + (int64_t)h * D;
                    for (int d = 0; d < D; ++d) o[d] += pr * bf16_to_f32(v[d]);
                }
                uint16_t* dst = out + ((int64_t)b * T + i) * H * D + (int64_t)h * D;
                for (int d = 0; d < D; ++d) dst[d] = f32_to_bf16(o[d] / l);
                lse[((int64_t)b * H + h) * T + i] = mx + std::log(l);
            }
}

}  // extern "C"
