// tcgen05 / TMEM / TMA GEMM with fused epilogues (sm_100a).
//
//   D[M,N] = epilogue( A[M,K] * B[N,K]^T )      bf16 in, fp32 accumulate
//
// Replaces the cuBLAS calls behind nn.Linear forward and its two backward GEMMs
// (ref train.py:39-45,138) and fuses what torch runs as
// separate kernels around them: bias add, ReLU, dropout, their backward gates,
// GELU fwd/bwd, residual add, bias-gradient column sums and, for the classifier
// layer, the whole softmax cross-entropy forward+backward.
//
// Three kernel families share the pieces below (selection: select_kernel / b200_gemm_plan):
//   gemm_kernel<BN, FAST>   128 x BN tiles, one SM.  General form (4 epilogue warps, every
//                           epilogue) and lean form (8 epilogue warps, light epilogues, bf16 out).
//   gemm_kernel_2sm<BN>     256 x BN tiles per SM pair: tcgen05 cta_group::2 on a 2-CTA cluster.
//   gemm_kernel_m256        256 x 256 tiles, one SM, two accumulators sharing each B tile.
//
// Structure of each (persistent, one CTA per SM, 192 or 320 threads):
//   warp 0      TMA producer: cp.async.bulk.tensor tiles of A and B into a
//               STAGES-deep ring of 128B-swizzled shared memory, mbarrier tx-count.
//   warp 1      MMA issuer: tcgen05.mma (M=128, N=BN, K=16) from shared-memory descriptors
//               into a TMEM accumulator; tcgen05.commit frees ring slots and publishes
//               finished accumulators.  Both warps walk their loops whole, one elected lane
//               issues (keeps every operand in uniform registers; see gemm_kernel).
//   warps 2..   epilogue: tcgen05.ld the accumulator (one row per thread; the lean kernels put
//               two warps on each TMEM lane quarter, splitting the columns), apply the fused
//               epilogue, store.  TMEM holds two accumulators so the epilogue of tile i
//               overlaps the MMAs of tile i+1.  The tile's bias slice is staged in shared
//               memory and the aux operand is fetched one chunk ahead: with one or two warps
//               per scheduler every exposed load latency is paid in full (ncu: the bias
//               loads were 38 % of all stall samples at K = 768).
//
// Operand layouts: K-major (reduction index contiguous: activations [M,K],
// nn.Linear weights [N,K]) and MN-major (stored [K,M] / [K,N]) are both native —
// the UMMA instruction descriptor carries a_major/b_major, the shared-memory
// descriptor the matching LBO/SBO — so dgrad (needs W as [K,N]) and wgrad (needs
// dY^T and X^T) read the tensors that already exist; nothing is transposed.
// Ragged edges in M, N and K are zero-filled by TMA; stores are masked.
#include "common.cuh"

#include <cstdlib>
#include <mutex>

void b200_count_launch(int n);

namespace b200 {

constexpr int BM = 128;
constexpr int BK = 64;  // 64 bf16 = 128 B = one swizzle row
// TMA warp + MMA warp + epilogue warps.  The general kernels run 4 epilogue warps (one per TMEM
// lane quarter, up to 255 registers each for the dropout / cross-entropy / column-sum paths); the
// FAST kernels (plain, bias, residual, GELU forward/backward with bf16 output) run 8 lean ones.
template <bool FAST> struct EpiWarps { static constexpr int value = FAST ? 8 : 4; };
template <bool FAST> constexpr int gemm_threads() { return 64 + 32 * EpiWarps<FAST>::value; }
constexpr int BIAS_SMEM_BYTES = 2 * 256 * 4;  // per-tile bias slice, double buffered
constexpr uint32_t A_STAGE_BYTES = BM * BK * 2;  // 16 KB

struct GemmParams {
    int M, N, K;
    int a_mn, b_mn;
    void* D; int64_t ldd; int d_fp32;
    const float* bias;
    int epilogue;
    const __nv_bfloat16* aux; int64_t ldaux;
    __nv_bfloat16* aux_out;
    float* colsum;
    float dropout_p, gate_scale;
    uint64_t seed;
    const int* step_dev;
    uint32_t layer_id;
    const int64_t* targets;
    float ce_grad_scale;
    float* loss_out;
    float loss_scale;
    int ce_write_grad;
    uint32_t idesc;
    int num_m_blocks, num_n_blocks, num_k_blocks;
    int batch_inner, batch_outer;          // two-level batch (e.g. heads, then sequences)
    int64_t d_stride_inner, d_stride_outer;  // output offsets per batch index, in elements
    int causal;                            // 0 none, 1 skip tiles above the diagonal, 2 k < m0+BM, 3 k >= m0
    int k_splits, kb_per_split;  // split-K: tile index also selects a K slice; partials are
                                 // combined with vector atomics into a zeroed fp32 D
    // Implicit-GEMM convolution: the operand is an NHWC activation read through a TMA im2col-mode
    // tensor map (rows = output pixels, columns = (filter tap, channel)); nothing is gathered in HBM.
    int a_conv, b_conv;
    int conv_cblocks, conv_S, conv_taps, conv_Q, conv_PQ, conv_stride, conv_pad, conv_C;
};

// (image, top-left input row, top-left input column) of output pixel `m` — the im2col coordinate of the
// first row of a tile; the TMA unit walks the following output pixels itself.
__device__ __forceinline__ void conv_origin(const GemmParams& p, int m, int& n, int& h, int& w) {
    n = m / p.conv_PQ;
    const int rem = m - n * p.conv_PQ;
    const int pp = rem / p.conv_Q;
    h = pp * p.conv_stride - p.conv_pad;
    w = (rem - pp * p.conv_Q) * p.conv_stride - p.conv_pad;
}
// column block `nb` (64 columns of the im2col matrix) -> channel offset and filter tap (s, r); blocks
// past the last tap read channel C.. (out of range: zero-filled, full byte count)
__device__ __forceinline__ void conv_tap(const GemmParams& p, int nb, int& c, uint16_t& fs, uint16_t& fr) {
    const int tap = nb / p.conv_cblocks;
    c = (nb - tap * p.conv_cblocks) * 64;
    if (tap >= p.conv_taps) { c = p.conv_C; fs = 0; fr = 0; return; }
    const int r = tap / p.conv_S;
    fr = (uint16_t)r;
    fs = (uint16_t)(tap - r * p.conv_S);
}
__device__ __forceinline__ void tma_load_im2col(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c, int w,
                                                int h, int n, uint16_t off_w, uint16_t off_h) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};" ::"r"(smem_u32(smem_dst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c), "r"(w), "r"(h), "r"(n), "h"(off_w), "h"(off_h)
        : "memory");
}

template <int BN> struct StagesFor { static constexpr int value = BN >= 256 ? 4 : (BN >= 128 ? 6 : 8); };

template <int BN>
constexpr uint32_t gemm_smem_bytes() {
    return StagesFor<BN>::value * (A_STAGE_BYTES + BN * BK * 2) + 1024 /*align*/ + 256 /*barriers*/ + BIAS_SMEM_BYTES;
}

// MUFU tanh (one instruction, ~2^-11 relative error: below bf16 resolution).  tanhf() costs
// ~30 instructions per element and made the GELU epilogues slower than their mainloops.
__device__ __forceinline__ float tanh_fast(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float gelu_tanh(float z) {
    const float u = 0.7978845608028654f * fmaf(0.044715f * z * z, z, z);
    return 0.5f * z * (1.f + tanh_fast(u));
}
__device__ __forceinline__ float gelu_tanh_grad(float z) {
    const float z2 = z * z;
    const float u = 0.7978845608028654f * fmaf(0.044715f * z2, z, z);
    const float t = tanh_fast(u);
    return 0.5f * (1.f + t) + 0.5f * z * (1.f - t * t) * 0.7978845608028654f * fmaf(3.f * 0.044715f, z2, 1.f);
}

// Store NC consecutive columns of one row; vector path when aligned and in range.
template <int NC>
__device__ __forceinline__ void store_row(const GemmParams& p, int64_t d_off, int row, int n, const float (&f)[NC]) {
    if (p.d_fp32) {
        float* dst = reinterpret_cast<float*>(p.D) + d_off + (int64_t)row * p.ldd + n;
        if (n + NC <= p.N && (p.ldd & 3) == 0 && ((uintptr_t)p.D & 15) == 0) {
#pragma unroll
            for (int j = 0; j < NC; j += 4)
                *reinterpret_cast<float4*>(dst + j) = make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]);
        } else {
#pragma unroll
            for (int j = 0; j < NC; ++j)
                if (n + j < p.N) dst[j] = f[j];
        }
    } else {
        __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.D) + d_off + (int64_t)row * p.ldd + n;
        if (n + NC <= p.N && (p.ldd & 7) == 0 && ((uintptr_t)dst & 15) == 0) {
#pragma unroll
            for (int j = 0; j < NC; j += 8) {
                uint4 o;
                o.x = pack_bf16(f[j], f[j + 1]);
                o.y = pack_bf16(f[j + 2], f[j + 3]);
                o.z = pack_bf16(f[j + 4], f[j + 5]);
                o.w = pack_bf16(f[j + 6], f[j + 7]);
                *reinterpret_cast<uint4*>(dst + j) = o;
            }
        } else {
#pragma unroll
            for (int j = 0; j < NC; ++j)
                if (n + j < p.N) dst[j] = __float2bfloat16_rn(f[j]);
        }
    }
}

// Split-K partial: D (fp32, zeroed by the caller) += f, 16-byte vector atomics.
template <int NC>
__device__ __forceinline__ void add_row(const GemmParams& p, int64_t d_off, int row, int n, const float (&f)[NC]) {
    float* dst = reinterpret_cast<float*>(p.D) + d_off + (int64_t)row * p.ldd + n;
    if (n + NC <= p.N && (p.ldd & 3) == 0 && ((uintptr_t)p.D & 15) == 0) {
#pragma unroll
        for (int j = 0; j < NC; j += 4)
            atomicAdd(reinterpret_cast<float4*>(dst + j), make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]));
    } else {
#pragma unroll
        for (int j = 0; j < NC; ++j)
            if (n + j < p.N) atomicAdd(dst + j, f[j]);
    }
}

template <int NC>
__device__ __forceinline__ void load_aux_row(const __nv_bfloat16* aux, int64_t ld, int N, int row,
                                             int n, float (&a)[NC]) {
    const __nv_bfloat16* src = aux + (int64_t)row * ld + n;
    if (n + NC <= N && (ld & 7) == 0 && ((uintptr_t)aux & 15) == 0) {
#pragma unroll
        for (int j = 0; j < NC; j += 8) {
            const uint4 raw = *reinterpret_cast<const uint4*>(src + j);
            const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float2 t = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[q]));
                a[j + 2 * q] = t.x;
                a[j + 2 * q + 1] = t.y;
            }
        }
    } else {
#pragma unroll
        for (int j = 0; j < NC; ++j) a[j] = (n + j < N) ? __bfloat162float(src[j]) : 0.f;
    }
}

// One NC-column chunk of one accumulator row through the fused epilogue.
// The aux operand (residual / pre-activation) of one chunk: issued one chunk ahead of its use so
// the row-strided global loads are in flight while the previous chunk is processed.
template <int NC>
__device__ __forceinline__ void prefetch_aux(const GemmParams& p, bool vec_ok, int row, int n, uint4 (&raw)[NC / 8]) {
    if (vec_ok && n + NC <= p.N) {
        const uint4* src = reinterpret_cast<const uint4*>(p.aux + (int64_t)row * p.ldaux + n);
#pragma unroll
        for (int j = 0; j < NC / 8; ++j) raw[j] = src[j];
    }
}
template <int NC>
__device__ __forceinline__ void aux_values(const GemmParams& p, bool vec_ok, int row, int n,
                                           const uint4 (&raw)[NC / 8], float (&a)[NC]) {
    if (vec_ok && n + NC <= p.N) {
#pragma unroll
        for (int j = 0; j < NC / 8; ++j) {
            const uint32_t w[4] = {raw[j].x, raw[j].y, raw[j].z, raw[j].w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float2 t = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[q]));
                a[8 * j + 2 * q] = t.x;
                a[8 * j + 2 * q + 1] = t.y;
            }
        }
    } else {
        load_aux_row<NC>(p.aux, p.ldaux, p.N, row, n, a);
    }
}

// `sbias` points at this chunk's slice of the tile's bias staged in shared memory (zero beyond N).
template <int NC>
__device__ __forceinline__ void epilogue_chunk(const GemmParams& p, int row, bool row_ok, int n,
                                               float (&f)[NC], uint32_t step, const float* sbias,
                                               bool aux_vec, const uint4 (&araw)[NC / 8]) {
    if (sbias) {
#pragma unroll
        for (int j = 0; j < NC; j += 4) {
            const float4 b4 = *reinterpret_cast<const float4*>(sbias + j);
            f[j] += b4.x; f[j + 1] += b4.y; f[j + 2] += b4.z; f[j + 3] += b4.w;
        }
    }
    switch (p.epilogue) {
        case B200_EPI_NONE:
            if (p.aux && row_ok) {  // residual add
                float a[NC];
                aux_values<NC>(p, aux_vec, row, n, araw, a);
#pragma unroll
                for (int j = 0; j < NC; ++j) f[j] += a[j];
            }
            break;
        case B200_EPI_RELU_DROPOUT: {
#pragma unroll
            for (int j = 0; j < NC; ++j) f[j] = fmaxf(f[j], 0.f);
            if (p.dropout_p > 0.f) {
                const Philox rng(p.seed);
                const float keep_scale = 1.f / (1.f - p.dropout_p);
                const uint32_t thresh = (uint32_t)fminf(p.dropout_p * 4294967296.f, 4294967295.f);
                const uint64_t base = ((uint64_t)row * (uint64_t)p.N + (uint64_t)n) >> 2;
#pragma unroll
                for (int j = 0; j < NC; j += 4) {
                    // counter = element index / 4, stream = (step, layer): the mask is a pure
                    // function of (seed, step, layer, element) so no mask tensor is stored.
                    uint4 r = rng(base + (j >> 2), step * 64u + p.layer_id);
                    f[j] = r.x >= thresh ? f[j] * keep_scale : 0.f;
                    f[j + 1] = r.y >= thresh ? f[j + 1] * keep_scale : 0.f;
                    f[j + 2] = r.z >= thresh ? f[j + 2] * keep_scale : 0.f;
                    f[j + 3] = r.w >= thresh ? f[j + 3] * keep_scale : 0.f;
                }
            }
            break;
        }
        case B200_EPI_GATE_POSITIVE: {
            float a[NC];
            if (row_ok) aux_values<NC>(p, aux_vec, row, n, araw, a);
#pragma unroll
            for (int j = 0; j < NC; ++j) f[j] = (row_ok && a[j] > 0.f) ? f[j] * p.gate_scale : 0.f;
            break;
        }
        case B200_EPI_GELU: {
            if (p.aux_out && row_ok) {
                __nv_bfloat16* dst = p.aux_out + (int64_t)row * p.ldd + n;
                if (n + NC <= p.N && (p.ldd & 7) == 0 && ((uintptr_t)p.aux_out & 15) == 0) {
#pragma unroll
                    for (int j = 0; j < NC; j += 8) {
                        uint4 o;
                        o.x = pack_bf16(f[j], f[j + 1]);
                        o.y = pack_bf16(f[j + 2], f[j + 3]);
                        o.z = pack_bf16(f[j + 4], f[j + 5]);
                        o.w = pack_bf16(f[j + 6], f[j + 7]);
                        *reinterpret_cast<uint4*>(dst + j) = o;
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < NC; ++j)
                        if (n + j < p.N) dst[j] = __float2bfloat16_rn(f[j]);
                }
            }
#pragma unroll
            for (int j = 0; j < NC; ++j) f[j] = gelu_tanh(f[j]);
            break;
        }
        case B200_EPI_GELU_BWD: {
            float a[NC];
            if (row_ok) aux_values<NC>(p, aux_vec, row, n, araw, a);
#pragma unroll
            for (int j = 0; j < NC; ++j) f[j] = row_ok ? f[j] * gelu_tanh_grad(a[j]) : 0.f;
            break;
        }
        default:
            break;
    }
}

// Softmax cross-entropy on a full logits row held by one thread (N <= NC).
template <int NC>
__device__ __forceinline__ void epilogue_ce(const GemmParams& p, int row, bool row_ok, float (&f)[NC]) {
    float mx = -INFINITY;
    int arg = 0;
#pragma unroll
    for (int j = 0; j < NC; ++j) {
        if (j < p.N) {
            f[j] += p.bias ? __ldg(p.bias + j) : 0.f;
            if (f[j] > mx) { mx = f[j]; arg = j; }
        }
    }
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < NC; ++j)
        if (j < p.N) sum += __expf(f[j] - mx);
    const int tgt = row_ok ? (int)p.targets[row] : 0;
    float zt = 0.f;
#pragma unroll
    for (int j = 0; j < NC; ++j)
        if (j == tgt) zt = f[j];
    float loss = row_ok ? (mx + __logf(sum) - zt) : 0.f;
    float ok = (row_ok && arg == tgt) ? 1.f : 0.f;
    loss = warp_sum(loss);
    ok = warp_sum(ok);
    if ((threadIdx.x & 31) == 0 && p.loss_out) {
        atomicAdd(p.loss_out, loss * p.loss_scale);
        atomicAdd(p.loss_out + 1, ok);
    }
    if (p.ce_write_grad) {
        const float inv = 1.f / sum;
#pragma unroll
        for (int j = 0; j < NC; ++j) {
            float g = (j < p.N) ? __expf(f[j] - mx) * inv : 0.f;
            if (j == tgt) g -= 1.f;
            f[j] = row_ok ? g * p.ce_grad_scale : 0.f;
        }
    }
}

template <int NC>
__device__ __forceinline__ void colsum_chunk(const GemmParams& p, bool row_ok, int n, const float (&f)[NC]) {
    // Sum over the warp's 32 rows with a transposing butterfly: NC-1 shuffles instead of
    // 5*NC, after which lane L holds the total of column L; then ONE coalesced atomic
    // instruction per warp.
    const int lane = threadIdx.x & 31;
    float v[NC];
#pragma unroll
    for (int j = 0; j < NC; ++j) v[j] = row_ok ? f[j] : 0.f;
#pragma unroll
    for (int o = NC / 2; o >= 1; o >>= 1) {
        const bool upper = (lane & o) != 0;
#pragma unroll
        for (int k = 0; k < o; ++k) {
            const float send = upper ? v[k] : v[k + o];
            const float keep = upper ? v[k + o] : v[k];
            v[k] = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
    }
#pragma unroll
    for (int o = NC; o < 32; o <<= 1) v[0] += __shfl_xor_sync(0xffffffffu, v[0], o);
    const int col = lane % NC;
    if (lane < NC && n + col < p.N) atomicAdd(p.colsum + n + col, v[0]);
}

// Epilogue of columns [c0, c1) x NC of one 128-row accumulator: TMEM -> registers -> fused
// epilogue -> global.  One thread owns one row.
template <int NC>
__device__ __forceinline__ void drain_columns(const GemmParams& p, uint32_t taddr, int row, bool row_ok, int n0,
                                              int c0, int c1, int64_t d_off, uint32_t step, const float* sbias) {
    const bool aux_vec = p.aux && row_ok && (p.ldaux & 7) == 0 && ((uintptr_t)p.aux & 15) == 0;
    uint4 araw[NC / 8], anext[NC / 8];
#pragma unroll
    for (int j = 0; j < NC / 8; ++j) araw[j] = anext[j] = make_uint4(0, 0, 0, 0);
    if (c0 < c1) prefetch_aux<NC>(p, aux_vec, row, n0 + c0 * NC, araw);
#pragma unroll 1
    for (int c = c0; c < c1; ++c) {
        uint32_t v[NC];
        tmem_ld_chunk(taddr + c * NC, v);
        if (c + 1 < c1) prefetch_aux<NC>(p, aux_vec, row, n0 + (c + 1) * NC, anext);
        tmem_ld_wait();
        float f[NC];
#pragma unroll
        for (int j = 0; j < NC; ++j) f[j] = __uint_as_float(v[j]);
        const int n = n0 + c * NC;
        if (n < p.N) {
            if (p.k_splits > 1) {  // partial sum of one K slice
                if (row_ok) add_row<NC>(p, d_off, row, n, f);
            } else {
                if (p.epilogue == B200_EPI_SOFTMAX_CE) epilogue_ce<NC>(p, row, row_ok, f);
                else epilogue_chunk<NC>(p, row, row_ok, n, f, step, sbias ? sbias + c * NC : nullptr, aux_vec, araw);
                if (p.colsum) colsum_chunk<NC>(p, row_ok, n, f);
                if (row_ok) store_row<NC>(p, d_off, row, n, f);
            }
        }
#pragma unroll
        for (int j = 0; j < NC / 8; ++j) araw[j] = anext[j];
    }
}

// The 256 epilogue threads stage the tile's bias slice (zero beyond N) and meet on a named
// barrier.  Done before they wait for the accumulator, so the loads hide behind the tile's MMAs.
template <int BN, int EPI_THREADS>
__device__ __forceinline__ void stage_bias(const GemmParams& p, float* sb, int n0, int et) {
    for (int j = et; j < BN; j += EPI_THREADS) sb[j] = (n0 + j < p.N) ? __ldg(p.bias + n0 + j) : 0.f;
    asm volatile("bar.sync 1, %0;" ::"n"(EPI_THREADS) : "memory");
}

// ---- FAST epilogue: everything per-tile is hoisted into registers by the caller (row pointers,
// mode), the chunk loop touches only TMEM, the staged bias and the row's own global addresses.
__device__ __forceinline__ void unpack8(const uint4& raw, float* a) {
    const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const float2 t = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[q]));
        a[2 * q] = t.x;
        a[2 * q + 1] = t.y;
    }
}
// 32 columns of one row as bf16; N is a multiple of 8 or the tail is written element-wise.
__device__ __forceinline__ void store32_bf16(__nv_bfloat16* dst, int n, int N, const float (&f)[32]) {
#pragma unroll
    for (int j = 0; j < 32; j += 8) {
        if (n + j + 8 <= N) {
            uint4 o;
            o.x = pack_bf16(f[j], f[j + 1]);
            o.y = pack_bf16(f[j + 2], f[j + 3]);
            o.z = pack_bf16(f[j + 4], f[j + 5]);
            o.w = pack_bf16(f[j + 6], f[j + 7]);
            *reinterpret_cast<uint4*>(dst + n + j) = o;
        } else {
#pragma unroll
            for (int q = 0; q < 8; ++q)
                if (n + j + q < N) dst[n + j + q] = __float2bfloat16_rn(f[j + q]);
        }
    }
}
// The same 32 columns, written so that every instruction covers whole 32-byte sectors: a thread's
// row segment is four 16-byte pieces, and one piece per lane per instruction leaves each sector
// half-written (measured: output-heavy GEMMs capped near 1.3 TB/s of stores).  Lanes 2i and 2i+1
// swap two pieces each, then write the two halves of one sector of row 2i, then of row 2i+1.
// `prow` is the partner lane's row pointer.  Whole chunk in range only (n + 32 <= N).
__device__ __forceinline__ uint4 shfl_xor1(const uint4& v) {
    uint4 r;
    r.x = __shfl_xor_sync(0xffffffffu, v.x, 1);
    r.y = __shfl_xor_sync(0xffffffffu, v.y, 1);
    r.z = __shfl_xor_sync(0xffffffffu, v.z, 1);
    r.w = __shfl_xor_sync(0xffffffffu, v.w, 1);
    return r;
}
__device__ __forceinline__ void store32_pair(__nv_bfloat16* drow, __nv_bfloat16* prow, int n, bool odd,
                                             const float (&f)[32]) {
    uint4 c[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        c[j].x = pack_bf16(f[8 * j], f[8 * j + 1]);
        c[j].y = pack_bf16(f[8 * j + 2], f[8 * j + 3]);
        c[j].z = pack_bf16(f[8 * j + 4], f[8 * j + 5]);
        c[j].w = pack_bf16(f[8 * j + 6], f[8 * j + 7]);
    }
    // even lane keeps pieces 0,2 of its row and receives 0,2 of the odd row;
    // odd lane keeps pieces 1,3 of its row and receives 1,3 of the even row
    const uint4 ra = shfl_xor1(odd ? c[0] : c[1]);
    const uint4 rb = shfl_xor1(odd ? c[2] : c[3]);
    __nv_bfloat16* row_lo = odd ? prow : drow;   // row 2i
    __nv_bfloat16* row_hi = odd ? drow : prow;   // row 2i+1
    const int off = n + (odd ? 8 : 0);
    const uint4 lo_a = odd ? ra : c[0], lo_b = odd ? rb : c[2];
    const uint4 hi_a = odd ? c[1] : ra, hi_b = odd ? c[3] : rb;
    if (row_lo) {
        *reinterpret_cast<uint4*>(row_lo + off) = lo_a;
        *reinterpret_cast<uint4*>(row_lo + off + 16) = lo_b;
    }
    if (row_hi) {
        *reinterpret_cast<uint4*>(row_hi + off) = hi_a;
        *reinterpret_cast<uint4*>(row_hi + off + 16) = hi_b;
    }
}
__device__ __forceinline__ __nv_bfloat16* partner_ptr(__nv_bfloat16* ptr) {
    const unsigned long long v = __shfl_xor_sync(0xffffffffu, (unsigned long long)ptr, 1);
    return reinterpret_cast<__nv_bfloat16*>(v);
}
__device__ __forceinline__ void fetch32(const __nv_bfloat16* arow, int n, int N, uint4 (&r)[4]) {
    if (arow) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (n + 8 * j + 8 <= N) r[j] = *reinterpret_cast<const uint4*>(arow + n + 8 * j);
    }
}
// drow / arow / zrow: this thread's row of D, aux and aux_out (null when the row is out of range or
// the operand is absent).  mode: B200_EPI_NONE (+ residual when arow), _GELU, _GELU_BWD.
__device__ __forceinline__ void fast_drain(uint32_t taddr, int n0, int N, int c0, int c1, int mode,
                                           __nv_bfloat16* drow, const __nv_bfloat16* arow,
                                           __nv_bfloat16* zrow, const float* sbias) {
    const bool odd = (threadIdx.x & 1) != 0;
    __nv_bfloat16* dprow = partner_ptr(drow);
    __nv_bfloat16* zprow = partner_ptr(zrow);
    uint4 araw[4], anext[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) araw[j] = anext[j] = make_uint4(0, 0, 0, 0);
    if (c0 < c1) fetch32(arow, n0 + c0 * 32, N, araw);
#pragma unroll 1
    for (int c = c0; c < c1; ++c) {
        const int n = n0 + c * 32;
        uint32_t v[32];
        tmem_ld_32x32(taddr + c * 32, v);
        if (c + 1 < c1) fetch32(arow, n + 32, N, anext);
        tmem_ld_wait();
        if (n < N) {
            float f[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
            if (sbias) {
#pragma unroll
                for (int j = 0; j < 32; j += 4) {
                    const float4 b4 = *reinterpret_cast<const float4*>(sbias + c * 32 + j);
                    f[j] += b4.x; f[j + 1] += b4.y; f[j + 2] += b4.z; f[j + 3] += b4.w;
                }
            }
            if (mode == B200_EPI_GELU) {
                // pre-activation, for the backward gate
                if (n + 32 <= N) store32_pair(zrow, zprow, n, odd, f);
                else if (zrow) store32_bf16(zrow, n, N, f);
#pragma unroll
                for (int j = 0; j < 32; ++j) f[j] = gelu_tanh(f[j]);
            } else if (arow) {
                float a[32];
#pragma unroll
                for (int j = 0; j < 4; ++j) unpack8(araw[j], a + 8 * j);
                if (mode == B200_EPI_GELU_BWD) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) f[j] *= gelu_tanh_grad(a[j]);
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) f[j] += a[j];
                }
            }
            if (n + 32 <= N) store32_pair(drow, dprow, n, odd, f);
            else if (drow) store32_bf16(drow, n, N, f);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) araw[j] = anext[j];
    }
}

struct TileInfo {
    int m0, n0, kb0, kb1, hb, bb;
    bool skip;
};
// tile index -> (batch, K slice, output tile) and the k-block range it covers.  Causal modes
// (attention): 1 = the output is lower-triangular, tiles strictly above the diagonal are not
// computed; 2 = A is lower-triangular [rows, k <= row]; 3 = A^T is (rows k >= m0).
template <int BN>
__device__ __forceinline__ TileInfo decode_tile(const GemmParams& p, int tile) {
    TileInfo t;
    const int mn_tiles = p.num_m_blocks * p.num_n_blocks;
    const int per_batch = mn_tiles * p.k_splits;
    const int bt = tile / per_batch;
    const int r = tile - bt * per_batch;
    const int ks = r / mn_tiles, mn = r - ks * mn_tiles;
    t.m0 = (mn % p.num_m_blocks) * BM;
    t.n0 = (mn / p.num_m_blocks) * BN;
    t.hb = bt % p.batch_inner;
    t.bb = bt / p.batch_inner;
    t.kb0 = ks * p.kb_per_split;
    t.kb1 = min(p.num_k_blocks, t.kb0 + p.kb_per_split);
    t.skip = false;
    if (p.causal == 1) t.skip = t.n0 >= t.m0 + BM;
    else if (p.causal == 2) t.kb1 = min(t.kb1, (t.m0 + BM + BK - 1) / BK);
    else if (p.causal == 3) t.kb0 = max(t.kb0, t.m0 / BK);
    return t;
}

template <int BN, bool FAST>
__global__ void __launch_bounds__(gemm_threads<FAST>(), 1)
gemm_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
            const GemmParams p) {
    constexpr int STAGES = StagesFor<BN>::value;
    constexpr uint32_t B_STAGE_BYTES = BN * BK * 2;
    constexpr uint32_t TMEM_COLS = (2 * BN) < 32 ? 32 : 2 * BN;
    constexpr int NC = BN < 32 ? BN : 32;  // epilogue chunk width

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + STAGES * A_STAGE_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_b + STAGES * B_STAGE_BYTES);
    uint64_t* full_bar = bars;
    uint64_t* empty_bar = bars + STAGES;
    uint64_t* tmem_full = bars + 2 * STAGES;
    uint64_t* tmem_empty = bars + 2 * STAGES + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);
    float* sbias = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 256);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int num_tiles = p.num_m_blocks * p.num_n_blocks * p.k_splits * p.batch_inner * p.batch_outer;
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tma_a);
        tma_prefetch_desc(&tma_b);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tmem_full[s], 1);
            mbar_init(&tmem_empty[s], EpiWarps<FAST>::value);
        }
        fence_mbar_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, TMEM_COLS);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // Everything above touched only shared memory, TMEM and kernel parameters.  From here
    // on we read what earlier kernels wrote: wait for them (no-op without PDL).
    asm volatile("griddepcontrol.wait;" ::: "memory");

    // Producer and MMA loops are walked by their WHOLE warp with one elected lane issuing: control
    // flow and every address/descriptor stay warp-uniform, so the compiler keeps them in uniform
    // registers and issues UTMALDG / UTCHMMA directly.  (A `lane == 0` branch made it wrap each such
    // instruction in an elect/broadcast loop: ~75 instructions per k-block on one thread, and the
    // mainloop was bound by MMA *issue*, ~800 cycles per k-block against 540 of tensor work.)
    if (warp == 0) {
        // ===================== TMA producer =====================
        uint32_t stage = 0, phase = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            const TileInfo t = decode_tile<BN>(p, tile);
            if (t.skip) continue;
            const int m0 = t.m0, n0 = t.n0;
            for (int kb = t.kb0; kb < t.kb1; ++kb) {
                mbar_wait(&empty_bar[stage], phase ^ 1);
                if (elect_one()) {
                    mbar_arrive_expect_tx(&full_bar[stage], A_STAGE_BYTES + B_STAGE_BYTES);
                    uint8_t* sa = smem_a + stage * A_STAGE_BYTES;
                    uint8_t* sb = smem_b + stage * B_STAGE_BYTES;
                    const int k0 = kb * BK;
                    if (p.a_conv) {          // rows = 128 output pixels from m0, k-block = 64 channels of one tap
                        int cn, ch, cw, cc;
                        uint16_t fs, fr;
                        conv_origin(p, m0, cn, ch, cw);
                        conv_tap(p, kb, cc, fs, fr);
                        tma_load_im2col(sa, &tma_a, &full_bar[stage], cc, cw, ch, cn, fs, fr);
                    } else if (!p.a_mn) {
                        tma_load_4d(sa, &tma_a, &full_bar[stage], k0, m0, t.hb, t.bb);
                    } else {
#pragma unroll
                        for (int j = 0; j < BM / 64; ++j)
                            tma_load_4d(sa + j * 8192, &tma_a, &full_bar[stage], m0 + j * 64, k0, t.hb, t.bb);
                    }
                    if (p.b_conv) {          // MN-major B: 64-column chunks of the im2col matrix x 64 pixels from k0
                        int cn, ch, cw;
                        conv_origin(p, k0, cn, ch, cw);
#pragma unroll
                        for (int j = 0; j < (BN >= 64 ? BN / 64 : 1); ++j) {
                            int cc;
                            uint16_t fs, fr;
                            conv_tap(p, n0 / 64 + j, cc, fs, fr);
                            tma_load_im2col(sb + j * 8192, &tma_b, &full_bar[stage], cc, cw, ch, cn, fs, fr);
                        }
                    } else if (!p.b_mn) {
                        tma_load_4d(sb, &tma_b, &full_bar[stage], k0, n0, t.hb, t.bb);
                    } else {
#pragma unroll
                        for (int j = 0; j < (BN >= 64 ? BN / 64 : 1); ++j)
                            tma_load_4d(sb + j * 8192, &tma_b, &full_bar[stage], n0 + j * 64, k0, t.hb, t.bb);
                    }
                }
                __syncwarp();
                if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        uint32_t stage = 0, phase = 0;
        int it = 0;
        // K-major: 8-row groups 1024 B apart, K advances 32 B inside the swizzle row.
        // MN-major: 64-element chunks 8192 B apart (LBO), 8-k groups 1024 B apart (SBO),
        //           K advances 16 rows = 2048 B.
        // Descriptors differ only in their 14-bit start-address field (bytes >> 4): build one per
        // operand and add stage / k-step offsets to it.
        const uint64_t da0 = umma_smem_desc(smem_u32(smem_a), p.a_mn ? 8192u : 16u, 1024u);
        const uint64_t db0 = umma_smem_desc(smem_u32(smem_b), p.b_mn ? 8192u : 16u, 1024u);
        const uint32_t a_k16 = (p.a_mn ? 2048u : 32u) >> 4, b_k16 = (p.b_mn ? 2048u : 32u) >> 4;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            const TileInfo t = decode_tile<BN>(p, tile);
            if (t.skip) continue;
            const uint32_t as = it & 1;
            mbar_wait(&tmem_empty[as], ((it >> 1) & 1) ^ 1);
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + as * BN;
            const int kb0 = t.kb0, kb1 = t.kb1;
            for (int kb = kb0; kb < kb1; ++kb) {
                mbar_wait(&full_bar[stage], phase);
                tc_fence_after();
                const uint64_t da = da0 + (uint64_t)(stage * (A_STAGE_BYTES >> 4));
                const uint64_t db = db0 + (uint64_t)(stage * (B_STAGE_BYTES >> 4));
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k)
                        umma_bf16(d_tmem, da + (uint64_t)(k * a_k16), db + (uint64_t)(k * b_k16), p.idesc,
                                  (kb > kb0 || k > 0) ? 1u : 0u);
                    umma_commit(&empty_bar[stage]);  // slot reusable once these MMAs retire
                    if (kb == kb1 - 1) umma_commit(&tmem_full[as]);  // accumulator complete
                }
                __syncwarp();
                if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
            ++it;
        }
    } else {
        // ===================== epilogue =====================
        // General kernels: warps 2..5, one per TMEM lane quarter.  FAST kernels: warps 2..9, two
        // per quarter, each draining half of the accumulator's columns.
        constexpr int EPI_THREADS = 32 * EpiWarps<FAST>::value;
        const int quarter = warp & 3;  // TMEM lanes this warp may read
        const int half = (warp - 2) >> 2;
        const int et = threadIdx.x - 64;
        constexpr int CH = BN / NC;
        const int c0 = FAST ? half * (CH / 2) : 0;
        const int c1 = FAST ? c0 + CH / 2 : CH;
        const bool use_sbias = p.bias && p.epilogue != B200_EPI_SOFTMAX_CE;
        const int mode = p.epilogue, N = p.N;
        uint32_t step = 0;
        if (p.step_dev) step = (uint32_t)__ldg(p.step_dev);
        int it = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            const TileInfo t = decode_tile<BN>(p, tile);
            if (t.skip) continue;
            const int m0 = t.m0, n0 = t.n0;
            const int64_t d_off = t.hb * p.d_stride_inner + t.bb * p.d_stride_outer;
            const uint32_t as = it & 1;
            float* sb = sbias + as * 256;
            if (use_sbias) stage_bias<BN, EPI_THREADS>(p, sb, n0, et);
            const int row = m0 + quarter * 32 + lane;
            const bool row_ok = row < p.M;
            const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + as * BN;
            if constexpr (FAST) {
                __nv_bfloat16* drow = row_ok ? reinterpret_cast<__nv_bfloat16*>(p.D) + d_off + (int64_t)row * p.ldd : nullptr;
                const __nv_bfloat16* arow = (row_ok && p.aux) ? p.aux + (int64_t)row * p.ldaux : nullptr;
                __nv_bfloat16* zrow = (row_ok && p.aux_out) ? p.aux_out + (int64_t)row * p.ldd : nullptr;
                mbar_wait(&tmem_full[as], (it >> 1) & 1);
                tc_fence_after();
                fast_drain(taddr, n0, N, c0, c1, mode, drow, arow, zrow, use_sbias ? sb : nullptr);
            } else {
                mbar_wait(&tmem_full[as], (it >> 1) & 1);
                tc_fence_after();
                drain_columns<NC>(p, taddr, row, row_ok, n0, c0, c1, d_off, step, use_sbias ? sb : nullptr);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty[as]);
            ++it;
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// ---------------------------------------------------------------- 256 x 256 tiles
// Large-GEMM variant.  Measured on B200 the 128x256 kernel is bound by what one SM can pull
// from L2 (48 KB of operands per 540 MMA cycles = 89 B/clk; 128-wide tiles are slower still),
// not by the tensor pipe.  This kernel gives each CTA TWO M=128 accumulators that share every B
// tile — a 256x256 output tile, 64 KB per 1080 MMA cycles = 59 B/clk — without clusters: both
// accumulators live in TMEM (2 x 256 columns = all of it), the MMA thread interleaves them per
// k-block, and the epilogue drains the upper half while the lower half's last MMAs retire and
// the next tile's MMAs refill the upper half.  Plain (non-batched, non-causal) problems only.
constexpr int BM2 = 256;
constexpr int STAGES2 = 3;
constexpr uint32_t A2_STAGE_BYTES = BM2 * BK * 2;   // 32 KB
constexpr uint32_t B2_STAGE_BYTES = 256 * BK * 2;   // 32 KB
constexpr uint32_t GEMM2_SMEM = STAGES2 * (A2_STAGE_BYTES + B2_STAGE_BYTES) + 1024 + 256 + BIAS_SMEM_BYTES;

__global__ void __launch_bounds__(gemm_threads<true>(), 1)
gemm_kernel_m256(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                 const GemmParams p) {
    constexpr int BN = 256;
    constexpr int NC = 32;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + STAGES2 * A2_STAGE_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_b + STAGES2 * B2_STAGE_BYTES);
    uint64_t* full_bar = bars;
    uint64_t* empty_bar = bars + STAGES2;
    uint64_t* tmem_full = bars + 2 * STAGES2;       // [2]: upper / lower accumulator complete
    uint64_t* tmem_empty = bars + 2 * STAGES2 + 2;  // [2]: upper / lower accumulator drained
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES2 + 4);
    float* sbias = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 256);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int num_tiles = p.num_m_blocks * p.num_n_blocks;
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tma_a);
        tma_prefetch_desc(&tma_b);
        for (int s = 0; s < STAGES2; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tmem_full[s], 1);
            mbar_init(&tmem_empty[s], 8);
        }
        fence_mbar_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    asm volatile("griddepcontrol.wait;" ::: "memory");

    if (warp == 0) {
        // TMA producer: whole warp walks the loop, one elected lane issues (see gemm_kernel)
        uint32_t stage = 0, phase = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            const int m0 = (tile % p.num_m_blocks) * BM2;
            const int n0 = (tile / p.num_m_blocks) * BN;
            for (int kb = 0; kb < p.num_k_blocks; ++kb) {
                mbar_wait(&empty_bar[stage], phase ^ 1);
                if (elect_one()) {
                    mbar_arrive_expect_tx(&full_bar[stage], A2_STAGE_BYTES + B2_STAGE_BYTES);
                    uint8_t* sa = smem_a + stage * A2_STAGE_BYTES;
                    uint8_t* sb = smem_b + stage * B2_STAGE_BYTES;
                    const int k0 = kb * BK;
                    if (!p.a_mn) {
                        tma_load_4d(sa, &tma_a, &full_bar[stage], k0, m0, 0, 0);  // one 256-row box
                    } else {
#pragma unroll
                        for (int j = 0; j < BM2 / 64; ++j)
                            tma_load_4d(sa + j * 8192, &tma_a, &full_bar[stage], m0 + j * 64, k0, 0, 0);
                    }
                    if (!p.b_mn) {
                        tma_load_4d(sb, &tma_b, &full_bar[stage], k0, n0, 0, 0);
                    } else {
#pragma unroll
                        for (int j = 0; j < BN / 64; ++j)
                            tma_load_4d(sb + j * 8192, &tma_b, &full_bar[stage], n0 + j * 64, k0, 0, 0);
                    }
                }
                __syncwarp();
                if (++stage == STAGES2) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // MMA issuer: whole warp, one elected lane
        uint32_t stage = 0, phase = 0;
        int it = 0;
        const uint64_t da0 = umma_smem_desc(smem_u32(smem_a), p.a_mn ? 8192u : 16u, 1024u);
        const uint64_t db0 = umma_smem_desc(smem_u32(smem_b), p.b_mn ? 8192u : 16u, 1024u);
        const uint32_t a_k16 = (p.a_mn ? 2048u : 32u) >> 4, b_k16 = (p.b_mn ? 2048u : 32u) >> 4;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
            const uint32_t par = (uint32_t)(it & 1);
            for (int kb = 0; kb < p.num_k_blocks; ++kb) {
                mbar_wait(&full_bar[stage], phase);
                tc_fence_after();
                const uint64_t da = da0 + (uint64_t)(stage * (A2_STAGE_BYTES >> 4));
                const uint64_t db = db0 + (uint64_t)(stage * (B2_STAGE_BYTES >> 4));
                const bool last = kb == p.num_k_blocks - 1;
                if (elect_one()) {
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        if (kb == 0) {  // this half of TMEM must have been drained (previous tile)
                            mbar_wait(&tmem_empty[h], par ^ 1);
                            tc_fence_after();
                        }
                        // rows 128..255 of the A stage start 16 KB in (K-major: 128 rows x 128 B;
                        // MN-major: two 8 KB chunks)
#pragma unroll
                        for (int k = 0; k < BK / 16; ++k)
                            umma_bf16(tmem_base + h * BN, da + (uint64_t)(h * (16384 >> 4) + k * a_k16),
                                      db + (uint64_t)(k * b_k16), p.idesc, (kb | k) != 0 ? 1u : 0u);
                        if (last) umma_commit(&tmem_full[h]);  // upper half completes first
                    }
                    umma_commit(&empty_bar[stage]);
                }
                __syncwarp();
                if (++stage == STAGES2) { stage = 0; phase ^= 1; }
            }
        }
    } else {
        // FAST epilogue only (the host routes everything else to the 128-row kernel)
        const int quarter = warp & 3;
        const int half = (warp - 2) >> 2;   // which 128 columns of each accumulator this warp drains
        const int et = threadIdx.x - 64;
        constexpr int CH = BN / NC;
        const int mode = p.epilogue, N = p.N;
        int it = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
            const int m0 = (tile % p.num_m_blocks) * BM2;
            const int n0 = (tile / p.num_m_blocks) * BN;
            const uint32_t par = (uint32_t)(it & 1);
            float* sb = sbias + par * 256;
            if (p.bias) stage_bias<BN, 256>(p, sb, n0, et);
#pragma unroll 1
            for (int h = 0; h < 2; ++h) {
                const int row = m0 + h * 128 + quarter * 32 + lane;
                const bool row_ok = row < p.M;
                __nv_bfloat16* drow = row_ok ? reinterpret_cast<__nv_bfloat16*>(p.D) + (int64_t)row * p.ldd : nullptr;
                const __nv_bfloat16* arow = (row_ok && p.aux) ? p.aux + (int64_t)row * p.ldaux : nullptr;
                __nv_bfloat16* zrow = (row_ok && p.aux_out) ? p.aux_out + (int64_t)row * p.ldd : nullptr;
                mbar_wait(&tmem_full[h], par);
                tc_fence_after();
                const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + h * BN;
                fast_drain(taddr, n0, N, half * (CH / 2), (half + 1) * (CH / 2), mode, drow, arow, zrow,
                           p.bias ? sb : nullptr);
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&tmem_empty[h]);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ---------------------------------------------------------------- two-SM tiles (cta_group::2)
// One 256x256 output tile per CTA PAIR (a 2-CTA cluster = the two SMs of a TPC).  Each CTA stages
// its own 128 rows of A and only HALF of B (128 of the 256 columns): the pair's tensor cores read
// both halves, so an SM pulls 32 KB per k-block from L2 instead of the 48 KB of the single-SM
// 128x256 kernel -- the mainloop of that kernel is bound by exactly this operand feed -- while each
// SM keeps a double-buffered 128x256 fp32 accumulator (2 x 256 TMEM columns) so the epilogue of
// tile i still overlaps the MMAs of tile i+1 (the 256x256 single-SM kernel cannot: its two
// accumulators fill TMEM).
//   * both CTAs run a TMA producer; every load completes on the LEADER's (rank 0) full barrier
//   * the leader's MMA thread issues tcgen05.mma.cta_group::2 (M = 256) for both SMs; its commits
//     are multicast to both CTAs' barriers (ring slot free / accumulator complete)
//   * every epilogue warp of both CTAs arrives on the leader's tmem_empty barrier
// Pair tiles are 256 x BN with BN = 256 or 128 (the narrower one for problems with few tiles:
// weight gradients, N = 768 outputs).
template <int BN> struct Stages2sm { static constexpr int value = BN >= 256 ? 6 : 8; };
template <int BN>
constexpr uint32_t gemm_2sm_smem() {
    return Stages2sm<BN>::value * (A_STAGE_BYTES + (BN / 2) * BK * 2) + 1024 + 256 + BIAS_SMEM_BYTES;
}

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `smem_addr` (a shared::cta address) in CTA `rank` of this cluster
__device__ __forceinline__ uint32_t mapa_u32(uint32_t smem_addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void tma_load_4d_pair(void* smem_dst, const CUtensorMap* map, uint32_t leader_bar,
                                                 int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(smem_u32(smem_dst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(leader_bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
__device__ __forceinline__ void tma_load_im2col_pair(void* smem_dst, const CUtensorMap* map, uint32_t leader_bar, int c,
                                                     int w, int h, int n, uint16_t off_w, uint16_t off_h) {
    asm volatile(
        "cp.async.bulk.tensor.4d.im2col.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};" ::"r"(smem_u32(smem_dst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(leader_bar), "r"(c), "r"(w), "r"(h), "r"(n), "h"(off_w), "h"(off_h)
        : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc,
                                               uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrives on the barrier at this offset in BOTH CTAs of the pair once all earlier MMAs retired
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
    asm volatile(
        "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
            smem_u32(bar)),
        "h"((uint16_t)3)
        : "memory");
}

template <int BN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(gemm_threads<true>(), 1)
gemm_kernel_2sm(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                const GemmParams p) {
    constexpr int STAGES = Stages2sm<BN>::value;
    constexpr uint32_t B_HALF_BYTES = (BN / 2) * BK * 2;   // this CTA's half of the B tile
    constexpr uint32_t TMEM_COLS = 2 * BN;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + STAGES * A_STAGE_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_b + STAGES * B_HALF_BYTES);
    uint64_t* full_bar = bars;                   // used in the leader only
    uint64_t* empty_bar = bars + STAGES;         // per CTA (multicast commit)
    uint64_t* tmem_full = bars + 2 * STAGES;     // per CTA (multicast commit)
    uint64_t* tmem_empty = bars + 2 * STAGES + 2;  // leader only: 16 epilogue warps arrive
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);
    float* sbias = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 256);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
    const int num_tiles = p.num_m_blocks * p.num_n_blocks;   // 256 x 256 tiles
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tma_a);
        tma_prefetch_desc(&tma_b);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tmem_full[s], 1);
            mbar_init(&tmem_empty[s], 2 * EpiWarps<true>::value);
        }
        fence_mbar_init();
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(TMEM_COLS)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();   // the peer's barriers are initialised before anything is signalled across
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    asm volatile("griddepcontrol.wait;" ::: "memory");

    if (warp == 0) {
        // ===================== TMA producer (both CTAs; whole warp, one elected lane) =====================
        uint32_t stage = 0, phase = 0;
        for (int tile = pair; tile < num_tiles; tile += num_pairs) {
            const int m0 = (tile % p.num_m_blocks) * 256 + (int)rank * 128;   // this CTA's A rows
            const int n0 = (tile / p.num_m_blocks) * BN + (int)rank * (BN / 2);  // this CTA's half of B
            for (int kb = 0; kb < p.num_k_blocks; ++kb) {
                mbar_wait(&empty_bar[stage], phase ^ 1);
                if (elect_one()) {
                    const uint32_t lbar = mapa_u32(smem_u32(&full_bar[stage]), 0);
                    if (rank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * (A_STAGE_BYTES + B_HALF_BYTES));
                    uint8_t* sa = smem_a + stage * A_STAGE_BYTES;
                    uint8_t* sb = smem_b + stage * B_HALF_BYTES;
                    const int k0 = kb * BK;
                    if (p.a_conv) {
                        int cn, ch, cw, cc;
                        uint16_t fs, fr;
                        conv_origin(p, m0, cn, ch, cw);
                        conv_tap(p, kb, cc, fs, fr);
                        tma_load_im2col_pair(sa, &tma_a, lbar, cc, cw, ch, cn, fs, fr);
                    } else if (!p.a_mn) {
                        tma_load_4d_pair(sa, &tma_a, lbar, k0, m0, 0, 0);
                    } else {
#pragma unroll
                        for (int j = 0; j < 2; ++j) tma_load_4d_pair(sa + j * 8192, &tma_a, lbar, m0 + j * 64, k0, 0, 0);
                    }
                    if (!p.b_mn) {
                        tma_load_4d_pair(sb, &tma_b, lbar, k0, n0, 0, 0);
                    } else {
#pragma unroll
                        for (int j = 0; j < BN / 128; ++j) tma_load_4d_pair(sb + j * 8192, &tma_b, lbar, n0 + j * 64, k0, 0, 0);
                    }
                }
                __syncwarp();
                if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (leader CTA only; whole warp, one elected lane) =====================
        if (rank == 0) {
            uint32_t stage = 0, phase = 0;
            int it = 0;
            const uint64_t da0 = umma_smem_desc(smem_u32(smem_a), p.a_mn ? 8192u : 16u, 1024u);
            const uint64_t db0 = umma_smem_desc(smem_u32(smem_b), p.b_mn ? 8192u : 16u, 1024u);
            const uint32_t a_k16 = (p.a_mn ? 2048u : 32u) >> 4, b_k16 = (p.b_mn ? 2048u : 32u) >> 4;
            for (int tile = pair; tile < num_tiles; tile += num_pairs, ++it) {
                const uint32_t as = it & 1;
                mbar_wait(&tmem_empty[as], ((it >> 1) & 1) ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * BN;
                for (int kb = 0; kb < p.num_k_blocks; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint64_t da = da0 + (uint64_t)(stage * (A_STAGE_BYTES >> 4));
                    const uint64_t db = db0 + (uint64_t)(stage * (B_HALF_BYTES >> 4));
                    if (elect_one()) {
#pragma unroll
                        for (int k = 0; k < BK / 16; ++k)
                            umma_bf16_pair(d_tmem, da + (uint64_t)(k * a_k16), db + (uint64_t)(k * b_k16), p.idesc,
                                           (kb | k) != 0 ? 1u : 0u);
                        umma_commit_pair(&empty_bar[stage]);
                        if (kb == p.num_k_blocks - 1) umma_commit_pair(&tmem_full[as]);
                    }
                    __syncwarp();
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else {
        // ===================== epilogue (both CTAs: own 128 rows x 256 columns) =====================
        const int quarter = warp & 3;
        const int half = (warp - 2) >> 2;
        const int et = threadIdx.x - 64;
        constexpr int CH = BN / 32;
        const int mode = p.epilogue, N = p.N;
        int it = 0;
        for (int tile = pair; tile < num_tiles; tile += num_pairs, ++it) {
            const int m0 = (tile % p.num_m_blocks) * 256 + (int)rank * 128;
            const int n0 = (tile / p.num_m_blocks) * BN;
            const uint32_t as = it & 1;
            float* sb = sbias + as * 256;
            if (p.bias) stage_bias<BN, 256>(p, sb, n0, et);
            const int row = m0 + quarter * 32 + lane;
            const bool row_ok = row < p.M;
            __nv_bfloat16* drow = row_ok ? reinterpret_cast<__nv_bfloat16*>(p.D) + (int64_t)row * p.ldd : nullptr;
            const __nv_bfloat16* arow = (row_ok && p.aux) ? p.aux + (int64_t)row * p.ldaux : nullptr;
            __nv_bfloat16* zrow = (row_ok && p.aux_out) ? p.aux_out + (int64_t)row * p.ldd : nullptr;
            mbar_wait(&tmem_full[as], (it >> 1) & 1);  // multicast commit from the leader
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + as * BN;
            fast_drain(taddr, n0, N, half * (CH / 2), (half + 1) * (CH / 2), mode, drow, arow, zrow,
                       p.bias ? sb : nullptr);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(mapa_u32(smem_u32(&tmem_empty[as]), 0));
        }
    }

    tc_fence_before();
    __syncthreads();
    cluster_sync_all();   // the leader's MMAs read the peer's shared memory until the last commit
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
    }
}

// ---------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(ptr);
    });
    return fn;
}

// 2-D bf16 tensor map: `inner` contiguous elements per row, `outer` rows, row pitch ld elements.
// Rank-4 map: [batch_outer][batch_inner][outer rows][inner contiguous]; plain GEMMs use 1x1 batches.
static int make_map(CUtensorMap* map, const void* base, int64_t inner, int64_t outer, int64_t ld,
                    int box_inner, int box_outer, int nbi = 1, int64_t sbi = 0, int nbo = 1,
                    int64_t sbo = 0) {
    EncodeTiledFn enc = get_encode_fn();
    B200_REQUIRE(enc != nullptr, "gemm: cuTensorMapEncodeTiled is unavailable from this driver");
    if (nbi <= 1) { nbi = 1; sbi = ld * outer; }
    if (nbo <= 1) { nbo = 1; sbo = sbi * nbi; }
    B200_REQUIRE((sbi % 8) == 0 && (sbo % 8) == 0, "gemm: batch strides must be multiples of 8 elements");
    cuuint64_t dims[4] = {(cuuint64_t)inner, (cuuint64_t)outer, (cuuint64_t)nbi, (cuuint64_t)nbo};
    cuuint64_t strides[3] = {(cuuint64_t)ld * 2, (cuuint64_t)sbi * 2, (cuuint64_t)sbo * 2};
    cuuint32_t box[4] = {(cuuint32_t)box_inner, (cuuint32_t)box_outer, 1, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides,
                     box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B200_REQUIRE(r == CUDA_SUCCESS, "gemm: cuTensorMapEncodeTiled failed (%d) inner=%lld outer=%lld ld=%lld box=%dx%d",
                 (int)r, (long long)inner, (long long)outer, (long long)ld, box_inner, box_outer);
    return 0;
}

typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeIm2colFn get_im2col_fn() {
    static EncodeIm2colFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeIm2colFn>(ptr);
    });
    return fn;
}

// im2col-mode tensor map over an NHWC bf16 activation: one load fetches `pixels` consecutive output
// pixels (walking q, then p, then the image) x 64 channels at filter-tap offset (s, r), zero-filling
// the padding halo.  Bounding box per the driver's convention: lower corner = -pad, upper corner =
// pad - (filter extent - 1); traversal stride = convolution stride.
static int make_im2col_map(CUtensorMap* map, const void* base, const b200_gemm_desc* d, int pixels) {
    EncodeIm2colFn enc = get_im2col_fn();
    B200_REQUIRE(enc != nullptr, "gemm: cuTensorMapEncodeIm2col is unavailable from this driver");
    cuuint64_t dims[4] = {(cuuint64_t)d->conv_c, (cuuint64_t)d->conv_w, (cuuint64_t)d->conv_h, (cuuint64_t)d->conv_n};
    cuuint64_t strides[3] = {(cuuint64_t)d->conv_c * 2, (cuuint64_t)d->conv_w * d->conv_c * 2,
                             (cuuint64_t)d->conv_h * d->conv_w * d->conv_c * 2};
    int lower[2] = {-d->conv_pad, -d->conv_pad};
    int upper[2] = {d->conv_pad - (d->conv_s - 1), d->conv_pad - (d->conv_r - 1)};
    cuuint32_t estr[4] = {1, (cuuint32_t)d->conv_stride, (cuuint32_t)d->conv_stride, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, lower, upper,
                     64, (cuuint32_t)pixels, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B200_REQUIRE(r == CUDA_SUCCESS, "gemm: cuTensorMapEncodeIm2col failed (%d) n=%d h=%d w=%d c=%d r=%d s=%d stride=%d pad=%d",
                 (int)r, d->conv_n, d->conv_h, d->conv_w, d->conv_c, d->conv_r, d->conv_s, d->conv_stride, d->conv_pad);
    return 0;
}

template <int BN, bool FAST>
static int launch(const CUtensorMap& ma, const CUtensorMap& mb, const GemmParams& p, int ctas, cudaStream_t s) {
    constexpr uint32_t smem = gemm_smem_bytes<BN>();
    // Programmatic dependent launch: the next GEMM's prologue (barrier init, TMEM
    // allocation, descriptor prefetch) overlaps this one's tail; every global access in
    // the kernel sits behind griddepcontrol.wait.  B200DDP_NO_PDL=1 disables it.
    static const bool pdl = getenv("B200DDP_NO_PDL") == nullptr;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)ctas);
    cfg.blockDim = dim3(gemm_threads<FAST>());
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    B200_CUDA_OK(cudaLaunchKernelEx(&cfg, gemm_kernel<BN, FAST>, ma, mb, p));
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}

static int launch_m256(const CUtensorMap& ma, const CUtensorMap& mb, const GemmParams& p, int ctas, cudaStream_t s) {
    static const bool pdl = getenv("B200DDP_NO_PDL") == nullptr;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)ctas);
    cfg.blockDim = dim3(gemm_threads<true>());
    cfg.dynamicSmemBytes = GEMM2_SMEM;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    B200_CUDA_OK(cudaLaunchKernelEx(&cfg, gemm_kernel_m256, ma, mb, p));
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}

template <int BN>
static int launch_2sm(const CUtensorMap& ma, const CUtensorMap& mb, const GemmParams& p, int ctas, cudaStream_t s) {
    static const bool pdl = getenv("B200DDP_NO_PDL") == nullptr;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)ctas);   // even; the kernel carries __cluster_dims__(2, 1, 1)
    cfg.blockDim = dim3(gemm_threads<true>());
    cfg.dynamicSmemBytes = gemm_2sm_smem<BN>();
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    B200_CUDA_OK(cudaLaunchKernelEx(&cfg, gemm_kernel_2sm<BN>, ma, mb, p));
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}

static int sm_count() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
    }
    return n;
}

static int pick_tile_n(const b200_gemm_desc* d, int sms) {
    if (d->epilogue == B200_EPI_SOFTMAX_CE) return d->N <= 16 ? 16 : 32;
    if (d->N <= 16 && !d->b_mn_major) return 16;
    if (d->N <= 32 && !d->b_mn_major) return 32;
    if (d->N <= 64) return 64;
    // split-K fills the SMs with K slices, so take the widest tile the output allows (an N = 64
    // MMA runs the tensor pipe at a quarter of its N = 256 rate)
    if (d->split_k) return d->N >= 192 ? 256 : (d->N >= 96 ? 128 : 64);
    const int64_t mblk = (d->M + BM - 1) / BM;
    int best = 64;
    int64_t best_cost = INT64_MAX;
    const int cands[3] = {256, 128, 64};
    for (int c : cands) {
        const int64_t tiles = mblk * ((d->N + c - 1) / c);
        const int64_t waves = (tiles + sms - 1) / sms;
        // measured: tiles are bound by operand bytes pulled from L2, (128 + BN) * 128 B per
        // k-block, so the cost of a wave scales with 128 + BN; ties go to the wider tile
        const int64_t cost = waves * (128 + c);
        if (cost < best_cost) { best_cost = cost; best = c; }
    }
    return best;
}

// Pair-tile width by a two-term cost: waves of pair tiles x (k-blocks x time per k-block + epilogue).
// Units of 10 ns, fitted to the GPT-2 shapes on B200 (tools/gpt2_gemm_shapes.py).
static int pick_pair_tile_n(const b200_gemm_desc* d, int sms) {
    if (d->N <= 128) return 128;
    const int64_t pairs = sms / 2 > 0 ? sms / 2 : 1;
    const int64_t mblk = (d->M + 255) / 256, kb = (d->K + BK - 1) / BK;
    const int64_t w256 = (mblk * ((d->N + 255) / 256) + pairs - 1) / pairs;
    const int64_t w128 = (mblk * ((d->N + 127) / 128) + pairs - 1) / pairs;
    // per k-block 0.34 us (256 wide) / 0.28 us (128 wide) per SM, plus the exposed epilogue
    const int64_t c256 = w256 * (kb * 34 + 150), c128 = w128 * (kb * 28 + 75);
    return c128 < c256 ? 128 : 256;
}

}  // namespace b200

int b200_gemm_init();
static int ensure_init() {
    static int rc = b200_gemm_init();
    return rc;
}

// Which kernel family and N tile a problem runs on.  Pure host logic (no CUDA calls), also exported
// as b200_gemm_plan so the policy is unit-tested on boxes without a GPU.
enum { GEMM_K_GENERAL = 0, GEMM_K_FAST = 1, GEMM_K_M256 = 2, GEMM_K_PAIR = 3 };
struct GemmPlan {
    int kernel, bn;
};
static int select_kernel(const b200_gemm_desc* d, int sms, GemmPlan* out) {
    using namespace b200;
    const int nbi = d->batch_inner > 1 ? d->batch_inner : 1, nbo = d->batch_outer > 1 ? d->batch_outer : 1;
    if (nbi * nbo > 1)
        B200_REQUIRE(d->epilogue == B200_EPI_NONE && !d->aux && !d->bias && !d->colsum && !d->split_k,
                     "gemm: batched mode supports the plain epilogue only");
    B200_REQUIRE(d->causal >= 0 && d->causal <= 3, "gemm: causal mode %d unknown", d->causal);
    B200_REQUIRE(d->tile_m == 0 || d->tile_m == 128 || d->tile_m == 256, "gemm: tile_m must be 0, 128 or 256");
    if (d->epilogue == B200_EPI_GATE_POSITIVE || d->epilogue == B200_EPI_GELU_BWD)
        B200_REQUIRE(d->aux != nullptr, "gemm: epilogue %d needs aux", d->epilogue);

    // ---- kernel selection
    // FAST epilogue kernels: bf16 output with 16-byte-aligned rows and one of the light epilogues
    const bool light = d->epilogue == B200_EPI_NONE || d->epilogue == B200_EPI_GELU || d->epilogue == B200_EPI_GELU_BWD;
    const bool vec_d = !d->d_fp32 && (d->ldd % 8) == 0 && ((uintptr_t)d->D % 16) == 0 &&
                       (d->d_stride_inner % 8) == 0 && (d->d_stride_outer % 8) == 0;
    const bool vec_aux = !d->aux || ((d->ldaux % 8) == 0 && ((uintptr_t)d->aux % 16) == 0 && (d->N % 8) == 0);
    const bool vec_z = !d->aux_out || (((uintptr_t)d->aux_out % 16) == 0 && (d->N % 8) == 0);
    static const bool fast_off = getenv("B200DDP_NO_FAST_EPILOGUE") != nullptr;
    static const bool m256_off = getenv("B200DDP_NO_M256") != nullptr;
    static const bool pair_off = getenv("B200DDP_NO_PAIR") != nullptr;
    const bool fast_cfg = !fast_off && light && vec_d && vec_aux && vec_z && !d->colsum && !d->split_k;
    // Large-tile kernels for plain non-batched problems with a light epilogue:
    //   256x256 dual accumulator, one SM   long reductions (K >= 4096) with at least a wave of tiles:
    //                                      1481 vs 1288 TFLOP/s at 8192^3; with short K its
    //                                      single-buffered accumulators expose the epilogue
    //   256xBN per SM pair (cta_group::2)  everything else: double-buffered accumulators, and each SM
    //                                      stages only half of B
    const bool conv = d->a_conv || d->b_conv;
    if (conv) {
        B200_REQUIRE(nbi * nbo == 1 && !d->causal, "gemm: implicit convolution operands are not batched");
        B200_REQUIRE(!(d->a_conv && d->b_conv), "gemm: only one operand can be the im2col view");
        B200_REQUIRE(d->conv_c > 0 && d->conv_c % 64 == 0, "gemm: implicit convolution needs channels %% 64 == 0 (got %d)", d->conv_c);
        B200_REQUIRE(d->conv_r >= 1 && d->conv_s >= 1 && d->conv_stride >= 1 && d->conv_pad >= 0 && d->conv_n > 0 &&
                     d->conv_h > 0 && d->conv_w > 0, "gemm: bad convolution geometry");
        const int P = (d->conv_h + 2 * d->conv_pad - d->conv_r) / d->conv_stride + 1;
        const int Q = (d->conv_w + 2 * d->conv_pad - d->conv_s) / d->conv_stride + 1;
        const int64_t pix = (int64_t)d->conv_n * P * Q, kk = (int64_t)d->conv_r * d->conv_s * d->conv_c;
        if (d->a_conv) B200_REQUIRE(d->M == pix && d->K == kk && !d->a_mn_major, "gemm: a_conv wants M = N*P*Q (%lld) and K = R*S*C (%lld)", (long long)pix, (long long)kk);
        if (d->b_conv) B200_REQUIRE(d->N == kk && d->K == pix, "gemm: b_conv wants N = R*S*C (%lld) and K = N*P*Q (%lld)", (long long)kk, (long long)pix);
    }
    // (the im2col view as B operand is read by the one-SM kernels only)
    const bool big_cfg = fast_cfg && nbi * nbo == 1 && !d->causal && !d->b_conv;
    const int64_t t2 = (int64_t)((d->M + BM2 - 1) / BM2) * ((d->N + 255) / 256);
    bool m256 = big_cfg && (d->tile_n == 0 || d->tile_n == 256) &&
                !conv && (d->tile_m == 256 || (d->tile_m == 0 && d->tile_n == 0 && d->cta_pair != 1 && !m256_off &&
                                                d->M >= 512 && d->K >= 4096 && t2 >= sms));
    int pair_bn = 0;   // 0: not the pair kernel
    if (big_cfg && !m256 && d->tile_m == 0 && d->M > 128 && d->N > 64 &&
        (d->cta_pair == 1 || (d->cta_pair == 0 && d->tile_n == 0 && !pair_off))) {
        if (d->tile_n == 128 || d->tile_n == 256) pair_bn = d->tile_n;
        else if (d->tile_n == 0) pair_bn = pick_pair_tile_n(d, sms);
    }
    const bool pair = pair_bn != 0;
    B200_REQUIRE(d->cta_pair != 1 || pair,
                 "gemm: cta_pair needs tile_n 0/128/256, tile_m 0, M > 128, N > 64, a non-batched problem and a light "
                 "epilogue with bf16 D");
    B200_REQUIRE(d->tile_m != 256 || m256,
                 "gemm: tile_m 256 needs tile_n 256, a non-batched problem and a light epilogue with bf16 output");
    int bn = pair ? pair_bn : (m256 ? 256 : (d->tile_n ? d->tile_n : pick_tile_n(d, sms)));
    B200_REQUIRE(bn == 16 || bn == 32 || bn == 64 || bn == 128 || bn == 256, "gemm: tile_n %d unsupported", bn);
    B200_REQUIRE(!((d->b_mn_major || d->b_conv) && bn < 64), "gemm: MN-major B needs tile_n >= 64");
    if (d->epilogue == B200_EPI_SOFTMAX_CE) {
        B200_REQUIRE(d->N <= bn && bn <= 32, "gemm: fused cross-entropy needs N <= 32 (got %d)", d->N);
        B200_REQUIRE(d->targets != nullptr, "gemm: fused cross-entropy needs targets");
    }
    out->bn = bn;
    out->kernel = pair ? GEMM_K_PAIR : (m256 ? GEMM_K_M256 : ((fast_cfg && bn >= 64) ? GEMM_K_FAST : GEMM_K_GENERAL));
    return 0;

}

extern "C" int b200_gemm(const b200_gemm_desc* d, void* stream) {
    using namespace b200;
    if (int rc = ensure_init()) return rc;
    // cuTensorMapEncodeTiled is a driver call and needs the primary context current on THIS
    // thread (CUDA_ERROR_INVALID_CONTEXT otherwise).  Autograd runs backward on worker threads
    // whose first CUDA work may be this GEMM, so bind the context once per thread.
    static thread_local bool ctx_bound = false;
    if (!ctx_bound) {
        int dev = 0;
        B200_CUDA_OK(cudaGetDevice(&dev));
        B200_CUDA_OK(cudaSetDevice(dev));  // binds the primary context; legal during stream capture
        ctx_bound = true;
    }
    B200_REQUIRE(d && d->A && d->B && d->D, "gemm: null pointer");
    B200_REQUIRE(d->M > 0 && d->N > 0 && d->K > 0, "gemm: bad shape %dx%dx%d", d->M, d->N, d->K);
    B200_REQUIRE((d->a_conv || (d->lda % 8) == 0) && (d->b_conv || (d->ldb % 8) == 0),
                 "gemm: lda/ldb must be multiples of 8 elements (got %lld, %lld)", (long long)d->lda, (long long)d->ldb);
    B200_REQUIRE(((uintptr_t)d->A % 16) == 0 && ((uintptr_t)d->B % 16) == 0, "gemm: A/B must be 16-byte aligned");
    const int sms = sm_count();
    const int nbi = d->batch_inner > 1 ? d->batch_inner : 1, nbo = d->batch_outer > 1 ? d->batch_outer : 1;
    GemmPlan plan;
    if (int rc = select_kernel(d, sms, &plan)) return rc;
    const int bn = plan.bn;
    const bool pair = plan.kernel == GEMM_K_PAIR, m256 = plan.kernel == GEMM_K_M256;
    const bool fast = plan.kernel == GEMM_K_FAST;

    CUtensorMap ma, mb;
    int rc;
    if (d->a_conv) rc = make_im2col_map(&ma, d->A, d, BM);   // 128 output pixels x 64 channels per load
    else if (!d->a_mn_major) rc = make_map(&ma, d->A, d->K, d->M, d->lda, BK, m256 ? BM2 : BM, nbi, d->a_stride_inner, nbo, d->a_stride_outer);
    else rc = make_map(&ma, d->A, d->M, d->K, d->lda, 64, BK, nbi, d->a_stride_inner, nbo, d->a_stride_outer);
    if (rc) return rc;
    if (d->b_conv) rc = make_im2col_map(&mb, d->B, d, BK);   // 64 pixels (one k-block) x 64 channels per load
    else if (!d->b_mn_major) rc = make_map(&mb, d->B, d->K, d->N, d->ldb, BK, pair ? bn / 2 : bn, nbi, d->b_stride_inner, nbo, d->b_stride_outer);
    else rc = make_map(&mb, d->B, d->N, d->K, d->ldb, 64, BK, nbi, d->b_stride_inner, nbo, d->b_stride_outer);
    if (rc) return rc;

    GemmParams p{};
    p.M = d->M; p.N = d->N; p.K = d->K;
    p.a_mn = d->a_mn_major; p.b_mn = d->b_mn_major || d->b_conv;
    p.a_conv = d->a_conv; p.b_conv = d->b_conv;
    if (d->a_conv || d->b_conv) {
        const int P = (d->conv_h + 2 * d->conv_pad - d->conv_r) / d->conv_stride + 1;
        const int Q = (d->conv_w + 2 * d->conv_pad - d->conv_s) / d->conv_stride + 1;
        p.conv_cblocks = d->conv_c / 64; p.conv_S = d->conv_s; p.conv_taps = d->conv_r * d->conv_s;
        p.conv_Q = Q; p.conv_PQ = P * Q; p.conv_stride = d->conv_stride; p.conv_pad = d->conv_pad; p.conv_C = d->conv_c;
    }
    p.D = d->D; p.ldd = d->ldd; p.d_fp32 = d->d_fp32;
    p.bias = d->bias; p.epilogue = d->epilogue;
    p.aux = (const __nv_bfloat16*)d->aux; p.ldaux = d->ldaux;
    p.aux_out = (__nv_bfloat16*)d->aux_out; p.colsum = d->colsum;
    p.dropout_p = d->dropout_p; p.gate_scale = d->gate_scale;
    p.seed = d->seed; p.step_dev = d->step_dev; p.layer_id = d->layer_id;
    p.targets = d->targets; p.ce_grad_scale = d->ce_grad_scale;
    p.loss_out = d->loss_out; p.loss_scale = d->loss_scale; p.ce_write_grad = d->ce_write_grad;
    p.num_m_blocks = (m256 || pair) ? (d->M + BM2 - 1) / BM2 : (d->M + BM - 1) / BM;
    p.num_n_blocks = (d->N + bn - 1) / bn;
    p.num_k_blocks = (d->K + BK - 1) / BK;
    // instruction descriptor: fp32 accumulate, bf16 A/B, majors, N>>3, M>>4
    p.idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(d->a_mn_major ? 1 : 0) << 15) |
              ((uint32_t)((d->b_mn_major || d->b_conv) ? 1 : 0) << 16) | ((uint32_t)(bn >> 3) << 17) |
              ((uint32_t)((pair ? 256 : BM) >> 4) << 24);
    p.batch_inner = nbi; p.batch_outer = nbo;
    p.d_stride_inner = d->d_stride_inner; p.d_stride_outer = d->d_stride_outer;
    p.causal = d->causal;
    // split-K (wgrad with a long reduction and few output tiles): enough K slices to give
    // every SM about two tiles, at least 4 k-blocks each
    p.k_splits = 1;
    p.kb_per_split = p.num_k_blocks;
    if (d->split_k) {
        B200_REQUIRE(d->d_fp32 && d->epilogue == B200_EPI_NONE && !d->bias && !d->aux && !d->colsum,
                     "gemm: split_k needs a plain fp32 output (zeroed by the caller)");
        const int mn = p.num_m_blocks * p.num_n_blocks;
        int want = d->split_k > 1 ? d->split_k : (2 * sms + mn - 1) / mn;
        const int max_by_k = p.num_k_blocks / 4 > 0 ? p.num_k_blocks / 4 : 1;
        if (want > max_by_k) want = max_by_k;
        if (want > 1) {
            p.kb_per_split = (p.num_k_blocks + want - 1) / want;
            p.k_splits = (p.num_k_blocks + p.kb_per_split - 1) / p.kb_per_split;
        }
    }
    const int tiles = p.num_m_blocks * p.num_n_blocks * p.k_splits * nbi * nbo;
    int ctas = d->max_ctas > 0 ? d->max_ctas : sms;
    if (ctas > tiles) ctas = tiles;
    cudaStream_t s = (cudaStream_t)stream;
    if (pair) {
        int c2 = 2 * tiles < sms ? 2 * tiles : sms;
        if (d->max_ctas > 0 && d->max_ctas < c2) c2 = d->max_ctas;
        c2 &= ~1;
        B200_REQUIRE(c2 >= 2, "gemm: the two-SM kernel needs at least 2 CTAs");
        return bn == 256 ? launch_2sm<256>(ma, mb, p, c2, s) : launch_2sm<128>(ma, mb, p, c2, s);
    }
    if (m256) return launch_m256(ma, mb, p, ctas, s);
    if (fast) {
        switch (bn) {
            case 64: return launch<64, true>(ma, mb, p, ctas, s);
            case 128: return launch<128, true>(ma, mb, p, ctas, s);
            default: return launch<256, true>(ma, mb, p, ctas, s);
        }
    }
    switch (bn) {
        case 16: return launch<16, false>(ma, mb, p, ctas, s);
        case 32: return launch<32, false>(ma, mb, p, ctas, s);
        case 64: return launch<64, false>(ma, mb, p, ctas, s);
        case 128: return launch<128, false>(ma, mb, p, ctas, s);
        default: return launch<256, false>(ma, mb, p, ctas, s);
    }
}

// out[0] = kernel family (0 general, 1 lean epilogue, 2 256x256 dual accumulator, 3 SM pair),
// out[1] = N tile.  `sms` <= 0 means the current device (148 without one).
extern "C" int b200_gemm_plan(const b200_gemm_desc* d, int sms, int* out) {
    B200_REQUIRE(d && out, "gemm_plan: null pointer");
    B200_REQUIRE(d->M > 0 && d->N > 0 && d->K > 0, "gemm: bad shape %dx%dx%d", d->M, d->N, d->K);
    GemmPlan plan;
    if (int rc = select_kernel(d, sms > 0 ? sms : b200::sm_count(), &plan)) return rc;
    out[0] = plan.kernel;
    out[1] = plan.bn;
    return 0;
}

// Load every kernel variant and set its shared-memory opt-in up front, so the first
// real launch can happen inside a CUDA-graph capture.
int b200_gemm_init() {
    using namespace b200;
    B200_REQUIRE(get_encode_fn() != nullptr, "init: cuTensorMapEncodeTiled is unavailable");
#define PREP(BN, FAST)                                                                                  \
    B200_CUDA_OK(cudaFuncSetAttribute(gemm_kernel<BN, FAST>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                      (int)gemm_smem_bytes<BN>()))
    PREP(16, false); PREP(32, false); PREP(64, false); PREP(128, false); PREP(256, false);
    PREP(64, true); PREP(128, true); PREP(256, true);
#undef PREP
    B200_CUDA_OK(cudaFuncSetAttribute(gemm_kernel_m256, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GEMM2_SMEM));
    B200_CUDA_OK(cudaFuncSetAttribute(gemm_kernel_2sm<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gemm_2sm_smem<256>()));
    B200_CUDA_OK(cudaFuncSetAttribute(gemm_kernel_2sm<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gemm_2sm_smem<128>()));
    return 0;
}
