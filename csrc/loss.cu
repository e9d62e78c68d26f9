// Fused softmax cross-entropy: forward value, argmax-correct count and the
// backward (softmax - onehot) * scale in one kernel.
//
// ref: train.py:137 criterion(output, target), train.py:138 its backward, and the
// argmax/eq/sum of validate() at train.py:168-169.  torch runs log_softmax,
// nll_loss, nll_loss_backward, _log_softmax_backward_data (+ argmax, eq, sum) as
// separate kernels and syncs the host for each scalar.
//
// Small class counts: one warp per row.  Large (LM head, 50257): one 256-thread
// block per row.  Row statistics are reduced without atomics on floats: every
// block writes its partial to scratch, the last block to finish (ticket) sums
// them in block order, so the mean loss is deterministic.
#include "common.cuh"

#include <cstdlib>
#include <mutex>

void b200_count_launch(int n);

namespace b200 {

constexpr int CE_MAX_BLOCKS = 4096;

template <bool BF16>
__device__ __forceinline__ float ld_logit(const void* base, int64_t i) {
    if (BF16) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(base)[i]);
    return reinterpret_cast<const float*>(base)[i];
}
template <bool BF16>
__device__ __forceinline__ void st_logit(void* base, int64_t i, float v) {
    if (BF16) reinterpret_cast<__nv_bfloat16*>(base)[i] = __float2bfloat16_rn(v);
    else reinterpret_cast<float*>(base)[i] = v;
}

struct RowStat {
    float mx, sum;
    int arg;
};
__device__ __forceinline__ RowStat merge(RowStat a, RowStat b) {
    RowStat r;
    if (b.mx > a.mx || (b.mx == a.mx && b.arg < a.arg)) { RowStat t = a; a = b; b = t; }
    r.mx = a.mx;
    r.arg = a.arg;
    r.sum = a.sum + (b.sum > 0.f ? b.sum * __expf(b.mx - a.mx) : 0.f);
    return r;
}

// TEAM = 32: warp per row, 8 rows per block.  TEAM = 256: block per row.
template <bool BF16, int TEAM>
__global__ void __launch_bounds__(256) ce_kernel(const void* __restrict__ logits, int64_t ld,
                                                 const int64_t* __restrict__ targets, int64_t rows,
                                                 int classes, void* dlogits, float grad_scale,
                                                 float* __restrict__ loss_out,
                                                 unsigned int* __restrict__ ticket,
                                                 float2* __restrict__ partials) {
    constexpr int TEAMS = 256 / TEAM;
    __shared__ RowStat s_stat[8];
    __shared__ float s_loss[8], s_ok[8];
    const int team = threadIdx.x / TEAM, lane = threadIdx.x % TEAM;
    const int warp = threadIdx.x >> 5;
    float acc_loss = 0.f, acc_ok = 0.f;

    for (int64_t row = (int64_t)blockIdx.x * TEAMS + team; row < rows;
         row += (int64_t)gridDim.x * TEAMS) {
        const int64_t base = row * ld;
        // Target logit first: dlogits may alias logits, and another warp of the team may already be
        // writing this row's gradient when lane 0 gets to the loss term (every thread of the team
        // passes the merge barrier below before the first write).  A target outside [0, classes)
        // (e.g. an ignore index) reads nothing, adds no loss and gets no -1.
        const int tgt_raw = (int)targets[row];
        const bool tgt_ok = tgt_raw >= 0 && tgt_raw < classes;
        const float x_tgt = tgt_ok ? ld_logit<BF16>(logits, base + tgt_raw) : 0.f;
        RowStat st{-INFINITY, 0.f, 0x7fffffff};
        for (int c = lane; c < classes; c += TEAM) {
            const float x = ld_logit<BF16>(logits, base + c);
            if (x > st.mx) {
                st.sum = st.sum * __expf(st.mx - x) + 1.f;
                st.mx = x;
                st.arg = c;
            } else {
                st.sum += __expf(x - st.mx);
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            RowStat other;
            other.mx = __shfl_xor_sync(0xffffffffu, st.mx, o);
            other.sum = __shfl_xor_sync(0xffffffffu, st.sum, o);
            other.arg = __shfl_xor_sync(0xffffffffu, st.arg, o);
            st = merge(st, other);
        }
        if (TEAM > 32) {
            if ((threadIdx.x & 31) == 0) s_stat[warp] = st;
            __syncthreads();
            st = s_stat[0];
#pragma unroll
            for (int w = 1; w < 8; ++w) st = merge(st, s_stat[w]);
            __syncthreads();
        }
        const int tgt = tgt_ok ? tgt_raw : -1;
        const float lse = st.mx + __logf(st.sum);
        if (lane == 0 && tgt_ok) {
            acc_loss += lse - x_tgt;
            acc_ok += (st.arg == tgt) ? 1.f : 0.f;
        }
        if (TEAM <= 32) __syncwarp();
        if (dlogits) {
            const float inv = 1.f / st.sum;
            for (int c = lane; c < classes; c += TEAM) {
                const float x = ld_logit<BF16>(logits, base + c);
                float g = __expf(x - st.mx) * inv;
                if (c == tgt) g -= 1.f;
                st_logit<BF16>(dlogits, base + c, g * grad_scale);
            }
        }
    }
    // block partial -> scratch; last block reduces in block order
    if ((threadIdx.x % TEAM) == 0) { s_loss[team] = acc_loss; s_ok[team] = acc_ok; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float l = 0.f, k = 0.f;
        for (int i = 0; i < TEAMS; ++i) { l += s_loss[i]; k += s_ok[i]; }
        partials[blockIdx.x] = make_float2(l, k);
        __threadfence();
        const unsigned int t = atomicAdd(ticket, 1u);
        if (t == gridDim.x - 1) {
            __threadfence();
            float L = 0.f, K = 0.f;
            for (unsigned int b = 0; b < gridDim.x; ++b) {
                L += __ldcg(&partials[b].x);
                K += __ldcg(&partials[b].y);
            }
            loss_out[0] = L / (float)rows;
            loss_out[1] = K;
            *ticket = 0u;
        }
    }
}

// Large vocabularies in bf16 (the LM head: 50257 classes, rows 16-byte aligned): one block per
// row with 16-byte loads and stores.  The scalar kernel above moves 2 bytes per thread per access
// (64 contiguous bytes per warp instruction) and reached 0.45 of the HBM roofline on the 823 MB of
// GPT-2 logits; here a warp instruction covers 512 contiguous bytes.  Two passes over the row (row
// statistics, then the gradient written in place); the grid is capped so that the rows in flight
// (100 KB each) stay in L2 between the passes.
__device__ __forceinline__ void ce_unpack8(const uint4& raw, float (&f)[8]) {
    const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const float2 t = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[q]));
        f[2 * q] = t.x;
        f[2 * q + 1] = t.y;
    }
}
__device__ __forceinline__ void ce_online(RowStat& st, float x, int c) {
    if (x > st.mx) {
        st.sum = st.sum * __expf(st.mx - x) + 1.f;
        st.mx = x;
        st.arg = c;
    } else {
        st.sum += __expf(x - st.mx);
    }
}

__global__ void __launch_bounds__(256) ce_row_vec_kernel(const __nv_bfloat16* logits, int64_t ld,
                                                         const int64_t* __restrict__ targets, int64_t rows, int classes,
                                                         __nv_bfloat16* dlogits, float grad_scale,
                                                         float* __restrict__ loss_out, unsigned int* __restrict__ ticket,
                                                         float2* __restrict__ partials) {
    __shared__ RowStat s_stat[8];
    const int lane = threadIdx.x, warp = threadIdx.x >> 5;
    const int nvec = classes >> 3, tail0 = nvec << 3;
    float acc_loss = 0.f, acc_ok = 0.f;
    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
        const __nv_bfloat16* src = logits + row * ld;
        const uint4* src4 = reinterpret_cast<const uint4*>(src);
        const int tgt_raw = (int)targets[row];
        const bool tgt_ok = tgt_raw >= 0 && tgt_raw < classes;   // out of range (ignore index): no loss, no -1
        const int tgt = tgt_ok ? tgt_raw : -1;
        // read before anything of this row is overwritten (dlogits may alias logits)
        const float x_tgt = (lane == 0 && tgt_ok) ? __bfloat162float(src[tgt]) : 0.f;
        RowStat st{-INFINITY, 0.f, 0x7fffffff};
        for (int v = lane; v < nvec; v += 256) {
            float x[8];
            ce_unpack8(src4[v], x);
            float m8 = x[0];
            int a8 = 0;
#pragma unroll
            for (int k = 1; k < 8; ++k)
                if (x[k] > m8) { m8 = x[k]; a8 = k; }
            if (m8 > st.mx) {
                st.sum *= __expf(st.mx - m8);
                st.mx = m8;
                st.arg = v * 8 + a8;
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) st.sum += __expf(x[k] - st.mx);
        }
        if (tail0 + lane < classes) ce_online(st, __bfloat162float(src[tail0 + lane]), tail0 + lane);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            RowStat other;
            other.mx = __shfl_xor_sync(0xffffffffu, st.mx, o);
            other.sum = __shfl_xor_sync(0xffffffffu, st.sum, o);
            other.arg = __shfl_xor_sync(0xffffffffu, st.arg, o);
            st = merge(st, other);
        }
        if ((threadIdx.x & 31) == 0) s_stat[warp] = st;
        __syncthreads();
        st = s_stat[0];
#pragma unroll
        for (int w = 1; w < 8; ++w) st = merge(st, s_stat[w]);
        __syncthreads();
        if (lane == 0 && tgt_ok) {
            acc_loss += st.mx + __logf(st.sum) - x_tgt;
            acc_ok += (st.arg == tgt) ? 1.f : 0.f;
        }
        if (dlogits) {
            const float inv = 1.f / st.sum;
            uint4* dst4 = reinterpret_cast<uint4*>(dlogits + row * ld);
            for (int v = lane; v < nvec; v += 256) {
                float x[8];
                ce_unpack8(src4[v], x);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    float g = __expf(x[k] - st.mx) * inv;
                    if (v * 8 + k == tgt) g -= 1.f;
                    x[k] = g * grad_scale;
                }
                dst4[v] = make_uint4(pack_bf16(x[0], x[1]), pack_bf16(x[2], x[3]), pack_bf16(x[4], x[5]),
                                     pack_bf16(x[6], x[7]));
            }
            if (tail0 + lane < classes) {
                const int c = tail0 + lane;
                float g = __expf(__bfloat162float(src[c]) - st.mx) * inv;
                if (c == tgt) g -= 1.f;
                dlogits[row * ld + c] = __float2bfloat16_rn(g * grad_scale);
            }
        }
    }
    if (threadIdx.x == 0) {   // block partial -> scratch; the last block reduces in block order
        partials[blockIdx.x] = make_float2(acc_loss, acc_ok);
        __threadfence();
        const unsigned int t = atomicAdd(ticket, 1u);
        if (t == gridDim.x - 1) {
            __threadfence();
            float L = 0.f, K = 0.f;
            for (unsigned int b = 0; b < gridDim.x; ++b) {
                L += __ldcg(&partials[b].x);
                K += __ldcg(&partials[b].y);
            }
            loss_out[0] = L / (float)rows;
            loss_out[1] = K;
            *ticket = 0u;
        }
    }
}

}  // namespace b200

extern "C" int64_t b200_ce_scratch_bytes(int64_t rows) {
    (void)rows;
    return 16 + (int64_t)sizeof(float2) * b200::CE_MAX_BLOCKS;
}

extern "C" int b200_cross_entropy(const void* logits, int logits_bf16, int64_t ld,
                                  const int64_t* targets, int64_t rows, int classes, void* dlogits,
                                  float grad_scale, float* loss_out, void* scratch, void* stream) {
    B200_REQUIRE(logits && targets && loss_out && scratch, "cross_entropy: null pointer");
    B200_REQUIRE(rows > 0 && classes > 0 && ld >= classes, "cross_entropy: bad shape rows=%lld classes=%d ld=%lld",
                 (long long)rows, classes, (long long)ld);
    B200_REQUIRE((uintptr_t)scratch % 16 == 0, "cross_entropy: scratch must be 16-byte aligned");
    unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);
    float2* partials = reinterpret_cast<float2*>(reinterpret_cast<char*>(scratch) + 16);
    cudaStream_t s = (cudaStream_t)stream;
    const bool big = classes > 2048;
    static const bool vec_off = getenv("B200DDP_NO_VEC_CE") != nullptr;
    if (big && logits_bf16 && !vec_off && (ld % 8) == 0 && ((uintptr_t)logits % 16) == 0 &&
        (!dlogits || ((uintptr_t)dlogits % 16) == 0)) {
        int64_t vb = rows < 148 * 4 ? rows : 148 * 4;   // rows in flight stay L2-resident between the passes
        b200::ce_row_vec_kernel<<<(int)vb, 256, 0, s>>>((const __nv_bfloat16*)logits, ld, targets, rows, classes,
                                                        (__nv_bfloat16*)dlogits, grad_scale, loss_out, ticket, partials);
        b200_count_launch(1);
        B200_CUDA_OK(cudaGetLastError());
        return 0;
    }
    const int teams = big ? 1 : 8;
    int64_t blocks = (rows + teams - 1) / teams;
    if (blocks > b200::CE_MAX_BLOCKS) blocks = b200::CE_MAX_BLOCKS;
#define LAUNCH(BF, TEAM)                                                                      \
    b200::ce_kernel<BF, TEAM><<<(int)blocks, 256, 0, s>>>(logits, ld, targets, rows, classes, \
                                                          dlogits, grad_scale, loss_out, ticket, partials)
    if (logits_bf16) { if (big) LAUNCH(true, 256); else LAUNCH(true, 32); }
    else             { if (big) LAUNCH(false, 256); else LAUNCH(false, 32); }
#undef LAUNCH
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------
// Allocation-free variant.
//
// Who needs it: code that sits inside somebody else's memory regime and is not
// allowed to grow it.  Typical cases are a plugin running in a serving process
// whose device memory was carved up front (a cudaMalloc there can fail, or stalls
// every stream while the driver maps memory), a latency-critical loop where the
// implicit device synchronisation of cudaMalloc/cudaFree is unacceptable, and
// host code that has no allocator handle at all (a C caller that only holds the
// tensors it was given).  Such callers cannot provide the partial-sum scratch
// that b200_cross_entropy asks for, so this entry point brings its own: a fixed
// array in the library's device image.
//
// Cost: every GPU that loads the library carries these B200_CE_STATIC_BYTES
// until the process exits, whether or not the entry point is ever called.
namespace b200 {
constexpr int CE_STATIC_MAX_BLOCKS = B200_CE_STATIC_MAX_BLOCKS;
constexpr size_t CE_STATIC_BYTES = 16 + sizeof(float2) * CE_STATIC_MAX_BLOCKS;
}  // namespace b200

// In host code this name is only a handle for the runtime, NOT a device address;
// the address a kernel may use comes from cudaGetSymbolAddress (below).
static __device__ __align__(256) unsigned char g_ce_static_scratch[b200::CE_STATIC_BYTES];

namespace b200 {
// Device address of the array on the current device, resolved once per device
// with that device current.
static int ce_static_scratch(void** out) {
    constexpr int MAX_DEVICES = 64;
    static std::mutex mu;
    static void* addr[MAX_DEVICES] = {};
    int dev = -1;
    B200_CUDA_OK(cudaGetDevice(&dev));
    B200_REQUIRE(dev >= 0 && dev < MAX_DEVICES, "cross_entropy_static: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (addr[dev] == nullptr) {
        void* p = nullptr;
        B200_CUDA_OK(cudaGetSymbolAddress(&p, g_ce_static_scratch));
        B200_REQUIRE(p != nullptr, "cross_entropy_static: cudaGetSymbolAddress returned null");
        addr[dev] = p;
    }
    *out = addr[dev];
    return 0;
}
}  // namespace b200

extern "C" int b200_cross_entropy_static(const void* logits, int logits_bf16, int64_t ld,
                                         const int64_t* targets, int64_t rows, int classes,
                                         void* dlogits, float grad_scale, float* loss_out,
                                         void* stream) {
    using namespace b200;
    B200_REQUIRE(logits && targets && loss_out, "cross_entropy_static: null pointer");
    B200_REQUIRE(rows > 0 && classes > 0 && ld >= classes,
                 "cross_entropy_static: bad shape rows=%lld classes=%d ld=%lld", (long long)rows, classes,
                 (long long)ld);
    const bool big = classes > 2048;
    const int teams = big ? 1 : 8;  // rows handled per block
    const int64_t blocks = (rows + teams - 1) / teams;
    B200_REQUIRE(blocks <= CE_STATIC_MAX_BLOCKS,
                 "cross_entropy_static: %lld rows of %d classes need %lld partial slots but the built-in "
                 "scratch holds %d (limit: %d rows up to 2048 classes, %d rows above); use "
                 "b200_cross_entropy with caller-provided scratch",
                 (long long)rows, classes, (long long)blocks, CE_STATIC_MAX_BLOCKS,
                 CE_STATIC_MAX_BLOCKS * 8, CE_STATIC_MAX_BLOCKS);
    void* scratch = nullptr;
    if (int rc = ce_static_scratch(&scratch)) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    // The array still holds whatever the previous call left.  Partials are fully
    // rewritten by this launch; the ticket counter is the one word that must be zero.
    B200_CUDA_OK(cudaMemsetAsync(scratch, 0, 16, s));
    unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);
    float2* partials = reinterpret_cast<float2*>(reinterpret_cast<char*>(scratch) + 16);
#define LAUNCH(BF, TEAM)                                                                      \
    b200::ce_kernel<BF, TEAM><<<(int)blocks, 256, 0, s>>>(logits, ld, targets, rows, classes, \
                                                          dlogits, grad_scale, loss_out, ticket, partials)
    if (logits_bf16) { if (big) LAUNCH(true, 256); else LAUNCH(true, 32); }
    else             { if (big) LAUNCH(false, 256); else LAUNCH(false, 32); }
#undef LAUNCH
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}
