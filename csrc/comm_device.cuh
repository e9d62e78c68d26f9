// Device side of the peer-memory collectives (flag barriers, in-kernel reduction, parameter push),
// shared by comm.cu (one kernel per bucket) and mlp_fused.cu (update CTAs inside the step kernel).
// The protocol is described at the top of comm.cu.
#pragma once
#include "common.cuh"
#include "optim_update.cuh"

namespace b200 {

constexpr int MAX_RANKS = B200_MAX_RANKS;
// flag block (uint32 words) in every rank's symmetric memory:
//   [0, R)      ready[j]  written by rank j
//   [R, 2R)     done[j]   written by rank j
//   [2R]        epoch     local, advanced by the last block of each kernel
//   [2R+1]      ticket    local block counter
constexpr int F_READY = 0, F_DONE = MAX_RANKS, F_EPOCH = 2 * MAX_RANKS, F_TICKET = 2 * MAX_RANKS + 1;

struct CommArgs {
    int rank, world;
    const void* grad[MAX_RANKS];  // peer-mapped gradient arenas (same layout on every rank)
    void* push[MAX_RANKS];        // peer-mapped parameter arenas written by two-shot (or null)
    uint32_t* flags[MAX_RANKS];   // peer-mapped flag blocks
    const void* grad_mc;          // NVLS: multicast address of the gradient arenas (or null)
    void* push_mc;                // NVLS: multicast address of the parameter arenas (or null)
    float* master_peer[MAX_RANKS];  // peer-mapped fp32 master arenas (masked fp32 push) or null
    const uint8_t* f32_mask;      // one byte per 8 elements: non-zero = also push the fp32 master
    int64_t parity_stride;        // bytes between the two gradient buffers (0 = single-buffered)
    int no_exit_wait;             // one-shot on double-buffered gradients: skip the exit barrier
    uint64_t timeout_ns;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void fence_acq_rel_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
// volatile + "memory": the loads must stay behind the entry barrier that makes the peers'
// gradients valid (a pure asm could legally be hoisted above it).
__device__ __forceinline__ uint4 ld_peer_16(const void* p) {
    uint4 v;
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p)
                 : "memory");
    return v;
}

// Polls with relaxed loads (cheap: the flag lives in local memory), then takes the value once
// more with an acquire load — epochs only grow, so it passes — which orders everything after it
// behind the peer's release store.  Bounded so a dead peer traps instead of hanging the GPU.
__device__ __forceinline__ void spin_until(const uint32_t* flag, uint32_t epoch, uint64_t timeout_ns,
                                           int who) {
    if ((int32_t)(ld_relaxed_sys(flag) - epoch) < 0) {
        const uint64_t t0 = globaltimer_ns();
        while ((int32_t)(ld_relaxed_sys(flag) - epoch) < 0) {
            if (globaltimer_ns() - t0 > timeout_ns) {
                printf("b200ddp: peer flag wait timed out (waiting on rank %d, epoch %u, block %d)\n",
                       who, epoch, (int)blockIdx.x);
                __trap();
            }
        }
    }
    (void)ld_acquire_sys(flag);
}

// All blocks: wait until every rank has published `epoch` in our ready flags.
// Block 0 first publishes ours to every rank.  The gradients being announced were
// written by earlier kernels of this stream, so they are already visible at the
// point of coherence; the release store orders the announcement after them.
__device__ __forceinline__ void begin_barrier(const CommArgs& a, uint32_t epoch, int vb) {
    if (vb == 0 && threadIdx.x < a.world)
        st_release_sys(a.flags[threadIdx.x] + F_READY + a.rank, epoch);
    if (threadIdx.x < a.world)
        spin_until(a.flags[a.rank] + F_READY + threadIdx.x, epoch, a.timeout_ns, threadIdx.x);
    __syncthreads();
}

// Exit.  Every block takes a ticket; the last one advances the local epoch (and the optimizer's
// device step counter).  With WAIT it first tells every rank "I am done" and waits for the same
// from everyone.  REMOTE_WRITES: one thread per block drains the block's stores to peer memory
// with a .sys fence (cumulative over the CTA barrier in front of it) before taking the ticket,
// so the last block's release store publishes all of them.
template <bool REMOTE_WRITES>
__device__ __forceinline__ void end_barrier(const CommArgs& a, uint32_t epoch, int* step_dev, int new_step,
                                            bool wait, int nvb) {
    __shared__ int s_last;
    __syncthreads();
    uint32_t* mine = a.flags[a.rank];
    if (threadIdx.x == 0) {
        if (REMOTE_WRITES) fence_acq_rel_sys(); else __threadfence();
        const unsigned t = atomicAdd(mine + F_TICKET, 1u);
        s_last = (t == (unsigned)nvb - 1);
        if (s_last) __threadfence();   // acquire side of the ticket chain
    }
    __syncthreads();
    if (!s_last) return;
    if (wait) {
        if (threadIdx.x < a.world) {
            st_release_sys(a.flags[threadIdx.x] + F_DONE + a.rank, epoch);
            spin_until(mine + F_DONE + threadIdx.x, epoch, a.timeout_ns, threadIdx.x);
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        mine[F_TICKET] = 0;
        mine[F_EPOCH] = epoch;
        if (step_dev) step_dev[0] = new_step;
    }
}

// One "vector" is 16 bytes of gradient: 4 fp32 or 8 bf16 elements.
template <bool BF16> struct Vec { static constexpr int N = BF16 ? 8 : 4; };

template <bool BF16>
__device__ __forceinline__ void unpack(const uint4& raw, float (&f)[Vec<BF16>::N]) {
    if constexpr (BF16) {
        const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const float2 t = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[q]));
            f[2 * q] = t.x;
            f[2 * q + 1] = t.y;
        }
    } else {
        f[0] = __uint_as_float(raw.x); f[1] = __uint_as_float(raw.y);
        f[2] = __uint_as_float(raw.z); f[3] = __uint_as_float(raw.w);
    }
}

// Fixed-order (rank 0..W-1) fp32 sum of vector i across all ranks' arenas.
// `poff` is the byte offset of the gradient buffer in use (step parity; 0 when single-buffered).
template <bool BF16>
__device__ __forceinline__ void reduce_vec(const CommArgs& a, int64_t poff, int64_t i, float (&s)[Vec<BF16>::N]) {
    constexpr int N = Vec<BF16>::N;
    uint4 raw[MAX_RANKS];
#pragma unroll
    for (int j = 0; j < MAX_RANKS; ++j)
        if (j < a.world)   // all in flight
            raw[j] = ld_peer_16(reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(a.grad[j]) + poff) + i);
    unpack<BF16>(raw[0], s);
#pragma unroll
    for (int j = 1; j < MAX_RANKS; ++j)
        if (j < a.world) {
            float t[N];
            unpack<BF16>(raw[j], t);
#pragma unroll
            for (int k = 0; k < N; ++k) s[k] += t[k];
        }
}

// NVLS: one load returns the sum over every rank's copy, reduced inside the NVSwitch
// (fp32 accumulation for bf16 inputs).  The address is the multicast mapping of the arenas.
template <bool BF16>
__device__ __forceinline__ void reduce_vec_mc(const void* mc_base, int64_t i, float (&s)[Vec<BF16>::N]) {
    const uint4* p = reinterpret_cast<const uint4*>(mc_base) + i;
    if constexpr (BF16) {
        uint4 raw;
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
                     : "=r"(raw.x), "=r"(raw.y), "=r"(raw.z), "=r"(raw.w)
                     : "l"(p)
                     : "memory");
        unpack<true>(raw, s);
    } else {
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(s[0]), "=f"(s[1]), "=f"(s[2]), "=f"(s[3])
                     : "l"(p)
                     : "memory");
    }
}
// NVLS: one store lands in every rank's copy (the switch replicates it).
__device__ __forceinline__ void st_mc_16(void* p, const uint4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p),
                 "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)),
                 "f"(__uint_as_float(v.w))
                 : "memory");
}
__device__ __forceinline__ void st_mc_8(void* p, const uint2& v) {
    asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1, %2};" ::"l"(p),
                 "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y))
                 : "memory");
}

// Range of vectors this rank handles: the whole bucket (one-shot) or its shard (two-shot).
template <int ALGO>
__device__ __forceinline__ void my_range(const CommArgs& a, int64_t lo, int64_t hi, int64_t& mlo, int64_t& mhi) {
    mlo = lo; mhi = hi;
    if (ALGO >= 1) {
        const int64_t per = (((hi - lo) + a.world - 1) / a.world + 15) & ~int64_t(15);
        mlo = lo + per * a.rank;
        const int64_t e = mlo + per;
        mhi = e < hi ? e : hi;
    }
}

// The reduce + update + push loop of one bucket, for the blocks (vb of nvb) that share it: the body of
// allreduce_optimizer_kernel (comm.cu) and of the update CTAs of the fused SimpleNet step
// (mlp_fused.cu).  algo 0: one-shot, full replica update.  algo 1: two-shot, shard update +
// parameter push.  algo 2: the same through the NVSwitch multicast mapping.
// [start, start+n) is the bucket, in elements; start and n are multiples of the vector width.
template <bool BF16, int ALGO, bool PUSH_BF16>
__device__ __forceinline__ void allreduce_optimizer_loop(
    const CommArgs& a, const b200_opt_hparams& hp, const OptScalars& s, int64_t start, int64_t n,
    float* __restrict__ master, float* __restrict__ m, float* __restrict__ v, __nv_bfloat16* __restrict__ shadow,
    int vb, int nvb) {
    constexpr int N = Vec<BF16>::N;
    // gradient buffer of this step: the producer alternates between two on step parity
    const int64_t poff = ((s.t - 1) & 1) ? a.parity_stride : 0;
    const void* grad_mc = reinterpret_cast<const char*>(a.grad_mc) + poff;
    int64_t lo, hi;
    my_range<ALGO>(a, start / N, (start + n) / N, lo, hi);
    const int64_t stride = (int64_t)nvb * blockDim.x;
    for (int64_t i = lo + (int64_t)vb * blockDim.x + threadIdx.x; i < hi; i += stride) {
        float g[N];
        if (ALGO == 2) reduce_vec_mc<BF16>(grad_mc, i, g);
        else reduce_vec<BF16>(a, poff, i, g);
        float pp[N], mm[N], vv[N];
#pragma unroll
        for (int q = 0; q < N / 4; ++q) {
            const float4 p4 = reinterpret_cast<const float4*>(master)[i * (N / 4) + q];
            const float4 m4 = reinterpret_cast<const float4*>(m)[i * (N / 4) + q];
            float4 v4 = make_float4(0, 0, 0, 0);
            if (hp.kind != 2) v4 = reinterpret_cast<const float4*>(v)[i * (N / 4) + q];
            pp[4 * q] = p4.x; pp[4 * q + 1] = p4.y; pp[4 * q + 2] = p4.z; pp[4 * q + 3] = p4.w;
            mm[4 * q] = m4.x; mm[4 * q + 1] = m4.y; mm[4 * q + 2] = m4.z; mm[4 * q + 3] = m4.w;
            vv[4 * q] = v4.x; vv[4 * q + 1] = v4.y; vv[4 * q + 2] = v4.z; vv[4 * q + 3] = v4.w;
        }
#pragma unroll
        for (int k = 0; k < N; ++k) opt_update(hp.kind, s, pp[k], mm[k], vv[k], g[k] * hp.grad_scale);
#pragma unroll
        for (int q = 0; q < N / 4; ++q) {
            reinterpret_cast<float4*>(m)[i * (N / 4) + q] = make_float4(mm[4 * q], mm[4 * q + 1], mm[4 * q + 2], mm[4 * q + 3]);
            if (hp.kind != 2)
                reinterpret_cast<float4*>(v)[i * (N / 4) + q] = make_float4(vv[4 * q], vv[4 * q + 1], vv[4 * q + 2], vv[4 * q + 3]);
            // fp32 master: local in every mode (two-shot fp32 push overwrites it below, same value)
            reinterpret_cast<float4*>(master)[i * (N / 4) + q] = make_float4(pp[4 * q], pp[4 * q + 1], pp[4 * q + 2], pp[4 * q + 3]);
        }
        uint32_t packed[N / 2];
#pragma unroll
        for (int k = 0; k < N / 2; ++k) packed[k] = pack_bf16(pp[2 * k], pp[2 * k + 1]);
        if (ALGO == 0) {
            if (shadow) {
                if constexpr (N == 8) reinterpret_cast<uint4*>(shadow)[i] = make_uint4(packed[0], packed[1], packed[2], packed[3]);
                else reinterpret_cast<uint2*>(shadow)[i] = make_uint2(packed[0], packed[1]);
            }
        } else if (ALGO == 2) {
            if (PUSH_BF16) {
                if constexpr (N == 8) st_mc_16(reinterpret_cast<uint4*>(a.push_mc) + i, make_uint4(packed[0], packed[1], packed[2], packed[3]));
                else st_mc_8(reinterpret_cast<uint2*>(a.push_mc) + i, make_uint2(packed[0], packed[1]));
            } else {
#pragma unroll
                for (int q = 0; q < N / 4; ++q)
                    st_mc_16(reinterpret_cast<uint4*>(a.push_mc) + i * (N / 4) + q,
                             make_uint4(__float_as_uint(pp[4 * q]), __float_as_uint(pp[4 * q + 1]),
                                        __float_as_uint(pp[4 * q + 2]), __float_as_uint(pp[4 * q + 3])));
            }
        } else if (PUSH_BF16) {
#pragma unroll
            for (int j = 0; j < MAX_RANKS; ++j)
                if (j < a.world) {
                    if constexpr (N == 8) reinterpret_cast<uint4*>(a.push[j])[i] = make_uint4(packed[0], packed[1], packed[2], packed[3]);
                    else reinterpret_cast<uint2*>(a.push[j])[i] = make_uint2(packed[0], packed[1]);
                }
        } else {
#pragma unroll
            for (int j = 0; j < MAX_RANKS; ++j)
                if (j < a.world && j != a.rank) {
#pragma unroll
                    for (int q = 0; q < N / 4; ++q)
                        reinterpret_cast<float4*>(a.push[j])[i * (N / 4) + q] =
                            make_float4(pp[4 * q], pp[4 * q + 1], pp[4 * q + 2], pp[4 * q + 3]);
                }
        }
        // Sharded update with a bf16 push: the fp32 master is current only on the shard owner.
        // Parameters that forward reads in fp32 (biases, norm scales / shifts) are marked in
        // f32_mask; their fp32 values go to every rank's master as well, so all replicas compute
        // with the same numbers (they are a fraction of a percent of the bytes).
        if (ALGO >= 1 && PUSH_BF16 && a.f32_mask && a.f32_mask[(i * N) >> 3]) {
#pragma unroll
            for (int j = 0; j < MAX_RANKS; ++j)
                if (j < a.world && j != a.rank) {
#pragma unroll
                    for (int q = 0; q < N / 4; ++q)
                        reinterpret_cast<float4*>(a.master_peer[j])[i * (N / 4) + q] =
                            make_float4(pp[4 * q], pp[4 * q + 1], pp[4 * q + 2], pp[4 * q + 3]);
                }
        }
    }
}

// Host: b200_comm (public descriptor) -> CommArgs, with the argument checks (comm.cu).
int fill_comm_args(const b200_comm* c, CommArgs* a);

}  // namespace b200
