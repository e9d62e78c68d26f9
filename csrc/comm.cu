// Peer-memory collectives over NVLink 5 / NVSwitch, fused with the optimizer.
//
// Replaces torch DDP's Reducer + ProcessGroup all-reduce + optim.Adam for the
// gradient path (ref train.py:138 backward hooks -> all-reduce, train.py:139
// optimizer.step).  One kernel per bucket:
//
//   one-shot  (small buckets)  every rank loads the bucket from all W ranks'
//             gradient arenas through peer-mapped pointers, sums in rank order in
//             fp32 registers, scales by 1/W and applies Adam/AdamW/SGD to its own
//             full replica.  One NVLink round trip; optimizer work replicated.
//   two-shot  (large buckets)  rank r pull-reduces shard r, updates only that
//             shard of master/m/v (optimizer state and work divided by W), and
//             pushes the *updated parameters* to every rank's parameter arena —
//             the all-gather moves new weights (bf16: 2 B) instead of gradients.
//
// The reduction order is fixed (rank 0..W-1), so every replica computes bit-
// identical sums and replicas never drift.
//
// Synchronisation: flag barriers over peer memory with monotonically increasing
// epochs (never reset, so no reset race), release stores / acquire loads at .sys
// scope, and a bounded spin: a dead peer raises a trap on the survivors instead of
// wedging seven GPUs.
//   entry  "my gradients are ready": one release store per peer, then every block
//          polls the local copies of the peers' flags (relaxed loads; the last one
//          is an acquire load).  No stand-alone fence: the gradients were written by
//          earlier kernels of the stream and are already at the point of coherence.
//   exit   "I am done reading / my pushes have landed": only where a peer could
//          otherwise be overtaken — kernels that write peer memory (two-shot, NVLS)
//          and single-buffered one-shot.  One .sys fence per block drains that
//          block's remote stores, the last block publishes and waits.
//   When the producer double-buffers the gradient arena on step parity
//   (B200_COMM_NO_EXIT_WAIT), one-shot needs no exit barrier at all: a peer can only
//   overwrite buffer t%2 in step t+2, after its step t+1 kernel saw OUR step t+1
//   entry flag, which we store after this kernel's reads have completed.
//
//   NVLS      (algo 2, when the arenas are bound to an NVSwitch multicast object)  same
//             sharding as two-shot, but the reduce-scatter is ONE multimem.ld_reduce per
//             vector (summed in the switch) and the parameter all-gather ONE multimem.st
//             (replicated by the switch): link traffic per rank drops from
//             2 S (W-1)/W to about S/W + S_param/W.
//
// Also here: the symmetric allocator (cudaMalloc + CUDA IPC handles exchanged by
// the Python process group) and a plain in-kernel all-reduce for benchmarking.
#include "comm_device.cuh"

#include <cstdlib>

void b200_count_launch(int n);


namespace b200 {

// algo 0: one-shot, full replica update.  algo 1: two-shot, shard update + parameter push.
// [start, start+n) is the bucket, in elements; start and n are multiples of the vector width.
template <bool BF16, int ALGO, bool PUSH_BF16>
__global__ void __launch_bounds__(256) allreduce_optimizer_kernel(
    const CommArgs a, b200_opt_hparams hp, int64_t start, int64_t n, float* __restrict__ master,
    float* __restrict__ m, float* __restrict__ v, __nv_bfloat16* __restrict__ shadow, int* step_dev,
    int step_host, int advance, const OptScalars* __restrict__ scalars) {
    __shared__ OptScalars sc;
    // Programmatic dependent launch (no-ops on an ordinary launch): become resident while the
    // backward kernel is still running, start when it completes.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const uint32_t epoch = *(volatile uint32_t*)(a.flags[a.rank] + F_EPOCH) + 1u;
    // thread 0 also publishes the "ready" flag: with precomputed `scalars` (csrc/mlp_fused.cu) the
    // two double-precision pow() of make_opt_scalars no longer sit in front of the barrier
    if (threadIdx.x == 0) {
        if (scalars) {
            sc = *scalars;
        } else {
            const int t = step_dev ? (*(volatile int*)step_dev + 1) : step_host;
            sc = make_opt_scalars(hp, t);
        }
    }
    begin_barrier(a, epoch, (int)blockIdx.x);  // includes __syncthreads
    const OptScalars s = sc;

    allreduce_optimizer_loop<BF16, ALGO, PUSH_BF16>(a, hp, s, start, n, master, m, v, shadow, (int)blockIdx.x,
                                                    (int)gridDim.x);
    end_barrier<ALGO >= 1>(a, epoch, advance ? step_dev : nullptr, s.t, ALGO >= 1 || !a.no_exit_wait, (int)gridDim.x);
}

// Plain all-reduce (sum * scale), for the bandwidth sweep and generic use.
// algo 0: one-shot into `out` (local).  algo 1: two-shot in place (pull shard, push to all).
template <bool BF16, int ALGO>
__global__ void __launch_bounds__(256) allreduce_kernel(const CommArgs a, int64_t start, int64_t n,
                                                        void* __restrict__ out, float scale) {
    constexpr int N = Vec<BF16>::N;
    const uint32_t epoch = *(volatile uint32_t*)(a.flags[a.rank] + F_EPOCH) + 1u;
    begin_barrier(a, epoch, (int)blockIdx.x);
    int64_t lo, hi;
    my_range<ALGO>(a, start / N, (start + n) / N, lo, hi);
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += stride) {
        float g[N];
        if (ALGO == 2) reduce_vec_mc<BF16>(a.grad_mc, i, g);
        else reduce_vec<BF16>(a, 0, i, g);
        uint4 o;
        if constexpr (BF16) {
            o = make_uint4(pack_bf16(g[0] * scale, g[1] * scale), pack_bf16(g[2] * scale, g[3] * scale),
                           pack_bf16(g[4] * scale, g[5] * scale), pack_bf16(g[6] * scale, g[7] * scale));
        } else {
            o = make_uint4(__float_as_uint(g[0] * scale), __float_as_uint(g[1] * scale),
                           __float_as_uint(g[2] * scale), __float_as_uint(g[3] * scale));
        }
        if (ALGO == 0) reinterpret_cast<uint4*>(out)[i] = o;
        else if (ALGO == 2) st_mc_16(reinterpret_cast<uint4*>(const_cast<void*>(a.grad_mc)) + i, o);
        else {
#pragma unroll
            for (int j = 0; j < MAX_RANKS; ++j)
                if (j < a.world) reinterpret_cast<uint4*>(const_cast<void*>(a.grad[j]))[i] = o;
        }
    }
    end_barrier<ALGO >= 1>(a, epoch, nullptr, 0, true, (int)gridDim.x);
}

// Rank `root` replicates [off, off + nvec) 16-byte vectors of its `push` arena into every rank's
// (DDP's per-forward buffer broadcast, SURVEY K4: ref train.py:136 -> torch _sync_buffers).  Entry
// barrier: nobody is still reading the old values; exit barrier: the new ones have landed.
// One multimem.st per vector when the arena has a multicast mapping, else one store per peer.
__global__ void __launch_bounds__(256) broadcast_kernel(const CommArgs a, int root, int64_t off, int64_t nvec) {
    const uint32_t epoch = *(volatile uint32_t*)(a.flags[a.rank] + F_EPOCH) + 1u;
    begin_barrier(a, epoch, (int)blockIdx.x);
    if (a.rank == root) {
        const uint4* src = reinterpret_cast<const uint4*>(a.push[root]) + off;
        const int64_t stride = (int64_t)gridDim.x * blockDim.x;
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
            const uint4 v = src[i];
            if (a.push_mc) {
                st_mc_16(reinterpret_cast<uint4*>(a.push_mc) + off + i, v);
            } else {
#pragma unroll
                for (int j = 0; j < MAX_RANKS; ++j)
                    if (j < a.world && j != root) (reinterpret_cast<uint4*>(a.push[j]) + off)[i] = v;
            }
        }
    }
    end_barrier<true>(a, epoch, nullptr, 0, true, (int)gridDim.x);
}

int fill_comm_args(const b200_comm* c, CommArgs* a) {
    B200_REQUIRE(c && c->world >= 1 && c->world <= MAX_RANKS, "comm: world size must be in [1, %d]", MAX_RANKS);
    B200_REQUIRE(c->rank >= 0 && c->rank < c->world, "comm: bad rank %d", c->rank);
    a->rank = c->rank;
    a->world = c->world;
    for (int j = 0; j < MAX_RANKS; ++j) {
        a->grad[j] = j < c->world ? c->grad[j] : nullptr;
        a->push[j] = j < c->world ? c->push[j] : nullptr;
        a->flags[j] = j < c->world ? (uint32_t*)c->flags[j] : nullptr;
        if (j < c->world) B200_REQUIRE(a->grad[j] && a->flags[j], "comm: missing peer pointer for rank %d", j);
    }
    a->grad_mc = c->grad_mc;
    a->push_mc = c->push_mc;
    a->f32_mask = (const uint8_t*)c->f32_mask;
    for (int j = 0; j < MAX_RANKS; ++j) {
        a->master_peer[j] = (j < c->world && c->f32_mask) ? (float*)c->master_peer[j] : nullptr;
        if (j < c->world && c->f32_mask)
            B200_REQUIRE(a->master_peer[j], "comm: f32_mask needs the peer-mapped master of rank %d", j);
    }
    a->parity_stride = c->parity_stride;
    a->no_exit_wait = c->no_exit_wait;
    B200_REQUIRE(c->parity_stride >= 0 && (c->parity_stride % 16) == 0, "comm: parity_stride must be a multiple of 16");
    B200_REQUIRE(!c->no_exit_wait || c->parity_stride > 0,
                 "comm: no_exit_wait is only safe on a double-buffered gradient arena (parity_stride > 0)");
    a->timeout_ns = c->timeout_ms > 0 ? (uint64_t)c->timeout_ms * 1000000ull : 30000000000ull;
    return 0;
}

static int grid_for(int64_t vecs, int max_blocks) {
    int64_t b = (vecs + 255) / 256;
    if (b < 1) b = 1;
    if (b > max_blocks) b = max_blocks;
    return (int)b;
}

}  // namespace b200

// ------------------------------------------------------------------ symmetric memory
extern "C" int b200_symm_alloc(int64_t bytes, void** ptr, void* handle64) {
    B200_REQUIRE(bytes > 0 && ptr && handle64, "symm_alloc: bad arguments");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    B200_CUDA_OK(cudaMalloc(ptr, (size_t)bytes));
    B200_CUDA_OK(cudaMemset(*ptr, 0, (size_t)bytes));
    cudaIpcMemHandle_t h;
    B200_CUDA_OK(cudaIpcGetMemHandle(&h, *ptr));
    memcpy(handle64, &h, 64);
    return 0;
}

extern "C" int b200_symm_open(const void* handle64, void** ptr) {
    B200_REQUIRE(handle64 && ptr, "symm_open: bad arguments");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    B200_CUDA_OK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}

extern "C" int b200_symm_close(void* ptr) {
    B200_CUDA_OK(cudaIpcCloseMemHandle(ptr));
    return 0;
}

extern "C" int b200_symm_free(void* ptr) {
    B200_CUDA_OK(cudaFree(ptr));
    return 0;
}

// ------------------------------------------------------------------ collectives
static int allreduce_optimizer_impl(const b200_comm* c, int algo, const b200_opt_hparams* hp,
                                    int64_t start, int64_t n, int grad_is_bf16, float* master,
                                    float* exp_avg, float* exp_avg_sq, void* shadow_bf16,
                                    int push_is_bf16, int* step_dev, int step_host, int advance,
                                    void* stream, bool pdl, const void* scalars = nullptr) {
    using namespace b200;
    CommArgs a;
    int rc = fill_comm_args(c, &a);
    if (rc) return rc;
    B200_REQUIRE(hp && master && exp_avg, "allreduce_optimizer: null pointer");
    B200_REQUIRE(algo >= 0 && algo <= 2, "allreduce_optimizer: unknown algo %d", algo);
    B200_REQUIRE(algo != 2 || (c->grad_mc && c->push_mc), "allreduce_optimizer: NVLS needs multicast addresses");
    B200_REQUIRE((start % 8) == 0 && (n % 8) == 0 && n > 0, "allreduce_optimizer: start/n must be multiples of 8");
    if (algo == 1)
        for (int j = 0; j < c->world; ++j) B200_REQUIRE(a.push[j], "allreduce_optimizer: two-shot needs push pointers");
    const int64_t nv = n / (grad_is_bf16 ? 8 : 4);
    const int64_t vecs = (algo >= 1 ? (nv + c->world - 1) / c->world : nv);
    const int blocks = grid_for(vecs, c->max_blocks > 0 ? c->max_blocks : 296);
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
#define GO(BF, AL, PB)                                                                             \
    B200_CUDA_OK(cudaLaunchKernelEx(&cfg, allreduce_optimizer_kernel<BF, AL, PB>, a, h, start, n,  \
                                    master, exp_avg, exp_avg_sq, sh, step_dev, step_host, advance,  \
                                    (const OptScalars*)scalars))
    if (algo == 0) { if (grad_is_bf16) GO(true, 0, false); else GO(false, 0, false); }
    else if (algo == 2) {
        if (push_is_bf16) { if (grad_is_bf16) GO(true, 2, true); else GO(false, 2, true); }
        else { if (grad_is_bf16) GO(true, 2, false); else GO(false, 2, false); }
    }
    else if (push_is_bf16) { if (grad_is_bf16) GO(true, 1, true); else GO(false, 1, true); }
    else { if (grad_is_bf16) GO(true, 1, false); else GO(false, 1, false); }
#undef GO
    b200_count_launch(1);
    return 0;
}

extern "C" int b200_allreduce_optimizer(const b200_comm* c, int algo, const b200_opt_hparams* hp,
                                        int64_t start, int64_t n, int grad_is_bf16, float* master,
                                        float* exp_avg, float* exp_avg_sq, void* shadow_bf16,
                                        int push_is_bf16, int* step_dev, int step_host, int advance,
                                        void* stream) {
    return allreduce_optimizer_impl(c, algo, hp, start, n, grad_is_bf16, master, exp_avg, exp_avg_sq, shadow_bf16,
                                    push_is_bf16, step_dev, step_host, advance, stream, false);
}

// Launched with programmatic stream serialization (used behind the fused MLP step kernel).
int b200_allreduce_optimizer_pdl(const b200_comm* c, int algo, const b200_opt_hparams* hp, int64_t start, int64_t n,
                                 int grad_is_bf16, float* master, float* exp_avg, float* exp_avg_sq,
                                 void* shadow_bf16, int push_is_bf16, int* step_dev, int step_host, int advance,
                                 const void* scalars, void* stream) {
    static const bool pdl = getenv("B200DDP_NO_PDL") == nullptr;
    return allreduce_optimizer_impl(c, algo, hp, start, n, grad_is_bf16, master, exp_avg, exp_avg_sq, shadow_bf16,
                                    push_is_bf16, step_dev, step_host, advance, stream, pdl, scalars);
}

extern "C" int b200_allreduce(const b200_comm* c, int algo, int64_t start, int64_t n, int is_bf16,
                              void* out, float scale, void* stream) {
    using namespace b200;
    CommArgs a;
    int rc = fill_comm_args(c, &a);
    if (rc) return rc;
    B200_REQUIRE(algo >= 0 && algo <= 2, "allreduce: unknown algo %d", algo);
    B200_REQUIRE(algo != 2 || c->grad_mc, "allreduce: NVLS needs the multicast address");
    B200_REQUIRE((start % 8) == 0 && (n % 8) == 0 && n > 0, "allreduce: start/n must be multiples of 8");
    B200_REQUIRE(algo >= 1 || out, "allreduce: one-shot needs an output buffer");
    const int64_t nv = n / (is_bf16 ? 8 : 4);
    const int64_t vecs = (algo >= 1 ? (nv + c->world - 1) / c->world : nv);
    const int blocks = grid_for(vecs, c->max_blocks > 0 ? c->max_blocks : 296);
    cudaStream_t s = (cudaStream_t)stream;
    if (is_bf16) {
        if (algo == 0) allreduce_kernel<true, 0><<<blocks, 256, 0, s>>>(a, start, n, out, scale);
        else if (algo == 1) allreduce_kernel<true, 1><<<blocks, 256, 0, s>>>(a, start, n, out, scale);
        else allreduce_kernel<true, 2><<<blocks, 256, 0, s>>>(a, start, n, out, scale);
    } else {
        if (algo == 0) allreduce_kernel<false, 0><<<blocks, 256, 0, s>>>(a, start, n, out, scale);
        else if (algo == 1) allreduce_kernel<false, 1><<<blocks, 256, 0, s>>>(a, start, n, out, scale);
        else allreduce_kernel<false, 2><<<blocks, 256, 0, s>>>(a, start, n, out, scale);
    }
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int b200_broadcast(const b200_comm* c, int root, int64_t offset_bytes, int64_t nbytes, void* stream) {
    using namespace b200;
    CommArgs a;
    int rc = fill_comm_args(c, &a);
    if (rc) return rc;
    B200_REQUIRE(root >= 0 && root < c->world, "broadcast: bad root %d", root);
    B200_REQUIRE(offset_bytes >= 0 && nbytes > 0 && (offset_bytes % 16) == 0 && (nbytes % 16) == 0,
                 "broadcast: offset and size must be multiples of 16 bytes");
    for (int j = 0; j < c->world; ++j) B200_REQUIRE(a.push[j], "broadcast: missing arena pointer of rank %d", j);
    const int64_t nvec = nbytes / 16;
    const int blocks = c->rank == root ? grid_for(nvec, c->max_blocks > 0 ? c->max_blocks : 148) : 1;
    broadcast_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(a, root, offset_bytes / 16, nvec);
    b200_count_launch(1);
    B200_CUDA_OK(cudaGetLastError());
    return 0;
}
