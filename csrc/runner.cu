// Host-side step runner: one C call per training step for the end-to-end path.
//
// The reference's loop does, per batch, `.to(device)` x2, the step, and `loss.item()`
// (ref train.py:133,141).  Through Python even our graph-replayed step costs ~70 us of
// interpreter and dispatcher time per step (two H2D copies, four events, the staging launch, the
// replay, the loss read-back), more than the 50 us the GPU needs: the end-to-end rate was host
// bound.  This runner issues the same work from C++:
//
//   copy stream   H2D of this step's inputs from pinned host memory into a ring slot
//   main stream   wait(copied) -> staging (fp32->bf16 + label copy + accumulator clears for the MLP
//                 program, or D2D into a captured step's static placeholders) -> cudaGraphLaunch
//                 -> a one-thread kernel stores the loss and a sequence number into pinned host
//                 memory (a DMA copy here would cost an engine switch on the critical stream:
//                 measured 60 vs 53 us per step)
//   host          polls for the PREVIOUS step's sequence number (so one step is always queued
//                 behind the one running) and hands its loss back
//
// Nothing is allocated after b200_runner_create.
#include "common.cuh"

#include <chrono>
#include <new>

namespace {

constexpr int MAX_SLOTS = 4;
constexpr int LOSS_RING = 8;

struct LossSlot {
    float loss;
    unsigned seq;   // step number + 1, written after the loss
};

struct Runner {
    b200_runner_desc d;
    cudaStream_t main, copy;
    void* dev_in[MAX_SLOTS][2];
    cudaEvent_t copied[MAX_SLOTS], consumed[MAX_SLOTS];
    LossSlot* host_loss;              // pinned + mapped, LOSS_RING entries
    int64_t step;
    float* zero_ptrs[B200_STAGE_MAX_ZERO];
    int zero_counts[B200_STAGE_MAX_ZERO];
};

__global__ void publish_loss_kernel(const float* __restrict__ loss_dev, LossSlot* slot, unsigned seq) {
    if (threadIdx.x == 0) {
        slot->loss = *loss_dev;
        __threadfence_system();
        *reinterpret_cast<volatile unsigned*>(&slot->seq) = seq;
    }
}

// Spins until step `step`'s loss has landed in host memory (bounded: a dead GPU must not hang the host).
int wait_loss(Runner* r, int64_t step, float* out) {
    volatile LossSlot* slot = r->host_loss + (step % LOSS_RING);
    const unsigned want = (unsigned)(step + 1);
    const auto t0 = std::chrono::steady_clock::now();
    for (uint64_t spins = 0; slot->seq != want; ++spins) {
        if ((spins & 0xffff) == 0xffff) {
            const cudaError_t q = cudaStreamQuery(r->main);   // surfaces a trapped kernel
            if (q != cudaSuccess && q != cudaErrorNotReady) {
                b200_set_error("runner: stream failed while waiting for step %lld: %s", (long long)step,
                               cudaGetErrorString(q));
                return (int)q;
            }
            if (std::chrono::steady_clock::now() - t0 > std::chrono::seconds(60)) {
                b200_set_error("runner: step %lld did not publish its loss within 60 s", (long long)step);
                return B200_ERR_INVALID;
            }
        }
    }
    *out = slot->loss;
    return 0;
}

}  // namespace

extern "C" int b200_runner_create(void** out, const b200_runner_desc* d) {
    B200_REQUIRE(out && d, "runner_create: null argument");
    // stream may be 0: torch's default stream is the legacy default stream
    B200_REQUIRE(d->graph_exec && d->loss_dev, "runner_create: graph_exec and loss_dev are required");
    B200_REQUIRE(d->slots >= 2 && d->slots <= MAX_SLOTS, "runner_create: slots must be 2..%d", MAX_SLOTS);
    B200_REQUIRE(d->in_bytes[0] > 0 && d->in_bytes[1] >= 0, "runner_create: bad input sizes");
    B200_REQUIRE(d->mode == 0 || d->mode == 1, "runner_create: mode must be 0 (copy) or 1 (MLP staging)");
    B200_REQUIRE(d->dst[0] && (d->in_bytes[1] == 0 || d->dst[1]), "runner_create: destination pointers missing");
    B200_REQUIRE(d->nzero >= 0 && d->nzero <= B200_STAGE_MAX_ZERO, "runner_create: at most %d zero regions", B200_STAGE_MAX_ZERO);
    if (d->mode == 1)
        B200_REQUIRE(d->in_bytes[0] % 16 == 0 && d->in_bytes[1] % 8 == 0, "runner_create: MLP staging needs fp32 x / int64 y");
    Runner* r = new (std::nothrow) Runner();
    B200_REQUIRE(r != nullptr, "runner_create: out of host memory");
    r->d = *d;
    r->main = (cudaStream_t)d->stream;
    r->step = 0;
    for (int k = 0; k < d->nzero; ++k) { r->zero_ptrs[k] = d->zero_ptrs[k]; r->zero_counts[k] = d->zero_counts[k]; }
    B200_CUDA_OK(cudaStreamCreateWithFlags(&r->copy, cudaStreamNonBlocking));
    for (int s = 0; s < d->slots; ++s) {
        for (int k = 0; k < 2; ++k) {
            r->dev_in[s][k] = nullptr;
            if (d->in_bytes[k] > 0) B200_CUDA_OK(cudaMalloc(&r->dev_in[s][k], (size_t)d->in_bytes[k]));
        }
        B200_CUDA_OK(cudaEventCreateWithFlags(&r->copied[s], cudaEventDisableTiming));
        B200_CUDA_OK(cudaEventCreateWithFlags(&r->consumed[s], cudaEventDisableTiming));
    }
    B200_CUDA_OK(cudaHostAlloc((void**)&r->host_loss, LOSS_RING * sizeof(LossSlot), cudaHostAllocMapped));
    memset(r->host_loss, 0, LOSS_RING * sizeof(LossSlot));
    *out = r;
    return 0;
}

// Runs one step on the pinned host inputs `h0` (and `h1`).  `*prev_valid` tells whether
// `*prev_loss` holds the loss of the step before this one (false on the first call).
extern "C" int b200_runner_step(void* handle, const void* h0, const void* h1, float* prev_loss, int* prev_valid) {
    Runner* r = (Runner*)handle;
    B200_REQUIRE(r && h0 && (r->d.in_bytes[1] == 0 || h1) && prev_loss && prev_valid, "runner_step: null argument");
    const b200_runner_desc& d = r->d;
    const int s = (int)(r->step % d.slots);
    if (r->step >= d.slots) B200_CUDA_OK(cudaStreamWaitEvent(r->copy, r->consumed[s], 0));
    B200_CUDA_OK(cudaMemcpyAsync(r->dev_in[s][0], h0, (size_t)d.in_bytes[0], cudaMemcpyHostToDevice, r->copy));
    if (d.in_bytes[1] > 0)
        B200_CUDA_OK(cudaMemcpyAsync(r->dev_in[s][1], h1, (size_t)d.in_bytes[1], cudaMemcpyHostToDevice, r->copy));
    B200_CUDA_OK(cudaEventRecord(r->copied[s], r->copy));
    B200_CUDA_OK(cudaStreamWaitEvent(r->main, r->copied[s], 0));
    if (d.mode == 1) {
        if (int rc = b200_stage_batch((const float*)r->dev_in[s][0], (const int64_t*)r->dev_in[s][1], d.dst[0],
                                      (int64_t*)d.dst[1], d.in_bytes[0] / 4, d.in_bytes[1] / 8, r->zero_ptrs,
                                      r->zero_counts, d.nzero, r->main))
            return rc;
    } else {
        B200_CUDA_OK(cudaMemcpyAsync(d.dst[0], r->dev_in[s][0], (size_t)d.in_bytes[0], cudaMemcpyDeviceToDevice, r->main));
        if (d.in_bytes[1] > 0)
            B200_CUDA_OK(cudaMemcpyAsync(d.dst[1], r->dev_in[s][1], (size_t)d.in_bytes[1], cudaMemcpyDeviceToDevice, r->main));
    }
    B200_CUDA_OK(cudaEventRecord(r->consumed[s], r->main));
    B200_CUDA_OK(cudaGraphLaunch((cudaGraphExec_t)d.graph_exec, r->main));
    publish_loss_kernel<<<1, 32, 0, r->main>>>(d.loss_dev, r->host_loss + (r->step % LOSS_RING), (unsigned)(r->step + 1));
    B200_CUDA_OK(cudaGetLastError());
    *prev_valid = 0;
    if (r->step > 0) {
        if (int rc = wait_loss(r, r->step - 1, prev_loss)) return rc;
        *prev_valid = 1;
    }
    r->step += 1;
    return 0;
}

// Waits for the last submitted step and returns its loss.
extern "C" int b200_runner_finish(void* handle, float* last_loss, int* valid) {
    Runner* r = (Runner*)handle;
    B200_REQUIRE(r && last_loss && valid, "runner_finish: null argument");
    *valid = 0;
    if (r->step > 0) {
        if (int rc = wait_loss(r, r->step - 1, last_loss)) return rc;
        *valid = 1;
    }
    return 0;
}

extern "C" int b200_runner_destroy(void* handle) {
    Runner* r = (Runner*)handle;
    if (!r) return 0;
    cudaStreamSynchronize(r->main);
    cudaStreamSynchronize(r->copy);
    for (int s = 0; s < r->d.slots; ++s) {
        for (int k = 0; k < 2; ++k)
            if (r->dev_in[s][k]) cudaFree(r->dev_in[s][k]);
        cudaEventDestroy(r->copied[s]);
        cudaEventDestroy(r->consumed[s]);
    }
    cudaFreeHost(r->host_loss);
    cudaStreamDestroy(r->copy);
    delete r;
    return 0;
}
