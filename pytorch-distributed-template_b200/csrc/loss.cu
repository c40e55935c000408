// nll_loss (reference model/loss.py:4-5) fused with the correct-counts of accuracy / top_k_acc
// (model/metric.py:4-20): one pass over logp, warp-shuffle reductions, integer counts.
// HBM bound by construction (48 B read per row for 10 classes); at n <= 4096 it is a
// single-wave, latency-bound launch.
#include "common.cuh"

namespace pdt {

constexpr int NLL_THREADS = 256;

// scratch layout: [0] ticket (uint32), [16 ...] nll_partial[grid]
__global__ void __launch_bounds__(NLL_THREADS)
nll_metrics_kernel(const float* __restrict__ logp, const int64_t* __restrict__ target, int64_t n, int classes,
                   int topk, int64_t ignore_index, pdt_stats* __restrict__ out, unsigned* __restrict__ ticket,
                   nll_partial* __restrict__ partial, float* __restrict__ ring, int32_t* __restrict__ ring_count,
                   int ring_cap) {
    pdl_wait();       // logp is the previous kernel's output
    float loss = 0.f;
    int c1 = 0, ck = 0, nv = 0;
    for (int64_t r = blockIdx.x * (int64_t)NLL_THREADS + threadIdx.x; r < n; r += (int64_t)gridDim.x * NLL_THREADS) {
        const int64_t t = target[r];
        if (t >= 0 && t < classes) {
            const float* row = logp + r * classes;
            const float ot = row[t];
            int ahead = 0;
            for (int j = 0; j < classes; ++j) {
                const float o = row[j];
                ahead += (o > ot) || (o == ot && j < t);
            }
            c1 += ahead < 1;
            ck += ahead < topk;
            if (t != ignore_index) {
                loss -= ot;
                nv += 1;
            }
        }
    }
    pdl_trigger();
    loss = warp_sum(loss);
    c1 = warp_sum_i(c1);
    ck = warp_sum_i(ck);
    nv = warp_sum_i(nv);
    __shared__ nll_partial sp[NLL_THREADS / 32];
    __shared__ bool last;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) sp[warp] = {loss, c1, ck, nv};
    __syncthreads();
    if (threadIdx.x == 0) {
        nll_partial p = sp[0];
        for (int w = 1; w < NLL_THREADS / 32; ++w) {
            p.loss += sp[w].loss; p.top1 += sp[w].top1; p.topk += sp[w].topk; p.valid += sp[w].valid;
        }
        partial[blockIdx.x] = p;
        __threadfence();
        last = atomicInc(ticket, gridDim.x - 1) == gridDim.x - 1;
    }
    __syncthreads();
    if (last && threadIdx.x == 0) {
        __threadfence();
        nll_partial p = {0.f, 0, 0, 0};
        for (unsigned g = 0; g < gridDim.x; ++g) {  // fixed order: deterministic loss
            const volatile nll_partial* q = partial + g;
            p.loss += q->loss; p.top1 += q->top1; p.topk += q->topk; p.valid += q->valid;
        }
        out->loss_sum = p.loss;
        out->loss_mean = p.loss / (float)p.valid;  // 0/0 = nan, like F.nll_loss on all-ignored input
        out->top1 = p.top1;
        out->topk = p.topk;
        out->n_valid = p.valid;
        out->n_rows = (int32_t)n;
        if (ring != nullptr) {      // the trainer's per-step loss log (same effect as pdt_ring_push)
            const int32_t c = *ring_count;
            ring[c % ring_cap] = p.loss / (float)p.valid;
            *ring_count = c + 1;
        }
    }
}

__global__ void nll_bwd_kernel(const int64_t* __restrict__ target, const pdt_stats* __restrict__ stats,
                               const float* __restrict__ grad_loss, float* __restrict__ dlogp, int64_t n,
                               int classes, int64_t ignore_index) {
    pdl_wait();
    pdl_trigger();
    const float g = (grad_loss ? *grad_loss : 1.f) / (float)stats->n_valid;
    const int64_t total = n * classes;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / classes;
        const int c = (int)(i - r * classes);
        const int64_t t = target[r];
        dlogp[i] = (t == c && t != ignore_index) ? -g : 0.f;
    }
}

// Scratch for pdt_nll_metrics_fwd_static: ticket (first 16 bytes) + one nll_partial per block.
// Lives in the module image, so every GPU that loads the library carries these 2 KB until exit.
constexpr int NLL_STATIC_MAX_BLOCKS = 64;
constexpr int64_t NLL_STATIC_MAX_ROWS = (int64_t)NLL_STATIC_MAX_BLOCKS * NLL_THREADS;   // 16,384
static __device__ __align__(256) unsigned char g_nll_static_scratch[2048];
static_assert(16 + NLL_STATIC_MAX_BLOCKS * sizeof(nll_partial) <= sizeof(g_nll_static_scratch), "scratch too small");
static_assert(NLL_STATIC_MAX_ROWS == PDT_NLL_STATIC_MAX_ROWS, "the limit in pdt.h and the one enforced here must agree");

static int nll_grid(int64_t n) {
    int64_t g = (n + NLL_THREADS - 1) / NLL_THREADS;
    return (int)(g < 1 ? 1 : (g > 592 ? 592 : g));
}

}  // namespace pdt

using namespace pdt;

extern "C" int64_t pdt_nll_scratch_bytes(int64_t n) { return 16 + (int64_t)nll_grid(n) * sizeof(nll_partial); }

extern "C" int pdt_nll_metrics_fwd(const float* logp, const int64_t* target, int64_t n, int classes, int topk,
                                   int64_t ignore_index, pdt_stats* out, void* scratch, void* stream) {
    PDT_REQUIRE(n > 0 && logp && target && out && scratch, "null pointer or n <= 0");
    PDT_REQUIRE(classes > 0 && topk > 0, "classes and topk must be positive");
    unsigned* ticket = reinterpret_cast<unsigned*>(scratch);
    nll_partial* partial = reinterpret_cast<nll_partial*>(reinterpret_cast<char*>(scratch) + 16);
    launch_pdl(4, nll_metrics_kernel, dim3(nll_grid(n)), dim3(NLL_THREADS), 0, stream, logp, target, n, classes, topk,
               ignore_index, out, ticket, partial, (float*)nullptr, (int32_t*)nullptr, 0);
    return check_launch("nll_metrics_fwd");
}

extern "C" int pdt_nll_metrics_fwd_log(const float* logp, const int64_t* target, int64_t n, int classes, int topk,
                                       int64_t ignore_index, pdt_stats* out, void* scratch, float* ring,
                                       int32_t* ring_count, int ring_capacity, void* stream) {
    PDT_REQUIRE(n > 0 && logp && target && out && scratch && ring && ring_count, "null pointer or n <= 0");
    PDT_REQUIRE(classes > 0 && topk > 0 && ring_capacity > 0, "classes, topk and ring_capacity must be positive");
    unsigned* ticket = reinterpret_cast<unsigned*>(scratch);
    nll_partial* partial = reinterpret_cast<nll_partial*>(reinterpret_cast<char*>(scratch) + 16);
    launch_pdl(4, nll_metrics_kernel, dim3(nll_grid(n)), dim3(NLL_THREADS), 0, stream, logp, target, n, classes, topk,
               ignore_index, out, ticket, partial, ring, ring_count, ring_capacity);
    return check_launch("nll_metrics_fwd_log");
}

extern "C" int pdt_nll_bwd(const int64_t* target, const pdt_stats* stats, const float* grad_loss, float* dlogp,
                           int64_t n, int classes, int64_t ignore_index, void* stream) {
    PDT_REQUIRE(n > 0 && target && stats && dlogp && classes > 0, "null pointer or bad size");
    const int64_t total = n * classes;
    int64_t g = (total + 255) / 256;
    if (g > 1184) g = 1184;
    launch_pdl(5, nll_bwd_kernel, dim3((unsigned)g), dim3(256), 0, stream, target, stats, grad_loss, dlogp, n, classes,
               ignore_index);
    return check_launch("nll_bwd");
}

// No-allocation variant.  In host code `g_nll_static_scratch` is only a handle for the symbol, not
// an address the GPU can use; the device address is looked up once per device (with that device
// current) and cached.  The cache is plain process-wide state: the first call on a device must not run
// concurrently with another host thread's first call on it (later calls only read the cache), and the
// cached address does not survive a cudaDeviceReset -- a process that resets the device must not call
// this entry point again.
extern "C" int pdt_nll_metrics_fwd_static(const float* logp, const int64_t* target, int64_t n, int classes,
                                          int topk, int64_t ignore_index, pdt_stats* out, void* stream) {
    PDT_REQUIRE(n > 0 && logp && target && out, "null pointer or n <= 0");
    PDT_REQUIRE(classes > 0 && topk > 0, "classes and topk must be positive");
    if (n > NLL_STATIC_MAX_ROWS) {
        set_error("pdt_nll_metrics_fwd_static: n = %lld exceeds the built-in scratch (at most %lld rows per call); "
                  "use pdt_nll_metrics_fwd with caller-provided scratch",
                  (long long)n, (long long)NLL_STATIC_MAX_ROWS);
        return PDT_EINVAL;
    }
    static void* dev_ptr[64] = {nullptr};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return check_launch("cudaGetDevice");
    PDT_REQUIRE(dev >= 0 && dev < 64, "device index out of range");
    if (dev_ptr[dev] == nullptr) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_nll_static_scratch);
        if (e != cudaSuccess || p == nullptr) {
            set_error("pdt_nll_metrics_fwd_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
            return PDT_ECUDA;
        }
        dev_ptr[dev] = p;
    }
    unsigned char* base = reinterpret_cast<unsigned char*>(dev_ptr[dev]);
    // the array keeps whatever the previous call left: the ticket has to start from zero
    if (cudaMemsetAsync(base, 0, 16, (cudaStream_t)stream) != cudaSuccess) return check_launch("cudaMemsetAsync");
    nll_metrics_kernel<<<nll_grid(n), NLL_THREADS, 0, (cudaStream_t)stream>>>(
        logp, target, n, classes, topk, ignore_index, out, reinterpret_cast<unsigned*>(base),
        reinterpret_cast<nll_partial*>(base + 16), (float*)nullptr, (int32_t*)nullptr, 0);
    return check_launch("nll_metrics_fwd_static");
}
