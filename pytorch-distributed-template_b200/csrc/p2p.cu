// One-shot all-reduce over NVLink peer memory for the small flat gradient buffer.
//
// The MNIST step's collective is 87,376 bytes: pure latency.  NCCL's floor on 8 GPUs is ~35 us
// (profiles/r01_allreduce_sweep_w8.jsonl); this kernel does the same sum in one launch with plain
// loads/stores on peer-mapped memory through NVSwitch:
//   1. every rank copies its contribution into its own exchange buffer (double buffered by call parity),
//   2. the last block to finish publishes "call v is ready" into every peer's flag array (release, system scope),
//   3. every block waits until all W flags in its OWN memory reached v (acquire), then sums the W peer
//      buffers in rank order (bit-identical result on every rank) and writes the sum in place.
// A buffer is rewritten two calls later, after its owner has seen every peer's flag of the call in
// between, which a peer only sets after it finished reading -- so one flag phase per call suffices.
// Waits are bounded (pdt_p2p_set_timeout_ms, default 30 s; PDT_P2P_TIMEOUT_MS): on time-out the rank sets the
// error word in its own control block AND in every peer's, and the call becomes a no-op for the data -- nothing is
// summed, no gradient is written, no weight is updated (a sum over stale buffers would silently corrupt the
// replicas).  The error is sticky: every later call on any rank is a no-op too, until the host reads it
// (pdt_p2p_error; the Trainer does at every loss flush) and raises.  Larger messages stay on NCCL.
#include "common.cuh"
#include <stdlib.h>
#include <string.h>

namespace pdt {

constexpr int P2P_THREADS = 128;
constexpr int P2P_MAX_WORLD = 8;

struct p2p_ctrl {            // lives at the start of every rank's exchange allocation
    unsigned flags[16];      // flags[q] = last call published by rank q   (written by peers)
    unsigned seq;            // calls completed locally
    unsigned ticket_in;      // block tickets of the copy-in phase
    unsigned ticket_out;     // block tickets of the final phase
    unsigned error;          // 1 = a wait timed out
    unsigned pad[12];
};
static_assert(sizeof(p2p_ctrl) == 128, "control block is one 128-byte line");

struct p2p_peers {
    char* base[P2P_MAX_WORLD];   // peer-mapped base address of every rank's allocation (own included)
};

__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// Wait until flag >= v.  Returns false when the wait was abandoned: the time-out passed (this rank then raises the
// error word everywhere) or some rank had raised it already.
__device__ __forceinline__ bool wait_flag(const unsigned* flag, unsigned v, p2p_ctrl* ctrl, const p2p_peers& peers, int world,
                                          unsigned long long timeout_ns) {
    if (ld_acquire_sys(flag) >= v) return true;
    const unsigned long long t0 = global_ns();
    for (unsigned spin = 1;; ++spin) {
        if (ld_acquire_sys(flag) >= v) return true;
        if ((spin & 63u) == 0) {
            if (*reinterpret_cast<volatile unsigned*>(&ctrl->error) != 0) return false;
            if (global_ns() - t0 > timeout_ns) {
                for (int q = 0; q < world; ++q)
                    st_release_sys(&reinterpret_cast<p2p_ctrl*>(peers.base[q])->error, 1u);
                return false;
            }
        }
    }
}
__device__ __forceinline__ float4 ld_peer_f4(const float4* p) {
    float4 v;
    asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(p));
    return v;
}

__global__ void __launch_bounds__(P2P_THREADS)
p2p_allreduce_kernel(p2p_peers peers, int rank, int world, float* __restrict__ buf, int64_t count4, int64_t cap4,
                     unsigned long long timeout_ns) {
    p2p_ctrl* ctrl = reinterpret_cast<p2p_ctrl*>(peers.base[rank]);
    __shared__ unsigned s_seq;
    __shared__ bool s_last;
    __shared__ int s_bad;
    if (threadIdx.x == 0) {
        s_seq = *reinterpret_cast<volatile unsigned*>(&ctrl->seq);
        s_bad = *reinterpret_cast<volatile unsigned*>(&ctrl->error) != 0;      // sticky: raised by an earlier call
    }
    __syncthreads();
    const unsigned seq = s_seq, v = seq + 1;
    const int64_t data_off = 128 + (int64_t)(seq & 1) * cap4 * 16;   // bytes: control line, then two buffers

    // 1. copy-in
    float4* mine = reinterpret_cast<float4*>(peers.base[rank] + data_off);
    const float4* src = reinterpret_cast<const float4*>(buf);
    for (int64_t i = blockIdx.x * (int64_t)P2P_THREADS + threadIdx.x; i < count4; i += (int64_t)gridDim.x * P2P_THREADS)
        mine[i] = src[i];
    __syncthreads();
    // 2. the last block publishes the call number into every rank's flag array.  One gpu-scope fence by one
    //    thread per block orders the block's stores before its ticket; the release store below is system scope
    //    and cumulative over everything the ticket chain made visible.
    if (threadIdx.x == 0) {
        __threadfence();
        s_last = atomicAdd(&ctrl->ticket_in, 1u) == gridDim.x - 1;
        if (s_last) {
            __threadfence();
            ctrl->ticket_in = 0;
        }
    }
    __syncthreads();
    if (s_last && threadIdx.x < world) {
        p2p_ctrl* peer = reinterpret_cast<p2p_ctrl*>(peers.base[threadIdx.x]);
        st_release_sys(&peer->flags[rank], v);
    }
    // 3. wait for every rank's flag (bounded), then reduce in rank order
    if (threadIdx.x < world && !s_bad) {
        if (!wait_flag(&ctrl->flags[threadIdx.x], v, ctrl, peers, world, timeout_ns)) atomicExch(&s_bad, 1);
    }
    __syncthreads();
    const bool bad = s_bad != 0;       // a peer never arrived: leave buf as it is (see the header comment)
    float4* dst = reinterpret_cast<float4*>(buf);
    for (int64_t i = blockIdx.x * (int64_t)P2P_THREADS + threadIdx.x; i < count4 && !bad;
         i += (int64_t)gridDim.x * P2P_THREADS) {
        // all W peer loads are issued before the first add: one NVLink round trip, not W of them
        float4 t[P2P_MAX_WORLD];
#pragma unroll
        for (int q = 0; q < P2P_MAX_WORLD; ++q)
            if (q < world) t[q] = ld_peer_f4(reinterpret_cast<const float4*>(peers.base[q] + data_off) + i);
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int q = 0; q < P2P_MAX_WORLD; ++q)
            if (q < world) { acc.x += t[q].x; acc.y += t[q].y; acc.z += t[q].z; acc.w += t[q].w; }
        dst[i] = acc;
    }
    // 4. the last block to finish advances the call counter
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(&ctrl->ticket_out, 1u) == gridDim.x - 1) {
            ctrl->ticket_out = 0;
            *reinterpret_cast<volatile unsigned*>(&ctrl->seq) = v;
        }
    }
}

// ------------------------------------------------------------------------------------------
// The data-parallel step's tail as ONE kernel: per-tile gradient partials -> flat gradient (x 1/W, x the
// loss gradient) -> sum over the W ranks through peer memory -> Adam.  It is pdt_grad_reduce +
// pdt_p2p_allreduce + pdt_adam_step_flat with the flat gradient going straight from the reducing thread into
// the exchange buffer and from the summing thread into the update: two kernel boundaries and two passes
// over the gradient less.  Same protocol as p2p_allreduce_kernel (same control block, so the two can be
// mixed call by call).
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(P2P_THREADS)
p2p_reduce_adam_kernel(p2p_peers peers, int rank, int world, const float* __restrict__ scratch, int64_t T, float beta,
                       float scale, const pdt_stats* __restrict__ stats, const float* __restrict__ gscale,
                       float* __restrict__ grad, float* __restrict__ p, float* __restrict__ m, float* __restrict__ v,
                       float* __restrict__ vmax, pdt_adam_hyper* __restrict__ hyper, int64_t cap4,
                       unsigned long long timeout_ns) {
    constexpr int64_t COUNT4 = (PDT_NUM_PARAMS + PDT_TAIL) / 4;      // 5,461 float4
    static_assert((PDT_NUM_PARAMS + PDT_TAIL) % 4 == 0 && PDT_NUM_PARAMS % 4 == 0, "float4 granularity");
    p2p_ctrl* ctrl = reinterpret_cast<p2p_ctrl*>(peers.base[rank]);
    __shared__ unsigned s_seq;
    __shared__ bool s_last;
    __shared__ float s_neg_step, s_bc2_sqrt;
    __shared__ int s_bad;
    const int step = hyper->step;
    if (threadIdx.x == 0) {
        s_seq = *reinterpret_cast<volatile unsigned*>(&ctrl->seq);
        s_bad = *reinterpret_cast<volatile unsigned*>(&ctrl->error) != 0;
        const double t = (double)(step + 1);
        s_neg_step = (float)(-(hyper->lr / (1.0 - pow(hyper->beta1, t))));
        s_bc2_sqrt = (float)sqrt(1.0 - pow(hyper->beta2, t));
    }
    __syncthreads();
    const unsigned seq = s_seq, vcall = seq + 1;
    const int64_t data_off = 128 + (int64_t)(seq & 1) * cap4 * 16;
    // block = 16 float4 of the flat vector x 8 slot lanes: slot lane j sums slots j, j+8, ... (independent loads
    // in flight), the 8 lane sums are folded in a fixed order, and slot lane 0 carries the float4 from there on
    const int q4 = threadIdx.x & 15, sl = threadIdx.x >> 4;
    const int64_t i4 = blockIdx.x * (int64_t)16 + q4;
    const bool owner = sl == 0 && i4 < COUNT4;
    const float mult = gscale != nullptr ? scale * *gscale : scale;
    __shared__ float4 red[8][16];

    // 1. this rank's flat gradient (fixed-order sum of the per-tile partials) into its exchange buffer
    float4* mine = reinterpret_cast<float4*>(peers.base[rank] + data_off);
    {
        const int i = (int)i4 * 4;
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
        if (i4 < COUNT4 && i < PDT_NUM_PARAMS) {
            const float* base;
            int64_t slots, stride;
            if (i < 260) {
                base = scratch + part1_off(T) + i; slots = T * 6; stride = PART1;
            } else if (i < 5280) {
                base = scratch + part2_off(T) + (i - 260); slots = T; stride = PART2;
            } else {
                base = scratch + part3_off(T) + (i - 5280); slots = T; stride = PART3;
            }
            int64_t k = sl;
#pragma unroll 2
            for (; k + 8 < slots; k += 16) {
                const float4 t0 = *reinterpret_cast<const float4*>(base + k * stride);
                const float4 t1 = *reinterpret_cast<const float4*>(base + (k + 8) * stride);
                a.x += t0.x; a.y += t0.y; a.z += t0.z; a.w += t0.w;
                b.x += t1.x; b.y += t1.y; b.z += t1.z; b.w += t1.w;
            }
            if (k < slots) {
                const float4 t0 = *reinterpret_cast<const float4*>(base + k * stride);
                a.x += t0.x; a.y += t0.y; a.z += t0.z; a.w += t0.w;
            }
        }
        red[sl][q4] = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
    }
    __syncthreads();
    if (owner) {
        const int i = (int)i4 * 4;
        float4 g4;
        if (i < PDT_NUM_PARAMS) {
            float4 r[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) r[j] = red[j][q4];
            g4.x = (((r[0].x + r[1].x) + (r[2].x + r[3].x)) + ((r[4].x + r[5].x) + (r[6].x + r[7].x))) * mult;
            g4.y = (((r[0].y + r[1].y) + (r[2].y + r[3].y)) + ((r[4].y + r[5].y) + (r[6].y + r[7].y))) * mult;
            g4.z = (((r[0].z + r[1].z) + (r[2].z + r[3].z)) + ((r[4].z + r[5].z) + (r[6].z + r[7].z))) * mult;
            g4.w = (((r[0].w + r[1].w) + (r[2].w + r[3].w)) + ((r[4].w + r[5].w) + (r[6].w + r[7].w))) * mult;
            if (beta != 0.f) {
                const float4 old = *reinterpret_cast<const float4*>(grad + i);
                g4.x = fmaf(beta, old.x, g4.x); g4.y = fmaf(beta, old.y, g4.y);
                g4.z = fmaf(beta, old.z, g4.z); g4.w = fmaf(beta, old.w, g4.w);
            }
        } else {      // the tail: mean loss (x 1/W) and the metric counts ride in the same exchange
            g4 = make_float4(0.f, 0.f, 0.f, 0.f);
            if (stats != nullptr)
                g4 = make_float4(stats->loss_mean * scale, (float)stats->top1, (float)stats->topk, (float)stats->n_valid);
        }
        mine[i4] = g4;
    }
    __syncthreads();
    // 2. the last block publishes the call number into every rank's flag array
    if (threadIdx.x == 0) {
        __threadfence();
        s_last = atomicAdd(&ctrl->ticket_in, 1u) == gridDim.x - 1;
        if (s_last) {
            __threadfence();
            ctrl->ticket_in = 0;
        }
    }
    __syncthreads();
    if (s_last && threadIdx.x < world) {
        p2p_ctrl* peer = reinterpret_cast<p2p_ctrl*>(peers.base[threadIdx.x]);
        st_release_sys(&peer->flags[rank], vcall);
    }
    // 3. wait for every rank's flag (bounded), then sum in rank order and update
    if (threadIdx.x < world && !s_bad) {
        if (!wait_flag(&ctrl->flags[threadIdx.x], vcall, ctrl, peers, world, timeout_ns)) atomicExch(&s_bad, 1);
    }
    __syncthreads();
    // a peer never arrived (or an earlier call failed): no gradient is written, no weight moves, the step counter
    // stays -- the host finds the error word at its next check and raises
    const bool bad = s_bad != 0;
    if (owner && !bad) {
        float4 t[P2P_MAX_WORLD];
#pragma unroll
        for (int q = 0; q < P2P_MAX_WORLD; ++q)
            if (q < world) t[q] = ld_peer_f4(reinterpret_cast<const float4*>(peers.base[q] + data_off) + i4);
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int q = 0; q < P2P_MAX_WORLD; ++q)
            if (q < world) { acc.x += t[q].x; acc.y += t[q].y; acc.z += t[q].z; acc.w += t[q].w; }
        const int i = (int)i4 * 4;
        *reinterpret_cast<float4*>(grad + i) = acc;        // .grad (and the tail) stay observable
        if (i < PDT_NUM_PARAMS) {
            const float w1 = (float)(1.0 - hyper->beta1), b2 = (float)hyper->beta2, w2 = (float)(1.0 - hyper->beta2);
            const float eps = (float)hyper->eps, wd = (float)hyper->weight_decay;
            const bool amsgrad = hyper->amsgrad != 0 && vmax != nullptr;
            const float gq[4] = {acc.x, acc.y, acc.z, acc.w};
            float4 p4 = *reinterpret_cast<const float4*>(p + i), m4 = *reinterpret_cast<const float4*>(m + i);
            float4 v4 = *reinterpret_cast<const float4*>(v + i);
            float4 x4 = amsgrad ? *reinterpret_cast<const float4*>(vmax + i) : make_float4(0.f, 0.f, 0.f, 0.f);
            float pq[4] = {p4.x, p4.y, p4.z, p4.w}, mq[4] = {m4.x, m4.y, m4.z, m4.w};
            float vq[4] = {v4.x, v4.y, v4.z, v4.w}, xq[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                float gv = gq[e];
                if (wd != 0.f) gv = fmaf(wd, pq[e], gv);
                mq[e] = fmaf(w1, gv - mq[e], mq[e]);
                vq[e] = fmaf(w2 * gv, gv, vq[e] * b2);
                float d = vq[e];
                if (amsgrad) {
                    d = fmaxf(xq[e], vq[e]);
                    xq[e] = d;
                }
                pq[e] = fmaf(s_neg_step, mq[e] / (sqrtf(d) / s_bc2_sqrt + eps), pq[e]);
            }
            *reinterpret_cast<float4*>(p + i) = make_float4(pq[0], pq[1], pq[2], pq[3]);
            *reinterpret_cast<float4*>(m + i) = make_float4(mq[0], mq[1], mq[2], mq[3]);
            *reinterpret_cast<float4*>(v + i) = make_float4(vq[0], vq[1], vq[2], vq[3]);
            if (amsgrad) *reinterpret_cast<float4*>(vmax + i) = make_float4(xq[0], xq[1], xq[2], xq[3]);
        }
    }
    // 4. the last block to finish advances the call counter; then the optimizer's step counter
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(&ctrl->ticket_out, 1u) == gridDim.x - 1) {
            ctrl->ticket_out = 0;
            *reinterpret_cast<volatile unsigned*>(&ctrl->seq) = vcall;
        }
    }
    advance_step(hyper, bad ? step - 1 : step);
}

}  // namespace pdt

using namespace pdt;

static unsigned long long g_timeout_ns = 0;
static unsigned long long timeout_ns() {
    if (g_timeout_ns == 0) {
        const char* e = getenv("PDT_P2P_TIMEOUT_MS");
        const long long ms = e ? atoll(e) : 30000;
        g_timeout_ns = (unsigned long long)(ms > 0 ? ms : 30000) * 1000000ull;
    }
    return g_timeout_ns;
}

extern "C" int pdt_p2p_set_timeout_ms(int64_t ms) {
    PDT_REQUIRE(ms > 0, "timeout must be positive");
    g_timeout_ns = (unsigned long long)ms * 1000000ull;
    return PDT_OK;
}

extern "C" int64_t pdt_p2p_bytes(int64_t capacity_floats) {
    const int64_t cap4 = (capacity_floats + 3) / 4;
    return 128 + 2 * cap4 * 16;
}

extern "C" int pdt_p2p_alloc(int64_t capacity_floats, void** dev_ptr, void* ipc_handle64) {
    PDT_REQUIRE(capacity_floats > 0 && dev_ptr && ipc_handle64, "bad arguments");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
    void* p = nullptr;
    const int64_t bytes = pdt_p2p_bytes(capacity_floats);
    if (cudaMalloc(&p, (size_t)bytes) != cudaSuccess) return check_launch("cudaMalloc(p2p)");
    if (cudaMemset(p, 0, (size_t)bytes) != cudaSuccess) return check_launch("cudaMemset(p2p)");
    cudaIpcMemHandle_t h;
    if (cudaIpcGetMemHandle(&h, p) != cudaSuccess) {
        cudaFree(p);
        return check_launch("cudaIpcGetMemHandle");
    }
    memcpy(ipc_handle64, &h, 64);
    *dev_ptr = p;
    return PDT_OK;
}

extern "C" int pdt_p2p_open(const void* ipc_handle64, void** peer_ptr) {
    PDT_REQUIRE(ipc_handle64 && peer_ptr, "bad arguments");
    cudaIpcMemHandle_t h;
    memcpy(&h, ipc_handle64, 64);
    if (cudaIpcOpenMemHandle(peer_ptr, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess)
        return check_launch("cudaIpcOpenMemHandle");
    return PDT_OK;
}

extern "C" int pdt_p2p_close(void* peer_ptr) {
    PDT_REQUIRE(peer_ptr, "null pointer");
    return cudaIpcCloseMemHandle(peer_ptr) == cudaSuccess ? PDT_OK : check_launch("cudaIpcCloseMemHandle");
}

extern "C" int pdt_p2p_free(void* dev_ptr) {
    PDT_REQUIRE(dev_ptr, "null pointer");
    return cudaFree(dev_ptr) == cudaSuccess ? PDT_OK : check_launch("cudaFree(p2p)");
}

extern "C" int pdt_p2p_allreduce(void* const* peer_bases, int rank, int world, float* buf, int64_t count,
                                 int64_t capacity_floats, void* stream) {
    PDT_REQUIRE(peer_bases && buf && count > 0, "bad arguments");
    PDT_REQUIRE(world >= 1 && world <= P2P_MAX_WORLD && rank >= 0 && rank < world, "world must be 1..8");
    PDT_REQUIRE(count <= capacity_floats, "count exceeds the exchange buffer capacity");
    PDT_REQUIRE((reinterpret_cast<uintptr_t>(buf) & 15) == 0, "buffer must be 16-byte aligned");
    p2p_peers peers = {};
    for (int q = 0; q < world; ++q) peers.base[q] = reinterpret_cast<char*>(peer_bases[q]);
    const int64_t count4 = (count + 3) / 4, cap4 = (capacity_floats + 3) / 4;
    int64_t g = (count4 + P2P_THREADS - 1) / P2P_THREADS;
    if (g > 64) g = 64;
    p2p_allreduce_kernel<<<(unsigned)g, P2P_THREADS, 0, (cudaStream_t)stream>>>(peers, rank, world, buf, count4, cap4,
                                                                                timeout_ns());
    return check_launch("p2p_allreduce");
}

extern "C" int pdt_p2p_error(const void* own_base, int* error_out) {
    PDT_REQUIRE(own_base && error_out, "bad arguments");
    unsigned e = 0;
    if (cudaMemcpy(&e, reinterpret_cast<const char*>(own_base) + offsetof(p2p_ctrl, error), 4, cudaMemcpyDeviceToHost) !=
        cudaSuccess)
        return check_launch("cudaMemcpy(p2p error)");
    *error_out = (int)e;
    return PDT_OK;
}

extern "C" int pdt_p2p_error_async(const void* own_base, int* pinned_out, void* stream) {
    PDT_REQUIRE(own_base && pinned_out, "bad arguments");
    if (cudaMemcpyAsync(pinned_out, reinterpret_cast<const char*>(own_base) + offsetof(p2p_ctrl, error), 4,
                        cudaMemcpyDeviceToHost, (cudaStream_t)stream) != cudaSuccess)
        return check_launch("cudaMemcpyAsync(p2p error)");
    return PDT_OK;
}

extern "C" int pdt_p2p_reduce_adam(void* const* peer_bases, int rank, int world, const void* scratch, int64_t n,
                                   float beta, float scale, const pdt_stats* stats, const float* grad_scale_dev,
                                   float* flat_grad, float* param, float* exp_avg, float* exp_avg_sq,
                                   float* max_exp_avg_sq, pdt_adam_hyper* hyper, int64_t capacity_floats,
                                   void* stream) {
    PDT_REQUIRE(peer_bases && scratch && flat_grad && param && exp_avg && exp_avg_sq && hyper && n > 0, "bad arguments");
    PDT_REQUIRE(world >= 1 && world <= P2P_MAX_WORLD && rank >= 0 && rank < world, "world must be 1..8");
    PDT_REQUIRE(capacity_floats >= PDT_NUM_PARAMS + PDT_TAIL, "exchange buffer smaller than the flat gradient + tail");
    PDT_REQUIRE(((reinterpret_cast<uintptr_t>(flat_grad) | reinterpret_cast<uintptr_t>(param) |
                  reinterpret_cast<uintptr_t>(exp_avg) | reinterpret_cast<uintptr_t>(exp_avg_sq) |
                  reinterpret_cast<uintptr_t>(max_exp_avg_sq)) & 15) == 0, "buffers must be 16-byte aligned");
    p2p_peers peers = {};
    for (int q = 0; q < world; ++q) peers.base[q] = reinterpret_cast<char*>(peer_bases[q]);
    const int64_t count4 = (PDT_NUM_PARAMS + PDT_TAIL) / 4, cap4 = (capacity_floats + 3) / 4;
    const unsigned g = (unsigned)((count4 + 15) / 16);     // 342 blocks of 128 threads: all resident at once
    p2p_reduce_adam_kernel<<<g, P2P_THREADS, 0, (cudaStream_t)stream>>>(
        peers, rank, world, reinterpret_cast<const float*>(scratch), num_tiles(n), beta, scale, stats, grad_scale_dev,
        flat_grad, param, exp_avg, exp_avg_sq, max_exp_avg_sq, hyper, cap4, timeout_ns());
    return check_launch("p2p_reduce_adam");
}
