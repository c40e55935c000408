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
// Waits are bounded (about a second): on time-out the error word is set and the kernel returns instead of
// hanging the GPU.  Larger messages stay on NCCL (bandwidth-bound there; see the sweep).
#include "common.cuh"

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
__device__ __forceinline__ float4 ld_peer_f4(const float4* p) {
    float4 v;
    asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(p));
    return v;
}

__global__ void __launch_bounds__(P2P_THREADS)
p2p_allreduce_kernel(p2p_peers peers, int rank, int world, float* __restrict__ buf, int64_t count4, int64_t cap4) {
    p2p_ctrl* ctrl = reinterpret_cast<p2p_ctrl*>(peers.base[rank]);
    __shared__ unsigned s_seq;
    __shared__ bool s_last;
    if (threadIdx.x == 0) s_seq = *reinterpret_cast<volatile unsigned*>(&ctrl->seq);
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
    if (threadIdx.x < world) {
        const long long t0 = clock64();
        while (ld_acquire_sys(&ctrl->flags[threadIdx.x]) < v) {
            if (clock64() - t0 > 4000000000LL) {   // ~2 s at 1.9 GHz: give up instead of hanging the GPU
                ctrl->error = 1;
                break;
            }
        }
    }
    __syncthreads();
    float4* dst = reinterpret_cast<float4*>(buf);
    for (int64_t i = blockIdx.x * (int64_t)P2P_THREADS + threadIdx.x; i < count4; i += (int64_t)gridDim.x * P2P_THREADS) {
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

}  // namespace pdt

using namespace pdt;

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
    p2p_allreduce_kernel<<<(unsigned)g, P2P_THREADS, 0, (cudaStream_t)stream>>>(peers, rank, world, buf, count4, cap4);
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
