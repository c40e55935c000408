// Shared device/host helpers for the pdt kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "pdt.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "pdt kernels are written for sm_100a (B200) only"
#endif

namespace pdt {

constexpr int TILE = PDT_TILE;
constexpr int P1F = PDT_P1_FEATS;
constexpr int P2F = PDT_P2_FEATS;
constexpr int HF = PDT_H_FEATS;
constexpr int NCLS = 10;

// per-tile partial-gradient scratch, in floats
constexpr int PART1 = 260;    // conv1.weight 250 + conv1.bias 10; 6 slots per tile (one per row band)
constexpr int PART2 = 5020;   // conv2.weight 5000 + conv2.bias 20; 1 slot per tile
constexpr int PART3 = 16560;  // fc1.weight 16000, fc1.bias 50, fc2.weight 500, fc2.bias 10
constexpr int PART_PER_TILE = 6 * PART1 + PART2 + PART3;  // 23,140

__host__ __device__ inline int64_t num_tiles(int64_t n) { return (n + TILE - 1) / TILE; }
__host__ __device__ inline int64_t part1_off(int64_t) { return 0; }
__host__ __device__ inline int64_t part2_off(int64_t T) { return T * 6 * PART1; }
__host__ __device__ inline int64_t part3_off(int64_t T) { return T * 6 * PART1 + T * PART2; }

// "lane4" layout of p1: four consecutive features of one sample are one float4, [tile][feature/4][32 lanes][4].
// A lane's run of columns is then ONE 128-bit load (a shared-memory load costs ~4 dispatch cycles whatever its
// width), while a warp-wide access is still contiguous (512 B) and rows of 4k columns keep their byte offsets.
__host__ __device__ inline int lane4(int f, int lane) { return (((f >> 2) * TILE + lane) << 2) + (f & 3); }

// one block's / one tile's share of the loss + metric record (loss.cu; the fused head kernels in fwd.cu)
struct nll_partial {
    float loss;
    int top1, topk, valid;
};

void set_error(const char* fmt, ...);
int check_launch(const char* what);

// ---- mbarrier + 1-D bulk TMA (cp.async.bulk -> UBLKCP) ---------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// global -> shared bulk copy; dst/src 16-byte aligned, bytes a multiple of 16
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// ---- programmatic dependent launch (PDL) ------------------------------------------------
// Every kernel of the step is launched with programmaticStreamSerialization allowed and follows
// one discipline: [prologue] pdl_wait(); pdl_trigger(); [main].  The prologue may only read data
// produced at least two kernels earlier in the stream (weights, saved activations, inputs) and
// may only write shared memory; because each kernel triggers AFTER its own wait, a dependent's
// prologue overlaps nothing older than its immediate predecessor.  Effect: the launch latency and
// the operand staging of kernel N+1 hide under the tail of kernel N.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

bool pdl_enabled(int which);
int carveout_percent();   // PDT_CARVEOUT: preferred shared-memory carve-out for every kernel (-1 = leave default)

// All kernels of the step ask for the SAME shared-memory carve-out: consecutive kernels with different
// carve-outs make every SM drain and repartition its L1/shared array between them.
template <typename K>
inline void prefer_carveout(K kernel) {
    const int pct = carveout_percent();
    if (pct < 0) return;
    int dev = 0;
    cudaGetDevice(&dev);
    static bool done[64] = {false};
    if (dev >= 0 && dev < 64 && !done[dev]) {
        cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
        done[dev] = true;
    }
}

// Raise a kernel's dynamic shared-memory limit once per (kernel, device).  SLOT names the kernel: two
// instantiations of one kernel template have the same pointer TYPE, so the type alone cannot key the cache.
template <int SLOT, typename K>
inline int ensure_smem(K kernel, int bytes) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return check_launch("cudaGetDevice");
    static bool done[64] = {false};
    if (dev < 64 && done[dev]) return PDT_OK;
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess)
        return check_launch("cudaFuncSetAttribute");
    if (dev < 64) done[dev] = true;
    return PDT_OK;
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(int which, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, void* stream,
                              Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled(which) ? 1 : 0;
    prefer_carveout(kernel);
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int warp_sum_i(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// the last block to pass here advances hyper->step (atomicInc wraps the ticket back to 0)
template <typename H>
__device__ __forceinline__ void advance_step(H* hyper, int step) {
    __shared__ bool last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) last = atomicInc(&hyper->ticket, gridDim.x - 1) == gridDim.x - 1;
    __syncthreads();
    if (last && threadIdx.x == 0) hyper->step = step + 1;
}

// 2x2 max-pool in torch's scan order ((0,0),(0,1),(1,0),(1,1)); first maximum wins.
__device__ __forceinline__ float pool4(float a, float b, float c, float d, int& k) {
    float m = a;
    k = 0;
    if (b > m) { m = b; k = 1; }
    if (c > m) { m = c; k = 2; }
    if (d > m) { m = d; k = 3; }
    return m;
}

// ---- the input batch: fp32 (already normalised by the data loader) or raw bytes --------------------------------
// The reference's MnistDataLoader turns each MNIST byte v into (v / 255 - 0.1307) / 0.3081 on the CPU
// (data_loader/data_loaders.py:13-16: ToTensor, then Normalize, both fp32) and ships 4 bytes per pixel to the GPU
// (trainer/trainer.py:46).  The *_u8 entry points take the bytes themselves -- a quarter of the host-to-device
// traffic, which is what bounds the 8-GPU end-to-end loop -- and apply the same two fp32 operations when a pixel is
// loaded, through a 256-entry table in shared memory: the kernels see bit-identical values.
struct x_norm {
    float mean, std;
};
__device__ __forceinline__ float normalize_byte(int v, x_norm nm) {
    return __fdiv_rn(__fsub_rn(__fdiv_rn((float)v, 255.f), nm.mean), nm.std);
}
// fills lut[256]; the caller synchronises the block before the first load_x4<true>
__device__ __forceinline__ void fill_xlut(float* lut, x_norm nm) {
    for (int i = threadIdx.x; i < 256; i += blockDim.x) lut[i] = normalize_byte(i, nm);
}
// four consecutive pixels starting at element `off` (a multiple of 4) of the batch
template <bool U8>
__device__ __forceinline__ float4 load_x4(const void* __restrict__ x, int64_t off, const float* lut) {
    if (U8) {
        const uchar4 v = __ldg(reinterpret_cast<const uchar4*>(static_cast<const uint8_t*>(x) + off));
        return make_float4(lut[v.x], lut[v.y], lut[v.z], lut[v.w]);
    }
    return __ldg(reinterpret_cast<const float4*>(static_cast<const float*>(x) + off));
}

// Load an 8-row x (NC4*4)-column window of 32 consecutive samples of x [n,784] into shared
// memory laid out [row*RS + col][33], transposing on the way so that lane == sample reads are
// bank-conflict free.  RS (row stride in positions) must be 8 mod 32 or 24 mod 32 so the
// transposing stores are conflict free too.  Executed by one warp.  Samples >= n read as zero.
template <int NC4, int RS, bool U8>
__device__ __forceinline__ void load_x_window(float* __restrict__ dst, const void* __restrict__ x, const float* lut,
                                              int64_t n, int64_t sample0, int row0, int col0, int lane) {
    static_assert(RS >= NC4 * 4 && (RS % 32 == 8 || RS % 32 == 24), "row stride");
    const int s8 = lane >> 2, rq = lane & 3;
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        const int s = g * 8 + s8;
        const int64_t sample = sample0 + s;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int row = rq + 4 * h;
            const int64_t src = sample * 784 + (row0 + row) * 28 + col0;
#pragma unroll
            for (int c4 = 0; c4 < NC4; ++c4) {
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (sample < n) v = load_x4<U8>(x, src + 4 * c4, lut);
                float* d = dst + (row * RS + c4 * 4) * 33 + s;
                d[0] = v.x;
                d[33] = v.y;
                d[66] = v.z;
                d[99] = v.w;
            }
        }
    }
}

}  // namespace pdt

#define PDT_REQUIRE(cond, msg)                \
    do {                                      \
        if (!(cond)) {                        \
            pdt::set_error("%s: %s", __func__, msg); \
            return PDT_EINVAL;                \
        }                                     \
    } while (0)
