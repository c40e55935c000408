// Forward kernels of the MNIST CNN (reference model/model.py:15-22), FP32 on the CUDA cores.
//
// Mapping used by every kernel here: one LANE = one SAMPLE.  A warp works on a tile of 32
// samples with identical control flow, weights are warp-uniform (broadcast shared-memory
// loads) and activations live in the lane layout [tile][feature][32], so every activation
// load/store of a warp is one fully coalesced 128-byte line and every shared-memory access is
// conflict free.  The arithmetic is bound by the FP32 FMA pipe: the inner
// loops keep input windows and accumulators in registers and issue >= 10 FFMA per shared load.
#include "common.cuh"
#include "conv2_tc.cuh"
#include <stdlib.h>

namespace pdt {

// ------------------------------------------------------------------------------------------
// conv1 (1->10, 5x5) + max_pool2d(2) + relu           model/model.py:16
// warp task = (tile, 4x4 pre-pool block, channel group): all 10 channels (4,000 FFMA per lane) or, while the
// grid is a single wave, 5 channels per warp and two warps per block (conv1_channel_split).
// ------------------------------------------------------------------------------------------
constexpr int C1_WPC = 4;

template <bool U8>
__device__ __forceinline__ void
conv1_fwd_body(const void* __restrict__ x, x_norm nm, const float* __restrict__ w, const float* __restrict__ b,
               float* __restrict__ p1, uint8_t* __restrict__ idx1, int64_t n, int64_t n_tasks, int csplit) {
    __shared__ __align__(16) float ws[10][28];
    __shared__ float patch[C1_WPC][64 * 33];
    __shared__ float xlut[U8 ? 256 : 1];
    if (U8) {
        fill_xlut(xlut, nm);
        __syncthreads();
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // n_tasks counts (tile, block, channel group) triples; csplit = 1 or 2 channel groups per block.  Two warps
    // per block (5 channels each) double the warps per scheduler while the batch is a single wave of CTAs.
    const int64_t task2 = (int64_t)blockIdx.x * C1_WPC + warp;
    const bool active = task2 < n_tasks;
    const int64_t task = task2 / csplit;
    const int c_lo = (int)(task2 % csplit) * (10 / csplit), c_hi = c_lo + 10 / csplit;
    const int64_t tile = task / 36;
    const int blk = (int)(task % 36), by = blk / 6, bx = blk % 6;

    // Prologue (before the dependency wait): only the input batch is touched -- it is never written by
    // a kernel of the step.  The weights are NOT read here: this kernel can directly follow the optimizer.
    float in[64];
    if (active) {
        float* mine = patch[warp];
        load_x_window<2, 8, U8>(mine, x, xlut, n, tile * TILE, 4 * by, 4 * bx, lane);
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 64; ++i) in[i] = mine[i * 33 + lane];
    }
    pdl_wait();
    for (int i = threadIdx.x; i < 280; i += blockDim.x) {
        const int c = i / 28, k = i % 28;
        ws[c][k] = k < 25 ? w[c * 25 + k] : (k == 25 ? b[c] : 0.f);
    }
    __syncthreads();
    if (!active) {
        pdl_trigger();
        return;
    }

    float* p1t = p1 + tile * (int64_t)(P1F * TILE) + lane * 4;
    uint8_t* i1t = idx1 + tile * (int64_t)(PDT_IDX1_BYTES * TILE) + lane;
#pragma unroll 1
    for (int c = c_lo; c < c_hi; ++c) {
        if (c == c_hi - 2) pdl_trigger();   // late: the next kernel's launch + prologue hide under our tail only
        float wv[28];
#pragma unroll
        for (int q = 0; q < 7; ++q) {
            const float4 t = *reinterpret_cast<const float4*>(&ws[c][q * 4]);
            wv[q * 4 + 0] = t.x; wv[q * 4 + 1] = t.y; wv[q * 4 + 2] = t.z; wv[q * 4 + 3] = t.w;
        }
        float acc[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i] = wv[25];
#pragma unroll
        for (int ky = 0; ky < 5; ++ky)
#pragma unroll
            for (int kx = 0; kx < 5; ++kx)
#pragma unroll
                for (int oy = 0; oy < 4; ++oy)
#pragma unroll
                    for (int ox = 0; ox < 4; ++ox)
                        acc[oy * 4 + ox] = fmaf(wv[ky * 5 + kx], in[(oy + ky) * 8 + ox + kx], acc[oy * 4 + ox]);
        unsigned bits = 0;
#pragma unroll
        for (int py = 0; py < 2; ++py) {
            float o[2];
#pragma unroll
            for (int px = 0; px < 2; ++px) {
                int k;
                const float m = pool4(acc[(2 * py) * 4 + 2 * px], acc[(2 * py) * 4 + 2 * px + 1],
                                      acc[(2 * py + 1) * 4 + 2 * px], acc[(2 * py + 1) * 4 + 2 * px + 1], k);
                bits |= (unsigned)k << (2 * (py * 2 + px));
                o[px] = fmaxf(m, 0.f);
            }
            // the pooled pair (2bx, 2bx+1) is half of one float4 of the lane4 layout
            *reinterpret_cast<float2*>(p1t + lane4(c * 144 + (2 * by + py) * 12 + 2 * bx, 0)) = make_float2(o[0], o[1]);
        }
        i1t[(c * 36 + blk) * TILE] = (uint8_t)bits;
    }
}

// ------------------------------------------------------------------------------------------
// conv2 (10->20, 5x5) + Dropout2d + max_pool2d(2) + relu       model/model.py:17
// CTA = (tile, 4x4 pre-pool block sb): the 8x8x10 input window of 32 samples (80 KB) arrives
// by 1-D bulk TMA copies (the lane layout makes each window row one contiguous 1 KB run).
// 8 warps = 4 channel groups of 5 x 2 row halves; 10,000 FFMA per lane, 200 per 7 LDS.128.
// ------------------------------------------------------------------------------------------
constexpr int C2_THREADS = 256;
constexpr int C2_SMEM_PATCH = 10 * 64 * TILE;   // floats
constexpr int C2_SMEM_W = 10 * 4 * 5 * 28;      // floats
constexpr int C2_SMEM_BYTES = (C2_SMEM_PATCH + C2_SMEM_W) * 4 + 10 * 8;   // + one mbarrier per input channel

// weights -> ws[ci][cg][ky][kx*5 + j] (28-float rows), co = cg*5 + j; one mbarrier per input channel.
// Prologue work: the weights are older than the previous kernel, so this runs before the dependency wait.
template <int NT>
__device__ __forceinline__ void conv2_stage_weights(const float* __restrict__ w, float* ws, uint64_t* bar) {
    if (threadIdx.x < 10) {
        mbar_init(&bar[threadIdx.x], 1);
        mbar_fence_init();
    }
    // all global loads of a thread are issued before the first store (one L2 latency, not 20)
    constexpr int PER = (5000 + NT - 1) / NT;
    float tmp[PER];
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        const int e = threadIdx.x + i * NT;
        tmp[i] = e < 5000 ? __ldg(w + e) : 0.f;
    }
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        const int e = threadIdx.x + i * NT;
        if (e < 5000) {
            const int co = e / 250, rem = e % 250, ci = rem / 25, k = rem % 25, ky = k / 5, kx = k % 5;
            ws[((ci * 4 + co / 5) * 5 + ky) * 28 + kx * 5 + (co % 5)] = tmp[i];
        }
    }
}

// the 8x8x10 window of the tile: 80 bulk copies of 1 KB; channel ci completes on bar[ci], so the main loop
// starts on channel 0 while the other nine stream in.  Executed by warp 0 after the dependency wait.
__device__ __forceinline__ void conv2_issue_window(const float* __restrict__ p1, float* patch, uint64_t* bar,
                                                   int64_t tile, int by, int bx, int lane) {
    if (lane < 10) mbar_arrive_expect_tx(&bar[lane], 64 * TILE * 4);
    __syncwarp();
    const float* src0 = p1 + tile * (int64_t)(P1F * TILE);
    for (int i = lane; i < 80; i += 32) {
        const int ci = i >> 3, row = i & 7;
        bulk_g2s(patch + (ci * 64 + row * 8) * TILE, src0 + (ci * 144 + (4 * by + row) * 12 + 4 * bx) * TILE,
                 8 * TILE * 4, &bar[ci]);
    }
}

// 8 warps: channel group cg (5 channels) x row half (2 pre-pool rows x 4 columns = two pooling windows).
// p1 is in the lane4 layout, so a window row of 8 columns is TWO 128-bit shared loads per lane:
// 12 + 35 shared loads per 1,000 FFMA (it was 48 + 35 with one load per value).
__device__ __forceinline__ void conv2_main(const float* patch, const float* ws, uint64_t* bar, int cg, int half,
                                           int lane, float (&acc)[5][8]) {
#pragma unroll
    for (int j = 0; j < 5; ++j)
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[j][i] = 0.f;
#pragma unroll 1
    for (int ci = 0; ci < 10; ++ci) {
        mbar_wait(&bar[ci], 0);
        float in[6][8];
        const float4* pc = reinterpret_cast<const float4*>(patch + (ci * 64 + (2 * half) * 8) * TILE) + lane;
#pragma unroll
        for (int r = 0; r < 6; ++r)
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const float4 t = pc[(r * 2 + q) * TILE];
                in[r][q * 4 + 0] = t.x; in[r][q * 4 + 1] = t.y; in[r][q * 4 + 2] = t.z; in[r][q * 4 + 3] = t.w;
            }
#pragma unroll
        for (int ky = 0; ky < 5; ++ky) {
            float wv[28];
            const float4* wp = reinterpret_cast<const float4*>(ws + ((ci * 4 + cg) * 5 + ky) * 28);
#pragma unroll
            for (int q = 0; q < 7; ++q) {
                const float4 t = wp[q];
                wv[q * 4 + 0] = t.x; wv[q * 4 + 1] = t.y; wv[q * 4 + 2] = t.z; wv[q * 4 + 3] = t.w;
            }
#pragma unroll
            for (int kx = 0; kx < 5; ++kx)
#pragma unroll
                for (int j = 0; j < 5; ++j)
#pragma unroll
                    for (int oy = 0; oy < 2; ++oy)
#pragma unroll
                        for (int ox = 0; ox < 4; ++ox)
                            acc[j][oy * 4 + ox] = fmaf(wv[kx * 5 + j], in[oy + ky][ox + kx], acc[j][oy * 4 + ox]);
        }
    }
}

// bias, Dropout2d, 2x2 max-pool, relu of this warp's 5 channels x 2 windows; writes p2 / idx2 and returns the
// pooled values in v[j][px]
__device__ __forceinline__ void conv2_epilogue(const float (&acc)[5][8], const float* __restrict__ b,
                                               const uint8_t* __restrict__ keep2d, float scale,
                                               float* __restrict__ p2, uint8_t* __restrict__ idx2, int64_t n,
                                               int64_t tile, int by, int bx, int cg, int half, int lane,
                                               float (&v)[5][2]) {
    const int64_t sample = tile * TILE + lane;
    float* p2t = p2 + tile * (int64_t)(P2F * TILE) + lane;
    uint8_t* i2t = idx2 + tile * (int64_t)(PDT_IDX2_BYTES * TILE) + lane;
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        const int co = cg * 5 + j;
        const float bias = __ldg(b + co);
        float s = 1.f;
        if (keep2d != nullptr) s = (sample < n && keep2d[sample * 20 + co]) ? scale : 0.f;
#pragma unroll
        for (int px = 0; px < 2; ++px) {
            int k;
            const float m = pool4((acc[j][2 * px] + bias) * s, (acc[j][2 * px + 1] + bias) * s,
                                  (acc[j][4 + 2 * px] + bias) * s, (acc[j][4 + 2 * px + 1] + bias) * s, k);
            const int f = co * 16 + (2 * by + half) * 4 + 2 * bx + px;     // index in view(-1, 320) order
            v[j][px] = fmaxf(m, 0.f);
            p2t[f * TILE] = v[j][px];
            i2t[f * TILE] = (uint8_t)k;
        }
    }
}

__global__ void __launch_bounds__(C2_THREADS)
conv2_fwd_kernel(const float* __restrict__ p1, const float* __restrict__ w, const float* __restrict__ b,
                 const uint8_t* __restrict__ keep2d, float scale, float* __restrict__ p2,
                 uint8_t* __restrict__ idx2, int64_t n) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float* patch = reinterpret_cast<float*>(smem_raw);
    float* ws = patch + C2_SMEM_PATCH;
    uint64_t* bar = reinterpret_cast<uint64_t*>(ws + C2_SMEM_W);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t tile = blockIdx.x >> 2;
    const int sb = blockIdx.x & 3, by = sb >> 1, bx = sb & 1;

    conv2_stage_weights<C2_THREADS>(w, ws, bar);
    pdl_wait();
    __syncthreads();
    if (warp == 0) conv2_issue_window(p1, patch, bar, tile, by, bx, lane);
    __syncthreads();

    const int cg = warp >> 1, half = warp & 1;
    float acc[5][8];
    conv2_main(patch, ws, bar, cg, half, lane, acc);

    pdl_trigger();   // main loop done: let the next kernel's launch and prologue overlap our epilogue
    float v[5][2];
    conv2_epilogue(acc, b, keep2d, scale, p2, idx2, n, tile, by, bx, cg, half, lane, v);
}

// ------------------------------------------------------------------------------------------
// fc1 + relu + dropout + fc2 + log_softmax                      model/model.py:18-22
// One thread-block CLUSTER of 5 CTAs per tile (so 5x more SMs than tiles work on this small
// layer): CTA r computes fc1 outputs [10r, 10r+10) -- its 10 weight rows (12.8 KB) and the tile's
// p2 (40 KB) arrive by two bulk TMA copies -- and writes them into CTA 0's shared memory through
// distributed shared memory; after a cluster barrier CTA 0 does fc2 + log_softmax.
// ------------------------------------------------------------------------------------------
constexpr int HD_THREADS = 320;
constexpr int HD_CLUSTER = 5;
constexpr int HD_W1 = 0;                          // [10][320] this CTA's rows of fc1.weight
constexpr int HD_ACT = HD_W1 + 3200;              // [320][32] p2 tile
constexpr int HD_RED = HD_ACT + P2F * TILE;       // [10 warps][10 outputs][32] K-split partials
constexpr int HD_W2 = HD_RED + 10 * 10 * TILE;    // [10 classes][10 hidden] this CTA's columns of fc2.weight
constexpr int HD_H = HD_W2 + 100;                 // [10][32] this CTA's hidden rows
constexpr int HD_Z = HD_H + 10 * TILE;            // [5 ranks][10 classes][32] partial logits (filled in CTA 0)
constexpr int HD_B = HD_Z + 5 * 10 * TILE;        // fc1b[10] (own rows) + fc2b[10]
constexpr int HD_SMEM_FLOATS = HD_B + 24;
constexpr int HD_SMEM_BYTES = HD_SMEM_FLOATS * 4 + 16;

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void st_cluster_f32(const float* local_smem, uint32_t rank, float v) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(local_smem)), "r"(rank));
    asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(remote), "f"(v) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// What the NLL variant appends to the tail of CTA 0: nll_loss + metric counts of the tile (loss.cu's
// arithmetic, per tile instead of per 256 rows), and the last tile to finish folds all tiles in a fixed order.
struct head_nll_args {
    const int64_t* target;
    int64_t ignore_index;
    pdt_stats* out;
    unsigned* ticket;
    nll_partial* partial;
    float* ring;
    int32_t* ring_count;
    int topk, ring_cap;
};

template <bool NLL>
__global__ void __cluster_dims__(HD_CLUSTER, 1, 1) __launch_bounds__(HD_THREADS)
head_fwd_kernel(const float* __restrict__ p2, const float* __restrict__ fc1w, const float* __restrict__ fc1b,
                const float* __restrict__ fc2w, const float* __restrict__ fc2b,
                const uint8_t* __restrict__ keep1d, float scale, float* __restrict__ hd,
                float* __restrict__ logp, int64_t n, head_nll_args la) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float* sm = reinterpret_cast<float*>(smem_raw);
    float* w1s = sm + HD_W1;
    float* acts = sm + HD_ACT;
    float* red = sm + HD_RED;
    float* w2s = sm + HD_W2;
    float* hs = sm + HD_H;
    float* zs = sm + HD_Z;
    float* bs = sm + HD_B;
    uint64_t* bar = reinterpret_cast<uint64_t*>(sm + HD_SMEM_FLOATS);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t r = cluster_ctarank();
    const int64_t tile = blockIdx.y;
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(bar, (3200 + P2F * TILE) * 4);
        bulk_g2s(w1s, fc1w + r * 3200, 3200 * 4, bar);
    }
    if (threadIdx.x < 100) w2s[threadIdx.x] = fc2w[(threadIdx.x / 10) * 50 + r * 10 + threadIdx.x % 10];
    else if (threadIdx.x < 110) bs[threadIdx.x - 100] = fc1b[r * 10 + threadIdx.x - 100];
    else if (threadIdx.x < 120) bs[threadIdx.x - 100] = fc2b[threadIdx.x - 110];
    const int64_t sample = tile * TILE + lane;
    // the dropout bit of hidden unit (r*10 + warp) for this lane's sample, fetched while TMA is in flight
    bool keep = true;
    if (keep1d != nullptr) keep = sample < n && keep1d[sample * 50 + r * 10 + warp] != 0;
    int64_t tgt = -1;   // the criterion's target of this lane's sample (part of the input batch, not of the previous kernel)
    if (NLL && r == 0 && warp == 0 && sample < n) tgt = la.target[sample];
    pdl_wait();       // p2 is the previous kernel's output: only now may it be fetched
    if (threadIdx.x == 0) bulk_g2s(acts, p2 + tile * (int64_t)(P2F * TILE), P2F * TILE * 4, bar);
    __syncthreads();
    cluster_sync_all();   // every CTA of the cluster is resident before anyone writes into CTA 0
    mbar_wait(bar, 0);

    // fc1, K split over the 10 warps: warp w covers inputs [32w, 32w+32) for all 10 outputs of this CTA
    {
        float acc[10];
#pragma unroll
        for (int o = 0; o < 10; ++o) acc[o] = 0.f;
#pragma unroll 2
        for (int q = 0; q < 8; ++q) {
            const int f = warp * 32 + q * 4;
            const float x0 = acts[(f + 0) * TILE + lane], x1 = acts[(f + 1) * TILE + lane];
            const float x2 = acts[(f + 2) * TILE + lane], x3 = acts[(f + 3) * TILE + lane];
#pragma unroll
            for (int o = 0; o < 10; ++o) {
                const float4 t = *reinterpret_cast<const float4*>(w1s + o * 320 + f);
                acc[o] = fmaf(t.x, x0, acc[o]);
                acc[o] = fmaf(t.y, x1, acc[o]);
                acc[o] = fmaf(t.z, x2, acc[o]);
                acc[o] = fmaf(t.w, x3, acc[o]);
            }
        }
#pragma unroll
        for (int o = 0; o < 10; ++o) red[(warp * 10 + o) * TILE + lane] = acc[o];
    }
    pdl_trigger();
    __syncthreads();
    {   // warp o folds the 10 K-partials of output o in a fixed order
        const int o = warp;
        float v = bs[o];
#pragma unroll
        for (int w = 0; w < 10; ++w) v += red[(w * 10 + o) * TILE + lane];
        float h = fmaxf(v, 0.f);
        if (keep1d != nullptr) h = keep ? h * scale : 0.f;
        hs[o * TILE + lane] = h;
        hd[tile * (int64_t)(HF * TILE) + ((int)r * 10 + o) * TILE + lane] = h;
    }
    __syncthreads();
    {   // this CTA's share of fc2: class c = warp, hidden units r*10 .. r*10+9  -> CTA 0
        const int c = warp;
        float z = 0.f;
#pragma unroll
        for (int j = 0; j < 10; ++j) z = fmaf(w2s[c * 10 + j], hs[j * TILE + lane], z);
        st_cluster_f32(zs + ((int)r * 10 + c) * TILE + lane, 0, z);
    }
    cluster_sync_all();   // all partial logits have landed in CTA 0
    if (r == 0 && warp == 0) {
        float z[10];
#pragma unroll
        for (int c = 0; c < 10; ++c) {
            float v = bs[10 + c];
#pragma unroll
            for (int q = 0; q < HD_CLUSTER; ++q) v += zs[(q * 10 + c) * TILE + lane];
            z[c] = v;
        }
        float m = z[0];
#pragma unroll
        for (int c = 1; c < 10; ++c) m = fmaxf(m, z[c]);
        float s = 0.f;
#pragma unroll
        for (int c = 0; c < 10; ++c) s += expf(z[c] - m);
        const float lse = m + logf(s);
        unsigned prev = 0;
        const unsigned T = gridDim.y;
        if (NLL) {   // before the logp stores, so the fence below does not have to wait for them
            float loss = 0.f;
            int c1 = 0, ck = 0, nv = 0;
            const int64_t t = tgt;
            if (t >= 0 && t < 10) {
                float ot = 0.f;
#pragma unroll
                for (int c = 0; c < 10; ++c) ot = c == (int)t ? z[c] - lse : ot;
                int ahead = 0;
#pragma unroll
                for (int c = 0; c < 10; ++c) {
                    const float o = z[c] - lse;
                    ahead += (o > ot) || (o == ot && c < (int)t);
                }
                c1 = ahead < 1;
                ck = ahead < la.topk;
                if (t != la.ignore_index) {
                    loss = -ot;
                    nv = 1;
                }
            }
            loss = warp_sum(loss);
            c1 = warp_sum_i(c1);
            ck = warp_sum_i(ck);
            nv = warp_sum_i(nv);
            if (lane == 0) {
                la.partial[tile] = {loss, c1, ck, nv};
                __threadfence();
                prev = atomicInc(la.ticket, T - 1);
            }
        }
        if (sample < n) {
#pragma unroll
            for (int c = 0; c < 10; ++c) logp[sample * 10 + c] = z[c] - lse;
        }
        if (NLL) {
            prev = __shfl_sync(0xffffffffu, prev, 0);
            if (prev == T - 1) {     // last tile: lane l folds tiles l, l+32, ... then a fixed shuffle tree
                __threadfence();
                nll_partial p = {0.f, 0, 0, 0};
                for (unsigned g = lane; g < T; g += 32) {
                    const volatile nll_partial* q = la.partial + g;
                    p.loss += q->loss; p.top1 += q->top1; p.topk += q->topk; p.valid += q->valid;
                }
                p.loss = warp_sum(p.loss);
                p.top1 = warp_sum_i(p.top1);
                p.topk = warp_sum_i(p.topk);
                p.valid = warp_sum_i(p.valid);
                if (lane == 0) {
                    la.out->loss_sum = p.loss;
                    la.out->loss_mean = p.loss / (float)p.valid;
                    la.out->top1 = p.top1;
                    la.out->topk = p.topk;
                    la.out->n_valid = p.valid;
                    la.out->n_rows = (int32_t)n;
                    if (la.ring != nullptr) {
                        const int32_t c = *la.ring_count;
                        la.ring[c % la.ring_cap] = p.loss / (float)p.valid;
                        *la.ring_count = c + 1;
                    }
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// Training step, middle section in ONE kernel: conv2 forward -> head forward -> criterion -> head
// backward (model/model.py:17-22, model/loss.py:4-5 and their autograd; trainer/trainer.py:49-57).
// The four conv2 CTAs of a tile form a thread-block cluster.  Each ends conv2 holding 80 of the
// tile's 320 pooled features in shared memory, so fc1 is split over FEATURES: a CTA needs only its
// [50][80] slice of fc1.weight, forward (partial hidden sums, exchanged once through distributed
// shared memory) and backward (dq2 and the fc1.weight partial of the same 80 features) alike.  The
// small rest (fc2, log_softmax, nll, dz2, dz1: ~1,500 FMA per lane) is recomputed by every CTA.
// The loss gradient is taken for grad_loss == 1; the reduction kernel applies the real grad_loss,
// which only arrives with backward (all partials are linear in it).
// Replaces conv2_fwd + head_fwd + nll_metrics + nll_bwd + head_bwd: p2/hd/logp no longer make a
// round trip through memory between them and three kernel boundaries disappear.
// ------------------------------------------------------------------------------------------
constexpr int CT_W1S = C2_SMEM_PATCH + C2_SMEM_W;   // [50][80] this CTA's columns of fc1.weight
constexpr int CT_W2S = CT_W1S + 50 * 80;            // [10][50] fc2.weight
constexpr int CT_B1S = CT_W2S + 500;                // fc1.bias (50, padded to 52)
constexpr int CT_B2S = CT_B1S + 52;                 // fc2.bias (10, padded to 12)
constexpr int CT_TG = CT_B2S + 12;                  // [32] int: targets of the tile
constexpr int CT_NV = CT_TG + 32;                   // per-warp counts of valid targets + their total (padded to 20)
constexpr int CT_BAR = CT_NV + 20;                  // 10 mbarriers (8 bytes each)
constexpr int CT_SMEM_BYTES = (CT_BAR + 20) * 4;
// after the main loop the 80 KB window is dead; these live in its place
constexpr int CT_P2S = 0;                           // [80][33] this CTA's pooled features
constexpr int CT_HP = CT_P2S + 80 * 33;             // [4 ranks][50][32] partial hidden sums (written by all 4 CTAs)
constexpr int CT_HS = CT_HP + 4 * 50 * TILE;        // [50][33] hidden layer after relu + dropout
constexpr int CT_ZS = CT_HS + 50 * 33;              // [10][32] logits
constexpr int CT_DZ2 = CT_ZS + 10 * TILE;           // [10][33] (+2 pad)
constexpr int CT_DZ1S = CT_DZ2 + 332;               // [50][32] lane layout
constexpr int CT_DZ1T = CT_DZ1S + 50 * TILE;        // [32][52] sample major
static_assert(CT_DZ1T + 32 * 52 <= C2_SMEM_PATCH, "head buffers must fit in the dead window");
static_assert(CT_HP % 4 == 0 && CT_DZ1T % 2 == 0 && CT_W1S % 4 == 0 && CT_BAR % 2 == 0, "vector load / mbarrier alignment");

// Tensor-core variant (conv2 as in conv2_tc.cu): CTA q of the cluster computes conv2 for samples 8q..8q+7 of the tile
// (all 320 features) and the four CTAs then exchange 80-feature x 8-sample blocks through distributed shared memory,
// after which every CTA holds what the CUDA-core variant leaves it with -- its 80 features of all 32 samples -- and
// the head phases run unchanged.  Nothing is aliased: the tensor-core operands, the head's constants and the head's
// working buffers each have their own shared memory (214 KB in all).
constexpr int CTT_HEAD = (TC_SMEM_BYTES + 15) / 16 * 4;                    // floats: start of the head's constants
constexpr int CTT_SHIFT = CTT_HEAD - CT_W1S;                               // added to CT_W1S .. CT_BAR
constexpr int CTT_DEAD = CTT_HEAD + (CT_BAR + 20 - CT_W1S);                // floats: start of the head's working buffers
constexpr int CTT_SMEM_BYTES = (CTT_DEAD + CT_DZ1T + 32 * 52) * 4;
static_assert(CTT_SMEM_BYTES <= 227 * 1024 && CTT_HEAD % 4 == 0 && CTT_DEAD % 4 == 0, "shared memory of the tensor-core variant");

struct head_train_args {
    const float *fc1w, *fc1b, *fc2w, *fc2b;
    const uint8_t* keep1d;
    float scale1d;
    float* hd;
    float* logp;
    float* dq2;
    float* part3;
    head_nll_args la;
};

// feature fl in [0,80) of CTA (by,bx) -> index in view(-1,320) order; fl = co*4 + half*2 + px
__device__ __forceinline__ int ct_feature(int fl, int by, int bx) {
    return (fl >> 2) * 16 + (2 * by + ((fl >> 1) & 1)) * 4 + 2 * bx + (fl & 1);
}

// 16 warps: warps 0..7 run conv2 exactly as conv2_fwd_kernel does; all 16 stage the operands and share the
// head phases (twice the warps to hide the shared-memory latency of those short, dependent phases).
constexpr int CT_THREADS = 512;
constexpr int CT_WARPS = CT_THREADS / 32;

template <bool TC>
__global__ void __cluster_dims__(4, 1, 1) __launch_bounds__(CT_THREADS)
conv2_head_train_kernel(const float* __restrict__ p1, const float* __restrict__ w, const float* __restrict__ b,
                        const uint8_t* __restrict__ keep2d, float scale, float* __restrict__ p2,
                        uint8_t* __restrict__ idx2, int64_t n, head_train_args ha) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float* sm = reinterpret_cast<float*>(smem_raw);
    constexpr int HS = TC ? CTT_SHIFT : 0;       // where the head's constants start
    constexpr int DS = TC ? CTT_DEAD : 0;        // where the head's working buffers start
    float* patch = sm;
    float* ws = sm + C2_SMEM_PATCH;
    float* w1s = sm + HS + CT_W1S;
    float* w2s = sm + HS + CT_W2S;
    float* b1s = sm + HS + CT_B1S;
    float* b2s = sm + HS + CT_B2S;
    int* tg = reinterpret_cast<int*>(sm + HS + CT_TG);
    int* nvs = reinterpret_cast<int*>(sm + HS + CT_NV);
    uint64_t* bar = reinterpret_cast<uint64_t*>(sm + HS + CT_BAR);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t tile = blockIdx.x >> 2;
    const int sb = blockIdx.x & 3, by = sb >> 1, bx = sb & 1;   // sb == rank in the cluster
    const int64_t sample = tile * TILE + lane;
    const head_nll_args& la = ha.la;

    // tensor-core variant: barriers, the raw weights' bulk copy and the TMEM allocation come first
    uint64_t* mma_done = reinterpret_cast<uint64_t*>(smem_raw + TC_OFF_BAR);
    uint64_t* ready = mma_done + 2;
    uint64_t* wbar = mma_done + 4;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem_raw + TC_OFF_BAR + 48);
    float* wraw = reinterpret_cast<float*>(smem_raw + TC_OFF_A + A_BUF_PADDED);
    if (TC) {
        if (threadIdx.x == 0) {
            mbar_init(&mma_done[0], 1);
            mbar_init(&mma_done[1], 1);
            mbar_init(&ready[0], 1);
            mbar_init(&ready[1], 1);
            mbar_init(wbar, 1);
            mbar_fence_init();
            mbar_arrive_expect_tx(wbar, 5000 * 4);
            bulk_g2s(wraw, w, 5000 * 4, wbar);
        }
        if (warp == 8) tmem_alloc(tmem_slot, 2 * TC_PAIR_COLS);
    } else {
        conv2_stage_weights<CT_THREADS>(w, ws, bar);
    }
    {   // head operands: all of them are older than the previous kernel (weights, targets)
        if (threadIdx.x < 480) {      // thread = (feature fl, one of 6 row groups): no division in the loop
            const int fl = threadIdx.x % 80, j0 = threadIdx.x / 80;
            const float* src = ha.fc1w + ct_feature(fl, by, bx);
            float tmp[9];
#pragma unroll
            for (int i = 0; i < 9; ++i) {
                const int j = j0 + 6 * i;
                tmp[i] = j < HF ? __ldg(src + j * 320) : 0.f;
            }
#pragma unroll
            for (int i = 0; i < 9; ++i) {
                const int j = j0 + 6 * i;
                if (j < HF) w1s[j * 80 + fl] = tmp[i];
            }
        }
        if (threadIdx.x < 500) w2s[threadIdx.x] = __ldg(ha.fc2w + threadIdx.x);
        if (threadIdx.x < 50) b1s[threadIdx.x] = __ldg(ha.fc1b + threadIdx.x);
        else if (threadIdx.x < 60) b2s[threadIdx.x - 50] = __ldg(ha.fc2b + threadIdx.x - 50);
        if (warp == 0) {
            const int64_t t = sample < n ? la.target[sample] : -1;
            tg[lane] = (t >= 0 && t < NCLS) ? (int)t : -1;
        }
        // n_valid of the whole batch, needed for the loss gradient before any grid-wide result exists:
        // every CTA counts for itself (n <= 8192 on this path: at most 64 KB of targets from L2)
        int cnt = 0;
#pragma unroll 4
        for (int64_t i = threadIdx.x; i < n; i += CT_THREADS) {
            const int64_t t = la.target[i];
            cnt += (t >= 0 && t < NCLS && t != la.ignore_index) ? 1 : 0;
        }
        cnt = warp_sum_i(cnt);
        if (lane == 0) nvs[warp] = cnt;
    }
    const int cg = (warp & 7) >> 1, half = warp & 1;
    float v[5][2];
    float* stage_p2 = reinterpret_cast<float*>(smem_raw + TC_OFF_P2);
    uint8_t* stage_idx = smem_raw + TC_OFF_IDX;
    uint32_t tmem_base = 0;
    if (TC) {
        // conv2 of samples 8 sb .. 8 sb + 7 on the tensor cores: warps 0-7 build operands and run epilogues, lane 0 of
        // warp 8 issues the MMAs (see conv2_fwd_tc_kernel); warps 9-15 have nothing to do until the head phases
        tc_fence_before();
        __syncthreads();
        tc_fence_after();
        tmem_base = *tmem_slot;
        // all 16 warps expand the weights into the B operand (older than the previous kernel: before the wait)
        mbar_wait_bounded(wbar, 0);
        tc_stage_weights<CT_THREADS>(wraw, smem_raw);
        if (threadIdx.x < 96) {      // the zero rows behind each A buffer
            const int bufi = threadIdx.x / 48, k = threadIdx.x % 48;
            *reinterpret_cast<float4*>(smem_raw + TC_OFF_A + bufi * A_BUF_PADDED + 2 * A_PART + k * 16) =
                make_float4(0.f, 0.f, 0.f, 0.f);
        }
        fence_async_smem();
        __syncthreads();             // B complete; the raw weights' space is operand buffer 1 again
        const uint32_t a_addr = smem_u32(smem_raw + TC_OFF_A), b_addr = smem_u32(smem_raw + TC_OFF_B);
        const int64_t sample0 = tile * TILE + 8 * sb;
        if (warp == 8) {
            if (lane == 0) {
#pragma unroll 1
                for (int p = 0; p < TC_PAIRS; ++p) {
                    mbar_wait_bounded(&ready[p & 1], (p >> 1) & 1);
                    tc_fence_after();
                    tc_issue(a_addr + (p & 1) * A_BUF_PADDED, b_addr, tmem_base + (p & 1) * TC_PAIR_COLS, &mma_done[p & 1]);
                }
            }
        } else if (warp < 8) {
            float bias[10];
#pragma unroll
            for (int j = 0; j < 10; ++j) bias[j] = __ldg(b + 10 * (warp >> 2) + j);
            pdl_wait();
            tc_build_a(p1, smem_raw + TC_OFF_A, tile, 8 * sb, n);
            fence_async_smem();
            workers_sync();
            if (threadIdx.x == 0) mbar_arrive(&ready[0]);
#pragma unroll 1
            for (int p = 0; p < TC_PAIRS; ++p) {
                const int nb = (p + 1) & 1;
                float sc[10];
                tc_keep_scale(keep2d, scale, sample0, p, n, sc);
                if (p + 1 < TC_PAIRS) {
                    tc_build_a(p1, smem_raw + TC_OFF_A + nb * A_BUF_PADDED, tile, 8 * sb + 2 * (p + 1), n);
                    fence_async_smem();
                    tc_fence_before();
                    workers_sync();
                    if (threadIdx.x == 0) mbar_arrive(&ready[nb]);
                }
                mbar_wait_bounded(&mma_done[p & 1], (p >> 1) & 1);
                tc_fence_after();
                tc_epilogue(tmem_base + (p & 1) * TC_PAIR_COLS, bias, sc, stage_p2, stage_idx, p);
            }
            tc_fence_before();
        } else {
            pdl_wait();
        }
        __syncthreads();
        // this CTA's 8 samples of p2 / idx2 to memory (the backward kernels read them)
        float* p2t = p2 + tile * (int64_t)(P2F * TILE) + 8 * sb;
        uint8_t* i2t = idx2 + tile * (int64_t)(PDT_IDX2_BYTES * TILE) + 8 * sb;
        for (int i = threadIdx.x; i < P2F * 2; i += CT_THREADS) {
            const int f = i >> 1, h = i & 1;
            *reinterpret_cast<float4*>(p2t + f * TILE + 4 * h) = *reinterpret_cast<const float4*>(stage_p2 + f * TC_SAMPLES + 4 * h);
        }
        for (int f = threadIdx.x; f < P2F; f += CT_THREADS)
            *reinterpret_cast<uint2*>(i2t + f * TILE) = *reinterpret_cast<const uint2*>(stage_idx + f * TC_SAMPLES);
    } else {
        pdl_wait();
        __syncthreads();
        if (warp == 0) conv2_issue_window(p1, patch, bar, tile, by, bx, lane);
        __syncthreads();
        if (warp < 8) {
            float acc[5][8];
            conv2_main(patch, ws, bar, cg, half, lane, acc);
            conv2_epilogue(acc, b, keep2d, scale, p2, idx2, n, tile, by, bx, cg, half, lane, v);
        }
    }
    // dropout bits of this warp's hidden units (w, w+16, ...): fetched now, used after the two cluster barriers
    unsigned keepbits = 0xfu;
    if (ha.keep1d != nullptr) {
        keepbits = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int j = warp + CT_WARPS * i;
            if (j < HF && sample < n) keepbits |= (ha.keep1d[sample * 50 + j] ? 1u : 0u) << i;
        }
    }

    float* p2s = sm + DS + CT_P2S;
    float* hp = sm + DS + CT_HP;
    float* hs = sm + DS + CT_HS;
    float* zs = sm + DS + CT_ZS;
    float* dz2s = sm + DS + CT_DZ2;
    float* dz1s = sm + DS + CT_DZ1S;
    float* dz1t = sm + DS + CT_DZ1T;

    // every CTA of the cluster is past its main loop: the windows are dead, here and in the peers
    cluster_sync_all();
    if (TC) {
        // every CTA sends each peer (and itself) the 80 features that peer's head slice owns, for its own 8 samples
        for (int i = threadIdx.x; i < 4 * 80 * TC_SAMPLES; i += CT_THREADS) {
            const int peer = i / (80 * TC_SAMPLES), rem = i - peer * (80 * TC_SAMPLES), fl = rem >> 3, s8 = rem & 7;
            const float val = stage_p2[ct_feature(fl, peer >> 1, peer & 1) * TC_SAMPLES + s8];
            st_cluster_f32(p2s + fl * 33 + 8 * sb + s8, (uint32_t)peer, val);
        }
    }
    if (!TC && warp < 8) {
#pragma unroll
        for (int j = 0; j < 5; ++j)
#pragma unroll
            for (int px = 0; px < 2; ++px) p2s[(((cg * 5 + j) << 2) + half * 2 + px) * 33 + lane] = v[j][px];
    } else if (warp == 8) {
        if (lane == 0) {
            int t = 0;
#pragma unroll
            for (int k = 0; k < CT_WARPS; ++k) t += nvs[k];
            nvs[CT_WARPS] = t;
        }
    } else if (warp < 11) {
        const int i = threadIdx.x - 9 * 32;
        dz1t[(i >> 1) * 52 + 50 + (i & 1)] = 0.f;
    }
    if (TC) {
        cluster_sync_all();          // the peers' feature blocks have landed in p2s
        if (warp == 8) {             // every accumulator was read before the barriers above
            tc_fence_after();
            tmem_dealloc(tmem_base, 2 * TC_PAIR_COLS);
        }
    } else {
        __syncthreads();
    }
    // fc1 over this CTA's 80 features: warp w takes hidden units w, w+16, ... (4 or 3 of them)
    {
        float acc[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[i] = 0.f;
#pragma unroll 4
        for (int f4 = 0; f4 < 20; ++f4) {
            const float x0 = p2s[(4 * f4 + 0) * 33 + lane], x1 = p2s[(4 * f4 + 1) * 33 + lane];
            const float x2 = p2s[(4 * f4 + 2) * 33 + lane], x3 = p2s[(4 * f4 + 3) * 33 + lane];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int j = warp + CT_WARPS * i;
                if (j < HF) {
                    const float4 t = *reinterpret_cast<const float4*>(w1s + j * 80 + 4 * f4);
                    acc[i] = fmaf(t.x, x0, acc[i]);
                    acc[i] = fmaf(t.y, x1, acc[i]);
                    acc[i] = fmaf(t.z, x2, acc[i]);
                    acc[i] = fmaf(t.w, x3, acc[i]);
                }
            }
        }
        // all-gather of the partial sums: slot `sb` of every CTA of the cluster
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int j = warp + CT_WARPS * i;
            if (j < HF) {
#pragma unroll
                for (uint32_t peer = 0; peer < 4; ++peer)
                    st_cluster_f32(hp + (sb * HF + j) * TILE + lane, peer, acc[i]);
            }
        }
    }
    cluster_sync_all();   // all four partial sums of every hidden unit have landed here
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int j = warp + CT_WARPS * i;
        if (j < HF) {
            float s = b1s[j];
#pragma unroll
            for (int q = 0; q < 4; ++q) s += hp[(q * HF + j) * TILE + lane];     // fixed order: same in all 4 CTAs
            float h = fmaxf(s, 0.f);
            if (ha.keep1d != nullptr) h = ((keepbits >> i) & 1u) ? h * ha.scale1d : 0.f;
            hs[j * 33 + lane] = h;
            if (sb == 0) ha.hd[tile * (int64_t)(HF * TILE) + j * TILE + lane] = h;
        }
    }
    __syncthreads();
    // fc2: warp c < 10 takes class c
    if (warp < NCLS) {
        const int c = warp;
        float z0 = b2s[c], z1 = 0.f;
#pragma unroll
        for (int j = 0; j < HF; j += 2) {
            z0 = fmaf(w2s[c * HF + j], hs[j * 33 + lane], z0);
            z1 = fmaf(w2s[c * HF + j + 1], hs[(j + 1) * 33 + lane], z1);
        }
        zs[c * TILE + lane] = z0 + z1;
    }
    __syncthreads();
    // log_softmax and the loss gradient w.r.t. the logits: EVERY warp forms them for the tile's 32 samples
    // (20 exp + 1 log per lane) rather than wait for one warp to publish them
    float dz2[NCLS];
    {
        float z[NCLS];
#pragma unroll
        for (int c = 0; c < NCLS; ++c) z[c] = zs[c * TILE + lane];
        float m = z[0];
#pragma unroll
        for (int c = 1; c < NCLS; ++c) m = fmaxf(m, z[c]);
        float se = 0.f;
#pragma unroll
        for (int c = 0; c < NCLS; ++c) se += expf(z[c] - m);
        const float lse = m + logf(se);
        const int t = tg[lane];
        const bool counted = t >= 0 && (int64_t)t != la.ignore_index;
        // dz2 = dlogp - softmax * sum(dlogp), dlogp = -1/n_valid at the target (grad_loss taken as 1)
        const float g = counted ? 1.f / (float)nvs[CT_WARPS] : 0.f;
#pragma unroll
        for (int c = 0; c < NCLS; ++c) dz2[c] = ((c == t) ? -g : 0.f) + expf(z[c] - lse) * g;      // sum(dlogp) = -g
        if (warp == 0) {
#pragma unroll
            for (int c = 0; c < NCLS; ++c) dz2s[c * 33 + lane] = dz2[c];
            if (sb == 0 && sample < n) {
#pragma unroll
                for (int c = 0; c < NCLS; ++c) ha.logp[sample * NCLS + c] = z[c] - lse;
            }
        }
    }
    // ---- head backward (autograd of model.py:18-22), for this CTA's 80 features ------------------
    // dz1 = relu'/dropout of (W2^T dz2): warp w, hidden units w, w+16, ...
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int j = warp + CT_WARPS * i;
        if (j < HF) {
            float dh = 0.f;
#pragma unroll
            for (int c = 0; c < NCLS; ++c) dh = fmaf(dz2[c], w2s[c * HF + j], dh);
            const float sc1 = ha.keep1d != nullptr ? ha.scale1d : 1.f;
            const float dz1 = hs[j * 33 + lane] > 0.f ? dh * sc1 : 0.f;
            dz1s[j * TILE + lane] = dz1;
            dz1t[lane * 52 + j] = dz1;
        }
    }
    __syncthreads();
    // dq2[f] = (p2[f] > 0) * dropout2d scale * sum_j dz1[j] * W1[j][f]; lane == sample; 40 pairs of features
    {
        float d[HF];
#pragma unroll
        for (int j = 0; j < HF; ++j) d[j] = dz1s[j * TILE + lane];
        const float sc2 = keep2d != nullptr ? scale : 1.f;
        float* out = ha.dq2 + tile * (int64_t)(P2F * TILE) + lane;
#pragma unroll 1
        for (int g2 = warp; g2 < 40; g2 += CT_WARPS) {
            float a0 = 0.f, a1 = 0.f;
#pragma unroll
            for (int j = 0; j < HF; ++j) {
                const float2 t = *reinterpret_cast<const float2*>(w1s + j * 80 + 2 * g2);
                a0 = fmaf(t.x, d[j], a0);
                a1 = fmaf(t.y, d[j], a1);
            }
            const int fl = 2 * g2;     // even: the pair (fl, fl+1) is px = 0, 1 of one pooled row -> adjacent features
            const int f = ct_feature(fl, by, bx);
            out[f * TILE] = p2s[fl * 33 + lane] > 0.f ? a0 * sc2 : 0.f;
            out[(f + 1) * TILE] = p2s[(fl + 1) * 33 + lane] > 0.f ? a1 * sc2 : 0.f;
        }
    }
    pdl_trigger();
    // fc1.weight partial of the tile, columns of this CTA: lane == feature, 10 hidden units per task (15 tasks)
    float* part = ha.part3 + tile * (int64_t)PART3;
    if (warp < 15) {
        const int fg = warp % 3, jg = warp / 3;
        const int fl = fg * 32 + lane;
        const int flc = fl < 80 ? fl : 79;
        float acc[10];
#pragma unroll
        for (int j = 0; j < 10; ++j) acc[j] = 0.f;
#pragma unroll 4
        for (int s_ = 0; s_ < 32; ++s_) {
            const float a = p2s[flc * 33 + s_];
            const float2* dz = reinterpret_cast<const float2*>(dz1t + s_ * 52 + jg * 10);
#pragma unroll
            for (int q = 0; q < 5; ++q) {
                const float2 t = dz[q];
                acc[q * 2 + 0] = fmaf(t.x, a, acc[q * 2 + 0]);
                acc[q * 2 + 1] = fmaf(t.y, a, acc[q * 2 + 1]);
            }
        }
        if (fl < 80) {
            const int f = ct_feature(fl, by, bx);
#pragma unroll
            for (int j = 0; j < 10; ++j) part[(jg * 10 + j) * 320 + f] = acc[j];
        }
    }
    // fc1.bias, fc2.weight, fc2.bias partials (560 sums over the 32 samples): every CTA of the cluster holds the
    // same dz1 / dz2 / hidden values, so each takes a quarter of them
    {
        const int e = sb * 140 + threadIdx.x;
        if (threadIdx.x < 140) {
            float s_ = 0.f;
            if (e < 50) {
                for (int k = 0; k < 32; ++k) s_ += dz1t[k * 52 + e];
                part[16000 + e] = s_;
            } else if (e < 550) {
                const int c = (e - 50) / 50, j = (e - 50) % 50;
                for (int k = 0; k < 32; ++k) s_ = fmaf(dz2s[c * 33 + k], hs[j * 33 + k], s_);
                part[16050 + (e - 50)] = s_;
            } else {
                const int c = e - 550;
                for (int k = 0; k < 32; ++k) s_ += dz2s[c * 33 + k];
                part[16550 + c] = s_;
            }
        }
    }
    // criterion value + metric counts of the tile, then the last tile folds all tiles (as head_fwd_kernel<true>);
    // kept for the very end so that no other warp of this CTA waits behind its fences
    if (sb == 0 && warp == 0) {
        float z[NCLS];
#pragma unroll
        for (int c = 0; c < NCLS; ++c) z[c] = zs[c * TILE + lane];
        float m = z[0];
#pragma unroll
        for (int c = 1; c < NCLS; ++c) m = fmaxf(m, z[c]);
        float se = 0.f;
#pragma unroll
        for (int c = 0; c < NCLS; ++c) se += expf(z[c] - m);
        const float lse = m + logf(se);
        const int t = tg[lane];
        const bool counted = t >= 0 && (int64_t)t != la.ignore_index;
        {
            unsigned prev = 0;
            const unsigned T = gridDim.x >> 2;
            float loss = 0.f;
            int c1 = 0, ck = 0, nv = 0;
            if (t >= 0) {
                float ot = 0.f;
#pragma unroll
                for (int c = 0; c < NCLS; ++c) ot = c == t ? z[c] - lse : ot;
                int ahead = 0;
#pragma unroll
                for (int c = 0; c < NCLS; ++c) {
                    const float o = z[c] - lse;
                    ahead += (o > ot) || (o == ot && c < t);
                }
                c1 = ahead < 1;
                ck = ahead < la.topk;
                if (counted) {
                    loss = -ot;
                    nv = 1;
                }
            }
            loss = warp_sum(loss);
            c1 = warp_sum_i(c1);
            ck = warp_sum_i(ck);
            nv = warp_sum_i(nv);
            if (lane == 0) {
                la.partial[tile] = {loss, c1, ck, nv};
                __threadfence();
                prev = atomicInc(la.ticket, T - 1);
            }
            prev = __shfl_sync(0xffffffffu, prev, 0);
            if (prev == T - 1) {     // last tile: lane l folds tiles l, l+32, ... then a fixed shuffle tree
                __threadfence();
                nll_partial p = {0.f, 0, 0, 0};
                for (unsigned g = lane; g < T; g += 32) {
                    const volatile nll_partial* q = la.partial + g;
                    p.loss += q->loss; p.top1 += q->top1; p.topk += q->topk; p.valid += q->valid;
                }
                p.loss = warp_sum(p.loss);
                p.top1 = warp_sum_i(p.top1);
                p.topk = warp_sum_i(p.topk);
                p.valid = warp_sum_i(p.valid);
                if (lane == 0) {
                    la.out->loss_sum = p.loss;
                    la.out->loss_mean = p.loss / (float)p.valid;
                    la.out->top1 = p.top1;
                    la.out->topk = p.topk;
                    la.out->n_valid = p.valid;
                    la.out->n_rows = (int32_t)n;
                    if (la.ring != nullptr) {
                        const int32_t c = *la.ring_count;
                        la.ring[c % la.ring_cap] = p.loss / (float)p.valid;
                        *la.ring_count = c + 1;
                    }
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// Dropout keep masks from Philox4x32-10 (counter = unit index / 4 + offset, key = seed).
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, ctr.x), lo0 = 0xD2511F53u * ctr.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, ctr.z), lo1 = 0xCD9E8D57u * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += 0x9E3779B9u;
        key.y += 0xBB67AE85u;
    }
    return ctr;
}

__device__ __forceinline__ void
dropout_masks_body(uint8_t* __restrict__ keep2d, uint8_t* __restrict__ keep1d, int64_t n, uint32_t thresh,
                   uint64_t* __restrict__ rng_state, unsigned vblock, unsigned vgrid) {
    pdl_wait();
    const uint64_t seed = rng_state[0], offset = rng_state[1];
    const int64_t total4 = (n * 70 + 3) / 4;  // 20 + 50 units per sample, four per Philox call
    for (int64_t i = vblock * (int64_t)blockDim.x + threadIdx.x; i < total4;
         i += (int64_t)vgrid * blockDim.x) {
        const uint64_t c = offset + (uint64_t)i;
        const uint4 r = philox4x32_10(make_uint4((uint32_t)c, (uint32_t)(c >> 32), 0u, 0u),
                                      make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
        const uint32_t rv[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int64_t u = i * 4 + q;
            if (u < n * 70) {
                const uint8_t keep = rv[q] >= thresh ? 1 : 0;  // P(keep) = 1 - p
                if (u < n * 20) keep2d[u] = keep;
                else keep1d[u - n * 20] = keep;
            }
        }
    }
    pdl_trigger();
    // the last block to finish advances the offset; the ticket lives in rng_state[2]
    __shared__ bool last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned* ticket = reinterpret_cast<unsigned*>(rng_state + 2);
        last = atomicInc(ticket, vgrid - 1) == vgrid - 1;
    }
    __syncthreads();
    if (last && threadIdx.x == 0) rng_state[1] = offset + (uint64_t)total4;
}

__global__ void dropout_masks_kernel(uint8_t* __restrict__ keep2d, uint8_t* __restrict__ keep1d, int64_t n,
                                     uint32_t thresh, uint64_t* __restrict__ rng_state) {
    dropout_masks_body(keep2d, keep1d, n, thresh, rng_state, blockIdx.x, gridDim.x);
}

template <bool U8>
__global__ void __launch_bounds__(C1_WPC * 32)
conv1_fwd_kernel(const void* __restrict__ x, x_norm nm, const float* __restrict__ w, const float* __restrict__ b,
                 float* __restrict__ p1, uint8_t* __restrict__ idx1, int64_t n, int64_t n_tasks, int csplit) {
    conv1_fwd_body<U8>(x, nm, w, b, p1, idx1, n, n_tasks, csplit);
}

// conv1 and the dropout-mask generator in ONE launch: the masks are independent of conv1 and only needed
// by the kernels after it, so their CTAs ride at the end of conv1's grid (one launch and one graph edge less).
template <bool U8>
__global__ void __launch_bounds__(C1_WPC * 32)
conv1_fwd_masks_kernel(const void* __restrict__ x, x_norm nm, const float* __restrict__ w, const float* __restrict__ b,
                       float* __restrict__ p1, uint8_t* __restrict__ idx1, int64_t n, int64_t n_tasks, int csplit,
                       unsigned conv_ctas, uint8_t* __restrict__ keep2d, uint8_t* __restrict__ keep1d,
                       uint32_t thresh, uint64_t* __restrict__ rng_state) {
    if (blockIdx.x < conv_ctas) conv1_fwd_body<U8>(x, nm, w, b, p1, idx1, n, n_tasks, csplit);
    else dropout_masks_body(keep2d, keep1d, n, thresh, rng_state, blockIdx.x - conv_ctas, gridDim.x - conv_ctas);
}

}  // namespace pdt

using namespace pdt;

// channel groups per block: 2 while the single-split grid would leave schedulers with two warps (env PDT_C1_SPLIT)
static int conv1_channel_split(int64_t n) {
    static int forced = -1;
    if (forced < 0) {
        const char* e = getenv("PDT_C1_SPLIT");
        forced = e ? atoi(e) : 0;
    }
    if (forced == 1 || forced == 2) return forced;
    return num_tiles(n) * 72 <= 148 * 16 ? 2 : 1;      // both halves still fit one wave (16 warps per SM)
}

template <bool U8>
static int conv1_fwd_launch(const void* x, x_norm nm, const float* w, const float* b, float* p1_lane, uint8_t* idx1_lane,
                            int64_t n, void* stream) {
    const int csplit = conv1_channel_split(n);
    const int64_t tasks = num_tiles(n) * 36 * csplit;
    const unsigned grid = (unsigned)((tasks + C1_WPC - 1) / C1_WPC);
    launch_pdl(1, conv1_fwd_kernel<U8>, dim3(grid), dim3(C1_WPC * 32), 0, stream, x, nm, w, b, p1_lane, idx1_lane, n, tasks,
               csplit);
    return check_launch("conv1_fwd");
}

extern "C" int pdt_conv1_pool_relu_fwd(const float* x, const float* w, const float* b, float* p1_lane,
                                       uint8_t* idx1_lane, int64_t n, void* stream) {
    PDT_REQUIRE(n > 0 && x && w && b && p1_lane && idx1_lane, "null pointer or n <= 0");
    return conv1_fwd_launch<false>(x, x_norm{0.f, 1.f}, w, b, p1_lane, idx1_lane, n, stream);
}

extern "C" int pdt_conv1_pool_relu_fwd_u8(const uint8_t* x, float mean, float std, const float* w, const float* b,
                                          float* p1_lane, uint8_t* idx1_lane, int64_t n, void* stream) {
    PDT_REQUIRE(n > 0 && x && w && b && p1_lane && idx1_lane, "null pointer or n <= 0");
    PDT_REQUIRE(std != 0.f, "std must not be zero");
    return conv1_fwd_launch<true>(x, x_norm{mean, std}, w, b, p1_lane, idx1_lane, n, stream);
}

static uint32_t keep_threshold(float p) {
    const double t = (double)p * 4294967296.0;
    return t >= 4294967295.0 ? 4294967295u : (uint32_t)t;
}

template <bool U8>
static int conv1_fwd_masks_launch(const void* x, x_norm nm, const float* w, const float* b, float* p1_lane,
                                  uint8_t* idx1_lane, uint8_t* keep2d, uint8_t* keep1d, float p, uint64_t* rng_state,
                                  int64_t n, void* stream) {
    const int csplit = conv1_channel_split(n);
    const int64_t tasks = num_tiles(n) * 36 * csplit;
    const unsigned conv_ctas = (unsigned)((tasks + C1_WPC - 1) / C1_WPC);
    const int64_t total4 = (n * 70 + 3) / 4;
    int64_t mask_ctas = (total4 + C1_WPC * 32 - 1) / (C1_WPC * 32);
    if (mask_ctas > 148) mask_ctas = 148;
    launch_pdl(1, conv1_fwd_masks_kernel<U8>, dim3(conv_ctas + (unsigned)mask_ctas), dim3(C1_WPC * 32), 0, stream, x, nm,
               w, b, p1_lane, idx1_lane, n, tasks, csplit, conv_ctas, keep2d, keep1d, keep_threshold(p), rng_state);
    return check_launch("conv1_fwd_masks");
}

extern "C" int pdt_conv1_fwd_masks(const float* x, const float* w, const float* b, float* p1_lane, uint8_t* idx1_lane,
                                   uint8_t* keep2d, uint8_t* keep1d, float p, uint64_t* rng_state, int64_t n,
                                   void* stream) {
    PDT_REQUIRE(n > 0 && x && w && b && p1_lane && idx1_lane && keep2d && keep1d && rng_state, "null pointer or n <= 0");
    PDT_REQUIRE(p >= 0.f && p < 1.f, "p must be in [0, 1)");
    return conv1_fwd_masks_launch<false>(x, x_norm{0.f, 1.f}, w, b, p1_lane, idx1_lane, keep2d, keep1d, p, rng_state, n,
                                         stream);
}

extern "C" int pdt_conv1_fwd_masks_u8(const uint8_t* x, float mean, float std, const float* w, const float* b,
                                      float* p1_lane, uint8_t* idx1_lane, uint8_t* keep2d, uint8_t* keep1d, float p,
                                      uint64_t* rng_state, int64_t n, void* stream) {
    PDT_REQUIRE(n > 0 && x && w && b && p1_lane && idx1_lane && keep2d && keep1d && rng_state, "null pointer or n <= 0");
    PDT_REQUIRE(p >= 0.f && p < 1.f && std != 0.f, "p must be in [0, 1) and std non-zero");
    return conv1_fwd_masks_launch<true>(x, x_norm{mean, std}, w, b, p1_lane, idx1_lane, keep2d, keep1d, p, rng_state, n,
                                        stream);
}

extern "C" int pdt_conv2_drop_pool_relu_fwd(const float* p1_lane, const float* w, const float* b,
                                            const uint8_t* keep2d, float scale, float* p2_lane,
                                            uint8_t* idx2_lane, int64_t n, void* stream) {
    PDT_REQUIRE(n > 0 && p1_lane && w && b && p2_lane && idx2_lane, "null pointer or n <= 0");
    if (int rc = ensure_smem<10>(conv2_fwd_kernel, C2_SMEM_BYTES)) return rc;
    const unsigned grid = (unsigned)(num_tiles(n) * 4);
    launch_pdl(2, conv2_fwd_kernel, dim3(grid), dim3(C2_THREADS), C2_SMEM_BYTES, stream, p1_lane, w, b, keep2d, scale,
               p2_lane, idx2_lane, n);
    return check_launch("conv2_fwd");
}

extern "C" int pdt_head_fwd(const float* p2_lane, const float* fc1w, const float* fc1b, const float* fc2w,
                            const float* fc2b, const uint8_t* keep1d, float scale, float* hd_lane, float* logp,
                            int64_t n, void* stream) {
    PDT_REQUIRE(n > 0 && p2_lane && fc1w && fc1b && fc2w && fc2b && hd_lane && logp, "null pointer or n <= 0");
    if (int rc = ensure_smem<11>(head_fwd_kernel<false>, HD_SMEM_BYTES)) return rc;
    launch_pdl(3, head_fwd_kernel<false>, dim3(HD_CLUSTER, (unsigned)num_tiles(n)), dim3(HD_THREADS), HD_SMEM_BYTES,
               stream, p2_lane, fc1w, fc1b, fc2w, fc2b, keep1d, scale, hd_lane, logp, n, head_nll_args{});
    return check_launch("head_fwd");
}

extern "C" int64_t pdt_head_nll_scratch_bytes(int64_t n) { return 16 + num_tiles(n) * (int64_t)sizeof(nll_partial); }

extern "C" int pdt_head_fwd_nll(const float* p2_lane, const float* fc1w, const float* fc1b, const float* fc2w,
                                const float* fc2b, const uint8_t* keep1d, float scale, float* hd_lane, float* logp,
                                int64_t n, const int64_t* target, int topk, int64_t ignore_index, pdt_stats* out,
                                void* scratch, float* ring, int32_t* ring_count, int ring_capacity, void* stream) {
    PDT_REQUIRE(n > 0 && p2_lane && fc1w && fc1b && fc2w && fc2b && hd_lane && logp, "null pointer or n <= 0");
    PDT_REQUIRE(target && out && scratch && topk > 0, "target, out and scratch are required; topk must be positive");
    PDT_REQUIRE(ring == nullptr || (ring_count != nullptr && ring_capacity > 0), "ring needs a counter and a capacity");
    if (int rc = ensure_smem<12>(head_fwd_kernel<true>, HD_SMEM_BYTES)) return rc;
    head_nll_args la;
    la.target = target;
    la.ignore_index = ignore_index;
    la.out = out;
    la.ticket = reinterpret_cast<unsigned*>(scratch);
    la.partial = reinterpret_cast<nll_partial*>(reinterpret_cast<char*>(scratch) + 16);
    la.ring = ring;
    la.ring_count = ring_count;
    la.topk = topk;
    la.ring_cap = ring_capacity;
    launch_pdl(3, head_fwd_kernel<true>, dim3(HD_CLUSTER, (unsigned)num_tiles(n)), dim3(HD_THREADS), HD_SMEM_BYTES,
               stream, p2_lane, fc1w, fc1b, fc2w, fc2b, keep1d, scale, hd_lane, logp, n, la);
    return check_launch("head_fwd_nll");
}

template <bool TC>
static int conv2_head_train_launch(const float* p1_lane, const float* w, const float* b, const uint8_t* keep2d,
                                   float scale2d, float* p2_lane, uint8_t* idx2_lane, const float* fc1w,
                                   const float* fc1b, const float* fc2w, const float* fc2b, const uint8_t* keep1d,
                                   float scale1d, float* hd_lane, float* logp, const int64_t* target, int topk,
                                   int64_t ignore_index, pdt_stats* out, void* nll_scratch, float* ring,
                                   int32_t* ring_count, int ring_capacity, float* dq2_lane, void* scratch,
                                   int64_t n, void* stream) {
    PDT_REQUIRE(n > 0 && p1_lane && w && b && p2_lane && idx2_lane, "null pointer or n <= 0");
    PDT_REQUIRE(fc1w && fc1b && fc2w && fc2b && hd_lane && logp && dq2_lane && scratch, "null pointer");
    PDT_REQUIRE(target && out && nll_scratch && topk > 0, "target, out and nll_scratch are required; topk must be positive");
    PDT_REQUIRE(ring == nullptr || (ring_count != nullptr && ring_capacity > 0), "ring needs a counter and a capacity");
    if (n > PDT_TRAIN_FUSED_MAX_ROWS) {
        set_error("pdt_conv2_head_train: n = %lld exceeds %d rows (every CTA counts the valid targets itself); "
                  "use the separate kernels", (long long)n, PDT_TRAIN_FUSED_MAX_ROWS);
        return PDT_EINVAL;
    }
    constexpr int SMEM = TC ? CTT_SMEM_BYTES : CT_SMEM_BYTES;
    if (int rc = ensure_smem<TC ? 32 : 13>(conv2_head_train_kernel<TC>, SMEM)) return rc;
    const int64_t T = num_tiles(n);
    head_train_args ha;
    ha.fc1w = fc1w; ha.fc1b = fc1b; ha.fc2w = fc2w; ha.fc2b = fc2b;
    ha.keep1d = keep1d;
    ha.scale1d = scale1d;
    ha.hd = hd_lane;
    ha.logp = logp;
    ha.dq2 = dq2_lane;
    ha.part3 = reinterpret_cast<float*>(scratch) + part3_off(T);
    ha.la.target = target;
    ha.la.ignore_index = ignore_index;
    ha.la.out = out;
    ha.la.ticket = reinterpret_cast<unsigned*>(nll_scratch);
    ha.la.partial = reinterpret_cast<nll_partial*>(reinterpret_cast<char*>(nll_scratch) + 16);
    ha.la.ring = ring;
    ha.la.ring_count = ring_count;
    ha.la.topk = topk;
    ha.la.ring_cap = ring_capacity;
    launch_pdl(2, conv2_head_train_kernel<TC>, dim3((unsigned)(T * 4)), dim3(CT_THREADS), SMEM, stream,
               p1_lane, w, b, keep2d, scale2d, p2_lane, idx2_lane, n, ha);
    return check_launch("conv2_head_train");
}

extern "C" int pdt_conv2_head_train(const float* p1_lane, const float* w, const float* b, const uint8_t* keep2d,
                                    float scale2d, float* p2_lane, uint8_t* idx2_lane, const float* fc1w,
                                    const float* fc1b, const float* fc2w, const float* fc2b, const uint8_t* keep1d,
                                    float scale1d, float* hd_lane, float* logp, const int64_t* target, int topk,
                                    int64_t ignore_index, pdt_stats* out, void* nll_scratch, float* ring,
                                    int32_t* ring_count, int ring_capacity, float* dq2_lane, void* scratch,
                                    int64_t n, void* stream) {
    return conv2_head_train_launch<false>(p1_lane, w, b, keep2d, scale2d, p2_lane, idx2_lane, fc1w, fc1b, fc2w, fc2b, keep1d,
                                          scale1d, hd_lane, logp, target, topk, ignore_index, out, nll_scratch, ring,
                                          ring_count, ring_capacity, dq2_lane, scratch, n, stream);
}

extern "C" int pdt_conv2_head_train_tc(const float* p1_lane, const float* w, const float* b, const uint8_t* keep2d,
                                       float scale2d, float* p2_lane, uint8_t* idx2_lane, const float* fc1w,
                                       const float* fc1b, const float* fc2w, const float* fc2b, const uint8_t* keep1d,
                                       float scale1d, float* hd_lane, float* logp, const int64_t* target, int topk,
                                       int64_t ignore_index, pdt_stats* out, void* nll_scratch, float* ring,
                                       int32_t* ring_count, int ring_capacity, float* dq2_lane, void* scratch,
                                       int64_t n, void* stream) {
    return conv2_head_train_launch<true>(p1_lane, w, b, keep2d, scale2d, p2_lane, idx2_lane, fc1w, fc1b, fc2w, fc2b, keep1d,
                                         scale1d, hd_lane, logp, target, topk, ignore_index, out, nll_scratch, ring,
                                         ring_count, ring_capacity, dq2_lane, scratch, n, stream);
}

extern "C" int pdt_dropout_masks(uint8_t* keep2d, uint8_t* keep1d, int64_t n, float p, uint64_t* rng_state,
                                 void* stream) {
    PDT_REQUIRE(n > 0 && keep2d && keep1d && rng_state, "null pointer or n <= 0");
    PDT_REQUIRE(p >= 0.f && p < 1.f, "p must be in [0, 1)");
    const double t = (double)p * 4294967296.0;
    const uint32_t thresh = t >= 4294967295.0 ? 4294967295u : (uint32_t)t;
    const int64_t total4 = (n * 70 + 3) / 4;
    int grid = (int)((total4 + 255) / 256);
    if (grid > 296) grid = 296;
    launch_pdl(0, dropout_masks_kernel, dim3(grid), dim3(256), 0, stream, keep2d, keep1d, n, thresh, rng_state);
    return check_launch("dropout_masks");
}
