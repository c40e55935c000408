// Backward kernels of the MNIST CNN: autograd of reference model/model.py:15-22 as run by
// loss.backward() at trainer/trainer.py:57.  Same lane == sample mapping as fwd.cu.  Weight
// gradients are a reduction over samples, i.e. over lanes and tiles: every lane accumulates
// its own sample in registers, a warp-shuffle tree folds the 32 lanes, and the per-tile
// results go to a scratch buffer that pdt_grad_reduce sums in a fixed order (deterministic,
// pairwise-like accumulation).
#include "common.cuh"

#ifndef PDT_HB_EARLY_WAIT
#define PDT_HB_EARLY_WAIT 1
#endif

namespace pdt {

// ------------------------------------------------------------------------------------------
// head backward: log_softmax, fc2, dropout, relu, fc1
// 5 CTAs per tile, CTA r owns input features f in [64r, 64r+64): it needs only those 64 columns
// of fc1.weight (50 bulk TMA copies of 256 B) and those 64 rows of the tile's p2.  The cheap part
// every CTA needs in full (dz2, dz1: ~500 FMA per lane) is recomputed by each of them instead of
// being exchanged, so the 5 CTAs are independent.
// ------------------------------------------------------------------------------------------
constexpr int HB_THREADS = 320;
constexpr int HB_SPLIT = 5;
constexpr int HB_FW = 64;                     // features per CTA
constexpr int HB_W1 = 0;                      // [50][64]   columns of fc1.weight
constexpr int HB_P2 = HB_W1 + 50 * HB_FW;     // [64][33]
constexpr int HB_W2 = HB_P2 + HB_FW * 33;     // [10][50]
constexpr int HB_DZ1S = HB_W2 + 500;          // [50][32]   lane layout
constexpr int HB_DZ1T = HB_DZ1S + 1600;       // [32][52]   sample major
constexpr int HB_HDS = HB_DZ1T + 32 * 52;     // [50][33]
constexpr int HB_DZ2S = HB_HDS + 50 * 33;     // [10][33] (+2 pad)
constexpr int HB_FLOATS = HB_DZ2S + 332;
constexpr int HB_SMEM_BYTES = HB_FLOATS * 4 + 16;

// NLL variant: dlogp is not read but formed in place from the targets, nll_loss's backward
// (dlogp[n,c] = -(grad_loss / n_valid) at c == target != ignore_index, else 0; pdt_nll_bwd).
template <bool NLL>
__global__ void __launch_bounds__(HB_THREADS)
head_bwd_kernel(const float* __restrict__ dlogp, const float* __restrict__ logp, const float* __restrict__ hd,
                const float* __restrict__ p2, const float* __restrict__ fc1w, const float* __restrict__ fc2w,
                float scale1d, float scale2d, float* __restrict__ dq2, float* __restrict__ part3, int64_t n,
                const int64_t* __restrict__ target, const pdt_stats* __restrict__ stats,
                const float* __restrict__ grad_loss, int64_t ignore_index) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float* sm = reinterpret_cast<float*>(smem_raw);
    float* w1c = sm + HB_W1;
    float* p2s = sm + HB_P2;
    float* w2s = sm + HB_W2;
    float* dz1s = sm + HB_DZ1S;
    float* dz1t = sm + HB_DZ1T;
    float* hds = sm + HB_HDS;
    float* dz2s = sm + HB_DZ2S;
    uint64_t* bar = reinterpret_cast<uint64_t*>(sm + HB_FLOATS);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int r = blockIdx.x;                 // feature slice
    const int64_t tile = blockIdx.y;
    const int fbase = r * HB_FW;
    const int64_t sample = tile * TILE + lane;
    if (PDT_HB_EARLY_WAIT) pdl_wait();
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (warp == 0) {
        if (lane == 0) mbar_arrive_expect_tx(bar, 50 * HB_FW * 4);
        __syncwarp();
        for (int j = lane; j < 50; j += 32) bulk_g2s(w1c + j * HB_FW, fc1w + j * 320 + fbase, HB_FW * 4, bar);
    }
    float hv[5];
    {
        // every global load of this prologue is issued before the first use (one latency, not seven)
        const float* src = p2 + (tile * (int64_t)P2F + fbase) * TILE;
        float t[7], u[2];
#pragma unroll
        for (int k = 0; k < 7; ++k) {
            const int i = threadIdx.x + k * HB_THREADS;
            t[k] = i < HB_FW * TILE ? __ldcg(src + i) : 0.f;
        }
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int i = threadIdx.x + k * HB_THREADS;
            u[k] = i < 500 ? __ldcg(fc2w + i) : 0.f;
        }
#pragma unroll
        for (int jj = 0; jj < 5; ++jj) hv[jj] = __ldcg(hd + tile * (int64_t)(HF * TILE) + (warp * 5 + jj) * TILE + lane);
#pragma unroll
        for (int k = 0; k < 7; ++k) {
            const int i = threadIdx.x + k * HB_THREADS;
            if (i < HB_FW * TILE) p2s[(i >> 5) * 33 + (i & 31)] = t[k];
        }
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int i = threadIdx.x + k * HB_THREADS;
            if (i < 500) w2s[i] = u[k];
        }
        if (threadIdx.x < 64) dz1t[(threadIdx.x >> 1) * 52 + 50 + (threadIdx.x & 1)] = 0.f;
    }
    // phase A: dz2 = dlogp - softmax * sum(dlogp)   (log_softmax backward), per lane
    float dz2[10];
    {
        float dl[10], sm_[10], S = 0.f;
#pragma unroll
        for (int c = 0; c < 10; ++c) sm_[c] = expf(sample < n ? __ldcg(logp + sample * 10 + c) : 0.f);
        // everything so far (weights, saved activations, logp) is older than the previous kernel;
        // dlogp is the previous kernel's output
        pdl_wait();
        if (NLL) {
            const int64_t t = sample < n ? target[sample] : -1;
            const float g = (grad_loss ? *grad_loss : 1.f) / (float)stats->n_valid;
#pragma unroll
            for (int c = 0; c < 10; ++c) {
                dl[c] = (t == c && t != ignore_index) ? -g : 0.f;
                S += dl[c];
            }
        } else {
#pragma unroll
            for (int c = 0; c < 10; ++c) {
                dl[c] = sample < n ? dlogp[sample * 10 + c] : 0.f;
                S += dl[c];
            }
        }
#pragma unroll
        for (int c = 0; c < 10; ++c) dz2[c] = dl[c] - sm_[c] * S;
        if (warp == 0) {
#pragma unroll
            for (int c = 0; c < 10; ++c) dz2s[c * 33 + lane] = dz2[c];
        }
    }
    __syncthreads();  // w2s ready
#pragma unroll
    for (int jj = 0; jj < 5; ++jj) {
        const int j = warp * 5 + jj;
        float dh = 0.f;
#pragma unroll
        for (int c = 0; c < 10; ++c) dh = fmaf(dz2[c], w2s[c * 50 + j], dh);
        const float dz1 = hv[jj] > 0.f ? dh * scale1d : 0.f;
        dz1s[j * TILE + lane] = dz1;
        dz1t[lane * 52 + j] = dz1;
        hds[j * 33 + lane] = hv[jj];
    }
    __syncthreads();
    mbar_wait(bar, 0);

    // phase B: dq2[f] = (p2[f] > 0) * scale2d * sum_j dz1[j] * W1[j][f], lane == sample; 16 chunks of 4 f
    if (warp < 8) {
        float d[50];
#pragma unroll
        for (int j = 0; j < 50; ++j) d[j] = dz1s[j * TILE + lane];
        float* out = dq2 + (tile * (int64_t)P2F + fbase) * TILE + lane;
#pragma unroll 1
        for (int q = 0; q < 2; ++q) {
            const int f0 = (warp * 2 + q) * 4;
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
            for (int j = 0; j < 50; ++j) {
                const float4 t = *reinterpret_cast<const float4*>(w1c + j * HB_FW + f0);
                a0 = fmaf(t.x, d[j], a0);
                a1 = fmaf(t.y, d[j], a1);
                a2 = fmaf(t.z, d[j], a2);
                a3 = fmaf(t.w, d[j], a3);
            }
            out[(f0 + 0) * TILE] = p2s[(f0 + 0) * 33 + lane] > 0.f ? a0 * scale2d : 0.f;
            out[(f0 + 1) * TILE] = p2s[(f0 + 1) * 33 + lane] > 0.f ? a1 * scale2d : 0.f;
            out[(f0 + 2) * TILE] = p2s[(f0 + 2) * 33 + lane] > 0.f ? a2 * scale2d : 0.f;
            out[(f0 + 3) * TILE] = p2s[(f0 + 3) * 33 + lane] > 0.f ? a3 * scale2d : 0.f;
        }
    }
    pdl_trigger();
    // phase C: fc1.weight partial; warp = (feature half fg, block of 10 hidden units jg), lane == f
    float* part = part3 + tile * (int64_t)PART3;
    {
        const int fg = warp & 1, jg = warp >> 1;
        const int fl = fg * 32 + lane;
        float acc[10];
#pragma unroll
        for (int j = 0; j < 10; ++j) acc[j] = 0.f;
#pragma unroll 4
        for (int s = 0; s < 32; ++s) {
            const float a = p2s[fl * 33 + s];
            const float2* dz = reinterpret_cast<const float2*>(dz1t + s * 52 + jg * 10);
#pragma unroll
            for (int q = 0; q < 5; ++q) {
                const float2 t = dz[q];
                acc[q * 2 + 0] = fmaf(t.x, a, acc[q * 2 + 0]);
                acc[q * 2 + 1] = fmaf(t.y, a, acc[q * 2 + 1]);
            }
        }
#pragma unroll
        for (int j = 0; j < 10; ++j) part[(jg * 10 + j) * 320 + fbase + fl] = acc[j];
    }
    // phase D (slice 0 only): fc1.bias, fc2.weight, fc2.bias partials
    if (r == 0) {
        for (int e = threadIdx.x; e < 560; e += HB_THREADS) {
            float s = 0.f;
            if (e < 50) {
                for (int k = 0; k < 32; ++k) s += dz1t[k * 52 + e];
                part[16000 + e] = s;
            } else if (e < 550) {
                const int c = (e - 50) / 50, j = (e - 50) % 50;
                for (int k = 0; k < 32; ++k) s = fmaf(dz2s[c * 33 + k], hds[j * 33 + k], s);
                part[16050 + (e - 50)] = s;
            } else {
                const int c = e - 550;
                for (int k = 0; k < 32; ++k) s += dz2s[c * 33 + k];
                part[16550 + c] = s;
            }
        }
    }
}

// Recover the dense conv2 output gradient of one pooled window from its compact form: the
// gradient g reaching the pooled value and the 2-bit argmax code k (row-major in the window).
__device__ __forceinline__ float sel(float g, int k, int pos) { return k == pos ? g : 0.f; }

// ------------------------------------------------------------------------------------------
// conv2 data gradient.  CTA = (tile, 6x6 quadrant of the 12x12 input gradient); warp = input
// channel, 36 accumulators; the dense 6x6x20 output-gradient window is decoded once into
// shared memory (92 KB).  Only the (a,ky),(b,kx) pairs that land inside the quadrant are
// issued, so the FMA count is exactly the useful 320,000 per sample.
// ------------------------------------------------------------------------------------------
constexpr int DG_THREADS = 320;
constexpr int DG_DY = 20 * 36 * TILE;  // floats
constexpr int DG_W = 10 * 20 * 28;
constexpr int DG_SMEM_BYTES = (DG_DY + DG_W) * 4;

// KEEP: the masked 6x6 gradient of channel `warp` is not written to dq1 but left in this warp's private
// shared-memory patch gs [36][32] for the conv1 weight gradient that follows in the same CTA.
template <int QY, int QX, bool KEEP>
__device__ __forceinline__ void conv2_dgrad_body(const float* __restrict__ dq2, const uint8_t* __restrict__ idx2,
                                                 const float* __restrict__ p1, float* __restrict__ dq1,
                                                 float* dys, const float* wsd, int64_t tile, int warp, int lane,
                                                 float* gs = nullptr) {
    // decode: warp handles channels 2*warp, 2*warp+1; windows py in [QY,QY+3), px in [QX,QX+3)
    const float* g_base = dq2 + tile * (int64_t)(P2F * TILE) + lane;
    const uint8_t* i_base = idx2 + tile * (int64_t)(PDT_IDX2_BYTES * TILE) + lane;
#pragma unroll
    for (int cc = 0; cc < 2; ++cc) {
        const int co = 2 * warp + cc;
#pragma unroll
        for (int wy = 0; wy < 3; ++wy)
#pragma unroll
            for (int wx = 0; wx < 3; ++wx) {
                const int py = QY + wy, px = QX + wx;
                const float g = g_base[(co * 16 + py * 4 + px) * TILE];
                const int k = i_base[(co * 16 + py * 4 + px) * TILE] & 3;
                // dys is [co][9][32][4]: the 36 positions of a channel are nine float4 per lane (lane4 order), so
                // the main loop reads them with 9 wide loads instead of 36
                const int p0 = (2 * wy) * 6 + 2 * wx;
                float* d = dys + co * 36 * TILE + lane * 4;
                d[((p0 + 0) >> 2) * (TILE * 4) + ((p0 + 0) & 3)] = sel(g, k, 0);
                d[((p0 + 1) >> 2) * (TILE * 4) + ((p0 + 1) & 3)] = sel(g, k, 1);
                d[((p0 + 6) >> 2) * (TILE * 4) + ((p0 + 6) & 3)] = sel(g, k, 2);
                d[((p0 + 7) >> 2) * (TILE * 4) + ((p0 + 7) & 3)] = sel(g, k, 3);
            }
    }
    __syncthreads();
    const int ci = warp;
    float acc[6][6];
#pragma unroll
    for (int r = 0; r < 6; ++r)
#pragma unroll
        for (int c = 0; c < 6; ++c) acc[r][c] = 0.f;
#pragma unroll 1
    for (int co = 0; co < 20; ++co) {
        float wv[28];
        const float4* wp = reinterpret_cast<const float4*>(wsd + (ci * 20 + co) * 28);
#pragma unroll
        for (int q = 0; q < 7; ++q) {
            const float4 t = wp[q];
            wv[q * 4 + 0] = t.x; wv[q * 4 + 1] = t.y; wv[q * 4 + 2] = t.z; wv[q * 4 + 3] = t.w;
        }
        float d[6][6];
        const float4* dp = reinterpret_cast<const float4*>(dys + co * 36 * TILE) + lane;
#pragma unroll
        for (int q = 0; q < 9; ++q) {
            const float4 t = dp[q * TILE];
            d[(4 * q + 0) / 6][(4 * q + 0) % 6] = t.x;
            d[(4 * q + 1) / 6][(4 * q + 1) % 6] = t.y;
            d[(4 * q + 2) / 6][(4 * q + 2) % 6] = t.z;
            d[(4 * q + 3) / 6][(4 * q + 3) % 6] = t.w;
        }
#pragma unroll
        for (int a = 0; a < 6; ++a)
#pragma unroll
            for (int ky = 0; ky < 5; ++ky) {
                const int lr = a + ky - 4 * QY;
                if (lr < 0 || lr > 5) continue;
#pragma unroll
                for (int b = 0; b < 6; ++b)
#pragma unroll
                    for (int kx = 0; kx < 5; ++kx) {
                        const int lc = b + kx - 4 * QX;
                        if (lc < 0 || lc > 5) continue;
                        acc[lr][lc] = fmaf(d[a][b], wv[ky * 5 + kx], acc[lr][lc]);
                    }
            }
    }
    if (!KEEP) pdl_trigger();   // main loop done
    const float* pm4 = p1 + tile * (int64_t)(P1F * TILE);   // lane4 layout
    float* out = KEEP ? gs + lane : dq1 + tile * (int64_t)(P1F * TILE) + lane;
#pragma unroll
    for (int r = 0; r < 6; ++r)
#pragma unroll
        for (int c = 0; c < 6; ++c) {
            const int f = ci * 144 + (6 * QY + r) * 12 + 6 * QX + c;
            const float v = pm4[lane4(f, lane)] > 0.f ? acc[r][c] : 0.f;
            if (KEEP) out[(r * 6 + c) * TILE] = v;
            else out[f * TILE] = v;
        }
}

__global__ void __launch_bounds__(DG_THREADS)
conv2_bwd_data_kernel(const float* __restrict__ dq2, const uint8_t* __restrict__ idx2, const float* __restrict__ w,
                      const float* __restrict__ p1, float* __restrict__ dq1) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float* dys = reinterpret_cast<float*>(smem_raw);
    float* wsd = dys + DG_DY;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t tile = blockIdx.x >> 2;
    const int q = blockIdx.x & 3;
    {
        float tmp[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const int e = threadIdx.x + i * DG_THREADS;
            tmp[i] = e < 5000 ? __ldg(w + e) : 0.f;
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const int e = threadIdx.x + i * DG_THREADS;
            if (e < 5000) {
                const int co = e / 250, rem = e % 250, ci = rem / 25, k = rem % 25;
                wsd[(ci * 20 + co) * 28 + k] = tmp[i];
            }
        }
    }
    pdl_wait();       // dq2 is the previous kernel's output; the weights above are older
    switch (q) {
        case 0: conv2_dgrad_body<0, 0, false>(dq2, idx2, p1, dq1, dys, wsd, tile, warp, lane); break;
        case 1: conv2_dgrad_body<0, 1, false>(dq2, idx2, p1, dq1, dys, wsd, tile, warp, lane); break;
        case 2: conv2_dgrad_body<1, 0, false>(dq2, idx2, p1, dq1, dys, wsd, tile, warp, lane); break;
        default: conv2_dgrad_body<1, 1, false>(dq2, idx2, p1, dq1, dys, wsd, tile, warp, lane); break;
    }
}

// ------------------------------------------------------------------------------------------
// conv2 weight gradient.  CTA = (tile, input channel ci): the 12x12 input plane of 32 samples
// arrives by one bulk TMA copy (18 KB); warp = pair of output channels, 50 accumulators.
// ------------------------------------------------------------------------------------------
constexpr int WG2_THREADS = 320;

__global__ void __launch_bounds__(WG2_THREADS, 2)
conv2_bwd_weight_kernel(const float* __restrict__ dq2, const uint8_t* __restrict__ idx2,
                        const float* __restrict__ p1, float* __restrict__ part2) {
    __shared__ __align__(128) float slab[144 * TILE];
    __shared__ __align__(8) uint64_t bar;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t tile = blockIdx.x / 10;
    const int ci = blockIdx.x % 10;
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(&bar, 144 * TILE * 4);
        bulk_g2s(slab, p1 + (tile * (int64_t)P1F + ci * 144) * TILE, 144 * TILE * 4, &bar);
    }
    const float* g_base = dq2 + tile * (int64_t)(P2F * TILE) + lane;
    const uint8_t* i_base = idx2 + tile * (int64_t)(PDT_IDX2_BYTES * TILE) + lane;
    float acc[2][25];
    float bsum[2] = {0.f, 0.f};
#pragma unroll
    for (int cc = 0; cc < 2; ++cc)
#pragma unroll
        for (int k = 0; k < 25; ++k) acc[cc][k] = 0.f;
    float gn[2][4];
    unsigned bn[2][4];
    auto fetch = [&](int py) {
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
            const int co = 2 * warp + cc;
#pragma unroll
            for (int px = 0; px < 4; ++px) {
                gn[cc][px] = g_base[(co * 16 + py * 4 + px) * TILE];
                bn[cc][px] = i_base[(co * 16 + py * 4 + px) * TILE];
            }
        }
    };
    fetch(0);
    mbar_wait(&bar, 0);
#pragma unroll 1
    for (int py = 0; py < 4; ++py) {
        float gc[2][4];
        unsigned bc[2][4];
#pragma unroll
        for (int cc = 0; cc < 2; ++cc)
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                gc[cc][q] = gn[cc][q];
                bc[cc][q] = bn[cc][q];
            }
        if (py + 1 < 4) fetch(py + 1);
        float P[6][12];
#pragma unroll
        for (int r = 0; r < 6; ++r)
#pragma unroll
            for (int q = 0; q < 3; ++q) {   // lane4 layout: a 12-column row is three 128-bit loads
                const float4 t = reinterpret_cast<const float4*>(slab)[((2 * py + r) * 3 + q) * TILE + lane];
                P[r][q * 4 + 0] = t.x; P[r][q * 4 + 1] = t.y; P[r][q * 4 + 2] = t.z; P[r][q * 4 + 3] = t.w;
            }
        float dy[2][2][8];
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
#pragma unroll
            for (int px = 0; px < 4; ++px) {
                const float g = gc[cc][px];
                const int k = bc[cc][px] & 3;
                bsum[cc] += g;
                dy[cc][0][2 * px] = sel(g, k, 0);
                dy[cc][0][2 * px + 1] = sel(g, k, 1);
                dy[cc][1][2 * px] = sel(g, k, 2);
                dy[cc][1][2 * px + 1] = sel(g, k, 3);
            }
        }
#pragma unroll
        for (int ky = 0; ky < 5; ++ky)
#pragma unroll
            for (int kx = 0; kx < 5; ++kx)
#pragma unroll
                for (int cc = 0; cc < 2; ++cc)
#pragma unroll
                    for (int a = 0; a < 2; ++a)
#pragma unroll
                        for (int ox = 0; ox < 8; ++ox)
                            acc[cc][ky * 5 + kx] = fmaf(dy[cc][a][ox], P[a + ky][ox + kx], acc[cc][ky * 5 + kx]);
    }
    float* part = part2 + tile * (int64_t)PART2;
#pragma unroll
    for (int cc = 0; cc < 2; ++cc) {
        const int co = 2 * warp + cc;
        float mine = 0.f;
#pragma unroll
        for (int k = 0; k < 25; ++k) {
            const float s = warp_sum(acc[cc][k]);
            mine = lane == k ? s : mine;
        }
        if (lane < 25) part[co * 250 + ci * 25 + lane] = mine;
        if (ci == 0) {
            const float bs = warp_sum(bsum[cc]);
            if (lane == 0) part[5000 + co] = bs;
        }
    }
}

// ------------------------------------------------------------------------------------------
// conv1 weight gradient.  CTA = (tile, band of 4 pre-pool rows): the 8x28 input band of 32
// samples is staged transposed in shared memory; warp = pair of channels, 50 accumulators.
// ------------------------------------------------------------------------------------------
constexpr int WG1_THREADS = 160;
constexpr int WG1_RS = 40;

template <bool U8>
__global__ void __launch_bounds__(WG1_THREADS)
conv1_bwd_weight_kernel(const float* __restrict__ dq1, const uint8_t* __restrict__ idx1,
                        const void* __restrict__ x, x_norm nm, float* __restrict__ part1, int64_t n) {
    __shared__ float band[8 * WG1_RS * 33];
    __shared__ float xlut[U8 ? 256 : 1];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t tile = blockIdx.x / 6;
    const int by = blockIdx.x % 6;
    if (U8) {
        fill_xlut(xlut, nm);
        __syncthreads();
    }
    for (int c4 = warp; c4 < 7; c4 += WG1_THREADS / 32)
        load_x_window<1, WG1_RS, U8>(band + c4 * 4 * 33, x, xlut, n, tile * TILE, 4 * by, c4 * 4, lane);
    pdl_wait();       // dq1 is the previous kernel's output; the input batch staged above is not
    __syncthreads();
    const float* g_base = dq1 + tile * (int64_t)(P1F * TILE) + lane;
    const uint8_t* i_base = idx1 + tile * (int64_t)(PDT_IDX1_BYTES * TILE) + lane;
    float acc[2][25];
    float bsum[2] = {0.f, 0.f};
#pragma unroll
    for (int cc = 0; cc < 2; ++cc)
#pragma unroll
        for (int k = 0; k < 25; ++k) acc[cc][k] = 0.f;
    // the compact gradient (4 values + 1 code byte per channel and block) is fetched one block ahead
    float gn[2][4];
    unsigned bn[2];
    auto fetch = [&](int bx) {
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
            const int c = 2 * warp + cc;
            bn[cc] = i_base[(c * 36 + by * 6 + bx) * TILE];
#pragma unroll
            for (int py = 0; py < 2; ++py)
#pragma unroll
                for (int px = 0; px < 2; ++px)
                    gn[cc][py * 2 + px] = g_base[(c * 144 + (2 * by + py) * 12 + 2 * bx + px) * TILE];
        }
    };
    fetch(0);
#pragma unroll 1
    for (int bx = 0; bx < 6; ++bx) {
        float gc[2][4];
        unsigned bc[2];
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
            bc[cc] = bn[cc];
#pragma unroll
            for (int q = 0; q < 4; ++q) gc[cc][q] = gn[cc][q];
        }
        if (bx + 1 < 6) fetch(bx + 1);
        float in[64];
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int c = 0; c < 8; ++c) in[r * 8 + c] = band[(r * WG1_RS + 4 * bx + c) * 33 + lane];
        float dy[2][16];
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
#pragma unroll
            for (int py = 0; py < 2; ++py)
#pragma unroll
                for (int px = 0; px < 2; ++px) {
                    const float g = gc[cc][py * 2 + px];
                    const int k = (bc[cc] >> (2 * (py * 2 + px))) & 3;
                    bsum[cc] += g;
                    dy[cc][(2 * py) * 4 + 2 * px] = sel(g, k, 0);
                    dy[cc][(2 * py) * 4 + 2 * px + 1] = sel(g, k, 1);
                    dy[cc][(2 * py + 1) * 4 + 2 * px] = sel(g, k, 2);
                    dy[cc][(2 * py + 1) * 4 + 2 * px + 1] = sel(g, k, 3);
                }
        }
#pragma unroll
        for (int ky = 0; ky < 5; ++ky)
#pragma unroll
            for (int kx = 0; kx < 5; ++kx)
#pragma unroll
                for (int cc = 0; cc < 2; ++cc)
#pragma unroll
                    for (int oy = 0; oy < 4; ++oy)
#pragma unroll
                        for (int ox = 0; ox < 4; ++ox)
                            acc[cc][ky * 5 + kx] =
                                fmaf(dy[cc][oy * 4 + ox], in[(oy + ky) * 8 + ox + kx], acc[cc][ky * 5 + kx]);
    }
    pdl_trigger();
    float* part = part1 + (tile * 6 + by) * (int64_t)PART1;
#pragma unroll
    for (int cc = 0; cc < 2; ++cc) {
        const int c = 2 * warp + cc;
        float mine = 0.f;
#pragma unroll
        for (int k = 0; k < 25; ++k) {
            const float s = warp_sum(acc[cc][k]);
            mine = lane == k ? s : mine;
        }
        if (lane < 25) part[c * 25 + lane] = mine;
        const float bs = warp_sum(bsum[cc]);
        if (lane == 0) part[250 + c] = bs;
    }
}

// ------------------------------------------------------------------------------------------
// conv2 data gradient + conv1 weight gradient in ONE kernel.  The data-gradient CTA of (tile, 6x6
// quadrant) ends with warp c holding channel c's masked gradient for its 32 samples: exactly what
// conv1's weight gradient of channel c consumes.  So instead of writing dq1 (5.8 KB/sample) and
// launching a second kernel that reads it back, each warp goes on to correlate its patch with the
// 16x16 window of the input image under the quadrant (staged at kernel start, before the
// dependency wait: the batch is older than the previous kernel).  Partials land in slots 0..3 of the
// tile's six conv1 slots; slots 4 and 5 are zeroed.
// ------------------------------------------------------------------------------------------
constexpr int DW_XS = 16 * 4 * TILE * 4;     // floats: x window [16 rows][4 column quads][32 lanes][4]
constexpr int DW_GS = 10 * 36 * TILE;        // floats: per-warp gradient patch [36][32]
constexpr int DW_SMEM_BYTES = (DG_DY + DG_W + DW_XS + DW_GS) * 4;

template <bool U8>
__global__ void __launch_bounds__(DG_THREADS)
conv2_bwd_data_wgrad1_kernel(const float* __restrict__ dq2, const uint8_t* __restrict__ idx2,
                             const float* __restrict__ w, const float* __restrict__ p1,
                             const void* __restrict__ x, x_norm nm, const uint8_t* __restrict__ idx1,
                             float* __restrict__ part1, int64_t n) {
    __shared__ float xlut[U8 ? 256 : 1];
    if (U8) {
        fill_xlut(xlut, nm);
        __syncthreads();
    }
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float* dys = reinterpret_cast<float*>(smem_raw);
    float* wsd = dys + DG_DY;
    float4* xs4 = reinterpret_cast<float4*>(wsd + DG_W);
    float* gs = wsd + DG_W + DW_XS;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t tile = blockIdx.x >> 2;
    const int q = blockIdx.x & 3;
    const int qy = q >> 1, qx = q & 1;
    {
        float tmp[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const int e = threadIdx.x + i * DG_THREADS;
            tmp[i] = e < 5000 ? __ldg(w + e) : 0.f;
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const int e = threadIdx.x + i * DG_THREADS;
            if (e < 5000) {
                const int co = e / 250, rem = e % 250, ci = rem / 25, k = rem % 25;
                wsd[(ci * 20 + co) * 28 + k] = tmp[i];
            }
        }
    }
    {   // input window: image rows 12qy .. 12qy+15, columns 12qx .. 12qx+15; lane == sample, one float4 per store
        const int64_t sample = tile * TILE + lane;
        const int64_t src = sample * 784 + (12 * qy) * 28 + 12 * qx;
        for (int i = warp; i < 64; i += DG_THREADS / 32) {
            const int row = i >> 2, c4 = i & 3;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (sample < n) v = load_x4<U8>(x, src + row * 28 + 4 * c4, xlut);
            xs4[(row * 4 + c4) * TILE + lane] = v;
        }
    }
    const int c = warp;   // conv1 channel of this warp == conv2 input channel
    const uint8_t* i1 = idx1 + tile * (int64_t)(PDT_IDX1_BYTES * TILE) + (c * 36 + 3 * qy * 6 + 3 * qx) * TILE + lane;
    unsigned cnext = i1[0];
    float* g_own = gs + warp * 36 * TILE;
    pdl_wait();       // dq2 is the previous kernel's output; weights, input batch and pool codes are older
    switch (q) {
        case 0: conv2_dgrad_body<0, 0, true>(dq2, idx2, p1, nullptr, dys, wsd, tile, warp, lane, g_own); break;
        case 1: conv2_dgrad_body<0, 1, true>(dq2, idx2, p1, nullptr, dys, wsd, tile, warp, lane, g_own); break;
        case 2: conv2_dgrad_body<1, 0, true>(dq2, idx2, p1, nullptr, dys, wsd, tile, warp, lane, g_own); break;
        default: conv2_dgrad_body<1, 1, true>(dq2, idx2, p1, nullptr, dys, wsd, tile, warp, lane, g_own); break;
    }
    pdl_trigger();
    // conv1 weight gradient of channel c over the quadrant: nine 2x2 blocks of pooled positions, each an
    // 8x8 input window (16 x 128-bit shared loads) against the decoded 4x4 gradient; 25 accumulators
    float acc[25];
    float bsum = 0.f;
#pragma unroll
    for (int k = 0; k < 25; ++k) acc[k] = 0.f;
    const float* g = g_own + lane;
#pragma unroll 1
    for (int blk = 0; blk < 9; ++blk) {
        const int by = blk / 3, bx = blk - 3 * by;
        const unsigned code = cnext;
        if (blk + 1 < 9) {
            const int nb = blk + 1, nby = nb / 3, nbx = nb - 3 * nby;
            cnext = i1[(nby * 6 + nbx) * TILE];
        }
        float in[64];
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const float4 t = xs4[((4 * by + r) * 4 + bx + h) * TILE + lane];
                in[r * 8 + 4 * h + 0] = t.x; in[r * 8 + 4 * h + 1] = t.y;
                in[r * 8 + 4 * h + 2] = t.z; in[r * 8 + 4 * h + 3] = t.w;
            }
        float dy[16];
#pragma unroll
        for (int py = 0; py < 2; ++py)
#pragma unroll
            for (int px = 0; px < 2; ++px) {
                const float gv = g[((2 * by + py) * 6 + 2 * bx + px) * TILE];
                const int k = (code >> (2 * (py * 2 + px))) & 3;
                bsum += gv;
                dy[(2 * py) * 4 + 2 * px] = sel(gv, k, 0);
                dy[(2 * py) * 4 + 2 * px + 1] = sel(gv, k, 1);
                dy[(2 * py + 1) * 4 + 2 * px] = sel(gv, k, 2);
                dy[(2 * py + 1) * 4 + 2 * px + 1] = sel(gv, k, 3);
            }
#pragma unroll
        for (int ky = 0; ky < 5; ++ky)
#pragma unroll
            for (int kx = 0; kx < 5; ++kx)
#pragma unroll
                for (int oy = 0; oy < 4; ++oy)
#pragma unroll
                    for (int ox = 0; ox < 4; ++ox)
                        acc[ky * 5 + kx] = fmaf(dy[oy * 4 + ox], in[(oy + ky) * 8 + ox + kx], acc[ky * 5 + kx]);
    }
    float* part = part1 + (tile * 6 + q) * (int64_t)PART1;
    float mine = 0.f;
#pragma unroll
    for (int k = 0; k < 25; ++k) {
        const float sk = warp_sum(acc[k]);
        mine = lane == k ? sk : mine;
    }
    if (lane < 25) part[c * 25 + lane] = mine;
    const float bs = warp_sum(bsum);
    if (lane == 0) part[250 + c] = bs;
    if (q == 0)      // the slots of the band decomposition that the quadrant decomposition does not use
        for (int e = threadIdx.x; e < 2 * PART1; e += DG_THREADS) part1[(tile * 6 + 4) * (int64_t)PART1 + e] = 0.f;
}

// ------------------------------------------------------------------------------------------
// Fixed-order reduction of the per-tile partials into the flat gradient vector (+ tail).
// ------------------------------------------------------------------------------------------
// block = 32 consecutive gradient elements x 8 slot lanes: every warp reads one coalesced 128-byte
// row per slot, slot lane j sums slots j, j+8, ... (independent loads in flight), and the 8 lane
// sums are folded in a fixed order, so the result does not depend on scheduling.
__global__ void __launch_bounds__(256)
grad_reduce_kernel(const float* __restrict__ scratch, int64_t T, float beta, float scale,
                   const pdt_stats* __restrict__ stats, float* __restrict__ grad, const float* __restrict__ gscale) {
    __shared__ float red[8][33];
    const int ex = threadIdx.x & 31, sl = threadIdx.x >> 5;
    const int i = blockIdx.x * 32 + ex;
    float s0 = 0.f, s1 = 0.f;
    if (i < PDT_NUM_PARAMS) {
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
        for (; k + 8 < slots; k += 16) {
            s0 += base[k * stride];
            s1 += base[(k + 8) * stride];
        }
        if (k < slots) s0 += base[k * stride];
    }
    red[sl][ex] = s0 + s1;
    __syncthreads();
    if (sl == 0 && i < PDT_NUM_PARAMS) {
        float s = ((red[0][ex] + red[1][ex]) + (red[2][ex] + red[3][ex])) +
                  ((red[4][ex] + red[5][ex]) + (red[6][ex] + red[7][ex]));
        s *= gscale != nullptr ? scale * *gscale : scale;
        grad[i] = beta != 0.f ? fmaf(beta, grad[i], s) : s;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0 && stats != nullptr) {
        grad[PDT_NUM_PARAMS + 0] = stats->loss_mean * scale;
        grad[PDT_NUM_PARAMS + 1] = (float)stats->top1;
        grad[PDT_NUM_PARAMS + 2] = (float)stats->topk;
        grad[PDT_NUM_PARAMS + 3] = (float)stats->n_valid;
    }
}

}  // namespace pdt

using namespace pdt;

extern "C" int64_t pdt_mnist_scratch_bytes(int64_t n) { return num_tiles(n) * (int64_t)PART_PER_TILE * 4; }

extern "C" int pdt_head_bwd(const float* dlogp, const float* logp, const float* hd_lane, const float* p2_lane,
                            const float* fc1w, const float* fc2w, float scale1d, float scale2d, float* dq2_lane,
                            void* scratch, int64_t n, void* stream) {
    PDT_REQUIRE(n > 0 && dlogp && logp && hd_lane && p2_lane && fc1w && fc2w && dq2_lane && scratch,
                "null pointer or n <= 0");
    int rc = ensure_smem<0>(head_bwd_kernel<false>, HB_SMEM_BYTES);
    if (rc) return rc;
    const int64_t T = num_tiles(n);
    launch_pdl(6, head_bwd_kernel<false>, dim3(HB_SPLIT, (unsigned)T), dim3(HB_THREADS), HB_SMEM_BYTES, stream,
               dlogp, logp, hd_lane, p2_lane, fc1w, fc2w, scale1d, scale2d, dq2_lane,
               reinterpret_cast<float*>(scratch) + part3_off(T), n, (const int64_t*)nullptr,
               (const pdt_stats*)nullptr, (const float*)nullptr, (int64_t)0);
    return check_launch("head_bwd");
}

extern "C" int pdt_head_bwd_nll(const int64_t* target, const pdt_stats* stats, const float* grad_loss,
                                int64_t ignore_index, const float* logp, const float* hd_lane, const float* p2_lane,
                                const float* fc1w, const float* fc2w, float scale1d, float scale2d, float* dq2_lane,
                                void* scratch, int64_t n, void* stream) {
    PDT_REQUIRE(n > 0 && target && stats && logp && hd_lane && p2_lane && fc1w && fc2w && dq2_lane && scratch,
                "null pointer or n <= 0");
    int rc = ensure_smem<1>(head_bwd_kernel<true>, HB_SMEM_BYTES);
    if (rc) return rc;
    const int64_t T = num_tiles(n);
    launch_pdl(6, head_bwd_kernel<true>, dim3(HB_SPLIT, (unsigned)T), dim3(HB_THREADS), HB_SMEM_BYTES, stream,
               (const float*)nullptr, logp, hd_lane, p2_lane, fc1w, fc2w, scale1d, scale2d, dq2_lane,
               reinterpret_cast<float*>(scratch) + part3_off(T), n, target, stats, grad_loss, ignore_index);
    return check_launch("head_bwd_nll");
}

extern "C" int pdt_conv2_bwd_data(const float* dq2_lane, const uint8_t* idx2_lane, const float* w,
                                  const float* p1_lane, float* dq1_lane, int64_t n, void* stream) {
    PDT_REQUIRE(n > 0 && dq2_lane && idx2_lane && w && p1_lane && dq1_lane, "null pointer or n <= 0");
    int rc = ensure_smem<2>(conv2_bwd_data_kernel, DG_SMEM_BYTES);
    if (rc) return rc;
    launch_pdl(7, conv2_bwd_data_kernel, dim3((unsigned)(num_tiles(n) * 4)), dim3(DG_THREADS), DG_SMEM_BYTES, stream,
               dq2_lane, idx2_lane, w, p1_lane, dq1_lane);
    return check_launch("conv2_bwd_data");
}

template <bool U8>
static int dgrad_wgrad1_launch(const float* dq2_lane, const uint8_t* idx2_lane, const float* w, const float* p1_lane,
                               const void* x, x_norm nm, const uint8_t* idx1_lane, void* scratch, int64_t n, void* stream) {
    int rc = ensure_smem<U8 ? 31 : 3>(conv2_bwd_data_wgrad1_kernel<U8>, DW_SMEM_BYTES);
    if (rc) return rc;
    const int64_t T = num_tiles(n);
    launch_pdl(7, conv2_bwd_data_wgrad1_kernel<U8>, dim3((unsigned)(T * 4)), dim3(DG_THREADS), DW_SMEM_BYTES, stream,
               dq2_lane, idx2_lane, w, p1_lane, x, nm, idx1_lane, reinterpret_cast<float*>(scratch) + part1_off(T), n);
    return check_launch("conv2_bwd_data_conv1_bwd_weight");
}

extern "C" int pdt_conv2_bwd_data_conv1_bwd_weight(const float* dq2_lane, const uint8_t* idx2_lane, const float* w,
                                                   const float* p1_lane, const float* x, const uint8_t* idx1_lane,
                                                   void* scratch, int64_t n, void* stream) {
    PDT_REQUIRE(n > 0 && dq2_lane && idx2_lane && w && p1_lane && x && idx1_lane && scratch, "null pointer or n <= 0");
    return dgrad_wgrad1_launch<false>(dq2_lane, idx2_lane, w, p1_lane, x, x_norm{0.f, 1.f}, idx1_lane, scratch, n, stream);
}

extern "C" int pdt_conv2_bwd_data_conv1_bwd_weight_u8(const float* dq2_lane, const uint8_t* idx2_lane, const float* w,
                                                      const float* p1_lane, const uint8_t* x, float mean, float std,
                                                      const uint8_t* idx1_lane, void* scratch, int64_t n, void* stream) {
    PDT_REQUIRE(n > 0 && dq2_lane && idx2_lane && w && p1_lane && x && idx1_lane && scratch, "null pointer or n <= 0");
    PDT_REQUIRE(std != 0.f, "std must not be zero");
    return dgrad_wgrad1_launch<true>(dq2_lane, idx2_lane, w, p1_lane, x, x_norm{mean, std}, idx1_lane, scratch, n, stream);
}

extern "C" int pdt_conv2_bwd_weight(const float* dq2_lane, const uint8_t* idx2_lane, const float* p1_lane,
                                    void* scratch, int64_t n, void* stream) {
    PDT_REQUIRE(n > 0 && dq2_lane && idx2_lane && p1_lane && scratch, "null pointer or n <= 0");
    const int64_t T = num_tiles(n);
    prefer_carveout(conv2_bwd_weight_kernel);
    conv2_bwd_weight_kernel<<<(unsigned)(T * 10), WG2_THREADS, 0, (cudaStream_t)stream>>>(
        dq2_lane, idx2_lane, p1_lane, reinterpret_cast<float*>(scratch) + part2_off(T));
    return check_launch("conv2_bwd_weight");
}

extern "C" int pdt_conv1_bwd_weight(const float* dq1_lane, const uint8_t* idx1_lane, const float* x, void* scratch,
                                    int64_t n, void* stream) {
    PDT_REQUIRE(n > 0 && dq1_lane && idx1_lane && x && scratch, "null pointer or n <= 0");
    const int64_t T = num_tiles(n);
    launch_pdl(8, conv1_bwd_weight_kernel<false>, dim3((unsigned)(T * 6)), dim3(WG1_THREADS), 0, stream,
               dq1_lane, idx1_lane, x, x_norm{0.f, 1.f}, reinterpret_cast<float*>(scratch) + part1_off(T), n);
    return check_launch("conv1_bwd_weight");
}

extern "C" int pdt_conv1_bwd_weight_u8(const float* dq1_lane, const uint8_t* idx1_lane, const uint8_t* x, float mean,
                                       float std, void* scratch, int64_t n, void* stream) {
    PDT_REQUIRE(n > 0 && dq1_lane && idx1_lane && x && scratch, "null pointer or n <= 0");
    PDT_REQUIRE(std != 0.f, "std must not be zero");
    const int64_t T = num_tiles(n);
    launch_pdl(8, conv1_bwd_weight_kernel<true>, dim3((unsigned)(T * 6)), dim3(WG1_THREADS), 0, stream,
               dq1_lane, idx1_lane, x, x_norm{mean, std}, reinterpret_cast<float*>(scratch) + part1_off(T), n);
    return check_launch("conv1_bwd_weight");
}

extern "C" int pdt_grad_reduce(const void* scratch, int64_t n, float beta, float scale, const pdt_stats* stats,
                               float* flat_grad, const float* grad_scale_dev, void* stream) {
    PDT_REQUIRE(n > 0 && scratch && flat_grad, "null pointer or n <= 0");
    prefer_carveout(grad_reduce_kernel);
    grad_reduce_kernel<<<(PDT_NUM_PARAMS + 31) / 32, 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const float*>(scratch), num_tiles(n), beta, scale, stats, flat_grad, grad_scale_dev);
    return check_launch("grad_reduce");
}
