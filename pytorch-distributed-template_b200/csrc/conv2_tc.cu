// conv2 (10 -> 20 channels, 5x5) + Dropout2d + 2x2 max-pool + relu on the 5th-generation tensor cores
// (tcgen05.mma, accumulators in tensor memory), as an implicit GEMM in split-TF32 ("3xTF32") arithmetic.
// Replaces model/model.py:17 of the reference, like conv2_fwd_kernel in fwd.cu (the CUDA-core version).
//
// GEMM view.  M = (sample, output position): one MMA covers TWO samples x 64 positions = 128 rows.  N = 32
// columns (20 output channels; channels 0-9 in columns 0-9, channels 10-19 in columns 16-25, the rest zero).
// K = (input channel, tap) = 10 x 25, walked in steps of 8 = two 16-byte chunks of four input channels each.
//
// No im2col is materialised.  The pair's pooled conv1 output is laid out in shared memory as
//     A[part hi|lo][chunk c = ci/4][iy 0..11][s 0..1][ix 0..11 + one pad][ci%4]   (16 bytes per (.., ix))
// so that, for a tap (ky, kx), the 8 rows (oy, s, ox = 0..7) of one core matrix are 8 CONSECUTIVE 16-byte
// groups starting at (iy = oy + ky, s, ix = kx), the next 8-row group ((oy, s) -> (oy, s+1) -> (oy+1, 0) ...)
// lies a constant 208 bytes further (stride-dimension byte offset), and a tap shift is nothing but a change of
// the descriptor's start address (+416 bytes per ky, +16 per kx).  Channels 8 and 9 share their chunk with the
// NEXT column: plane 2 holds {ci8(ix), ci9(ix), ci8(ix+1), ci9(ix+1)}, so one of its chunks covers the taps
// (ky, kx) and (ky, kx+1) and only kx = 0, 2, 4 are visited.  The two K-chunks of one MMA are either planes 0
// and 1 of one tap (leading-dimension byte offset = one plane, 4,992 bytes), or plane 2 at kx = 0 and kx = 2 of
// one ky (offset 32 bytes), or plane 2 at kx = 4 of two vertically adjacent taps (offset = one input row, 416
// bytes): 25 + 5 + 3 = 33 K-steps for the 250 real products of a window (264 slots).
//
// Precision.  kind::tf32 keeps 10 mantissa bits of each operand.  Both operands are split x = hi + lo with hi
// rounded to TF32 and lo = x - hi (exact in fp32).  B holds [hi | lo] side by side as 64 columns, so ONE pass
// A_hi x [B_hi | B_lo] (N = 64) produces hi*hi in accumulator columns 0-31 and hi*lo in columns 32-63, and a
// second pass A_lo x B_hi (N = 32, the same B buffer) adds lo*hi to columns 0-31; the epilogue adds the two
// column sets.  A_hi -- 4 KB per MMA, the operand that bounds the MMA rate (measured: ~50 cycles per 128x32x8
// MMA from shared memory, i.e. shared-memory bandwidth, not the tensor pipe) -- is read once instead of twice.
// The dropped lo*lo term and the truncation of lo are each ~2^-21 relative per product, below the rounding noise
// of a 250-term fp32 sum; measured against the oracle in tests/test_gpu_conv2_tc.py.
//
// One CTA = 8 consecutive samples of a tile (4 pairs).  Per pair: all 8 warps build A (hi and lo) from p1's lane4
// layout, one thread issues the 66 MMAs and commits them to an mbarrier; the epilogue (bias, Dropout2d scale, 2x2
// max-pool with torch's first-maximum rule, relu, arg-max code) reads the accumulator with tcgen05.ld -- the four
// positions of a pooling window are lanes l, l^1, l^16, l^17 of one warp -- while the next pair's MMAs run on the
// second accumulator and the pair after that is being built.
#include "conv2_tc.cuh"

namespace pdt {

// -DPDT_TC_STAMPS: thread 0 of CTA 0 records clock64() at the phase boundaries (scripts/time_conv2.py prints them)
__device__ long long g_tc_stamps[64];
#ifdef PDT_TC_STAMPS
#define TC_STAMP(i) do { if (blockIdx.x == 0 && threadIdx.x == 0) g_tc_stamps[i] = clock64(); } while (0)
#else
#define TC_STAMP(i) do { } while (0)
#endif


__global__ void __launch_bounds__(TC_THREADS, 1)
conv2_fwd_tc_kernel(const float* __restrict__ p1, const float* __restrict__ w, const float* __restrict__ b,
                    const uint8_t* __restrict__ keep2d, float scale, float* __restrict__ p2,
                    uint8_t* __restrict__ idx2, int64_t n) {
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t* mma_done = reinterpret_cast<uint64_t*>(smem + TC_OFF_BAR);      // [2] accumulator of pair p complete
    uint64_t* ready = mma_done + 2;                                            // [2] operand buffer built, accumulator free
    uint64_t* wbar = mma_done + 4;                                             // raw weights landed
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + TC_OFF_BAR + 48);
    float* stage_p2 = reinterpret_cast<float*>(smem + TC_OFF_P2);
    uint8_t* stage_idx = smem + TC_OFF_IDX;
    float* wraw = reinterpret_cast<float*>(smem + TC_OFF_A + A_BUF_PADDED);    // borrows operand buffer 1 until pair 1 is built
    const int warp = threadIdx.x >> 5;
    const int64_t tile = blockIdx.x >> 2;
    const int q = blockIdx.x & 3;
    const int64_t sample0 = tile * TILE + 8 * q;

    // prologue: nothing here reads data of this step (the weights are older than the previous kernel)
    TC_STAMP(0);
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
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t a_addr = smem_u32(smem + TC_OFF_A), b_addr = smem_u32(smem + TC_OFF_B);

    if (warp == 8) {
        // ---- the issuing warp: lane 0 feeds the tensor core, one pair after the other ------------------------
        if ((threadIdx.x & 31) == 0) {
#pragma unroll 1
            for (int p = 0; p < TC_PAIRS; ++p) {
                mbar_wait_bounded(&ready[p & 1], (p >> 1) & 1);
                tc_fence_after();
                tc_issue(a_addr + (p & 1) * A_BUF_PADDED, b_addr, tmem_base + (p & 1) * TC_PAIR_COLS, &mma_done[p & 1]);
            }
        }
    } else {
        // ---- the eight worker warps ---------------------------------------------------------------------------
        float bias[10];
#pragma unroll
        for (int j = 0; j < 10; ++j) bias[j] = __ldg(b + 10 * (warp >> 2) + j);
        if (threadIdx.x < 96) {      // the zero rows behind each A buffer (read by the padding halves of the last K-steps)
            const int bufi = threadIdx.x / 48, k = threadIdx.x % 48;
            *reinterpret_cast<float4*>(smem + TC_OFF_A + bufi * A_BUF_PADDED + 2 * A_PART + k * 16) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        mbar_wait_bounded(wbar, 0);
        tc_stage_weights(wraw, smem);
        TC_STAMP(1);
        pdl_wait();
        tc_build_a(p1, smem + TC_OFF_A, tile, 8 * q, n);
        fence_async_smem();
        workers_sync();              // B and A(0) complete; the raw weights are no longer needed
        if (threadIdx.x == 0) mbar_arrive(&ready[0]);
        TC_STAMP(2);
#pragma unroll 1
        for (int p = 0; p < TC_PAIRS; ++p) {
            const int nb = (p + 1) & 1;
            float sc[10];
            tc_keep_scale(keep2d, scale, sample0, p, n, sc);
            if (p + 1 < TC_PAIRS) {
                // buffer nb was last read by pair p-1's MMAs (awaited in the previous iteration); accumulator nb was
                // last read by the previous epilogue, complete on every thread before the barrier below
                tc_build_a(p1, smem + TC_OFF_A + nb * A_BUF_PADDED, tile, 8 * q + 2 * (p + 1), n);
                fence_async_smem();
                tc_fence_before();
                workers_sync();
                if (threadIdx.x == 0) mbar_arrive(&ready[nb]);
            } else {
                pdl_trigger();
            }
            TC_STAMP(4 + 5 * p);
            mbar_wait_bounded(&mma_done[p & 1], (p >> 1) & 1);
            tc_fence_after();
            TC_STAMP(7 + 5 * p);
            tc_epilogue(tmem_base + (p & 1) * TC_PAIR_COLS, bias, sc, stage_p2, stage_idx, p);
            TC_STAMP(8 + 5 * p);
        }
        tc_fence_before();
        workers_sync();
        // the CTA's 8 samples of p2 / idx2, feature-major: 32 (8) contiguous bytes per feature
        float* p2t = p2 + tile * (int64_t)(P2F * TILE) + 8 * q;
        uint8_t* i2t = idx2 + tile * (int64_t)(PDT_IDX2_BYTES * TILE) + 8 * q;
        for (int i = threadIdx.x; i < P2F * 2; i += TC_WORKERS) {
            const int f = i >> 1, h = i & 1;
            *reinterpret_cast<float4*>(p2t + f * TILE + 4 * h) = *reinterpret_cast<const float4*>(stage_p2 + f * TC_SAMPLES + 4 * h);
        }
        for (int f = threadIdx.x; f < P2F; f += TC_WORKERS)
            *reinterpret_cast<uint2*>(i2t + f * TILE) = *reinterpret_cast<const uint2*>(stage_idx + f * TC_SAMPLES);
        TC_STAMP(24);
    }
    __syncthreads();
    if (warp == 8) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 2 * TC_PAIR_COLS);
    }
}

}  // namespace pdt

using namespace pdt;

extern "C" int pdt_debug_tc_stamps(long long* host64) {
    return cudaMemcpyFromSymbol(host64, g_tc_stamps, sizeof(long long) * 64) == cudaSuccess ? PDT_OK : check_launch("stamps");
}

extern "C" int pdt_conv2_drop_pool_relu_fwd_tc(const float* p1_lane, const float* w, const float* b,
                                               const uint8_t* keep2d, float scale, float* p2_lane,
                                               uint8_t* idx2_lane, int64_t n, void* stream) {
    PDT_REQUIRE(n > 0 && p1_lane && w && b && p2_lane && idx2_lane, "null pointer or n <= 0");
    if (int rc = ensure_smem<30>(conv2_fwd_tc_kernel, TC_SMEM_BYTES)) return rc;
    const unsigned grid = (unsigned)(num_tiles(n) * 4);
    launch_pdl(2, conv2_fwd_tc_kernel, dim3(grid), dim3(TC_THREADS), TC_SMEM_BYTES, stream, p1_lane, w, b, keep2d, scale,
               p2_lane, idx2_lane, n);
    return check_launch("conv2_fwd_tc");
}
