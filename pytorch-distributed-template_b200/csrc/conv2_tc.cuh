// Device-side building blocks of conv2 on the tensor cores (tcgen05.mma, split TF32): shared by the stand-alone forward
// kernel (conv2_tc.cu) and the fused training kernel (fwd.cu).  The design is described at the top of conv2_tc.cu.
#pragma once
#include "common.cuh"

namespace pdt {


// ---- tcgen05 / TMEM primitives (inline PTX; the SASS shows UTCHMMA / LDTM / UTCBAR) --------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// generic-proxy shared-memory writes -> visible to the tensor core's (async proxy) operand reads
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem], 128 x N x 8, TF32 inputs, fp32 accumulate; issued by ONE thread
__device__ __forceinline__ void mma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// all MMAs issued so far by this thread arrive (once) on the mbarrier when they have completed
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
// 16 consecutive fp32 columns of this thread's TMEM lane (lane = 32 * (warp % 4) + laneid)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// mbarrier wait that cannot hang the GPU: a commit that never arrives (a malformed descriptor) traps after ~1 s
__device__ __forceinline__ void mbar_wait_bounded(uint64_t* bar, uint32_t parity) {
    const long long t0 = clock64();
    for (;;) {
        uint32_t done;
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
        if (done) return;
        if (clock64() - t0 > 2000000000LL) __trap();
    }
}

// shared-memory matrix descriptor, K-major, no swizzle: 8 rows x 16 bytes per core matrix (rows 16 bytes apart),
// `sbo` bytes between 8-row groups, `lbo` bytes between the two 16-byte K-chunks of one MMA
__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((smem_addr >> 4) & 0x3fffu) | ((uint64_t)((lbo >> 4) & 0x3fffu) << 16) |
           ((uint64_t)((sbo >> 4) & 0x3fffu) << 32) | (1ull << 46);      // bits 46-47: descriptor version 1 (sm_100)
}
// instruction descriptor: D fp32, A and B TF32, both K-major, N columns, M = 128 rows
__host__ __device__ constexpr uint32_t umma_idesc_tf32(int n_cols) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n_cols >> 3) << 17) | ((128u >> 4) << 24);
}

// x = hi + lo, hi = x rounded to TF32 (10 mantissa bits, ties away from zero), lo exact
__device__ __forceinline__ void split_tf32(float x, float& hi, float& lo) {
    hi = __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
    lo = x - hi;
}

// ---- geometry ------------------------------------------------------------------------------------------------
constexpr int TC_WORKERS = 256;                        // 8 warps build the operand and run the epilogue
constexpr int TC_THREADS = TC_WORKERS + 32;           // + one warp whose lane 0 issues the MMAs
constexpr int TC_SAMPLES = 8;                         // per CTA: lanes 8q .. 8q+7 of a tile
constexpr int TC_PAIRS = TC_SAMPLES / 2;
constexpr int TC_KSTEPS = 33;                         // MMAs per pass
constexpr int TC_NCOLS = 32;                          // columns of one of B's halves (hi or lo)
constexpr int TC_ACC_COLS = 2 * TC_NCOLS;             // accumulator columns of one group: [hi*hi + lo*hi | hi*lo]
#ifndef TC_GROUPS
#define TC_GROUPS 4
#endif
// The K-steps of a pair are dealt to TC_GROUPS separate accumulators that the epilogue adds up in fp32
// (round-to-nearest): the tensor core's own accumulation rounds toward zero, a bias that grows with the length of
// one accumulation chain.  Measured at batch 1024, distance of conv2's pooled output from its fp64 value relative
// to its maximum: one chain of 66 MMAs 2.4e-6, two chains 1.1e-6, four chains 7.5e-7 -- the CUDA-core kernel's fp32
// sums sit at 6.5e-7 (scripts/time_conv2.py).  2 pairs x 4 groups x 64 columns = all 512 TMEM columns.
constexpr int TC_PAIR_COLS = TC_GROUPS * TC_ACC_COLS;
static_assert(2 * TC_PAIR_COLS <= 512 && (TC_GROUPS == 1 || TC_GROUPS == 2 || TC_GROUPS == 4), "TMEM columns");
__host__ __device__ constexpr int tc_group(int j) { return j * TC_GROUPS / TC_KSTEPS; }
// a row of 12 columns is padded to 13 groups of 16 bytes: with 192-byte rows every lane of a warp that builds the
// operand stores into one of TWO 16-byte bank groups (16-way conflicts: ncu counted 482 k conflicts in 892 k
// shared-memory wavefronts); with 208-byte rows the stores spread over all eight
constexpr int A_ROW = 16, A_S = 13 * 16, A_IY = 2 * A_S, A_PLANE = 12 * A_IY;     // 16, 208, 416, 4992 bytes
constexpr int A_PART = 3 * A_PLANE;                                                // 14,976 (hi, then lo)
constexpr int A_BUF = 2 * A_PART + 2 * A_IY;                                       // + two zero rows behind lo's last plane
constexpr int A_BUF_PADDED = (A_BUF + 127) / 128 * 128;
constexpr int B_KSTEP = 2 * TC_ACC_COLS * 16;                                      // 2,048 bytes: [chunk 2][n 64][4 floats]
constexpr int B_BYTES = TC_KSTEPS * B_KSTEP;                                       // 67,584
constexpr int TC_OFF_B = 0;
constexpr int TC_OFF_A = TC_OFF_B + B_BYTES;
constexpr int TC_OFF_P2 = TC_OFF_A + 2 * A_BUF_PADDED;                             // staging of the CTA's p2 [320][8] floats
constexpr int TC_OFF_IDX = TC_OFF_P2 + P2F * TC_SAMPLES * 4;                       // and idx2 [320][8] bytes
constexpr int TC_OFF_BAR = TC_OFF_IDX + P2F * TC_SAMPLES;                          // mbarriers + the TMEM base address
constexpr int TC_SMEM_BYTES = TC_OFF_BAR + 64;       // mma_done[2], ready[2], weights-landed, TMEM base
static_assert(5000 * 4 <= A_BUF_PADDED, "the raw weights borrow the second operand buffer");
static_assert(TC_SMEM_BYTES <= 227 * 1024, "shared memory");

// element e of half h of K-step j -> (input channel, ky, kx); false for padding (zero weight)
__device__ __forceinline__ bool tc_kelem(int j, int h, int e, int& ci, int& ky, int& kx) {
    if (j < 25) {                        // planes 0 | 1 of tap j
        ci = 4 * h + e; ky = j / 5; kx = j % 5;
        return true;
    }
    ci = 8 + (e & 1);
    if (j < 30) {                        // plane 2 at kx = 0 | 2 of row ky: taps kx and kx + 1
        ky = j - 25; kx = 2 * h + (e >> 1);
        return true;
    }
    ky = 2 * (j - 30) + h;               // plane 2 at kx = 4 of rows ky0 | ky0 + 1: tap kx = 4 (kx = 5 does not exist)
    kx = 4;
    return ky < 5 && (e >> 1) == 0;
}

// conv2.weight [20][10][5][5] (a raw copy `wraw` in shared memory, brought in by one bulk copy) -> B operand
// [kstep][half][n 64][4]: columns 0-31 the TF32-rounded weights (channels 0-9 in columns 0-9, 10-19 in columns
// 16-25), columns 32-63 their remainders.  One task = one 16-byte group.
template <int NT = TC_WORKERS>      // threads 0 .. NT-1 of the block take part
__device__ __forceinline__ void tc_stage_weights(const float* wraw, unsigned char* smem) {
    float4* bs = reinterpret_cast<float4*>(smem + TC_OFF_B);
    constexpr int TASKS = TC_KSTEPS * 2 * TC_ACC_COLS;      // 4,224
#pragma unroll 2
    for (int task = threadIdx.x; task < TASKS; task += NT) {
        const int j = task >> 7, h = (task >> 6) & 1, nn = task & 31;
        const int co = nn < 10 ? nn : (nn >= 16 && nn < 26 ? nn - 6 : -1);
        float v[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            int ci, ky, kx;
            const bool live = tc_kelem(j, h, e, ci, ky, kx);
            v[e] = (live && co >= 0 && ci < 10) ? wraw[co * 250 + ci * 25 + ky * 5 + kx] : 0.f;
        }
        float4 hi, lo;
        split_tf32(v[0], hi.x, lo.x);
        split_tf32(v[1], hi.y, lo.y);
        split_tf32(v[2], hi.z, lo.z);
        split_tf32(v[3], hi.w, lo.w);
        bs[task] = (task & 32) ? lo : hi;
    }
}

// one pair's pooled conv1 output, lane4 layout in global memory -> A operand (hi and lo) in `abuf`
__device__ __forceinline__ void tc_build_a(const float* __restrict__ p1, unsigned char* abuf, int64_t tile, int lane0,
                                           int64_t n) {
    const int id = threadIdx.x;
    if (id >= 216) return;      // (workers 216-255 have no build task)
    const int s = id & 1, g = (id >> 1) % 3, iy = (id / 6) % 12, c = id / 72;
    const int lane = lane0 + s;
    const bool valid = tile * TILE + lane < n;
    const float4* src = reinterpret_cast<const float4*>(p1 + tile * (int64_t)(P1F * TILE)) + lane;
    const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
    unsigned char* dst = abuf + c * A_PLANE + iy * A_IY + s * A_S + (4 * g) * A_ROW;
    float col[4][4];
    if (c < 2) {                 // planes 0, 1: four channels of one column per 16 bytes
        float4 v[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) v[e] = valid ? __ldg(src + ((4 * c + e) * 36 + iy * 3 + g) * TILE) : zero;
        const float t[4][4] = {{v[0].x, v[1].x, v[2].x, v[3].x}, {v[0].y, v[1].y, v[2].y, v[3].y},
                               {v[0].z, v[1].z, v[2].z, v[3].z}, {v[0].w, v[1].w, v[2].w, v[3].w}};
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int e = 0; e < 4; ++e) col[j][e] = t[j][e];
    } else {                     // plane 2: channels 8, 9 of this column and of the next one
        const float4 a8 = valid ? __ldg(src + (8 * 36 + iy * 3 + g) * TILE) : zero;
        const float4 a9 = valid ? __ldg(src + (9 * 36 + iy * 3 + g) * TILE) : zero;
        const float4 n8 = (valid && g < 2) ? __ldg(src + (8 * 36 + iy * 3 + g + 1) * TILE) : zero;
        const float4 n9 = (valid && g < 2) ? __ldg(src + (9 * 36 + iy * 3 + g + 1) * TILE) : zero;
        const float r8[5] = {a8.x, a8.y, a8.z, a8.w, n8.x}, r9[5] = {a9.x, a9.y, a9.z, a9.w, n9.x};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            col[j][0] = r8[j]; col[j][1] = r9[j]; col[j][2] = r8[j + 1]; col[j][3] = r9[j + 1];
        }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        float4 hi, lo;
        split_tf32(col[j][0], hi.x, lo.x);
        split_tf32(col[j][1], hi.y, lo.y);
        split_tf32(col[j][2], hi.z, lo.z);
        split_tf32(col[j][3], hi.w, lo.w);
        *reinterpret_cast<float4*>(dst + j * A_ROW) = hi;
        *reinterpret_cast<float4*>(dst + A_PART + j * A_ROW) = lo;
    }
}

// the 66 MMAs of one pair, issued by one thread, then one commit: 33 of 128x64x8 (A_hi x [B_hi | B_lo]) that
// overwrite / accumulate all 64 accumulator columns, then 33 of 128x32x8 (A_lo x B_hi) onto columns 0-31.  The
// K-step loop is fully unrolled so that every descriptor is the pass's base descriptor plus a compile-time
// constant (computing the offsets at run time cost ~120 cycles per MMA, more than the MMA itself).
__device__ __forceinline__ void tc_issue(uint32_t abuf_addr, uint32_t b_addr, uint32_t d_tmem, uint64_t* bar) {
#pragma unroll 1
    for (int pass = 0; pass < 2; ++pass) {
        const uint32_t idesc = pass == 0 ? umma_idesc_tf32(TC_ACC_COLS) : umma_idesc_tf32(TC_NCOLS);
        const uint32_t a_part = abuf_addr + (pass == 0 ? 0 : A_PART);
        const uint64_t a_tap = umma_desc(a_part, A_PLANE, A_S);                          // planes 0 | 1 of one tap
        const uint64_t a_row = umma_desc(a_part + 2 * A_PLANE, 2 * A_ROW, A_S);          // plane 2 at kx = 0 | 2
        const uint64_t a_col = umma_desc(a_part + 2 * A_PLANE + 4 * A_ROW, A_IY, A_S);   // plane 2 at kx = 4, rows ky0 | ky0+1
        const uint64_t b0 = umma_desc(b_addr, TC_ACC_COLS * 16, 128);
#pragma unroll
        for (int j = 0; j < TC_KSTEPS; ++j) {
            const uint64_t a_desc = j < 25 ? a_tap + (uint64_t)(((j / 5) * A_IY + (j % 5) * A_ROW) >> 4)
                                  : j < 30 ? a_row + (uint64_t)(((j - 25) * A_IY) >> 4)
                                           : a_col + (uint64_t)((2 * (j - 30) * A_IY) >> 4);
            // the first K-step of a group overwrites its accumulator in pass 0; everything else accumulates
            const bool first = j == 0 || tc_group(j) != tc_group(j - 1);
            mma_tf32(d_tmem + tc_group(j) * TC_ACC_COLS, a_desc, b0 + (uint64_t)((j * B_KSTEP) >> 4), idesc,
                     first ? (uint32_t)pass : 1u);
        }
    }
    mma_commit(bar);
}

// bias, Dropout2d, 2x2 max-pool (first maximum in scan order), relu, arg-max code of one pair -> staging.
// bias[j] / sc[j]: this thread's 10 channels (fetched before the wait for the accumulator).
__device__ __forceinline__ void tc_epilogue(uint32_t tmem_acc, const float (&bias)[10], const float (&sc)[10],
                                            float* stage_p2, uint8_t* stage_idx, int pair) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int quad = warp & 3, colh = warp >> 2;
    float v[16];
    {
        float part[TC_GROUPS][16];
#pragma unroll
        for (int g = 0; g < TC_GROUPS; ++g) {
            float u[16];
            const uint32_t col = tmem_acc + ((uint32_t)(32 * quad) << 16) + g * TC_ACC_COLS + 16 * colh;
            tmem_ld16(col, part[g]);                     // hi*hi + lo*hi
            tmem_ld16(col + TC_NCOLS, u);                // hi*lo
#pragma unroll
            for (int j = 0; j < 10; ++j) part[g][j] += u[j];
        }
#pragma unroll
        for (int j = 0; j < 10; ++j) {
            if (TC_GROUPS == 4) v[j] = (part[0][j] + part[1][j]) + (part[2 % TC_GROUPS][j] + part[3 % TC_GROUPS][j]);
            else if (TC_GROUPS == 2) v[j] = part[0][j] + part[1 % TC_GROUPS][j];
            else v[j] = part[0][j];
        }
    }
    const int s = (lane >> 3) & 1, ox = lane & 7;
    const bool writer = (lane & 17) == 0;
#pragma unroll
    for (int j = 0; j < 10; ++j) {
        const float val = (v[j] + bias[j]) * sc[j];
        const float x1 = __shfl_xor_sync(0xffffffffu, val, 1);
        const float x16 = __shfl_xor_sync(0xffffffffu, val, 16);
        const float x17 = __shfl_xor_sync(0xffffffffu, val, 17);
        if (writer) {
            int k;
            const float m = pool4(val, x1, x16, x17, k);
            const int f = (10 * colh + j) * 16 + quad * 4 + (ox >> 1);
            stage_p2[f * TC_SAMPLES + 2 * pair + s] = fmaxf(m, 0.f);
            stage_idx[f * TC_SAMPLES + 2 * pair + s] = (uint8_t)k;
        }
    }
}

// Dropout2d factors of this thread's sample and 10 channels for one pair
__device__ __forceinline__ void tc_keep_scale(const uint8_t* __restrict__ keep2d, float scale, int64_t sample0, int pair,
                                              int64_t n, float (&sc)[10]) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t sample = sample0 + 2 * pair + ((lane >> 3) & 1);
#pragma unroll
    for (int j = 0; j < 10; ++j) sc[j] = 1.f;
    if (keep2d == nullptr) return;
    uint8_t k[10];
#pragma unroll
    for (int j = 0; j < 10; ++j) k[j] = sample < n ? keep2d[sample * 20 + 10 * (warp >> 2) + j] : (uint8_t)0;
#pragma unroll
    for (int j = 0; j < 10; ++j) sc[j] = k[j] ? scale : 0.f;
}

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void workers_sync() { asm volatile("bar.sync 1, %0;" ::"n"(TC_WORKERS) : "memory"); }

}  // namespace pdt
