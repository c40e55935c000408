// Fused flat-buffer optimizers (reference: optimizer.step() at trainer/trainer.py:58, i.e.
// torch/optim/sgd.py and torch/optim/adam.py selected by config/config.json:38-45), the
// flat scale/cast kernel, and the FP32-FMA peak probe.  All HBM-bound streaming kernels:
// one launch over the whole parameter vector, float4 accesses, grid sized to the SM count.
#include "common.cuh"
#include <cuda_bf16.h>

namespace pdt {

constexpr int OPT_THREADS = 256;

static int stream_grid(int64_t count4) {
    int64_t g = (count4 + OPT_THREADS - 1) / OPT_THREADS;
    const int64_t cap = 148 * 8;
    return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

__global__ void __launch_bounds__(OPT_THREADS)
sgd_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ buf, int64_t count,
           float gscale, pdt_sgd_hyper* __restrict__ hyper) {
    pdl_wait();
    pdl_trigger();
    const float lr = (float)hyper->lr, mom = (float)hyper->momentum, wd = (float)hyper->weight_decay;
    const float damp1 = (float)(1.0 - hyper->dampening);
    const bool nesterov = hyper->nesterov != 0;
    const int step = hyper->step;
    const bool use_mom = hyper->momentum != 0.0 && buf != nullptr;
    for (int64_t i = blockIdx.x * (int64_t)OPT_THREADS + threadIdx.x; i < count;
         i += (int64_t)gridDim.x * OPT_THREADS) {
        float pv = p[i];
        float gv = g[i] * gscale;
        if (wd != 0.f) gv = fmaf(wd, pv, gv);
        if (use_mom) {
            float b = step == 0 ? gv : fmaf(buf[i], mom, damp1 * gv);
            buf[i] = b;
            gv = nesterov ? fmaf(mom, b, gv) : b;
        }
        p[i] = fmaf(-lr, gv, pv);
    }
    advance_step(hyper, step);
}

__global__ void __launch_bounds__(OPT_THREADS)
adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
            float* __restrict__ vmax, int64_t count, float gscale, pdt_adam_hyper* __restrict__ hyper) {
    __shared__ float s_neg_step, s_bc2_sqrt;
    pdl_wait();
    pdl_trigger();
    const int step = hyper->step;
    if (threadIdx.x == 0) {
        // bias corrections in double, as torch computes them in Python (optim/adam.py)
        const double t = (double)(step + 1);
        const double bc1 = 1.0 - pow(hyper->beta1, t);
        const double bc2 = 1.0 - pow(hyper->beta2, t);
        s_neg_step = (float)(-(hyper->lr / bc1));
        s_bc2_sqrt = (float)sqrt(bc2);
    }
    __syncthreads();
    const float w1 = (float)(1.0 - hyper->beta1), b2 = (float)hyper->beta2, w2 = (float)(1.0 - hyper->beta2);
    const float eps = (float)hyper->eps, wd = (float)hyper->weight_decay;
    const bool amsgrad = hyper->amsgrad != 0 && vmax != nullptr;
    const float neg_step = s_neg_step, bc2_sqrt = s_bc2_sqrt;
    for (int64_t i = blockIdx.x * (int64_t)OPT_THREADS + threadIdx.x; i < count;
         i += (int64_t)gridDim.x * OPT_THREADS) {
        const float pv = p[i];
        float gv = g[i] * gscale;
        if (wd != 0.f) gv = fmaf(wd, pv, gv);
        float mv = m[i];
        mv = fmaf(w1, gv - mv, mv);                 // exp_avg.lerp_(grad, 1 - beta1)
        float vv = v[i] * b2;
        vv = fmaf(w2 * gv, gv, vv);                 // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2)
        m[i] = mv;
        v[i] = vv;
        float d = vv;
        if (amsgrad) {
            d = fmaxf(vmax[i], vv);
            vmax[i] = d;
        }
        const float denom = sqrtf(d) / bc2_sqrt + eps;
        p[i] = fmaf(neg_step, mv / denom, pv);      // param.addcdiv_(exp_avg, denom, value=-step_size)
    }
    advance_step(hyper, step);
}

__global__ void __launch_bounds__(OPT_THREADS)
scale_cast_kernel(const float* __restrict__ src, void* __restrict__ dst, int64_t count, float scale, int dtype) {
    const int64_t n4 = count / 4;
    for (int64_t i = blockIdx.x * (int64_t)OPT_THREADS + threadIdx.x; i < n4; i += (int64_t)gridDim.x * OPT_THREADS) {
        float4 t = reinterpret_cast<const float4*>(src)[i];
        t.x *= scale; t.y *= scale; t.z *= scale; t.w *= scale;
        if (dtype == 0) {
            reinterpret_cast<float4*>(dst)[i] = t;
        } else {
            __nv_bfloat162 lo = __floats2bfloat162_rn(t.x, t.y), hi = __floats2bfloat162_rn(t.z, t.w);
            uint2 o;
            o.x = *reinterpret_cast<unsigned*>(&lo);
            o.y = *reinterpret_cast<unsigned*>(&hi);
            reinterpret_cast<uint2*>(dst)[i] = o;
        }
    }
    if (blockIdx.x == 0 && threadIdx.x < (count & 3)) {
        const int64_t i = n4 * 4 + threadIdx.x;
        const float t = src[i] * scale;
        if (dtype == 0) reinterpret_cast<float*>(dst)[i] = t;
        else reinterpret_cast<__nv_bfloat16*>(dst)[i] = __float2bfloat16_rn(t);
    }
}

__global__ void fma_peak_kernel(float* out, int iters) {
    float a[8];
    const float x = 1.0000001f, y = 1e-9f * (float)threadIdx.x;
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = (float)(threadIdx.x + k);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int k = 0; k < 8; ++k) a[k] = fmaf(a[k], x, y);
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += a[k];
    if (s == 12345.678f) out[0] = s;  // never true in practice; keeps the chain alive
}

// same chains with the packed instruction of sm_100 (fma.rn.f32x2 -> FFMA2): 2 FMAs per lane per issue
__global__ void fma2_peak_kernel(float* out, int iters) {
    float2 a[8];
    const float2 x = make_float2(1.0000001f, 0.9999999f);
    const float2 y = make_float2(1e-9f * (float)threadIdx.x, 2e-9f);
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = make_float2((float)(threadIdx.x + k), (float)k);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int k = 0; k < 8; ++k) a[k] = __ffma2_rn(a[k], x, y);
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += a[k].x + a[k].y;
    if (s == 12345.678f) out[0] = s;
}

// Register-operand patterns as in the conv kernels: 16 accumulators, 8 weights and 8 inputs, all distinct
// registers (acc[o] += w[k] * in[o + k]), so the operands are not trivially in the reuse cache.
__global__ void fma_regs_kernel(float* out, int iters) {
    float acc[16], w[8], in[24];
#pragma unroll
    for (int k = 0; k < 16; ++k) acc[k] = (float)(threadIdx.x + k);
#pragma unroll
    for (int k = 0; k < 8; ++k) w[k] = 1.0f + 1e-7f * (float)(threadIdx.x + k);
#pragma unroll
    for (int k = 0; k < 24; ++k) in[k] = 1e-9f * (float)(threadIdx.x * 3 + k);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < 8; ++k)
#pragma unroll
            for (int o = 0; o < 16; ++o) acc[o] = fmaf(w[k], in[o + k], acc[o]);
#pragma unroll
        for (int k = 0; k < 8; ++k) w[k] += 1e-9f;      // keep the loop from being hoisted
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 16; ++k) s += acc[k];
    if (s == 12345.678f) out[0] = s;
}

// the same 128 FMAs per iteration as 64 packed FFMA2: acc2[o] += (w[k], w[k]) * (in[2o + k], in[2o + k + 1])
__global__ void fma2_regs_kernel(float* out, int iters) {
    float2 acc2[8], w2[8];
    float in[24];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc2[k] = make_float2((float)(threadIdx.x + k), (float)k);
#pragma unroll
    for (int k = 0; k < 8; ++k) w2[k] = make_float2(1.0f + 1e-7f * (float)(threadIdx.x + k), 1.0f + 1e-7f * (float)(threadIdx.x + k));
#pragma unroll
    for (int k = 0; k < 24; ++k) in[k] = 1e-9f * (float)(threadIdx.x * 3 + k);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < 8; k += 2)          // even k: pairs (in[2o+k], in[2o+k+1]) are register-aligned
#pragma unroll
            for (int o = 0; o < 8; ++o) {
                acc2[o] = __ffma2_rn(w2[k], make_float2(in[2 * o + k], in[2 * o + k + 1]), acc2[o]);
                acc2[o] = __ffma2_rn(w2[k + 1], make_float2(in[2 * o + k], in[2 * o + k + 1]), acc2[o]);
            }
#pragma unroll
        for (int k = 0; k < 8; ++k) w2[k].x += 1e-9f, w2[k].y += 1e-9f;
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += acc2[k].x + acc2[k].y;
    if (s == 12345.678f) out[0] = s;
}

// mixed: per 8 FFMA2 also 4 shared-memory loads, to see how much issue bandwidth FFMA2 leaves free
__global__ void fma2_lds_peak_kernel(float* out, int iters) {
    __shared__ float sm[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) sm[i] = 1e-9f * i;
    __syncthreads();
    float2 a[8];
    const float2 x = make_float2(1.0000001f, 0.9999999f);
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = make_float2((float)(threadIdx.x + k), (float)k);
    int idx = threadIdx.x & 31;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const float y0 = sm[(idx + r * 32) & 1023], y1 = sm[(idx + r * 32 + 128) & 1023];
            const float y2 = sm[(idx + r * 32 + 256) & 1023], y3 = sm[(idx + r * 32 + 384) & 1023];
#pragma unroll
            for (int k = 0; k < 8; ++k) a[k] = __ffma2_rn(a[k], x, make_float2((k & 1) ? y0 : y1, (k & 2) ? y2 : y3));
        }
        idx += 7;
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += a[k].x + a[k].y;
    if (s == 12345.678f) out[0] = s;
}

// grad_reduce_kernel + adam_kernel in one launch (single-GPU steps: nothing sits between the two).
// Same block shape and summation order as grad_reduce_kernel (bwd.cu), same update as adam_kernel.
__global__ void __launch_bounds__(256)
reduce_adam_kernel(const float* __restrict__ scratch, int64_t T, float beta, float scale,
                   const pdt_stats* __restrict__ stats, float* __restrict__ grad, float* __restrict__ p,
                   float* __restrict__ m, float* __restrict__ v, float* __restrict__ vmax,
                   pdt_adam_hyper* __restrict__ hyper, const float* __restrict__ gscale) {
    __shared__ float red[8][33];
    __shared__ float s_neg_step, s_bc2_sqrt;
    pdl_wait();
    const int step = hyper->step;
    if (threadIdx.x == 0) {
        const double t = (double)(step + 1);
        s_neg_step = (float)(-(hyper->lr / (1.0 - pow(hyper->beta1, t))));
        s_bc2_sqrt = (float)sqrt(1.0 - pow(hyper->beta2, t));
    }
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
        float gv = beta != 0.f ? fmaf(beta, grad[i], s) : s;
        grad[i] = gv;                                   // .grad stays observable, as after a separate reduce
        const float w1 = (float)(1.0 - hyper->beta1), b2 = (float)hyper->beta2, w2 = (float)(1.0 - hyper->beta2);
        const float eps = (float)hyper->eps, wd = (float)hyper->weight_decay;
        const float pv = p[i];
        if (wd != 0.f) gv = fmaf(wd, pv, gv);
        float mv = m[i];
        mv = fmaf(w1, gv - mv, mv);
        float vv = v[i] * b2;
        vv = fmaf(w2 * gv, gv, vv);
        m[i] = mv;
        v[i] = vv;
        float d = vv;
        if (hyper->amsgrad != 0 && vmax != nullptr) {
            d = fmaxf(vmax[i], vv);
            vmax[i] = d;
        }
        p[i] = fmaf(s_neg_step, mv / (sqrtf(d) / s_bc2_sqrt + eps), pv);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0 && stats != nullptr) {
        grad[PDT_NUM_PARAMS + 0] = stats->loss_mean * scale;
        grad[PDT_NUM_PARAMS + 1] = (float)stats->top1;
        grad[PDT_NUM_PARAMS + 2] = (float)stats->topk;
        grad[PDT_NUM_PARAMS + 3] = (float)stats->n_valid;
    }
    advance_step(hyper, step);
}

// one thread: ring[counter++ % capacity] = *value (the per-step loss log of the trainer)
__global__ void ring_push_kernel(float* ring, int32_t* counter, const float* value, int capacity) {
    pdl_wait();
    const int32_t c = *counter;
    ring[c % capacity] = *value;
    *counter = c + 1;
}

}  // namespace pdt

using namespace pdt;

extern "C" int pdt_sgd_step_flat(float* param, const float* grad, float* momentum_buf, int64_t count, float grad_scale,
                                 pdt_sgd_hyper* hyper, void* stream) {
    PDT_REQUIRE(count > 0 && param && grad && hyper, "null pointer or count <= 0");
    launch_pdl(9, sgd_kernel, dim3(stream_grid(count)), dim3(OPT_THREADS), 0, stream, param, grad, momentum_buf, count,
               grad_scale, hyper);
    return check_launch("sgd_step_flat");
}

extern "C" int pdt_adam_step_flat(float* param, const float* grad, float* exp_avg, float* exp_avg_sq,
                                  float* max_exp_avg_sq, int64_t count, float grad_scale, pdt_adam_hyper* hyper,
                                  void* stream) {
    PDT_REQUIRE(count > 0 && param && grad && exp_avg && exp_avg_sq && hyper, "null pointer or count <= 0");
    launch_pdl(9, adam_kernel, dim3(stream_grid(count)), dim3(OPT_THREADS), 0, stream, param, grad, exp_avg, exp_avg_sq,
               max_exp_avg_sq, count, grad_scale, hyper);
    return check_launch("adam_step_flat");
}

extern "C" int pdt_flat_scale_cast(const float* src, void* dst, int64_t count, float scale, int dst_dtype,
                                   void* stream) {
    PDT_REQUIRE(count > 0 && src && dst, "null pointer or count <= 0");
    PDT_REQUIRE(dst_dtype == 0 || dst_dtype == 1, "dst_dtype must be 0 (fp32) or 1 (bf16)");
    scale_cast_kernel<<<stream_grid(count / 4 + 1), OPT_THREADS, 0, (cudaStream_t)stream>>>(src, dst, count, scale,
                                                                                           dst_dtype);
    return check_launch("flat_scale_cast");
}

extern "C" int pdt_fma_peak(float* out, int blocks, int threads, int iters, int mode, void* stream) {
    PDT_REQUIRE(out && blocks > 0 && threads > 0 && threads <= 1024 && iters > 0, "bad arguments");
    PDT_REQUIRE(mode >= 0 && mode <= 4, "mode: 0 scalar FFMA, 1 packed FFMA2, 2 FFMA2 + shared loads, 3/4 register patterns");
    if (mode == 3) fma_regs_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(out, iters / 2);   // 128 FMA per iteration
    else if (mode == 4) fma2_regs_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(out, iters / 2);
    else if (mode == 0) fma_peak_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(out, iters);
    else if (mode == 1) fma2_peak_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(out, iters);
    else fma2_lds_peak_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(out, iters);
    return check_launch("fma_peak");
}

extern "C" int pdt_ring_push(float* ring, int32_t* counter, const float* value, int capacity, void* stream) {
    PDT_REQUIRE(ring && counter && value && capacity > 0, "bad arguments");
    prefer_carveout(ring_push_kernel);
    ring_push_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(ring, counter, value, capacity);
    return check_launch("ring_push");
}

extern "C" int pdt_memcpy_async(void* dst, const void* src, int64_t bytes, void* stream) {
    PDT_REQUIRE(dst && src && bytes > 0, "bad arguments");
    if (cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDefault, (cudaStream_t)stream) != cudaSuccess)
        return check_launch("memcpy_async");
    return PDT_OK;
}

extern "C" int pdt_adam_step_reduce(const void* scratch, int64_t n, float beta, float scale, const pdt_stats* stats,
                                    float* flat_grad, float* param, float* exp_avg, float* exp_avg_sq,
                                    float* max_exp_avg_sq, pdt_adam_hyper* hyper, const float* grad_scale_dev, void* stream) {
    PDT_REQUIRE(n > 0 && scratch && flat_grad && param && exp_avg && exp_avg_sq && hyper, "null pointer or n <= 0");
    prefer_carveout(reduce_adam_kernel);
    reduce_adam_kernel<<<(PDT_NUM_PARAMS + 31) / 32, 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const float*>(scratch), num_tiles(n), beta, scale, stats, flat_grad, param, exp_avg,
        exp_avg_sq, max_exp_avg_sq, hyper, grad_scale_dev);
    return check_launch("adam_step_reduce");
}
