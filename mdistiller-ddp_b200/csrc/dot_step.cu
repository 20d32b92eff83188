// Distillation-Oriented Trainer (dual momentum) and plain momentum-SGD updates on flat
// fp32 ranges -- one launch per range instead of 7 tiny launches per parameter tensor.
//
// Replaces dot() at mdistiller/engine/dot.py:15-55 (driven by step_kd/step, dot.py:96-174)
// and torch.optim.SGD as configured at mdistiller/engine/trainer.py:86-91.
// Recurrence: SURVEY.md Appendix A.3.  Rounding follows torch's kernels: `mul_` is a
// separate multiply, every `add(alpha=)` is one fused multiply-add.
//
// Pure streaming: 5 reads + 3 writes of 4 bytes per element (32 B; 20 B for SGD), no
// reuse, so the only job is to keep enough 128-bit requests in flight on every SM.

#include "common.cuh"

namespace mdb200 {

struct StepCoef {
    float neg_lr;
    float coef_task, coef_kd;  // momentum coefficients after the membership rules
    float wd_task, wd_kd;      // weight decay entering the task / KD branch
    float keep;                // 1 - dampening
    float gscale;              // multiplies both gradients (1/world_size after a sum reduce-scatter)
    int first_task, first_kd;  // buffer is created this step: v = d (no dampening factor)
    int bypass_task;           // momentum == 0: task update uses d directly (dot.py:31,40)
};

template <bool TASK, bool KD>
__device__ __forceinline__ void step_one(float &p, float gt, float gk, float &vt, float &vk, const StepCoef &c) {
    if (TASK) {
        float d = gt * c.gscale;
        d = fmaf(c.wd_task, p, d);
        if (!c.bypass_task) {
            vt = c.first_task ? d : fmaf(c.keep, d, __fmul_rn(c.coef_task, vt));
            d = vt;
        }
        p = fmaf(c.neg_lr, d, p);
    }
    if (KD) {
        float d = gk * c.gscale;
        d = fmaf(c.wd_kd, p, d);
        vk = c.first_kd ? d : fmaf(c.keep, d, __fmul_rn(c.coef_kd, vk));
        p = fmaf(c.neg_lr, vk, p);
    }
}

constexpr int kStepThreads = 256;
constexpr int kStepUnroll = 2;  // float4 slots per thread per trip: 10 independent 16 B loads in flight

template <bool TASK, bool KD>
__global__ void __launch_bounds__(kStepThreads) flat_step_vec4(float4 *__restrict__ p, const float4 *__restrict__ gt,
                                                               const float4 *__restrict__ gk, float4 *__restrict__ vt,
                                                               float4 *__restrict__ vk, long long n4, StepCoef c,
                                                               const float *__restrict__ gs_dev) {
    if (gs_dev) {                            // device-side factor: clipping coefficient / AMP unscale ...
        const float f = __ldg(gs_dev);
        if (f < 0.f) return;                 // ... or the AMP overflow verdict: skip the whole update (mdb200_amp_step_scale)
        c.gscale *= f;
    }
    const long long tile = (long long)kStepThreads * kStepUnroll;
    const bool use_vt = TASK && !c.bypass_task;
    for (long long t0 = (long long)blockIdx.x * tile; t0 < n4; t0 += (long long)gridDim.x * tile) {
        float4 P[kStepUnroll], GT[kStepUnroll], GK[kStepUnroll], VT[kStepUnroll], VK[kStepUnroll];
#pragma unroll
        for (int u = 0; u < kStepUnroll; ++u) {
            const long long i = t0 + u * kStepThreads + threadIdx.x;
            if (i < n4) {
                P[u] = ld_stream_rw(p + i);
                if (TASK) GT[u] = ld_stream(gt + i);
                if (KD) GK[u] = ld_stream(gk + i);
                if (use_vt && !c.first_task) VT[u] = ld_stream_rw(vt + i);
                if (KD && !c.first_kd) VK[u] = ld_stream_rw(vk + i);
            }
        }
#pragma unroll
        for (int u = 0; u < kStepUnroll; ++u) {
            const long long i = t0 + u * kStepThreads + threadIdx.x;
            if (i < n4) {
                step_one<TASK, KD>(P[u].x, GT[u].x, GK[u].x, VT[u].x, VK[u].x, c);
                step_one<TASK, KD>(P[u].y, GT[u].y, GK[u].y, VT[u].y, VK[u].y, c);
                step_one<TASK, KD>(P[u].z, GT[u].z, GK[u].z, VT[u].z, VK[u].z, c);
                step_one<TASK, KD>(P[u].w, GT[u].w, GK[u].w, VT[u].w, VK[u].w, c);
                st_stream(p + i, P[u]);
                if (use_vt) st_stream(vt + i, VT[u]);
                if (KD) st_stream(vk + i, VK[u]);
            }
        }
    }
}

// Scalar path: unaligned ranges and the < 4 element tail of the vector path.
template <bool TASK, bool KD>
__global__ void __launch_bounds__(kStepThreads) flat_step_scalar(float *p, const float *gt, const float *gk, float *vt,
                                                                 float *vk, long long n, StepCoef c, const float *gs_dev) {
    if (gs_dev) {
        const float f = __ldg(gs_dev);
        if (f < 0.f) return;
        c.gscale *= f;
    }
    const bool use_vt = TASK && !c.bypass_task;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        float P = p[i], GT = TASK ? gt[i] : 0.f, GK = KD ? gk[i] : 0.f;
        float VT = (use_vt && !c.first_task) ? vt[i] : 0.f;
        float VK = (KD && !c.first_kd) ? vk[i] : 0.f;
        step_one<TASK, KD>(P, GT, GK, VT, VK, c);
        p[i] = P;
        if (use_vt) vt[i] = VT;
        if (KD) vk[i] = VK;
    }
}

template <bool TASK, bool KD>
static cudaError_t launch_step(float *p, const float *gt, const float *gk, float *vt, float *vk, long long n,
                               const StepCoef &c, cudaStream_t st, const float *gs_dev = nullptr) {
    const bool vec = aligned16(p) && (!TASK || aligned16(gt)) && (!KD || aligned16(gk)) &&
                     (!(TASK && !c.bypass_task) || aligned16(vt)) && (!KD || aligned16(vk));
    const int sms = sm_count();
    long long done = 0;
    if (vec && n >= 4) {
        const long long n4 = n >> 2;
        const long long tile = (long long)kStepThreads * kStepUnroll;
        long long blocks = (n4 + tile - 1) / tile;
        const long long cap = (long long)sms * 32;
        if (blocks > cap) blocks = cap;
        flat_step_vec4<TASK, KD><<<(unsigned)blocks, kStepThreads, 0, st>>>(
            reinterpret_cast<float4 *>(p), reinterpret_cast<const float4 *>(gt), reinterpret_cast<const float4 *>(gk),
            reinterpret_cast<float4 *>(vt), reinterpret_cast<float4 *>(vk), n4, c, gs_dev);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
        done = n4 << 2;
    }
    if (done < n) {
        const long long rem = n - done;
        long long blocks = (rem + kStepThreads - 1) / kStepThreads;
        const long long cap = (long long)sms * 16;
        if (blocks > cap) blocks = cap;
        flat_step_scalar<TASK, KD><<<(unsigned)blocks, kStepThreads, 0, st>>>(
            p + done, TASK ? gt + done : nullptr, KD ? gk + done : nullptr, vt ? vt + done : nullptr,
            vk ? vk + done : nullptr, rem, c, gs_dev);
        return cudaGetLastError();
    }
    return cudaSuccess;
}

}  // namespace mdb200

using namespace mdb200;

extern "C" int mdb200_dot_step(float *p, const float *g_task, const float *g_kd, float *v_task, float *v_kd, int64_t n,
                               float lr, float momentum, float momentum_kd, float weight_decay, float dampening,
                               float grad_scale, int first_task, int first_kd, int membership, void *stream) {
    if (!p || n < 0) return MDB200_EINVAL;
    if (membership < MDB200_DOT_BOTH || membership > MDB200_DOT_KD_ONLY) return MDB200_EINVAL;
    const bool task = membership != MDB200_DOT_KD_ONLY, kd = membership != MDB200_DOT_TASK_ONLY;
    const bool bypass = task && momentum == 0.f;
    if (task && (!g_task || (!bypass && !v_task))) return MDB200_EINVAL;
    if (kd && (!g_kd || !v_kd)) return MDB200_EINVAL;
    if (!aligned4(p) || (task && !aligned4(g_task)) || (kd && !aligned4(g_kd))) return MDB200_EALIGN;
    if (n == 0) return 0;
    // (m_kd + m_task)/2 is formed in double like the reference's Python floats, then rounded once
    const float blend = (float)(((double)momentum_kd + (double)momentum) / 2.0);
    StepCoef c;
    c.neg_lr = -lr;
    c.coef_task = (membership == MDB200_DOT_BOTH) ? momentum : blend;
    c.coef_kd = (membership == MDB200_DOT_BOTH) ? momentum_kd : blend;
    c.wd_task = weight_decay;
    c.wd_kd = (membership == MDB200_DOT_KD_ONLY) ? weight_decay : 0.f;  // dot.py:49-53
    c.keep = 1.0f - dampening;
    c.gscale = grad_scale;
    c.first_task = first_task != 0;
    c.first_kd = first_kd != 0;
    c.bypass_task = bypass;
    if (c.first_kd) c.wd_kd = 0.f;  // a fresh KD buffer is a plain clone of the gradient (dot.py:46-48)
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    cudaError_t e;
    if (task && kd) e = launch_step<true, true>(p, g_task, g_kd, v_task, v_kd, n, c, st);
    else if (task) e = launch_step<true, false>(p, g_task, nullptr, v_task, nullptr, n, c, st);
    else e = launch_step<false, true>(p, nullptr, g_kd, nullptr, v_kd, n, c, st);
    return (int)e;
}

extern "C" int mdb200_sgd_step(float *p, const float *g, float *v, int64_t n, float lr, float momentum,
                               float weight_decay, float dampening, float grad_scale, const float *grad_scale_dev,
                               int first_step, void *stream) {
    if (!p || !g || n < 0) return MDB200_EINVAL;
    if (momentum != 0.f && !v) return MDB200_EINVAL;
    if (!aligned4(p) || !aligned4(g)) return MDB200_EALIGN;
    if (n == 0) return 0;
    StepCoef c;
    c.neg_lr = -lr;
    c.coef_task = momentum;
    c.coef_kd = 0.f;
    c.wd_task = weight_decay;
    c.wd_kd = 0.f;
    c.keep = 1.0f - dampening;
    c.gscale = grad_scale;
    c.first_task = first_step != 0;
    c.first_kd = 0;
    c.bypass_task = momentum == 0.f;
    return (int)launch_step<true, false>(p, g, nullptr, v, nullptr, n, c, reinterpret_cast<cudaStream_t>(stream),
                                         grad_scale_dev);
}
