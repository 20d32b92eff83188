// Optimizer-path variants of the base trainer on flat fp32 ranges:
//   * Adam / AdamW  -- torch.optim.Adam / AdamW as built at mdistiller/engine/trainer.py:92-107
//                      (configs/imagenet/optim/{adam,adamw}.yaml), stepped at trainer.py:321;
//   * global-norm gradient clipping -- nn.utils.clip_grad_norm_ at trainer.py:316-320, split into a
//     sum-of-squares reduction over the rank's shard, a scalar all-reduce done by the host code, and
//     a device-side coefficient the step kernels multiply into the gradient (no host sync).
// All streaming, HBM-bound: 28 B/element for Adam (read p, g, m, v; write p, m, v), 4 B/element for
// the norm.

#include "common.cuh"

namespace mdb200 {

struct AdamCoef {
    float one_minus_b1, b2, one_minus_b2;
    float neg_step_size;    // -lr / (1 - beta1^t)
    float inv_bc2_sqrt;     // 1 / sqrt(1 - beta2^t)
    float eps, wd, decay;   // decay = 1 - lr*wd (AdamW); wd enters the gradient for Adam
    float gscale;
    int decoupled;
};

__device__ __forceinline__ void adam_one(float &p, float g, float &m, float &v, const AdamCoef &c, float gs) {
    g *= gs;
    if (c.decoupled) p *= c.decay;                 // param.mul_(1 - lr*wd)
    else g = fmaf(c.wd, p, g);                     // grad.add(param, alpha=wd)
    m = fmaf(g - m, c.one_minus_b1, m);            // exp_avg.lerp_(grad, 1 - beta1)
    v = fmaf(c.one_minus_b2 * g, g, __fmul_rn(v, c.b2));   // exp_avg_sq.mul_(beta2).addcmul_(g, g, 1 - beta2)
    const float denom = sqrtf(v) * c.inv_bc2_sqrt + c.eps;
    p = fmaf(c.neg_step_size, m / denom, p);       // param.addcdiv_(exp_avg, denom, value=-step_size)
}

constexpr int kAdamThreads = 256;

__global__ void __launch_bounds__(kAdamThreads) adam_vec4(float4 *__restrict__ p, const float4 *__restrict__ g,
                                                          float4 *__restrict__ m, float4 *__restrict__ v, long long n4,
                                                          const AdamCoef c, const float *__restrict__ gs_dev) {
    const float gs_f = gs_dev ? __ldg(gs_dev) : 1.0f;
    if (gs_f < 0.f) return;                  // AMP overflow verdict: skip the whole update (mdb200_amp_step_scale)
    const float gs = c.gscale * gs_f;
    const long long tile = 2LL * kAdamThreads;
    for (long long t0 = (long long)blockIdx.x * tile; t0 < n4; t0 += (long long)gridDim.x * tile) {
        float4 P[2], G[2], M[2], V[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const long long i = t0 + u * kAdamThreads + threadIdx.x;
            if (i < n4) {
                P[u] = ld_stream_rw(p + i);
                G[u] = ld_stream(g + i);
                M[u] = ld_stream_rw(m + i);
                V[u] = ld_stream_rw(v + i);
            }
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const long long i = t0 + u * kAdamThreads + threadIdx.x;
            if (i < n4) {
                adam_one(P[u].x, G[u].x, M[u].x, V[u].x, c, gs);
                adam_one(P[u].y, G[u].y, M[u].y, V[u].y, c, gs);
                adam_one(P[u].z, G[u].z, M[u].z, V[u].z, c, gs);
                adam_one(P[u].w, G[u].w, M[u].w, V[u].w, c, gs);
                st_stream(p + i, P[u]);
                st_stream(m + i, M[u]);
                st_stream(v + i, V[u]);
            }
        }
    }
}

__global__ void __launch_bounds__(kAdamThreads) adam_scalar(float *p, const float *g, float *m, float *v, long long n,
                                                            const AdamCoef c, const float *gs_dev) {
    const float gs_f = gs_dev ? __ldg(gs_dev) : 1.0f;
    if (gs_f < 0.f) return;                  // AMP overflow verdict: skip the whole update (mdb200_amp_step_scale)
    const float gs = c.gscale * gs_f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        float P = p[i], M = m[i], V = v[i];
        adam_one(P, g[i], M, V, c, gs);
        p[i] = P;
        m[i] = M;
        v[i] = V;
    }
}

// sum of (scale*x)^2 over a flat range: fp64 per-CTA partials, last CTA adds them in index order
constexpr int kNormThreads = 256;
constexpr int kNormMaxCtas = 1024;

__global__ void __launch_bounds__(kNormThreads) sum_squares_kernel(const float *__restrict__ x, long long n, float scale,
                                                                   double *out, double *part, unsigned *ticket) {
    double acc = 0.0;
    const bool vec = (reinterpret_cast<uintptr_t>(x) & 15u) == 0;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (vec) {
        const long long n4 = n >> 2;
        const float4 *x4 = reinterpret_cast<const float4 *>(x);
        for (long long i = i0; i < n4; i += stride) {
            const float4 a = ld_stream(x4 + i);
            const float b0 = a.x * scale, b1 = a.y * scale, b2 = a.z * scale, b3 = a.w * scale;
            acc += (double)(b0 * b0 + b1 * b1) + (double)(b2 * b2 + b3 * b3);
        }
        for (long long i = (n4 << 2) + i0; i < n; i += stride) {
            const float b = x[i] * scale;
            acc += (double)(b * b);
        }
    } else {
        for (long long i = i0; i < n; i += stride) {
            const float b = x[i] * scale;
            acc += (double)(b * b);
        }
    }
    __shared__ double sh[kNormThreads / 32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) sh[warp] = acc;
    __syncthreads();
    if (warp != 0) return;
    unsigned t = 0;
    if (lane == 0) {
        double s = 0.0;
        for (int w = 0; w < kNormThreads / 32; ++w) s += sh[w];
        part[blockIdx.x] = s;
        asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(t) : "l"(ticket) : "memory");
    }
    t = __shfl_sync(0xffffffffu, t, 0);
    if (t != gridDim.x - 1) return;
    __threadfence();
    double s = 0.0;
    for (unsigned i = lane; i < gridDim.x; i += 32) s += __ldcg(part + i);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) {
        *out = s;
        *ticket = 0u;
    }
}

__global__ void clip_coefficient_kernel(const double *total_sumsq, float max_norm, float *coef) {
    // clip_coef = max_norm / (total_norm + 1e-6), clamped to 1 (torch.nn.utils.clip_grad_norm_)
    const float norm = (float)sqrt(*total_sumsq);
    *coef = fminf(max_norm / (norm + 1e-6f), 1.0f);
}

// One thread: everything torch's GradScaler decides per step (unscale_ + found_inf + update) plus the clipping
// coefficient, from the fp64 sum of squares of the SCALED gradient.  fp64 sums of fp32 squares cannot overflow, so the
// sum is non-finite exactly when some gradient element is inf or nan.
__global__ void amp_step_scale_kernel(const double *total_sumsq, float *scale, int *growth_tracker, float growth_factor,
                                      float backoff_factor, int growth_interval, float max_norm, float *coef) {
    const double ss = *total_sumsq;
    const float sc = *scale;
    if (!isfinite(ss)) {                     // overflow: skip this step, back the scale off
        *coef = -1.0f;
        *scale = sc * backoff_factor;
        *growth_tracker = 0;
        return;
    }
    const float inv = (float)(1.0 / (double)sc);          // GradScaler: scale.double().reciprocal().float()
    float c = inv;
    if (max_norm > 0.f) {
        const float norm = (float)sqrt(ss) * inv;         // norm of the unscaled, rank-averaged gradient
        c *= fminf(max_norm / (norm + 1e-6f), 1.0f);
    }
    *coef = c;
    int t = *growth_tracker + 1;
    if (t == growth_interval) {
        *scale = sc * growth_factor;
        t = 0;
    }
    *growth_tracker = t;
}

}  // namespace mdb200

using namespace mdb200;

extern "C" int mdb200_amp_step_scale(const double *total_sumsq, float *scale, int *growth_tracker, float growth_factor,
                                     float backoff_factor, int growth_interval, float max_norm, float *coef, void *stream) {
    if (!total_sumsq || !scale || !growth_tracker || !coef) return MDB200_EINVAL;
    if (!(growth_factor >= 1.f) || !(backoff_factor > 0.f && backoff_factor <= 1.f) || growth_interval < 1) return MDB200_EINVAL;
    amp_step_scale_kernel<<<1, 1, 0, reinterpret_cast<cudaStream_t>(stream)>>>(total_sumsq, scale, growth_tracker, growth_factor,
                                                                           backoff_factor, growth_interval, max_norm, coef);
    return (int)cudaGetLastError();
}

extern "C" int mdb200_adam_step(float *p, const float *g, float *m, float *v, int64_t n, float lr, double beta1,
                                double beta2, float eps, float weight_decay, int decoupled, int64_t step,
                                float grad_scale, const float *grad_scale_dev, void *stream) {
    if (!p || !g || !m || !v || n < 0 || step < 1) return MDB200_EINVAL;
    if (!aligned4(p) || !aligned4(g) || !aligned4(m) || !aligned4(v)) return MDB200_EALIGN;
    if (n == 0) return 0;
    // betas arrive as doubles: 1 - beta2 = 0.001 must not inherit float(0.999)'s 1.3e-5 relative error
    const double bc1 = 1.0 - pow(beta1, (double)step), bc2 = 1.0 - pow(beta2, (double)step);
    AdamCoef c;
    c.one_minus_b1 = (float)(1.0 - beta1);
    c.b2 = (float)beta2;
    c.one_minus_b2 = (float)(1.0 - beta2);
    c.neg_step_size = (float)(-(double)lr / bc1);
    c.inv_bc2_sqrt = (float)(1.0 / sqrt(bc2));
    c.eps = eps;
    c.wd = weight_decay;
    c.decay = (float)(1.0 - (double)lr * (double)weight_decay);
    c.gscale = grad_scale;
    c.decoupled = (decoupled != 0) && weight_decay != 0.f;
    if (!decoupled && weight_decay == 0.f) c.wd = 0.f;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const int sms = sm_count();
    long long done = 0;
    if (aligned16(p) && aligned16(g) && aligned16(m) && aligned16(v) && n >= 4) {
        const long long n4 = n >> 2;
        long long blocks = (n4 + 2 * kAdamThreads - 1) / (2 * kAdamThreads);
        if (blocks > (long long)sms * 32) blocks = (long long)sms * 32;
        adam_vec4<<<(unsigned)blocks, kAdamThreads, 0, st>>>(reinterpret_cast<float4 *>(p), reinterpret_cast<const float4 *>(g),
                                                             reinterpret_cast<float4 *>(m), reinterpret_cast<float4 *>(v), n4, c,
                                                             grad_scale_dev);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return (int)e;
        done = n4 << 2;
    }
    if (done < n) {
        const long long rem = n - done;
        long long blocks = (rem + kAdamThreads - 1) / kAdamThreads;
        if (blocks > (long long)sms * 16) blocks = (long long)sms * 16;
        adam_scalar<<<(unsigned)blocks, kAdamThreads, 0, st>>>(p + done, g + done, m + done, v + done, rem, c, grad_scale_dev);
        return (int)cudaGetLastError();
    }
    return 0;
}

extern "C" int mdb200_sum_squares(const float *x, int64_t n, float scale, double *out, void *workspace,
                                  size_t workspace_bytes, void *stream) {
    if (!x || !out || n < 0) return MDB200_EINVAL;
    if (!workspace || workspace_bytes < 64 + (size_t)kNormMaxCtas * sizeof(double)) return MDB200_EWORKSPC;
    if (!aligned4(x) || !aligned16(workspace) || (reinterpret_cast<uintptr_t>(out) & 7u)) return MDB200_EALIGN;
    long long blocks = (n / 4 + kNormThreads - 1) / kNormThreads;
    const long long cap = (long long)sm_count() * 4;
    if (blocks > cap) blocks = cap;
    if (blocks > kNormMaxCtas) blocks = kNormMaxCtas;
    if (blocks < 1) blocks = 1;
    sum_squares_kernel<<<(unsigned)blocks, kNormThreads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        x, n, scale, out, reinterpret_cast<double *>(reinterpret_cast<char *>(workspace) + 64),
        reinterpret_cast<unsigned *>(workspace));
    return (int)cudaGetLastError();
}

extern "C" int mdb200_clip_coefficient(const double *total_sumsq, float max_norm, float *coef, void *stream) {
    if (!total_sumsq || !coef || !(max_norm > 0.f)) return MDB200_EINVAL;
    clip_coefficient_kernel<<<1, 1, 0, reinterpret_cast<cudaStream_t>(stream)>>>(total_sumsq, max_norm, coef);
    return (int)cudaGetLastError();
}
