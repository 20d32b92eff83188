// Fused DKD / KD (+ cross-entropy) logit loss: value and d/dz_student in one pass.
//
// Replaces mdistiller/distillers/DKD.py:8-51, KD.py:8-13, the F.cross_entropy term at
// DKD.py:71 / KD.py:31 and their autograd backward.  Mathematics:
// SURVEY.md Appendix A.1/A.2.
//
// Shape of the computation (bandwidth bound, no tensor cores): every row of the two
// [rows, classes] logit matrices is read ONCE with 128-bit streaming loads into
// registers, reduced with warp shuffles inside a group of LPR lanes, and the gradient
// row is written once.  12 algorithmic bytes per element (16 with a split CE gradient).
//
// Per row, with u = z/T and g the label:
//   m  = max_{j != g} u_j                      (label excluded for DKD, included for KD)
//   e_j = exp(u_j - m),  E = sum e_j           (for student and teacher)
//   W  = sum et_j * ((ut_j - mt) - (us_j - ms))
//   KL over the kept classes = W/Et - (log Et - log Es)           -> NCKD (DKD) / KL (KD)
//   pi = exp(u_g - m) / (exp(u_g - m) + E)                        -> TCKD from log-space terms
// The label term and the rest are accumulated separately, which is what the reference's
// cat_mask does and what keeps (1 - pi) accurate for confident rows (Appendix A.4).
// Everything stays finite where the reference's torch.log(0) would give inf/nan.

#include <cstdlib>
#include <mutex>

#include "common.cuh"

namespace mdb200 {

constexpr int kLossThreads = 256;
constexpr int kLossWarps = kLossThreads / 32;
constexpr int kMaxCtas = 2048;             // partial-sum slots per loss term in the workspace
constexpr size_t kWorkspaceHeader = 64;    // ticket counter lives here, zero between launches
constexpr float kNegBig = -1.0e30f;        // padding / masked label: exp2 -> 0, differences stay finite
// launcher verdict "this grid needs more partial-sum slots than the scratch has" (not a CUDA error)
constexpr cudaError_t kGridExceedsSlots = static_cast<cudaError_t>(-1000);

struct LossArgs {
    const float *zs, *zt;
    const long long *target;
    float *gkd, *gce;
    long long rows, classes, ld_s, ld_t, ld_g;
    float alpha, beta, inv_t, k2;  // k2 = log2(e) / T
    float c_kd;                    // kd_scale * T / rows      (gradient prefactor)
    float c_ce;                    // ce_scale / rows
    double l_kd;                   // kd_scale * T^2 / rows    (loss prefactor)
    double l_ce;                   // ce_scale / rows
    double *part;                  // [2][slots] per-CTA partial sums
    int slots;                     // CTA slots per loss term in `part` (kMaxCtas for a caller workspace)
    unsigned *ticket;
    float *loss_out;
};

struct RowCoef {
    float cs, ct;      // grad_kd_j = cs*es_j - ct*et_j           (kept classes)
    float gt_kd;       // added at the label column
    float cce, gt_ce;  // grad_ce_j = cce*e_j (+ gt_ce at the label)
    float loss_kd, loss_ce;
};

// Row-level scalar algebra shared by the register-resident and the generic kernel.
//   CE: 0 = none, 1 = T == 1 (softmax(z_s) falls out of the student terms), 2 = any T.
// Every lane of a row group runs this redundantly, so it is kept to a handful of SFU
// operations: 4 reciprocals, 4 lg2, 2 ex2 (+1/+1 for the general-T cross entropy).  Only the
// loss VALUE depends on the logarithms; the gradient coefficients use E, Z and the two
// exponentials alone.
__device__ __forceinline__ float fast_rcp(float x) { return __fdividef(1.0f, x); }

// log(a) - log(b) given both logs, or log1p((a-b)/b) when a and b are close (the difference
// `a_minus_b` is supplied separately because the caller can form it without cancellation).
__device__ __forceinline__ float log_ratio(float log_a, float log_b, float a_minus_b, float b, float rb) {
    const float x = a_minus_b * rb;
    // |x| < 1/8: alternating series of log1p through x^8 (truncation < 1e-8 relative)
    if (fabsf(a_minus_b) < 0.125f * b) {
        const float x2 = x * x;
        float p = fmaf(x, -1.0f / 8.0f, 1.0f / 7.0f);
        p = fmaf(x, p, -1.0f / 6.0f);
        p = fmaf(x, p, 1.0f / 5.0f);
        p = fmaf(x, p, -1.0f / 4.0f);
        p = fmaf(x, p, 1.0f / 3.0f);
        p = fmaf(x, p, -0.5f);
        return fmaf(x2, p, x);
    }
    return log_a - log_b;
}

template <bool DKD, int CE>
__device__ __forceinline__ RowCoef row_coefficients(const LossArgs &A, float ms, float mt, float Es,
                                                    float Et, float W, float zsg, float ztg, float m1,
                                                    float Z1) {
    RowCoef r;
    r.cce = r.gt_ce = r.loss_ce = 0.f;
    const float rEs = fast_rcp(Es), rEt = fast_rcp(Et);
    const float logEs = __logf(Es), logEt = __logf(Et);
    // log(Et/Es): series when the two sums are within 12.5 % (small-KL rows), log difference otherwise
    const float kl_kept = W * A.inv_t * rEt - log_ratio(logEt, logEs, Et - Es, Es, rEs);
    if (DKD) {
        // label logit relative to the kept-class maximum; exactly one of {a, sh} is 1
        const float dgs = (zsg - ms) * A.inv_t, dgt = (ztg - mt) * A.inv_t;
        const float qs = __expf(-fabsf(dgs)), qt = __expf(-fabsf(dgt));
        const bool ps = dgs >= 0.f, ptq = dgt >= 0.f;
        const float as = ps ? 1.f : qs, shs = ps ? qs : 1.f;   // exp(dgs - hs), exp(-hs), hs = max(dgs, 0)
        const float at = ptq ? 1.f : qt, sht = ptq ? qt : 1.f;
        const float bs = Es * shs, bt = Et * sht;
        const float Zs = as + bs, Zt = at + bt;
        const float rZs = fast_rcp(Zs), rZt = fast_rcp(Zt);
        const float logZs = __logf(Zs), logZt = __logf(Zt);
        const float hs = fmaxf(dgs, 0.f), ht = fmaxf(dgt, 0.f);
        const float lpi_s = (dgs - hs) - logZs, lpi_t = (dgt - ht) - logZt;      // log pi
        const float lom_s = (logEs - hs) - logZs, lom_t = (logEt - ht) - logZt;  // log (1 - pi)
        const float pi_s = as * rZs, pi_t = at * rZt;
        const float om_s = bs * rZs, om_t = bt * rZt;
        const float d = (pi_s > 0.5f) ? (om_t - om_s) : (pi_s - pi_t);  // pi_s - pi_t without cancellation
        // TCKD = pi_t*log(pi_t/pi_s) + (1-pi_t)*log((1-pi_t)/(1-pi_s)); near-equal arguments go through log1p
        const float l1 = log_ratio(lpi_t, lpi_s, -d, pi_s, fast_rcp(fmaxf(pi_s, 1e-37f)));
        const float l2 = log_ratio(lom_t, lom_s, d, om_s, fast_rcp(fmaxf(om_s, 1e-37f)));
        const float tckd = pi_t * l1 + om_t * l2;
        r.loss_kd = A.alpha * tckd + A.beta * kl_kept;
        r.cs = A.c_kd * (A.beta - A.alpha * d) * rEs;
        r.ct = A.c_kd * A.beta * rEt;
        r.gt_kd = A.c_kd * A.alpha * d;
        if (CE == 1) {
            r.cce = A.c_ce * shs * rZs;
            r.gt_ce = -A.c_ce * om_s;
            r.loss_ce = -lpi_s;
        }
    } else {
        r.loss_kd = kl_kept;
        r.cs = A.c_kd * rEs;
        r.ct = A.c_kd * rEt;
        r.gt_kd = 0.f;
        if (CE == 1) {
            r.cce = A.c_ce * rEs;
            r.gt_ce = -A.c_ce;
            r.loss_ce = logEs - (zsg - ms);
        }
    }
    if (CE == 2) {
        const float e1g = __expf(zsg - m1);
        const float Zf = DKD ? Z1 + e1g : Z1;  // DKD masked the label out of the register row
        const float rZf = fast_rcp(Zf);
        r.cce = A.c_ce * rZf;
        r.gt_ce = DKD ? -A.c_ce * (Z1 * rZf) : -A.c_ce;
        r.loss_ce = __logf(Zf) - (zsg - m1);
    }
    return r;
}

template <int W>
__device__ __forceinline__ float group_max(float v) {
#pragma unroll
    for (int o = W / 2; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
template <int W>
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
    for (int o = W / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// CTA partials -> workspace; the last CTA to arrive adds them in a fixed order, so the two
// loss values are reproducible bit for bit (fixed grid => fixed summation tree).  Only warp 0
// takes part after the one block barrier: lane 0 publishes with an acq_rel ticket (release of
// its partial-sum stores, acquire of everybody else's), the verdict travels by shuffle, and the
// other seven warps retire immediately instead of waiting for the ticket round trip.
__device__ __forceinline__ unsigned ticket_acq_rel(unsigned *p) {
    unsigned old;
    asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(old) : "l"(p) : "memory");
    return old;
}

__device__ __forceinline__ void cta_reduce_and_publish(const LossArgs &A, double acc_kd, double acc_ce) {
    __shared__ double sh[2][kLossWarps];
    acc_kd = warp_sum(acc_kd);
    acc_ce = warp_sum(acc_ce);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
        sh[0][warp] = acc_kd;
        sh[1][warp] = acc_ce;
    }
    __syncthreads();
    if (warp != 0) return;
    unsigned t = 0;
    if (lane == 0) {
        double cta_kd = 0.0, cta_ce = 0.0;
        for (int w = 0; w < kLossWarps; ++w) {
            cta_kd += sh[0][w];
            cta_ce += sh[1][w];
        }
        A.part[blockIdx.x] = cta_kd;
        A.part[A.slots + blockIdx.x] = cta_ce;
        t = ticket_acq_rel(A.ticket);
    }
    t = __shfl_sync(0xffffffffu, t, 0);
    if (t != gridDim.x - 1) return;
    __threadfence();  // the other 31 lanes read the partials too: order their loads after the ticket
    double s_kd = 0.0, s_ce = 0.0;
#pragma unroll 4
    for (unsigned i = lane; i < gridDim.x; i += 32) {
        s_kd += __ldcg(A.part + i);
        s_ce += __ldcg(A.part + A.slots + i);
    }
    s_kd = warp_sum(s_kd);
    s_ce = warp_sum(s_ce);
    if (lane == 0) {
        A.loss_out[0] = (float)(s_kd * A.l_kd);
        A.loss_out[1] = (float)(s_ce * A.l_ce);
        *A.ticket = 0u;  // ready for the next launch on this scratch
    }
}

#define MDB200_FOR4(V, BODY) \
    { float &x = V.x; BODY } { float &x = V.y; BODY } { float &x = V.z; BODY } { float &x = V.w; BODY }

// ---------------------------------------------------------------------------------------
// Register-resident kernel: a group of LPR lanes owns one row, NV float4 per lane per matrix.
// Requires classes % 4 == 0, 16-byte aligned rows, classes <= LPR*NV*4.
// ---------------------------------------------------------------------------------------
// STAGED: the rows of the NEXT sweep travel global -> shared memory as TMA bulk copies
// (cp.async.bulk, completion on a per-warp mbarrier) while the warp works on the current rows in
// registers, so every warp always has a tile in flight without a second register copy.  Needs
// contiguous rows (ld == classes); one shared-memory tile per warp (the registers are the second buffer).
template <int LPR, int NV, bool DKD, int CE, bool SPLIT, bool STAGED = false>
__global__ void __launch_bounds__(kLossThreads, (NV >= 16 ? 1 : 2)) loss_rows_in_registers(const LossArgs A) {
    constexpr int GPW = 32 / LPR;  // rows per warp
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = lane % LPR, grp = lane / LPR;
    const int nvec = (int)(A.classes >> 2);
    const long long sweep = (long long)gridDim.x * kLossWarps * GPW;
    double acc_kd = 0.0, acc_ce = 0.0;

    extern __shared__ __align__(128) unsigned char staged_smem[];
    unsigned long long *bar = nullptr;
    float *tile_s = nullptr, *tile_t = nullptr;
    unsigned parity = 0;
    long long g_staged = -1;
    const long long first = ((long long)blockIdx.x * kLossWarps + warp) * GPW;
    // lane 0: start the copies of the tile whose first row is `b` (rows past the end are simply not copied)
    auto issue_tile = [&](long long b) {
        const long long left = A.rows - b;
        const unsigned bytes = (unsigned)((left < GPW ? left : GPW) * A.classes * sizeof(float));
        mbar_expect_tx(bar, 2u * bytes);
        bulk_g2s(tile_s, A.zs + b * A.classes, bytes, bar);
        bulk_g2s(tile_t, A.zt + b * A.classes, bytes, bar);
    };
    if (STAGED) {
        const unsigned region = (unsigned)((2 * GPW * A.classes * sizeof(float) + 127) & ~127ull);
        bar = reinterpret_cast<unsigned long long *>(staged_smem) + warp;
        tile_s = reinterpret_cast<float *>(staged_smem + 128 + (size_t)warp * region);
        tile_t = tile_s + (size_t)GPW * A.classes;
        if (threadIdx.x == 0) {
            for (int w = 0; w < kLossWarps; ++w) mbar_init(reinterpret_cast<unsigned long long *>(staged_smem) + w, 1);
            mbar_init_fence();
        }
        __syncthreads();
        if (first < A.rows) {
            if (lane == 0) issue_tile(first);
            if (DKD || CE != 0) g_staged = A.target[first + grp < A.rows ? first + grp : A.rows - 1];
        }
    }

    for (long long base = first; base < A.rows; base += sweep) {
        const long long row = base + grp;
        const bool live = row < A.rows;           // idle groups of the last warp redo the last row
        const long long r = live ? row : A.rows - 1;

        float4 s[NV], t[NV];
        long long g = -1;
        if (STAGED) {
            mbar_wait(bar, parity);
            parity ^= 1u;
            const float4 *ps = reinterpret_cast<const float4 *>(tile_s + (size_t)grp * A.classes);
            const float4 *pt = reinterpret_cast<const float4 *>(tile_t + (size_t)grp * A.classes);
#pragma unroll
            for (int i = 0; i < NV; ++i) {
                const int v = sub + i * LPR;
                if (v < nvec) {
                    s[i] = ps[v];
                    t[i] = pt[v];
                } else {
                    s[i] = t[i] = make_float4(kNegBig, kNegBig, kNegBig, kNegBig);
                }
            }
            g = g_staged;
            __syncwarp();                          // every lane has its copy: the tile may be overwritten
            const long long nb = base + sweep;
            if (nb < A.rows) {
                if (lane == 0) {
                    fence_proxy_async_smem();
                    issue_tile(nb);
                }
                if (DKD || CE != 0) g_staged = A.target[nb + grp < A.rows ? nb + grp : A.rows - 1];
            }
        } else {
            const float4 *ps = reinterpret_cast<const float4 *>(A.zs + r * A.ld_s);
            const float4 *pt = reinterpret_cast<const float4 *>(A.zt + r * A.ld_t);
#pragma unroll
            for (int i = 0; i < NV; ++i) {
                const int v = sub + i * LPR;
                if (v < nvec) {
                    s[i] = ld_stream(ps + v);
                    t[i] = ld_stream(pt + v);
                } else {
                    s[i] = t[i] = make_float4(kNegBig, kNegBig, kNegBig, kNegBig);
                }
            }
            if (DKD || CE != 0) g = A.target[r];
        }

        // label logits: picked up (and, for DKD, masked out of the row) by the owning lane
        float zsg = kNegBig, ztg = kNegBig;
        int gv = -1, gc = 0;
        if (DKD || CE != 0) {
            const bool ok = g >= 0 && g < A.classes;
            gv = ok ? (int)(g >> 2) : -1;
            gc = (int)(g & 3);
#pragma unroll
            for (int i = 0; i < NV; ++i) {
                if (sub + i * LPR == gv) {
                    if (gc == 0) { zsg = s[i].x; ztg = t[i].x; if (DKD) s[i].x = t[i].x = kNegBig; }
                    else if (gc == 1) { zsg = s[i].y; ztg = t[i].y; if (DKD) s[i].y = t[i].y = kNegBig; }
                    else if (gc == 2) { zsg = s[i].z; ztg = t[i].z; if (DKD) s[i].z = t[i].z = kNegBig; }
                    else { zsg = s[i].w; ztg = t[i].w; if (DKD) s[i].w = t[i].w = kNegBig; }
                }
            }
            const int owner = ok ? (gv % LPR) : 0;
            zsg = __shfl_sync(0xffffffffu, zsg, owner, LPR);
            ztg = __shfl_sync(0xffffffffu, ztg, owner, LPR);
        }

        // pass 1: maxima of the kept classes
        float ms = kNegBig, mt = kNegBig;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            ms = fmaxf(ms, fmaxf(fmaxf(s[i].x, s[i].y), fmaxf(s[i].z, s[i].w)));
            mt = fmaxf(mt, fmaxf(fmaxf(t[i].x, t[i].y), fmaxf(t[i].z, t[i].w)));
        }
        ms = group_max<LPR>(ms);
        mt = group_max<LPR>(mt);
        const float m1 = fmaxf(ms, zsg);  // T = 1 row maximum for the general CE

        // pass 2: exponentials replace the logits in place; sums and the cross term
        float Es = 0.f, Et = 0.f, W = 0.f, Z1 = 0.f;
        float4 c1[CE == 2 ? NV : 1];
        constexpr float kLog2e = 1.4426950408889634f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            float4 e1 = make_float4(0.f, 0.f, 0.f, 0.f);
#define MDB200_P2(c)                                                   \
    {                                                                  \
        const float ds = s[i].c - ms, dt = t[i].c - mt;                \
        if (CE == 2) { e1.c = ex2_approx((s[i].c - m1) * kLog2e); Z1 += e1.c; } \
        const float es = ex2_approx(ds * A.k2), et = ex2_approx(dt * A.k2);     \
        Es += es; Et += et;                                            \
        W = fmaf(et, dt - ds, W);                                      \
        s[i].c = es; t[i].c = et;                                      \
    }
            MDB200_P2(x) MDB200_P2(y) MDB200_P2(z) MDB200_P2(w)
#undef MDB200_P2
            if (CE == 2) c1[i] = e1;
        }
        Es = group_sum<LPR>(Es);
        Et = group_sum<LPR>(Et);
        W = group_sum<LPR>(W);
        if (CE == 2) Z1 = group_sum<LPR>(Z1);

        const RowCoef k = row_coefficients<DKD, CE>(A, ms, mt, Es, Et, W, zsg, ztg, m1, Z1);
        if (live && sub == 0) {
            acc_kd += (double)k.loss_kd;
            if (CE != 0) acc_ce += (double)k.loss_ce;
        }

        // pass 3: gradient rows
        float4 *pg = reinterpret_cast<float4 *>(A.gkd + r * A.ld_g);
        float4 *pc = SPLIT ? reinterpret_cast<float4 *>(A.gce + r * A.ld_g) : nullptr;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int v = sub + i * LPR;
            float4 gk, gc4;
#define MDB200_P3(c)                                                   \
    {                                                                  \
        gk.c = fmaf(k.cs, s[i].c, -k.ct * t[i].c);                     \
        if (CE != 0) {                                                 \
            const float ce = k.cce * (CE == 2 ? c1[CE == 2 ? i : 0].c : s[i].c); \
            if (SPLIT) gc4.c = ce; else gk.c += ce;                    \
        }                                                              \
    }
            MDB200_P3(x) MDB200_P3(y) MDB200_P3(z) MDB200_P3(w)
#undef MDB200_P3
            if (v == gv) {
                const float add_kd = k.gt_kd + ((CE != 0 && !SPLIT) ? k.gt_ce : 0.f);
                if (gc == 0) { gk.x += add_kd; if (SPLIT) gc4.x += k.gt_ce; }
                else if (gc == 1) { gk.y += add_kd; if (SPLIT) gc4.y += k.gt_ce; }
                else if (gc == 2) { gk.z += add_kd; if (SPLIT) gc4.z += k.gt_ce; }
                else { gk.w += add_kd; if (SPLIT) gc4.w += k.gt_ce; }
            }
            if (live && v < nvec) {
                st_stream(pg + v, gk);
                if (SPLIT) st_stream(pc + v, gc4);
            }
        }
    }
    cta_reduce_and_publish(A, acc_kd, acc_ce);
}

// ---------------------------------------------------------------------------------------
// Lean staged kernel (round 2): same tile pipeline as the STAGED variant above, with the per-sweep
// overhead that made narrow rows issue-bound taken out (ncu r1: 8.07 M warp instructions for 79 MB at
// 65536x100, 52 % issue-active at 23 % occupancy):
//   * the label column is handled THROUGH the shared-memory tile: every lane of a row group reads the two
//     label logits with one broadcast LDS each, lane 0 of the group then overwrites them with the mask
//     value (DKD), and only then do the lanes pull their vectors -- no per-vector compare/select chains;
//   * the label's gradient is one scalar store by lane 0 of the group after the row's vector stores
//     (__syncwarp orders the two stores to that address), instead of a predicated add in every vector;
//   * CFIX: the class count as a compile-time constant for the two standard datasets (100, 1000), so the
//     tail guards, the row offsets inside the tile and the vector count fold away;
//   * MINB: resident blocks per SM asked of the register allocator.
// Measured and dropped (B200, profiles/r2_loss_ab.txt): 3-4 resident blocks per SM (80 / 64 registers: spills, no
// gain -- ncu shows the SMs streaming at 6.4 TB/s while they are active, the time goes to the launch ramp and the
// drain); 8 lanes per row; and drawing tiles from an atomic counter instead of the fixed stride (+20 %: 2 368 warps
// on one counter cost more than the uneven last round they remove).
// Arithmetic and summation order are those of loss_rows_in_registers; only the label column's gradient is
// formed from scalars (same formula).
// ---------------------------------------------------------------------------------------
template <int LPR, int NV, int CFIX, bool DKD, int CE, bool SPLIT, int MINB>
__global__ void __launch_bounds__(kLossThreads, MINB) loss_rows_lean(const LossArgs A) {
    constexpr int GPW = 32 / LPR;  // rows per warp
    const int C = CFIX ? CFIX : (int)A.classes;
    const int nvec = C >> 2;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = lane % LPR, grp = lane / LPR;
    const long long sweep = (long long)gridDim.x * kLossWarps * GPW;
    double acc_kd = 0.0, acc_ce = 0.0;

    extern __shared__ __align__(128) unsigned char staged_smem[];
    const unsigned region = (unsigned)((2u * GPW * C * sizeof(float) + 127u) & ~127u);
    unsigned long long *bar = reinterpret_cast<unsigned long long *>(staged_smem) + warp;
    float *tile_s = reinterpret_cast<float *>(staged_smem + 128 + (size_t)warp * region);
    float *tile_t = tile_s + GPW * C;
    unsigned parity = 0;
    const long long first = ((long long)blockIdx.x * kLossWarps + warp) * GPW;
    auto issue_tile = [&](long long b) {
        const long long left = A.rows - b;
        const unsigned bytes = (unsigned)((left < GPW ? left : GPW) * C * sizeof(float));
        mbar_expect_tx(bar, 2u * bytes);
        bulk_g2s(tile_s, A.zs + b * C, bytes, bar);
        bulk_g2s(tile_t, A.zt + b * C, bytes, bar);
    };
    if (threadIdx.x == 0) {
        for (int w = 0; w < kLossWarps; ++w) mbar_init(reinterpret_cast<unsigned long long *>(staged_smem) + w, 1);
        mbar_init_fence();
    }
    __syncthreads();
    long long g_next = -1;
    if (first < A.rows) {
        if (lane == 0) issue_tile(first);
        if (DKD || CE != 0) g_next = A.target[first + grp < A.rows ? first + grp : A.rows - 1];
    }

    const float4 *ps = reinterpret_cast<const float4 *>(tile_s + grp * C) + sub;
    const float4 *pt = reinterpret_cast<const float4 *>(tile_t + grp * C) + sub;
    constexpr float kLog2e = 1.4426950408889634f;

    for (long long base = first; base < A.rows; base += sweep) {
        const long long row = base + grp;
        const bool live = row < A.rows;
        const long long r = live ? row : A.rows - 1;
        const long long g = g_next;
        const bool has_label = (DKD || CE != 0) && g >= 0 && g < C;

        mbar_wait(bar, parity);
        parity ^= 1u;
        float zsg = kNegBig, ztg = kNegBig;
        if (DKD || CE != 0) {
            if (has_label) {               // same address for the LPR lanes of a group: one broadcast read each
                zsg = tile_s[grp * C + (int)g];
                ztg = tile_t[grp * C + (int)g];
            }
            if (DKD) {
                __syncwarp();              // every lane holds the label logits: now take them out of the row
                if (has_label && sub == 0) tile_s[grp * C + (int)g] = tile_t[grp * C + (int)g] = kNegBig;
                __syncwarp();
            }
        }
        float4 s[NV], t[NV];
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            if (i * LPR + (LPR - 1) < (CFIX ? CFIX / 4 : 0) || sub + i * LPR < nvec) {
                s[i] = ps[i * LPR];
                t[i] = pt[i * LPR];
            } else {
                s[i] = t[i] = make_float4(kNegBig, kNegBig, kNegBig, kNegBig);
            }
        }
        __syncwarp();                      // every lane has its copy: the tile may be overwritten
        const long long nb = base + sweep;
        if (nb < A.rows) {
            if (lane == 0) {
                fence_proxy_async_smem();
                issue_tile(nb);
            }
            if (DKD || CE != 0) g_next = A.target[nb + grp < A.rows ? nb + grp : A.rows - 1];
        }

        // pass 1: maxima of the kept classes
        float ms = kNegBig, mt = kNegBig;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            ms = fmaxf(ms, fmaxf(fmaxf(s[i].x, s[i].y), fmaxf(s[i].z, s[i].w)));
            mt = fmaxf(mt, fmaxf(fmaxf(t[i].x, t[i].y), fmaxf(t[i].z, t[i].w)));
        }
        ms = group_max<LPR>(ms);
        mt = group_max<LPR>(mt);
        const float m1 = fmaxf(ms, zsg);

        // pass 2: exponentials replace the logits in place; sums and the cross term
        float Es = 0.f, Et = 0.f, W = 0.f, Z1 = 0.f;
        float4 c1[CE == 2 ? NV : 1];
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            float4 e1 = make_float4(0.f, 0.f, 0.f, 0.f);
#define MDB200_P2(c)                                                   \
    {                                                                  \
        const float ds = s[i].c - ms, dt = t[i].c - mt;                \
        if (CE == 2) { e1.c = ex2_approx((s[i].c - m1) * kLog2e); Z1 += e1.c; } \
        const float es = ex2_approx(ds * A.k2), et = ex2_approx(dt * A.k2);     \
        Es += es; Et += et;                                            \
        W = fmaf(et, dt - ds, W);                                      \
        s[i].c = es; t[i].c = et;                                      \
    }
            MDB200_P2(x) MDB200_P2(y) MDB200_P2(z) MDB200_P2(w)
#undef MDB200_P2
            if (CE == 2) c1[i] = e1;
        }
        Es = group_sum<LPR>(Es);
        Et = group_sum<LPR>(Et);
        W = group_sum<LPR>(W);
        if (CE == 2) Z1 = group_sum<LPR>(Z1);

        const RowCoef k = row_coefficients<DKD, CE>(A, ms, mt, Es, Et, W, zsg, ztg, m1, Z1);
        if (live && sub == 0) {
            acc_kd += (double)k.loss_kd;
            if (CE != 0) acc_ce += (double)k.loss_ce;
        }

        // pass 3: gradient rows (the label column is patched afterwards)
        float4 *pg = reinterpret_cast<float4 *>(A.gkd + r * A.ld_g) + sub;
        float4 *pc = SPLIT ? reinterpret_cast<float4 *>(A.gce + r * A.ld_g) + sub : nullptr;
        const float nct = -k.ct;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            float4 gk, gc4;
#define MDB200_P3(c)                                                   \
    {                                                                  \
        gk.c = fmaf(k.cs, s[i].c, nct * t[i].c);                       \
        if (CE != 0) {                                                 \
            const float ce = k.cce * (CE == 2 ? c1[CE == 2 ? i : 0].c : s[i].c); \
            if (SPLIT) gc4.c = ce; else gk.c += ce;                    \
        }                                                              \
    }
            MDB200_P3(x) MDB200_P3(y) MDB200_P3(z) MDB200_P3(w)
#undef MDB200_P3
            if (live && (i * LPR + (LPR - 1) < (CFIX ? CFIX / 4 : 0) || sub + i * LPR < nvec)) {
                st_stream(pg + i * LPR, gk);
                if (SPLIT) st_stream(pc + i * LPR, gc4);
            }
        }
        if (DKD || CE != 0) {
            __syncwarp();                  // the row's vector stores come first: the patch below must land last
            if (live && has_label && sub == 0) {
                // the label column from scalars; for DKD its exponentials were masked to exactly 0
                const float esg = DKD ? 0.f : ex2_approx((zsg - ms) * A.k2);
                const float etg = DKD ? 0.f : ex2_approx((ztg - mt) * A.k2);
                float vk = fmaf(k.cs, esg, nct * etg) + k.gt_kd;
                if (CE != 0) {
                    const float e_ce = DKD ? 0.f : (CE == 2 ? ex2_approx((zsg - m1) * kLog2e) : esg);
                    const float vc = k.cce * e_ce + k.gt_ce;
                    if (SPLIT) A.gce[r * A.ld_g + g] = vc; else vk += vc;
                }
                A.gkd[r * A.ld_g + g] = vk;
            }
        }
    }
    cta_reduce_and_publish(A, acc_kd, acc_ce);
}

// ---------------------------------------------------------------------------------------
// Generic kernel: any class count / alignment.  One CTA per row, three sweeps over the
// row (the 2nd and 3rd hit L1/L2), scalar coalesced accesses.
// ---------------------------------------------------------------------------------------
template <typename Op>
__device__ __forceinline__ float cta_allreduce(float v, float *sh, Op op, float ident) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = op(v, __shfl_xor_sync(0xffffffffu, v, o));
    __syncthreads();  // protect sh from the previous use
    if (lane == 0) sh[warp] = v;
    __syncthreads();
    float r = ident;
    for (int w = 0; w < kLossWarps; ++w) r = op(r, sh[w]);
    return r;
}

template <bool DKD, int CE, bool SPLIT>
__global__ void __launch_bounds__(kLossThreads) loss_rows_generic(const LossArgs A) {
    __shared__ float sh[kLossWarps];
    const auto fmax_op = [](float a, float b) { return fmaxf(a, b); };
    const auto add_op = [](float a, float b) { return a + b; };
    constexpr float kLog2e = 1.4426950408889634f;
    const int C = (int)A.classes;
    double acc_kd = 0.0, acc_ce = 0.0;
    for (long long r = blockIdx.x; r < A.rows; r += gridDim.x) {
        const float *zs = A.zs + r * A.ld_s;
        const float *zt = A.zt + r * A.ld_t;
        const long long gl = (DKD || CE != 0) ? A.target[r] : -1;   // KD without CE may come with target == NULL
        const int g = (gl >= 0 && gl < A.classes) ? (int)gl : -1;
        const float zsg = g >= 0 ? __ldg(zs + g) : kNegBig;
        const float ztg = g >= 0 ? __ldg(zt + g) : kNegBig;
        float ms = kNegBig, mt = kNegBig;
        for (int j = threadIdx.x; j < C; j += kLossThreads) {
            if (DKD && j == g) continue;
            ms = fmaxf(ms, __ldg(zs + j));
            mt = fmaxf(mt, __ldg(zt + j));
        }
        ms = cta_allreduce(ms, sh, fmax_op, kNegBig);
        mt = cta_allreduce(mt, sh, fmax_op, kNegBig);
        const float m1 = fmaxf(ms, zsg);
        float Es = 0.f, Et = 0.f, W = 0.f, Z1 = 0.f;
        for (int j = threadIdx.x; j < C; j += kLossThreads) {
            if (DKD && j == g) continue;
            const float a = __ldg(zs + j), b = __ldg(zt + j);
            const float ds = a - ms, dt = b - mt;
            const float es = ex2_approx(ds * A.k2), et = ex2_approx(dt * A.k2);
            Es += es;
            Et += et;
            W = fmaf(et, dt - ds, W);
            if (CE == 2) Z1 += ex2_approx((a - m1) * kLog2e);
        }
        Es = cta_allreduce(Es, sh, add_op, 0.f);
        Et = cta_allreduce(Et, sh, add_op, 0.f);
        W = cta_allreduce(W, sh, add_op, 0.f);
        if (CE == 2) Z1 = cta_allreduce(Z1, sh, add_op, 0.f);
        const RowCoef k = row_coefficients<DKD, CE>(A, ms, mt, Es, Et, W, zsg, ztg, m1, Z1);
        if (threadIdx.x == 0) {
            acc_kd += (double)k.loss_kd;
            if (CE != 0) acc_ce += (double)k.loss_ce;
        }
        float *gk = A.gkd + r * A.ld_g;
        float *gc = SPLIT ? A.gce + r * A.ld_g : nullptr;
        for (int j = threadIdx.x; j < C; j += kLossThreads) {
            const bool at_label = (j == g);
            float vk, vc = 0.f;
            const float a = __ldg(zs + j), b = __ldg(zt + j);
            if (DKD && at_label) {
                vk = k.gt_kd;
                if (CE != 0) vc = k.gt_ce;
            } else {
                const float es = ex2_approx((a - ms) * A.k2), et = ex2_approx((b - mt) * A.k2);
                vk = fmaf(k.cs, es, -k.ct * et);
                if (CE == 1) vc = k.cce * es;
                if (CE == 2) vc = k.cce * ex2_approx((a - m1) * kLog2e);
                if (at_label) {
                    vk += k.gt_kd;
                    vc += k.gt_ce;
                }
            }
            if (CE != 0 && !SPLIT) vk += vc;
            gk[j] = vk;
            if (SPLIT) gc[j] = vc;
        }
    }
    cta_reduce_and_publish(A, acc_kd, acc_ce);
}

__global__ void scale_by_scalar_kernel(float *out, const float *g, const float *up, long long n) {
    const float s = __ldg(up);
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool vec = ((reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(g)) & 15u) == 0;
    if (vec) {
        const long long n4 = n >> 2;
        for (long long i = i0; i < n4; i += stride) {
            float4 v = reinterpret_cast<const float4 *>(g)[i];
            v.x *= s; v.y *= s; v.z *= s; v.w *= s;
            reinterpret_cast<float4 *>(out)[i] = v;
        }
        for (long long i = (n4 << 2) + i0; i < n; i += stride) out[i] = g[i] * s;
    } else {
        for (long long i = i0; i < n; i += stride) out[i] = g[i] * s;
    }
}

// ---------------------------------------------------------------------------------------
// host-side dispatch
// ---------------------------------------------------------------------------------------
// Launch context of one entry-point call: the current device, its SM count, the stream, and whether the
// scratch's arrival counter has to be cleared before the kernel (static scratch: it may hold leftovers).
struct LaunchCtx {
    int dev, sms;
    cudaStream_t st;
    bool clear_ticket;
};

// Everything cached about a kernel is cached PER DEVICE: the dynamic shared-memory opt-in
// (cudaFuncSetAttribute) and the occupancy answer both belong to the device that was current when they
// were made, and one process may drive several GPUs.  One mutex guards all the small tables.
constexpr int kMaxDevices = 64;
static std::mutex s_kernel_cache_mutex;

struct PerDeviceInt {
    int v[kMaxDevices];   // 0 = not looked up yet
};

// Resident CTAs per SM of a kernel at kLossThreads threads and `smem` dynamic bytes; looked up once per device
// (and again whenever `smem` grows, for the staged kernels).
template <typename Kernel>
static int resident_ctas_per_sm(Kernel kernel, int dev, size_t smem, PerDeviceInt &cache, PerDeviceInt *smem_seen = nullptr) {
    const int d = (dev >= 0 && dev < kMaxDevices) ? dev : 0;
    std::lock_guard<std::mutex> lock(s_kernel_cache_mutex);
    if (cache.v[d] > 0 && (!smem_seen || smem_seen->v[d] == (int)smem)) return cache.v[d];
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, kLossThreads, smem) != cudaSuccess || n < 1) n = 1;
    cache.v[d] = n;
    if (smem_seen) smem_seen->v[d] = (int)smem;
    return n;
}

// Dynamic shared memory above 48 KB needs an opt-in per kernel AND per device.
template <typename Kernel>
static cudaError_t ensure_dynamic_smem(Kernel kernel, int dev, size_t smem, PerDeviceInt &configured) {
    const int d = (dev >= 0 && dev < kMaxDevices) ? dev : 0;
    std::lock_guard<std::mutex> lock(s_kernel_cache_mutex);
    if ((size_t)configured.v[d] >= smem) return cudaSuccess;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess) configured.v[d] = (int)smem;
    return e;
}

// Grid fits the scratch?  Then (static scratch only) clear the arrival counter on the stream, right before the
// kernel -- a refused call must not have queued anything.
static cudaError_t before_launch(const LossArgs &A, const LaunchCtx &ctx, long long grid) {
    if (grid > A.slots) return kGridExceedsSlots;
    if (ctx.clear_ticket) return cudaMemsetAsync(A.ticket, 0, kWorkspaceHeader, ctx.st);
    return cudaSuccess;
}

// Persistent grid: never more CTAs than the chip holds at once, so every CTA loops over an
// (almost) equal share of the rows and there is no partial last wave.
template <int LPR, int NV, bool DKD, int CE, bool SPLIT>
static cudaError_t launch_regs(const LossArgs &A, const LaunchCtx &ctx) {
    constexpr int rows_per_cta = kLossWarps * (32 / LPR);
    static PerDeviceInt occ_cache;
    const int occ = resident_ctas_per_sm(loss_rows_in_registers<LPR, NV, DKD, CE, SPLIT>, ctx.dev, 0, occ_cache);
    long long want = (A.rows + rows_per_cta - 1) / rows_per_cta;
    const long long cap = (long long)ctx.sms * occ;
    if (want > cap) want = cap;
    if (want > kMaxCtas) want = kMaxCtas;
    if (cudaError_t e = before_launch(A, ctx, want)) return e;
    loss_rows_in_registers<LPR, NV, DKD, CE, SPLIT><<<(unsigned)want, kLossThreads, 0, ctx.st>>>(A);
    return cudaGetLastError();
}

template <int LPR, int NV, bool DKD, int CE, bool SPLIT>
static cudaError_t launch_staged(const LossArgs &A, const LaunchCtx &ctx) {
    constexpr int GPW = 32 / LPR;
    constexpr int rows_per_cta = kLossWarps * GPW;
    auto kernel = loss_rows_in_registers<LPR, NV, DKD, CE, SPLIT, true>;
    const size_t region = (2 * (size_t)GPW * A.classes * sizeof(float) + 127) & ~(size_t)127;
    const size_t smem = 128 + kLossWarps * region;
    static PerDeviceInt configured, occ_cache, occ_smem;
    if (cudaError_t e = ensure_dynamic_smem(kernel, ctx.dev, smem, configured)) return e;
    const int occ = resident_ctas_per_sm(kernel, ctx.dev, smem, occ_cache, &occ_smem);
    long long want = (A.rows + rows_per_cta - 1) / rows_per_cta;
    const long long cap = (long long)ctx.sms * occ;
    if (want > cap) want = cap;
    if (want > kMaxCtas) want = kMaxCtas;
    if (cudaError_t e = before_launch(A, ctx, want)) return e;
    kernel<<<(unsigned)want, kLossThreads, smem, ctx.st>>>(A);
    return cudaGetLastError();
}

template <int LPR, int NV, int CFIX, bool DKD, int CE, bool SPLIT, int MINB>
static cudaError_t launch_lean(const LossArgs &A, const LaunchCtx &ctx) {
    constexpr int GPW = 32 / LPR;
    constexpr int rows_per_cta = kLossWarps * GPW;
    auto kernel = loss_rows_lean<LPR, NV, CFIX, DKD, CE, SPLIT, MINB>;
    const size_t region = (2 * (size_t)GPW * A.classes * sizeof(float) + 127) & ~(size_t)127;
    const size_t smem = 128 + kLossWarps * region;
    static PerDeviceInt configured, occ_cache, occ_smem;
    if (cudaError_t e = ensure_dynamic_smem(kernel, ctx.dev, smem, configured)) return e;
    const int occ = resident_ctas_per_sm(kernel, ctx.dev, smem, occ_cache, &occ_smem);
    long long want = (A.rows + rows_per_cta - 1) / rows_per_cta;
    const long long cap = (long long)ctx.sms * occ;
    if (want > cap) want = cap;
    if (want > kMaxCtas) want = kMaxCtas;
    if (cudaError_t e = before_launch(A, ctx, want)) return e;
    kernel<<<(unsigned)want, kLossThreads, smem, ctx.st>>>(A);
    return cudaGetLastError();
}

// MDB200_LOSS_LEAN: 0 = round-1 staged kernels, 1 (default) = lean kernels (A/B timing knob)
static int lean_variant() {
    static const int v = getenv("MDB200_LOSS_LEAN") ? atoi(getenv("MDB200_LOSS_LEAN")) : 1;
    return v;
}

template <bool DKD, int CE, bool SPLIT>
static cudaError_t launch_shape(const LossArgs &A, bool vec_ok, const LaunchCtx &ctx) {
    const long long nvec = A.classes >> 2;
    // TMA-staged variants where they measured faster (B200): narrow rows at any batch (65536x100: -9 %), 1000-class
    // rows up to mid-size batches (4096x1000: -10 %, 16384x1000: -3 %; +4.5 % at 65536x1000, so not there).
    // MDB200_LOSS_STAGED=0 switches them off for A/B timing.
    static const bool staged = !(getenv("MDB200_LOSS_STAGED") && atoi(getenv("MDB200_LOSS_STAGED")) == 0);
    if (staged && vec_ok && A.ld_s == A.classes && A.ld_t == A.classes) {
        // lean kernels where they measured faster on B200 (profiles/r2_loss_ab.txt): -8 % at 65536x100 (DKD, KD), -3..-5 % on
        // 1000-class rows up to 32 K rows; the general-T cross entropy next to DKD keeps a third register row and
        // spills there, so that one combination stays on the round-1 kernel
        const int lean = lean_variant();
        constexpr bool third_row = DKD && CE == 2;
        if constexpr (!third_row) {
            if (lean >= 1 && A.classes == 100) return launch_lean<4, 7, 100, DKD, CE, SPLIT, 2>(A, ctx);
            if (lean >= 1 && nvec <= 28) return launch_lean<4, 7, 0, DKD, CE, SPLIT, 2>(A, ctx);
            if (lean >= 1 && nvec <= 32) return launch_lean<8, 4, 0, DKD, CE, SPLIT, 2>(A, ctx);
        }
        if constexpr (CE != 2) {
            if (lean >= 1 && A.classes == 1000 && A.rows <= 32768) return launch_lean<32, 8, 1000, DKD, CE, SPLIT, 2>(A, ctx);
        }
        if (nvec <= 28) return launch_staged<4, 7, DKD, CE, SPLIT>(A, ctx);
        if (nvec <= 32) return launch_staged<8, 4, DKD, CE, SPLIT>(A, ctx);
        if (nvec > 128 && nvec <= 256 && A.rows <= 32768) return launch_staged<32, 8, DKD, CE, SPLIT>(A, ctx);
    }
    if (vec_ok && nvec <= 28) return launch_regs<4, 7, DKD, CE, SPLIT>(A, ctx);     // classes <= 112 (CIFAR-100: 25 of 28 slots)
    if (vec_ok && nvec <= 32) return launch_regs<8, 4, DKD, CE, SPLIT>(A, ctx);
    if (vec_ok && nvec <= 128) return launch_regs<32, 4, DKD, CE, SPLIT>(A, ctx);
    if (vec_ok && nvec <= 256) return launch_regs<32, 8, DKD, CE, SPLIT>(A, ctx);   // ImageNet-1k: 250 of 256 slots
    if constexpr (CE != 2) {  // a third 64-float register row would spill: general-T CE goes generic there
        if (vec_ok && nvec <= 512) return launch_regs<32, 16, DKD, CE, SPLIT>(A, ctx);
    }
    static PerDeviceInt occ_cache;
    const int occ = resident_ctas_per_sm(loss_rows_generic<DKD, CE, SPLIT>, ctx.dev, 0, occ_cache);
    long long want = A.rows;
    const long long cap = (long long)ctx.sms * occ;
    if (want > cap) want = cap;
    if (want > kMaxCtas) want = kMaxCtas;
    if (cudaError_t e = before_launch(A, ctx, want)) return e;
    loss_rows_generic<DKD, CE, SPLIT><<<(unsigned)want, kLossThreads, 0, ctx.st>>>(A);
    return cudaGetLastError();
}

template <bool DKD>
static cudaError_t launch_ce(const LossArgs &A, int ce, bool split, bool vec_ok, const LaunchCtx &ctx) {
    if (ce == 0) return launch_shape<DKD, 0, false>(A, vec_ok, ctx);
    if (ce == 1) return split ? launch_shape<DKD, 1, true>(A, vec_ok, ctx)
                              : launch_shape<DKD, 1, false>(A, vec_ok, ctx);
    return split ? launch_shape<DKD, 2, true>(A, vec_ok, ctx)
                 : launch_shape<DKD, 2, false>(A, vec_ok, ctx);
}

}  // namespace mdb200

using namespace mdb200;

extern "C" size_t mdb200_loss_workspace_bytes(void) {
    return kWorkspaceHeader + 2 * (size_t)kMaxCtas * sizeof(double);
}

static LaunchCtx make_launch_ctx(void *stream, bool clear_ticket) {
    LaunchCtx ctx;
    ctx.dev = 0;
    if (cudaGetDevice(&ctx.dev) != cudaSuccess) ctx.dev = 0;
    ctx.sms = sm_count();
    ctx.st = reinterpret_cast<cudaStream_t>(stream);
    ctx.clear_ticket = clear_ticket;
    return ctx;
}

// Argument checks and LossArgs set-up shared by the two loss entry points; the scratch
// pointers are filled in by the caller.
static int prepare_loss(LossArgs &A, bool &vec_ok, int &ce, bool &split, const float *z_student,
                        const float *z_teacher, const int64_t *target, int64_t rows, int64_t classes,
                        int64_t ld_s, int64_t ld_t, int kind, float alpha, float beta, float temperature,
                        float kd_scale, int ce_mode, float ce_scale, float *grad_kd, float *grad_ce,
                        int64_t ld_g, float *loss_out) {
    if (!z_student || !z_teacher || !grad_kd || !loss_out) return MDB200_EINVAL;
    if (rows <= 0 || classes <= 0 || classes > 0x7fffffffLL || !(temperature > 0.f)) return MDB200_EINVAL;
    if (kind != MDB200_LOSS_DKD && kind != MDB200_LOSS_KD) return MDB200_EINVAL;
    if (ce_mode < MDB200_CE_NONE || ce_mode > MDB200_CE_COMBINED) return MDB200_EINVAL;
    if (ce_mode == MDB200_CE_SPLIT && !grad_ce) return MDB200_EINVAL;
    if ((kind == MDB200_LOSS_DKD || ce_mode != MDB200_CE_NONE) && !target) return MDB200_EINVAL;
    if (kind == MDB200_LOSS_DKD && classes < 2) return MDB200_EINVAL;
    if (ld_s < classes || ld_t < classes || ld_g < classes) return MDB200_ESHAPE;
    if (!aligned4(z_student) || !aligned4(z_teacher) || !aligned4(grad_kd) || (grad_ce && !aligned4(grad_ce)))
        return MDB200_EALIGN;
    A.zs = z_student;
    A.zt = z_teacher;
    A.target = reinterpret_cast<const long long *>(target);
    A.gkd = grad_kd;
    A.gce = grad_ce;
    A.rows = rows;
    A.classes = classes;
    A.ld_s = ld_s;
    A.ld_t = ld_t;
    A.ld_g = ld_g;
    A.alpha = alpha;
    A.beta = beta;
    A.inv_t = 1.0f / temperature;
    A.k2 = (float)(1.4426950408889634 / (double)temperature);
    A.c_kd = (float)((double)kd_scale * (double)temperature / (double)rows);
    A.c_ce = (float)((double)ce_scale / (double)rows);
    A.l_kd = (double)kd_scale * (double)temperature * (double)temperature / (double)rows;
    A.l_ce = (ce_mode == MDB200_CE_NONE) ? 0.0 : (double)ce_scale / (double)rows;
    A.loss_out = loss_out;
    split = ce_mode == MDB200_CE_SPLIT;
    vec_ok = (classes % 4 == 0) && (ld_s % 4 == 0) && (ld_t % 4 == 0) && (ld_g % 4 == 0) &&
             aligned16(z_student) && aligned16(z_teacher) && aligned16(grad_kd) && (!split || aligned16(grad_ce));
    ce = ce_mode == MDB200_CE_NONE ? 0 : (temperature == 1.0f ? 1 : 2);
    return 0;
}

extern "C" int mdb200_distill_loss(const float *z_student, const float *z_teacher, const int64_t *target,
                                   int64_t rows, int64_t classes, int64_t ld_s, int64_t ld_t, int kind,
                                   float alpha, float beta, float temperature, float kd_scale, int ce_mode,
                                   float ce_scale, float *grad_kd, float *grad_ce, int64_t ld_g,
                                   float *loss_out, void *workspace, size_t workspace_bytes, void *stream) {
    LossArgs A;
    bool vec_ok, split;
    int ce;
    const int rc = prepare_loss(A, vec_ok, ce, split, z_student, z_teacher, target, rows, classes, ld_s, ld_t, kind,
                                alpha, beta, temperature, kd_scale, ce_mode, ce_scale, grad_kd, grad_ce, ld_g, loss_out);
    if (rc != 0) return rc;
    if (!workspace || workspace_bytes < mdb200_loss_workspace_bytes()) return MDB200_EWORKSPC;
    if (!aligned16(workspace)) return MDB200_EALIGN;
    A.ticket = reinterpret_cast<unsigned *>(workspace);
    A.part = reinterpret_cast<double *>(reinterpret_cast<char *>(workspace) + kWorkspaceHeader);
    A.slots = kMaxCtas;
    const LaunchCtx ctx = make_launch_ctx(stream, false);
    cudaError_t e = (kind == MDB200_LOSS_DKD) ? launch_ce<true>(A, ce, split, vec_ok, ctx)
                                              : launch_ce<false>(A, ce, split, vec_ok, ctx);
    return (int)e;
}

// ---------------------------------------------------------------------------------------
// Allocation-free variant: the reduction scratch lives inside the library image.
//
// Some hosts of this library are handed device pointers by another runtime and own no
// allocator of their own: a custom-call plugin inside a framework that has reserved all of
// HBM for its arena, or a latency-sensitive thread that must not issue cudaMalloc because
// it serialises against every stream on the device.  They cannot produce the workspace the
// regular entry point asks for, so this one brings its own, fixed at build time.
//
// The array below is a DEVICE object; in host code its name is only a handle for
// cudaGetSymbolAddress (passing `g_static_scratch` itself to a kernel would hand it a host
// address).  Each device that runs this entry point resolves its own copy once.
// ---------------------------------------------------------------------------------------
constexpr int kStaticSlots = MDB200_STATIC_SCRATCH_CTAS;
constexpr size_t kStaticScratchBytes = kWorkspaceHeader + 2 * (size_t)kStaticSlots * sizeof(double);
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

static void *s_static_scratch_ptr[kMaxDevices];   // per-device address of g_static_scratch, 0 = not resolved yet
static std::mutex s_static_scratch_mutex;

// Device address of this device's copy of the scratch (resolved on first use, then cached).
static int static_scratch_for_current_device(void **out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= kMaxDevices) return MDB200_EINVAL;
    std::lock_guard<std::mutex> lock(s_static_scratch_mutex);
    if (s_static_scratch_ptr[dev] == nullptr) {
        void *ptr = nullptr;
        e = cudaGetSymbolAddress((void **)&ptr, g_static_scratch);   // `dev` is current here
        if (e != cudaSuccess) return (int)e;
        if (ptr == nullptr) return (int)cudaErrorInvalidSymbol;
        s_static_scratch_ptr[dev] = ptr;
    }
    *out = s_static_scratch_ptr[dev];
    return 0;
}

extern "C" int mdb200_distill_loss_static_scratch(const float *z_student, const float *z_teacher,
                                                  const int64_t *target, int64_t rows, int64_t classes,
                                                  int64_t ld_s, int64_t ld_t, int kind, float alpha, float beta,
                                                  float temperature, float kd_scale, int ce_mode, float ce_scale,
                                                  float *grad_kd, float *grad_ce, int64_t ld_g, float *loss_out,
                                                  void *stream) {
    LossArgs A;
    bool vec_ok, split;
    int ce;
    int rc = prepare_loss(A, vec_ok, ce, split, z_student, z_teacher, target, rows, classes, ld_s, ld_t, kind, alpha,
                          beta, temperature, kd_scale, ce_mode, ce_scale, grad_kd, grad_ce, ld_g, loss_out);
    if (rc != 0) return rc;
    void *scratch = nullptr;
    rc = static_scratch_for_current_device(&scratch);
    if (rc != 0) return rc;
    // The array keeps whatever the previous call (or an aborted one) left behind; only the
    // arrival counter in the header has to start at zero, the partial sums are overwritten.
    // The launcher clears it on `stream` once it knows the grid fits (clear_ticket), so a
    // refused call has queued nothing and touched nothing.
    A.ticket = reinterpret_cast<unsigned *>(scratch);
    A.part = reinterpret_cast<double *>(reinterpret_cast<char *>(scratch) + kWorkspaceHeader);
    A.slots = kStaticSlots;
    const LaunchCtx ctx = make_launch_ctx(stream, true);
    cudaError_t e = (kind == MDB200_LOSS_DKD) ? launch_ce<true>(A, ce, split, vec_ok, ctx)
                                              : launch_ce<false>(A, ce, split, vec_ok, ctx);
    if (e == kGridExceedsSlots) return MDB200_ESCRATCH;
    return (int)e;
}

extern "C" int mdb200_scale_by_device_scalar(float *out, const float *grad, const float *upstream, int64_t n,
                                             void *stream) {
    if (!out || !grad || !upstream || n < 0) return MDB200_EINVAL;
    if (n == 0) return 0;
    if (!aligned4(out) || !aligned4(grad)) return MDB200_EALIGN;
    long long blocks = (n / 4 + 255) / 256;
    const long long cap = (long long)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    scale_by_scalar_kernel<<<(unsigned)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(out, grad, upstream, n);
    return (int)cudaGetLastError();
}
