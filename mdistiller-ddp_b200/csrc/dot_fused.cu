// Reduce-scatter + DOT update + all-gather in ONE kernel over NVLink peer memory.
//
// Replaces, for the data-parallel step, the sequence  ncclReduceScatter(G_task) [+ G_kd] ->
// flat_step_vec4 on the shard -> ncclAllGather(P)  (itself the replacement of DDP's bucketed
// all-reduce created at tools/train.py:86 plus dot() at mdistiller/engine/dot.py:15-55).
//
// Every rank owns the element range [lo, lo+n) of the flat parameter space.  The flat
// buffers P, G_task, G_kd are symmetric allocations (same size on every GPU, mapped into
// every peer; optionally bound to an NVSwitch multicast object).  For each 16-byte slot of
// its range a thread
//   1. fetches the gradient SUMMED OVER RANKS:  one `multimem.ld_reduce.add.v4.f32` (the
//      NVSwitch adds the W replicas in flight, "NVLS"), or W peer loads when no multicast
//      address is available;
//   2. applies the dual-momentum update in registers (same arithmetic as flat_step_vec4,
//      momentum state is rank-local for the shard);
//   3. writes the new parameters to EVERY rank: one `multimem.st.v4.f32`, or W peer stores.
// Gradient traffic therefore never makes the HBM round trip NCCL's reduce-scatter output and
// all-gather input cost, and transfer and arithmetic overlap slot by slot.
//
// The kernel assumes the gradients of all ranks are complete and visible when it starts and
// makes no statement about when peers may read P: the caller brackets it with the symmetric
// memory barrier (see optim.py), one rank per GPU.

#include "common.cuh"

namespace mdb200 {

constexpr int kMaxPeers = 16;
constexpr int kFusedThreads = 256;

struct FusedCoef {
    float neg_lr, m_task, m_kd, wd, keep, gscale;
    int first_task, first_kd, bypass_task;
};

struct PeerTable {
    float *ptr[kMaxPeers];   // base of the same flat buffer on every rank, already offset to `lo`
};

enum GradSource { kLocal = 0, kMulticast = 1, kPeers = 2 };

__device__ __forceinline__ float4 mc_ld_reduce(const float4 *mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(mc)
                 : "memory");
    return v;
}

__device__ __forceinline__ void mc_st(float4 *mc, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}

__device__ __forceinline__ float4 add4(float4 a, float4 b) {
    return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
}

template <int SRC>
__device__ __forceinline__ float4 fetch_grad(const float4 *local_or_mc, const PeerTable &peers, int world, long long i) {
    if (SRC == kLocal) return ld_stream(local_or_mc + i);
    if (SRC == kMulticast) return mc_ld_reduce(local_or_mc + i);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
    for (int r = 0; r < world; ++r) acc = add4(acc, ld_stream_rw(reinterpret_cast<const float4 *>(peers.ptr[r]) + i));
    return acc;
}

__device__ __forceinline__ void fused_one(float &p, float gt, float gk, float &vt, float &vk, const FusedCoef &c) {
    float d = fmaf(c.wd, p, gt * c.gscale);
    if (!c.bypass_task) {
        vt = c.first_task ? d : fmaf(c.keep, d, __fmul_rn(c.m_task, vt));
        d = vt;
    }
    p = fmaf(c.neg_lr, d, p);
    const float k = gk * c.gscale;
    vk = c.first_kd ? k : fmaf(c.keep, k, __fmul_rn(c.m_kd, vk));
    p = fmaf(c.neg_lr, vk, p);
}

// SRC_T / SRC_K: where the task / KD gradient sums come from; MC_OUT: multicast store of P.
template <int SRC_T, int SRC_K, bool MC_OUT, int kFusedUnroll>
__global__ void __launch_bounds__(kFusedThreads)
dot_fused_shard(float4 *p_local, float4 *p_mc, const PeerTable p_peers,
                const float4 *gt_src, const PeerTable gt_peers, const float4 *gk_src, const PeerTable gk_peers,
                float4 *__restrict__ vt, float4 *__restrict__ vk, long long n4, int world, const FusedCoef c) {
    const long long tile = (long long)kFusedThreads * kFusedUnroll;
    const bool use_vt = !c.bypass_task;
    for (long long t0 = (long long)blockIdx.x * tile; t0 < n4; t0 += (long long)gridDim.x * tile) {
        float4 P[kFusedUnroll], GT[kFusedUnroll], GK[kFusedUnroll], VT[kFusedUnroll], VK[kFusedUnroll];
#pragma unroll
        for (int u = 0; u < kFusedUnroll; ++u) {
            const long long i = t0 + u * kFusedThreads + threadIdx.x;
            if (i < n4) {   // NVLink requests first: they have the longest way to go
                GT[u] = fetch_grad<SRC_T>(gt_src, gt_peers, world, i);
                GK[u] = fetch_grad<SRC_K>(gk_src, gk_peers, world, i);
            }
        }
#pragma unroll
        for (int u = 0; u < kFusedUnroll; ++u) {
            const long long i = t0 + u * kFusedThreads + threadIdx.x;
            if (i < n4) {
                P[u] = ld_stream_rw(p_local + i);
                if (use_vt && !c.first_task) VT[u] = ld_stream_rw(vt + i);
                if (!c.first_kd) VK[u] = ld_stream_rw(vk + i);
            }
        }
#pragma unroll
        for (int u = 0; u < kFusedUnroll; ++u) {
            const long long i = t0 + u * kFusedThreads + threadIdx.x;
            if (i < n4) {
                fused_one(P[u].x, GT[u].x, GK[u].x, VT[u].x, VK[u].x, c);
                fused_one(P[u].y, GT[u].y, GK[u].y, VT[u].y, VK[u].y, c);
                fused_one(P[u].z, GT[u].z, GK[u].z, VT[u].z, VK[u].z, c);
                fused_one(P[u].w, GT[u].w, GK[u].w, VT[u].w, VK[u].w, c);
                if (MC_OUT) {
                    mc_st(p_mc + i, P[u]);
                } else {
                    for (int r = 0; r < world; ++r) st_stream(reinterpret_cast<float4 *>(p_peers.ptr[r]) + i, P[u]);
                }
                if (use_vt) st_stream(vt + i, VT[u]);
                st_stream(vk + i, VK[u]);
            }
        }
    }
}

template <int SRC_T, int SRC_K, int U>
static cudaError_t launch_fused_u(bool mc_out, float4 *p_local, float4 *p_mc, const PeerTable &pp, const float4 *gt,
                                  const PeerTable &gtp, const float4 *gk, const PeerTable &gkp, float4 *vt, float4 *vk,
                                  long long n4, int world, const FusedCoef &c, cudaStream_t st) {
    const long long tile = (long long)kFusedThreads * U;
    long long blocks = (n4 + tile - 1) / tile;
    const long long cap = (long long)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    if (mc_out)
        dot_fused_shard<SRC_T, SRC_K, true, U><<<(unsigned)blocks, kFusedThreads, 0, st>>>(p_local, p_mc, pp, gt, gtp, gk,
                                                                                           gkp, vt, vk, n4, world, c);
    else
        dot_fused_shard<SRC_T, SRC_K, false, U><<<(unsigned)blocks, kFusedThreads, 0, st>>>(p_local, p_mc, pp, gt, gtp, gk,
                                                                                            gkp, vt, vk, n4, world, c);
    return cudaGetLastError();
}

// One 16-byte slot per thread per trip measured best (N = 2, peer loads: 79 us vs 84 / 96 / 147 us for
// 2 / 4 / 8 slots): the NVLink round trip is hidden by resident warps (~40 registers -> full occupancy),
// not by per-thread unrolling.
template <int SRC_T, int SRC_K>
static cudaError_t launch_fused(bool mc_out, float4 *p_local, float4 *p_mc, const PeerTable &pp, const float4 *gt,
                                const PeerTable &gtp, const float4 *gk, const PeerTable &gkp, float4 *vt, float4 *vk,
                                long long n4, int world, const FusedCoef &c, cudaStream_t st) {
    return launch_fused_u<SRC_T, SRC_K, 1>(mc_out, p_local, p_mc, pp, gt, gtp, gk, gkp, vt, vk, n4, world, c, st);
}

}  // namespace mdb200

using namespace mdb200;

extern "C" int mdb200_dot_step_fused(float *p_local, float *p_multicast, float *const *p_peers,
                                     const float *g_task_local, const float *g_task_multicast,
                                     float *const *g_task_peers, const float *g_kd_local,
                                     const float *g_kd_multicast, float *const *g_kd_peers, int kd_is_reduced,
                                     float *v_task, float *v_kd, int64_t shard_lo, int64_t shard_len, int world,
                                     float lr, float momentum, float momentum_kd, float weight_decay, float dampening,
                                     float grad_scale, int first_task, int first_kd, void *stream) {
    if (!p_local || !g_task_local || !g_kd_local || !v_kd) return MDB200_EINVAL;
    if (world < 1 || world > kMaxPeers || shard_lo < 0 || shard_len < 0) return MDB200_EINVAL;
    if ((shard_lo & 3) || (shard_len & 3)) return MDB200_EALIGN;   // shards are whole float4 slots by construction
    const bool bypass = momentum == 0.f;
    if (!bypass && !v_task) return MDB200_EINVAL;
    const bool mc = p_multicast && g_task_multicast && (kd_is_reduced || g_kd_multicast);
    if (!mc && (!p_peers || !g_task_peers || (!kd_is_reduced && !g_kd_peers))) return MDB200_EINVAL;
    if (!aligned16(p_local) || !aligned16(g_task_local) || !aligned16(g_kd_local) || !aligned16(v_kd) ||
        (v_task && !aligned16(v_task)) || (mc && (!aligned16(p_multicast) || !aligned16(g_task_multicast))))
        return MDB200_EALIGN;
    if (shard_len == 0) return 0;

    PeerTable pp{}, gtp{}, gkp{};
    if (!mc) {
        for (int r = 0; r < world; ++r) {
            if (!p_peers[r] || !g_task_peers[r] || (!kd_is_reduced && !g_kd_peers[r])) return MDB200_EINVAL;
            pp.ptr[r] = p_peers[r] + shard_lo;
            gtp.ptr[r] = g_task_peers[r] + shard_lo;
            gkp.ptr[r] = kd_is_reduced ? nullptr : g_kd_peers[r] + shard_lo;
        }
    }
    FusedCoef c;
    c.neg_lr = -lr;
    c.m_task = momentum;
    c.m_kd = momentum_kd;
    c.wd = weight_decay;
    c.keep = 1.0f - dampening;
    c.gscale = grad_scale;
    c.first_task = first_task != 0;
    c.first_kd = first_kd != 0;
    c.bypass_task = bypass;

    float4 *pl = reinterpret_cast<float4 *>(p_local + shard_lo);
    float4 *pm = mc ? reinterpret_cast<float4 *>(p_multicast + shard_lo) : nullptr;
    float4 *vt4 = reinterpret_cast<float4 *>(v_task), *vk4 = reinterpret_cast<float4 *>(v_kd);
    const long long n4 = shard_len >> 2;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    // the KD sum is either already in this rank's bucket (reduce-scattered earlier, overlapping the
    // second backward) or fetched like the task sum
    const float4 *gk_loc = reinterpret_cast<const float4 *>(g_kd_local + shard_lo);
    cudaError_t e;
    if (mc) {
        const float4 *gt_mc = reinterpret_cast<const float4 *>(g_task_multicast + shard_lo);
        if (kd_is_reduced)
            e = launch_fused<kMulticast, kLocal>(true, pl, pm, pp, gt_mc, gtp, gk_loc, gkp, vt4, vk4, n4, world, c, st);
        else
            e = launch_fused<kMulticast, kMulticast>(true, pl, pm, pp, gt_mc, gtp,
                                                     reinterpret_cast<const float4 *>(g_kd_multicast + shard_lo), gkp,
                                                     vt4, vk4, n4, world, c, st);
    } else {
        if (kd_is_reduced)
            e = launch_fused<kPeers, kLocal>(false, pl, pm, pp, nullptr, gtp, gk_loc, gkp, vt4, vk4, n4, world, c, st);
        else
            e = launch_fused<kPeers, kPeers>(false, pl, pm, pp, nullptr, gtp, nullptr, gkp, vt4, vk4, n4, world, c, st);
    }
    return (int)e;
}
