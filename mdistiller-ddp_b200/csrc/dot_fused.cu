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
// Cross-rank ordering.  The kernel needs every rank's gradients complete before its first fetch, and no
// rank may read P before every shard owner has written it.  Either the caller brackets the launch with two
// symmetric-memory barrier launches (sync pad == NULL), or -- round 2 -- the kernel does both handshakes
// itself through a SYNC PAD, 64 uint32 of symmetric memory per rank:
//   pad[0..15]   "start" flags, pad[r] written by rank r:  r's gradients are complete      (epoch number)
//   pad[16..31]  "done" flags,  pad[16+r] written by rank r: r's writes to this rank's P have landed
//   pad[32]      local count of finished thread blocks (ticket);  pad[33] local "go" word
// Start: block 0 releases its rank's start flag into every peer's pad and waits (acquire, system scope, `world`
// threads) until its OWN pad shows the current epoch from every rank -- the waiting reads are local; it then
// publishes the epoch in the local "go" word, which one thread of every other block polls at device scope (so a
// whole grid never hammers the system-scope flags).  End: each
// block fences its peer / multicast stores at system scope and takes a ticket; the last block releases the
// rank's done flag into every peer's pad and waits for every peer's done flag before it exits, so kernel
// completion on this stream implies "all of P is final here".  Epochs only grow, a flag is never reset, and a
// wait gives up with a trap after 20 s instead of hanging the GPU.  One rank per GPU (ranks wait on each other).
// The epoch is either a launch argument or -- epoch argument 0 -- counted by the kernel itself in pad[34] (every
// rank's count advances in lockstep, one per launch), which is what lets the launch sit inside a CUDA graph.

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

struct SyncPad {
    unsigned *local;              // this rank's pad, or nullptr: no in-kernel handshake
    unsigned *peer[kMaxPeers];    // every rank's pad base (own rank included)
    unsigned epoch;
    int rank;
};
constexpr int kPadDone = 16, kPadTicket = 32, kPadGo = 33, kPadEpoch = 34;

__device__ __forceinline__ void st_release_sys(unsigned *p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
// spin until *flag has reached `epoch` (wrap-safe); a peer that never arrives ends in a trap, not a hang
__device__ __forceinline__ void pad_wait(const unsigned *flag, unsigned epoch) {
    const unsigned long long t0 = global_ns();
    while ((int)(ld_acquire_sys(flag) - epoch) < 0) {
        __nanosleep(64);
        if (global_ns() - t0 > 20ull * 1000 * 1000 * 1000) __trap();
    }
}

// every rank's inputs are complete (called by all threads of all blocks at kernel start)
__device__ __forceinline__ unsigned pad_start(SyncPad &s, int world) {
    if (s.local == nullptr) return 0u;
    if (s.epoch == 0u) s.epoch = s.local[kPadEpoch] + 1u;     // self-counted: only the last block of a launch advances it
    if (blockIdx.x == 0) {
        if (threadIdx.x < world) {
            st_release_sys(s.peer[threadIdx.x] + s.rank, s.epoch);
            pad_wait(s.local + threadIdx.x, s.epoch);
        }
        __syncthreads();
        if (threadIdx.x == 0)
            asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(s.local + kPadGo), "r"(s.epoch) : "memory");
    } else if (threadIdx.x == 0) {
        const unsigned long long t0 = global_ns();
        unsigned v;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(s.local + kPadGo) : "memory");
            if ((int)(v - s.epoch) >= 0) break;
            __nanosleep(200);
            if (global_ns() - t0 > 25ull * 1000 * 1000 * 1000) __trap();
        } while (true);
    }
    __syncthreads();
    return s.epoch;
}

// this rank's remote writes are visible everywhere AND every peer's writes to this rank have landed
__device__ __forceinline__ void pad_finish(const SyncPad &s, int world) {
    if (s.local == nullptr) return;
    __shared__ int is_last;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned t = atomicAdd(s.local + kPadTicket, 1u);
        is_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!is_last) return;
    if (threadIdx.x == 0) {
        s.local[kPadTicket] = 0u;                         // ready for the next launch
        s.local[kPadEpoch] = s.epoch;                     // every block of this launch has read the old value long ago
    }
    __threadfence_system();
    if (threadIdx.x < world) {
        st_release_sys(s.peer[threadIdx.x] + kPadDone + s.rank, s.epoch);
        pad_wait(s.local + kPadDone + threadIdx.x, s.epoch);
    }
}

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
    // runtime index into the peer table: the compiler keeps a copy of the table in a small stack frame (128-384 B,
    // L1-resident).  Unrolling over all 16 possible peers instead keeps it in the constant bank but hoists 16 loads
    // per bucket (128 registers, a quarter of the occupancy); the measured-best shape of round 1 is kept.
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
    for (int r = 0; r < world; ++r) acc = add4(acc, ld_stream_rw(reinterpret_cast<const float4 *>(peers.ptr[r]) + i));
    return acc;
}

template <bool HAS_KD>
__device__ __forceinline__ void fused_one(float &p, float gt, float gk, float &vt, float &vk, const FusedCoef &c) {
    float d = fmaf(c.wd, p, gt * c.gscale);
    if (!c.bypass_task) {
        vt = c.first_task ? d : fmaf(c.keep, d, __fmul_rn(c.m_task, vt));
        d = vt;
    }
    p = fmaf(c.neg_lr, d, p);
    if (HAS_KD) {
        const float k = gk * c.gscale;
        vk = c.first_kd ? k : fmaf(c.keep, k, __fmul_rn(c.m_kd, vk));
        p = fmaf(c.neg_lr, vk, p);
    }
}

// SRC_T / SRC_K: where the task / KD gradient sums come from; MC_OUT: multicast store of P; HAS_KD = false is the
// single-bucket momentum-SGD step of the base trainer (the task branch alone).
template <int SRC_T, int SRC_K, bool MC_OUT, int kFusedUnroll, bool HAS_KD = true>
__global__ void __launch_bounds__(kFusedThreads)
dot_fused_shard(float4 *p_local, float4 *p_mc, const PeerTable p_peers,
                const float4 *gt_src, const PeerTable gt_peers, const float4 *gk_src, const PeerTable gk_peers,
                float4 *__restrict__ vt, float4 *__restrict__ vk, long long n4, int world, const FusedCoef c,
                SyncPad sync) {
    const long long tile = (long long)kFusedThreads * kFusedUnroll;
    const bool use_vt = !c.bypass_task;
    pad_start(sync, world);
    for (long long t0 = (long long)blockIdx.x * tile; t0 < n4; t0 += (long long)gridDim.x * tile) {
        float4 P[kFusedUnroll], GT[kFusedUnroll], GK[kFusedUnroll], VT[kFusedUnroll], VK[kFusedUnroll];
#pragma unroll
        for (int u = 0; u < kFusedUnroll; ++u) {
            const long long i = t0 + u * kFusedThreads + threadIdx.x;
            if (i < n4) {   // NVLink requests first: they have the longest way to go
                GT[u] = fetch_grad<SRC_T>(gt_src, gt_peers, world, i);
                if (HAS_KD) GK[u] = fetch_grad<SRC_K>(gk_src, gk_peers, world, i);
            }
        }
#pragma unroll
        for (int u = 0; u < kFusedUnroll; ++u) {
            const long long i = t0 + u * kFusedThreads + threadIdx.x;
            if (i < n4) {
                P[u] = ld_stream_rw(p_local + i);
                if (use_vt && !c.first_task) VT[u] = ld_stream_rw(vt + i);
                if (HAS_KD && !c.first_kd) VK[u] = ld_stream_rw(vk + i);
            }
        }
#pragma unroll
        for (int u = 0; u < kFusedUnroll; ++u) {
            const long long i = t0 + u * kFusedThreads + threadIdx.x;
            if (i < n4) {
                fused_one<HAS_KD>(P[u].x, GT[u].x, GK[u].x, VT[u].x, VK[u].x, c);
                fused_one<HAS_KD>(P[u].y, GT[u].y, GK[u].y, VT[u].y, VK[u].y, c);
                fused_one<HAS_KD>(P[u].z, GT[u].z, GK[u].z, VT[u].z, VK[u].z, c);
                fused_one<HAS_KD>(P[u].w, GT[u].w, GK[u].w, VT[u].w, VK[u].w, c);
                if (MC_OUT) {
                    mc_st(p_mc + i, P[u]);
                } else {
                    for (int r = 0; r < world; ++r) st_stream(reinterpret_cast<float4 *>(p_peers.ptr[r]) + i, P[u]);
                }
                if (use_vt) st_stream(vt + i, VT[u]);
                if (HAS_KD) st_stream(vk + i, VK[u]);
            }
        }
    }
    pad_finish(sync, world);
}

template <int SRC_T, int SRC_K, int U, bool HAS_KD = true>
static cudaError_t launch_fused_u(bool mc_out, float4 *p_local, float4 *p_mc, const PeerTable &pp, const float4 *gt,
                                  const PeerTable &gtp, const float4 *gk, const PeerTable &gkp, float4 *vt, float4 *vk,
                                  long long n4, int world, const FusedCoef &c, const SyncPad &sync, cudaStream_t st) {
    const long long tile = (long long)kFusedThreads * U;
    long long blocks = (n4 + tile - 1) / tile;
    // With the in-kernel handshake every block waits at the start, so all of them must be resident at once:
    // ask the device how many fit (40-register kernel: 8 per SM) instead of assuming.
    int per_sm = 8;
    if (sync.local != nullptr) {
        auto k = mc_out ? dot_fused_shard<SRC_T, SRC_K, true, U, HAS_KD> : dot_fused_shard<SRC_T, SRC_K, false, U, HAS_KD>;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, kFusedThreads, 0) != cudaSuccess || per_sm < 1)
            per_sm = 1;
        if (per_sm > 8) per_sm = 8;
    } else {
        per_sm = 16;
    }
    const long long cap = (long long)sm_count() * per_sm;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    if (mc_out)
        dot_fused_shard<SRC_T, SRC_K, true, U, HAS_KD><<<(unsigned)blocks, kFusedThreads, 0, st>>>(
            p_local, p_mc, pp, gt, gtp, gk, gkp, vt, vk, n4, world, c, sync);
    else
        dot_fused_shard<SRC_T, SRC_K, false, U, HAS_KD><<<(unsigned)blocks, kFusedThreads, 0, st>>>(
            p_local, p_mc, pp, gt, gtp, gk, gkp, vt, vk, n4, world, c, sync);
    return cudaGetLastError();
}

// One 16-byte slot per thread per trip measured best (N = 2, peer loads: 79 us vs 84 / 96 / 147 us for
// 2 / 4 / 8 slots): the NVLink round trip is hidden by resident warps (~40 registers -> full occupancy),
// not by per-thread unrolling.
template <int SRC_T, int SRC_K, bool HAS_KD = true>
static cudaError_t launch_fused(bool mc_out, float4 *p_local, float4 *p_mc, const PeerTable &pp, const float4 *gt,
                                const PeerTable &gtp, const float4 *gk, const PeerTable &gkp, float4 *vt, float4 *vk,
                                long long n4, int world, const FusedCoef &c, const SyncPad &sync, cudaStream_t st) {
    return launch_fused_u<SRC_T, SRC_K, 1, HAS_KD>(mc_out, p_local, p_mc, pp, gt, gtp, gk, gkp, vt, vk, n4, world, c, sync, st);
}

// Host description of the sync pad -> kernel argument (pads == NULL: the caller brackets the launch itself).
static int make_sync_pad(SyncPad &s, unsigned *const *pads, int rank, int world, unsigned epoch) {
    s = SyncPad{};
    if (pads == nullptr) return 0;
    if (rank < 0 || rank >= world) return MDB200_EINVAL;
    for (int r = 0; r < world; ++r) {
        if (!pads[r]) return MDB200_EINVAL;
        s.peer[r] = pads[r];
    }
    s.local = pads[rank];
    s.epoch = epoch;
    s.rank = rank;
    return 0;
}

}  // namespace mdb200

using namespace mdb200;

extern "C" int mdb200_dot_step_fused(float *p_local, float *p_multicast, float *const *p_peers,
                                     const float *g_task_local, const float *g_task_multicast,
                                     float *const *g_task_peers, const float *g_kd_local,
                                     const float *g_kd_multicast, float *const *g_kd_peers, int kd_is_reduced,
                                     float *v_task, float *v_kd, int64_t shard_lo, int64_t shard_len, int world,
                                     float lr, float momentum, float momentum_kd, float weight_decay, float dampening,
                                     float grad_scale, int first_task, int first_kd, uint32_t *const *sync_pads,
                                     int rank, uint32_t epoch, void *stream) {
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
    if (shard_len == 0 && sync_pads == nullptr) return 0;     // with a sync pad the handshake still has to happen

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

    SyncPad sync;
    if (int rc = make_sync_pad(sync, reinterpret_cast<unsigned *const *>(sync_pads), rank, world, epoch)) return rc;
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
            e = launch_fused<kMulticast, kLocal>(true, pl, pm, pp, gt_mc, gtp, gk_loc, gkp, vt4, vk4, n4, world, c, sync, st);
        else
            e = launch_fused<kMulticast, kMulticast>(true, pl, pm, pp, gt_mc, gtp,
                                                     reinterpret_cast<const float4 *>(g_kd_multicast + shard_lo), gkp,
                                                     vt4, vk4, n4, world, c, sync, st);
    } else {
        if (kd_is_reduced)
            e = launch_fused<kPeers, kLocal>(false, pl, pm, pp, nullptr, gtp, gk_loc, gkp, vt4, vk4, n4, world, c, sync, st);
        else
            e = launch_fused<kPeers, kPeers>(false, pl, pm, pp, nullptr, gtp, nullptr, gkp, vt4, vk4, n4, world, c, sync, st);
    }
    return (int)e;
}

// Single-bucket variant: gradient sum over ranks -> momentum SGD on the shard -> parameters to every rank.
extern "C" int mdb200_sgd_step_fused(float *p_local, float *p_multicast, float *const *p_peers, const float *g_local,
                                     const float *g_multicast, float *const *g_peers, float *v, int64_t shard_lo,
                                     int64_t shard_len, int world, float lr, float momentum, float weight_decay,
                                     float dampening, float grad_scale, int first_step, uint32_t *const *sync_pads,
                                     int rank, uint32_t epoch, void *stream) {
    if (!p_local || !g_local) return MDB200_EINVAL;
    if (world < 1 || world > kMaxPeers || shard_lo < 0 || shard_len < 0) return MDB200_EINVAL;
    if (shard_len == 0 && sync_pads == nullptr) return 0;
    if ((shard_lo & 3) || (shard_len & 3)) return MDB200_EALIGN;
    const bool bypass = momentum == 0.f;
    if (!bypass && !v) return MDB200_EINVAL;
    const bool mc = p_multicast && g_multicast;
    if (!mc && (!p_peers || !g_peers)) return MDB200_EINVAL;
    if (!aligned16(p_local) || !aligned16(g_local) || (v && !aligned16(v)) ||
        (mc && (!aligned16(p_multicast) || !aligned16(g_multicast))))
        return MDB200_EALIGN;
    PeerTable pp{}, gp{}, none{};
    if (!mc) {
        for (int r = 0; r < world; ++r) {
            if (!p_peers[r] || !g_peers[r]) return MDB200_EINVAL;
            pp.ptr[r] = p_peers[r] + shard_lo;
            gp.ptr[r] = g_peers[r] + shard_lo;
        }
    }
    FusedCoef c;
    c.neg_lr = -lr;
    c.m_task = momentum;
    c.m_kd = 0.f;
    c.wd = weight_decay;
    c.keep = 1.0f - dampening;
    c.gscale = grad_scale;
    c.first_task = first_step != 0;
    c.first_kd = 0;
    c.bypass_task = bypass;
    SyncPad sync;
    if (int rc = make_sync_pad(sync, reinterpret_cast<unsigned *const *>(sync_pads), rank, world, epoch)) return rc;
    float4 *pl = reinterpret_cast<float4 *>(p_local + shard_lo);
    float4 *pm = mc ? reinterpret_cast<float4 *>(p_multicast + shard_lo) : nullptr;
    float4 *v4 = reinterpret_cast<float4 *>(v);
    const long long n4 = shard_len >> 2;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    cudaError_t e;
    if (mc)
        e = launch_fused<kMulticast, kLocal, false>(true, pl, pm, pp, reinterpret_cast<const float4 *>(g_multicast + shard_lo),
                                                    gp, nullptr, none, v4, nullptr, n4, world, c, sync, st);
    else
        e = launch_fused<kPeers, kLocal, false>(false, pl, pm, pp, nullptr, gp, nullptr, none, v4, nullptr, n4, world, c, sync, st);
    return (int)e;
}
