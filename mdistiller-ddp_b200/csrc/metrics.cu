// On-device counters for the per-iteration progress line: loss values, top-1 / top-5 hits and the
// batch size packed into one 5-float record by ONE launch.
//
// Replaces, on the metrics tail of DOT.train_iter / BaseTrainer.train_iter
// (mdistiller/engine/trainer.py:277-290, 436-450), `accuracy()` (engine/utils.py:125-136: topk, eq,
// sums) and the per-scalar reductions: the record is all-reduced once and read back once.
// A label is in the top-k iff fewer than k classes rank before it; equal logits rank by class index
// (torch.topk leaves the order of equal values unspecified, so any consistent rule is within its contract).

#include "common.cuh"

namespace mdb200 {

constexpr int kMetricThreads = 256;

// record = {loss_ce, loss_kd, top-1 hits, top-5 hits, rows}; record[2] and [3] must be zero on entry
__global__ void __launch_bounds__(kMetricThreads) metrics_record_kernel(const float *__restrict__ logits,
                                                                        const long long *__restrict__ target,
                                                                        long long rows, int classes, long long ld,
                                                                        const float *loss_ce, const float *loss_kd,
                                                                        float *record) {
    const int lane = threadIdx.x & 31;
    const long long warp0 = ((long long)blockIdx.x * kMetricThreads + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * kMetricThreads) >> 5;
    int hit1 = 0, hit5 = 0;
    for (long long r = warp0; r < rows; r += nwarps) {
        const float *z = logits + r * ld;
        const long long g = target[r];
        if (g < 0 || g >= classes) continue;   // never in the top-k
        const float zg = __ldg(z + g);
        int before = 0;
        for (int j = lane; j < classes; j += 32) {
            const float v = __ldg(z + j);
            before += (v > zg) || (v == zg && j < g);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) before += __shfl_xor_sync(0xffffffffu, before, o);
        hit1 += before < 1;
        hit5 += before < 5;
    }
    if (lane == 0) {   // counts are small integers: float atomics stay exact and order-independent
        if (hit1) atomicAdd(record + 2, (float)hit1);
        if (hit5) atomicAdd(record + 3, (float)hit5);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        record[0] = loss_ce ? *loss_ce : 0.f;
        record[1] = loss_kd ? *loss_kd : 0.f;
        record[4] = (float)rows;
    }
}

}  // namespace mdb200

using namespace mdb200;

extern "C" int mdb200_metrics_record(const float *logits, const int64_t *target, int64_t rows, int64_t classes,
                                     int64_t ld, const float *loss_ce, const float *loss_kd, float *record,
                                     void *stream) {
    if (!logits || !target || !record || rows <= 0 || classes <= 0 || classes > 0x7fffffff) return MDB200_EINVAL;
    if (ld < classes) return MDB200_ESHAPE;
    if (!aligned4(logits) || !aligned4(record)) return MDB200_EALIGN;
    long long blocks = (rows * 32 + kMetricThreads - 1) / kMetricThreads;
    const long long cap = (long long)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    metrics_record_kernel<<<(unsigned)blocks, kMetricThreads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        logits, reinterpret_cast<const long long *>(target), rows, (int)classes, ld, loss_ce, loss_kd, record);
    return (int)cudaGetLastError();
}
