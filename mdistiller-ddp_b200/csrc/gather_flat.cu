// Multi-tensor gather: per-parameter gradient tensors (wherever autograd allocated them) -> one flat
// fp32 bucket, in ONE launch.
//
// Replaces, on this path, what the reference gets from `zero_grad(set_to_none=True)` + keeping the
// `p.grad` tensors themselves (mdistiller/engine/dot.py:111-123, trainer.py:427-431) and what torch DDP
// does when it copies gradients into its buckets (tools/train.py:86): the flat update kernels and the
// reduce-scatter need contiguous buckets.  Letting autograd accumulate into bucket views instead costs
// one `add_` launch per parameter per backward (124 per step for ResNet-18); this kernel costs one.
//
// The chunk table (which tensor, which slice, where in the bucket) depends only on the parameter
// shapes, so the caller builds it once and keeps it on the device; only the source pointers change from
// step to step and travel as kernel arguments.  A null source pointer means "this parameter received
// no gradient in this backward": its slice of the bucket is zero-filled, so the bucket never needs a
// separate memset.  8 algorithmic bytes per element (4 for zero-filled slices).

#include "common.cuh"

namespace mdb200 {

constexpr int kGatherMaxTensors = 448;   // 8-byte pointers by value: 3.5 KB of the 4 KB argument space
constexpr int kGatherThreads = 256;

struct SourceTable {
    const float *src[kGatherMaxTensors];
};

// chunk record, 4 x int64: tensor index (relative to the launch), element count, source offset, bucket offset
__global__ void __launch_bounds__(kGatherThreads) gather_flat_kernel(float *__restrict__ dst, const SourceTable tab,
                                                                     const long long *__restrict__ chunks, int tensor_lo,
                                                                     int tensor_hi) {
    const long long *c = chunks + 4LL * blockIdx.x;
    const int t = (int)c[0];
    if (t < tensor_lo || t >= tensor_hi) return;   // chunk belongs to another launch of a > 448-tensor model
    const long long n = c[1];
    const float *src = tab.src[t - tensor_lo];
    float *out = dst + c[3];
    if (src == nullptr) {
        for (long long i = threadIdx.x; i < n; i += kGatherThreads) out[i] = 0.f;
        return;
    }
    src += c[2];
    const bool vec = ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(out)) & 15u) == 0;
    if (vec) {
        const long long n4 = n >> 2;
        const float4 *s4 = reinterpret_cast<const float4 *>(src);
        float4 *o4 = reinterpret_cast<float4 *>(out);
        for (long long i = threadIdx.x; i < n4; i += kGatherThreads) st_stream(o4 + i, ld_stream(s4 + i));
        for (long long i = (n4 << 2) + threadIdx.x; i < n; i += kGatherThreads) out[i] = src[i];
    } else {
        for (long long i = threadIdx.x; i < n; i += kGatherThreads) out[i] = src[i];
    }
}

}  // namespace mdb200

using namespace mdb200;

extern "C" int mdb200_gather_flat(float *dst_flat, const float *const *src_ptrs, int n_tensors,
                                  const int64_t *chunk_table_dev, int64_t n_chunks, void *stream) {
    if (!dst_flat || !src_ptrs || !chunk_table_dev || n_tensors < 0 || n_chunks < 0) return MDB200_EINVAL;
    if (!aligned4(dst_flat) || (reinterpret_cast<uintptr_t>(chunk_table_dev) & 7u)) return MDB200_EALIGN;
    if (n_tensors == 0 || n_chunks == 0) return 0;
    if (n_chunks > 0x7fffffffLL) return MDB200_EINVAL;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    for (int lo = 0; lo < n_tensors; lo += kGatherMaxTensors) {
        const int hi = (lo + kGatherMaxTensors < n_tensors) ? lo + kGatherMaxTensors : n_tensors;
        SourceTable tab;
        for (int i = lo; i < hi; ++i) {
            if (src_ptrs[i] && !aligned4(src_ptrs[i])) return MDB200_EALIGN;
            tab.src[i - lo] = src_ptrs[i];
        }
        gather_flat_kernel<<<(unsigned)n_chunks, kGatherThreads, 0, st>>>(
            dst_flat, tab, reinterpret_cast<const long long *>(chunk_table_dev), lo, hi);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return (int)e;
    }
    return 0;
}
