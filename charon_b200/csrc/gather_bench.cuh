// charon_b200/csrc/gather_bench.cuh — the "achievable random-gather bandwidth" microbenchmark of SURVEY 8(d).
// Same addressing as the filter query (3 x hash_and_fit -> one 32-byte sector each) on the SAME filter allocation, but
// with nothing else in the kernel: values come from a counter-based generator, the AND of the three rows is popcounted
// into one word per thread.  bytes moved (algorithmic) = 96 B per value; this is the denominator the fused count
// kernel's gather rate is compared with.
#pragma once
#include "count_kernel.cuh"

namespace cgk {

__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

// mode 0: full query (3 hash_and_fit + 3 row loads per value);  mode 1: one precomputed-address 32-bit load per value
// (pure random sector gather, address = splitmix % words)
template <int MODE>
__global__ void __launch_bounds__(256) gather_bench_kernel(const CountParams P, uint64_t n_values, uint64_t filter_words, int mode,
                                                           unsigned long long *sink) {
    const uint64_t pol = l2_policy_evict_last();
    unsigned long long acc = 0;
    const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
    for (uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; i < n_values; i += stride * 4) {
        // four independent values per iteration: 12 loads in flight per thread
        Mask<ModeTraits<MODE>::NW> m[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const uint64_t idx = i + uint64_t(u) * stride;
            const uint64_t v = splitmix64(idx);
            if (mode == 0) m[u] = bulk_contains<MODE>(P, v, pol);
            else {
                const uint64_t w = __umul64hi(v, filter_words);
                const uint32_t *ap = reinterpret_cast<const uint32_t *>(P.filter) + w;
                uint32_t x;
                if (mode == 1) x = ldg_na_u32(ap, pol);
                else if (mode == 2) asm volatile("ld.global.u32 %0, [%1];" : "=r"(x) : "l"(ap));
                else if (mode == 3) asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(x) : "l"(ap));
                else if (mode == 4) asm volatile("ld.global.cv.u32 %0, [%1];" : "=r"(x) : "l"(ap));
                else if (mode == 5) asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(x) : "l"(ap));
                else if (mode == 6) asm volatile("ld.global.nc.L1::no_allocate.L2::64B.u32 %0, [%1];" : "=r"(x) : "l"(ap));
                else if (mode == 7) asm volatile("ld.global.L1::evict_first.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(x) : "l"(ap), "l"(l2_policy_evict_first()));
                else asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(x) : "l"(ap));
                m[u].w[0] = x;
#pragma unroll
                for (int j = 1; j < ModeTraits<MODE>::NW; ++j) m[u].w[j] = 0;
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (i + uint64_t(u) * stride < n_values) {
#pragma unroll
                for (int j = 0; j < ModeTraits<MODE>::NW; ++j) acc += __popcll(m[u].w[j]);
            }
    }
    if (acc == 0x7FFFFFFFFFFFFFFFull) *sink = acc;   // keeps the loads alive without a store on the common path
    else if (threadIdx.x == 0 && blockIdx.x == 0) atomicAdd(sink, acc & 1ull);
}

}  // namespace cgk
