// tools/gather_gran.cu — stand-alone probe: how many DRAM bytes does one random 4-byte gather cost on B200, and can the
// program change it?  (profiles/gather_granularity_r01.md found ~116 B whatever the load instruction.)
//   gather_gran <limit_bytes|-1> <mode> [MB=858] [loads=60000000]
// limit: cudaLimitMaxL2FetchGranularity set BEFORE the context exists (-1: untouched).
// mode 0: ld.global.nc.u32   1: tex1Dfetch through a texture object   2: cp.async.bulk (TMA 1-D, 16 B) into shared memory
//      3: 16-byte cp.async.cg   4: ld.global.nc.L2::64B  5: ld.global.nc.u8
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

__device__ __forceinline__ uint64_t mix(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull; x = (x ^ (x >> 27)) * 0x94D049BB133111EBull; return x ^ (x >> 31);
}

template <int MODE>
__global__ void __launch_bounds__(256) gather(const uint32_t *buf, cudaTextureObject_t tex, uint64_t words, uint64_t n, unsigned long long *sink) {
    __shared__ __align__(16) unsigned char land[256 * 4 * 16];
    __shared__ __align__(8) unsigned long long bar;
    unsigned long long acc = 0;
    const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
    if (MODE == 2) {
        if (threadIdx.x == 0) {
            const uint32_t b = uint32_t(__cvta_generic_to_shared(&bar));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(b), "r"(1));
        }
        __syncthreads();
    }
    uint32_t phase = 0;
    for (uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; i < n; i += stride * 4) {
        uint32_t x[4];
        if (MODE == 2) {
            const uint32_t b = uint32_t(__cvta_generic_to_shared(&bar));
            if (threadIdx.x == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(256 * 4 * 16) : "memory");
            __syncthreads();
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const uint64_t w = __umul64hi(mix(i + u * stride), words) & ~3ull;
                const uint32_t d = uint32_t(__cvta_generic_to_shared(land + (threadIdx.x * 4 + u) * 16));
                asm volatile("cp.async.bulk.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1], 16, [%2];" ::"r"(d), "l"(buf + w), "r"(b) : "memory");
            }
            uint32_t done = 0;
            while (!done) asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(b), "r"(phase) : "memory");
            phase ^= 1;
#pragma unroll
            for (int u = 0; u < 4; ++u) x[u] = *reinterpret_cast<const uint32_t *>(land + (threadIdx.x * 4 + u) * 16);
            __syncthreads();
        } else if (MODE == 3) {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const uint64_t w = __umul64hi(mix(i + u * stride), words) & ~3ull;
                const uint32_t d = uint32_t(__cvta_generic_to_shared(land + (threadIdx.x * 4 + u) * 16));
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(buf + w) : "memory");
            }
            asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
#pragma unroll
            for (int u = 0; u < 4; ++u) x[u] = *reinterpret_cast<const uint32_t *>(land + (threadIdx.x * 4 + u) * 16);
        } else {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const uint64_t w = __umul64hi(mix(i + u * stride), words);
                if (MODE == 0) asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(x[u]) : "l"(buf + w));
                else if (MODE == 1) x[u] = tex1Dfetch<unsigned int>(tex, int(w));
                else if (MODE == 4) asm volatile("ld.global.nc.L1::no_allocate.L2::64B.u32 %0, [%1];" : "=r"(x[u]) : "l"(buf + w));
                else asm volatile("ld.global.nc.L1::no_allocate.u8 %0, [%1];" : "=r"(x[u]) : "l"(reinterpret_cast<const unsigned char *>(buf + w)));
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) acc += __popc(x[u]);
    }
    if (acc == 0x7FFFFFFFFFFFFFFFull) *sink = acc;
}

int main(int argc, char **argv) {
    const int limit = argc > 1 ? std::atoi(argv[1]) : -1;
    const int mode = argc > 2 ? std::atoi(argv[2]) : 0;
    const uint64_t mb = argc > 3 ? std::strtoull(argv[3], nullptr, 10) : 858;
    const uint64_t n = argc > 4 ? std::strtoull(argv[4], nullptr, 10) : 60000000ull;
    if (limit >= 0) CK(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, size_t(limit)));
    size_t got = 0;
    CK(cudaDeviceGetLimit(&got, cudaLimitMaxL2FetchGranularity));
    const uint64_t words = (mb << 20) / 4;
    uint32_t *buf = nullptr; unsigned long long *sink = nullptr;
    CK(cudaMalloc(&buf, words * 4)); CK(cudaMalloc(&sink, 8));
    CK(cudaMemset(buf, 0x5A, words * 4)); CK(cudaMemset(sink, 0, 8));
    cudaTextureObject_t tex = 0;
    if (mode == 1) {
        // a linear texture is limited to 2^27 texels: bind uint4 texels? keep u32 texels and cap the span
        cudaResourceDesc rd{}; rd.resType = cudaResourceTypeLinear; rd.res.linear.devPtr = buf; rd.res.linear.desc = cudaCreateChannelDesc<unsigned int>();
        rd.res.linear.sizeInBytes = words * 4;
        cudaTextureDesc td{}; td.readMode = cudaReadModeElementType;
        CK(cudaCreateTextureObject(&tex, &rd, &td, nullptr));
    }
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    const int grid = 148 * 8;
    float best = 1e9f;
    for (int r = 0; r < 4; ++r) {
        CK(cudaEventRecord(a));
        switch (mode) {
            case 0: gather<0><<<grid, 256>>>(buf, tex, words, n, sink); break;
            case 1: gather<1><<<grid, 256>>>(buf, tex, words < (1ull << 27) ? words : (1ull << 27), n, sink); break;
            case 2: gather<2><<<grid, 256>>>(buf, tex, words, n, sink); break;
            case 3: gather<3><<<grid, 256>>>(buf, tex, words, n, sink); break;
            case 4: gather<4><<<grid, 256>>>(buf, tex, words, n, sink); break;
            default: gather<5><<<grid, 256>>>(buf, tex, words, n, sink); break;
        }
        CK(cudaGetLastError());
        CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
        float ms = 0; CK(cudaEventElapsedTime(&ms, a, b));
        if (r && ms < best) best = ms;
    }
    std::printf("limit_req %d limit_got %zu mode %d MB %llu loads %llu : %.3f ms  %.2f G loads/s\n", limit, got, mode, (unsigned long long)mb,
                (unsigned long long)n, best, double(n) / best / 1e6);
    return 0;
}
