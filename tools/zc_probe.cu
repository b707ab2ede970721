// tools/zc_probe.cu — stand-alone probe for the end-to-end path: how should 0.27 GB of packed bases per step reach a kernel that
// is busy with random filter gathers?
//   (1) cudaMemcpyAsync H2D alone                      (2) the same next to the gather kernel (what cg_submit_batch does today)
//   (3) gather kernel alone                            (4) ONE kernel: every warp streams its own 2 KB pieces straight from pinned
//                                                          host memory (cp.async, one piece ahead) while it gathers
//   (5) the streaming part of (4) alone
//   build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/bin/zc_probe tools/zc_probe.cu
//   zc_probe [filter_MB=858] [host_MB=266] [gathers_per_piece=2200]   (result of round 2: profiles/zc_probe_r02.log)
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

__device__ __forceinline__ uint64_t mix(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull; x = (x ^ (x >> 27)) * 0x94D049BB133111EBull; return x ^ (x >> 31);
}

constexpr int PIECE = 2048;          // bytes per piece = one 8 kb read at 2 bits per base
constexpr int WARPS = 12;

// STREAM: fetch pieces from `src` (host or device memory) one ahead into shared memory; GATHER: g random 4-byte loads per piece
template <bool STREAM, bool GATHER>
__global__ void __launch_bounds__(WARPS * 32, 1) probe(const uint4 *src, uint64_t n_pieces, const uint32_t *filt, uint64_t words, uint32_t g,
                                                      unsigned int *ticket, unsigned long long *sink) {
    __shared__ __align__(16) unsigned char buf[WARPS][2][PIECE];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    unsigned long long acc = 0;
    auto take = [&]() -> uint64_t { unsigned int t = 0; if (lane == 0) t = atomicAdd(ticket, 1u); return __shfl_sync(0xFFFFFFFFu, t, 0); };
    auto fetch = [&](uint64_t p, int slot) {
        if (STREAM && p < n_pieces) {
            const uint32_t d = uint32_t(__cvta_generic_to_shared(&buf[warp][slot][0]));
#pragma unroll
            for (int q = 0; q < PIECE / 16 / 32; ++q)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d + (q * 32 + lane) * 16), "l"(src + p * (PIECE / 16) + q * 32 + lane) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    uint64_t cur = take(), nxt = take();
    int slot = 0;
    fetch(cur, 0);
    while (cur < n_pieces) {
        fetch(nxt, slot ^ 1);
        if (GATHER) {
            for (uint32_t i = lane; i < g; i += 32 * 4) {
                uint32_t x[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const uint64_t w = __umul64hi(mix(cur * 4096 + i + u * 32), words);
                    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(x[u]) : "l"(filt + w));
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) acc += __popc(x[u]);
            }
        }
        asm volatile("cp.async.wait_group 1;" ::: "memory");
        __syncwarp();
        if (STREAM) {
            const uint4 *b = reinterpret_cast<const uint4 *>(&buf[warp][slot][0]);
#pragma unroll
            for (int q = 0; q < PIECE / 16 / 32; ++q) { const uint4 v = b[q * 32 + lane]; acc += v.x ^ v.y ^ v.z ^ v.w; }
        }
        __syncwarp();
        cur = nxt; nxt = take(); slot ^= 1;
    }
    if (acc == 0x7FFFFFFFFFFFFFFFull) *sink = acc;
}

int main(int argc, char **argv) {
    const uint64_t fmb = argc > 1 ? std::strtoull(argv[1], nullptr, 10) : 858, hmb = argc > 2 ? std::strtoull(argv[2], nullptr, 10) : 266;
    const uint32_t g = argc > 3 ? uint32_t(std::atoi(argv[3])) : 2200;
    const uint64_t words = (fmb << 20) / 4, hbytes = hmb << 20, n_pieces = hbytes / PIECE;
    int sms = 0; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    uint32_t *filt; unsigned char *hsrc, *dsrc; unsigned int *ticket; unsigned long long *sink;
    CK(cudaMalloc(&filt, words * 4)); CK(cudaMemset(filt, 0x5A, words * 4));
    CK(cudaHostAlloc(&hsrc, hbytes, cudaHostAllocDefault)); for (uint64_t i = 0; i < hbytes; i += 4096) hsrc[i] = (unsigned char)i;
    CK(cudaMalloc(&dsrc, hbytes)); CK(cudaMalloc(&ticket, 4)); CK(cudaMalloc(&sink, 8));
    cudaStream_t s1, s2; CK(cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking)); CK(cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking));
    cudaEvent_t a, b, c, d; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b)); CK(cudaEventCreate(&c)); CK(cudaEventCreate(&d));
    float ms = 0, ms2 = 0;
    for (int rep = 0; rep < 3; ++rep) {
        // (1) copy alone
        CK(cudaEventRecord(a, s1)); CK(cudaMemcpyAsync(dsrc, hsrc, hbytes, cudaMemcpyHostToDevice, s1)); CK(cudaEventRecord(b, s1));
        CK(cudaStreamSynchronize(s1)); CK(cudaEventElapsedTime(&ms, a, b));
        std::printf("rep %d (1) H2D alone                      %.3f ms  %.1f GB/s\n", rep, ms, hbytes / ms / 1e6);
        // (3) gathers alone (bases from device memory, as today)
        CK(cudaMemsetAsync(ticket, 0, 4, s2)); CK(cudaEventRecord(c, s2));
        probe<true, true><<<sms, WARPS * 32, 0, s2>>>(reinterpret_cast<const uint4 *>(dsrc), n_pieces, filt, words, g, ticket, sink);
        CK(cudaEventRecord(d, s2)); CK(cudaStreamSynchronize(s2)); CK(cudaEventElapsedTime(&ms, c, d));
        std::printf("rep %d (3) gathers + device-resident pieces %.3f ms\n", rep, ms);
        // (2) both at once on two streams
        CK(cudaMemsetAsync(ticket, 0, 4, s2)); CK(cudaStreamSynchronize(s2));
        CK(cudaEventRecord(c, s2));
        probe<true, true><<<sms, WARPS * 32, 0, s2>>>(reinterpret_cast<const uint4 *>(dsrc), n_pieces, filt, words, g, ticket, sink);
        CK(cudaEventRecord(d, s2));
        CK(cudaEventRecord(a, s1)); CK(cudaMemcpyAsync(dsrc, hsrc, hbytes, cudaMemcpyHostToDevice, s1)); CK(cudaEventRecord(b, s1));
        CK(cudaDeviceSynchronize()); CK(cudaEventElapsedTime(&ms, a, b)); CK(cudaEventElapsedTime(&ms2, c, d));
        std::printf("rep %d (2) together: H2D %.3f ms (%.1f GB/s), kernel %.3f ms\n", rep, ms, hbytes / ms / 1e6, ms2);
        // (4) one kernel, pieces straight from pinned host memory
        CK(cudaMemsetAsync(ticket, 0, 4, s2)); CK(cudaEventRecord(c, s2));
        probe<true, true><<<sms, WARPS * 32, 0, s2>>>(reinterpret_cast<const uint4 *>(hsrc), n_pieces, filt, words, g, ticket, sink);
        CK(cudaEventRecord(d, s2)); CK(cudaStreamSynchronize(s2)); CK(cudaEventElapsedTime(&ms, c, d));
        std::printf("rep %d (4) gathers + pieces from host     %.3f ms  (%.1f GB/s over PCIe)\n", rep, ms, hbytes / ms / 1e6);
        // (5) the stream from host alone
        CK(cudaMemsetAsync(ticket, 0, 4, s2)); CK(cudaEventRecord(c, s2));
        probe<true, false><<<sms, WARPS * 32, 0, s2>>>(reinterpret_cast<const uint4 *>(hsrc), n_pieces, filt, words, g, ticket, sink);
        CK(cudaEventRecord(d, s2)); CK(cudaStreamSynchronize(s2)); CK(cudaEventElapsedTime(&ms, c, d));
        std::printf("rep %d (5) pieces from host alone          %.3f ms  (%.1f GB/s over PCIe)\n", rep, ms, hbytes / ms / 1e6);
    }
    return 0;
}
