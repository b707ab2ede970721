// charon_b200/csrc/host_pack.cpp — ASCII -> 4-bit rank packing on host threads (see host_pack.hpp).  Plain C++17 + AVX2 /
// AVX-512BW intrinsics behind a run-time CPU check; compiled by the host compiler and linked into libcharon_gpu.so.
#include "host_pack.hpp"

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstdlib>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#if defined(__x86_64__)
#include <immintrin.h>
#endif

namespace cgh {

namespace {

#ifndef PACK_NT_DEFAULT
#define PACK_NT_DEFAULT true
#endif

void pack_scalar(const uint8_t *src, uint8_t *dst, uint64_t begin, uint64_t end) {
    uint64_t i = begin;
    for (; i + 1 < end; i += 2) dst[i >> 1] = uint8_t(rank_of(src[i]) | (rank_of(src[i + 1]) << 4));
    if (i < end) dst[i >> 1] = rank_of(src[i]);
}

// Streaming (non-temporal) stores for the packed buffer: CHARON_PACK_NT=1/0.
bool use_nt_stores() {
    static const bool v = [] { const char *e = std::getenv("CHARON_PACK_NT"); return e ? std::atoi(e) != 0 : PACK_NT_DEFAULT; }();
    return v;
}

#if defined(__x86_64__)
// 64 characters -> 32 bytes.  Classification by two nibble look-ups (vpshufb): the low-nibble table says which of A/C/G/T|U
// the character could be, the high-nibble table which of them its column of the ASCII chart allows (4x/6x: A C G, 5x/7x:
// T U); their AND is one-hot or zero and indexes the rank table (0 -> N).  Adjacent ranks are fused with one multiply-add
// (even * 1 + odd * 16) and narrowed to bytes.
__attribute__((target("avx2"))) void pack_avx2(const uint8_t *src, uint8_t *dst, uint64_t begin, uint64_t end) {
    const __m256i lo_tbl = _mm256_broadcastsi128_si256(_mm_setr_epi8(0, 1, 0, 2, 8, 8, 0, 4, 0, 0, 0, 0, 0, 0, 0, 0));
    const __m256i hi_tbl = _mm256_broadcastsi128_si256(_mm_setr_epi8(0, 0, 0, 0, 7, 8, 7, 8, 0, 0, 0, 0, 0, 0, 0, 0));
    const __m256i rk_tbl = _mm256_broadcastsi128_si256(_mm_setr_epi8(3, 0, 1, 3, 2, 3, 3, 3, 4, 3, 3, 3, 3, 3, 3, 3));
    const __m256i nib = _mm256_set1_epi8(0x0F), mul = _mm256_set1_epi16(0x1001);
    auto ranks = [&](const uint8_t *p) __attribute__((target("avx2"))) {
        const __m256i c = _mm256_loadu_si256(reinterpret_cast<const __m256i *>(p));
        const __m256i lo = _mm256_shuffle_epi8(lo_tbl, _mm256_and_si256(c, nib));
        const __m256i hi = _mm256_shuffle_epi8(hi_tbl, _mm256_and_si256(_mm256_srli_epi16(c, 4), nib));
        return _mm256_maddubs_epi16(_mm256_shuffle_epi8(rk_tbl, _mm256_and_si256(lo, hi)), mul);
    };
    uint64_t i = begin;
    // the packed buffer is written once and next read by the copy engine: streaming stores when the alignment allows
    if (use_nt_stores() && (reinterpret_cast<uintptr_t>(dst + (i >> 1)) & 31u) == 0) {
        for (; i + 64 <= end; i += 64) {
            const __m256i q = _mm256_permute4x64_epi64(_mm256_packus_epi16(ranks(src + i), ranks(src + i + 32)), 0xD8);
            _mm256_stream_si256(reinterpret_cast<__m256i *>(dst + (i >> 1)), q);
        }
        _mm_sfence();
    }
    for (; i + 64 <= end; i += 64) {
        const __m256i q = _mm256_permute4x64_epi64(_mm256_packus_epi16(ranks(src + i), ranks(src + i + 32)), 0xD8);
        _mm256_storeu_si256(reinterpret_cast<__m256i *>(dst + (i >> 1)), q);
    }
    if (i < end) pack_scalar(src, dst, i, end);
}

// Same algorithm on 512-bit registers: 128 characters -> 64 bytes per iteration.
__attribute__((target("avx512f,avx512bw"))) void pack_avx512(const uint8_t *src, uint8_t *dst, uint64_t begin, uint64_t end) {
    const __m512i lo_tbl = _mm512_broadcast_i32x4(_mm_setr_epi8(0, 1, 0, 2, 8, 8, 0, 4, 0, 0, 0, 0, 0, 0, 0, 0));
    const __m512i hi_tbl = _mm512_broadcast_i32x4(_mm_setr_epi8(0, 0, 0, 0, 7, 8, 7, 8, 0, 0, 0, 0, 0, 0, 0, 0));
    const __m512i rk_tbl = _mm512_broadcast_i32x4(_mm_setr_epi8(3, 0, 1, 3, 2, 3, 3, 3, 4, 3, 3, 3, 3, 3, 3, 3));
    const __m512i nib = _mm512_set1_epi8(0x0F), mul = _mm512_set1_epi16(0x1001);
    const __m512i order = _mm512_setr_epi64(0, 2, 4, 6, 1, 3, 5, 7);   // packus interleaves the two sources per 128-bit lane
    auto ranks = [&](const uint8_t *p) __attribute__((target("avx512f,avx512bw"))) {
        const __m512i c = _mm512_loadu_si512(reinterpret_cast<const void *>(p));
        const __m512i lo = _mm512_shuffle_epi8(lo_tbl, _mm512_and_si512(c, nib));
        const __m512i hi = _mm512_shuffle_epi8(hi_tbl, _mm512_and_si512(_mm512_srli_epi16(c, 4), nib));
        return _mm512_maddubs_epi16(_mm512_shuffle_epi8(rk_tbl, _mm512_and_si512(lo, hi)), mul);
    };
    uint64_t i = begin;
    if (use_nt_stores() && (reinterpret_cast<uintptr_t>(dst + (i >> 1)) & 63u) == 0) {
        for (; i + 128 <= end; i += 128) {
            const __m512i q = _mm512_permutexvar_epi64(order, _mm512_packus_epi16(ranks(src + i), ranks(src + i + 64)));
            _mm512_stream_si512(reinterpret_cast<__m512i *>(dst + (i >> 1)), q);
        }
        _mm_sfence();
    }
    for (; i + 128 <= end; i += 128) {
        const __m512i q = _mm512_permutexvar_epi64(order, _mm512_packus_epi16(ranks(src + i), ranks(src + i + 64)));
        _mm512_storeu_si512(reinterpret_cast<void *>(dst + (i >> 1)), q);
    }
    if (i < end) pack_avx2(src, dst, i, end);
}
#endif

bool have_avx512() {
#if defined(__x86_64__)
    static const bool v = __builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw") && !std::getenv("CHARON_NO_AVX512");
    return v;
#else
    return false;
#endif
}

bool have_avx2() {
#if defined(__x86_64__)
    static const bool v = __builtin_cpu_supports("avx2");
    return v;
#else
    return false;
#endif
}

constexpr uint64_t BLOCK = 256 * 1024;   // bases per work item (a multiple of 64)

struct Job {
    const uint8_t *src = nullptr;
    uint8_t *dst = nullptr;
    uint64_t n = 0, n_blocks = 0, gen = 0;
    std::atomic<uint8_t> *done = nullptr;
};

// Work items are handed out through one counter tagged with the job generation (high 24 bits), so that a worker that wakes
// up late with a stale job can neither take nor burn a block of the next job.
struct Pool {
    std::vector<std::thread> workers;
    std::mutex mu, job_mu;
    std::condition_variable cv;
    uint64_t generation = 0;
    bool stop = false;
    Job job;
    std::atomic<uint64_t> next{0};
    std::unique_ptr<std::atomic<uint8_t>[]> done;
    uint64_t done_cap = 0;

    explicit Pool(int nthreads) {
        for (int t = 1; t < nthreads; ++t) workers.emplace_back([this] { loop(); });
    }
    ~Pool() {
        { std::lock_guard<std::mutex> g(mu); stop = true; ++generation; }
        cv.notify_all();
        for (auto &w : workers) w.join();
    }
    bool take_one(const Job &j) {
        uint64_t v = next.load(std::memory_order_acquire);
        for (;;) {
            if ((v >> 40) != (j.gen & 0xFFFFFFu) || (v & 0xFFFFFFFFFFull) >= j.n_blocks) return false;
            if (next.compare_exchange_weak(v, v + 1, std::memory_order_acq_rel, std::memory_order_acquire)) break;
        }
        const uint64_t b = v & 0xFFFFFFFFFFull, lo = b * BLOCK, hi = std::min(j.n, lo + BLOCK);
        pack_range(j.src, j.dst, lo, hi);
        j.done[b].store(1, std::memory_order_release);
        return true;
    }
    void loop() {
        uint64_t seen = 0;
        for (;;) {
            Job j;
            {
                std::unique_lock<std::mutex> g(mu);
                cv.wait(g, [&] { return generation != seen; });
                seen = generation;
                if (stop) return;
                j = job;
            }
            while (take_one(j)) {}
        }
    }
};

std::mutex g_pool_mu;
std::unique_ptr<Pool> g_pool;
int g_threads = 0;

int resolve_threads() {
    if (g_threads > 0) return g_threads;
    int n = 0;
    if (const char *e = std::getenv("CHARON_HOST_THREADS")) n = std::atoi(e);
    if (n <= 0) n = int(std::min<unsigned>(16u, std::max(1u, std::thread::hardware_concurrency())));
    g_threads = n;
    return n;
}

}  // namespace

void pack_range(const uint8_t *src, uint8_t *dst, uint64_t begin, uint64_t end) {
#if defined(__x86_64__)
    if (have_avx512()) { pack_avx512(src, dst, begin, end); return; }
    if (have_avx2()) { pack_avx2(src, dst, begin, end); return; }
#endif
    pack_scalar(src, dst, begin, end);
}

int threads() { std::lock_guard<std::mutex> g(g_pool_mu); return resolve_threads(); }

void set_threads(int n) {
    std::lock_guard<std::mutex> g(g_pool_mu);
    if (n <= 0 || n == g_threads) return;
    g_threads = n;
    g_pool.reset();
}

void pack_chunked(const uint8_t *src, uint8_t *dst, uint64_t n, uint64_t chunk, const std::function<void(uint64_t, uint64_t)> &on_chunk) {
    if (n == 0) return;
    chunk = std::max<uint64_t>(BLOCK, (chunk + BLOCK - 1) / BLOCK * BLOCK);
    Pool *p;
    {
        std::lock_guard<std::mutex> g(g_pool_mu);
        if (!g_pool) g_pool.reset(new Pool(resolve_threads()));
        p = g_pool.get();
    }
    std::lock_guard<std::mutex> job(p->job_mu);
    const uint64_t n_blocks = (n + BLOCK - 1) / BLOCK, per_chunk = chunk / BLOCK;
    // a block's done flag is only written by the thread that took it, and every block of a finished job has been taken and
    // completed before pack_chunked returned: the array can be reset / replaced here
    if (n_blocks > p->done_cap) { p->done.reset(new std::atomic<uint8_t>[n_blocks]); p->done_cap = n_blocks; }
    for (uint64_t b = 0; b < n_blocks; ++b) p->done[b].store(0, std::memory_order_relaxed);
    Job j;
    {
        std::lock_guard<std::mutex> g(p->mu);
        ++p->generation;
        j.src = src; j.dst = dst; j.n = n; j.n_blocks = n_blocks; j.gen = p->generation; j.done = p->done.get();
        p->job = j;
        p->next.store((j.gen & 0xFFFFFFu) << 40, std::memory_order_release);
    }
    p->cv.notify_all();
    uint64_t issued = 0;   // blocks already handed to on_chunk
    auto try_issue = [&](bool final) {
        for (;;) {
            const uint64_t hi = std::min(n_blocks, issued + per_chunk);
            if (issued >= n_blocks) return;
            if (!final && hi - issued < per_chunk && hi < n_blocks) return;
            for (uint64_t b = issued; b < hi; ++b)
                if (!p->done[b].load(std::memory_order_acquire)) return;
            on_chunk(issued * BLOCK, std::min(n, hi * BLOCK));
            issued = hi;
        }
    };
    while (p->take_one(j)) try_issue(false);
    while (issued < n_blocks) { try_issue(true); if (issued < n_blocks) std::this_thread::yield(); }
}

}  // namespace cgh
