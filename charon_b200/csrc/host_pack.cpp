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

// ---- two bits per base -------------------------------------------------------------------------------------------------
inline uint8_t code2_of(uint8_t c, bool &exc) {
    const uint8_t r = rank_of(c);
    exc = r == 3;
    return r == 4 ? 3 : (r == 3 ? 0 : r);
}
uint64_t pack2_scalar(const uint8_t *src, uint64_t begin, uint64_t end, uint8_t *dst, uint64_t pos0, uint64_t *exc, uint64_t cap, uint64_t n_exc) {
    for (uint64_t i = begin; i < end; i += 4) {
        uint8_t b = 0;
        for (uint64_t j = 0; j < 4 && i + j < end; ++j) {
            bool x;
            b |= uint8_t(code2_of(src[i + j], x) << (2 * j));
            if (x) { if (n_exc < cap) exc[n_exc] = pos0 + i + j; ++n_exc; }
        }
        dst[i >> 2] = b;
    }
    return n_exc;
}
#if defined(__x86_64__)
// 64 characters -> 16 bytes.  Ranks as in pack_avx512 (A0 C1 G2 N3 T4); code = rank - (rank >> 2) maps T to 3, and the lanes
// whose rank is 3 (N) are the exceptions: cleared in the stream, reported through the mask.
__attribute__((target("avx512f,avx512bw"))) uint64_t pack2_avx512(const uint8_t *src, uint64_t n, uint8_t *dst, uint64_t pos0, uint64_t *exc,
                                                                  uint64_t cap, uint64_t &i_out) {
    const __m512i lo_tbl = _mm512_broadcast_i32x4(_mm_setr_epi8(0, 1, 0, 2, 8, 8, 0, 4, 0, 0, 0, 0, 0, 0, 0, 0));
    const __m512i hi_tbl = _mm512_broadcast_i32x4(_mm_setr_epi8(0, 0, 0, 0, 7, 8, 7, 8, 0, 0, 0, 0, 0, 0, 0, 0));
    const __m512i c2_tbl = _mm512_broadcast_i32x4(_mm_setr_epi8(0, 0, 1, 0, 2, 0, 0, 0, 3, 0, 0, 0, 0, 0, 0, 0));   // one-hot class -> 2-bit code (0 for none)
    const __m512i nib = _mm512_set1_epi8(0x0F), m1 = _mm512_set1_epi16(0x0401), m2 = _mm512_set1_epi32(0x00100001);
    uint64_t n_exc = 0, i = 0;
    for (; i + 64 <= n; i += 64) {
        const __m512i c = _mm512_loadu_si512(reinterpret_cast<const void *>(src + i));
        const __m512i lo = _mm512_shuffle_epi8(lo_tbl, _mm512_and_si512(c, nib));
        const __m512i hi = _mm512_shuffle_epi8(hi_tbl, _mm512_and_si512(_mm512_srli_epi16(c, 4), nib));
        const __m512i cls = _mm512_and_si512(lo, hi);                       // 1 A, 2 C, 4 G, 8 T/U, 0 anything else
        uint64_t bad = _mm512_cmpeq_epi8_mask(cls, _mm512_setzero_si512());
        const __m512i code = _mm512_shuffle_epi8(c2_tbl, cls);
        const __m512i p4 = _mm512_maddubs_epi16(code, m1);                  // 16-bit lanes: c0 + 4 c1
        const __m512i p8 = _mm512_madd_epi16(p4, m2);                       // 32-bit lanes: (c0 + 4 c1) + 16 (c2 + 4 c3)
        _mm_storeu_si128(reinterpret_cast<__m128i *>(dst + (i >> 2)), _mm512_cvtepi32_epi8(p8));
        while (bad) {
            const unsigned j = unsigned(__builtin_ctzll(bad));
            bad &= bad - 1;
            if (n_exc < cap) exc[n_exc] = pos0 + i + j;
            ++n_exc;
        }
    }
    i_out = i;
    return n_exc;
}
__attribute__((target("avx2"))) uint64_t pack2_avx2(const uint8_t *src, uint64_t n, uint8_t *dst, uint64_t pos0, uint64_t *exc, uint64_t cap,
                                                    uint64_t &i_out) {
    const __m256i lo_tbl = _mm256_broadcastsi128_si256(_mm_setr_epi8(0, 1, 0, 2, 8, 8, 0, 4, 0, 0, 0, 0, 0, 0, 0, 0));
    const __m256i hi_tbl = _mm256_broadcastsi128_si256(_mm_setr_epi8(0, 0, 0, 0, 7, 8, 7, 8, 0, 0, 0, 0, 0, 0, 0, 0));
    const __m256i c2_tbl = _mm256_broadcastsi128_si256(_mm_setr_epi8(0, 0, 1, 0, 2, 0, 0, 0, 3, 0, 0, 0, 0, 0, 0, 0));
    const __m256i nib = _mm256_set1_epi8(0x0F), m1 = _mm256_set1_epi16(0x0401), m2 = _mm256_set1_epi32(0x00100001);
    const __m256i gather = _mm256_setr_epi8(0, 4, 8, 12, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, 0, 4, 8, 12, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1);
    uint64_t n_exc = 0, i = 0;
    for (; i + 32 <= n; i += 32) {
        const __m256i c = _mm256_loadu_si256(reinterpret_cast<const __m256i *>(src + i));
        const __m256i lo = _mm256_shuffle_epi8(lo_tbl, _mm256_and_si256(c, nib));
        const __m256i hi = _mm256_shuffle_epi8(hi_tbl, _mm256_and_si256(_mm256_srli_epi16(c, 4), nib));
        const __m256i cls = _mm256_and_si256(lo, hi);
        uint32_t bad = uint32_t(_mm256_movemask_epi8(_mm256_cmpeq_epi8(cls, _mm256_setzero_si256())));
        const __m256i p8 = _mm256_madd_epi16(_mm256_maddubs_epi16(_mm256_shuffle_epi8(c2_tbl, cls), m1), m2);
        const __m256i q = _mm256_shuffle_epi8(p8, gather);                  // 4 bytes at the bottom of each 128-bit half
        const uint32_t a = uint32_t(_mm256_extract_epi32(q, 0)), b = uint32_t(_mm256_extract_epi32(q, 4));
        const uint64_t w = uint64_t(a) | (uint64_t(b) << 32);
        __builtin_memcpy(dst + (i >> 2), &w, 8);
        while (bad) {
            const unsigned j = unsigned(__builtin_ctz(bad));
            bad &= bad - 1;
            if (n_exc < cap) exc[n_exc] = pos0 + i + j;
            ++n_exc;
        }
    }
    i_out = i;
    return n_exc;
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
    std::vector<uint64_t> *exc = nullptr;   // two bits per base when set: one exception list per block
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
        if (j.exc) {
            std::vector<uint64_t> &e = j.exc[b];
            e.resize(64);
            uint64_t k = pack2_range(j.src + lo, hi - lo, j.dst + (lo >> 2), lo, e.data(), e.size());
            if (k > e.size()) { e.resize(k); k = pack2_range(j.src + lo, hi - lo, j.dst + (lo >> 2), lo, e.data(), e.size()); }
            e.resize(k);
        } else
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

uint64_t pack2_range(const uint8_t *src, uint64_t n, uint8_t *dst, uint64_t pos0, uint64_t *exc, uint64_t cap) {
    uint64_t done = 0, n_exc = 0;
#if defined(__x86_64__)
    if (have_avx512()) n_exc = pack2_avx512(src, n, dst, pos0, exc, cap, done);
    else if (have_avx2()) n_exc = pack2_avx2(src, n, dst, pos0, exc, cap, done);
#endif
    return pack2_scalar(src, done, n, dst, pos0, exc, cap, n_exc);
}

int threads() { std::lock_guard<std::mutex> g(g_pool_mu); return resolve_threads(); }

void set_threads(int n) {
    std::lock_guard<std::mutex> g(g_pool_mu);
    if (n <= 0 || n == g_threads) return;
    g_threads = n;
    g_pool.reset();
}

static void pack_job(const uint8_t *src, uint8_t *dst, uint64_t n, uint64_t chunk, const std::function<void(uint64_t, uint64_t)> &on_chunk,
                     std::vector<uint64_t> *exc_out) {
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
    std::vector<std::vector<uint64_t>> block_exc(exc_out ? n_blocks : 0);
    Job j;
    {
        std::lock_guard<std::mutex> g(p->mu);
        ++p->generation;
        j.src = src; j.dst = dst; j.n = n; j.n_blocks = n_blocks; j.gen = p->generation; j.done = p->done.get();
        j.exc = exc_out ? block_exc.data() : nullptr;
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
    if (exc_out) {
        exc_out->clear();
        for (auto &e : block_exc) exc_out->insert(exc_out->end(), e.begin(), e.end());
    }
}

void pack_chunked(const uint8_t *src, uint8_t *dst, uint64_t n, uint64_t chunk, const std::function<void(uint64_t, uint64_t)> &on_chunk) {
    pack_job(src, dst, n, chunk, on_chunk, nullptr);
}
void pack2_chunked(const uint8_t *src, uint8_t *dst, uint64_t n, uint64_t chunk, const std::function<void(uint64_t, uint64_t)> &on_chunk,
                   std::vector<uint64_t> &exc) {
    pack_job(src, dst, n, chunk, on_chunk, &exc);
}

}  // namespace cgh
