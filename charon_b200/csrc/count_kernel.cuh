// charon_b200/csrc/count_kernel.cuh — the fused hot-path kernel (sm_100a).
//
// One CTA owns one read at a time (dynamic work queue) and walks it in tiles that live entirely in
// shared memory:
//   phase 0  ASCII -> dna5 rank (+complement) codes                    [replaces seqan3 char->dna5, SURVEY A.1]
//   phase 1  canonical base-5 rolling k-mer hash ^ seed, min(fwd, rc)   [seqan3::views::kmer_hash x2, A.1/A.2]
//   phase 2  window minimiser: van Herk prefix/suffix rightmost-argmin  [seqan3::detail::minimiser_view, A.2]
//            + emission = "tracked position changed"; exact sequential fix-up only for tiles with ties
//   phase 3  3x hash_and_fit -> row gather from the column-compacted filter in HBM -> AND -> bin mask
//            -> per-bin totals / unique counts                          [bulk_contains A.3 + read_entry.hpp:86-138]
// Reference call sites replaced: src/dehost_main.cpp:366-373, src/classify_main.cpp:173-180.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace cgk {

constexpr int NT = 128;          // threads per CTA
constexpr int TK_CAP = 3072;     // k-mers per tile held in shared memory
constexpr uint32_t INV5 = 0xCCCCCCCDu;  // 5^-1 mod 2^32 (exact division of multiples of 5)
#define CG_MAX_PARTS 16

struct CountParams {
    const uint8_t *bases;       // ASCII, one base per byte — or, when `packed`, rank nibbles (16-byte aligned buffer)
    const uint64_t *seg_off;
    const uint32_t *read_seg;   // may be null: one segment per read
    const uint64_t *seg_len;    // may be null: segment s = [seg_off[s], seg_off[s+1]); else [seg_off[s], seg_off[s]+seg_len[s])
    uint32_t n_reads;
    // minimiser parameters
    uint32_t K, Wv;             // k-mer size, window in k-mers (w - k + 1)
    uint32_t kl, kh;            // digits in the low / high 32-bit limb (kl = min(K,13))
    uint32_t p_lo_top, p_hi_top, p_kl;   // 5^(kl-1), 5^(kh-1), 5^kl
    uint64_t seed;
    // filter
    const void *filter;
    uint64_t bin_size;
    uint32_t hash_shift, hash_funs, row_shift;   // compact row width = 1 << row_shift bits
    uint32_t bins, ncat, single_pass;
    const uint8_t *bin_to_cat;  // device, 256 entries
    // outputs
    uint32_t *num_hashes, *counts, *unique_counts;
    uint32_t *work_counter;
    // read-length dispatch between the warp-per-read and the CTA-per-read kernels (0 = no dispatch)
    unsigned long long long_read_min;   // warp kernel skips reads with >= this many bases ...
    uint32_t long_only;                 // ... and the CTA kernel, when set, processes only those
    // emission capture (index construction + test hook): emitted minimiser values appended to dbg_out;
    // filter == nullptr skips the membership query entirely (capture-only launch)
    uint64_t *dbg_out;
    unsigned long long dbg_cap;
    unsigned long long *dbg_n;
    // row-range partitioned filter: part p holds rows [p*rows_per_part, (p+1)*rows_per_part) in its own
    // allocation on GPU p; remote parts are NVLink peer mappings.  n_parts == 0: `filter` is the whole (replicated) filter.
    uint32_t n_parts, my_part;
    uint32_t rows_per_part;             // rows are < 2^32 (bin_size <= 2^32 - 3)
    uint32_t async_gather;              // host-side choice of the warp kernel's AG instantiation (filter larger than L2); not read on the device
    uint32_t packed;                    // transport of `bases` (enum cg_bases_format): 0 ASCII or 2 two bits per base; 2 only in the
                                        // (19, 41) kernels.  Rank nibbles (1), and packed input of the generic kernel, are unpacked by the library.
    // CG_BASES_2BIT: ascending batch positions of the bases that are not A/C/G/T (they hash as N), and for every segment
    // the index range [exc_seg_off[s], exc_seg_off[s + 1]) of its entries.  n_exc == 0: no such base in the batch.
    const unsigned long long *exc_pos;
    const uint32_t *exc_seg_off;
    unsigned long long n_exc;
    const void *part_base[CG_MAX_PARTS];
    // One-pass counting for indexes where only a few bins belong to categories that own several bins (warp-per-read kernel,
    // more than 4 bins).  ReadEntry::get_counts needs, per category, the bin with the most hits before it can tell which
    // hits are unique; a category with ONE bin has no choice to make, so its bin is in the best-bin mask from the start
    // (`static_best`), and only the minimisers that hit a bin of a multi-bin category (`multibin`) have to wait for the end
    // of the read: their masks are parked in shared memory and resolved then.  A read that parks more masks than fit is
    // redone the two-pass way.
    uint32_t amb_mode;
    unsigned long long multibin[4], static_best[4];
};

__constant__ uint64_t c_ibf_seeds[5] = {13572355802537770549ULL, 13043817825332782213ULL, 10650232656628343401ULL,
                                        16499269484942379435ULL, 4893150838803335377ULL};

template <int NW> struct Mask { uint64_t w[NW]; };

// L2 residency control: filter rows are the hot, randomly re-referenced working set (evict_last), read bases are
// touched exactly once (evict_first) — otherwise the 1 B/base stream pushes the filter out of the L2.
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
    uint64_t p; asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p)); return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t p; asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p)); return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_normal() {
    uint64_t p; asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(p)); return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_unchanged() {
    uint64_t p; asm volatile("createpolicy.fractional.L2::evict_unchanged.b64 %0, 1.0;" : "=l"(p)); return p;
}
__device__ __forceinline__ uint32_t ldg_na_u32(const uint32_t *p, uint64_t pol) {
    uint32_t v; asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol)); return v;
}
__device__ __forceinline__ uint64_t ldg_na_u64(const uint64_t *p, uint64_t pol) {
    uint64_t v; asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.u64 %0, [%1], %2;" : "=l"(v) : "l"(p), "l"(pol)); return v;
}
__device__ __forceinline__ void ldg_na_v2u64(const uint64_t *p, uint64_t &a, uint64_t &b, uint64_t pol) {
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.u64 {%0, %1}, [%2], %3;" : "=l"(a), "=l"(b) : "l"(p), "l"(pol));
}
__device__ __forceinline__ uint4 ldg_stream_v4(const uint4 *p, uint64_t pol) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0, %1, %2, %3}, [%4], %5;"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p), "l"(pol));
    return v;
}

// remote (peer-mapped) rows: plain read-only loads, no local-L2 residency hint
template <int MODE>
__device__ __forceinline__ Mask<(MODE == 2 ? 2 : (MODE == 3 ? 4 : 1))> fetch_row_remote(const void *part, uint64_t row, uint32_t row_shift) {
    Mask<(MODE == 2 ? 2 : (MODE == 3 ? 4 : 1))> m;
    if (MODE == 0) {
        const uint64_t bit = row << row_shift;
        const uint32_t word = __ldg(reinterpret_cast<const uint32_t *>(part) + (bit >> 5));
        const uint32_t rb = 1u << row_shift;
        const uint32_t msk = rb >= 32 ? 0xFFFFFFFFu : ((1u << rb) - 1u);
        m.w[0] = (word >> (uint32_t(bit) & 31u)) & msk;
    } else if (MODE == 1) {
        m.w[0] = __ldg(reinterpret_cast<const unsigned long long *>(part) + row);
    } else if (MODE == 2) {
        const ulonglong2 a = __ldg(reinterpret_cast<const ulonglong2 *>(part) + row);
        m.w[0] = a.x; m.w[1] = a.y;
    } else {
        const ulonglong2 a = __ldg(reinterpret_cast<const ulonglong2 *>(part) + row * 2);
        const ulonglong2 b = __ldg(reinterpret_cast<const ulonglong2 *>(part) + row * 2 + 1);
        m.w[0] = a.x; m.w[1] = a.y; m.w[2] = b.x; m.w[3] = b.y;
    }
    return m;
}

// seqan3 interleaved_bloom_filter::hash_and_fit without the final * technical_bins
// bin_size < 2^32 always (include/index_arguments.hpp:20 caps it at 2^32 - 3; cg_create / cg_build_index reject more), so
// floor(h * bin_size / 2^64) needs two 32 x 32 -> 64 products instead of a full 64 x 64 high product, and a row is 32 bits.
__device__ __forceinline__ uint32_t ibf_row32(uint64_t h, uint64_t seed, uint32_t hash_shift, uint32_t bin_size) {
    h *= seed;
    h ^= h >> hash_shift;
    h *= 11400714819323198485ULL;
    const uint64_t t = uint64_t(uint32_t(h)) * bin_size;
    const uint64_t u = uint64_t(uint32_t(h >> 32)) * bin_size + (t >> 32);   // <= (2^32-1)^2 + 2^32 - 1 < 2^64
    return uint32_t(u >> 32);
}
__device__ __forceinline__ uint64_t ibf_row(uint64_t h, uint64_t seed, uint32_t hash_shift, uint64_t bin_size) {
    return ibf_row32(h, seed, hash_shift, uint32_t(bin_size));
}

// MODE 0: row width <= 32 bits (runtime 1<<row_shift), MODE 1: 64, MODE 2: 128, MODE 3: 256
template <int MODE> struct ModeTraits;
template <> struct ModeTraits<0> { static constexpr int NW = 1; };
template <> struct ModeTraits<1> { static constexpr int NW = 1; };
template <> struct ModeTraits<2> { static constexpr int NW = 2; };
template <> struct ModeTraits<3> { static constexpr int NW = 4; };

template <int MODE>
__device__ __forceinline__ Mask<ModeTraits<MODE>::NW> fetch_row(const void *filter, uint64_t row, uint32_t row_shift, uint64_t pol) {
    Mask<ModeTraits<MODE>::NW> m;
    if (MODE == 0) {
        const uint32_t r = uint32_t(row);   // 32-bit arithmetic: word index = (r << row_shift) >> 5 without forming the 64-bit bit index
        uint32_t word = ldg_na_u32(reinterpret_cast<const uint32_t *>(filter) + (r >> (5u - row_shift)), pol);
        uint32_t rb = 1u << row_shift;
        uint32_t msk = rb >= 32 ? 0xFFFFFFFFu : ((1u << rb) - 1u);
        m.w[0] = (word >> ((r << row_shift) & 31u)) & msk;
    } else if (MODE == 1) {
        m.w[0] = ldg_na_u64(reinterpret_cast<const uint64_t *>(filter) + row, pol);
    } else if (MODE == 2) {
        ldg_na_v2u64(reinterpret_cast<const uint64_t *>(filter) + row * 2, m.w[0], m.w[1], pol);
    } else {
        ldg_na_v2u64(reinterpret_cast<const uint64_t *>(filter) + row * 4, m.w[0], m.w[1], pol);
        ldg_na_v2u64(reinterpret_cast<const uint64_t *>(filter) + row * 4 + 2, m.w[2], m.w[3], pol);
    }
    return m;
}

// Partitioned filter: rows owned by other GPUs are read through their NVLink peer mappings.  Kept out of line so that
// the replicated hot path keeps its register allocation.  bulk_contains is an AND over the hash rows (include/index.hpp:
// 110-112), so the rows this GPU owns are fetched first and a remote row is only requested while some bin is still alive:
// a k-mer broken by a read error usually dies on its first row, and every request that is not made is one NVLink round trip
// (the remote 4-byte gathers are request-bound).
template <int MODE>
__device__ __noinline__ Mask<ModeTraits<MODE>::NW> bulk_contains_partitioned(const CountParams &P, uint64_t value, uint64_t pol) {
    constexpr int NW = ModeTraits<MODE>::NW;
    Mask<NW> acc;
#pragma unroll
    for (int j = 0; j < NW; ++j) acc.w[j] = ~uint64_t(0);
    uint32_t owner[5], local[5];
#pragma unroll
    for (int i = 0; i < 5; ++i)
        if (i < int(P.hash_funs)) {
            const uint32_t row = uint32_t(ibf_row(value, c_ibf_seeds[i], P.hash_shift, P.bin_size));   // rows < 2^32
            owner[i] = row / P.rows_per_part;
            local[i] = row - owner[i] * P.rows_per_part;
        }
    bool alive = true;
#pragma unroll
    for (int pass = 0; pass < 2; ++pass) {          // pass 0: rows in this GPU's own partition; pass 1: the others, one at a time
#pragma unroll
        for (int i = 0; i < 5; ++i)
            if (i < int(P.hash_funs) && alive && (owner[i] == P.my_part) == (pass == 0)) {
                const Mask<NW> r = pass == 0 ? fetch_row<MODE>(P.filter, local[i], P.row_shift, pol) : fetch_row_remote<MODE>(P.part_base[owner[i]], local[i], P.row_shift);
                uint64_t any = 0;
#pragma unroll
                for (int j = 0; j < NW; ++j) { acc.w[j] &= r.w[j]; any |= acc.w[j]; }
                alive = any != 0;
            }
    }
    return acc;
}

template <int MODE, int NH>
__device__ __forceinline__ Mask<ModeTraits<MODE>::NW> bulk_contains_n(const CountParams &P, uint64_t value, uint64_t pol) {
    constexpr int NW = ModeTraits<MODE>::NW;
    uint64_t rows[NH];
#pragma unroll
    for (int i = 0; i < NH; ++i) rows[i] = ibf_row(value, c_ibf_seeds[i], P.hash_shift, P.bin_size);
    Mask<NW> r[NH];
#pragma unroll
    for (int i = 0; i < NH; ++i) r[i] = fetch_row<MODE>(P.filter, rows[i], P.row_shift, pol);   // all loads in flight before the AND
    Mask<NW> acc = r[0];
#pragma unroll
    for (int i = 1; i < NH; ++i) {
#pragma unroll
        for (int j = 0; j < NW; ++j) acc.w[j] &= r[i].w[j];
    }
    return acc;
}

template <int MODE>
__device__ __forceinline__ Mask<ModeTraits<MODE>::NW> bulk_contains(const CountParams &P, uint64_t value, uint64_t pol) {
    if (P.n_parts) return bulk_contains_partitioned<MODE>(P, value, pol);
    // `charon index` always writes 3 hash functions (include/index_arguments.hpp:21); the others only exist for foreign indexes
    switch (P.hash_funs) {
        case 3: return bulk_contains_n<MODE, 3>(P, value, pol);
        case 1: return bulk_contains_n<MODE, 1>(P, value, pol);
        case 2: return bulk_contains_n<MODE, 2>(P, value, pol);
        case 4: return bulk_contains_n<MODE, 4>(P, value, pol);
        default: return bulk_contains_n<MODE, 5>(P, value, pol);
    }
}

struct Smem {
    uint64_t *v;        // TK_CAP seeded canonical k-mer values of the tile
    uint16_t *ppos;     // TK_CAP prefix rightmost-argmin (phase 2a/2b); reused as the emission list (phase 2c/3)
    uint16_t *rw;       // TK_CAP rightmost argmin of each window
    uint8_t *rk;        // TK_CAP + 64 base codes: rank | complement << 4
    uint8_t *lut;       // 256 ASCII -> code
    uint32_t *cnt;      // 256 per-bin totals
    uint32_t *uniq;     // 256 per-category unique counts
    unsigned long long *best;  // 256 per-category (total << 8 | 255 - bin) keys
    uint8_t *b2c;       // 256
    unsigned long long *bestmask;  // 4
};
constexpr size_t SMEM_BYTES = size_t(TK_CAP) * 8 + size_t(TK_CAP) * 2 * 2 + (TK_CAP + 64) + 256 + 256 * 4 * 2 + 256 * 8 + 256 + 4 * 8;

__device__ __forceinline__ Smem carve(unsigned char *raw) {
    Smem s;
    s.v = reinterpret_cast<uint64_t *>(raw);
    s.ppos = reinterpret_cast<uint16_t *>(s.v + TK_CAP);
    s.rw = s.ppos + TK_CAP;
    s.rk = reinterpret_cast<uint8_t *>(s.rw + TK_CAP);
    s.lut = s.rk + TK_CAP + 64;
    s.cnt = reinterpret_cast<uint32_t *>(s.lut + 256);
    s.uniq = s.cnt + 256;
    s.best = reinterpret_cast<unsigned long long *>(s.uniq + 256);
    s.b2c = reinterpret_cast<uint8_t *>(s.best + 256);
    s.bestmask = reinterpret_cast<unsigned long long *>(s.b2c + 256);
    return s;
}
static_assert((size_t(TK_CAP) * 8 + size_t(TK_CAP) * 4 + TK_CAP + 64 + 256 + 2048) % 8 == 0, "best[] must be 8-byte aligned");

__device__ __forceinline__ uint8_t base_code(unsigned c) {
    // dna5 rank: A0 C1 G2 N3 T4 (U->T, everything else -> N); complement rank in the high nibble
    unsigned u = c & 0xDFu;  // upper-case
    unsigned r = 3;
    if (u == 'A') r = 0; else if (u == 'C') r = 1; else if (u == 'G') r = 2; else if (u == 'T' || u == 'U') r = 4;
    unsigned comp = (0x03124u >> (r * 4)) & 0xFu;
    return uint8_t(r | (comp << 4));
}

template <int MODE>
__global__ void __launch_bounds__(NT) count_kernel(const __grid_constant__ CountParams P) {
    constexpr int NW = ModeTraits<MODE>::NW;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const Smem S = carve(smem_raw);
    __shared__ uint32_t s_read, s_nlist, s_susp, s_nhash;
    __shared__ long long s_carry_pos;
    __shared__ uint64_t s_carry_val;
    __shared__ unsigned long long s_dbg_base;

    const uint32_t tid = threadIdx.x;
    const uint32_t lane = tid & 31u;
    const uint32_t K = P.K, kl = P.kl, kh = P.kh;
    const bool small_bins = (NW == 1) && (P.bins <= 4);
    const uint64_t pol_filter = l2_policy_evict_last();

    for (uint32_t i = tid; i < 256; i += NT) { S.lut[i] = base_code(i); S.b2c[i] = P.bin_to_cat[i]; }

    for (;;) {
        __syncthreads();
        if (tid == 0) s_read = atomicAdd(P.work_counter, 1u);
        __syncthreads();
        const uint32_t r = s_read;
        if (r >= P.n_reads) break;
        for (uint32_t i = tid; i < 256; i += NT) { S.cnt[i] = 0; S.uniq[i] = 0; S.best[i] = 0; }
        if (tid < 4) S.bestmask[tid] = P.single_pass ? ~0ull : 0ull;
        if (tid == 0) s_nhash = 0;
        uint32_t rc_cnt[4] = {0, 0, 0, 0}, rc_unq[4] = {0, 0, 0, 0};   // register counters (bins <= 4)
        const uint32_t seg0 = P.read_seg ? P.read_seg[r] : r;
        const uint32_t seg1 = P.read_seg ? P.read_seg[r + 1] : r + 1;
        const int npass = P.single_pass ? 1 : 2;

        for (int pass = 0; pass < npass; ++pass) {
            const bool do_tot = (pass == 0);
            const bool do_unq = P.single_pass || pass == 1;
            __syncthreads();
            unsigned long long bm[NW];
#pragma unroll
            for (int j = 0; j < NW; ++j) bm[j] = S.bestmask[j];

            for (uint32_t seg = seg0; seg < seg1; ++seg) {
                const uint64_t sbase = P.seg_off[seg];
                const uint64_t slen = P.seg_len ? P.seg_len[seg] : P.seg_off[seg + 1] - sbase;
                if (slen < K) continue;
                const uint64_t nkm = slen - K + 1;
                const uint32_t Wp = uint32_t(nkm < P.Wv ? nkm : P.Wv);
                const uint64_t nwin = nkm - Wp + 1;
                uint32_t Tmax = (NT - 1) * Wp;
                if (Tmax > TK_CAP - Wp + 1) Tmax = TK_CAP - Wp + 1;
                bool have_prev = false;
                long long prev_pos = 0;
                uint64_t prev_val = 0;

                for (uint64_t g0 = 0; g0 < nwin; g0 += Tmax) {
                    const uint32_t Tw = uint32_t(nwin - g0 < Tmax ? nwin - g0 : Tmax);
                    const uint32_t nk = Tw + Wp - 1;
                    const uint32_t nb = nk + K - 1;
                    // ---- phase 0: load + encode -------------------------------------------------
                    const uint8_t *src = P.bases + sbase + g0;
                    for (uint32_t i = tid; i < nb; i += NT) S.rk[i] = S.lut[src[i]];
                    if (tid == 0) { s_nlist = 0; s_susp = 0; }
                    __syncthreads();
                    // ---- phase 1: rolling canonical hash ----------------------------------------
                    {
                        const uint32_t CH = ((nk + NT - 1) / NT) | 1u;   // odd stride: conflict-free 64-bit smem stores
                        uint32_t i = tid * CH;
                        if (i < nk) {
                            const uint32_t i1 = (i + CH < nk) ? i + CH : nk;
                            uint32_t Hi = 0, Lo = 0, RHi = 0, RLo = 0;
                            for (uint32_t j = 0; j < kh; ++j) Hi = Hi * 5u + (S.rk[i + j] & 15u);
                            for (uint32_t j = kh; j < K; ++j) Lo = Lo * 5u + (S.rk[i + j] & 15u);
                            for (uint32_t j = kl; j-- > 0;) RLo = RLo * 5u + (S.rk[i + j] >> 4);
                            for (uint32_t j = K; j-- > kl;) RHi = RHi * 5u + (S.rk[i + j] >> 4);
                            for (;;) {
                                uint64_t f = (uint64_t(Hi) * P.p_kl + Lo) ^ P.seed;
                                uint64_t rc = (uint64_t(RHi) * P.p_kl + RLo) ^ P.seed;
                                S.v[i] = f < rc ? f : rc;
                                if (++i >= i1) break;
                                const uint32_t b0 = S.rk[i - 1], bh = S.rk[i - 1 + kh], bl = S.rk[i - 1 + kl], bK = S.rk[i - 1 + K];
                                Lo = (Lo - (bh & 15u) * P.p_lo_top) * 5u + (bK & 15u);
                                RLo = (RLo - (b0 >> 4)) * INV5 + (bl >> 4) * P.p_lo_top;
                                if (kh) {
                                    Hi = (Hi - (b0 & 15u) * P.p_hi_top) * 5u + (bh & 15u);
                                    RHi = (RHi - (bl >> 4)) * INV5 + (bK >> 4) * P.p_hi_top;
                                }
                            }
                        }
                    }
                    __syncthreads();
                    // ---- phase 2a: prefix rightmost-argmin per block of Wp k-mers ---------------
                    for (uint32_t b = tid; b * Wp < nk; b += NT) {
                        uint32_t i = b * Wp;
                        const uint32_t e = (i + Wp < nk) ? i + Wp : nk;
                        uint64_t mv = S.v[i];
                        uint32_t mp = i;
                        S.ppos[i] = uint16_t(i);
                        for (++i; i < e; ++i) {
                            const uint64_t x = S.v[i];
                            if (x <= mv) { mv = x; mp = i; }
                            S.ppos[i] = uint16_t(mp);
                        }
                    }
                    __syncthreads();
                    // ---- phase 2b: suffix argmin, combine with next block's prefix -> window argmin
                    for (uint32_t b = tid; b * Wp < Tw; b += NT) {
                        const uint32_t lo = b * Wp;
                        const uint32_t e = (lo + Wp < nk) ? lo + Wp : nk;
                        uint64_t sv = 0;
                        uint32_t sp = 0;
                        for (uint32_t i = e; i-- > lo;) {
                            const uint64_t x = S.v[i];
                            if (i == e - 1 || x < sv) { sv = x; sp = i; }   // strict '<': keeps the rightmost
                            if (i < Tw) {
                                uint32_t rr = sp;
                                const uint32_t q = i + Wp - 1;
                                if (q >= e) {
                                    const uint32_t pp = S.ppos[q];
                                    if (S.v[pp] <= sv) rr = pp;
                                }
                                S.rw[i] = uint16_t(rr);
                            }
                        }
                    }
                    __syncthreads();
                    // ---- phase 2c: emission = tracked position changed; detect tie hazards -------
                    for (uint32_t w0 = 0; w0 < Tw; w0 += NT) {
                        const uint32_t w = w0 + tid;
                        bool em = false;
                        uint32_t rr = 0;
                        if (w < Tw) {
                            rr = S.rw[w];
                            if (w > 0) {
                                const uint32_t pr = S.rw[w - 1];
                                em = (rr != pr);
                                if (em && pr >= w && S.v[rr] == S.v[pr]) s_susp = 1;
                            } else if (!have_prev) {
                                em = true;
                            } else {
                                const long long lp = prev_pos - (long long)g0;
                                if (lp >= 0 && prev_val == S.v[rr]) { if (uint32_t(lp) != rr) s_susp = 1; }
                                else em = true;
                            }
                            if (w == Tw - 1) { s_carry_pos = (long long)g0 + rr; s_carry_val = S.v[rr]; }
                        }
                        const uint32_t bal = __ballot_sync(0xFFFFFFFFu, em);
                        if (bal) {
                            uint32_t basei = 0;
                            if (lane == 0) basei = atomicAdd(&s_nlist, (uint32_t)__popc(bal));
                            basei = __shfl_sync(0xFFFFFFFFu, basei, 0);
                            if (em) S.ppos[basei + __popc(bal & ((1u << lane) - 1u))] = uint16_t(rr);
                        }
                    }
                    __syncthreads();
                    if (s_susp) {   // exact sequential replay of the stateful rule (rare: ties at the window minimum)
                        if (tid == 0) {
                            uint32_t n = 0;
                            long long p = have_prev ? prev_pos - (long long)g0 : -1;
                            uint64_t pv = prev_val;
                            for (uint32_t w = 0; w < Tw; ++w) {
                                const uint32_t rr = S.rw[w];
                                if (p >= (long long)w && (w > 0 || have_prev) && pv == S.v[rr]) continue;
                                p = rr; pv = S.v[rr];
                                S.ppos[n++] = uint16_t(rr);
                            }
                            s_nlist = n;
                            s_carry_pos = (long long)g0 + p;
                            s_carry_val = pv;
                        }
                        __syncthreads();
                    }
                    have_prev = true;
                    prev_pos = s_carry_pos;
                    prev_val = s_carry_val;
                    // ---- phase 3: membership query + counting ------------------------------------
                    const uint32_t nl = s_nlist;
                    if (P.dbg_out && do_tot) {   // capture: one global atomic per tile
                        if (tid == 0) s_dbg_base = atomicAdd(P.dbg_n, (unsigned long long)nl);
                        __syncthreads();
                        for (uint32_t i = tid; i < nl; i += NT) {
                            const unsigned long long slot = s_dbg_base + i;
                            if (slot < P.dbg_cap) P.dbg_out[slot] = S.v[S.ppos[i]];
                        }
                    }
                    if (P.filter)
                    for (uint32_t i = tid; i < nl; i += NT) {
                        const uint64_t hv = S.v[S.ppos[i]];
                        const Mask<NW> m = bulk_contains<MODE>(P, hv, pol_filter);
                        if (small_bins) {
                            const uint32_t m0 = uint32_t(m.w[0]);
                            if (do_tot) {
#pragma unroll
                                for (int b = 0; b < 4; ++b) rc_cnt[b] += (m0 >> b) & 1u;
                            }
                            if (do_unq) {
                                const uint32_t mm = m0 & uint32_t(bm[0]);
#pragma unroll
                                for (int b = 0; b < 4; ++b) rc_unq[b] += (mm == (1u << b)) ? 1u : 0u;
                            }
                        } else {
                            int pc = 0;
                            int ub = -1;
#pragma unroll
                            for (int j = 0; j < NW; ++j) {
                                uint64_t x = m.w[j];
                                const uint64_t y = x & bm[j];
                                if (do_unq && y) { pc += __popcll(y); ub = j * 64 + (__ffsll((long long)y) - 1); }
                                if (do_tot)
                                    while (x) { const int b = __ffsll((long long)x) - 1; x &= x - 1; atomicAdd(&S.cnt[j * 64 + b], 1u); }
                            }
                            if (do_unq && pc == 1) atomicAdd(&S.uniq[S.b2c[ub]], 1u);
                        }
                    }
                    if (tid == 0 && do_tot) s_nhash += nl;
                    __syncthreads();
                }  // tiles
            }      // segments

            if (small_bins) {   // flush register counters
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    const uint32_t c = __reduce_add_sync(0xFFFFFFFFu, rc_cnt[b]);
                    const uint32_t u = __reduce_add_sync(0xFFFFFFFFu, rc_unq[b]);
                    if (lane == 0) { if (c) atomicAdd(&S.cnt[b], c); if (u) atomicAdd(&S.uniq[S.b2c[b]], u); }
                    rc_cnt[b] = 0; rc_unq[b] = 0;
                }
            }
            __syncthreads();
            if (do_tot) {   // best bin per category: greatest total, lowest bin on ties (read_entry.hpp:102-114)
                for (uint32_t b = tid; b < P.bins; b += NT)
                    atomicMax(&S.best[S.b2c[b]], ((unsigned long long)S.cnt[b] << 8) | (255ull - b));
                __syncthreads();
                if (!P.single_pass)
                    for (uint32_t c = tid; c < P.ncat; c += NT) {
                        const uint32_t bb = 255u - uint32_t(S.best[c] & 255ull);
                        if (bb < P.bins) atomicOr(&S.bestmask[bb >> 6], 1ull << (bb & 63u));
                    }
            }
        }  // passes
        __syncthreads();
        if (tid == 0) P.num_hashes[r] = s_nhash;
        for (uint32_t c = tid; c < P.ncat; c += NT) {
            P.counts[size_t(r) * P.ncat + c] = uint32_t(S.best[c] >> 8);
            P.unique_counts[size_t(r) * P.ncat + c] = S.uniq[c];
        }
    }
}

}  // namespace cgk
