// charon_b200/csrc/count_kernel_v2.cuh — register-resident fused hot-path kernel, compile-time (k, w).
//
// Same contract as count_kernel.cuh (the generic any-(k,w) kernel) but specialised so that the inner loops are
// fully unrolled and the k-mer values of a thread never leave its registers:
//   * thread b of a 128-thread CTA owns block b = WW consecutive k-mers of the tile (WW = w-k+1 = 23);
//   * phase 1: 41 base codes -> rolling canonical hash (two 32-bit limbs per strand, exact /5 by modular inverse)
//              -> v[0..22] in registers, running prefix minimum -> shared memory (only array the neighbours read);
//   * phase 2: descending suffix minimum from registers, combined with the right neighbour's prefix minimum
//              = minimum of every window (van Herk / Gil-Werman), VALUES only;
//   * emission rule (exact restatement of seqan3's stateful minimiser_view): with p(w) the tracked
//              position,  p(w) = p(w-1) if min(w) == min(w-1) and p(w-1) is still inside window w, else the
//              rightmost argmin of window w; a value is emitted whenever p changes.  When no window of the tile
//              attains its minimum twice, that is simply "min(w) != min(w-1)" — the fast path.  Possible double
//              minima are detected conservatively (equal low words in any of the three min steps); such tiles are
//              re-done exactly by walking, per run of equal window minima, the chain p -> rightmost-argmin(p+1).
//   * phase 3: emitted values -> 3 row gathers -> AND -> counting (identical to the generic kernel).
// Tiles overlap by one window: local window 0 of a non-first tile is context only (the previous tile's last window).
#pragma once
#include "count_kernel.cuh"

namespace cgk {

constexpr int NT2 = 128;
#ifndef CG_V2_MINB
#define CG_V2_MINB 4
#endif
#ifndef CG_V2_LIST_CAP
#define CG_V2_LIST_CAP 1024
#endif
constexpr int LIST_CAP = CG_V2_LIST_CAP;   // compacted emission list of a tile (expected ~265 entries)
#ifndef CG_PRIV_CAP
#define CG_PRIV_CAP 8
#endif
constexpr int PRIV_CAP = CG_PRIV_CAP;      // per-thread emission slots (expected ~2 per 23 windows); overrun => exact path
constexpr uint64_t VMAX = ~0ull;   // never a real value: k <= 26 keeps the top 3 bits of v equal to the seed's (100b)

constexpr uint32_t cpow5(int e) { uint32_t r = 1; for (int i = 0; i < e; ++i) r *= 5u; return r; }

template <int KK, int WW> struct V2Cfg {
    static constexpr int KL = KK < 13 ? KK : 13;
    static constexpr int KH = KK - KL;
    static constexpr uint32_t P_LO_TOP = cpow5(KL - 1);
    static constexpr uint32_t P_HI_TOP = KH ? cpow5(KH - 1) : 0u;
    static constexpr uint32_t P_KL = cpow5(KL);
    static constexpr int NK = NT2 * WW;           // k-mers per tile
    static constexpr int TW = (NT2 - 1) * WW;     // windows per tile (incl. the context window)
    static constexpr int NB = NK + KK - 1;        // bases per tile
    static constexpr int NCODE = WW + KK;         // codes per thread: one extra k-mer to the left (window k0-1)
    static constexpr size_t SMEM = size_t(NK) * 8 + size_t(LIST_CAP) * 8 + size_t(NT2) * PRIV_CAP * 8 + (NB + 128) + 256 + 256 * 4 * 2 + 256 * 8 + 256 + 4 * 8;
};

struct Smem2 {
    uint64_t *arr;      // NK: prefix minima (fast path) / raw values (exact path)
    uint64_t *priv;     // NT2 * PRIV_CAP per-thread emission slots (overrun spills into list: such tiles are redone)
    uint64_t *list;     // LIST_CAP emitted minimiser values
    uint32_t *cnt, *uniq;
    unsigned long long *best, *bestmask;
    uint8_t *codes;     // 32 bytes of left slack + NB + slack ; codes[i] for tile base i lives at codes[32 + i]
    uint8_t *lut, *b2c;
    // one-pass counting (CountParams::amb_mode, warp-per-read kernel only): parked masks of minimisers that hit a bin of a
    // multi-bin category, amb_cap of them; *amb_n counts every attempt, so amb_n > amb_cap = overflow.  nullptr elsewhere.
    unsigned long long *amb;
    uint32_t *amb_n, amb_cap;
};

template <int KK, int WW>
__device__ __forceinline__ Smem2 carve2(unsigned char *raw) {
    using C = V2Cfg<KK, WW>;
    Smem2 s;
    s.arr = reinterpret_cast<uint64_t *>(raw);
    s.priv = s.arr + C::NK;
    s.list = s.priv + NT2 * PRIV_CAP;
    s.best = reinterpret_cast<unsigned long long *>(s.list + LIST_CAP);
    s.bestmask = s.best + 256;
    s.cnt = reinterpret_cast<uint32_t *>(s.bestmask + 4);
    s.uniq = s.cnt + 256;
    s.lut = reinterpret_cast<uint8_t *>(s.uniq + 256);
    s.b2c = s.lut + 256;
    s.codes = s.b2c + 256;
    s.amb = nullptr; s.amb_n = nullptr; s.amb_cap = 0;
    return s;
}

__device__ __forceinline__ uint32_t lo32(uint64_t x) { return uint32_t(x); }
__device__ __forceinline__ uint32_t encode4(const uint8_t *lut, uint32_t w) {   // 4 ASCII bases -> 4 codes
    return uint32_t(lut[w & 0xFFu]) | (uint32_t(lut[(w >> 8) & 0xFFu]) << 8) | (uint32_t(lut[(w >> 16) & 0xFFu]) << 16) |
           (uint32_t(lut[w >> 24]) << 24);
}
__device__ __forceinline__ uint64_t umin64(uint64_t a, uint64_t b) { return a < b ? a : b; }
// The same minimum with ONE compare (DSETP + 2 SEL instead of ISETP + ISETP.EX + 2 SEL): every real value carries the seed's
// top bits (1 00.. for k <= 26), i.e. read as a double it is negative and finite (or a negative denormal), and for those
// the IEEE order is the reverse of the unsigned order of the bit patterns.  VMAX (all ones) reads as NaN: it may only be
// passed as `a` (an unordered compare is false and selects b).  Used on the hot min chains; CG_UMIN_INT=1 switches back.
#ifndef CG_UMIN_INT
#define CG_UMIN_INT 0
#endif
__device__ __forceinline__ uint64_t umin64h(uint64_t a, uint64_t b) {
#if CG_UMIN_INT
    return a < b ? a : b;
#else
    return __longlong_as_double((long long)a) > __longlong_as_double((long long)b) ? a : b;
#endif
}
// The two SERIAL chains (running prefix / suffix minimum) may use either flavour: CG_UMIN_CHAIN_INT=1 keeps the integer
// compare there (shorter dependent latency than a DSETP through the short scoreboard, one more instruction).
// Measured A/B: 4.52 vs 4.54 ms bacterial, 6.48 vs 6.43 ms human-scale: a wash, the one-compare flavour stays.
#ifndef CG_UMIN_CHAIN_INT
#define CG_UMIN_CHAIN_INT 0
#endif
__device__ __forceinline__ uint64_t umin64c(uint64_t a, uint64_t b) {
#if CG_UMIN_CHAIN_INT
    return a < b ? a : b;
#else
    return umin64h(a, b);
#endif
}
// a != b with one compare (DSETP.NEU): distinct real values are distinct doubles (all negative, never -0 vs +0).
__device__ __forceinline__ bool neq64h(uint64_t a, uint64_t b) {
#if CG_UMIN_INT
    return a != b;
#else
    return __longlong_as_double((long long)a) != __longlong_as_double((long long)b);
#endif
}
#ifndef CG_TIE_LEAVING_ONLY
#define CG_TIE_LEAVING_ONLY 1   // see p2_block.  On its own (round 1's last commit) this was NOT exact: a fast tile hands no tracked
#endif                          // position to the next one; exact_tile now reconstructs it (seam_tracked_position, tests/test_tile_seams.py).
// Store through a 32-bit shared-space address.  A moving generic pointer costs a 64-bit add next to the address the STS uses;
// deliberately no "memory" clobber (the window-minimum loads around it must stay free to move): the caller fences once.
__device__ __forceinline__ void st_shared_u64(uint32_t addr, uint64_t v) { asm volatile("st.shared.u64 [%0], %1;" ::"r"(addr), "l"(v)); }

// Two-bit transport (CG_BASES_2BIT): A0 C1 G2 T3, sixteen bases per 32-bit word, base i at bits 2i.  The even and the odd
// fields, masked apart, are nibble selectors into the four code bytes; two more PRMT per eight bases interleave them.
constexpr uint32_t PK2_T = 0x04122140u;   // code bytes (rank | complement rank << 4) of A, C, G, T
__device__ __forceinline__ uint4 decode_2bit16(uint32_t w) {   // 16 bases -> 16 code bytes
    const uint32_t e = w & 0x33333333u, o = (w >> 2) & 0x33333333u;
    const uint32_t el = __byte_perm(PK2_T, 0u, e), eh = __byte_perm(PK2_T, 0u, e >> 16);
    const uint32_t ol = __byte_perm(PK2_T, 0u, o), oh = __byte_perm(PK2_T, 0u, o >> 16);
    uint4 r;
    r.x = __byte_perm(el, ol, 0x5140u); r.y = __byte_perm(el, ol, 0x7362u);
    r.z = __byte_perm(eh, oh, 0x5140u); r.w = __byte_perm(eh, oh, 0x7362u);
    return r;
}
// first index in [lo, hi) whose exception position is >= pos
__device__ __forceinline__ uint32_t exc_lower_bound(const unsigned long long *exc, uint32_t lo, uint32_t hi, unsigned long long pos) {
    while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (exc[mid] < pos) lo = mid + 1; else hi = mid; }
    return lo;
}
// The non-ACGT bases of batch positions [G, G + nb) of a two-bit tile become N in the code bytes (codes[0] = base G).
// [lo, hi) = the segment's exception entries; threads t0, t0 + stride, ... share the work.  Cold: reads rarely hold an N.
__device__ __noinline__ void patch_exceptions(const unsigned long long *exc, uint32_t lo, uint32_t hi, unsigned long long G, uint32_t nb,
                                              uint8_t *codes, uint32_t t0, uint32_t stride) {
    for (uint32_t e = exc_lower_bound(exc, lo, hi, G) + t0; e < hi; e += stride) {
        const unsigned long long p = exc[e];
        if (p >= G + nb) break;
        codes[p - G] = 0x33;
    }
}
// warp flavour: looks the segment's range up itself, so that the hot loop only spends one test on P.n_exc
__device__ __noinline__ void patch_tile_warp(const CountParams &P, uint32_t seg, unsigned long long G, uint32_t nb, uint8_t *codes, uint32_t lane) {
    const uint32_t lo = P.exc_seg_off[seg], hi = P.exc_seg_off[seg + 1];
    if (hi > lo) {
        __syncwarp();   // every lane's decoded vectors are in place before single bytes of them are overwritten
        patch_exceptions(P.exc_pos, lo, hi, G, nb, codes, lane, 32u);
    }
}
// code byte of base i of the batch (cold paths only: segments shorter than one window)
__device__ __noinline__ uint32_t code_at_2bit(const CountParams &P, uint64_t i) {
    if (P.n_exc) {
        const uint32_t e = exc_lower_bound(P.exc_pos, 0u, uint32_t(P.n_exc), i);
        if (e < P.n_exc && P.exc_pos[e] == i) return 0x33u;
    }
    const uint32_t r = (uint32_t(P.bases[i >> 2]) >> ((uint32_t(i) & 3u) * 2u)) & 3u;
    return (PK2_T >> (8u * r)) & 0xFFu;
}
template <int FMT>
__device__ __forceinline__ uint32_t code_at(const CountParams &P, const uint8_t *lut, uint64_t i) {
    return FMT == 2 ? code_at_2bit(P, i) : uint32_t(lut[P.bases[i]]);
}
// seeded canonical value of the k-mer that starts at base i (cold paths only)
template <int FMT, int KK>
__device__ __noinline__ uint64_t kmer_value_slow(const CountParams &P, const uint8_t *lut, uint64_t i) {
    uint64_t f = 0, rc = 0;
#pragma unroll 1
    for (int j = 0; j < KK; ++j) f = f * 5u + (code_at<FMT>(P, lut, i + j) & 15u);
#pragma unroll 1
    for (int j = KK; j-- > 0;) rc = rc * 5u + (code_at<FMT>(P, lut, i + j) >> 4);
    f ^= P.seed; rc ^= P.seed;
    return f < rc ? f : rc;
}

// per-read counters shared by the list phase and the inline (exact path) emission
// one-pass counting: park the mask if it touches a bin of a multi-bin category (true = parked, or lost to overflow)
template <int NW>
__device__ __forceinline__ bool amb_park(const Mask<NW> &m, const unsigned long long *mb, const Smem2 &S) {
    bool touch = false;
#pragma unroll
    for (int j = 0; j < NW; ++j) touch = touch || (m.w[j] & mb[j]) != 0;
    if (touch) {
        const uint32_t slot = atomicAdd(S.amb_n, 1u);
        if (slot < S.amb_cap) {
#pragma unroll
            for (int j = 0; j < NW; ++j) S.amb[slot * NW + j] = m.w[j];
        }
    }
    return touch;
}

template <int MODE>
struct Tally {
    static constexpr int NW = ModeTraits<MODE>::NW;
    uint32_t rc_cnt[4], rc_unq[4];
    __device__ __forceinline__ void zero() {
#pragma unroll
        for (int b = 0; b < 4; ++b) { rc_cnt[b] = 0; rc_unq[b] = 0; }
    }
    __device__ __forceinline__ void add(const Mask<NW> &m, const unsigned long long *bm, bool small_bins, bool do_tot, bool do_unq,
                                        const Smem2 &S, const unsigned long long *mb = nullptr) {
        if (!small_bins && mb && amb_park<NW>(m, mb, S)) do_unq = false;   // its uniqueness is decided at the end of the read
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
            int pc = 0, ub = -1;
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
};

// rightmost argmin of arr[w .. w+WW-1] (exact path)
template <int WW>
__device__ __forceinline__ void scan_window(const uint64_t *arr, uint32_t w, uint64_t &m, uint32_t &r) {
    m = arr[w]; r = w;
#pragma unroll
    for (int j = 1; j < WW; ++j) {
        const uint64_t x = arr[w + j];
        if (x <= m) { m = x; r = w + j; }
    }
}

// Phase 2 for one full block.  v[] = the block's 23 values; on return v[0] holds the minimum of window k0.
// next = prefix minima of the right neighbour block (shared memory).  EDGE: some windows of the block lie beyond Tw.
// Emission: window w+1 starts a run iff min(w+1) != min(w).  That rule equals seqan3's "tracked position changed" unless
// the k-mer that has just left window w carries the value of min(w): if the tracked element is still inside, a newcomer
// takes over iff it is strictly smaller (min changes) and nothing happens otherwise (min unchanged); if the tracked
// element left, seqan3 always emits the new rightmost minimum, and the value rule misses that exactly when
// leaving == min(w).  So `tie` only needs that one test (on low words: a superset).  CG_TIE_LEAVING_ONLY=0 restores the
// earlier scheme (run starts compared on low words, every min step flagged on equal low words: 3 more ISETP per window).
template <int WW, bool EDGE>
__device__ __forceinline__ void p2_block(uint64_t (&v)[WW], uint64_t vleft, const uint64_t *next, uint32_t k0, uint32_t Tw, uint32_t &pa,
                                         bool &tie) {
    uint64_t sv = v[WW - 1], mw_next = 0;
#pragma unroll
    for (int j = WW - 1; j >= 0; --j) {
        const bool wv = !EDGE || (k0 + j < Tw);
        if (j < WW - 1) {
#if !CG_TIE_LEAVING_ONLY
            tie = tie | (lo32(v[j]) == lo32(sv));
#endif
            sv = umin64c(sv, v[j]);
        }
        uint64_t mw = sv;
        if (j >= 1) {
            const uint64_t npm = (EDGE && !wv) ? VMAX : next[j - 1];
#if !CG_TIE_LEAVING_ONLY
            tie = tie | ((lo32(npm) == lo32(sv)) & wv);
#endif
            mw = umin64h(npm, sv);   // npm may be VMAX: first operand
        }
        const uint64_t leaving = j >= 1 ? v[j >= 1 ? j - 1 : 0] : vleft;
        tie = tie | ((lo32(leaving) == lo32(mw)) & wv);
        if (j < WW - 1) {
#if CG_TIE_LEAVING_ONLY
            const bool start = neq64h(mw_next, mw) & (!EDGE || (k0 + j + 1 < Tw));
#else
            // low words suffice: two DIFFERENT adjacent window minima with equal low words always trip one of the tie flags
            // (old minimum left: `leaving` check; newcomer took over: it met the old minimum in a prefix / suffix / combine step)
            const bool start = (lo32(mw_next) != lo32(mw)) & (!EDGE || (k0 + j + 1 < Tw));
#endif
            if (start) { st_shared_u64(pa, mw_next); pa += 8u; }   // one predicated STS + one 32-bit add per window
        }
        mw_next = mw;
    }
    v[0] = mw_next;
}

// One emitted minimiser handled on the spot (exact path / short segments): capture, query, count with shared-memory
// atomics (the register tallies of the list phase are flushed into the same counters at the end of the read).
template <int MODE>
__device__ __noinline__ void process_exact(const CountParams &P, const Smem2 &S, uint32_t *s_nhash, uint64_t hv, bool do_tot, bool do_unq) {
    constexpr int NW = ModeTraits<MODE>::NW;
    if (P.dbg_out && do_tot) {
        const unsigned long long slot = atomicAdd(P.dbg_n, 1ull);
        if (slot < P.dbg_cap) P.dbg_out[slot] = hv;
    }
    if (do_tot) atomicAdd(s_nhash, 1u);
    if (!P.filter) return;
    const Mask<NW> m = bulk_contains<MODE>(P, hv, l2_policy_evict_last());
    if (P.amb_mode && S.amb && do_tot && do_unq && amb_park<NW>(m, P.multibin, S)) do_unq = false;   // one-pass counting, pass 0
    int pc = 0, ub = -1;
#pragma unroll
    for (int j = 0; j < NW; ++j) {
        uint64_t x = m.w[j];
        const uint64_t y = x & S.bestmask[j];
        if (do_unq && y) { pc += __popcll(y); ub = j * 64 + (__ffsll((long long)y) - 1); }
        if (do_tot)
            while (x) { const int b = __ffsll((long long)x) - 1; x &= x - 1; atomicAdd(&S.cnt[j * 64 + b], 1u); }
    }
    if (do_unq && pc == 1) atomicAdd(&S.uniq[S.b2c[ub]], 1u);
}

// Tracked position (tile-local k-mer index) in the context window of a tile that has to be redone exactly although the tile
// before it went down the fast path and so handed nothing over (its last window = this tile's window 0, which holds its
// minimum m more than once).  seqan3's tracked element is the rightmost minimum of the window in which the current run of
// equal window minima STARTED (either the previous one fell out and the window was rescanned, or a strictly smaller newcomer
// took over — then it is that window's only minimum), and it is still inside this window, otherwise the fast tile would have
// been flagged when it left (leaving value == window minimum, p2_block).  Hence the run started at most WW - 1 windows ago:
// the values of the WW - 1 k-mers in front of the tile are recomputed from the bases, the run start ws is the first window,
// going backwards, whose minimum differs from m, and the answer is the rightmost m at or left of k-mer ws + WW - 1.
template <int WW>
__device__ __noinline__ uint32_t seam_tracked_position(const CountParams &P, const Smem2 &S, uint64_t m, uint64_t sbase, uint64_t g0) {
    uint64_t prev[WW > 1 ? WW - 1 : 1];   // prev[i] = value of tile-local k-mer i - (WW - 1)
    for (int i = 0; i < WW - 1; ++i) {
        const uint64_t at = sbase + g0 - uint64_t(WW - 1) + uint64_t(i);
        uint64_t f = 0, rc = 0;
        for (uint32_t j = 0; j < P.K; ++j) f = f * 5u + ((P.packed == 2 ? code_at<2>(P, S.lut, at + j) : code_at<0>(P, S.lut, at + j)) & 15u);
        for (uint32_t j = P.K; j-- > 0;) rc = rc * 5u + ((P.packed == 2 ? code_at<2>(P, S.lut, at + j) : code_at<0>(P, S.lut, at + j)) >> 4);
        f ^= P.seed; rc ^= P.seed;
        prev[i] = f < rc ? f : rc;
    }
    int ws = -(WW - 1);                   // run start, relative to this tile's window 0
    for (int d = 1; d <= WW - 1; ++d) {   // window -d = k-mers -d .. WW - 1 - d
        uint64_t mw = ~0ull;
        for (int x = -d; x <= WW - 1 - d; ++x) { const uint64_t v = x < 0 ? prev[x + WW - 1] : S.arr[x]; mw = v < mw ? v : mw; }
        if (mw != m) { ws = 1 - d; break; }
    }
    for (int x = ws + WW - 1; x >= 0; --x)
        if (x < WW && S.arr[x] == m) return uint32_t(x);
    return 0xFFFFFFFFu;                   // not reached for a window that holds m
}

// Exact minimiser emission for one tile whose raw values sit in S.arr: per run of equal window minima walk the chain
// p -> rightmost-argmin(window p+1) (see the file header).  `carry` = tracked position (tile-local) entering the tile.
template <int MODE, int WW>
__device__ __noinline__ void exact_tile(const CountParams &P, const Smem2 &S, uint32_t *s_nhash, long long *s_carry_pos, uint32_t *s_carry_valid,
                                        uint32_t Tw, bool first, long long carry, uint64_t sbase, uint64_t g0, bool do_tot, bool do_unq,
                                        uint32_t t0 = threadIdx.x, uint32_t stride = NT2) {
    for (uint32_t w = t0; w < Tw; w += stride) {
        uint64_t m; uint32_t p;
        scan_window<WW>(S.arr, w, m, p);
        bool walk = false, emit = false;
        if (w == 0) {
            walk = true;
            if (first) emit = true;
            else if (carry >= 0 && carry < WW && S.arr[carry] == m) p = uint32_t(carry);   // context window: restore, no emission
            else if (carry < 0) {   // the tile before this one took the fast path: which of several equal minima was being tracked?
                uint32_t occ = 0;
#pragma unroll 1
                for (int j = 0; j < WW; ++j) occ += S.arr[j] == m ? 1u : 0u;
                if (occ > 1) { const uint32_t q = seam_tracked_position<WW>(P, S, m, sbase, g0); if (q < uint32_t(WW)) p = q; }
            }
        } else {
            uint64_t mp; uint32_t rp;
            scan_window<WW>(S.arr, w - 1, mp, rp);
            if (mp != m) { emit = true; walk = true; }
        }
        while (walk) {
            if (emit) process_exact<MODE>(P, S, s_nhash, m, do_tot, do_unq);
            const uint32_t nw = p + 1;
            if (nw >= Tw) {
                // The chain leaves the tile.  Only the run that is still alive at the tile's last window owns the carried
                // position (an earlier run whose tracked element merely sits near the tile end must not overwrite it).
                uint64_t ml; uint32_t rl;
                scan_window<WW>(S.arr, Tw - 1, ml, rl);
                if (ml == m) { *s_carry_pos = (long long)g0 + p; *s_carry_valid = 1; }
                break;
            }
            uint64_t m2; uint32_t r2;
            scan_window<WW>(S.arr, nw, m2, r2);
            if (m2 != m) break;
            p = r2;
            emit = true;
        }
    }
}

template <int MODE, int KK, int WW>
__global__ void __launch_bounds__(NT2, CG_V2_MINB) count_kernel_v2(const __grid_constant__ CountParams P) {
    using C = V2Cfg<KK, WW>;
    constexpr int NW = ModeTraits<MODE>::NW;
    constexpr int KL = C::KL, KH = C::KH;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const Smem2 S = carve2<KK, WW>(smem_raw);
    __shared__ uint32_t s_read, s_nlist, s_nhash, s_carry_valid;
    __shared__ long long s_carry_pos;
    __shared__ unsigned long long s_dbg_base;

    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    const bool small_bins = (NW == 1) && (P.bins <= 4);
    const uint64_t seed = P.seed;
    const uint64_t pol_filter = l2_policy_evict_last(), pol_stream = l2_policy_evict_first();
    for (uint32_t i = tid; i < 256; i += NT2) { S.lut[i] = base_code(i); S.b2c[i] = P.bin_to_cat[i]; }

    Tally<MODE> T;

    for (;;) {
        __syncthreads();
        if (tid == 0) s_read = atomicAdd(P.work_counter, 1u);
        __syncthreads();
        const uint32_t r = s_read;
        if (r >= P.n_reads) break;
        if (P.long_only) {   // companion launch of the warp-per-read kernel: only the reads it skipped
            const uint32_t a = P.read_seg ? P.read_seg[r] : r, b = P.read_seg ? P.read_seg[r + 1] : r + 1;
            const uint64_t total = b > a ? (P.seg_len ? P.seg_off[b - 1] + P.seg_len[b - 1] : P.seg_off[b]) - P.seg_off[a] : 0;
            if (total < P.long_read_min) continue;
        }
        for (uint32_t i = tid; i < 256; i += NT2) { S.cnt[i] = 0; S.uniq[i] = 0; S.best[i] = 0; }
        if (tid < 4) S.bestmask[tid] = P.single_pass ? ~0ull : 0ull;
        if (tid == 0) s_nhash = 0;
        T.zero();
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
                if (slen < KK) continue;
                const uint64_t nkm = slen - KK + 1;
                if (nkm < WW) {
                    // ---- fewer k-mers than one window: a single window, a single emission (its minimum) ----
                    if (tid < 32) {
                        uint64_t best = VMAX;
                        for (uint32_t i = lane; i < nkm; i += 32)
                            best = umin64(best, P.packed == 2 ? kmer_value_slow<2, KK>(P, S.lut, sbase + i) : kmer_value_slow<0, KK>(P, S.lut, sbase + i));
#pragma unroll
                        for (int o = 16; o > 0; o >>= 1) best = umin64(best, __shfl_xor_sync(0xFFFFFFFFu, best, o));
                        if (lane == 0) { const Smem2 Sx = S; process_exact<MODE>(P, Sx, &s_nhash, best, do_tot, do_unq); }
                    }
                    continue;
                }
                const uint64_t nwin = nkm - WW + 1;
                uint64_t g0 = 0;
                bool first = true;
                for (;;) {
                    const uint32_t Tw = uint32_t(nwin - g0 < uint64_t(C::TW) ? nwin - g0 : uint64_t(C::TW));
                    const uint32_t nk = Tw + WW - 1;
                    const uint32_t nb = nk + KK - 1;
                    // ---- phase 0: load + encode.  16-byte vectors aligned in GLOBAL memory, all of a thread's loads in
                    //      flight before the first use; codes[32 + head + i] = code of tile base i -----------------
                    const uint64_t G = sbase + g0;   // index of the tile's first base in the batch
                    const uint8_t *gsrc = P.bases + G;
                    const uint32_t head = P.packed == 2 ? uint32_t(G & 63u) : uint32_t(reinterpret_cast<uintptr_t>(gsrc) & 15u);
                    if (P.packed == 2) {   // 16-byte vectors of 64 two-bit bases: half a vector (32 bases) per thread
                        const uint2 *vsrc = reinterpret_cast<const uint2 *>(P.bases) + ((G >> 6) << 1);
                        const uint32_t nhalf = (head + nb + 31u) >> 5;
                        if (tid < nhalf) {
                            uint2 x;
                            asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.u32 {%0, %1}, [%2], %3;" : "=r"(x.x), "=r"(x.y) : "l"(vsrc + tid), "l"(pol_stream));
                            *reinterpret_cast<uint4 *>(S.codes + 32 + 32 * tid) = decode_2bit16(x.x);
                            *reinterpret_cast<uint4 *>(S.codes + 32 + 32 * tid + 16) = decode_2bit16(x.y);
                        }
                        if (P.n_exc) {
                            const uint32_t elo = P.exc_seg_off[seg], ehi = P.exc_seg_off[seg + 1];
                            if (ehi > elo) {
                                __syncthreads();
                                patch_exceptions(P.exc_pos, elo, ehi, G, nb, S.codes + 32 + head, tid, NT2);
                            }
                        }
                    } else {
                        const uint4 *vsrc = reinterpret_cast<const uint4 *>(gsrc - head);
                        const uint32_t nvec = (head + nb + 15u) >> 4;
                        constexpr int NV = (C::NB + 15 + 15) / 16 / NT2 + 1;   // vectors per thread (2 for k=19, w=41)
                        uint4 x[NV];
#pragma unroll
                        for (int q = 0; q < NV; ++q) {
                            const uint32_t i = tid + q * NT2;
                            if (i < nvec) x[q] = ldg_stream_v4(vsrc + i, pol_stream);
                        }
#pragma unroll
                        for (int q = 0; q < NV; ++q) {
                            const uint32_t i = tid + q * NT2;
                            if (i < nvec) {
                                uint4 y;
                                y.x = encode4(S.lut, x[q].x); y.y = encode4(S.lut, x[q].y);
                                y.z = encode4(S.lut, x[q].z); y.w = encode4(S.lut, x[q].w);
                                *reinterpret_cast<uint4 *>(S.codes + 32 + 16 * i) = y;
                            }
                        }
                    }
                    __syncthreads();
                    if (tid == 0) s_nlist = 0;
                    // ---- phase 1: hash own block (+ the k-mer left of it), prefix minima -> smem ---------------
                    const uint32_t k0 = tid * WW;
                    const bool active = k0 < nk;
                    const bool full = (k0 + WW <= nk);
                    uint64_t v[WW];
                    uint64_t vleft = VMAX;          // value of k-mer k0-1 (exists when k0 > 0)
                    uint64_t m_prevwin = VMAX;      // minimum of window k0-1
                    bool tie = false;
                    if (active) {
                        const uint8_t *cp = S.codes + 32 + head + k0 - 1;   // cp[t] = code of base k0-1+t, t = 0..NCODE-1
                        uint32_t Hi = 0, Lo = 0, RHi = 0, RLo = 0;
                        // k-mer k0-1 (garbage for k0 == 0: reads the zeroed slack, never used)
#pragma unroll
                        for (int j = 0; j < KH; ++j) Hi = Hi * 5u + (cp[j] & 15u);
#pragma unroll
                        for (int j = KH; j < KK; ++j) Lo = Lo * 5u + (cp[j] & 15u);
#pragma unroll
                        for (int j = KL; j-- > 0;) RLo = RLo * 5u + (cp[j] >> 4);
#pragma unroll
                        for (int j = KK; j-- > KL;) RHi = RHi * 5u + (cp[j] >> 4);
                        {
                            const uint64_t f = (uint64_t(Hi) * C::P_KL + Lo) ^ seed, rc = (uint64_t(RHi) * C::P_KL + RLo) ^ seed;
                            vleft = k0 > 0 ? umin64(f, rc) : VMAX;
                        }
#pragma unroll
                        for (int j = 0; j < WW; ++j) {
                            // roll from k-mer (k0-1+j) to (k0+j): outgoing base cp[j], incoming cp[j+KK]
                            const uint32_t b0 = cp[j], bh = cp[j + KH], bl = cp[j + KL], bK = cp[j + KK];
                            Lo = (Lo - (bh & 15u) * C::P_LO_TOP) * 5u + (bK & 15u);
                            RLo = (RLo - (b0 >> 4)) * INV5 + (bl >> 4) * C::P_LO_TOP;
                            if (KH) {
                                Hi = (Hi - (b0 & 15u) * C::P_HI_TOP) * 5u + (bh & 15u);
                                RHi = (RHi - (bl >> 4)) * INV5 + (bK >> 4) * C::P_HI_TOP;
                            }
                            const uint64_t f = (uint64_t(Hi) * C::P_KL + Lo) ^ seed, rc = (uint64_t(RHi) * C::P_KL + RLo) ^ seed;
                            v[j] = umin64(f, rc);
                        }
                        if (!full) {
#pragma unroll
                            for (int j = 0; j < WW; ++j) if (k0 + j >= nk) v[j] = VMAX;
                        }
                        uint64_t pm = v[0];
                        S.arr[k0] = pm;
#pragma unroll
                        for (int j = 1; j < WW; ++j) {
                            if (j == WW - 1) {   // window k0-1 = {k-mer k0-1} U prefix[0..WW-2]
                                tie = tie | (lo32(vleft) == lo32(pm));
                                m_prevwin = umin64(vleft, pm);
                            }
                            tie = tie | ((lo32(v[j]) == lo32(pm)) & (full || k0 + j < nk));
                            pm = umin64(pm, v[j]);
                            S.arr[k0 + j] = pm;
                        }
                        if (WW == 1) m_prevwin = vleft;
                    } else {
#pragma unroll
                        for (int j = 0; j < WW; ++j) v[j] = VMAX;
                    }
                    __syncthreads();
                    // ---- phase 2: suffix minima + neighbour prefix = window minima; a window whose minimum differs from
                    //      its predecessor's starts a run = one emission, appended to the thread's private slots ----------
                    const uint64_t v0_raw = v[0];
                    uint64_t *const pslot0 = S.priv + tid * PRIV_CAP;
                    const uint32_t pa0 = uint32_t(__cvta_generic_to_shared(pslot0));
                    uint32_t pa = pa0;   // shared-space address of the next free slot
                    if (active && k0 < Tw) {
                        const bool interior = (k0 + WW <= Tw);   // every window of the block is valid (=> next block's prefix exists)
                        if (interior) p2_block<WW, false>(v, vleft, S.arr + k0 + WW, k0, Tw, pa, tie);
                        else p2_block<WW, true>(v, vleft, S.arr + k0 + WW, k0, Tw, pa, tie);
                        // window k0 against window k0-1 (computed locally in phase 1, no neighbour needed); v[0] now holds min(k0)
                        if (k0 > 0 ? (v[0] != m_prevwin) : first) { st_shared_u64(pa, v[0]); pa += 8u; }
                    }
                    {   // compaction: warp scan of the per-thread counts, one shared atomic per warp
                        asm volatile("" ::: "memory");   // the slots were written by st_shared_u64 (no memory clobber of its own): order the reads below
                    const uint32_t cnt = (pa - pa0) >> 3;
                        tie |= (cnt > PRIV_CAP);            // private slots overrun: redo the tile exactly (results below are discarded)
                        uint32_t inc = cnt;
#pragma unroll
                        for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, o); if (lane >= uint32_t(o)) inc += t; }
                        uint32_t wbase = 0;
                        if (lane == 31) wbase = atomicAdd(&s_nlist, inc);
                        wbase = __shfl_sync(0xFFFFFFFFu, wbase, 31);
                        uint32_t dst = wbase + inc - cnt;
                        for (uint32_t q = 0; q < cnt && q < PRIV_CAP; ++q, ++dst) if (dst < LIST_CAP) S.list[dst] = pslot0[q];
                    }
                    const int redo = __syncthreads_or(int(tie));
                    uint32_t nl = s_nlist;
                    if (redo || nl > LIST_CAP) {
                        // ---- exact path: raw values to smem, per-run chain walk, inline emission -------------------
                        const long long carry = s_carry_valid ? s_carry_pos - (long long)g0 : -1;   // read before anyone rewrites it
                        if (active) {
#pragma unroll
                            for (int j = 0; j < WW; ++j) S.arr[k0 + j] = j == 0 ? v0_raw : v[j];
                        }
                        __syncthreads();
                        if (tid == 0) s_carry_valid = 0;     // re-armed by the walker of the run alive at the last window
                        __syncthreads();
                        { const Smem2 Sx = S; exact_tile<MODE, WW>(P, Sx, &s_nhash, &s_carry_pos, &s_carry_valid, Tw, first, carry, sbase, g0, do_tot, do_unq); }
                        nl = 0;
                        __syncthreads();
                    } else if (tid == 0) {
                        s_carry_valid = 0;
                    }
                    // ---- phase 3: membership query + counting over the emission list -------------------------
                    if (P.dbg_out && do_tot && nl) {
                        if (tid == 0) s_dbg_base = atomicAdd(P.dbg_n, (unsigned long long)nl);
                        __syncthreads();
                        for (uint32_t i = tid; i < nl; i += NT2) {
                            const unsigned long long slot = s_dbg_base + i;
                            if (slot < P.dbg_cap) P.dbg_out[slot] = S.list[i];
                        }
                    }
                    if (P.filter)
                        for (uint32_t i = tid; i < nl; i += NT2) {
                            const Mask<NW> m = bulk_contains<MODE>(P, S.list[i], pol_filter);
                            T.add(m, bm, small_bins, do_tot, do_unq, S);
                        }
                    if (tid == 0 && do_tot) atomicAdd(&s_nhash, nl);
                    if (g0 + Tw >= nwin) break;
                    g0 += Tw - 1;
                    first = false;   // no barrier needed: the next tile's writes to codes/arr/list are all behind its own barriers
                }  // tiles
            }      // segments

            if (small_bins) {
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    const uint32_t c = __reduce_add_sync(0xFFFFFFFFu, T.rc_cnt[b]);
                    const uint32_t u = __reduce_add_sync(0xFFFFFFFFu, T.rc_unq[b]);
                    if (lane == 0) { if (c) atomicAdd(&S.cnt[b], c); if (u) atomicAdd(&S.uniq[S.b2c[b]], u); }
                    T.rc_cnt[b] = 0; T.rc_unq[b] = 0;
                }
            }
            __syncthreads();
            if (do_tot) {
                for (uint32_t b = tid; b < P.bins; b += NT2)
                    atomicMax(&S.best[S.b2c[b]], ((unsigned long long)S.cnt[b] << 8) | (255ull - b));
                __syncthreads();
                if (!P.single_pass)
                    for (uint32_t c = tid; c < P.ncat; c += NT2) {
                        const uint32_t bb = 255u - uint32_t(S.best[c] & 255ull);
                        if (bb < P.bins) atomicOr(&S.bestmask[bb >> 6], 1ull << (bb & 63u));
                    }
            }
        }  // passes
        __syncthreads();
        if (tid == 0) P.num_hashes[r] = s_nhash;
        for (uint32_t c = tid; c < P.ncat; c += NT2) {
            P.counts[size_t(r) * P.ncat + c] = uint32_t(S.best[c] >> 8);
            P.unique_counts[size_t(r) * P.ncat + c] = S.uniq[c];
        }
    }
}

}  // namespace cgk
