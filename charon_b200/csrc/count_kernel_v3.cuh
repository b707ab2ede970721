// charon_b200/csrc/count_kernel_v3.cuh — warp-per-read variant of the register-resident kernel (count_kernel_v2.cuh).
//
// Same arithmetic, same exactness argument, but the unit of work is a WARP: each warp owns one read at a time and walks
// it in tiles of 31*23 = 713 windows, so there is no CTA-wide barrier anywhere on the hot path (only __syncwarp).  Warps
// drift apart in phase, which lets one warp's load / gather latency hide behind the others' arithmetic, and the next
// tile's bases are prefetched with cp.async into a per-warp staging buffer (tagged by address) while the current tile is processed.
// Reads longer than P.long_read_min are left to the CTA-per-read kernel (count_kernel_v2), launched afterwards on the
// same stream, so that one very long read cannot serialise on a single warp.
#pragma once
#include "count_kernel_v2.cuh"

namespace cgk {

#ifndef CG_V3_WARPS
#define CG_V3_WARPS 4
#endif
#ifndef CG_V3_DQ
#define CG_V3_DQ 3      // gather groups (of 32 minimisers, 3 row loads each) kept in flight across a tile, in registers
                        // (re-measured on the final instruction mix: 2 -> 4.536 ms, 3 -> 4.520, 4 -> 4.697 bacterial)
#endif
#ifndef CG_V3_RC_LEFT
#define CG_V3_RC_LEFT 0     // reverse strand rolled right to left (2 IMAD per limb) instead of left to right (exact division by
#endif                      // 5: 3 per limb).  Fewer instructions, but all 23 values stay live across two passes and ptxas spills
#ifndef CG_V3_RC_LEFT_AG    // more.  Measured A/B (variants/, two runs each): register-carrier kernel (L2-resident filter) 4.524 ->
#define CG_V3_RC_LEFT_AG 1  // 4.56 ms = off; cp.async-carrier kernel (filter beyond L2) 6.55 -> 6.48 ms = on.
#endif
#ifndef CG_V3_DQA
#define CG_V3_DQA 3     // the same, when they fly as cp.async into shared memory (template parameter AG)
                        // (2 -> 6.536 ms, 3 -> 6.502, 4 -> 6.694 human-scale)
#endif
#ifndef CG_V3_MINB
#define CG_V3_MINB 3
#endif
#ifndef CG_V3_WARPS_AG
#define CG_V3_WARPS_AG 12   // the cp.async-carrier kernel (filter beyond L2) runs best as ONE 12-warp CTA per SM (measured: 4x3 6.25 ms,
#endif                      // 12x1 5.95 ms human-scale; the register-carrier kernel the other way round: 4.70 vs 4.81 ms bacterial)
#ifndef CG_V3_MINB_AG
#define CG_V3_MINB_AG 1
#endif
// warps per CTA (independent of each other) and resident CTAs per SM, per gather carrier.  With more than 4 bins every warp
// carries 4 KB of counters: 10 warps are what fits next to the landing slots of the gathers.
template <bool AG, bool SMALL = true> struct V3Shape {
    static constexpr int WARPS = AG ? (SMALL ? CG_V3_WARPS_AG : (CG_V3_WARPS_AG < 10 ? CG_V3_WARPS_AG : 10)) : CG_V3_WARPS;
    static constexpr int MINB = AG ? CG_V3_MINB_AG : CG_V3_MINB;
};
#ifndef CG_V3_WLIST_CAP
#define CG_V3_WLIST_CAP 256
#endif
#ifndef CG_V3_AMB_BYTES
#define CG_V3_AMB_BYTES 1536                       // per warp: 48 parked 256-bit masks (192 of 64 bits)
#endif
constexpr int WLIST_CAP = CG_V3_WLIST_CAP;         // emission ring of a warp (power of two; a tile emits ~60 entries, up to two tiles' worth wait in it)

template <int KK, int WW, bool SMALL, bool AG = false> struct V3Cfg {
    static constexpr int NK = 32 * WW;            // k-mers per warp tile
    static constexpr int TW = 31 * WW;            // windows per warp tile (incl. the context window)
    static constexpr int NB = NK + KK - 1;        // bases per warp tile
    static constexpr int NVEC = (NB + 15 + 15) / 16;      // 16-byte vectors that can cover a tile at any alignment
    static constexpr int NVL = (NVEC + 31) / 32;          // vectors per lane
    static constexpr int NCNT = SMALL ? 8 : 256;
    static constexpr int CODES = ((32 + 16 * NVL * 32 + 64) + 15) / 16 * 16;
    static constexpr int GQ = AG ? CG_V3_DQA * 3 * 32 * 16 : 0;   // landing slots of the asynchronous gathers: 16 bytes per (group, hash function, lane)
    static constexpr int RAW = NVL * 32 * 16;     // cp.async staging of the next tile's bases (ASCII, or 32 rank nibbles per vector)
    static constexpr int AMB = SMALL ? 0 : CG_V3_AMB_BYTES;   // one-pass counting: parked masks (CountParams::amb_mode)
    static_assert((63 + NB + 31) / 32 <= 32, "packed transports: 32 bases per lane must cover a tile");
    // a lane that overruns its PRIV_CAP slots (such a tile is redone exactly) spills into its right neighbour's slots; lane 31
    // spills into PRIV_GUARD, never into the list, whose head holds minimisers still waiting for a full gather group
    static constexpr int PRIV_GUARD = WW + 1 - PRIV_CAP > 0 ? WW + 1 - PRIV_CAP : 0;
    static constexpr size_t WARP_BYTES = size_t(NK) * 8 + (size_t(32) * PRIV_CAP + PRIV_GUARD) * 8 + size_t(WLIST_CAP) * 8 + size_t(NCNT) * 8 + 32 +
                                         size_t(NCNT) * 4 * 2 + CODES + RAW + GQ + AMB;
    static constexpr size_t SMEM = 512 + V3Shape<AG, SMALL>::WARPS * WARP_BYTES;
    static_assert(SMEM <= 227 * 1024, "a CTA of the warp-per-read kernel must fit the 227 KB of shared memory of an SM");
    static_assert(((size_t(32) * PRIV_CAP + PRIV_GUARD) * 8) % 16 == 0 && CODES % 16 == 0 && WARP_BYTES % 16 == 0,
                  "the staging vectors behind the slots need 16-byte alignment");
};

// FMT: transport of the bases the kernel decodes (enum cg_bases_format: 0 ASCII, 2 two bits per base) — a template parameter:
// with the format a run-time value the three-way address arithmetic costs the hot loop registers (+5 % kernel time measured).
template <int MODE, int KK, int WW, bool SMALL, bool AG = false, int FMT = 0>
#ifdef CG_V3_MAXNREG
__global__ void __maxnreg__(CG_V3_MAXNREG)
#else
__global__ void __launch_bounds__(V3Shape<AG, SMALL>::WARPS * 32, V3Shape<AG, SMALL>::MINB)
#endif
count_kernel_v3(const __grid_constant__ CountParams P) {
    using C2 = V2Cfg<KK, WW>;
    using C = V3Cfg<KK, WW, SMALL, AG && MODE == 0>;
    constexpr int NW = ModeTraits<MODE>::NW;
    constexpr int KL = C2::KL, KH = C2::KH;
    constexpr int WARPS3 = V3Shape<AG, SMALL>::WARPS;
    static_assert(FMT == 0 || FMT == 2, "the kernels decode ASCII or two-bit bases; rank nibbles are unpacked by the library");
    static_assert(!AG || MODE == 0, "the cp.async carrier only exists for narrow rows");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    Smem2 S;
    S.lut = smem_raw;
    S.b2c = smem_raw + 256;
    {
        unsigned char *w = smem_raw + 512 + size_t(warp) * C::WARP_BYTES;
        S.arr = reinterpret_cast<uint64_t *>(w);
        S.priv = S.arr + C::NK;
        S.list = S.priv + 32 * PRIV_CAP + C::PRIV_GUARD;
        S.best = reinterpret_cast<unsigned long long *>(S.list + WLIST_CAP);
        S.bestmask = S.best + C::NCNT;
        S.cnt = reinterpret_cast<uint32_t *>(S.bestmask + 4);
        S.uniq = S.cnt + C::NCNT;
        S.codes = reinterpret_cast<uint8_t *>(S.uniq + C::NCNT);
        S.amb = SMALL ? nullptr : reinterpret_cast<unsigned long long *>(S.codes + C::CODES + C::RAW + C::GQ);
        S.amb_cap = uint32_t(C::AMB / (8 * NW));
    }
    __shared__ uint32_t s_amb_n_w[WARPS3];
    S.amb_n = &s_amb_n_w[warp];
    const bool amb_mode = !SMALL && P.amb_mode != 0;
    unsigned long long mbw[NW];
#pragma unroll
    for (int j = 0; j < NW; ++j) mbw[j] = P.multibin[j];
    uint4 *const raw = reinterpret_cast<uint4 *>(S.codes + C::CODES);   // 16-byte aligned: every term of WARP_BYTES is a multiple of 16
    __shared__ uint32_t s_nhash_w[WARPS3], s_carry_valid_w[WARPS3];
    __shared__ long long s_carry_pos_w[WARPS3];
    uint32_t *const s_nhash = &s_nhash_w[warp];
    uint32_t *const s_carry_valid = &s_carry_valid_w[warp];
    long long *const s_carry_pos = &s_carry_pos_w[warp];

    const bool small_bins = SMALL;
    const uint64_t seed = P.seed;
    // filter rows: sticky in L2 while the filter can live there; a filter beyond L2 (AG) gains nothing from stickiness
#ifndef CG_V3_AG_POLICY
#define CG_V3_AG_POLICY 1   // measured on the 858 MB filter: evict_last 5.93 ms, evict_first 5.83, evict_normal / unchanged 5.93-5.94
#endif
    const uint64_t pol_filter = !AG || CG_V3_AG_POLICY == 0 ? l2_policy_evict_last()
                                : (CG_V3_AG_POLICY == 1 ? l2_policy_evict_first() : (CG_V3_AG_POLICY == 2 ? l2_policy_evict_normal() : l2_policy_evict_unchanged()));
    const uint64_t pol_stream = l2_policy_evict_first();
    for (uint32_t i = tid; i < 256; i += WARPS3 * 32) { S.lut[i] = base_code(i); S.b2c[i] = P.bin_to_cat[i]; }
    __syncthreads();   // the only CTA-wide barrier

    auto ticket = [&]() -> uint32_t {
        uint32_t t = 0;
        if (lane == 0) t = atomicAdd(P.work_counter, 1u);
        return __shfl_sync(0xFFFFFFFFu, t, 0);
    };
    // prefetched vectors of the next tile, tagged by the batch index of the tile's first base (+1; 0 = nothing staged).
    // ASCII transport: 16 bases per vector, aligned in global memory; packed transport: 32 rank nibbles per vector.
    // (P is also passed by reference to out-of-line helpers, so it has a local-memory copy: the lambdas below must capture plain
    // values, not P, or every use turns into a local load inside the hot loop)
    constexpr uint32_t fmt = FMT;    // 0 ASCII (16 bases per staged vector), 2 two bits per base (64 per vector)
    const uint8_t *const gbases = P.bases;
    uint64_t pf_tag = 0;
    const uint32_t gq_s = uint32_t(__cvta_generic_to_shared(S.codes + C::CODES + C::RAW)) + lane * 16u;
    const uint32_t raw_s = uint32_t(__cvta_generic_to_shared(raw));
    auto tile_head = [&](uint64_t G) -> uint32_t {
        return fmt == 2 ? uint32_t(G & 63u) : uint32_t(reinterpret_cast<uintptr_t>(gbases + G) & 15u);
    };
    auto tile_nvec = [&](uint32_t head, uint32_t nb) -> uint32_t {
        return fmt == 2 ? (head + nb + 63u) >> 6 : (head + nb + 15u) >> 4;
    };
    auto prefetch = [&](uint64_t G, uint32_t nb) {
        const uint32_t head = tile_head(G);
        const uint4 *vsrc = fmt == 2 ? reinterpret_cast<const uint4 *>(gbases) + (G >> 6) : reinterpret_cast<const uint4 *>(gbases + G - head);
        const uint32_t nvec = tile_nvec(head, nb);
#pragma unroll
        for (int q = 0; q < C::NVL; ++q) {
            const uint32_t i = lane + q * 32;
            if (i < nvec)
                asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;" ::"r"(raw_s + 16u * i), "l"(vsrc + i), "l"(pol_stream) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        pf_tag = G + 1;
    };

    Tally<MODE> T;
    uint32_t cur = ticket();
    uint32_t nxt = ticket();
    while (cur < P.n_reads) {
        const uint32_t r = cur;
        const uint32_t seg0 = P.read_seg ? P.read_seg[r] : r;
        const uint32_t seg1 = P.read_seg ? P.read_seg[r + 1] : r + 1;
        // address of the first tile of the next read (prefetch target while this read's last tile is processed)
        uint64_t next_read_G = 0;
        uint32_t next_read_nb = 0;   // 0: no prediction
        if (nxt < P.n_reads) {
            const uint32_t ns = P.read_seg ? P.read_seg[nxt] : nxt;
            const uint64_t nbase = P.seg_off[ns];
            const uint64_t nlen = P.seg_len ? P.seg_len[ns] : P.seg_off[ns + 1] - nbase;
            if (nlen >= uint64_t(KK + WW - 1)) {
                const uint64_t nwin = nlen - KK + 1 - WW + 1;
                const uint32_t tw = uint32_t(nwin < uint64_t(C::TW) ? nwin : uint64_t(C::TW));
                next_read_G = nbase;
                next_read_nb = tw + WW - 1 + KK - 1;
            }
        }
        bool skip = false;
        if (P.long_read_min) {   // long reads are left to the CTA-per-read kernel
            const uint64_t total = (P.seg_len ? P.seg_off[seg1 - 1] + P.seg_len[seg1 - 1] : P.seg_off[seg1]) - P.seg_off[seg0];
            skip = seg1 > seg0 && total >= P.long_read_min;
        }
        if (!skip) {
        for (uint32_t i = lane; i < uint32_t(C::NCNT); i += 32) { S.cnt[i] = 0; S.uniq[i] = 0; S.best[i] = 0; }
        if (lane < 4) S.bestmask[lane] = P.single_pass ? ~0ull : (amb_mode ? P.static_best[lane] : 0ull);
        if (lane == 0) { *s_nhash = 0; *S.amb_n = 0; }
        T.zero();
        __syncwarp();
        const int npass = P.single_pass ? 1 : 2;
        for (int pass = 0; pass < npass; ++pass) {
            const bool do_tot = (pass == 0);
            // one-pass counting: pass 0 also decides uniqueness wherever no multi-bin category is involved and parks the rest
            const bool amb_now = amb_mode && pass == 0;
            const bool do_unq = P.single_pass || pass == 1 || amb_now;
            const unsigned long long *const mb = amb_now ? mbw : nullptr;
            __syncwarp();
            unsigned long long bm[NW];
#pragma unroll
            for (int j = 0; j < NW; ++j) bm[j] = S.bestmask[j];
            const bool last_pass = (pass == npass - 1) || amb_now;   // amb_now: a second pass only after an overflow

            // S.list is a ring: `pend` emitted values starting at `qh` wait to be queried (warp-uniform); only full groups of 32 are
            // queried per tile, the rest waits for the next tile's emissions.
            uint32_t pend = 0, qh = 0;
            uint64_t *const ring = S.list;
            auto LI = [&](uint32_t x) -> uint64_t & { return ring[(qh + x) & uint32_t(WLIST_CAP - 1)]; };
            // Deferred gathers (narrow rows, 3 hash functions, local filter): the row fetches of the first full groups are issued
            // at the end of a tile and consumed after the NEXT tile's hashing, so the HBM / L2 latency hides behind arithmetic.
            // Two carriers = two instantiations (template parameter AG; the host launches AG when the filter does not fit in L2 —
            // one kernel with a run-time switch between them was 20-50 % slower on both):
            //  * registers (plain loads, up to DQ groups): cheapest in instructions, best while the rows hit L2; but ptxas
            //    recycles the loads' scoreboard at the top of the next tile, so a DRAM-latency load is waited for there;
            //  * shared memory (16-byte cp.async.cg of the aligned chunk holding the row word, up to DQA groups): completion is
            //    tracked per commit group, nothing forces an early wait; +4 % on an L2-resident filter, -4.5 % on an 0.86 GB one.
            // With AG every tile commits exactly two cp.async groups, in this order: the next tile's bases (phase 0), the gathers
            // (end of phase 3); either may be empty.  Without AG only the bases are committed — the extra commits / wait_group 1
            // alone made the register carrier 20-50 % slower, so the two structures are kept apart at compile time.  (Measured and dropped: 4-byte cp.async.ca, prefetch.global.L2.)
            constexpr int DQ = (MODE == 0) ? CG_V3_DQ : 0;
            constexpr int DQA = (MODE == 0) ? CG_V3_DQA : 0;
            constexpr bool ag = AG && (MODE == 0);
            uint32_t dq_w[(DQ && !ag) ? DQ : 1][3], dq_sh[(DQ && !ag) ? DQ : 1];
            uint32_t ndq = 0;
            const bool can_defer = DQ && P.hash_funs == 3 && (ag || P.n_parts == 0);   // the cp.async carrier also reaches rows in peer partitions
            auto dq_finish = [&]() {   // register carrier: consume the rows fetched at the end of the previous tile
                if (DQ && !ag) {
                    const uint32_t rb = 1u << P.row_shift;
                    const uint32_t row_msk = rb >= 32 ? 0xFFFFFFFFu : ((1u << rb) - 1u);
#pragma unroll
                    for (int g = 0; g < DQ; ++g)
                        if (g < int(ndq)) {
                            Mask<NW> m;
#pragma unroll
                            for (int j = 1; j < NW; ++j) m.w[j] = 0;
                            m.w[0] = ((dq_w[g][0] >> (dq_sh[g] & 31u)) & (dq_w[g][1] >> ((dq_sh[g] >> 5) & 31u)) & (dq_w[g][2] >> ((dq_sh[g] >> 10) & 31u))) & row_msk;
                            T.add(m, bm, small_bins, do_tot, do_unq, S, mb);
                        }
                }
                ndq = 0;
            };
            // cp.async carrier (AG): a TWO-STAGE pipeline with early exit.  bulk_contains is the AND of the three hash rows
            // (include/index.hpp:110-112 via seqan3), so a minimiser whose first row holds no bin at all cannot be in any bin:
            // rows 1 and 2 are only fetched for the lanes whose row 0 came back non-zero.  On a filter beyond L2 every row
            // is a DRAM line, and with k-mers broken by read errors about half of all minimisers stop after one line.
            //   end of tile t   : cohort A(t) = the first DQA full groups: row 0 -> cp.async into slot (g, 0)           [stage 0]
            //   end of tile t+1 : A(t): m0 = row-0 bits; lanes with m0 != 0 fetch rows 1, 2 into slots (g, 1), (g, 2)  [stage 1]
            //   end of tile t+2 : A(t): mask = m0 & row 1 & row 2 -> tallies
            // The values of a stage-0 cohort stay in the ring (behind its head) until its rows 1, 2 have been derived.
            uint32_t s0_n = 0, s1_n = 0, s0_base = 0, s0_span = 0;         // warp-uniform: groups per stage; ring index of the stage-0 cohort's first value and how far the head has moved past it
            uint32_t s0_sh[DQA ? DQA : 1], s1_sh[DQA ? DQA : 1], s1_m0[DQA ? DQA : 1];
            // 16-byte chunk that holds a row's word, and the row's word / bit position inside it.  Row-range partitioned filter
            // (SURVEY 8e): the chunk lives in the partition of the GPU that owns the row — this GPU's own memory or a peer's,
            // mapped over NVLink; the asynchronous copy hides the round trip behind the next tile's hashing either way.
            auto row_chunk = [&](uint32_t row, uint32_t &word, uint32_t &bitpos) -> const uint32_t * {
                const uint32_t *base = reinterpret_cast<const uint32_t *>(P.filter);
                if (P.n_parts) {
                    const uint32_t owner = row / P.rows_per_part;
                    row -= owner * P.rows_per_part;
                    if (owner != P.my_part) base = reinterpret_cast<const uint32_t *>(P.part_base[owner]);
                }
                word = row >> (5u - P.row_shift);
                bitpos = (row << P.row_shift) & 31u;
                return base + (word & ~3u);
            };
            auto ag_step = [&](bool bases_group_behind, uint32_t new_groups) {
                if (ag) {
                    const uint32_t rb = 1u << P.row_shift;
                    const uint32_t row_msk = rb >= 32 ? 0xFFFFFFFFu : ((1u << rb) - 1u);
                    if (s0_n | s1_n) {
                        if (bases_group_behind) asm volatile("cp.async.wait_group 1;" ::: "memory");
                        else asm volatile("cp.async.wait_group 0;" ::: "memory");
                    }
                    // stage 1 -> done
#pragma unroll
                    for (int g = 0; g < DQA; ++g)
                        if (g < int(s1_n) && s1_m0[g]) {
                            const uint32_t sh = s1_sh[g];
                            uint32_t w1, w2;
                            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w1) : "r"(gq_s + uint32_t(g * 3 + 1) * 512u + (((sh >> 10) & 3u) << 2)) : "memory");
                            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w2) : "r"(gq_s + uint32_t(g * 3 + 2) * 512u + (((sh >> 12) & 3u) << 2)) : "memory");
                            Mask<NW> m;
#pragma unroll
                            for (int j = 1; j < NW; ++j) m.w[j] = 0;
                            m.w[0] = s1_m0[g] & (w1 >> (sh & 31u)) & (w2 >> ((sh >> 5) & 31u));
                            if (m.w[0]) T.add(m, bm, small_bins, do_tot, do_unq, S, mb);
                        }
                    // stage 0 -> stage 1
#pragma unroll
                    for (int g = 0; g < DQA; ++g)
                        if (g < int(s0_n)) {
                            uint32_t w0;
                            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w0) : "r"(gq_s + uint32_t(g * 3) * 512u + (((s0_sh[g] >> 5) & 3u) << 2)) : "memory");
                            const uint32_t m0 = (w0 >> (s0_sh[g] & 31u)) & row_msk;
                            uint32_t sh = 0;
                            if (m0) {
                                const uint64_t value = ring[(s0_base + uint32_t(g) * 32u + lane) & uint32_t(WLIST_CAP - 1)];
#pragma unroll
                                for (int i = 1; i < 3; ++i) {
                                    uint32_t word, bitpos;
                                    const uint32_t *src = row_chunk(ibf_row32(value, c_ibf_seeds[i], P.hash_shift, uint32_t(P.bin_size)), word, bitpos);
                                    asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;" ::"r"(gq_s + uint32_t(g * 3 + i) * 512u),
                                                 "l"(src), "l"(pol_filter)
                                                 : "memory");
                                    sh |= (bitpos << (5 * (i - 1))) | ((word & 3u) << (10 + 2 * (i - 1)));
                                }
                            }
                            s1_sh[g] = sh; s1_m0[g] = m0;
                        }
                    s1_n = s0_n;
                    // new cohort -> stage 0
#pragma unroll
                    for (int g = 0; g < DQA; ++g)
                        if (uint32_t(g) < new_groups) {
                            const uint64_t value = LI(g * 32 + lane);
                            uint32_t word, bitpos;
                            const uint32_t *src = row_chunk(ibf_row32(value, c_ibf_seeds[0], P.hash_shift, uint32_t(P.bin_size)), word, bitpos);
                            asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;" ::"r"(gq_s + uint32_t(g * 3) * 512u),
                                         "l"(src), "l"(pol_filter)
                                         : "memory");
                            s0_sh[g] = bitpos | ((word & 3u) << 5);
                        }
                    s0_n = new_groups;
                    s0_base = qh;
                }
            };
            (void)s0_span;
            for (uint32_t seg = seg0; seg < seg1; ++seg) {
                const uint64_t sbase = P.seg_off[seg];
                const uint64_t slen = P.seg_len ? P.seg_len[seg] : P.seg_off[seg + 1] - sbase;
                if (slen < KK) continue;
                const uint64_t nkm = slen - KK + 1;
                if (nkm < WW) {   // a single window: its minimum is the only emission
                    uint64_t best = VMAX;
                    for (uint32_t i = lane; i < nkm; i += 32) best = umin64(best, kmer_value_slow<FMT, KK>(P, S.lut, sbase + i));
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) best = umin64(best, __shfl_xor_sync(0xFFFFFFFFu, best, o));
                    if (lane == 0) { const Smem2 Sx = S; process_exact<MODE>(P, Sx, s_nhash, best, do_tot, do_unq); }   // a copy: S itself must never have its address taken
                    __syncwarp();
                    continue;
                }
                const uint64_t nwin = nkm - WW + 1;
                uint64_t g0 = 0;
                bool first = true;
                for (;;) {
                    const uint32_t Tw = uint32_t(nwin - g0 < uint64_t(C::TW) ? nwin - g0 : uint64_t(C::TW));
                    const uint32_t nk = Tw + WW - 1;
                    const uint32_t nb = nk + KK - 1;
                    const bool last_tile = (g0 + Tw >= nwin);
                    // ---- phase 0: (prefetched) 16-byte vectors -> code bytes in shared memory -----------------------
                    const uint64_t G = sbase + g0;
                    const uint32_t head = tile_head(G);
                    const uint32_t nvec = tile_nvec(head, nb);
                    if (AG && MODE == 0) {
                        if (pf_tag != G + 1) {                        // not predicted: load now, wait for everything
                            prefetch(G, nb);
                            asm volatile("cp.async.wait_group 0;" ::: "memory");
                        } else {
                            asm volatile("cp.async.wait_group 1;" ::: "memory");   // the bases; the previous tile's gathers (committed later) may still fly
                        }
                    } else {
                        if (pf_tag != G + 1) prefetch(G, nb);         // not predicted: load now
                        asm volatile("cp.async.wait_all;" ::: "memory");
                    }
                    __syncwarp();                                      // staging complete for every lane; previous tile's readers of codes are done
                    if (fmt == 2) {
                        if (lane < ((head + nb + 31u) >> 5)) {   // half a vector (32 bases) per lane
                            const uint2 x = reinterpret_cast<const uint2 *>(raw)[lane];
                            *reinterpret_cast<uint4 *>(S.codes + 32 + 32 * lane) = decode_2bit16(x.x);
                            *reinterpret_cast<uint4 *>(S.codes + 32 + 32 * lane + 16) = decode_2bit16(x.y);
                        }
                        if (P.n_exc) patch_tile_warp(P, seg, G, nb, S.codes + 32 + head, lane);   // the batch holds bases that are not A/C/G/T
                    } else {
#pragma unroll
                        for (int q = 0; q < C::NVL; ++q) {
                            const uint32_t i = lane + q * 32;
                            if (i < nvec) {
                                const uint4 x = raw[i];
                                uint4 y;
                                y.x = encode4(S.lut, x.x); y.y = encode4(S.lut, x.y);
                                y.z = encode4(S.lut, x.z); y.w = encode4(S.lut, x.w);
                                *reinterpret_cast<uint4 *>(S.codes + 32 + 16 * i) = y;
                            }
                        }
                    }
                    __syncwarp();   // every lane has consumed its staged vectors: the buffer may be refilled
                    // predict + prefetch the next tile: next tile of this segment, else the next read's first tile
                    if (!last_tile) {
                        const uint64_t g1 = g0 + Tw - 1;
                        const uint32_t tw1 = uint32_t(nwin - g1 < uint64_t(C::TW) ? nwin - g1 : uint64_t(C::TW));
                        prefetch(sbase + g1, tw1 + WW - 1 + KK - 1);
                    } else if (last_pass && seg + 1 == seg1 && next_read_nb) {
                        prefetch(next_read_G, next_read_nb);
                    } else {
                        pf_tag = 0;
                        if (AG && MODE == 0) asm volatile("cp.async.commit_group;" ::: "memory");   // empty: keeps the two-groups-per-tile order
                    }
                    __syncwarp();
                    // ---- phase 1: rolling canonical hash of the lane's block, prefix minima -> smem -------------------
                    const uint32_t k0 = lane * WW;
                    const bool active = k0 < nk;
                    const bool full = (k0 + WW <= nk);
                    uint64_t v[WW];
                    uint64_t vleft = VMAX, m_prevwin = VMAX;
                    bool tie = false;
                    if (active) {
                        const uint8_t *cp = S.codes + 32 + head + k0 - 1;
                        uint32_t Hi = 0, Lo = 0, RHi = 0, RLo = 0;
#pragma unroll
                        for (int j = 0; j < KH; ++j) Hi = Hi * 5u + (cp[j] & 15u);
#pragma unroll
                        for (int j = KH; j < KK; ++j) Lo = Lo * 5u + (cp[j] & 15u);
                        if (AG ? CG_V3_RC_LEFT_AG : CG_V3_RC_LEFT) {
                        // The reverse strand is rolled RIGHT TO LEFT (from the block's last k-mer down to k-mer k0-1): moving
                        // left, the complement digits enter at the bottom, so both strands roll with "drop the top digit,
                        // times 5, add" (2 IMAD per limb) and the exact division by 5 (3 per limb) is gone.  v[] holds the
                        // reverse-strand values until the forward pass folds the canonical minimum into them.
#pragma unroll
                        for (int j = KL; j-- > 0;) RLo = RLo * 5u + (cp[WW + j] >> 4);
#pragma unroll
                        for (int j = KK; j-- > KL;) RHi = RHi * 5u + (cp[WW + j] >> 4);
#pragma unroll
                        for (int m = WW; m >= 1; --m) {   // k-mer k0-1+m = bases cp[m .. m+KK-1]
                            v[m - 1] = (uint64_t(RHi) * C2::P_KL + RLo) ^ seed;
                            const uint32_t cn = cp[m - 1] >> 4, cl = cp[m - 1 + KL] >> 4;
                            RLo = (RLo - cl * C2::P_LO_TOP) * 5u + cn;
                            if (KH) RHi = (RHi - (cp[m - 1 + KK] >> 4) * C2::P_HI_TOP) * 5u + cl;
                        }
                        {
                            const uint64_t f = (uint64_t(Hi) * C2::P_KL + Lo) ^ seed, rc = (uint64_t(RHi) * C2::P_KL + RLo) ^ seed;
                            vleft = k0 > 0 ? umin64h(f, rc) : VMAX;
                        }
#pragma unroll
                        for (int j = 0; j < WW; ++j) {
                            const uint32_t b0 = cp[j], bh = cp[j + KH], bK = cp[j + KK];
                            Lo = (Lo - (bh & 15u) * C2::P_LO_TOP) * 5u + (bK & 15u);
                            if (KH) Hi = (Hi - (b0 & 15u) * C2::P_HI_TOP) * 5u + (bh & 15u);
                            const uint64_t f = (uint64_t(Hi) * C2::P_KL + Lo) ^ seed;
                            v[j] = umin64h(f, v[j]);
                        }
                        } else {
#pragma unroll
                        for (int j = KL; j-- > 0;) RLo = RLo * 5u + (cp[j] >> 4);
#pragma unroll
                        for (int j = KK; j-- > KL;) RHi = RHi * 5u + (cp[j] >> 4);
                        {
                            const uint64_t f = (uint64_t(Hi) * C2::P_KL + Lo) ^ seed, rc = (uint64_t(RHi) * C2::P_KL + RLo) ^ seed;
                            vleft = k0 > 0 ? umin64h(f, rc) : VMAX;
                        }
#pragma unroll
                        for (int j = 0; j < WW; ++j) {
                            const uint32_t b0 = cp[j], bh = cp[j + KH], bl = cp[j + KL], bK = cp[j + KK];
                            Lo = (Lo - (bh & 15u) * C2::P_LO_TOP) * 5u + (bK & 15u);
                            RLo = (RLo - (b0 >> 4)) * INV5 + (bl >> 4) * C2::P_LO_TOP;
                            if (KH) {
                                Hi = (Hi - (b0 & 15u) * C2::P_HI_TOP) * 5u + (bh & 15u);
                                RHi = (RHi - (bl >> 4)) * INV5 + (bK >> 4) * C2::P_HI_TOP;
                            }
                            const uint64_t f = (uint64_t(Hi) * C2::P_KL + Lo) ^ seed, rc = (uint64_t(RHi) * C2::P_KL + RLo) ^ seed;
                            v[j] = umin64h(f, rc);
                        }
                        }
                        // k-mers at or beyond nk are NOT masked: every valid window w < Tw = nk - WW + 1 only covers k-mers
                        // w .. w+WW-1 <= nk-1 (own suffix + the right neighbour's prefix up to that index), the tie flags below
                        // are guarded, and windows >= Tw are dropped by the EDGE variant of p2_block / by exact_tile's bound.
                        uint64_t pm = v[0];
                        S.arr[k0] = pm;
#pragma unroll
                        for (int j = 1; j < WW; ++j) {
                            if (j == WW - 1) {
#if !CG_TIE_LEAVING_ONLY
                                tie = tie | (lo32(vleft) == lo32(pm));
#endif
                                m_prevwin = umin64h(vleft, pm);
                            }
#if !CG_TIE_LEAVING_ONLY
                            tie = tie | (lo32(v[j]) == lo32(pm));   // k-mers past nk can only add a spurious exact-path tile
#endif
                            pm = umin64c(pm, v[j]);
                            S.arr[k0 + j] = pm;
                        }
                        if (WW == 1) m_prevwin = vleft;
                    } else {
#pragma unroll
                        for (int j = 0; j < WW; ++j) v[j] = VMAX;
                    }
                    __syncwarp();
                    // ---- phase 2: window minima, run starts -> private slots -> warp-compacted list -------------------
                    const uint64_t v0_raw = v[0];
                    uint64_t *const pslot0 = S.priv + lane * PRIV_CAP;
                    const uint32_t pa0 = uint32_t(__cvta_generic_to_shared(pslot0));
                    uint32_t pa = pa0;   // shared-space address of the next free slot
                    if (active && k0 < Tw) {
                        if (k0 + WW <= Tw) p2_block<WW, false>(v, vleft, S.arr + k0 + WW, k0, Tw, pa, tie);
                        else p2_block<WW, true>(v, vleft, S.arr + k0 + WW, k0, Tw, pa, tie);
                        if (k0 > 0 ? (CG_TIE_LEAVING_ONLY ? neq64h(v[0], m_prevwin) : (lo32(v[0]) != lo32(m_prevwin))) : first) { st_shared_u64(pa, v[0]); pa += 8u; }   // see p2_block
                    }
                    asm volatile("" ::: "memory");   // the slots were written by st_shared_u64 (no memory clobber of its own): order the reads below
                    const uint32_t cnt = (pa - pa0) >> 3;
                    tie = tie | (cnt > PRIV_CAP);
#ifdef CG_FORCE_EXACT
                    tie = true;
#endif
                    uint32_t inc = cnt;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, o); if (lane >= uint32_t(o)) inc += t; }
                    uint32_t nl = __shfl_sync(0xFFFFFFFFu, inc, 31);
                    const bool redo = __any_sync(0xFFFFFFFFu, tie) || s0_span + pend + nl > WLIST_CAP;   // the ring also holds the stage-0 cohort's values
                    if (!redo) {
                        uint32_t dst = pend + inc - cnt;
                        for (uint32_t q = 0; q < cnt; ++q, ++dst) LI(dst) = pslot0[q];
                        if (lane == 0) *s_carry_valid = 0;
                    } else {
                        // ---- exact path: raw values to smem, per-run chain walk, inline emission ------------------------
                        const long long carry = *s_carry_valid ? *s_carry_pos - (long long)g0 : -1;
                        __syncwarp();
                        if (active) {
#pragma unroll
                            for (int j = 0; j < WW; ++j) S.arr[k0 + j] = j == 0 ? v0_raw : v[j];
                        }
                        __syncwarp();
                        if (lane == 0) *s_carry_valid = 0;   // re-armed by the walker of the run alive at the last window
                        __syncwarp();
                        { const Smem2 Sx = S; exact_tile<MODE, WW>(P, Sx, s_nhash, s_carry_pos, s_carry_valid, Tw, first, carry, sbase, g0, do_tot, do_unq, lane, 32u); }
                        nl = 0;
                    }
                    __syncwarp();
                    // ---- phase 3: membership query + counting ----------------------------------------------------------
                    if (P.dbg_out && do_tot && nl) {
                        unsigned long long base = 0;
                        if (lane == 0) base = atomicAdd(P.dbg_n, (unsigned long long)nl);
                        base = __shfl_sync(0xFFFFFFFFu, base, 0);
                        for (uint32_t i = lane; i < nl; i += 32) if (base + i < P.dbg_cap) P.dbg_out[base + i] = LI(pend + i);
                    }
                    if (lane == 0 && do_tot && nl) atomicAdd(s_nhash, nl);
                    if (P.filter) {
                        // only full groups of 32 are queried now; the remainder waits for the next tile's emissions
                        const uint32_t fullg = (pend + nl) & ~31u;
                        uint32_t first_now = 0;   // groups below this index go through the deferred queue
                        if (can_defer) {
                            const uint32_t ngroups = fullg >> 5;
                            if (ag) {
                                ndq = ngroups < uint32_t(DQA) ? ngroups : uint32_t(DQA);
                                ag_step(true, ndq);   // this tile's bases group was committed after the gathers that are consumed now
                                s0_span = ndq ? fullg : 0u;
                            } else {
                                dq_finish();          // issued one tile ago
#pragma unroll
                                for (int g = 0; g < DQ; ++g)
                                    if (uint32_t(g) < ngroups) {
                                        const uint64_t value = LI(g * 32 + lane);
                                        uint32_t sh = 0;
#pragma unroll
                                        for (int i = 0; i < 3; ++i) {
                                            const uint32_t row = ibf_row32(value, c_ibf_seeds[i], P.hash_shift, uint32_t(P.bin_size));
                                            dq_w[g][i] = ldg_na_u32(reinterpret_cast<const uint32_t *>(P.filter) + (row >> (5u - P.row_shift)), pol_filter);
                                            sh |= ((row << P.row_shift) & 31u) << (5 * i);
                                        }
                                        dq_sh[g] = sh;
                                    }
                                ndq = ngroups < uint32_t(DQ) ? ngroups : uint32_t(DQ);
                            }
                            first_now = ndq * 32;
                        }
                        for (uint32_t i = first_now + lane; i < fullg; i += 32) {
                            const Mask<NW> m = bulk_contains<MODE>(P, LI(i), pol_filter);
                            T.add(m, bm, small_bins, do_tot, do_unq, S, mb);
                        }
                        qh = (qh + fullg) & uint32_t(WLIST_CAP - 1);
                        pend = pend + nl - fullg;
                    }
                    if (AG && MODE == 0) asm volatile("cp.async.commit_group;" ::: "memory");   // the gathers of this tile (possibly none)
                    if (last_tile) break;
                    g0 += Tw - 1;
                    first = false;
                }  // tiles
            }      // segments
            if (ag) {              // drain both stages (each step commits its own, possibly empty, group)
                if (can_defer) {
                    ag_step(false, 0);
                    asm volatile("cp.async.commit_group;" ::: "memory");
                    ag_step(false, 0);
                }
            } else dq_finish();
            if (pend) {   // flush what is still waiting at the end of this pass
                __syncwarp();
                for (uint32_t i = lane; i < pend; i += 32) {
                    const Mask<NW> m = bulk_contains<MODE>(P, LI(i), pol_filter);
                    T.add(m, bm, small_bins, do_tot, do_unq, S, mb);
                }
                pend = 0;
            }

            if (small_bins) {
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    const uint32_t c = __reduce_add_sync(0xFFFFFFFFu, T.rc_cnt[b]);
                    const uint32_t u = __reduce_add_sync(0xFFFFFFFFu, T.rc_unq[b]);
                    if (lane == 0) { if (c) atomicAdd(&S.cnt[b], c); if (u) atomicAdd(&S.uniq[S.b2c[b]], u); }
                    T.rc_cnt[b] = 0; T.rc_unq[b] = 0;
                }
            }
            __syncwarp();
            if (do_tot) {
                for (uint32_t b = lane; b < P.bins; b += 32)
                    atomicMax(&S.best[S.b2c[b]], ((unsigned long long)S.cnt[b] << 8) | (255ull - b));
                __syncwarp();
                if (!P.single_pass)
                    for (uint32_t c = lane; c < P.ncat; c += 32) {
                        const uint32_t bb = 255u - uint32_t(S.best[c] & 255ull);
                        if (bb < P.bins) atomicOr(&S.bestmask[bb >> 6], 1ull << (bb & 63u));
                    }
            }
            if (amb_now) {   // the best bins are known: decide the parked minimisers, or redo the read's uniqueness in a second pass
                __syncwarp();
                const uint32_t na = *S.amb_n;
                if (na <= S.amb_cap) {
                    for (uint32_t i = lane; i < na; i += 32) {
                        int pc = 0, ub = -1;
#pragma unroll
                        for (int j = 0; j < NW; ++j) {
                            const uint64_t y = S.amb[i * NW + j] & S.bestmask[j];
                            if (y) { pc += __popcll(y); ub = j * 64 + (__ffsll((long long)y) - 1); }
                        }
                        if (pc == 1) atomicAdd(&S.uniq[S.b2c[ub]], 1u);
                    }
                    break;
                }
                for (uint32_t i = lane; i < uint32_t(C::NCNT); i += 32) S.uniq[i] = 0;
            }
        }  // passes
        __syncwarp();
        if (lane == 0) P.num_hashes[r] = *s_nhash;
        for (uint32_t c = lane; c < P.ncat; c += 32) {
            P.counts[size_t(r) * P.ncat + c] = uint32_t(S.best[c] >> 8);
            P.unique_counts[size_t(r) * P.ncat + c] = S.uniq[c];
        }
        __syncwarp();
        }  // !skip
        cur = nxt;
        nxt = ticket();
    }
}

}  // namespace cgk
