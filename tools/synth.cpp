// tools/synth.cpp — seeded synthetic references and ONT-like reads.  Bench/test tooling shared by
// both arms (the GPU path and the CPU oracle see the same bytes); not part of the product library.
//   refs : iid uniform ACGT from splitmix64(seed)
//   reads: fixed length L; source drawn from the references with the given probabilities (rest: pure random),
//          random strand, per-base edits: substitution / insertion / deletion each with rate `err`.
#include <cstdint>
#include <cstring>
#include <vector>
static inline uint64_t splitmix64(uint64_t &s) {
    uint64_t z = (s += 0x9E3779B97F4A7C15ULL);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
static inline double u01(uint64_t &s) { return double(splitmix64(s) >> 11) * (1.0 / 9007199254740992.0); }
extern "C" {
void synth_reference(uint64_t seed, uint64_t n, char *out) {
    const uint64_t CH = 1 << 20;
#pragma omp parallel for schedule(static)
    for (int64_t c = 0; c < int64_t((n + CH - 1) / CH); ++c) {
        uint64_t s0 = seed ^ 0x5851F42D4C957F2DULL, s1 = uint64_t(c) + 0x14057B7EF767814FULL;
        uint64_t s = splitmix64(s0) ^ (splitmix64(s1) * 0xD1B54A32D192ED03ULL);   // decorrelate seeds and chunks
        uint64_t lo = uint64_t(c) * CH, hi = lo + CH < n ? lo + CH : n;
        for (uint64_t i = lo; i < hi;) {
            uint64_t r = splitmix64(s);
            for (int j = 0; j < 32 && i < hi; ++j, ++i) { out[i] = "ACGT"[r & 3]; r >>= 2; }
        }
    }
}
// refs: concatenated references, ref_off[nref+1].  frac[nref]: probability of drawing from each ref (rest: random).
// out: n_reads*L bases.  src_out[n_reads] (may be NULL): source ref index, or nref for random.
void synth_reads(uint64_t seed, const char *refs, const uint64_t *ref_off, uint32_t nref, const double *frac, uint32_t n_reads,
                 uint32_t L, double err, char *out, uint8_t *src_out, uint64_t first_read_index) {
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < int64_t(n_reads); ++r) {
        uint64_t s = seed ^ ((first_read_index + uint64_t(r) + 1) * 0xA0761D6478BD642FULL);
        splitmix64(s);
        char *dst = out + uint64_t(r) * L;
        double u = u01(s), acc = 0;
        uint32_t src = nref;
        for (uint32_t i = 0; i < nref; ++i) { acc += frac[i]; if (u < acc) { src = i; break; } }
        if (src_out) src_out[r] = uint8_t(src);
        if (src == nref) {
            for (uint32_t i = 0; i < L;) { uint64_t x = splitmix64(s); for (int j = 0; j < 32 && i < L; ++j, ++i) { dst[i] = "ACGT"[x & 3]; x >>= 2; } }
            continue;
        }
        uint64_t rl = ref_off[src + 1] - ref_off[src];
        uint64_t need = uint64_t(L) + L / 8 + 64;
        uint64_t start = rl > need ? uint64_t(u01(s) * double(rl - need)) : 0;
        const char *p = refs + ref_off[src] + start, *pend = refs + ref_off[src + 1];
        uint32_t i = 0;
        while (i < L) {
            double e = u01(s);
            if (p >= pend) { dst[i++] = "ACGT"[splitmix64(s) & 3]; continue; }
            if (e < err) { char c = *p++; char n; do { n = "ACGT"[splitmix64(s) & 3]; } while (n == c); dst[i++] = n; }
            else if (e < 2 * err) { dst[i++] = "ACGT"[splitmix64(s) & 3]; }
            else if (e < 3 * err) { ++p; }
            else dst[i++] = *p++;
        }
        if (splitmix64(s) & 1) {  // reverse strand
            for (uint32_t a = 0, b = L - 1; a < b; ++a, --b) { char t = dst[a]; dst[a] = dst[b]; dst[b] = t; }
            for (uint32_t a = 0; a < L; ++a) { char c = dst[a]; dst[a] = c == 'A' ? 'T' : c == 'C' ? 'G' : c == 'G' ? 'C' : 'A'; }
        }
    }
}
}
