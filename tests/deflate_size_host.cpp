// tests/deflate_size_host.cpp — TEST INFRASTRUCTURE: host build of charon_b200/csrc/deflate_size.cuh (the sequential core of the
// GPU gzip-size kernel) with a scalar match search over the same occ[] / idx[] arrays the kernel uses.  Reads one sequence per
// line from stdin, prints the gzip size per line.  tests/test_deflate_size.py compares the output with the real zlib of the
// box.  Never linked into the product.
#include <cstdio>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>
#include "../charon_b200/csrc/deflate_size.cuh"

struct HostMatcher {
    const uint8_t *text; uint32_t n;
    std::vector<uint32_t> occ, idx, start;
    HostMatcher(const uint8_t *t, uint32_t n_) : text(t), n(n_), occ(n_), idx(n_), start(dsz::N_LETTER_CODES + 1, 0) {
        if (n < 3) return;
        const uint32_t np = n - 2;     // positions with three letters
        std::vector<uint32_t> cnt(dsz::N_LETTER_CODES, 0);
        for (uint32_t p = 0; p < np; ++p) cnt[code(p)]++;
        for (int b = 0; b < dsz::N_LETTER_CODES; ++b) start[b + 1] = start[b] + cnt[b];
        std::fill(cnt.begin(), cnt.end(), 0);
        for (uint32_t p = 0; p < np; ++p) { const uint32_t b = code(p); idx[p] = start[b] + cnt[b]++; occ[idx[p]] = p; }
    }
    uint32_t code(uint32_t p) const { return dsz::letter_code(text[p]) * 25 + dsz::letter_code(text[p + 1]) * 5 + dsz::letter_code(text[p + 2]); }
    uint8_t at(uint32_t p) const { return text[p]; }
    bool head(uint32_t p, uint32_t, uint32_t &q) const {
        const uint32_t b = code(p);
        if (idx[p] == start[b]) return false;
        q = occ[idx[p] - 1];
        return true;
    }
    uint32_t longest(uint32_t p, uint32_t prev_length, uint32_t &match_start) const {
        const uint32_t lookahead = n - p;
        uint32_t chain = dsz::MAX_CHAIN;
        if (prev_length >= dsz::GOOD_LENGTH) chain >>= 2;
        const uint32_t nice = dsz::NICE_LENGTH < lookahead ? dsz::NICE_LENGTH : lookahead;
        const uint32_t cap = uint32_t(dsz::MAX_MATCH) < lookahead ? uint32_t(dsz::MAX_MATCH) : lookahead;
        const uint32_t limit = p > dsz::MAX_DIST ? p - dsz::MAX_DIST : 0;
        const uint32_t bstart = start[code(p)];
        uint32_t best = prev_length, k = idx[p] - 1;
        for (;;) {
            const uint32_t q = occ[k];
            uint32_t len = 0;
            while (len < cap && text[q + len] == text[p + len]) ++len;
            if (len > best) { match_start = q; best = len; if (len >= nice) break; }
            if (k == bstart) break;
            --k;
            if (occ[k] <= limit) break;
            if (--chain == 0) break;
        }
        return best <= lookahead ? best : lookahead;
    }
};

int main() {
    std::string line;
    static dsz::Trees T;
    while (std::getline(std::cin, line)) {
        HostMatcher m(reinterpret_cast<const uint8_t *>(line.data()), uint32_t(line.size()));
        dsz::init_block(T);
        dsz::TreeSink sink{T};
        std::printf("%u\n", dsz::gzip_size(m, sink, uint32_t(line.size())));
    }
    return 0;
}
