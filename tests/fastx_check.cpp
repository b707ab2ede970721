// tests/fastx_check.cpp — TEST INFRASTRUCTURE: the FASTQ reader's SSE2 fast paths (plain A/C/G/T runs copied 16 letters at a time,
// quality range check) against a letter-by-letter normaliser on records that mix upper / lower case, IUPAC codes, U and N.
// usage: fastx_check <tmp.fq>
#include "../charon_b200/host/fastx.hpp"
#include <cstdio>
#include <fstream>
#include <random>
int main(int argc, char **argv) {
    const std::string path = argc > 1 ? argv[1] : "/tmp/fastx_check.fq";
    // write a FASTQ with assorted content and read it back; compare against a scalar normaliser
    std::mt19937 rng(1);
    const char *alpha = "ACGTacgtNnRYKMUu";
    std::vector<std::string> seqs;
    for (int r = 0; r < 3000; ++r) {
        int n = rng() % 300; std::string s;
        for (int i = 0; i < n; ++i) { int p = rng() % 100; s += p < 90 ? "ACGT"[rng() % 4] : alpha[rng() % 16]; }
        seqs.push_back(s);
    }
    { std::ofstream f(path); for (size_t i = 0; i < seqs.size(); ++i) { f << "@r" << i << "\n" << seqs[i] << "\n+\n" << std::string(seqs[i].size(), 'I') << "\n"; } }
    charon_host::FastxReader rd(path);
    charon_host::FastxRecord rec; size_t i = 0; int bad = 0;
    while (rd.next(rec)) {
        std::string want;
        for (char c : seqs[i]) { char u = c & ~0x20; want += (u=='A'||u=='C'||u=='G'||u=='T') ? u : (u=='U' ? 'T' : 'N'); }
        if (want != rec.seq) { ++bad; if (bad < 4) std::printf("DIFF %zu\n%s\n%s\n", i, want.c_str(), rec.seq.c_str()); }
        ++i;
    }
    std::printf("records %zu bad %d\n", i, bad);
    return bad != 0;
}
