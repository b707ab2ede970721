// tools/prep_microbench.cpp — what one host thread needs per read for the phred sum, the 2-bit packing and freeing the strings (no zlib):
// build from tools/: g++ -O2 -std=c++14 -fopenmp -o /tmp/prep_microbench prep_microbench.cpp -L../charon_b200 -lcharon_gpu -lz -Wl,-rpath,$PWD/../charon_b200
// (3.2 us per 8 kb read on the B200 box's host; tools/prep_probe.sh times the same inside the host program)
#include "../charon_b200/host/fastx.hpp"
#include "../include/charon_gpu.h"
#include <chrono>
#include <cstdio>
#include <vector>
using namespace charon_host;
#include <fstream>
#include <random>
int main() {
    { std::mt19937 rng(1); std::ofstream f("/tmp/t_prep.fq"); std::string s(8000, 'A'), q(8000, '5');
      for (int r = 0; r < 20000; ++r) { for (auto &c : s) c = "ACGT"[rng() & 3]; f << "@r" << r << "\n" << s << "\n+\n" << q << "\n"; } }
    FastxReader rd("/tmp/t_prep.fq");
    std::vector<FastxRecord> recs; FastxRecord r;
    while (rd.next(r)) recs.push_back(r);
    std::vector<uint8_t> packed(recs.size() * 2100);
    std::vector<std::vector<uint64_t>> read_exc(recs.size());
    auto t0 = std::chrono::steady_clock::now();
    double acc = 0;
    uint64_t tmp_exc[256];
    for (size_t i = 0; i < recs.size(); ++i) {
        FastxRecord &x = recs[i];
        int sum = 0; for (unsigned char q : x.qual) sum += int(q) - 33;
        acc += x.qual.size() ? float(sum) / float(x.qual.size()) : 0.f;
        uint64_t k = 0; const uint64_t at = i * 8000;
        cg_pack_bases_2bit(x.seq.data(), x.seq.size(), packed.data() + at / 4, at, tmp_exc, 256, &k);
        std::vector<uint64_t> &e = read_exc[i];
        if (k) e.insert(e.end(), tmp_exc, tmp_exc + k);
        FastxRecord().seq.swap(x.seq); FastxRecord().qual.swap(x.qual);
    }
    double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::printf("%zu reads: %.2f us per read (%f)\n", recs.size(), dt / recs.size() * 1e6, acc);
}
