// tests/select_support_dump.cpp — writes the two select supports host/index_writer.hpp appends to an `.idx` file for a given bit
// vector (stdin: <bits> then ceil(bits/64) words in hex), to the file named in argv[1].  tests/test_select_support.py
// re-derives select(i) from the written tables and compares it with a direct scan of the vector.
#include <cstdio>
#include <vector>
#include "../charon_b200/host/index_writer.hpp"
int main(int argc, char **argv) {
    unsigned long long bits = 0;
    if (argc < 2 || std::scanf("%llu", &bits) != 1) return 2;
    std::vector<uint64_t> w((bits + 63) / 64);
    for (auto &x : w) { unsigned long long v; if (std::scanf("%llx", &v) != 1) return 2; x = v; }
    charon_host::detail::Writer o(argv[1]);
    charon_host::detail::select_support(o, w, bits, 1);
    charon_host::detail::select_support(o, w, bits, 0);
    return 0;
}
