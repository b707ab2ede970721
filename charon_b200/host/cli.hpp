// charon_b200/host/cli.hpp — command line of `charon dehost` / `charon classify`.
// Same option names, aliases, types and defaults as the reference's CLI11 set-up
// (src/dehost_main.cpp:208-312, src/classify_main.cpp:25-119, src/main.cpp:40-62); `--gpus` and
// `--batch_reads` are additions of this implementation.
#pragma once
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "stats_model.hpp"

#ifndef CHARON_B200_VERSION
#define CHARON_B200_VERSION "charon-b200 0.1 (CLI of rmcolq/charon)"
#endif

namespace charon_host {

struct CliError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

inline std::string make_absolute(const std::string &p) {   // src/utils.cpp:7
    if (!p.empty() && p[0] == '/') return p;
    char buf[4096];
    if (!getcwd(buf, sizeof buf)) return p;
    return std::string(buf) + "/" + p;
}
inline bool path_exists(const std::string &p) { return access(p.c_str(), F_OK) == 0; }
inline bool ends_with(const std::string &s, const std::string &suf) { return s.size() >= suf.size() && s.compare(s.size() - suf.size(), suf.size(), suf) == 0; }

inline std::string usage(bool dehost) {
    std::string s;
    s += dehost ? "Dehost read file into host and other using index.\n" : "Classify read file using index.\n";
    s += std::string("Usage: charon ") + (dehost ? "dehost" : "classify") + " [OPTIONS] <fastaq> [<fastaq>]\n\n";
    s += "Positionals:\n  <fastaq> FILE REQUIRED        Fasta/q file\n  <fastaq> FILE                 Paired Fasta/q file\n\nOptions:\n";
    s += "  -h,--help                     Print this help message and exit\n";
    s += "  --db FILE REQUIRED            Prefix for the index.\n";
    s += "  -e,--extract STRING           Reads from this category in the index will be extracted to file.\n";
    s += "  -p,--prefix FILE              Prefix for the output files.\n";
    s += "  --chunk_size INT              Read file is read in chunks of this size, to be processed in parallel within a chunk.\n";
    s += "  --lo_hi_threshold FLOAT       Threshold used during model fitting stage to decide if read should be used to train lo or hi distribution.\n";
    s += "  --num_reads_to_fit INT        Number of reads to use to train each distribution in the model.\n";
    s += "  -d,--dist STRING              Probability distribution to use for modelling.\n";
    s += "  --min_length INT              Minimum read length to classify.\n";
    s += "  --min_quality INT             Minimum read quality to classify.\n";
    s += "  --min_compression FLOAT       Minimum read gzip compression ratio to classify (a measure of how much information is in the read.\n";
    s += "  --confidence INT              Minimum difference between the top 2 unique hit counts.\n";
    if (dehost) s += "  --host_unique_prop_lo_threshold INT\n                                Require non-host reads to have unique host proportion below this threshold for classification.\n";
    s += "  --min_proportion_diff FLOAT   Minimum difference between the proportion of (non-unique) kmers found in each category.\n";
    if (dehost) s += "  --min_probability_diff FLOAT  Minimum difference between the probability found in each category.\n";
    s += "  --log FILE                    File for log\n";
    s += "  -t,--threads INT              Maximum number of threads to use.\n";
    s += "  -v                            Verbosity of logging. Repeat for increased verbosity\n";
    s += "  --gpus INT                    [charon-b200] number of B200 GPUs to shard read batches over (default 1).\n";
    s += "  --gpu_compression             [charon-b200] compute the gzip compression ratio of every read on the GPU (bit-exact zlib level-6 size) instead of with zlib on the -t threads.\n";
    s += "  --partition_filter            [charon-b200] split the filter by row range over the GPUs (NVLink peer reads) instead of replicating it.\n";
    s += "  --batch_reads INT             [charon-b200] reads per GPU batch (default 16384).\n";
    return s;
}

namespace detail {
inline long to_int(const std::string &name, const std::string &v, long lo, long hi) {
    char *end = nullptr;
    const long x = std::strtol(v.c_str(), &end, 10);
    if (v.empty() || *end || x < lo || x > hi) throw CliError(name + ": Value " + v + " not in range [" + std::to_string(lo) + " - " + std::to_string(hi) + "]");
    return x;
}
inline float to_float(const std::string &name, const std::string &v) {
    char *end = nullptr;
    const float x = std::strtof(v.c_str(), &end);
    if (v.empty() || *end) throw CliError(name + ": Could not convert: " + v + " = " + v);
    return x;
}
}  // namespace detail

// Parses argv[first..] of one subcommand.  Returns false when help was printed.
inline bool parse_subcommand(bool dehost, int argc, char **argv, int first, Options &o) {
    o = dehost ? Options::dehost_defaults() : Options::classify_defaults();
    std::vector<std::string> positional;
    bool have_db = false;
    for (int i = first; i < argc; ++i) {
        std::string a = argv[i], val;
        bool has_val = false;
        if (a.size() > 2 && a[0] == '-' && a[1] == '-') {
            const size_t eq = a.find('=');
            if (eq != std::string::npos) { val = a.substr(eq + 1); a = a.substr(0, eq); has_val = true; }
        } else if (a.size() > 2 && a[0] == '-' && a[1] != '-' && a[1] != 'v') {   // -t4
            val = a.substr(2); a = a.substr(0, 2); has_val = true;
        }
        auto need = [&]() -> std::string {
            if (has_val) return val;
            if (i + 1 >= argc) throw CliError(a + ": 1 required");
            return std::string(argv[++i]);
        };
        if (a == "-h" || a == "--help") { std::fputs(usage(dehost).c_str(), stdout); return false; }
        else if (a == "--db") { o.db = need(); have_db = true; }
        else if (a == "-e" || a == "--extract") o.category_to_extract = need();
        else if (a == "-p" || a == "--prefix") { o.prefix = need(); if (path_exists(o.prefix)) throw CliError("--prefix: Path already exists: " + o.prefix); }
        else if (a == "--chunk_size") o.chunk_size = uint8_t(detail::to_int(a, need(), 0, 255));
        else if (a == "--lo_hi_threshold") o.lo_hi_threshold = detail::to_float(a, need());
        else if (a == "--num_reads_to_fit") o.num_reads_to_fit = uint16_t(detail::to_int(a, need(), 0, 65535));
        else if (a == "-d" || a == "--dist") o.dist = need();
        else if (a == "--min_length") o.min_length = uint32_t(detail::to_int(a, need(), 0, 4294967295L));
        else if (a == "--min_quality") o.min_quality = detail::to_float(a, need());
        else if (a == "--min_compression") o.min_compression = detail::to_float(a, need());
        else if (a == "--confidence") o.confidence_threshold = uint8_t(detail::to_int(a, need(), 0, 255));
        else if (dehost && a == "--host_unique_prop_lo_threshold") o.host_unique_prop_lo_threshold = detail::to_float(a, need());
        else if (a == "--min_proportion_diff") o.min_proportion_difference = detail::to_float(a, need());
        else if (dehost && a == "--min_probability_diff") o.min_prob_difference = detail::to_float(a, need());
        else if (a == "--log") o.log_file = make_absolute(need());
        else if (a == "-t" || a == "--threads") o.threads = uint8_t(detail::to_int(a, need(), 0, 255));
        else if (a == "--gpus") o.gpus = int(detail::to_int(a, need(), 1, 64));
        else if (a == "--partition_filter") o.partition_filter = true;
        else if (a == "--gpu_compression") o.gpu_compression = true;
        else if (a == "--batch_reads") o.batch_reads = uint32_t(detail::to_int(a, need(), 1, 1 << 22));
        else if (a.size() >= 2 && a[0] == '-' && a.find_first_not_of('v', 1) == std::string::npos) o.verbosity = uint8_t(o.verbosity + a.size() - 1);
        else if (!a.empty() && a[0] == '-' && a.size() > 1) throw CliError("The following argument was not expected: " + a);
        else positional.push_back(a);
    }
    if (positional.empty()) throw CliError("<fastaq> is required");
    if (positional.size() > 2) throw CliError("The following argument was not expected: " + positional[2]);
    if (!have_db) throw CliError("--db is required");
    o.read_file = make_absolute(positional[0]);
    if (!path_exists(o.read_file)) throw CliError("<fastaq>: File does not exist: " + positional[0]);
    if (positional.size() == 2) {
        o.read_file2 = make_absolute(positional[1]);
        if (!path_exists(o.read_file2)) throw CliError("<fastaq>: File does not exist: " + positional[1]);
    }
    // the reference checks --db with CLI::ExistingPath BEFORE appending ".idx" (dehost_main.cpp:228-231,489-491)
    if (!path_exists(o.db)) throw CliError("--db: Path does not exist: " + o.db);
    return true;
}

}  // namespace charon_host
