// tests/training_trace.cpp — drives ONE implementation of the --extract training state machine (the product's
// host/stats_model.hpp with -DIMPL_PRODUCT, the oracle's oracle.hpp with -DIMPL_ORACLE) over unique-proportion vectors read
// from stdin and prints its state after every read, then the fitted KDE datasets.  tests/test_training_twin.py compares both
// with the independent Python statement tests/pyref.py:TrainingTrace.
#include <cstdio>
#include <vector>
#ifdef IMPL_PRODUCT
#include "../charon_b200/host/stats_model.hpp"
using Model = charon_host::StatsModel;
using Opt = charon_host::Options;
#else
#include "../oracle/oracle.hpp"
using Model = orc::StatsModel;
using Opt = orc::Options;
#endif

int main() {
    unsigned n, ncat, cap, force;
    float thr;
    if (std::scanf("%u %u %u %f %u", &n, &ncat, &cap, &thr, &force) != 5) return 2;
    Opt o;
    o.dist = "kde"; o.lo_hi_threshold = thr; o.num_reads_to_fit = uint16_t(cap); o.run_extract = true;
    Model m(o, ncat);
    std::vector<float> p(ncat);
    for (unsigned r = 0; r < n; ++r) {
        for (unsigned c = 0; c < ncat; ++c) if (std::scanf("%f", &p[c]) != 1) return 2;
        const bool ready = m.add_read_to_training_data(p);
        std::printf("%d", ready ? 1 : 0);
#ifdef IMPL_PRODUCT
        for (unsigned c = 0; c < ncat; ++c) std::printf(" %zu %zu %d", m.training()[c].pos.size(), m.training()[c].neg.size(), m.models()[c].ready ? 1 : 0);
#else
        for (unsigned c = 0; c < ncat; ++c) std::printf(" %zu %zu %d", m.training[c].pos.size(), m.training[c].neg.size(), m.models[c].ready ? 1 : 0);
#endif
        std::printf("\n");
    }
    if (force) m.force_ready();
    for (unsigned c = 0; c < ncat; ++c) {
#ifdef IMPL_PRODUCT
        const auto &kp = m.models()[c].k_pos; const auto &kn = m.models()[c].k_neg;
#else
        const auto &kp = m.models[c].k_pos.dataset; const auto &kn = m.models[c].k_neg.dataset;
#endif
        std::printf("fit %u pos %zu", c, kp.size()); for (float v : kp) std::printf(" %.9g", v);
        std::printf("\nfit %u neg %zu", c, kn.size()); for (float v : kn) std::printf(" %.9g", v);
        std::printf("\n");
    }
    return 0;
}
