// charon_b200/host/stats_model.hpp — host side of the statistical model: options, the training state
// machine that only runs with --extract, and the hand-over of the (default or fitted) parameters to the
// GPU library (cg_set_model).  The per-read evaluation itself (Model::prob, call_host, call_category)
// runs on the GPU; nothing here evaluates a probability.
//
// Restates reference include/classify_stats.hpp:34-114 (TrainingData), :116-208 (Gamma/Beta fits),
// :210-254 (KDEParams ctor/fit), :341-368 (Model::train*), :395-584 (StatsModel) and the option
// defaults of include/dehost_arguments.hpp:22-43 / include/classify_arguments.hpp:16-39.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <numeric>
#include <string>
#include <vector>

#include "../../include/charon_gpu.h"
#include "../csrc/kde_defaults.inc"

namespace charon_host {

struct Options {
    bool dehost = true;
    // IO
    std::string read_file, read_file2, db;
    bool is_paired = false;
    // output
    bool run_extract = false;
    std::string category_to_extract, prefix;
    uint8_t chunk_size = 100;
    // stats
    float lo_hi_threshold = 0.15f;
    uint16_t num_reads_to_fit = 5000;
    std::string dist = "kde";
    // thresholds
    float min_quality = 15.0f;
    uint32_t min_length = 140;
    float min_compression = 0.0f;
    uint8_t confidence_threshold = 7;
    float confidence_probability_threshold = 0.0f;
    float host_unique_prop_lo_threshold = 0.05f;
    float min_proportion_difference = 0.04f;
    float min_prob_difference = 0.0f;
    // general
    std::string log_file = "charon.log";
    uint8_t threads = 1;
    uint8_t verbosity = 0;
    // B200 extensions (not in the reference CLI)
    int gpus = 1;
    uint32_t batch_reads = 16384;
    bool gpu_compression = false;    // compute the gzip compression ratio of the reads on the GPU (cg_read_scalars.compression = NULL)
    bool partition_filter = false;   // split the filter by row range over --gpus GPUs instead of replicating it

    static Options dehost_defaults() { return Options(); }
    static Options classify_defaults() {   // classify_arguments.hpp:16-39
        Options o; o.dehost = false; o.dist = "beta"; o.min_quality = 10.0f; o.min_compression = 0.15f;
        o.confidence_threshold = 2; o.min_proportion_difference = 0.0f;
        return o;
    }
};

namespace detail {
template <class V> double mean_of(const V &v) {   // classify_stats.hpp:20-23
    double sum = 0.0;
    for (auto x : v) sum += x;
    return v.empty() ? 0 : sum / double(v.size());
}
inline double variance_of(const std::vector<float> &v, double mean) {   // :25-31 (float accumulator lambda, double init)
    double acc = 0.0;
    for (float val : v) { const float a = float(acc); acc = double(a) + ((double(val) - mean) * (double(val) - mean)); }
    return v.size() <= 1 ? 0 : acc / double(v.size() - 1);
}
}  // namespace detail

struct TrainingData {   // classify_stats.hpp:34-114
    bool complete = false, pos_complete = false, neg_complete = false;
    uint16_t num_reads_to_fit = 0;
    std::vector<float> pos, neg;
    bool check_status() {
        if (pos.size() >= num_reads_to_fit) pos_complete = true;
        if (neg.size() >= num_reads_to_fit) neg_complete = true;
        if (pos_complete && neg_complete) complete = true;
        return complete;
    }
    bool add_pos(float v) { if (pos.size() < num_reads_to_fit) pos.push_back(v); else check_status(); return complete; }
    bool add_neg(float v) { if (neg.size() < num_reads_to_fit && v > 0) neg.push_back(v); else check_status(); return complete; }
    void clear() { pos.clear(); neg.clear(); }
};

struct CategoryModel {   // Model, classify_stats.hpp:261-271
    bool ready = false;
    float g_pos[3] = {25.f, 0.f, 0.02f}, g_neg[3] = {10.f, 0.f, 0.005f};   // shape, loc, scale
    float b_pos[2] = {6.f, 4.f}, b_neg[2] = {6.f, 40.f};                   // alpha, beta
    std::vector<float> k_pos, k_neg;                                       // datasets in summation order
    float h_pos = 0.1f, h_neg = 0.001f;
};

class StatsModel {
public:
    StatsModel() = default;
    StatsModel(const Options &o, size_t ncat) : opt_(o), models_(ncat), training_(ncat) {
        static const float dpos[CHARON_KDE_DEFAULT_N] = {CHARON_KDE_DEFAULT_POS};
        static const float dneg[CHARON_KDE_DEFAULT_N] = {CHARON_KDE_DEFAULT_NEG};
        for (size_t i = 0; i < ncat; ++i) {
            // KDEParams ctor sorts its dataset ascending (classify_stats.hpp:214-218)
            models_[i].k_pos.assign(dpos, dpos + CHARON_KDE_DEFAULT_N); std::sort(models_[i].k_pos.begin(), models_[i].k_pos.end());
            models_[i].k_neg.assign(dneg, dneg + CHARON_KDE_DEFAULT_N); std::sort(models_[i].k_neg.begin(), models_[i].k_neg.end());
            training_[i].num_reads_to_fit = o.num_reads_to_fit;
        }
    }
    bool ready() const { return ready_; }
    const std::vector<CategoryModel> &models() const { return models_; }
    const std::vector<TrainingData> &training() const { return training_; }   // read-only view for tests/training_trace.cpp

    void force_ready() {   // :463-473
        for (size_t i = 0; i < models_.size(); ++i) if (!models_[i].ready) train(i);
        ready_ = true;
    }
    bool add_read_to_training_data(const std::vector<float> &props) {   // :535-578
        uint8_t pos_i = 255; double max_val = 0.0; int above = 0;
        for (uint8_t i = 0; i < props.size(); ++i) {
            const float val = props[i];
            if (val > opt_.lo_hi_threshold) above += 1;
            if (val == max_val) pos_i = 255;
            else if (val > max_val) { pos_i = i; max_val = val; }
        }
        const bool add_pos = (pos_i != 255 && above == 1);
        const bool add_neg = add_pos || (above == 0);
        if (add_pos) { const bool r = training_.at(pos_i).add_pos(props[pos_i]); if (r && !models_[pos_i].ready) train_model_at(pos_i); }
        if (add_neg)
            for (uint8_t i = 0; i < props.size(); ++i) if (i != pos_i) {
                const bool r = training_.at(i).add_neg(props[i]);
                if (r && !models_[i].ready) train_model_at(i);
            }
        return ready_;
    }

    // Hand the current parameters to one GPU handle (replaces the in-process Model objects, :261-393).
    int upload(cg_handle *h, uint8_t host_index) const {
        std::vector<cg_category_model> cm(models_.size());
        const int dist = opt_.dist == "kde" ? CG_DIST_KDE : (opt_.dist == "beta" ? CG_DIST_BETA : CG_DIST_GAMMA);
        for (size_t i = 0; i < models_.size(); ++i) {
            const CategoryModel &m = models_[i];
            cg_category_model &c = cm[i];
            c.dist = dist;
            c.pos = m.k_pos.data(); c.npos = uint32_t(m.k_pos.size()); c.h_pos = m.h_pos;
            c.neg = m.k_neg.data(); c.nneg = uint32_t(m.k_neg.size()); c.h_neg = m.h_neg;
            for (int j = 0; j < 2; ++j) { c.beta_pos[j] = m.b_pos[j]; c.beta_neg[j] = m.b_neg[j]; }
            for (int j = 0; j < 3; ++j) { c.gamma_pos[j] = m.g_pos[j]; c.gamma_neg[j] = m.g_neg[j]; }
        }
        cg_thresholds t{};
        t.min_quality = opt_.min_quality; t.min_length = opt_.min_length; t.min_compression = opt_.min_compression;
        t.confidence_threshold = int8_t(opt_.confidence_threshold);
        t.confidence_probability_threshold = opt_.confidence_probability_threshold;
        t.min_hits = 0;   // uninitialised in the reference (classify_stats.hpp:406); SURVEY Appendix B-5
        t.host_unique_prop_lo_threshold = opt_.host_unique_prop_lo_threshold;
        t.min_proportion_difference = opt_.min_proportion_difference;
        t.min_prob_difference = opt_.min_prob_difference;
        t.host_index = host_index;
        // paired dehost calls call_category (src/dehost_main.cpp:470,475 -> add_paired_read(..., dehost=false))
        const int mode = (opt_.dehost && !opt_.is_paired) ? CG_MODE_DEHOST : CG_MODE_CLASSIFY;
        return cg_set_model(h, cm.data(), &t, mode);
    }

private:
    static void fit_gamma(float *g, const std::vector<float> &data) {   // GammaParams::fit :127-136
        const double mu = detail::mean_of(data);
        const double ln_mu = std::log(mu);
        std::vector<float> logs(data.size());
        for (size_t i = 0; i < data.size(); ++i) logs[i] = std::log(data[i]);
        const double s = ln_mu - detail::mean_of(logs);
        g[0] = float((3 - s + std::sqrt((s - 3) * (s - 3) + 24 * s)) / (12 * s));
        g[2] = float(mu / g[0]);
    }
    static void fit_beta(float *b, const std::vector<float> &data) {    // BetaParams::fit :171-190
        const double mu = detail::mean_of(data);
        const double var = detail::variance_of(data, mu);
        b[0] = float(mu * ((mu * (1 - mu) / var) - 1));
        b[1] = float((1 - mu) * ((mu * (1 - mu) / var) - 1));
    }
    void train(size_t i) {   // Model::train :358-368
        CategoryModel &m = models_[i];
        TrainingData &t = training_[i];
        if (opt_.dist == "kde") {          // train_kde :341-356: KDEParams::fit copies the data UNSORTED, h unchanged
            if (t.pos_complete) m.k_pos = t.pos;
            if (t.neg_complete) m.k_neg = t.neg;
        } else if (opt_.dist == "beta") {  // train_beta :315-339
            if (t.pos_complete) fit_beta(m.b_pos, t.pos);
            if (t.neg_complete) fit_beta(m.b_neg, t.neg);
        } else {                           // train_gamma :289-313
            if (t.pos_complete) fit_gamma(m.g_pos, t.pos);
            if (t.neg_complete) fit_gamma(m.g_neg, t.neg);
            else m.g_neg[1] = float(detail::mean_of(t.neg) - double(m.g_neg[0] * m.g_neg[2]));   // fit_loc :138-141
        }
        m.ready = true;
        t.clear();
    }
    void check_if_ready() { if (ready_) return; for (const auto &m : models_) if (!m.ready) return; ready_ = true; }
    void train_model_at(size_t i) { train(i); check_if_ready(); }   // :519-533

    Options opt_;
    bool ready_ = false;
    std::vector<CategoryModel> models_;
    std::vector<TrainingData> training_;
};

}  // namespace charon_host
