// hibf_layout.hpp — seqan::hibf's layout computation restated for the host layer (row a3 / the host half of N3).
//
// HIBFIndex::populate_index (include/index_hibf.h:110-129) hands its k-mer lists to
// seqan::hibf::hierarchical_interleaved_bloom_filter{config}, which (hibf = lib/hibf)
//   1. sketches every user bin with HyperLogLog            hibf/src/sketch/compute_sketches.cpp:27-43, hyperloglog.cpp
//   2. estimates the k-mer counts                          hibf/include/hibf/sketch/estimate_kmer_counts.hpp
//   3. lays the user bins out over a tree of IBFs          hibf/src/layout/compute_layout.cpp, hierarchical_binning.cpp,
//      (sort by size, cluster similar bins, DP over         simple_binning.cpp, compute_fpr_correction.cpp,
//       technical bins with union estimates)                compute_relaxed_fpr_correction.cpp, hibf/src/sketch/toolbox.cpp
//   4. turns the layout into a tree (hibf/src/layout/graph.cpp) and builds the IBFs bottom-up
//      (hibf/src/hierarchical_interleaved_bloom_filter.cpp:41-210, hibf/src/build/*.cpp).
// This header is steps 1-3 and the tree of step 4 as plain data; hibf_build.cpp does the build with K1 as the emplace
// step.  Everything that decides the layout is reproduced bit for bit — the float accumulation order of the HLL estimate,
// the size_t/double conversions of the DP scores, std::sort / std::priority_queue / std::mt19937_64 of the same libstdc++ —
// for the reference run with ONE thread (its clustering breaks distance ties by OpenMP thread id).
#pragma once

#include <cstddef>
#include <cstdint>
#include <optional>
#include <vector>

namespace trxh::hibf
{

// hibf/src/sketch/hyperloglog.cpp
class hyperloglog
{
public:
    explicit hyperloglog(uint8_t bits = 12);
    void add(uint64_t value);
    double estimate() const;
    void merge(hyperloglog const & other);
    double merge_and_estimate(hyperloglog const & other)
    {
        merge(other);
        return estimate();
    }
    std::vector<uint8_t> const & registers() const { return data_; }
    std::vector<uint8_t> & registers() { return data_; }

private:
    uint8_t bits_;
    uint64_t size_, rank_mask_;
    double normalization_factor_;
    std::vector<uint8_t> data_;
};

// the fields of seqan::hibf::config that matter for layout and build (hibf/include/hibf/config.hpp), with its defaults
struct config
{
    size_t number_of_user_bins{};
    size_t number_of_hash_functions{2};
    double maximum_fpr{0.05};
    double relaxed_fpr{0.3};
    size_t threads{1};
    uint8_t sketch_bits{12};
    size_t tmax{};
    double empty_bin_fraction{};
    double alpha{1.2};
    double max_rearrangement_ratio{0.5};
    bool disable_estimate_union{false};
    bool disable_rearrangement{false};
    void validate_and_set_defaults(); // hibf/src/config.cpp:69-127
};

// hibf/include/hibf/layout/layout.hpp
struct layout
{
    struct max_bin
    {
        std::vector<size_t> previous_TB_indices; // merged bins leading to this lower-level IBF
        size_t id{};                             // its fullest technical bin
    };
    struct user_bin
    {
        std::vector<size_t> previous_TB_indices;
        size_t storage_TB_id{};
        size_t number_of_technical_bins{};
        size_t idx{};
    };
    size_t top_level_max_bin_id{};
    std::vector<max_bin> max_bins;
    std::vector<user_bin> user_bins;
};

std::vector<double> compute_fpr_correction(double fpr, size_t hash_count, size_t t_max);
double compute_relaxed_fpr_correction(double fpr, double relaxed_fpr, size_t hash_count);
size_t bin_size_in_bits(double fpr, size_t hash_count, size_t elements); // hibf/include/hibf/build/bin_size_in_bits.hpp

// hibf/src/layout/compute_layout.cpp:25-66 (cfg already validated)
layout compute_layout(config const & cfg, std::vector<size_t> const & kmer_counts, std::vector<hyperloglog> const & sketches);

// hibf/include/hibf/layout/graph.hpp
struct graph_node
{
    std::vector<graph_node> children;
    size_t parent_bin_index{};
    size_t max_bin_index{};
    size_t number_of_technical_bins{};
    std::optional<size_t> favourite_child_idx;
    std::vector<layout::user_bin> remaining_records;
    bool max_bin_is_merged() const { return favourite_child_idx.has_value(); }
};
graph_node make_graph(layout const & l); // hibf/src/layout/graph.cpp:96-108

} // namespace trxh::hibf
