// hibf_layout.cpp — see hibf_layout.hpp.  Behavioural restatement of seqan::hibf's sketch + layout code on plain vectors;
// citations are relative to /root/reference/lib/hibf.  Compiled with -ffp-contract=off (host/Makefile): the floating-point
// expressions below must round exactly like the reference's.
#include "hibf_layout.hpp"

#include <algorithm>
#include <bit>
#include <cmath>
#include <limits>
#include <numeric>
#include <queue>
#include <random>
#include <stdexcept>
#include <atomic>
#include <string>
#include <thread>
#include <unordered_map>

namespace trxh::hibf
{

namespace
{
constexpr size_t none = std::numeric_limits<size_t>::max();

// fn(i) for i in [begin, end) on the host's cores, handed out one index at a time.  Only used where the iterations write
// to disjoint places and the result does not depend on their order (this toolchain has no OpenMP runtime).
template <class F>
void parallel_for(size_t begin, size_t end, F && fn)
{
    size_t const n = end > begin ? end - begin : 0;
    unsigned const workers = static_cast<unsigned>(std::min<size_t>(std::max(1u, std::thread::hardware_concurrency()), n));
    if (workers <= 1 || n < 8)
    {
        for (size_t i = begin; i < end; ++i)
            fn(i);
        return;
    }
    std::atomic<size_t> next{begin};
    std::vector<std::thread> pool;
    for (unsigned w = 0; w < workers; ++w)
        pool.emplace_back([&] {
            for (size_t i = next.fetch_add(1); i < end; i = next.fetch_add(1))
                fn(i);
        });
    for (std::thread & t : pool)
        t.join();
}

size_t next_multiple_of_64(size_t v) { return ((v + 63) >> 6) << 6; }          // include/hibf/misc/next_multiple_of_64.hpp
size_t divide_and_ceil(size_t a, size_t b) { return (a + (b - 1)) / b; }       // include/hibf/misc/divide_and_ceil.hpp
size_t subtract_empty_bins(size_t tmax, double fraction)                       // include/hibf/misc/subtract_empty_bins.hpp
{
    if (fraction == 0.0 || tmax <= 2u)
        return tmax;
    size_t const number_of_empty_bins = std::clamp<size_t>(tmax * fraction, 1, tmax - 2);
    return tmax - number_of_empty_bins;
}
} // namespace

// ------------------------------------------------------------------------------------------------ HyperLogLog
hyperloglog::hyperloglog(uint8_t bits) : bits_(bits), size_(1ULL << bits), data_(size_, 0u)
{
    if (bits < 5u || bits > 32u)
        throw std::invalid_argument("[HyperLogLog] bit width must be in the range [5,32].");
    double correction_factor;
    switch (size_)
    {
    case 32: correction_factor = 0.697; break;
    case 64: correction_factor = 0.709; break;
    default: correction_factor = 0.7213 / (1.0 + 1.079 / size_); break;
    }
    normalization_factor_ = correction_factor * size_ * size_;
    rank_mask_ = (1ULL << bits) - 1u;
}

void hyperloglog::add(uint64_t value)
{
    // src/sketch/hyperloglog.cpp:54-70: one 64x64→128 multiply folded to 64 bits; the top `bits` bits pick the register
    __uint128_t product = value;
    product *= 0x9E3779B97F4A7C15ULL;
    uint64_t const hash = static_cast<uint64_t>(product) ^ static_cast<uint64_t>(product >> 64);
    uint64_t const index = hash >> (64 - bits_);
    uint8_t const rank = static_cast<uint8_t>(std::countl_zero((hash << bits_) | rank_mask_) + 1);
    data_[index] = std::max(rank, data_[index]);
}

double hyperloglog::estimate() const
{
    // src/sketch/hyperloglog.cpp:72-125.  The reference sums 2^-register in eight float lanes (AVX): lane l of the packed
    // sum collects, from every group of eight consecutive registers, the one at offset 7 - l (simde_mm256_set_ps lists the
    // highest lane first); the lanes are then added up in lane order.  Float addition is not associative, so the lanes
    // are kept.
    static float const * const expectation = [] { // 2^-i, exact in float (include/hibf/sketch/hyperloglog.hpp:113-119)
        static float table[64];
        for (int i = 0; i < 64; ++i)
            table[i] = 1.0f / static_cast<float>(1ULL << i);
        return table;
    }();
    float lane[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    uint8_t const * const d = data_.data();
    for (size_t i = 0; i < size_; i += 8)
    {
        lane[7] += expectation[d[i + 0]];
        lane[6] += expectation[d[i + 1]];
        lane[5] += expectation[d[i + 2]];
        lane[4] += expectation[d[i + 3]];
        lane[3] += expectation[d[i + 4]];
        lane[2] += expectation[d[i + 5]];
        lane[1] += expectation[d[i + 6]];
        lane[0] += expectation[d[i + 7]];
    }
    float sum = 0;
    for (size_t l = 0; l < 8; ++l)
        sum += lane[l];
    double estimate = normalization_factor_ / static_cast<double>(sum);
    if (estimate <= 2.5 * size_) // small range: linear counting of empty registers
    {
        uint32_t zeros = 0;
        for (size_t i = 0; i < size_; ++i)
            zeros += (d[i] == 0u);
        if (zeros != 0u)
            estimate = size_ * std::log(static_cast<double>(size_) / zeros);
    }
    return estimate;
}

__attribute__((optimize("O3"))) void hyperloglog::merge(hyperloglog const & other)
{
    uint8_t * __restrict__ a = data_.data();
    uint8_t const * __restrict__ b = other.data_.data();
    for (size_t i = 0; i < size_; ++i) // byte-wise maximum: vectorises to pmaxub
        a[i] = a[i] > b[i] ? a[i] : b[i];
}

// ------------------------------------------------------------------------------------------------ config
void config::validate_and_set_defaults()
{
    if (number_of_user_bins == 0u)
        throw std::invalid_argument{"[HIBF CONFIG ERROR] You did not set the required config::number_of_user_bins."};
    if (number_of_hash_functions == 0u || number_of_hash_functions > 5u)
        throw std::invalid_argument{"[HIBF CONFIG ERROR] config::number_of_hash_functions must be in [1,5]."};
    if (maximum_fpr <= 0.0 || maximum_fpr >= 1.0)
        throw std::invalid_argument{"[HIBF CONFIG ERROR] config::maximum_fpr must be in (0.0,1.0)."};
    if (relaxed_fpr <= 0.0 || relaxed_fpr >= 1.0)
        throw std::invalid_argument{"[HIBF CONFIG ERROR] config::relaxed_fpr must be in (0.0,1.0)."};
    if (relaxed_fpr < maximum_fpr)
        throw std::invalid_argument{"[HIBF CONFIG ERROR] config::relaxed_fpr must be greater than or equal to config::maximum_fpr."};
    if (threads == 0u)
        throw std::invalid_argument{"[HIBF CONFIG ERROR] config::threads must be greater than 0."};
    if (sketch_bits < 5u || sketch_bits > 32u)
        throw std::invalid_argument{"[HIBF CONFIG ERROR] config::sketch_bits must be in [5,32]."};
    if (tmax == 0)
        tmax = next_multiple_of_64(std::ceil(std::sqrt(number_of_user_bins)));
    else if (tmax % 64 != 0)
        tmax = next_multiple_of_64(tmax);
    if (empty_bin_fraction < 0.0 || empty_bin_fraction >= 1.0)
        throw std::invalid_argument{"[HIBF CONFIG ERROR] config::empty_bin_fraction must be in [0.0,1.0)."};
    if (alpha < 0.0)
        throw std::invalid_argument{"[HIBF CONFIG ERROR] config::alpha must be positive."};
    if (max_rearrangement_ratio < 0.0 || max_rearrangement_ratio > 1.0)
        throw std::invalid_argument{"[HIBF CONFIG ERROR] config::max_rearrangement_ratio must be in [0.0,1.0]."};
    if (disable_estimate_union || max_rearrangement_ratio == 0.0)
        disable_rearrangement = true;
}

// ------------------------------------------------------------------------------------------------ corrections
std::vector<double> compute_fpr_correction(double fpr, size_t hash_count, size_t t_max)
{
    // src/layout/compute_fpr_correction.cpp:15-37: by how much a user bin split over `split` technical bins has to grow so
    // that the union of its parts keeps the false-positive rate
    size_t const max_tb = next_multiple_of_64(t_max);
    std::vector<double> correction(max_tb + 1u, 0.0);
    correction[1] = 1.0;
    double const numerator = std::log1p(-std::exp(std::log(fpr) / hash_count));
    for (size_t split = 2u; split <= max_tb; ++split)
    {
        double const log_target_fpr = std::log1p(-std::exp(std::log1p(-fpr) / split));
        correction[split] = numerator / std::log1p(-std::exp(log_target_fpr / hash_count));
    }
    return correction;
}

double compute_relaxed_fpr_correction(double fpr, double relaxed_fpr, size_t hash_count)
{
    // src/layout/compute_relaxed_fpr_correction.cpp:13-26: merged bins may be smaller (their hits are re-checked below)
    double const numerator = std::log1p(-std::exp(std::log(fpr) / hash_count));
    double const denominator = std::log1p(-std::exp(std::log(relaxed_fpr) / hash_count));
    return numerator / denominator;
}

size_t bin_size_in_bits(double fpr, size_t hash_count, size_t elements)
{
    double const numerator{-static_cast<double>(elements * hash_count)};
    double const denominator{std::log(1 - std::exp(std::log(fpr) / hash_count))};
    double const result{std::ceil(numerator / denominator)};
    return result;
}

// ------------------------------------------------------------------------------------------------ toolbox
namespace
{

// src/sketch/toolbox.cpp:23-34
void sort_by_cardinalities(std::vector<size_t> const & kmer_counts, std::vector<size_t> & positions)
{
    std::sort(positions.begin(), positions.end(), [&](size_t a, size_t b) { return kmer_counts[a] > kmer_counts[b]; });
}

// src/sketch/toolbox.cpp:36-52: estimates[j'] = |union of user bins j' .. j| for every j' <= j
void precompute_union_estimates_for(std::vector<uint64_t> & estimates, std::vector<hyperloglog> const & sketches,
                                    std::vector<size_t> const & counts, std::vector<size_t> const & positions, int64_t j)
{
    hyperloglog temp = sketches[positions[j]];
    estimates[j] = counts[positions[j]];
    for (int64_t j_prime = j - 1; j_prime >= 0; --j_prime)
        estimates[j_prime] = static_cast<uint64_t>(temp.merge_and_estimate(sketches[positions[j_prime]]));
}

// src/sketch/toolbox.cpp:54-71: estimates[j] = |union of user bins 0 .. j|
void precompute_initial_union_estimates(std::vector<uint64_t> & estimates, std::vector<hyperloglog> const & sketches,
                                        std::vector<size_t> const & counts, std::vector<size_t> const & positions)
{
    estimates.resize(sketches.size());
    hyperloglog temp = sketches[positions[0]];
    estimates[0] = counts[positions[0]];
    for (size_t j = 1; j < positions.size(); ++j)
        estimates[j] = static_cast<uint64_t>(temp.merge_and_estimate(sketches[positions[j]]));
}

struct neighbor
{
    size_t id;
    double dist;
    bool operator>(neighbor const & other) const { return dist > other.dist; }
};
using prio_queue = std::priority_queue<neighbor, std::vector<neighbor>, std::greater<neighbor>>;
struct entry
{
    size_t id;
    prio_queue pq;
};
using distance_matrix = std::vector<entry>;
using id_map = std::unordered_map<size_t, size_t>; // the reference's robin_hood map is only looked up, never iterated
                                                   // while it holds more than one element
struct clustering_node
{
    size_t left, right;
    hyperloglog hll;
};

// src/sketch/toolbox.cpp:352-372 (fixed seed)
void random_shuffle(distance_matrix & dist, id_map & remaining_ids)
{
    size_t const n = dist.size();
    std::mt19937_64 gen(0x7E1E5665D46800E5ULL);
    for (size_t i = 0; i + 1 < n; ++i)
    {
        std::uniform_int_distribution<size_t> distrib(i, n - 1);
        size_t const swap_i = distrib(gen);
        size_t const id = dist[i].id, swap_id = dist[swap_i].id;
        std::swap(dist[i], dist[swap_i]);
        std::swap(remaining_ids[id], remaining_ids[swap_id]);
    }
}

// src/sketch/toolbox.cpp:374-412: compacts the still-active queues to the front
void prune(distance_matrix & dist, id_map & remaining_ids)
{
    if (dist.empty())
        return;
    size_t valid_end = 0, invalid_start = dist.size() - 1;
    while (valid_end != invalid_start)
    {
        if (remaining_ids.count(dist[valid_end].id))
            ++valid_end;
        else if (!remaining_ids.count(dist[invalid_start].id))
            --invalid_start;
        else
        {
            std::swap(dist[valid_end], dist[invalid_start]);
            remaining_ids.at(dist[valid_end].id) = valid_end;
        }
    }
    if (remaining_ids.count(dist[valid_end].id))
        ++valid_end;
    dist.resize(valid_end);
}

// src/sketch/toolbox.cpp:414-441
bool rotate(std::vector<clustering_node> & clustering, size_t previous_rightmost, size_t first, size_t id)
{
    if (id == previous_rightmost)
        return true;
    clustering_node & curr = clustering[id - first];
    if (curr.left == none)
        return false;
    if (rotate(clustering, previous_rightmost, first, curr.left))
        return true;
    if (rotate(clustering, previous_rightmost, first, curr.right))
    {
        std::swap(curr.left, curr.right);
        return true;
    }
    return false;
}

// src/sketch/toolbox.cpp:443-460
void trace(std::vector<clustering_node> const & clustering, std::vector<size_t> & permutation, size_t previous_rightmost,
           size_t first, size_t id)
{
    clustering_node const & curr = clustering[id - first];
    if (curr.left == none)
    {
        if (id != previous_rightmost)
            permutation.push_back(id);
        return;
    }
    trace(clustering, permutation, previous_rightmost, first, curr.left);
    trace(clustering, permutation, previous_rightmost, first, curr.right);
}

// src/sketch/toolbox.cpp:120-350 as one thread runs it: agglomerative clustering of positions[first, last) by estimated
// Jaccard distance; the leaves of the tree, left to right, are appended to `permutation`
void cluster_bins(std::vector<hyperloglog> const & sketches, std::vector<size_t> & positions, std::vector<size_t> & permutation,
                  size_t first, size_t last)
{
    size_t const n = sketches.size();
    size_t const chunk_size = std::floor(std::sqrt(n));
    size_t const prune_steps = chunk_size;
    size_t steps_without_prune = 0;
    distance_matrix dist;
    dist.reserve(n + 1);
    id_map remaining_ids;
    std::vector<clustering_node> clustering;
    clustering.reserve(2 * n);
    std::vector<double> estimates;
    estimates.reserve(2 * n);
    size_t min_id_seen = none;
    size_t new_id = last - 1;

    for (size_t id = first; id < last; ++id)
    {
        clustering.push_back({none, none, sketches[positions[id]]});
        estimates.emplace_back(sketches[positions[id]].estimate());
    }
    size_t previous_rightmost = none;
    if (first != 0)
    {
        // one bin of overlap with the previous group: it takes part in the clustering under the id `last`
        size_t const actual_previous_rightmost = permutation.back();
        ++new_id;
        previous_rightmost = new_id;
        clustering.push_back({none, none, sketches[positions[actual_previous_rightmost]]});
        estimates.emplace_back(sketches[positions[actual_previous_rightmost]].estimate());
    }
    for (size_t id = first; id < first + clustering.size(); ++id)
    {
        dist.push_back({id, prio_queue{}});
        remaining_ids[id] = id - first;
    }

    double min_dist = std::numeric_limits<double>::max();
    // every queue is filled by one thread in the reference's push order, the minimum is then taken in index order: the
    // result is what one thread computes, the n^2/2 sketch merges just do not wait for each other
    parallel_for(0, clustering.size(), [&](size_t i) {
        for (size_t j = i + 1; j < clustering.size(); ++j)
        {
            hyperloglog temp = clustering[i].hll;
            double const estimate_ij = temp.merge_and_estimate(clustering[j].hll);
            double const distance = 2 - (estimates[i] + estimates[j]) / estimate_ij; // Jaccard distance estimate
            dist[i].pq.push({j + first, distance});
        }
    });
    for (size_t i = 0; i < clustering.size(); ++i)
    {
        if (dist[i].pq.empty())
            continue;
        neighbor const & curr = dist[i].pq.top();
        if (curr.dist < min_dist)
        {
            min_dist = curr.dist;
            min_id_seen = dist[i].id;
        }
    }
    random_shuffle(dist, remaining_ids);

    while (remaining_ids.size() > 1)
    {
        ++new_id;
        size_t const min_id = min_id_seen; // one thread: its candidate is the minimum
        size_t const min_index = remaining_ids.at(min_id);
        size_t const neighbor_id = dist[min_index].pq.top().id;
        clustering.push_back({min_id, neighbor_id, std::move(clustering[min_id - first].hll)});
        estimates.emplace_back(clustering.back().hll.merge_and_estimate(clustering[neighbor_id - first].hll));
        remaining_ids.erase(min_id);
        remaining_ids.erase(neighbor_id);
        remaining_ids[new_id] = min_index;
        dist[min_index] = {new_id, prio_queue{}};
        ++steps_without_prune;
        if (steps_without_prune > prune_steps)
        {
            prune(dist, remaining_ids);
            steps_without_prune = 0;
        }

        min_id_seen = none;
        min_dist = std::numeric_limits<double>::max();
        hyperloglog const new_hll = clustering.back().hll;
        // the queues are independent of each other (remaining_ids is only read here); the minimum is taken afterwards in
        // index order, like the one thread of the reference does
        parallel_for(0, dist.size(), [&](size_t i) {
            size_t const other_id = dist[i].id;
            if (other_id == new_id || !remaining_ids.count(other_id))
                return;
            hyperloglog temp = new_hll;
            double const estimate_ij = temp.merge_and_estimate(clustering[other_id - first].hll);
            double const distance = 2 - (estimates[other_id - first] + estimates.back()) / estimate_ij;
            dist[i].pq.push({new_id, distance});
            while (!remaining_ids.count(dist[i].pq.top().id)) // lazy deletion of merged neighbours
                dist[i].pq.pop();
        });
        for (size_t i = 0; i < dist.size(); ++i)
        {
            size_t const other_id = dist[i].id;
            if (other_id == new_id || !remaining_ids.count(other_id))
                continue;
            neighbor const & curr = dist[i].pq.top();
            if (curr.dist < min_dist)
            {
                min_dist = curr.dist;
                min_id_seen = other_id;
            }
        }
    }
    size_t const final_root_index = remaining_ids.begin()->second;
    size_t const final_root_id = dist[final_root_index].id;
    if (first != 0)
        rotate(clustering, previous_rightmost, first, final_root_id);
    trace(clustering, permutation, previous_rightmost, first, final_root_id);
}

// src/sketch/toolbox.cpp:86-118
void rearrange_bins(std::vector<hyperloglog> const & sketches, std::vector<size_t> const & kmer_counts,
                    std::vector<size_t> & positions, double max_ratio)
{
    std::vector<size_t> permutation;
    size_t first = 0, last = 1;
    while (first < positions.size())
    {
        // the group ends where the sizes differ too much (or the list ends)
        if (last == positions.size() || kmer_counts[positions[first]] * max_ratio > kmer_counts[positions[last]])
        {
            cluster_bins(sketches, positions, permutation, first, last);
            first = last;
        }
        ++last;
    }
    for (size_t i = 0; i < permutation.size(); ++i)
    {
        size_t swap_index = permutation[i];
        while (swap_index < i)
            swap_index = permutation[swap_index];
        std::swap(positions[i], positions[swap_index]);
    }
}

} // namespace

// ------------------------------------------------------------------------------------------------ binning
namespace
{

// include/hibf/layout/data_store.hpp: what one level of the recursion works on
struct data_store
{
    layout * hibf_layout{};
    std::vector<size_t> const * kmer_counts{};
    std::vector<hyperloglog> const * sketches{};
    std::vector<size_t> positions;
    std::vector<double> fpr_correction;
    double relaxed_fpr_correction{};
    std::vector<size_t> previous_bin_indices; // previous_level::bin_indices (empty = top level)
    std::vector<uint64_t> union_estimates;
    bool user_bins_arranged{false};
};

// src/layout/simple_binning.cpp:19-118: every user bin gets at least one technical bin of a lowest-level IBF, the extra
// technical bins go where they lower the maximum bin size most.  Returns the id of the fullest technical bin.
size_t simple_binning(data_store & data, size_t num_technical_bins)
{
    size_t const num_user_bins = data.positions.size();
    if (num_technical_bins == 0)
        num_technical_bins = next_multiple_of_64(num_user_bins);
    if (num_user_bins > num_technical_bins)
        throw std::runtime_error{"You cannot have less technical bins than user bins for this simple binning algorithm. "
                                 "Please see the hierarchical_binning algorithm or increase the number of technical bins."};
    auto const & counts = *data.kmer_counts;
    std::vector<std::vector<size_t>> matrix(num_technical_bins, std::vector<size_t>(num_user_bins, none));
    std::vector<std::vector<size_t>> trace(num_technical_bins, std::vector<size_t>(num_user_bins, none));
    size_t const extra_bins = num_technical_bins - num_user_bins + 1;

    double const init_ub_cardinality = static_cast<double>(counts[data.positions[0]]);
    for (size_t i = 0; i < extra_bins; ++i)
    {
        size_t const corrected = static_cast<size_t>(init_ub_cardinality * data.fpr_correction[i + 1]);
        matrix[i][0] = divide_and_ceil(corrected, i + 1u);
    }
    for (size_t j = 1; j < num_user_bins; ++j)
    {
        double const ub_cardinality = static_cast<double>(counts[data.positions[j]]);
        for (size_t i = j; i < j + extra_bins; ++i)
        {
            size_t minimum = none;
            for (size_t i_prime = j - 1; i_prime < i; ++i_prime)
            {
                size_t const corrected = static_cast<size_t>(ub_cardinality * data.fpr_correction[i - i_prime]);
                size_t const score = std::max<size_t>(divide_and_ceil(corrected, i - i_prime), matrix[i_prime][j - 1]);
                if (score < minimum)
                {
                    trace[i][j] = i_prime;
                    minimum = score;
                }
            }
            matrix[i][j] = minimum;
        }
    }

    size_t trace_i = num_technical_bins - 1, trace_j = num_user_bins - 1;
    size_t max_id = 0, max_size = 0, bin_id = 0;
    while (trace_j > 0)
    {
        size_t const next_i = trace[trace_i][trace_j];
        size_t const number_of_bins = trace_i - next_i;
        size_t const cardinality = counts[data.positions[trace_j]];
        size_t const corrected = static_cast<size_t>(cardinality * data.fpr_correction[number_of_bins]);
        size_t const per_bin = divide_and_ceil(corrected, number_of_bins);
        data.hibf_layout->user_bins.push_back({data.previous_bin_indices, bin_id, number_of_bins, data.positions[trace_j]});
        if (per_bin > max_size)
        {
            max_id = bin_id;
            max_size = per_bin;
        }
        bin_id += number_of_bins;
        trace_i = trace[trace_i][trace_j];
        --trace_j;
    }
    ++trace_i; // now the number of bins user bin 0 is split into
    size_t const cardinality = counts[data.positions[0]];
    size_t const corrected = static_cast<size_t>(cardinality * data.fpr_correction[trace_i]);
    size_t const per_bin = divide_and_ceil(corrected, trace_i);
    data.hibf_layout->user_bins.push_back({data.previous_bin_indices, bin_id, trace_i, data.positions[0]});
    if (per_bin > max_size)
        max_id = bin_id;
    return max_id;
}

// src/layout/hierarchical_binning.cpp: the DP over (technical bins x user bins) of one IBF — a user bin is split over
// several technical bins or merged with its neighbours into one (which opens a lower-level IBF) — minimising
// max-bin-size * #technical bins + alpha * k-mers that move to lower levels.
class hierarchical_binning
{
public:
    hierarchical_binning(data_store & d, config const & c) :
        cfg(c), data(d), num_user_bins(d.positions.size()),
        num_technical_bins(d.previous_bin_indices.empty() ? subtract_empty_bins(c.tmax, c.empty_bin_fraction)
                                                          : needed_technical_bins(d.positions.size()))
    {}

    size_t execute()
    {
        if (!data.user_bins_arranged)
        {
            sort_by_cardinalities(*data.kmer_counts, data.positions);
            if (!cfg.disable_estimate_union && !cfg.disable_rearrangement)
                rearrange_bins(*data.sketches, *data.kmer_counts, data.positions, cfg.max_rearrangement_ratio);
            data.user_bins_arranged = true;
        }
        matrix_t matrix(num_technical_bins, std::vector<size_t>(num_user_bins, none));
        matrix_t ll_matrix(num_technical_bins, std::vector<size_t>(num_user_bins, 0u));
        trace_t trace(num_technical_bins, std::vector<std::pair<size_t, size_t>>(num_user_bins, {none, none}));
        initialization(matrix, ll_matrix, trace);
        recursion(matrix, ll_matrix, trace);
        return backtracking(trace);
    }

private:
    using matrix_t = std::vector<std::vector<size_t>>;
    using trace_t = std::vector<std::vector<std::pair<size_t, size_t>>>;
    config const & cfg;
    data_store & data;
    size_t num_user_bins, num_technical_bins;

    struct maximum_bin_tracker
    {
        size_t max_id{}, max_size{};
        void update_max(size_t new_id, size_t new_size)
        {
            if (new_size > max_size)
            {
                max_id = new_id;
                max_size = new_size;
            }
        }
    };

    size_t needed_technical_bins(size_t requested_num_ub) const
    {
        size_t const needed = std::min<size_t>(next_multiple_of_64(requested_num_ub), cfg.tmax);
        return subtract_empty_bins(needed, cfg.empty_bin_fraction);
    }
    size_t max_merge_levels(size_t num_ubs_in_merge) const
    {
        size_t const lower_lvl_tbs = needed_technical_bins(num_ubs_in_merge);
        double const levels = std::log(num_ubs_in_merge) / std::log(lower_lvl_tbs);
        return static_cast<size_t>(std::ceil(levels));
    }

    void initialization(matrix_t & matrix, matrix_t & ll_matrix, trace_t & trace)
    {
        auto const & counts = *data.kmer_counts;
        // first column: user bin 0 split over 1, 2, ... technical bins
        double const ub_cardinality = static_cast<double>(counts[data.positions[0]]);
        for (size_t i = 0; i < num_technical_bins; ++i)
        {
            size_t const corrected = static_cast<size_t>(ub_cardinality * data.fpr_correction[i + 1]);
            matrix[i][0] = divide_and_ceil(corrected, i + 1u);
            trace[i][0] = {0u, 0u};
        }
        // first row: user bins 0 .. j merged into technical bin 0
        size_t sum = counts[data.positions[0]];
        if (!cfg.disable_estimate_union)
        {
            precompute_initial_union_estimates(data.union_estimates, *data.sketches, counts, data.positions);
            for (size_t j = 1; j < num_user_bins; ++j)
            {
                sum += counts[data.positions[j]];
                matrix[0][j] = data.union_estimates[j] * data.relaxed_fpr_correction;
                ll_matrix[0][j] = max_merge_levels(j + 1) * sum;
                trace[0][j] = {0u, j - 1};
            }
        }
        else
        {
            for (size_t j = 1; j < num_user_bins; ++j)
            {
                sum += counts[data.positions[j]];
                matrix[0][j] = sum * data.relaxed_fpr_correction;
                ll_matrix[0][j] = max_merge_levels(j + 1) * sum;
                trace[0][j] = {0u, j - 1};
            }
        }
    }

    void recursion(matrix_t & matrix, matrix_t & ll_matrix, trace_t & trace)
    {
        auto const & counts = *data.kmer_counts;
        // precompute_union_estimates_for(j) for every column at once: the chains of different columns do not depend on
        // each other (n^2/2 sketch merges and estimates — most of the layout's time), rows of a lower-triangular table
        std::vector<std::vector<uint64_t>> chains;
        if (!cfg.disable_estimate_union)
        {
            chains.resize(num_user_bins);
            parallel_for(1, num_user_bins, [&](size_t j) {
                chains[j].resize(j + 1);
                precompute_union_estimates_for(chains[j], *data.sketches, counts, data.positions, static_cast<int64_t>(j));
            });
        }
        for (size_t j = 1; j < num_user_bins; ++j)
        {
            size_t const current_weight = counts[data.positions[j]];
            double const ub_cardinality = static_cast<double>(current_weight);
            if (!cfg.disable_estimate_union)
            {
                std::copy(chains[j].begin(), chains[j].end(), data.union_estimates.begin());
                std::vector<uint64_t>().swap(chains[j]);
            }
            for (size_t i = 1; i < num_technical_bins; ++i)
            {
                size_t minimum = none, full_minimum = none;
                // vertical: user bin j split over i - i' technical bins
                for (size_t i_prime = 0; i_prime < i; ++i_prime)
                {
                    size_t const corrected = static_cast<size_t>(ub_cardinality * data.fpr_correction[i - i_prime]);
                    size_t score = std::max<size_t>(divide_and_ceil(corrected, i - i_prime), matrix[i_prime][j - 1]);
                    size_t full_score = score * (i + 1) + cfg.alpha * ll_matrix[i_prime][j - 1]; // size_t + double → size_t
                    if (full_score < full_minimum)
                    {
                        minimum = score;
                        full_minimum = full_score;
                        trace[i][j] = {i_prime, j - 1};
                        ll_matrix[i][j] = ll_matrix[i_prime][j - 1];
                    }
                }
                // horizontal: user bins j' + 1 .. j merged into technical bin i
                size_t j_prime = j - 1;
                size_t weight = current_weight;
                auto get_weight = [&]() -> size_t
                {
                    size_t const uncorrected = cfg.disable_estimate_union ? weight : data.union_estimates[j_prime + 1];
                    return data.relaxed_fpr_correction * uncorrected;
                };
                while (j_prime != 0 && ((i - trace[i][j_prime].first) < 2) && get_weight() < minimum)
                {
                    weight += counts[data.positions[j_prime]];
                    --j_prime;
                    size_t const score = std::max<size_t>(matrix[i - 1][j_prime], get_weight());
                    size_t const ll_kmers = ll_matrix[i - 1][j_prime] + max_merge_levels(j - j_prime) * weight;
                    size_t const full_score = score * (i + 1) + cfg.alpha * ll_kmers;
                    if (full_score < full_minimum)
                    {
                        minimum = score;
                        full_minimum = full_score;
                        trace[i][j] = {i - 1, j_prime};
                        ll_matrix[i][j] = ll_kmers;
                    }
                }
                matrix[i][j] = minimum;
            }
        }
    }

    void backtrack_merged_bin(size_t trace_j, size_t next_j, size_t bin_id, maximum_bin_tracker & max_tracker, bool is_first_row = false)
    {
        auto const & counts = *data.kmer_counts;
        size_t kmer_count = counts[data.positions[trace_j]];
        hyperloglog sketch = cfg.disable_estimate_union ? hyperloglog{} : (*data.sketches)[data.positions[trace_j]];
        data_store libf{data.hibf_layout, data.kmer_counts, data.sketches, {data.positions[trace_j]}, data.fpr_correction,
                        data.relaxed_fpr_correction};
        trace_j -= !is_first_row;
        while (trace_j != next_j)
        {
            trace_j -= is_first_row;
            if (!cfg.disable_estimate_union)
                sketch.merge((*data.sketches)[data.positions[trace_j]]);
            else
                kmer_count += counts[data.positions[trace_j]];
            libf.positions.push_back(data.positions[trace_j]);
            trace_j -= !is_first_row;
        }
        // process_merged_bin: the merged user bins get an IBF of their own one level down
        libf.previous_bin_indices = data.previous_bin_indices;
        libf.previous_bin_indices.push_back(bin_id);
        size_t const lower_max_bin = add_lower_level(libf);
        data.hibf_layout->max_bins.push_back({libf.previous_bin_indices, lower_max_bin});
        if (!cfg.disable_estimate_union)
            kmer_count = sketch.estimate();
        max_tracker.update_max(bin_id, kmer_count * data.relaxed_fpr_correction);
    }

    void backtrack_split_bin(size_t trace_j, size_t number_of_bins, size_t bin_id, maximum_bin_tracker & max_tracker)
    {
        data.hibf_layout->user_bins.push_back({data.previous_bin_indices, bin_id, number_of_bins, data.positions[trace_j]});
        size_t const cardinality = (*data.kmer_counts)[data.positions[trace_j]];
        size_t const corrected = static_cast<size_t>(cardinality * data.fpr_correction[number_of_bins]);
        max_tracker.update_max(bin_id, divide_and_ceil(corrected, number_of_bins));
    }

    size_t backtracking(trace_t const & trace)
    {
        size_t trace_i = num_technical_bins - 1, trace_j = num_user_bins - 1;
        maximum_bin_tracker max_tracker{};
        size_t bin_id = 0;
        while (trace_j > 0u && trace_i > 0u)
        {
            size_t const next_i = trace[trace_i][trace_j].first, next_j = trace[trace_i][trace_j].second;
            size_t const number_of_bins = trace_i - next_i;
            if (number_of_bins == 1 && next_j != trace_j - 1u) // merged bin
            {
                backtrack_merged_bin(trace_j, next_j, bin_id, max_tracker);
                trace_i = next_i;
                trace_j = next_j;
            }
            else // split bin
            {
                backtrack_split_bin(trace_j, number_of_bins, bin_id, max_tracker);
                trace_i = trace[trace_i][trace_j].first;
                --trace_j;
            }
            bin_id += number_of_bins;
        }
        if (trace_i == 0u && trace_j > 0u) // first row: the remaining user bins are merged into the last technical bin
            backtrack_merged_bin(trace_j, 0, bin_id, max_tracker, true);
        else if (trace_j == 0u) // first column: user bin 0 takes the remaining technical bins
            backtrack_split_bin(trace_j, trace_i + 1, bin_id, max_tracker);
        return max_tracker.max_id;
    }

    size_t add_lower_level(data_store & libf) const
    {
        size_t const number_of_user_bins = libf.positions.size();
        if (number_of_user_bins > cfg.tmax)
            return hierarchical_binning{libf, cfg}.execute();
        size_t const number_of_technical_bins = subtract_empty_bins(next_multiple_of_64(number_of_user_bins), cfg.empty_bin_fraction);
        return simple_binning(libf, number_of_technical_bins);
    }
};

} // namespace

layout compute_layout(config const & cfg, std::vector<size_t> const & kmer_counts, std::vector<hyperloglog> const & sketches)
{
    layout result;
    data_store store;
    store.hibf_layout = &result;
    store.kmer_counts = &kmer_counts;
    store.sketches = &sketches;
    store.positions.resize(cfg.number_of_user_bins);
    std::iota(store.positions.begin(), store.positions.end(), size_t{0});
    store.fpr_correction = compute_fpr_correction(cfg.maximum_fpr, cfg.number_of_hash_functions, cfg.tmax);
    store.relaxed_fpr_correction = compute_relaxed_fpr_correction(cfg.maximum_fpr, cfg.relaxed_fpr, cfg.number_of_hash_functions);
    result.top_level_max_bin_id = hierarchical_binning{store, cfg}.execute();
    // IBF levels ascending (src/layout/compute_layout.cpp:52-58; the same unstable sort on the same sequence)
    std::ranges::sort(result.max_bins, [](auto const & r, auto const & l) { return r.previous_TB_indices.size() < l.previous_TB_indices.size(); });
    return result;
}

// ------------------------------------------------------------------------------------------------ graph
graph_node make_graph(layout const & l)
{
    graph_node root;
    root.parent_bin_index = 0;
    root.max_bin_index = l.top_level_max_bin_id;
    // src/layout/graph.cpp:18-54: one node per lower-level IBF, hung below the merged bin it resolves
    for (auto const & mb : l.max_bins)
    {
        graph_node * parent = &root;
        for (size_t k = 0; k + 1 < mb.previous_TB_indices.size(); ++k)
            for (graph_node & child : parent->children)
                if (child.parent_bin_index == mb.previous_TB_indices[k])
                {
                    parent = &child;
                    break;
                }
        graph_node child;
        child.parent_bin_index = mb.previous_TB_indices.back();
        child.max_bin_index = mb.id;
        parent->children.push_back(child);
        if (parent->max_bin_index == mb.previous_TB_indices.back())
            parent->favourite_child_idx = parent->children.size() - 1;
    }
    // src/layout/graph.cpp:56-94: the user bins of every IBF, the one in the fullest technical bin first
    for (auto const & record : l.user_bins)
    {
        graph_node * current = &root;
        for (size_t bin : record.previous_TB_indices)
        {
            current->number_of_technical_bins = std::max(current->number_of_technical_bins, bin + 1);
            for (graph_node & child : current->children)
                if (child.parent_bin_index == bin)
                {
                    current = &child;
                    break;
                }
        }
        current->number_of_technical_bins = std::max(current->number_of_technical_bins, record.storage_TB_id + record.number_of_technical_bins);
        if (record.storage_TB_id == current->max_bin_index)
            current->remaining_records.insert(current->remaining_records.begin(), record);
        else
            current->remaining_records.push_back(record);
    }
    return root;
}

} // namespace trxh::hibf
