// rh_order.hpp — in which order does the reference's collector pop the items of one rank?
//
// OTFCollector keeps the items of a rank in a robin_hood::unordered_map<uint64_t, CollectorsItem> keyed by
// `kmer_ & submask_` and always takes `begin()` (include/otf_collector.h:190-202,355-357).  `begin()` is the first
// occupied slot of the table's array, erase(begin()) only shifts the rest of that cluster back by one
// (include/robin_hood.h:1398-1414), and nothing is inserted into a rank while it is being drained (arcs only lead to
// higher ranks), so the pop order of a rank is simply the ascending SLOT order of its table after the last insert.
// That order decides which of two items that collide on a later key was "first" — its kmer_/shift_count_ survive
// the absorb (otf_collector.h:204-208) — and with it which k-mers get probed, so it is part of the result.
//
// This is a restatement of where robin_hood (include/robin_hood.h, unordered_map with MaxLoadFactor100 = 80) puts a
// key, nothing else: hash_int (:748-759) and keyToIdx (:1347-1360), the info byte = (distance + 1) * inc + a few hash
// bits, insertion in front of the first poorer entry with shiftUp (:2345-2398, :1379-1396), the info halving of
// try_increase_info (:2399-2425), and growth by re-inserting the old array in slot order (increase_size :2428-2456,
// rehashPowerOfTwo :2216-2250, insert_move :1452-1490).  Payloads are 32-bit indices into the caller's item vector.
// tests/test_host_cpu.py drives it against the real container through oracle/_ref (ref_rh_pop_order).
#pragma once

#include <algorithm>
#include <cstdint>
#include <stdexcept>
#include <vector>

namespace trxh
{

class rh_slots
{
public:
    static constexpr uint32_t npos = UINT32_MAX;

    void clear()
    {
        info_.clear();
        keys_.clear();
        vals_.clear();
        mask_ = num_ = max_allowed_ = 0;
        inc_ = 32;
        shift_ = 0;
        mult_ = UINT64_C(0xc4ceb9fe1a85ec53);
    }
    // Table::clear() (include/robin_hood.h:1682-1700): the elements go, the array keeps its size, the info counters start
    // over; the element limit and the hash multiplier stay as they were — a set that was big once keeps iterating in the
    // order of a big table
    void clear_keep_capacity()
    {
        if (num_ == 0)
            return;
        num_ = 0;
        std::fill(info_.begin(), info_.end(), uint8_t{0});
        info_[with_buffer_] = 1;
        inc_ = 32;
        shift_ = 0;
    }
    bool empty() const { return num_ == 0; }
    size_t size() const { return num_; }

    uint32_t find(uint64_t key) const
    {
        if (mask_ == 0)
            return npos;
        size_t idx;
        uint32_t info;
        key_to_idx(key, idx, info);
        do
        {
            if (info == info_[idx] && keys_[idx] == key)
                return vals_[idx];
            ++idx;
            info += inc_;
        } while (info <= info_[idx]);
        return npos;
    }

    // operator[] for a key that find() did not see
    void insert_new(uint64_t key, uint32_t value)
    {
        for (int attempt = 0; attempt < 256; ++attempt)
        {
            size_t idx = 0;
            uint32_t info = 0;
            if (mask_ != 0)
            {
                key_to_idx(key, idx, info);
                while (info < info_[idx])
                    ++idx, info += inc_;
                while (info == info_[idx])
                    ++idx, info += inc_;
            }
            if (num_ >= max_allowed_)
            {
                increase_size();
                continue;
            }
            place(idx, info, key, value);
            return;
        }
        throw std::overflow_error("robin_hood::map overflow");
    }

    // f(value) for every entry in slot order = the order in which `begin()` hands them out
    template <class F>
    void for_each(F && f) const
    {
        for (size_t i = 0; i < with_buffer_; ++i)
            if (info_[i])
                f(vals_[i]);
    }

    // f(key) in slot order: iteration order of a robin_hood::unordered_set<size_t> filled by the same inserts
    template <class F>
    void for_each_key(F && f) const
    {
        for (size_t i = 0; i < with_buffer_; ++i)
            if (info_[i])
                f(keys_[i]);
    }
    // insert() of a set: a key that is already there is left alone
    void insert_key(uint64_t key)
    {
        if (find(key) == npos)
            insert_new(key, 0);
    }

private:
    static size_t max_allowed_for(size_t n) { return n * 80 / 100; }
    static size_t with_buffer_for(size_t n)
    {
        size_t const a = max_allowed_for(n);
        return n + (a < 0xFF ? a : 0xFF);
    }

    void key_to_idx(uint64_t key, size_t & idx, uint32_t & info) const
    {
        uint64_t h = key;
        h ^= h >> 33;
        h *= UINT64_C(0xff51afd7ed558ccd);
        h ^= h >> 33;
        h *= mult_;
        h ^= h >> 33;
        info = inc_ + static_cast<uint32_t>((h & 31) >> shift_);
        idx = static_cast<size_t>(h >> 5) & mask_;
    }

    void init_data(size_t n)
    {
        num_ = 0;
        mask_ = n - 1;
        max_allowed_ = max_allowed_for(n);
        with_buffer_ = with_buffer_for(n);
        info_.assign(with_buffer_ + 8, 0);
        info_[with_buffer_] = 1; // sentinel: ends every probe sequence
        keys_.assign(with_buffer_, 0);
        vals_.assign(with_buffer_, 0);
        inc_ = 32;
        shift_ = 0;
    }

    bool try_increase_info()
    {
        if (inc_ <= 2)
            return false;
        inc_ >>= 1;
        ++shift_;
        for (size_t i = 0; i < ((with_buffer_ + 7) & ~size_t{7}) && i < info_.size(); ++i)
            info_[i] = static_cast<uint8_t>((info_[i] >> 1) & 0x7f);
        info_[with_buffer_] = 1;
        max_allowed_ = max_allowed_for(mask_ + 1);
        return true;
    }

    void increase_size()
    {
        if (mask_ == 0)
        {
            init_data(8);
            return;
        }
        size_t const allowed = max_allowed_for(mask_ + 1);
        if (num_ < allowed && try_increase_info())
            return;
        if (num_ * 2 < allowed)
        {
            mult_ += UINT64_C(0xc4ceb9fe1a85ec54); // nextHashMultiplier: same size, different mixing
            rehash(mask_ + 1);
        }
        else
            rehash((mask_ + 1) * 2);
    }

    void rehash(size_t n)
    {
        std::vector<uint8_t> old_info;
        std::vector<uint64_t> old_keys;
        std::vector<uint32_t> old_vals;
        old_info.swap(info_);
        old_keys.swap(keys_);
        old_vals.swap(vals_);
        size_t const old_with_buffer = with_buffer_;
        init_data(n);
        for (size_t i = 0; i < old_with_buffer; ++i)
            if (old_info[i])
                insert_move(old_keys[i], old_vals[i]);
    }

    void insert_move(uint64_t key, uint32_t value)
    {
        if (max_allowed_ == 0 && !try_increase_info())
            throw std::overflow_error("robin_hood::map overflow");
        size_t idx;
        uint32_t info;
        key_to_idx(key, idx, info);
        while (info <= info_[idx])
            ++idx, info += inc_;
        place(idx, info, key, value);
    }

    // the entry goes to `idx`; whatever sits in [idx, next empty slot) moves up by one and gets one step poorer
    void place(size_t idx, uint32_t info, uint64_t key, uint32_t value)
    {
        size_t const insertion_idx = idx;
        uint8_t const insertion_info = static_cast<uint8_t>(info);
        if (static_cast<uint32_t>(insertion_info) + inc_ > 0xFF)
            max_allowed_ = 0;
        while (info_[idx] != 0)
            ++idx;
        for (size_t i = idx; i != insertion_idx; --i)
        {
            keys_[i] = keys_[i - 1];
            vals_[i] = vals_[i - 1];
            info_[i] = static_cast<uint8_t>(info_[i - 1] + inc_);
            if (static_cast<uint32_t>(info_[i]) + inc_ > 0xFF)
                max_allowed_ = 0;
        }
        keys_[insertion_idx] = key;
        vals_[insertion_idx] = value;
        info_[insertion_idx] = insertion_info;
        ++num_;
    }

    std::vector<uint8_t> info_;
    std::vector<uint64_t> keys_;
    std::vector<uint32_t> vals_;
    uint64_t mult_{UINT64_C(0xc4ceb9fe1a85ec53)};
    size_t mask_{0}, num_{0}, max_allowed_{0}, with_buffer_{0};
    uint32_t inc_{32}, shift_{0};
};

} // namespace trxh
