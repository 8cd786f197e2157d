// Log-linear latency histogram (HdrHistogram-style), header-only.
//
// Used by the load generator to report p50/p99 of GET /api/hola, the metric named in
// BASELINE.json. Values are recorded in nanoseconds. Each power of two is divided into 128
// equal sub-buckets, so any recorded value is reproduced with a relative error below 0.8 %.
// Recording is a handful of integer instructions and no allocation; histograms from several
// threads are merged by adding counts.
#pragma once

#include <algorithm>
#include <cstdint>
#include <vector>

namespace hola {

class Histogram {
 public:
    static constexpr int kSubBits = 7;
    static constexpr uint64_t kSub = 1ull << kSubBits;            // 128
    static constexpr size_t kBuckets = (64 - kSubBits + 1) * kSub;  // covers the whole uint64 range

    Histogram() : counts_(kBuckets, 0) {}

    static size_t index_of(uint64_t v) {
        if (v < kSub) return static_cast<size_t>(v);
        const int e = 63 - __builtin_clzll(v);  // floor(log2 v) >= kSubBits
        const int shift = e - kSubBits;
        return static_cast<size_t>((static_cast<uint64_t>(shift) + 1) * kSub + ((v >> shift) - kSub));
    }

    // Highest value that maps to bucket `index`.
    static uint64_t upper_edge(size_t index) {
        if (index < kSub) return index;
        const uint64_t shift = index / kSub - 1;
        const uint64_t base = (kSub + index % kSub) << shift;
        return base + ((1ull << shift) - 1);
    }

    void record(uint64_t v) {
        ++counts_[index_of(v)];
        ++total_;
        sum_ += v;
        if (v > max_) max_ = v;
        if (v < min_) min_ = v;
    }

    void merge(const Histogram& other) {
        for (size_t i = 0; i < kBuckets; ++i) counts_[i] += other.counts_[i];
        total_ += other.total_;
        sum_ += other.sum_;
        max_ = std::max(max_, other.max_);
        min_ = std::min(min_, other.min_);
    }

    void reset() {
        std::fill(counts_.begin(), counts_.end(), 0);
        total_ = 0;
        sum_ = 0;
        max_ = 0;
        min_ = UINT64_MAX;
    }

    uint64_t count() const { return total_; }
    uint64_t max() const { return total_ ? max_ : 0; }
    uint64_t min() const { return total_ ? min_ : 0; }
    double mean() const { return total_ ? static_cast<double>(sum_) / static_cast<double>(total_) : 0.0; }

    // Smallest recorded-bucket edge such that at least pct % of the samples are <= it.
    uint64_t percentile(double pct) const {
        if (total_ == 0) return 0;
        if (pct >= 100.0) return max_;
        uint64_t rank = static_cast<uint64_t>(pct / 100.0 * static_cast<double>(total_) + 0.5);
        if (rank < 1) rank = 1;
        if (rank > total_) rank = total_;
        uint64_t seen = 0;
        for (size_t i = 0; i < kBuckets; ++i) {
            seen += counts_[i];
            if (seen >= rank) return std::min(upper_edge(i), max_);
        }
        return max_;
    }

 private:
    std::vector<uint64_t> counts_;
    uint64_t total_ = 0;
    uint64_t sum_ = 0;
    uint64_t max_ = 0;
    uint64_t min_ = UINT64_MAX;
};

}  // namespace hola
