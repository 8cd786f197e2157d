// Circuit breaker, bulkhead and metrics for the downstream call, modelled on Hystrix 1.5.1.
//
// The reference wraps its `aloha` client in a Hystrix command with a static fallback
// (/root/reference/src/main/java/com/redhat/developers/msa/hola/HolaResource.java:104-112) and
// overrides two settings (/root/reference/src/main/resources/config.properties:17,19): a 600 ms
// execution timeout and a request-volume threshold of 5. Everything else is the library default,
// reproduced here as the defaults of BreakerConfig.
//
// Behaviour reproduced:
//   * health counts come from a rolling window made of `health_interval_ms` buckets and only
//     *closed* buckets are visible, so the breaker reacts up to one interval late - as the
//     stream-based metrics of Hystrix 1.5 do;
//   * the circuit opens when, within the window, requests >= request_volume_threshold and
//     error percentage >= error_threshold_pct; short-circuited calls do not count as requests;
//   * after sleep_window_ms one trial call is let through per sleep window; its success closes
//     the circuit and clears the health window, its failure leaves the circuit open;
//   * a bounded number of concurrent executions (thread-pool isolation with core size 10 and no
//     queue in the reference); calls beyond that are rejected straight to the fallback;
//   * every outcome is recorded as Hystrix event types, from which /hystrix.stream is rendered.
//
// The class is clock-agnostic: every method takes "now" in milliseconds, which lets the unit
// tests drive the state machine with a synthetic clock. All methods are thread-safe.
//
// Concurrency. Hystrix keeps one set of counters per command and every request thread writes to
// it; here the request path must not meet another core's cache lines, let alone a shared lock.
// Counters are therefore kept in *shards* - one per event-loop thread, each behind its own mutex
// that only its owner takes on the request path - and merged when somebody looks (the metrics
// stream, the trip decision). What is shared is small and lock-free: the open flag, the time of
// the last trial (one compare-and-swap decides who gets the single trial per sleep window) and
// the bulkhead's count of running calls. The trip decision itself depends only on *closed* health
// buckets, so it is re-evaluated once per health interval (500 ms) - under the command's lock,
// merging the shards - and not on every call. All shards cut their buckets at the same instants
// (a shared origin), which makes the merged window identical to the single window it replaces:
// for any single-threaded sequence of calls the behaviour is exactly what it was.
#pragma once

#include <array>
#include <atomic>
#include <cstdint>
#include <mutex>
#include <string>
#include <vector>

#include "core/properties.h"

namespace hola {

enum class HystrixEvent : int {
    kSuccess = 0,
    kFailure,
    kTimeout,
    kShortCircuited,
    kThreadPoolRejected,
    kSemaphoreRejected,
    kBadRequest,
    kFallbackSuccess,
    kFallbackFailure,
    kFallbackRejection,
    kFallbackMissing,
    kFallbackEmit,
    kExceptionThrown,
    kEmit,
    kResponseFromCache,
    kCollapsed,
    kCount
};

constexpr int kHystrixEventCount = static_cast<int>(HystrixEvent::kCount);

struct BreakerConfig {
    // hystrix.command.default.*
    bool breaker_enabled = true;           // circuitBreaker.enabled
    int request_volume_threshold = 20;     // circuitBreaker.requestVolumeThreshold (reference: 5)
    int sleep_window_ms = 5000;            // circuitBreaker.sleepWindowInMilliseconds
    int error_threshold_pct = 50;          // circuitBreaker.errorThresholdPercentage
    bool force_open = false;               // circuitBreaker.forceOpen
    bool force_closed = false;             // circuitBreaker.forceClosed
    int timeout_ms = 1000;                 // execution.isolation.thread.timeoutInMilliseconds (reference: 600)
    bool timeout_enabled = true;           // execution.timeout.enabled
    bool fallback_enabled = true;          // fallback.enabled
    int rolling_window_ms = 10000;         // metrics.rollingStats.timeInMilliseconds
    int rolling_buckets = 10;              // metrics.rollingStats.numBuckets
    int health_interval_ms = 500;          // metrics.healthSnapshot.intervalInMilliseconds
    int percentile_window_ms = 60000;      // metrics.rollingPercentile.timeInMilliseconds
    int percentile_buckets = 6;            // metrics.rollingPercentile.numBuckets
    int percentile_bucket_size = 100;      // metrics.rollingPercentile.bucketSize
    // hystrix.threadpool.default.*
    int pool_core_size = 10;               // coreSize
    int pool_max_queue_size = -1;          // maxQueueSize (-1: hand-off, no queueing)
    int pool_queue_rejection_threshold = 5;  // queueSizeRejectionThreshold

    // Applies Archaius-style keys (`hystrix.command.default.…`, `hystrix.threadpool.default.…`)
    // found in `props`; unknown keys are ignored. Returns the number of keys applied.
    int apply(const Properties& props);
    // Applies one key/value pair; returns false when the key is not a recognised setting.
    bool apply_key(const std::string& key, const std::string& value);

    // The settings of the reference: library defaults plus its two overrides.
    static BreakerConfig reference_defaults();
};

// Fixed-interval rolling window of event counts in which only closed buckets are visible.
class RollingEventWindow {
 public:
    RollingEventWindow(int bucket_ms, int buckets);
    void add(int64_t now_ms, HystrixEvent e, int64_t n = 1);
    // Max-tracking variant for "rolling max" style metrics: keeps the maximum per bucket.
    void update_max(int64_t now_ms, HystrixEvent slot, int64_t v);
    // Sum (or, for max windows, maximum) over the closed buckets still inside the window.
    int64_t sum(int64_t now_ms, HystrixEvent e);
    int64_t max(int64_t now_ms, HystrixEvent slot);
    void reset(int64_t now_ms);
    // Starts the window with its first bucket beginning at `origin_ms` unless it is running
    // already (an unstarted window otherwise starts at the time of its first use).
    void start_at(int64_t origin_ms) {
        if (!started_) reset(origin_ms);
    }

 private:
    using Bucket = std::array<int64_t, kHystrixEventCount>;
    void roll(int64_t now_ms);
    int bucket_ms_;
    int n_;
    bool started_ = false;
    int64_t current_start_ = 0;
    Bucket current_{};
    std::vector<Bucket> closed_;  // ring, newest at head_
    int head_ = 0;
    int filled_ = 0;
};

// Rolling latency distribution: a few coarse buckets of bounded sample reservoirs.
class RollingPercentile {
 public:
    RollingPercentile(int window_ms, int buckets, int bucket_size);
    void add(int64_t now_ms, int value);
    struct Snapshot {
        std::vector<int> sorted;
        int mean = 0;
        int percentile(double pct) const;
    };
    Snapshot snapshot(int64_t now_ms);
    // Appends the samples of the closed buckets to *out (for merging several reservoirs).
    void collect(int64_t now_ms, std::vector<int>* out);
    static Snapshot make_snapshot(std::vector<int>&& samples);
    void start_at(int64_t origin_ms) {
        if (!started_) {
            started_ = true;
            current_start_ = origin_ms;
        }
    }

 private:
    struct Bucket {
        std::vector<int> samples;
        size_t next = 0;
    };
    void roll(int64_t now_ms);
    int bucket_ms_;
    int n_;
    int bucket_size_;
    bool started_ = false;
    int64_t current_start_ = 0;
    Bucket current_;
    std::vector<Bucket> closed_;
    int head_ = 0;
    int filled_ = 0;
};

struct HealthCounts {
    int64_t total = 0;
    int64_t errors = 0;
    int error_pct = 0;
};

struct CommandSnapshot {
    int64_t now_ms = 0;
    bool circuit_open = false;
    HealthCounts health;
    std::array<int64_t, kHystrixEventCount> rolling{};
    std::array<int64_t, kHystrixEventCount> cumulative{};
    int current_concurrent = 0;
    int64_t rolling_max_concurrent = 0;
    RollingPercentile::Snapshot latency_execute;
    RollingPercentile::Snapshot latency_total;
};

struct PoolSnapshot {
    int64_t now_ms = 0;
    int active = 0;
    int64_t completed_tasks = 0;
    int64_t task_count = 0;
    int largest_pool_size = 0;
    int pool_size = 0;
    int64_t rolling_executed = 0;
    int64_t rolling_rejected = 0;
    int64_t rolling_max_active = 0;
};

// One command key + its thread pool, as a single unit (the reference has exactly one of each).
class HystrixCommand {
 public:
    static constexpr int kMaxShards = 64;

    HystrixCommand(std::string name, std::string group, std::string pool, const BreakerConfig& cfg);
    ~HystrixCommand();
    HystrixCommand(HystrixCommand&& other) noexcept;  // for factory functions; not while in use
    HystrixCommand(const HystrixCommand&) = delete;
    HystrixCommand& operator=(const HystrixCommand&) = delete;

    const std::string& name() const { return name_; }
    const std::string& group() const { return group_; }
    const std::string& pool_name() const { return pool_; }
    const BreakerConfig& config() const { return cfg_; }

    enum class Admission { kRun, kShortCircuited, kPoolRejected };

    // `shard` in every call below: which set of counters the event is booked to. Pass the index
    // of the calling event-loop thread (any value works - it is taken modulo kMaxShards - but
    // two threads that share a shard share its lock).

    // Decides whether a call may run. kRun reserves a pool slot that must be given back with
    // exactly one of complete_*(). The two rejections record their events (including the
    // successful fallback, when fallbacks are enabled) and need no further call.
    Admission admit(int64_t now_ms, int shard = 0);

    // Outcomes of an admitted call. `execute_ms` is the time spent in the call itself.
    void complete_success(int64_t now_ms, int execute_ms, int shard = 0);
    void complete_failure(int64_t now_ms, int execute_ms, int shard = 0);
    void complete_timeout(int64_t now_ms, int execute_ms, int shard = 0);

    // Views of the state machine, for tests and the metrics stream (they merge the shards).
    bool is_open(int64_t now_ms);
    HealthCounts health(int64_t now_ms);
    CommandSnapshot snapshot(int64_t now_ms);
    PoolSnapshot pool_snapshot(int64_t now_ms);
    bool has_executed() const { return executed_once_.load(std::memory_order_relaxed); }

 private:
    struct Shard;
    static constexpr int64_t kUnset = INT64_MIN;

    Shard& shard(int index);
    void touch(int64_t now_ms);                      // fixes the windows' origin at the first use of the command
    void sync_health(Shard& s);                      // s.mu held: applies a pending reset of the health window
    void record(Shard& s, int64_t now_ms, HystrixEvent e);  // s.mu held
    void finish(int64_t now_ms, HystrixEvent e, int execute_ms, int shard);
    bool allow_request(int64_t now_ms);
    bool evaluate_open_locked(int64_t now_ms);       // mu_ held: the trip decision on the merged health counts
    HealthCounts health_locked(int64_t now_ms);      // mu_ held
    void reset_health_locked(int64_t now_ms);        // mu_ held

    std::string name_, group_, pool_;
    BreakerConfig cfg_;

    // Shared and lock-free on the request path.
    std::atomic<bool> circuit_open_{false};
    std::atomic<int64_t> opened_or_tested_ms_{0};
    std::atomic<int> active_{0};
    std::atomic<int> largest_active_{0};
    std::atomic<bool> executed_once_{false};
    std::atomic<int64_t> origin_ms_{kUnset};          // bucket alignment of the rolling windows
    std::atomic<int64_t> health_origin_ms_{kUnset};   // ... of the health windows (moves when the circuit closes)
    std::atomic<int64_t> evaluated_bucket_{-1};       // last health bucket for which "stays closed" was established

    std::mutex mu_;  // trip evaluation, health reset, shard creation
    std::array<std::atomic<Shard*>, kMaxShards> shards_{};
};

// Rendering of the two `data:` payloads of /hystrix.stream with Hystrix 1.5.1's field names.
std::string hystrix_command_json(const HystrixCommand& cmd, const CommandSnapshot& s);
std::string hystrix_pool_json(const HystrixCommand& cmd, const PoolSnapshot& s);

}  // namespace hola
