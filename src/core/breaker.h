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
    HystrixCommand(std::string name, std::string group, std::string pool, const BreakerConfig& cfg);

    const std::string& name() const { return name_; }
    const std::string& group() const { return group_; }
    const std::string& pool_name() const { return pool_; }
    const BreakerConfig& config() const { return cfg_; }

    enum class Admission { kRun, kShortCircuited, kPoolRejected };

    // Decides whether a call may run. kRun reserves a pool slot that must be given back with
    // exactly one of complete_*(). The two rejections record their events (including the
    // successful fallback, when fallbacks are enabled) and need no further call.
    Admission admit(int64_t now_ms);

    // Outcomes of an admitted call. `execute_ms` is the time spent in the call itself.
    void complete_success(int64_t now_ms, int execute_ms);
    void complete_failure(int64_t now_ms, int execute_ms);
    void complete_timeout(int64_t now_ms, int execute_ms);

    // Pure views of the state machine, for tests and the metrics stream.
    bool is_open(int64_t now_ms);
    HealthCounts health(int64_t now_ms);
    CommandSnapshot snapshot(int64_t now_ms);
    PoolSnapshot pool_snapshot(int64_t now_ms);
    bool has_executed() const { return executed_once_.load(std::memory_order_relaxed); }

 private:
    bool allow_request_locked(int64_t now_ms);
    bool is_open_locked(int64_t now_ms);
    HealthCounts health_locked(int64_t now_ms);
    void record_locked(int64_t now_ms, HystrixEvent e);
    void finish_locked(int64_t now_ms, HystrixEvent e, int execute_ms);

    const std::string name_, group_, pool_;
    const BreakerConfig cfg_;
    std::mutex mu_;
    bool circuit_open_ = false;
    int64_t opened_or_tested_ms_ = 0;
    RollingEventWindow health_window_;
    RollingEventWindow rolling_window_;
    RollingEventWindow max_window_;       // slot kEmit: command concurrency, slot kSuccess: pool active
    RollingEventWindow pool_window_;      // slot kSuccess: executed, kThreadPoolRejected: rejected
    std::array<int64_t, kHystrixEventCount> cumulative_{};
    RollingPercentile latency_execute_;
    RollingPercentile latency_total_;
    int active_ = 0;
    int largest_active_ = 0;
    int64_t completed_tasks_ = 0;
    std::atomic<bool> executed_once_{false};  // read by the metrics stream without the lock
};

// Rendering of the two `data:` payloads of /hystrix.stream with Hystrix 1.5.1's field names.
std::string hystrix_command_json(const HystrixCommand& cmd, const CommandSnapshot& s);
std::string hystrix_pool_json(const HystrixCommand& cmd, const PoolSnapshot& s);

}  // namespace hola
