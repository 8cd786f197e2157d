#include "core/breaker.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>

#include "core/json.h"

namespace hola {

// ---------------------------------------------------------------------------------------------
// BreakerConfig

namespace {

bool parse_int(const std::string& s, int* out) {
    if (s.empty()) return false;
    char* end = nullptr;
    const long v = std::strtol(s.c_str(), &end, 10);
    while (end && (*end == ' ' || *end == '\t')) ++end;
    if (!end || *end != '\0') return false;
    *out = static_cast<int>(v);
    return true;
}

bool parse_bool(const std::string& s, bool* out) {
    if (s == "true" || s == "TRUE" || s == "True") {
        *out = true;
        return true;
    }
    if (s == "false" || s == "FALSE" || s == "False") {
        *out = false;
        return true;
    }
    return false;
}

}  // namespace

bool BreakerConfig::apply_key(const std::string& key, const std::string& value) {
    static const std::string kCmd = "hystrix.command.default.";
    static const std::string kPool = "hystrix.threadpool.default.";
    if (key.compare(0, kCmd.size(), kCmd) == 0) {
        const std::string k = key.substr(kCmd.size());
        if (k == "circuitBreaker.enabled") return parse_bool(value, &breaker_enabled);
        if (k == "circuitBreaker.requestVolumeThreshold") return parse_int(value, &request_volume_threshold);
        if (k == "circuitBreaker.sleepWindowInMilliseconds") return parse_int(value, &sleep_window_ms);
        if (k == "circuitBreaker.errorThresholdPercentage") return parse_int(value, &error_threshold_pct);
        if (k == "circuitBreaker.forceOpen") return parse_bool(value, &force_open);
        if (k == "circuitBreaker.forceClosed") return parse_bool(value, &force_closed);
        if (k == "execution.isolation.thread.timeoutInMilliseconds") return parse_int(value, &timeout_ms);
        if (k == "execution.timeout.enabled") return parse_bool(value, &timeout_enabled);
        if (k == "fallback.enabled") return parse_bool(value, &fallback_enabled);
        if (k == "metrics.rollingStats.timeInMilliseconds") return parse_int(value, &rolling_window_ms);
        if (k == "metrics.rollingStats.numBuckets") return parse_int(value, &rolling_buckets);
        if (k == "metrics.healthSnapshot.intervalInMilliseconds") return parse_int(value, &health_interval_ms);
        if (k == "metrics.rollingPercentile.timeInMilliseconds") return parse_int(value, &percentile_window_ms);
        if (k == "metrics.rollingPercentile.numBuckets") return parse_int(value, &percentile_buckets);
        if (k == "metrics.rollingPercentile.bucketSize") return parse_int(value, &percentile_bucket_size);
        return false;
    }
    if (key.compare(0, kPool.size(), kPool) == 0) {
        const std::string k = key.substr(kPool.size());
        if (k == "coreSize") return parse_int(value, &pool_core_size);
        if (k == "maxQueueSize") return parse_int(value, &pool_max_queue_size);
        if (k == "queueSizeRejectionThreshold") return parse_int(value, &pool_queue_rejection_threshold);
        return false;
    }
    return false;
}

int BreakerConfig::apply(const Properties& props) {
    int applied = 0;
    for (const auto& kv : props.entries()) {
        if (apply_key(kv.first, kv.second)) ++applied;
    }
    return applied;
}

BreakerConfig BreakerConfig::reference_defaults() {
    BreakerConfig c;
    c.timeout_ms = 600;              // config.properties:17
    c.request_volume_threshold = 5;  // config.properties:19
    return c;
}

// ---------------------------------------------------------------------------------------------
// RollingEventWindow

RollingEventWindow::RollingEventWindow(int bucket_ms, int buckets)
    : bucket_ms_(std::max(1, bucket_ms)), n_(std::max(1, buckets)), closed_(static_cast<size_t>(n_)) {}

void RollingEventWindow::reset(int64_t now_ms) {
    started_ = true;
    current_start_ = now_ms;
    current_.fill(0);
    for (auto& b : closed_) b.fill(0);
    head_ = 0;
    filled_ = 0;
}

void RollingEventWindow::roll(int64_t now_ms) {
    if (!started_) {
        reset(now_ms);
        return;
    }
    if (now_ms < current_start_ + bucket_ms_) return;
    const int64_t elapsed_buckets = (now_ms - current_start_) / bucket_ms_;
    if (elapsed_buckets > n_) {
        // Everything that was recorded has left the window.
        for (auto& b : closed_) b.fill(0);
        current_.fill(0);
        filled_ = n_;
        current_start_ += elapsed_buckets * bucket_ms_;
        return;
    }
    for (int64_t k = 0; k < elapsed_buckets; ++k) {
        head_ = (head_ + 1) % n_;
        closed_[static_cast<size_t>(head_)] = current_;
        current_.fill(0);
        if (filled_ < n_) ++filled_;
    }
    current_start_ += elapsed_buckets * bucket_ms_;
}

void RollingEventWindow::add(int64_t now_ms, HystrixEvent e, int64_t n) {
    roll(now_ms);
    current_[static_cast<size_t>(e)] += n;
}

void RollingEventWindow::update_max(int64_t now_ms, HystrixEvent slot, int64_t v) {
    roll(now_ms);
    int64_t& cell = current_[static_cast<size_t>(slot)];
    if (v > cell) cell = v;
}

int64_t RollingEventWindow::sum(int64_t now_ms, HystrixEvent e) {
    roll(now_ms);
    int64_t total = 0;
    for (const auto& b : closed_) total += b[static_cast<size_t>(e)];
    return total;
}

int64_t RollingEventWindow::max(int64_t now_ms, HystrixEvent slot) {
    roll(now_ms);
    int64_t m = 0;
    for (const auto& b : closed_) m = std::max(m, b[static_cast<size_t>(slot)]);
    return m;
}

// ---------------------------------------------------------------------------------------------
// RollingPercentile

RollingPercentile::RollingPercentile(int window_ms, int buckets, int bucket_size)
    : bucket_ms_(std::max(1, window_ms / std::max(1, buckets))),
      n_(std::max(1, buckets)),
      bucket_size_(std::max(1, bucket_size)),
      closed_(static_cast<size_t>(n_)) {}

void RollingPercentile::roll(int64_t now_ms) {
    if (!started_) {
        started_ = true;
        current_start_ = now_ms;
        return;
    }
    if (now_ms < current_start_ + bucket_ms_) return;
    const int64_t elapsed = (now_ms - current_start_) / bucket_ms_;
    const int64_t steps = std::min<int64_t>(elapsed, n_ + 1);
    for (int64_t k = 0; k < steps; ++k) {
        head_ = (head_ + 1) % n_;
        closed_[static_cast<size_t>(head_)] = std::move(current_);
        current_ = Bucket{};
        if (filled_ < n_) ++filled_;
    }
    current_start_ += elapsed * bucket_ms_;
}

void RollingPercentile::add(int64_t now_ms, int value) {
    roll(now_ms);
    if (current_.samples.size() < static_cast<size_t>(bucket_size_)) {
        current_.samples.push_back(value);
    } else {
        current_.samples[current_.next % static_cast<size_t>(bucket_size_)] = value;
    }
    ++current_.next;
}

void RollingPercentile::collect(int64_t now_ms, std::vector<int>* out) {
    roll(now_ms);
    for (const auto& b : closed_) out->insert(out->end(), b.samples.begin(), b.samples.end());
}

RollingPercentile::Snapshot RollingPercentile::make_snapshot(std::vector<int>&& samples) {
    Snapshot s;
    s.sorted = std::move(samples);
    std::sort(s.sorted.begin(), s.sorted.end());
    if (!s.sorted.empty()) {
        int64_t sum = 0;
        for (int v : s.sorted) sum += v;
        s.mean = static_cast<int>(sum / static_cast<int64_t>(s.sorted.size()));
    }
    return s;
}

RollingPercentile::Snapshot RollingPercentile::snapshot(int64_t now_ms) {
    std::vector<int> samples;
    collect(now_ms, &samples);
    return make_snapshot(std::move(samples));
}

int RollingPercentile::Snapshot::percentile(double pct) const {
    if (sorted.empty()) return 0;
    if (pct <= 0.0) return sorted.front();
    if (pct >= 100.0) return sorted.back();
    const double rank = (pct / 100.0) * static_cast<double>(sorted.size());
    const size_t lo = static_cast<size_t>(std::floor(rank));
    const size_t hi = static_cast<size_t>(std::ceil(rank));
    if (hi >= sorted.size()) return sorted.back();
    if (lo == hi) return sorted[lo];
    return sorted[lo] + static_cast<int>((rank - static_cast<double>(lo)) * (sorted[hi] - sorted[lo]));
}

// ---------------------------------------------------------------------------------------------
// HystrixCommand

struct alignas(64) HystrixCommand::Shard {
    explicit Shard(const BreakerConfig& cfg)
        : health(cfg.health_interval_ms, std::max(1, cfg.rolling_window_ms / std::max(1, cfg.health_interval_ms))),
          rolling(cfg.rolling_window_ms / std::max(1, cfg.rolling_buckets), cfg.rolling_buckets),
          max_window(cfg.rolling_window_ms / std::max(1, cfg.rolling_buckets), cfg.rolling_buckets),
          pool_window(cfg.rolling_window_ms / std::max(1, cfg.rolling_buckets), cfg.rolling_buckets),
          latency_execute(cfg.percentile_window_ms, cfg.percentile_buckets, cfg.percentile_bucket_size),
          latency_total(cfg.percentile_window_ms, cfg.percentile_buckets, cfg.percentile_bucket_size) {}

    std::mutex mu;  // the owning thread on the request path; anybody when the shards are merged
    int64_t health_origin_seen = kUnset;
    RollingEventWindow health;
    RollingEventWindow rolling;
    RollingEventWindow max_window;   // slot kEmit: command concurrency, slot kSuccess: pool active
    RollingEventWindow pool_window;  // slot kSuccess: executed, kThreadPoolRejected: rejected
    std::array<int64_t, kHystrixEventCount> cumulative{};
    RollingPercentile latency_execute;
    RollingPercentile latency_total;
    int64_t completed_tasks = 0;
};

HystrixCommand::HystrixCommand(std::string name, std::string group, std::string pool, const BreakerConfig& cfg)
    : name_(std::move(name)), group_(std::move(group)), pool_(std::move(pool)), cfg_(cfg) {}

HystrixCommand::HystrixCommand(HystrixCommand&& o) noexcept
    : name_(std::move(o.name_)), group_(std::move(o.group_)), pool_(std::move(o.pool_)), cfg_(o.cfg_) {
    circuit_open_.store(o.circuit_open_.load());
    opened_or_tested_ms_.store(o.opened_or_tested_ms_.load());
    active_.store(o.active_.load());
    largest_active_.store(o.largest_active_.load());
    executed_once_.store(o.executed_once_.load());
    origin_ms_.store(o.origin_ms_.load());
    health_origin_ms_.store(o.health_origin_ms_.load());
    evaluated_bucket_.store(o.evaluated_bucket_.load());
    for (int i = 0; i < kMaxShards; ++i) shards_[static_cast<size_t>(i)].store(o.shards_[static_cast<size_t>(i)].exchange(nullptr));
}

HystrixCommand::~HystrixCommand() {
    for (auto& s : shards_) delete s.load(std::memory_order_relaxed);
}

HystrixCommand::Shard& HystrixCommand::shard(int index) {
    std::atomic<Shard*>& slot = shards_[static_cast<size_t>(static_cast<unsigned>(index) % kMaxShards)];
    Shard* s = slot.load(std::memory_order_acquire);
    if (s) return *s;
    std::lock_guard<std::mutex> lock(mu_);
    s = slot.load(std::memory_order_relaxed);
    if (!s) {
        s = new Shard(cfg_);
        slot.store(s, std::memory_order_release);
    }
    return *s;
}

void HystrixCommand::touch(int64_t now_ms) {
    if (origin_ms_.load(std::memory_order_relaxed) != kUnset) return;
    int64_t expected = kUnset;
    if (origin_ms_.compare_exchange_strong(expected, now_ms, std::memory_order_relaxed)) {
        expected = kUnset;
        health_origin_ms_.compare_exchange_strong(expected, now_ms, std::memory_order_release);
    } else {
        // Another thread is fixing the origin this very moment; wait for the second store.
        while (health_origin_ms_.load(std::memory_order_acquire) == kUnset) {
        }
    }
}

// A shard's health window follows the command's health origin: when the origin has moved (the
// circuit closed and the history was wiped) the shard's window is wiped the next time it is used.
// Two wipes cannot fall into the same millisecond - re-opening takes at least one closed bucket -
// so the origin identifies the wipe.
void HystrixCommand::sync_health(Shard& s) {
    const int64_t origin = health_origin_ms_.load(std::memory_order_acquire);
    if (s.health_origin_seen != origin) {
        s.health.reset(origin);
        s.health_origin_seen = origin;
    }
}

void HystrixCommand::record(Shard& s, int64_t now_ms, HystrixEvent e) {
    sync_health(s);
    s.health.add(now_ms, e);
    s.rolling.start_at(origin_ms_.load(std::memory_order_relaxed));
    s.rolling.add(now_ms, e);
    ++s.cumulative[static_cast<size_t>(e)];
}

HealthCounts HystrixCommand::health_locked(int64_t now_ms) {
    HealthCounts h;
    int64_t success = 0;
    for (auto& slot : shards_) {
        Shard* s = slot.load(std::memory_order_acquire);
        if (!s) continue;
        std::lock_guard<std::mutex> lock(s->mu);
        sync_health(*s);
        success += s->health.sum(now_ms, HystrixEvent::kSuccess);
        h.errors += s->health.sum(now_ms, HystrixEvent::kFailure) + s->health.sum(now_ms, HystrixEvent::kTimeout) +
                    s->health.sum(now_ms, HystrixEvent::kThreadPoolRejected) +
                    s->health.sum(now_ms, HystrixEvent::kSemaphoreRejected);
    }
    h.total = success + h.errors;
    h.error_pct = h.total > 0 ? static_cast<int>(static_cast<double>(h.errors) / static_cast<double>(h.total) * 100.0) : 0;
    return h;
}

bool HystrixCommand::evaluate_open_locked(int64_t now_ms) {
    if (circuit_open_.load(std::memory_order_relaxed)) return true;
    const HealthCounts h = health_locked(now_ms);
    if (h.total >= cfg_.request_volume_threshold && h.error_pct >= cfg_.error_threshold_pct) {
        opened_or_tested_ms_.store(now_ms, std::memory_order_relaxed);
        circuit_open_.store(true, std::memory_order_release);
        return true;
    }
    // Closed, and it stays closed until another health bucket closes: only closed buckets count.
    const int64_t origin = health_origin_ms_.load(std::memory_order_relaxed);
    const int64_t bucket = (now_ms - origin) / std::max(1, cfg_.health_interval_ms);
    if (bucket > evaluated_bucket_.load(std::memory_order_relaxed)) evaluated_bucket_.store(bucket, std::memory_order_relaxed);
    return false;
}

void HystrixCommand::reset_health_locked(int64_t now_ms) {
    evaluated_bucket_.store(-1, std::memory_order_relaxed);
    health_origin_ms_.store(now_ms, std::memory_order_release);
}

bool HystrixCommand::allow_request(int64_t now_ms) {
    if (!cfg_.breaker_enabled) return true;
    if (cfg_.force_open) return false;
    bool open = circuit_open_.load(std::memory_order_acquire);
    if (!open) {
        const int64_t origin = health_origin_ms_.load(std::memory_order_relaxed);
        const int64_t bucket = (now_ms - origin) / std::max(1, cfg_.health_interval_ms);
        if (bucket > evaluated_bucket_.load(std::memory_order_relaxed)) {
            std::lock_guard<std::mutex> lock(mu_);
            open = evaluate_open_locked(now_ms);
        }
    }
    if (cfg_.force_closed) return true;  // the open flag is still kept current for the metrics, as Hystrix does
    if (!open) return true;
    // Open: let one trial call through per sleep window; whoever moves the time stamp gets it.
    int64_t tested = opened_or_tested_ms_.load(std::memory_order_relaxed);
    while (now_ms > tested + cfg_.sleep_window_ms) {
        if (opened_or_tested_ms_.compare_exchange_weak(tested, now_ms, std::memory_order_relaxed)) return true;
    }
    return false;
}

HystrixCommand::Admission HystrixCommand::admit(int64_t now_ms, int shard_index) {
    touch(now_ms);
    if (!executed_once_.load(std::memory_order_relaxed)) executed_once_.store(true, std::memory_order_relaxed);
    Shard& s = shard(shard_index);
    const HystrixEvent fallback = cfg_.fallback_enabled ? HystrixEvent::kFallbackSuccess : HystrixEvent::kFallbackMissing;
    if (!allow_request(now_ms)) {
        std::lock_guard<std::mutex> lock(s.mu);
        record(s, now_ms, HystrixEvent::kShortCircuited);
        record(s, now_ms, fallback);
        s.latency_total.start_at(origin_ms_.load(std::memory_order_relaxed));
        s.latency_total.add(now_ms, 0);
        return Admission::kShortCircuited;
    }
    // Take a slot of the bulkhead, if there is one (never counting past its size, not even
    // for a moment: the count is what the metrics stream shows).
    int running = active_.load(std::memory_order_relaxed);
    while (running < cfg_.pool_core_size && !active_.compare_exchange_weak(running, running + 1, std::memory_order_acq_rel)) {
    }
    if (running >= cfg_.pool_core_size) {
        std::lock_guard<std::mutex> lock(s.mu);
        record(s, now_ms, HystrixEvent::kThreadPoolRejected);
        record(s, now_ms, fallback);
        s.pool_window.start_at(origin_ms_.load(std::memory_order_relaxed));
        s.pool_window.add(now_ms, HystrixEvent::kThreadPoolRejected);
        s.latency_total.start_at(origin_ms_.load(std::memory_order_relaxed));
        s.latency_total.add(now_ms, 0);
        return Admission::kPoolRejected;
    }
    const int now_running = running + 1;
    int largest = largest_active_.load(std::memory_order_relaxed);
    while (now_running > largest && !largest_active_.compare_exchange_weak(largest, now_running, std::memory_order_relaxed)) {
    }
    std::lock_guard<std::mutex> lock(s.mu);
    const int64_t origin = origin_ms_.load(std::memory_order_relaxed);
    s.max_window.start_at(origin);
    s.max_window.update_max(now_ms, HystrixEvent::kEmit, now_running);
    s.max_window.update_max(now_ms, HystrixEvent::kSuccess, now_running);
    s.pool_window.start_at(origin);
    s.pool_window.add(now_ms, HystrixEvent::kSuccess);
    return Admission::kRun;
}

void HystrixCommand::finish(int64_t now_ms, HystrixEvent e, int execute_ms, int shard_index) {
    touch(now_ms);
    int running = active_.load(std::memory_order_relaxed);
    while (running > 0 && !active_.compare_exchange_weak(running, running - 1, std::memory_order_acq_rel)) {
    }
    Shard& s = shard(shard_index);
    std::lock_guard<std::mutex> lock(s.mu);
    ++s.completed_tasks;
    record(s, now_ms, e);
    if (e != HystrixEvent::kSuccess)
        record(s, now_ms, cfg_.fallback_enabled ? HystrixEvent::kFallbackSuccess : HystrixEvent::kFallbackMissing);
    const int64_t origin = origin_ms_.load(std::memory_order_relaxed);
    s.latency_execute.start_at(origin);
    s.latency_execute.add(now_ms, execute_ms);
    s.latency_total.start_at(origin);
    s.latency_total.add(now_ms, execute_ms);
}

void HystrixCommand::complete_success(int64_t now_ms, int execute_ms, int shard_index) {
    finish(now_ms, HystrixEvent::kSuccess, execute_ms, shard_index);
    if (circuit_open_.load(std::memory_order_acquire)) {
        // The trial call (or a straggler from before the trip) came back fine: close the circuit
        // and forget the history, this success included.
        std::lock_guard<std::mutex> lock(mu_);
        if (circuit_open_.load(std::memory_order_relaxed)) {
            reset_health_locked(now_ms);
            circuit_open_.store(false, std::memory_order_release);
        }
    }
}

void HystrixCommand::complete_failure(int64_t now_ms, int execute_ms, int shard_index) {
    finish(now_ms, HystrixEvent::kFailure, execute_ms, shard_index);
}

void HystrixCommand::complete_timeout(int64_t now_ms, int execute_ms, int shard_index) {
    finish(now_ms, HystrixEvent::kTimeout, execute_ms, shard_index);
}

bool HystrixCommand::is_open(int64_t now_ms) {
    touch(now_ms);
    if (!cfg_.breaker_enabled) return false;
    if (cfg_.force_open) return true;
    std::lock_guard<std::mutex> lock(mu_);
    return evaluate_open_locked(now_ms);
}

HealthCounts HystrixCommand::health(int64_t now_ms) {
    touch(now_ms);
    std::lock_guard<std::mutex> lock(mu_);
    return health_locked(now_ms);
}

CommandSnapshot HystrixCommand::snapshot(int64_t now_ms) {
    touch(now_ms);
    CommandSnapshot out;
    out.now_ms = now_ms;
    {
        std::lock_guard<std::mutex> lock(mu_);
        out.circuit_open = cfg_.breaker_enabled && (cfg_.force_open || (!cfg_.force_closed && evaluate_open_locked(now_ms)));
        out.health = health_locked(now_ms);
    }
    const int64_t origin = origin_ms_.load(std::memory_order_relaxed);
    std::vector<int> execute_samples, total_samples;
    for (auto& slot : shards_) {
        Shard* s = slot.load(std::memory_order_acquire);
        if (!s) continue;
        std::lock_guard<std::mutex> lock(s->mu);
        s->rolling.start_at(origin);
        s->max_window.start_at(origin);
        s->latency_execute.start_at(origin);
        s->latency_total.start_at(origin);
        for (int i = 0; i < kHystrixEventCount; ++i) {
            out.rolling[static_cast<size_t>(i)] += s->rolling.sum(now_ms, static_cast<HystrixEvent>(i));
            out.cumulative[static_cast<size_t>(i)] += s->cumulative[static_cast<size_t>(i)];
        }
        out.rolling_max_concurrent = std::max(out.rolling_max_concurrent, s->max_window.max(now_ms, HystrixEvent::kEmit));
        s->latency_execute.collect(now_ms, &execute_samples);
        s->latency_total.collect(now_ms, &total_samples);
    }
    out.current_concurrent = active_.load(std::memory_order_relaxed);
    out.latency_execute = RollingPercentile::make_snapshot(std::move(execute_samples));
    out.latency_total = RollingPercentile::make_snapshot(std::move(total_samples));
    return out;
}

PoolSnapshot HystrixCommand::pool_snapshot(int64_t now_ms) {
    touch(now_ms);
    PoolSnapshot p;
    p.now_ms = now_ms;
    const int64_t origin = origin_ms_.load(std::memory_order_relaxed);
    for (auto& slot : shards_) {
        Shard* s = slot.load(std::memory_order_acquire);
        if (!s) continue;
        std::lock_guard<std::mutex> lock(s->mu);
        s->pool_window.start_at(origin);
        s->max_window.start_at(origin);
        p.completed_tasks += s->completed_tasks;
        p.rolling_executed += s->pool_window.sum(now_ms, HystrixEvent::kSuccess);
        p.rolling_rejected += s->pool_window.sum(now_ms, HystrixEvent::kThreadPoolRejected);
        p.rolling_max_active = std::max(p.rolling_max_active, s->max_window.max(now_ms, HystrixEvent::kSuccess));
    }
    p.active = active_.load(std::memory_order_relaxed);
    p.task_count = p.completed_tasks + p.active;
    p.largest_pool_size = largest_active_.load(std::memory_order_relaxed);
    p.pool_size = std::min(cfg_.pool_core_size, p.largest_pool_size);
    return p;
}

// ---------------------------------------------------------------------------------------------
// /hystrix.stream payloads

namespace {

void field(std::string* out, const char* name) {
    if (out->back() != '{') out->push_back(',');
    out->push_back('"');
    out->append(name);
    out->append("\":");
}
void num(std::string* out, const char* name, int64_t v) {
    field(out, name);
    out->append(std::to_string(v));
}
void boolean(std::string* out, const char* name, bool v) {
    field(out, name);
    out->append(v ? "true" : "false");
}
void str(std::string* out, const char* name, std::string_view v) {
    field(out, name);
    json_append_string(out, v);
}
void latency(std::string* out, const char* name, const RollingPercentile::Snapshot& s) {
    static const struct { const char* label; double pct; } kPoints[] = {
        {"0", 0}, {"25", 25}, {"50", 50}, {"75", 75}, {"90", 90}, {"95", 95}, {"99", 99}, {"99.5", 99.5}, {"100", 100}};
    field(out, name);
    out->push_back('{');
    bool first = true;
    for (const auto& p : kPoints) {
        if (!first) out->push_back(',');
        first = false;
        out->push_back('"');
        out->append(p.label);
        out->append("\":");
        out->append(std::to_string(s.percentile(p.pct)));
    }
    out->push_back('}');
}

}  // namespace

std::string hystrix_command_json(const HystrixCommand& cmd, const CommandSnapshot& s) {
    const BreakerConfig& c = cmd.config();
    auto r = [&](HystrixEvent e) { return s.rolling[static_cast<size_t>(e)]; };
    std::string o = "{";
    str(&o, "type", "HystrixCommand");
    str(&o, "name", cmd.name());
    str(&o, "group", cmd.group());
    num(&o, "currentTime", s.now_ms);
    boolean(&o, "isCircuitBreakerOpen", s.circuit_open);
    num(&o, "errorPercentage", s.health.error_pct);
    num(&o, "errorCount", s.health.errors);
    num(&o, "requestCount", s.health.total);
    num(&o, "rollingCountBadRequests", r(HystrixEvent::kBadRequest));
    num(&o, "rollingCountCollapsedRequests", r(HystrixEvent::kCollapsed));
    num(&o, "rollingCountEmit", r(HystrixEvent::kEmit));
    num(&o, "rollingCountExceptionsThrown", r(HystrixEvent::kExceptionThrown));
    num(&o, "rollingCountFailure", r(HystrixEvent::kFailure));
    num(&o, "rollingCountFallbackEmit", r(HystrixEvent::kFallbackEmit));
    num(&o, "rollingCountFallbackFailure", r(HystrixEvent::kFallbackFailure));
    num(&o, "rollingCountFallbackMissing", r(HystrixEvent::kFallbackMissing));
    num(&o, "rollingCountFallbackRejection", r(HystrixEvent::kFallbackRejection));
    num(&o, "rollingCountFallbackSuccess", r(HystrixEvent::kFallbackSuccess));
    num(&o, "rollingCountResponsesFromCache", r(HystrixEvent::kResponseFromCache));
    num(&o, "rollingCountSemaphoreRejected", r(HystrixEvent::kSemaphoreRejected));
    num(&o, "rollingCountShortCircuited", r(HystrixEvent::kShortCircuited));
    num(&o, "rollingCountSuccess", r(HystrixEvent::kSuccess));
    num(&o, "rollingCountThreadPoolRejected", r(HystrixEvent::kThreadPoolRejected));
    num(&o, "rollingCountTimeout", r(HystrixEvent::kTimeout));
    num(&o, "currentConcurrentExecutionCount", s.current_concurrent);
    num(&o, "rollingMaxConcurrentExecutionCount", s.rolling_max_concurrent);
    num(&o, "latencyExecute_mean", s.latency_execute.mean);
    latency(&o, "latencyExecute", s.latency_execute);
    num(&o, "latencyTotal_mean", s.latency_total.mean);
    latency(&o, "latencyTotal", s.latency_total);
    num(&o, "propertyValue_circuitBreakerRequestVolumeThreshold", c.request_volume_threshold);
    num(&o, "propertyValue_circuitBreakerSleepWindowInMilliseconds", c.sleep_window_ms);
    num(&o, "propertyValue_circuitBreakerErrorThresholdPercentage", c.error_threshold_pct);
    boolean(&o, "propertyValue_circuitBreakerForceOpen", c.force_open);
    boolean(&o, "propertyValue_circuitBreakerForceClosed", c.force_closed);
    boolean(&o, "propertyValue_circuitBreakerEnabled", c.breaker_enabled);
    str(&o, "propertyValue_executionIsolationStrategy", "THREAD");
    num(&o, "propertyValue_executionIsolationThreadTimeoutInMilliseconds", c.timeout_ms);
    num(&o, "propertyValue_executionTimeoutInMilliseconds", c.timeout_ms);
    boolean(&o, "propertyValue_executionIsolationThreadInterruptOnTimeout", true);
    field(&o, "propertyValue_executionIsolationThreadPoolKeyOverride");
    o.append("null");
    num(&o, "propertyValue_executionIsolationSemaphoreMaxConcurrentRequests", 10);
    num(&o, "propertyValue_fallbackIsolationSemaphoreMaxConcurrentRequests", 10);
    num(&o, "propertyValue_metricsRollingStatisticalWindowInMilliseconds", c.rolling_window_ms);
    boolean(&o, "propertyValue_requestCacheEnabled", true);
    boolean(&o, "propertyValue_requestLogEnabled", true);
    num(&o, "reportingHosts", 1);
    str(&o, "threadPool", cmd.pool_name());
    o.push_back('}');
    return o;
}

std::string hystrix_pool_json(const HystrixCommand& cmd, const PoolSnapshot& s) {
    const BreakerConfig& c = cmd.config();
    std::string o = "{";
    str(&o, "type", "HystrixThreadPool");
    str(&o, "name", cmd.pool_name());
    num(&o, "currentTime", s.now_ms);
    num(&o, "currentActiveCount", s.active);
    num(&o, "currentCompletedTaskCount", s.completed_tasks);
    num(&o, "currentCorePoolSize", c.pool_core_size);
    num(&o, "currentLargestPoolSize", s.largest_pool_size);
    num(&o, "currentMaximumPoolSize", c.pool_core_size);
    num(&o, "currentPoolSize", s.pool_size);
    num(&o, "currentQueueSize", 0);
    num(&o, "currentTaskCount", s.task_count);
    num(&o, "rollingCountThreadsExecuted", s.rolling_executed);
    num(&o, "rollingMaxActiveThreads", s.rolling_max_active);
    num(&o, "rollingCountCommandRejections", s.rolling_rejected);
    num(&o, "propertyValue_queueSizeRejectionThreshold", c.pool_queue_rejection_threshold);
    num(&o, "propertyValue_metricsRollingStatisticalWindowInMilliseconds", c.rolling_window_ms);
    num(&o, "reportingHosts", 1);
    o.push_back('}');
    return o;
}

}  // namespace hola
