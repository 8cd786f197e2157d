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

RollingPercentile::Snapshot RollingPercentile::snapshot(int64_t now_ms) {
    roll(now_ms);
    Snapshot s;
    for (const auto& b : closed_) s.sorted.insert(s.sorted.end(), b.samples.begin(), b.samples.end());
    std::sort(s.sorted.begin(), s.sorted.end());
    if (!s.sorted.empty()) {
        int64_t sum = 0;
        for (int v : s.sorted) sum += v;
        s.mean = static_cast<int>(sum / static_cast<int64_t>(s.sorted.size()));
    }
    return s;
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

HystrixCommand::HystrixCommand(std::string name, std::string group, std::string pool, const BreakerConfig& cfg)
    : name_(std::move(name)),
      group_(std::move(group)),
      pool_(std::move(pool)),
      cfg_(cfg),
      health_window_(cfg.health_interval_ms, std::max(1, cfg.rolling_window_ms / std::max(1, cfg.health_interval_ms))),
      rolling_window_(cfg.rolling_window_ms / std::max(1, cfg.rolling_buckets), cfg.rolling_buckets),
      max_window_(cfg.rolling_window_ms / std::max(1, cfg.rolling_buckets), cfg.rolling_buckets),
      pool_window_(cfg.rolling_window_ms / std::max(1, cfg.rolling_buckets), cfg.rolling_buckets),
      latency_execute_(cfg.percentile_window_ms, cfg.percentile_buckets, cfg.percentile_bucket_size),
      latency_total_(cfg.percentile_window_ms, cfg.percentile_buckets, cfg.percentile_bucket_size) {}

HealthCounts HystrixCommand::health_locked(int64_t now_ms) {
    HealthCounts h;
    const int64_t success = health_window_.sum(now_ms, HystrixEvent::kSuccess);
    h.errors = health_window_.sum(now_ms, HystrixEvent::kFailure) + health_window_.sum(now_ms, HystrixEvent::kTimeout) +
               health_window_.sum(now_ms, HystrixEvent::kThreadPoolRejected) +
               health_window_.sum(now_ms, HystrixEvent::kSemaphoreRejected);
    h.total = success + h.errors;
    h.error_pct = h.total > 0 ? static_cast<int>(static_cast<double>(h.errors) / static_cast<double>(h.total) * 100.0) : 0;
    return h;
}

bool HystrixCommand::is_open_locked(int64_t now_ms) {
    if (circuit_open_) return true;
    const HealthCounts h = health_locked(now_ms);
    if (h.total < cfg_.request_volume_threshold) return false;
    if (h.error_pct < cfg_.error_threshold_pct) return false;
    circuit_open_ = true;
    opened_or_tested_ms_ = now_ms;
    return true;
}

bool HystrixCommand::allow_request_locked(int64_t now_ms) {
    if (!cfg_.breaker_enabled) return true;
    if (cfg_.force_open) return false;
    if (cfg_.force_closed) {
        is_open_locked(now_ms);  // keeps the open flag current for the metrics, as Hystrix does
        return true;
    }
    if (!is_open_locked(now_ms)) return true;
    // Open: let one trial call through per sleep window.
    if (now_ms > opened_or_tested_ms_ + cfg_.sleep_window_ms) {
        opened_or_tested_ms_ = now_ms;
        return true;
    }
    return false;
}

void HystrixCommand::record_locked(int64_t now_ms, HystrixEvent e) {
    health_window_.add(now_ms, e);
    rolling_window_.add(now_ms, e);
    ++cumulative_[static_cast<size_t>(e)];
}

HystrixCommand::Admission HystrixCommand::admit(int64_t now_ms) {
    std::lock_guard<std::mutex> lock(mu_);
    executed_once_.store(true, std::memory_order_relaxed);
    if (!allow_request_locked(now_ms)) {
        record_locked(now_ms, HystrixEvent::kShortCircuited);
        record_locked(now_ms, cfg_.fallback_enabled ? HystrixEvent::kFallbackSuccess : HystrixEvent::kFallbackMissing);
        latency_total_.add(now_ms, 0);
        return Admission::kShortCircuited;
    }
    if (active_ >= cfg_.pool_core_size) {
        record_locked(now_ms, HystrixEvent::kThreadPoolRejected);
        record_locked(now_ms, cfg_.fallback_enabled ? HystrixEvent::kFallbackSuccess : HystrixEvent::kFallbackMissing);
        pool_window_.add(now_ms, HystrixEvent::kThreadPoolRejected);
        latency_total_.add(now_ms, 0);
        return Admission::kPoolRejected;
    }
    ++active_;
    largest_active_ = std::max(largest_active_, active_);
    max_window_.update_max(now_ms, HystrixEvent::kEmit, active_);
    max_window_.update_max(now_ms, HystrixEvent::kSuccess, active_);
    pool_window_.add(now_ms, HystrixEvent::kSuccess);
    return Admission::kRun;
}

void HystrixCommand::finish_locked(int64_t now_ms, HystrixEvent e, int execute_ms) {
    if (active_ > 0) --active_;
    ++completed_tasks_;
    record_locked(now_ms, e);
    if (e != HystrixEvent::kSuccess)
        record_locked(now_ms, cfg_.fallback_enabled ? HystrixEvent::kFallbackSuccess : HystrixEvent::kFallbackMissing);
    latency_execute_.add(now_ms, execute_ms);
    latency_total_.add(now_ms, execute_ms);
}

void HystrixCommand::complete_success(int64_t now_ms, int execute_ms) {
    std::lock_guard<std::mutex> lock(mu_);
    finish_locked(now_ms, HystrixEvent::kSuccess, execute_ms);
    if (circuit_open_) {
        circuit_open_ = false;
        health_window_.reset(now_ms);
    }
}

void HystrixCommand::complete_failure(int64_t now_ms, int execute_ms) {
    std::lock_guard<std::mutex> lock(mu_);
    finish_locked(now_ms, HystrixEvent::kFailure, execute_ms);
}

void HystrixCommand::complete_timeout(int64_t now_ms, int execute_ms) {
    std::lock_guard<std::mutex> lock(mu_);
    finish_locked(now_ms, HystrixEvent::kTimeout, execute_ms);
}

bool HystrixCommand::is_open(int64_t now_ms) {
    std::lock_guard<std::mutex> lock(mu_);
    if (!cfg_.breaker_enabled) return false;
    if (cfg_.force_open) return true;
    return is_open_locked(now_ms);
}

HealthCounts HystrixCommand::health(int64_t now_ms) {
    std::lock_guard<std::mutex> lock(mu_);
    return health_locked(now_ms);
}

CommandSnapshot HystrixCommand::snapshot(int64_t now_ms) {
    std::lock_guard<std::mutex> lock(mu_);
    CommandSnapshot s;
    s.now_ms = now_ms;
    s.circuit_open = cfg_.breaker_enabled && (cfg_.force_open || (!cfg_.force_closed && is_open_locked(now_ms)));
    s.health = health_locked(now_ms);
    for (int i = 0; i < kHystrixEventCount; ++i)
        s.rolling[static_cast<size_t>(i)] = rolling_window_.sum(now_ms, static_cast<HystrixEvent>(i));
    s.cumulative = cumulative_;
    s.current_concurrent = active_;
    s.rolling_max_concurrent = max_window_.max(now_ms, HystrixEvent::kEmit);
    s.latency_execute = latency_execute_.snapshot(now_ms);
    s.latency_total = latency_total_.snapshot(now_ms);
    return s;
}

PoolSnapshot HystrixCommand::pool_snapshot(int64_t now_ms) {
    std::lock_guard<std::mutex> lock(mu_);
    PoolSnapshot p;
    p.now_ms = now_ms;
    p.active = active_;
    p.completed_tasks = completed_tasks_;
    p.task_count = completed_tasks_ + active_;
    p.largest_pool_size = largest_active_;
    p.pool_size = std::min(cfg_.pool_core_size, largest_active_);
    p.rolling_executed = pool_window_.sum(now_ms, HystrixEvent::kSuccess);
    p.rolling_rejected = pool_window_.sum(now_ms, HystrixEvent::kThreadPoolRejected);
    p.rolling_max_active = max_window_.max(now_ms, HystrixEvent::kSuccess);
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
