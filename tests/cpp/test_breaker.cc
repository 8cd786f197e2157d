// Circuit-breaker state machine driven by a synthetic clock.
#include <atomic>
#include <string>
#include <thread>
#include <vector>

#include "core/breaker.h"
#include "testing.h"

using namespace hola;

namespace {

HystrixCommand make(BreakerConfig cfg = BreakerConfig::reference_defaults()) {
    return HystrixCommand("AlohaService#aloha()", "http://aloha:8080/", "http://aloha:8080/", cfg);
}

// Runs one call that fails/succeeds instantly at time t. Returns the admission.
HystrixCommand::Admission fail_at(HystrixCommand& c, int64_t t) {
    const auto a = c.admit(t);
    if (a == HystrixCommand::Admission::kRun) c.complete_failure(t, 1);
    return a;
}
HystrixCommand::Admission succeed_at(HystrixCommand& c, int64_t t) {
    const auto a = c.admit(t);
    if (a == HystrixCommand::Admission::kRun) c.complete_success(t, 1);
    return a;
}

}  // namespace

TEST(breaker_reference_settings) {
    const BreakerConfig c = BreakerConfig::reference_defaults();
    CHECK_EQ(c.timeout_ms, 600);               // config.properties:17
    CHECK_EQ(c.request_volume_threshold, 5);   // config.properties:19
    CHECK_EQ(c.error_threshold_pct, 50);
    CHECK_EQ(c.sleep_window_ms, 5000);
    CHECK_EQ(c.pool_core_size, 10);
    Properties p;
    CHECK(p.load("hystrix.command.default.execution.isolation.thread.timeoutInMilliseconds=250\n"
                 "hystrix.command.default.circuitBreaker.requestVolumeThreshold=3\n"
                 "hystrix.command.default.circuitBreaker.sleepWindowInMilliseconds = 1000\n"
                 "hystrix.threadpool.default.coreSize=2\n"
                 "unrelated.key=1\n"
                 "hystrix.command.default.circuitBreaker.enabled=notabool\n"));
    BreakerConfig d = BreakerConfig::reference_defaults();
    CHECK_EQ(d.apply(p), 4);
    CHECK_EQ(d.timeout_ms, 250);
    CHECK_EQ(d.request_volume_threshold, 3);
    CHECK_EQ(d.sleep_window_ms, 1000);
    CHECK_EQ(d.pool_core_size, 2);
    CHECK(d.breaker_enabled);
}

TEST(breaker_trips_at_threshold_after_snapshot_interval) {
    HystrixCommand c = make();
    // Five failures inside one 500 ms health bucket: the breaker cannot see them yet.
    for (int i = 0; i < 5; ++i) CHECK(fail_at(c, 1000 + i * 10) == HystrixCommand::Admission::kRun);
    CHECK(!c.is_open(1100));
    CHECK(fail_at(c, 1200) == HystrixCommand::Admission::kRun);  // still admitted
    // Once the bucket closes the counts become visible and the circuit opens.
    CHECK(c.is_open(1500));
    CHECK(c.admit(1501) == HystrixCommand::Admission::kShortCircuited);
    const HealthCounts h = c.health(1501);
    CHECK_EQ(h.total, int64_t{6});  // short-circuited calls are not requests
    CHECK_EQ(h.error_pct, 100);
}

TEST(breaker_needs_volume_and_error_rate) {
    HystrixCommand c = make();
    for (int i = 0; i < 4; ++i) fail_at(c, 1000 + i);
    CHECK(!c.is_open(2000));  // 4 < 5 requests
    HystrixCommand d = make();
    for (int i = 0; i < 6; ++i) succeed_at(d, 1000 + i);
    for (int i = 0; i < 5; ++i) fail_at(d, 1010 + i);
    CHECK(!d.is_open(2000));  // 5/11 = 45 %
    fail_at(d, 2001);
    CHECK(d.is_open(2600));   // 6/12 = 50 %
}

TEST(breaker_half_open_probe) {
    HystrixCommand c = make();
    for (int i = 0; i < 5; ++i) fail_at(c, 1000 + i);
    CHECK(c.is_open(1500));  // opened at t=1500
    CHECK(c.admit(3000) == HystrixCommand::Admission::kShortCircuited);
    CHECK(c.admit(6500) == HystrixCommand::Admission::kShortCircuited);  // "now > opened + 5000" is strict
    // First call after the sleep window is the probe; the next one is refused again.
    CHECK(c.admit(6501) == HystrixCommand::Admission::kRun);
    CHECK(c.admit(6502) == HystrixCommand::Admission::kShortCircuited);
    c.complete_failure(6600, 99);  // probe failed: stays open for another window
    CHECK(c.is_open(6601));
    CHECK(c.admit(11000) == HystrixCommand::Admission::kShortCircuited);
    CHECK(c.admit(11502) == HystrixCommand::Admission::kRun);
    c.complete_success(11550, 48);  // probe succeeded: closed, history wiped
    CHECK(!c.is_open(11551));
    CHECK_EQ(c.health(12100).total, int64_t{0});
    CHECK(succeed_at(c, 12101) == HystrixCommand::Admission::kRun);
}

TEST(breaker_window_expiry) {
    HystrixCommand c = make();
    for (int i = 0; i < 4; ++i) fail_at(c, 1000 + i);
    // 11 s later the failures have left the 10 s window; one more failure does not trip.
    fail_at(c, 12000);
    CHECK(!c.is_open(12600));
    CHECK_EQ(c.health(12600).total, int64_t{1});
}

TEST(breaker_bulkhead) {
    HystrixCommand c = make();
    for (int i = 0; i < 10; ++i) CHECK(c.admit(1000) == HystrixCommand::Admission::kRun);
    CHECK(c.admit(1000) == HystrixCommand::Admission::kPoolRejected);
    c.complete_success(1010, 10);
    CHECK(c.admit(1011) == HystrixCommand::Admission::kRun);
    const PoolSnapshot p = c.pool_snapshot(2500);
    CHECK_EQ(p.active, 10);
    CHECK_EQ(p.rolling_rejected, int64_t{1});
    CHECK_EQ(p.rolling_executed, int64_t{11});
    CHECK_EQ(p.rolling_max_active, int64_t{10});
    // A pool rejection counts as an error towards the breaker.
    CHECK_EQ(c.health(2500).errors, int64_t{1});
}

TEST(breaker_timeouts_count_as_errors) {
    HystrixCommand c = make();
    for (int i = 0; i < 5; ++i) {
        CHECK(c.admit(1000 + i) == HystrixCommand::Admission::kRun);
        c.complete_timeout(1600 + i, 600);
    }
    CHECK(c.is_open(2200));
    const CommandSnapshot s = c.snapshot(3000);
    CHECK_EQ(s.rolling[static_cast<size_t>(HystrixEvent::kTimeout)], int64_t{5});
    CHECK_EQ(s.rolling[static_cast<size_t>(HystrixEvent::kFallbackSuccess)], int64_t{5});
}

TEST(breaker_force_flags_and_disable) {
    BreakerConfig cfg = BreakerConfig::reference_defaults();
    cfg.force_open = true;
    HystrixCommand open = make(cfg);
    CHECK(open.admit(1) == HystrixCommand::Admission::kShortCircuited);
    cfg = BreakerConfig::reference_defaults();
    cfg.force_closed = true;
    HystrixCommand closed = make(cfg);
    for (int i = 0; i < 20; ++i) fail_at(closed, 1000 + i);
    CHECK(closed.admit(3000) == HystrixCommand::Admission::kRun);
    cfg = BreakerConfig::reference_defaults();
    cfg.breaker_enabled = false;
    HystrixCommand off = make(cfg);
    for (int i = 0; i < 20; ++i) fail_at(off, 1000 + i);
    CHECK(!off.is_open(3000));
    CHECK(off.admit(3000) == HystrixCommand::Admission::kRun);
}

TEST(breaker_stream_json_shape) {
    HystrixCommand c = make();
    succeed_at(c, 1000);
    fail_at(c, 1001);
    const std::string cmd = hystrix_command_json(c, c.snapshot(12000));
    const char* fields[] = {"\"type\":\"HystrixCommand\"",
                            "\"name\":\"AlohaService#aloha()\"",
                            "\"group\":\"http://aloha:8080/\"",
                            "\"isCircuitBreakerOpen\":false",
                            "\"rollingCountSuccess\":",
                            "\"rollingCountFallbackSuccess\":",
                            "\"latencyExecute\":{\"0\":",
                            "\"99.5\":",
                            "\"latencyTotal_mean\":",
                            "\"propertyValue_circuitBreakerRequestVolumeThreshold\":5",
                            "\"propertyValue_executionIsolationThreadTimeoutInMilliseconds\":600",
                            "\"propertyValue_executionIsolationStrategy\":\"THREAD\"",
                            "\"propertyValue_executionIsolationThreadPoolKeyOverride\":null",
                            "\"reportingHosts\":1",
                            "\"threadPool\":\"http://aloha:8080/\""};
    for (const char* f : fields) {
        if (cmd.find(f) == std::string::npos) std::fprintf(stderr, "  missing %s in %s\n", f, cmd.c_str());
        CHECK(cmd.find(f) != std::string::npos);
    }
    const std::string pool = hystrix_pool_json(c, c.pool_snapshot(12000));
    CHECK(pool.find("\"type\":\"HystrixThreadPool\"") != std::string::npos);
    CHECK(pool.find("\"currentCorePoolSize\":10") != std::string::npos);
    CHECK(pool.find("\"rollingCountThreadsExecuted\":") != std::string::npos);
    CHECK(pool.find("\"propertyValue_queueSizeRejectionThreshold\":5") != std::string::npos);
}

TEST(rolling_percentile) {
    RollingPercentile p(60000, 6, 100);
    for (int i = 1; i <= 100; ++i) p.add(1000, i);
    CHECK_EQ(p.snapshot(5000).sorted.size(), size_t{0});  // bucket still open
    const auto s = p.snapshot(11000);
    CHECK_EQ(s.sorted.size(), size_t{100});
    CHECK_EQ(s.mean, 50);
    CHECK_EQ(s.percentile(0), 1);
    CHECK_EQ(s.percentile(50), 51);
    CHECK_EQ(s.percentile(100), 100);
    CHECK_EQ(p.snapshot(80000).sorted.size(), size_t{0});  // aged out
}

// The same schedule of calls booked to one shard and spread over many shards must be
// indistinguishable: shards cut their buckets at the same instants, so the merged window is the
// single window.
TEST(breaker_shards_merge_to_the_same_state_machine) {
    uint64_t rng = 0x9E3779B97F4A7C15ull;
    auto next = [&rng] {
        rng ^= rng << 13;
        rng ^= rng >> 7;
        rng ^= rng << 17;
        return rng;
    };
    for (int round = 0; round < 50; ++round) {
        HystrixCommand one = make(), many = make();
        int64_t now = 5000;
        struct Pending { int64_t due; int outcome; int shard; };
        std::vector<Pending> pending;
        for (int step = 0; step < 120; ++step) {
            now += static_cast<int64_t>(next() % 400);
            for (size_t i = 0; i < pending.size();) {
                if (pending[i].due <= now) {
                    const Pending p = pending[i];
                    pending.erase(pending.begin() + static_cast<long>(i));
                    if (p.outcome == 0) { one.complete_success(now, 3); many.complete_success(now, 3, p.shard); }
                    else if (p.outcome == 1) { one.complete_failure(now, 3); many.complete_failure(now, 3, p.shard); }
                    else { one.complete_timeout(now, 600); many.complete_timeout(now, 600, p.shard); }
                } else {
                    ++i;
                }
            }
            const int shard = static_cast<int>(next() % 7);
            const auto a = one.admit(now), b = many.admit(now, shard);
            CHECK(a == b);
            if (a == HystrixCommand::Admission::kRun) {
                pending.push_back(Pending{now + static_cast<int64_t>(next() % 900), static_cast<int>(next() % 4 == 0 ? 0 : 1 + next() % 2), shard});
            }
            CHECK_EQ(one.is_open(now), many.is_open(now));
            const HealthCounts h1 = one.health(now), h2 = many.health(now);
            CHECK_EQ(h1.total, h2.total);
            CHECK_EQ(h1.errors, h2.errors);
        }
        const CommandSnapshot s1 = one.snapshot(now + 1000), s2 = many.snapshot(now + 1000);
        for (int e = 0; e < kHystrixEventCount; ++e) {
            CHECK_EQ(s1.rolling[static_cast<size_t>(e)], s2.rolling[static_cast<size_t>(e)]);
            CHECK_EQ(s1.cumulative[static_cast<size_t>(e)], s2.cumulative[static_cast<size_t>(e)]);
        }
        CHECK_EQ(s1.rolling_max_concurrent, s2.rolling_max_concurrent);
        CHECK_EQ(s1.latency_total.sorted.size(), s2.latency_total.sorted.size());
        const PoolSnapshot p1 = one.pool_snapshot(now + 1000), p2 = many.pool_snapshot(now + 1000);
        CHECK_EQ(p1.rolling_executed, p2.rolling_executed);
        CHECK_EQ(p1.rolling_rejected, p2.rolling_rejected);
        CHECK_EQ(p1.completed_tasks, p2.completed_tasks);
        CHECK_EQ(p1.rolling_max_active, p2.rolling_max_active);
    }
}

// Eight threads, each on its own shard, against one command while a ninth takes snapshots: totals
// must add up and the bulkhead must never admit more than its size (run under TSan by `make tsan`).
TEST(breaker_concurrent_shards) {
    BreakerConfig cfg = BreakerConfig::reference_defaults();
    cfg.force_closed = true;  // keep admitting; the trip logic is exercised by the other tests
    cfg.pool_core_size = 4;
    HystrixCommand c = make(cfg);
    constexpr int kThreads = 8, kCalls = 20000;
    std::atomic<int> running{0}, max_running{0}, admitted{0}, rejected{0};
    std::atomic<bool> stop{false};
    std::thread watcher([&] {
        int64_t t = 1000;
        while (!stop.load()) {
            const CommandSnapshot s = c.snapshot(t);
            CHECK(s.current_concurrent >= 0 && s.current_concurrent <= 4);
            c.pool_snapshot(t);
            c.is_open(t);
            t += 7;
        }
    });
    std::vector<std::thread> threads;
    for (int k = 0; k < kThreads; ++k) {
        threads.emplace_back([&, k] {
            for (int i = 0; i < kCalls; ++i) {
                const int64_t now = 1000 + i / 10;
                if (c.admit(now, k) == HystrixCommand::Admission::kRun) {
                    const int r = running.fetch_add(1) + 1;
                    int m = max_running.load();
                    while (r > m && !max_running.compare_exchange_weak(m, r)) {
                    }
                    admitted.fetch_add(1);
                    running.fetch_sub(1);
                    if (i % 3 == 0) c.complete_failure(now, 1, k); else c.complete_success(now, 1, k);
                } else {
                    rejected.fetch_add(1);
                }
            }
        });
    }
    for (auto& t : threads) t.join();
    stop.store(true);
    watcher.join();
    CHECK(max_running.load() <= 4);
    CHECK_EQ(admitted.load() + rejected.load(), kThreads * kCalls);
    const CommandSnapshot s = c.snapshot(1000 + kCalls / 10 + 20000);
    const auto cum = [&](HystrixEvent e) { return s.cumulative[static_cast<size_t>(e)]; };
    CHECK_EQ(cum(HystrixEvent::kSuccess) + cum(HystrixEvent::kFailure), static_cast<int64_t>(admitted.load()));
    CHECK_EQ(cum(HystrixEvent::kThreadPoolRejected), static_cast<int64_t>(rejected.load()));
    CHECK_EQ(s.current_concurrent, 0);
    CHECK_EQ(c.pool_snapshot(1).completed_tasks, static_cast<int64_t>(admitted.load()));
}
