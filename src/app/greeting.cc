#include "app/greeting.h"

#include "core/formatter.h"

namespace hola {

GreetingService::GreetingService(ConfigResolver* resolver, std::string hostname, Logger logger, int64_t cache_ns)
    : hostname_(std::move(hostname)),
      value_(resolver, "hello", "Hola de %s", cache_ns, /*log_changes=*/true, std::move(logger)) {}

void GreetingService::refresh(int64_t now_ns) {
    std::unique_lock<std::mutex> lock(mu_, std::try_to_lock);
    if (!lock.owns_lock()) {
        // Someone else is refreshing. Only the very first request has nothing to serve yet and
        // has to wait for it.
        if (version_.load(std::memory_order_acquire) != 0) return;
        lock.lock();
        if (version_.load(std::memory_order_acquire) != 0) return;
    }
    if (now_ns <= reload_after_.load(std::memory_order_relaxed)) return;  // lost the race, all done
    const uint64_t before = value_.version();
    const std::string& pattern = value_.get(now_ns);
    if (value_.version() != before || !current_) {
        auto snap = std::make_shared<GreetingSnapshot>();
        snap->pattern = pattern;
        FormatResult r = java_format(pattern, hostname_);
        snap->ok = r.ok;
        snap->text = std::move(r.text);
        snap->error = std::move(r.error);
        snap->version = version_.load(std::memory_order_relaxed) + 1;
        current_ = snap;
        version_.store(snap->version, std::memory_order_release);
    }
    reload_after_.store(value_.reload_after_ns(), std::memory_order_relaxed);
}

std::shared_ptr<const GreetingSnapshot> GreetingService::snapshot() {
    std::lock_guard<std::mutex> lock(mu_);
    return current_;
}

}  // namespace hola
