// The greeting of GET /api/hola: `String.format(translation, hostname)`
// (/root/reference/src/main/java/com/redhat/developers/msa/hola/HolaResource.java:56-66).
//
// The reference builds a fresh resolver chain inside every request, so (as far as can be told
// without a JVM) its 5-second cache never carries over from one request to the next and, in
// -Dconf mode, the properties file is opened and parsed per request. The
// externally visible guarantee is only "an edit to the file shows up within 5 seconds".
//
// Here there is one process-wide resolver with a real 5 s cache. The formatted text is kept as
// an immutable snapshot with a version number; worker threads compare an atomic version on each
// request and re-copy the snapshot only when it has changed. The request path therefore does two
// relaxed atomic loads and, at most once per 5 s on one thread, a file read.
#pragma once

#include <atomic>
#include <cstdint>
#include <functional>
#include <memory>
#include <mutex>
#include <string>

#include "core/config.h"

namespace hola {

struct GreetingSnapshot {
    uint64_t version = 0;
    bool ok = true;      // false: the pattern is one String.format would throw on -> HTTP 500
    std::string text;    // response body when ok
    std::string error;   // JVM-style exception text when !ok
    std::string pattern;
};

class GreetingService {
 public:
    using Logger = std::function<void(const std::string&)>;

    GreetingService(ConfigResolver* resolver, std::string hostname, Logger logger,
                    int64_t cache_ns = 5000000000ll);

    // Cheap check, called per request with the loop's cached monotonic clock; re-resolves the
    // pattern when the cache period is over. Only one thread refreshes at a time; the others
    // keep serving the previous snapshot.
    void poll(int64_t now_ns) {
        if (now_ns > reload_after_.load(std::memory_order_relaxed)) refresh(now_ns);
    }
    uint64_t version() const { return version_.load(std::memory_order_acquire); }
    std::shared_ptr<const GreetingSnapshot> snapshot();

    const std::string& hostname() const { return hostname_; }

 private:
    void refresh(int64_t now_ns);

    std::string hostname_;
    std::mutex mu_;
    CachedConfigValue value_;
    std::shared_ptr<const GreetingSnapshot> current_;
    std::atomic<int64_t> reload_after_{INT64_MIN};
    std::atomic<uint64_t> version_{0};
};

}  // namespace hola
