// Span collection and shipping to Zipkin.
//
// Mirrors the reference's `Brave.Builder("hola").spanCollector(HttpSpanCollector.create(
// System.getenv("ZIPKIN_SERVER_URL"), …))`
// (/root/reference/src/main/java/com/redhat/developers/msa/hola/ZipkinResource.java:33-41):
// finished spans are queued in memory, a background thread wakes every second, encodes what has
// accumulated as one Zipkin v1 JSON array and POSTs it to <url>/api/v1/spans. When the queue is
// full spans are dropped; a dead or slow collector never touches a request.
//
// Differences from the reference, both deliberate:
//   * ZIPKIN_SERVER_URL unset means "tracing off" instead of a NullPointerException that takes
//     the whole deployment down;
//   * the queue is one single-producer ring per worker thread (no lock, no allocation on the
//     request path) instead of one shared blocking queue.
#pragma once

#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "core/trace.h"

namespace hola {

// Fixed-size record written by a worker; everything the reporter needs to build the JSON.
struct SpanRecord {
    uint64_t trace_id;
    uint64_t span_id;
    uint64_t parent_id;
    int64_t start_us;
    int64_t end_us;
    uint16_t status;      // HTTP status (0: none received)
    uint16_t url_len;
    bool has_parent;
    bool client;          // false: server span (sr/ss), true: client span (cs/cr)
    bool shared;          // ids came from the caller: no timestamp/duration of our own
    char url[217];
};
static_assert(sizeof(SpanRecord) == 264, "keep SpanRecord compact");

// Single-producer single-consumer ring of span records.
class SpanRing {
 public:
    explicit SpanRing(size_t capacity_pow2);
    // Producer side. Returns the slot to fill, or nullptr when the ring is full (span dropped).
    SpanRecord* begin_push();
    void commit_push();
    // Consumer side.
    const SpanRecord* peek() const;
    void pop();
    uint64_t dropped() const { return dropped_.load(std::memory_order_relaxed); }

 private:
    const size_t mask_;
    std::unique_ptr<SpanRecord[]> slots_;
    alignas(64) std::atomic<uint64_t> head_{0};  // next slot to write (producer)
    alignas(64) std::atomic<uint64_t> tail_{0};  // next slot to read (consumer)
    alignas(64) std::atomic<uint64_t> dropped_{0};
};

struct TracerStats {
    uint64_t spans_reported = 0;  // accepted by the collector (2xx)
    uint64_t spans_dropped = 0;   // ring overflow + batches the collector did not take
    uint64_t posts_ok = 0;
    uint64_t posts_failed = 0;
};

class Tracer {
 public:
    // `zipkin_url` empty: disabled. Otherwise e.g. "http://zipkin-query:9411".
    Tracer(std::string zipkin_url, std::string service_name, int flush_interval_ms = 1000,
           size_t ring_capacity = 1024, size_t max_batch = 1000);
    ~Tracer();

    bool enabled() const { return enabled_; }
    const std::string& error() const { return error_; }  // why a non-empty URL was rejected

    // Each worker thread calls this once and keeps the ring for its lifetime.
    SpanRing* register_producer();

    void start();
    void stop();  // flushes once more, then joins
    TracerStats stats() const;
    const Endpoint& endpoint() const { return endpoint_; }

 private:
    void thread_main();
    void flush();
    size_t flush_batch();  // one POST; returns the number of spans the collector took
    bool post(const std::string& body);

    bool enabled_ = false;
    std::string error_;
    std::string host_;
    uint16_t port_ = 80;
    std::string path_;
    Endpoint endpoint_;
    const int flush_interval_ms_;
    const size_t ring_capacity_;
    const size_t max_batch_;

    mutable std::mutex mu_;
    std::condition_variable cv_;
    bool stop_ = false;
    std::atomic<bool> stopping_{false};  // same as stop_, readable without the lock
    bool started_ = false;
    std::vector<std::unique_ptr<SpanRing>> rings_;
    std::thread thread_;
    int sock_ = -1;  // kept open between flushes when the collector allows it
    std::string body_;  // the batch being encoded (reporter thread only)

    std::atomic<uint64_t> spans_reported_{0};
    std::atomic<uint64_t> spans_failed_{0};
    std::atomic<uint64_t> posts_ok_{0};
    std::atomic<uint64_t> posts_failed_{0};
};

// Splits "http://host[:port][/prefix]" into its parts. Returns false for anything else.
bool parse_http_url(const std::string& url, std::string* host, uint16_t* port, std::string* path);

}  // namespace hola
