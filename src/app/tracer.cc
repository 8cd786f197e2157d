#include "app/tracer.h"

#include <arpa/inet.h>
#include <fcntl.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/socket.h>
#include <unistd.h>

#include <cerrno>
#include <chrono>
#include <cstring>

#include "core/http.h"
#include "net/dns.h"

namespace hola {

bool parse_http_url(const std::string& url, std::string* host, uint16_t* port, std::string* path) {
    static const std::string kScheme = "http://";
    if (url.compare(0, kScheme.size(), kScheme) != 0) return false;
    const size_t start = kScheme.size();
    size_t slash = url.find('/', start);
    std::string authority = url.substr(start, slash == std::string::npos ? std::string::npos : slash - start);
    if (authority.empty()) return false;
    *path = slash == std::string::npos ? "" : url.substr(slash);
    while (!path->empty() && path->back() == '/') path->pop_back();
    std::string p;
    if (authority[0] == '[') {  // IPv6 literal: http://[fd00::1]:8080/
        const size_t close = authority.find(']');
        if (close == std::string::npos || close == 1) return false;
        *host = authority.substr(1, close - 1);
        if (close + 1 == authority.size()) {
            *port = 80;
            return true;
        }
        if (authority[close + 1] != ':') return false;
        p = authority.substr(close + 2);
    } else {
        const size_t colon = authority.rfind(':');
        if (colon == std::string::npos) {
            *host = authority;
            *port = 80;
            return true;
        }
        *host = authority.substr(0, colon);
        p = authority.substr(colon + 1);
    }
    if (host->empty() || p.empty() || p.size() > 5) return false;
    unsigned v = 0;
    for (char c : p) {
        if (c < '0' || c > '9') return false;
        v = v * 10 + static_cast<unsigned>(c - '0');
    }
    if (v == 0 || v > 65535) return false;
    *port = static_cast<uint16_t>(v);
    return true;
}

// ---------------------------------------------------------------------------------------------
// SpanRing

namespace {
size_t round_up_pow2(size_t n) {
    size_t p = 2;
    while (p < n) p <<= 1;
    return p;
}
}  // namespace

SpanRing::SpanRing(size_t capacity_pow2)
    : mask_(round_up_pow2(capacity_pow2) - 1), slots_(new SpanRecord[round_up_pow2(capacity_pow2)]) {}

SpanRecord* SpanRing::begin_push() {
    const uint64_t head = head_.load(std::memory_order_relaxed);
    const uint64_t tail = tail_.load(std::memory_order_acquire);
    if (head - tail > mask_) {
        dropped_.store(dropped_.load(std::memory_order_relaxed) + 1, std::memory_order_relaxed);
        return nullptr;
    }
    return &slots_[head & mask_];
}

void SpanRing::commit_push() { head_.store(head_.load(std::memory_order_relaxed) + 1, std::memory_order_release); }

const SpanRecord* SpanRing::peek() const {
    const uint64_t tail = tail_.load(std::memory_order_relaxed);
    if (tail == head_.load(std::memory_order_acquire)) return nullptr;
    return &slots_[tail & mask_];
}

void SpanRing::pop() { tail_.store(tail_.load(std::memory_order_relaxed) + 1, std::memory_order_release); }

// ---------------------------------------------------------------------------------------------
// Tracer

Tracer::Tracer(std::string zipkin_url, std::string service_name, int flush_interval_ms, size_t ring_capacity,
               size_t max_batch)
    : flush_interval_ms_(flush_interval_ms), ring_capacity_(ring_capacity), max_batch_(max_batch) {
    endpoint_.service_name = std::move(service_name);
    endpoint_.ipv4 = local_ipv4();
    if (zipkin_url.empty()) return;
    std::string prefix;
    if (!parse_http_url(zipkin_url, &host_, &port_, &prefix)) {
        error_ = "ZIPKIN_SERVER_URL is not an http:// URL: " + zipkin_url;
        return;
    }
    path_ = prefix + "/api/v1/spans";
    enabled_ = true;
}

Tracer::~Tracer() { stop(); }

SpanRing* Tracer::register_producer() {
    std::lock_guard<std::mutex> lock(mu_);
    rings_.emplace_back(new SpanRing(ring_capacity_));
    return rings_.back().get();
}

void Tracer::start() {
    if (!enabled_ || started_) return;
    started_ = true;
    thread_ = std::thread([this] { thread_main(); });
}

void Tracer::stop() {
    if (!started_) return;
    {
        std::lock_guard<std::mutex> lock(mu_);
        stop_ = true;
        stopping_.store(true, std::memory_order_relaxed);
    }
    cv_.notify_all();
    if (thread_.joinable()) thread_.join();
    started_ = false;
    if (sock_ >= 0) {
        close(sock_);
        sock_ = -1;
    }
}

TracerStats Tracer::stats() const {
    TracerStats s;
    s.spans_reported = spans_reported_.load(std::memory_order_relaxed);
    s.spans_dropped = spans_failed_.load(std::memory_order_relaxed);
    s.posts_ok = posts_ok_.load(std::memory_order_relaxed);
    s.posts_failed = posts_failed_.load(std::memory_order_relaxed);
    // register_producer() appends to rings_ from the worker threads as they start; a host that
    // asks for statistics right after starting the service must not walk the vector meanwhile.
    std::lock_guard<std::mutex> lock(mu_);
    for (const auto& r : rings_) s.spans_dropped += r->dropped();
    return s;
}

void Tracer::thread_main() {
    pthread_setname_np(pthread_self(), "hola-zipkin");
    std::unique_lock<std::mutex> lock(mu_);
    while (!stop_) {
        cv_.wait_for(lock, std::chrono::milliseconds(flush_interval_ms_), [this] { return stop_; });
        lock.unlock();
        flush();
        lock.lock();
    }
}

// One reporting cycle: POST what has accumulated, in batches of at most max_batch_ spans, until
// the rings are empty or the next cycle is due. The reference's collector sends one batch per
// cycle because its queue holds no more than one batch (1000 spans); with rings that hold more
// (--zipkin-ring) a live collector is fed as fast as it accepts.
void Tracer::flush() {
    const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(flush_interval_ms_);
    while (flush_batch() == max_batch_ && !stopping_.load(std::memory_order_relaxed) && std::chrono::steady_clock::now() < deadline) {
    }
}

size_t Tracer::flush_batch() {
    std::vector<SpanRing*> rings;
    {
        std::lock_guard<std::mutex> lock(mu_);
        for (auto& r : rings_) rings.push_back(r.get());
    }
    std::string& body = body_;
    body.assign("[");
    size_t count = 0;
    // Take from the rings in turn so that one busy worker cannot starve the others' spans.
    bool any = true;
    while (any && count < max_batch_) {
        any = false;
        for (SpanRing* ring : rings) {
            if (count >= max_batch_) break;
            const SpanRecord* rec = ring->peek();
            if (!rec) continue;
            any = true;
            Span span;
            span.trace_id = rec->trace_id;
            span.id = rec->span_id;
            span.parent_id = rec->parent_id;
            span.has_parent = rec->has_parent;
            span.name = "get";
            span.first_annotation = rec->client ? "cs" : "sr";
            span.last_annotation = rec->client ? "cr" : "ss";
            span.first_us = rec->start_us;
            span.last_us = rec->end_us;
            if (!rec->shared) {
                span.timestamp_us = rec->start_us;
                span.duration_us = rec->end_us - rec->start_us;
            }
            span.tags.push_back(BinaryAnnotation{"http.url", std::string(rec->url, rec->url_len)});
            if (rec->status != 0 && (rec->status < 200 || rec->status > 299))
                span.tags.push_back(BinaryAnnotation{"http.status_code", std::to_string(rec->status)});
            ring->pop();
            if (count > 0) body.push_back(',');
            span_append_json(&body, span, endpoint_);
            ++count;
        }
    }
    if (count == 0) return 0;
    body.push_back(']');
    if (post(body)) {
        spans_reported_.fetch_add(count, std::memory_order_relaxed);
        posts_ok_.fetch_add(1, std::memory_order_relaxed);
    } else {
        spans_failed_.fetch_add(count, std::memory_order_relaxed);
        posts_failed_.fetch_add(1, std::memory_order_relaxed);
        return 0;  // do not hammer a collector that is not there: wait for the next cycle
    }
    // Rings that are full drop at the producer, which is the reference's behaviour when its
    // 1000-span queue is full.
    return count;
}

namespace {

bool wait_fd(int fd, short events, int timeout_ms) {
    pollfd p{fd, events, 0};
    while (true) {
        const int r = poll(&p, 1, timeout_ms);
        if (r > 0) return true;
        if (r == 0) return false;
        if (errno != EINTR) return false;
    }
}

int connect_with_timeout(const std::string& host, uint16_t port, int timeout_ms) {
    IpAddress address;
    if (resolve_address(host, DnsConfig::load_system(), &address, timeout_ms) != DnsStatus::kOk) return -1;
    sockaddr_storage addr;
    const socklen_t addr_len = address.to_sockaddr(port, &addr);
    const int fd = socket(address.family == AF_INET6 ? AF_INET6 : AF_INET, SOCK_STREAM | SOCK_NONBLOCK | SOCK_CLOEXEC, 0);
    if (fd < 0) return -1;
    const int one = 1;
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
    if (connect(fd, reinterpret_cast<sockaddr*>(&addr), addr_len) != 0) {
        if (errno != EINPROGRESS || !wait_fd(fd, POLLOUT, timeout_ms)) {
            close(fd);
            return -1;
        }
        int err = 0;
        socklen_t len = sizeof err;
        if (getsockopt(fd, SOL_SOCKET, SO_ERROR, &err, &len) != 0 || err != 0) {
            close(fd);
            return -1;
        }
    }
    return fd;
}

bool send_all(int fd, const char* data, size_t len, int timeout_ms) {
    size_t sent = 0;
    while (sent < len) {
        const ssize_t n = send(fd, data + sent, len - sent, MSG_NOSIGNAL);
        if (n > 0) {
            sent += static_cast<size_t>(n);
            continue;
        }
        if (n < 0 && errno == EINTR) continue;
        if (n < 0 && (errno == EAGAIN || errno == EWOULDBLOCK)) {
            if (!wait_fd(fd, POLLOUT, timeout_ms)) return false;
            continue;
        }
        return false;
    }
    return true;
}

}  // namespace

bool Tracer::post(const std::string& body) {
    const bool stopping = stopping_.load(std::memory_order_relaxed);
    const int connect_timeout_ms = stopping ? 500 : 10000;
    const int io_timeout_ms = stopping ? 1000 : 60000;
    const std::string host_for_header = host_.find(':') != std::string::npos ? "[" + host_ + "]" : host_;
    std::string head = "POST " + path_ + " HTTP/1.1\r\nHost: " + host_for_header + ":" + std::to_string(port_) +
                       "\r\nContent-Type: application/json\r\nContent-Length: " + std::to_string(body.size()) +
                       "\r\nConnection: keep-alive\r\n\r\n";
    for (int attempt = 0; attempt < 2; ++attempt) {
        const bool reused = sock_ >= 0;
        if (sock_ < 0) sock_ = connect_with_timeout(host_, port_, connect_timeout_ms);
        if (sock_ < 0) return false;
        bool ok = send_all(sock_, head.data(), head.size(), io_timeout_ms) &&
                  send_all(sock_, body.data(), body.size(), io_timeout_ms);
        std::string in;
        ResponseHead rh;
        bool have_head = false;
        while (ok && !have_head) {
            if (!wait_fd(sock_, POLLIN, io_timeout_ms)) {
                ok = false;
                break;
            }
            char buf[4096];
            const ssize_t n = recv(sock_, buf, sizeof buf, 0);
            if (n <= 0) {
                if (n < 0 && (errno == EINTR || errno == EAGAIN)) continue;
                ok = false;
                break;
            }
            in.append(buf, static_cast<size_t>(n));
            const ParseStatus st = parse_response_head(in.data(), in.size(), &rh);
            if (st == ParseStatus::kBad) ok = false;
            if (st == ParseStatus::kComplete) have_head = true;
        }
        if (!ok) {
            close(sock_);
            sock_ = -1;
            if (reused && in.empty()) continue;  // stale keep-alive connection: one more try
            return false;
        }
        // Drain a small, length-delimited body so the connection can be reused; otherwise drop it.
        const size_t have = in.size() - rh.head_len;
        bool reusable = rh.keep_alive && !rh.chunked && rh.content_length >= 0 && rh.content_length <= 65536;
        if (reusable) {
            size_t remaining = static_cast<size_t>(rh.content_length) > have ? static_cast<size_t>(rh.content_length) - have : 0;
            while (remaining > 0) {
                if (!wait_fd(sock_, POLLIN, io_timeout_ms)) {
                    reusable = false;
                    break;
                }
                char buf[4096];
                const ssize_t n = recv(sock_, buf, std::min(sizeof buf, remaining), 0);
                if (n <= 0) {
                    if (n < 0 && (errno == EINTR || errno == EAGAIN)) continue;
                    reusable = false;
                    break;
                }
                remaining -= static_cast<size_t>(n);
            }
        }
        if (!reusable) {
            close(sock_);
            sock_ = -1;
        }
        return rh.status >= 200 && rh.status < 300;
    }
    return false;
}

}  // namespace hola
