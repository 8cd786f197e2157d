// Non-blocking outbound HTTP/1.1 client bound to one event loop, with keep-alive pooling.
//
// Stands in for the Feign + Apache HttpClient stack the reference builds for every chained
// request (/root/reference/src/main/java/com/redhat/developers/msa/hola/HolaResource.java:94-113).
// The reference constructs a new HttpClient per call and never closes it (a connection leak,
// deliberately not reproduced); here each worker keeps a small pool of idle connections per
// destination, and a request that finds its pooled connection dead is retried once on a fresh
// one, as Apache HttpClient's default retry handler does for idempotent requests.
//
// The client has no timers of its own: the caller owns the deadline (the circuit breaker's
// 600 ms) and calls cancel() when it fires.
#pragma once

#include <cstdint>
#include <functional>
#include <memory>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "core/http.h"
#include "net/event_loop.h"
#include "net/resolver.h"

namespace hola {

struct ClientResult {
    enum class Code : uint8_t { kOk, kResolveFailed, kConnectFailed, kIoError, kProtocolError, kTooLarge };
    Code code = Code::kIoError;
    int status = 0;            // valid when code == kOk
    std::string content_type;  // valid when code == kOk
    std::string body;          // valid when code == kOk
    const char* describe() const;
};

class HttpClient {
 public:
    using Callback = std::function<void(ClientResult&&)>;
    using RequestId = uint64_t;

    HttpClient(EventLoop* loop, Resolver* resolver, size_t max_idle_per_host = 64,
               size_t max_response_bytes = 1 << 20);
    ~HttpClient();

    // Sends `request` (a complete serialised request head, no body) to host:port. The callback
    // runs exactly once on the loop thread unless the request is cancelled first. It may run
    // before send() returns (e.g. immediate connect failure).
    RequestId send(const std::string& host, uint16_t port, std::string request, Callback cb);

    // Abandons a request: its connection is closed and its callback will not run.
    void cancel(RequestId id);

    // Closes idle pooled connections (used at shutdown).
    void close_idle();

    size_t idle_connections() const;
    uint64_t connections_opened() const { return connections_opened_; }

 private:
    class Conn;
    friend class Conn;
    struct Pending;

    void dispatch(const std::shared_ptr<Pending>& p);
    void start_connect(const std::shared_ptr<Pending>& p, const IpAddress& address);
    void fail(const std::shared_ptr<Pending>& p, ClientResult::Code code);
    void finish(Conn* conn, ClientResult&& result, bool reusable);
    void retry_or_fail(Conn* conn, ClientResult::Code code);
    void retire(Conn* conn);
    void drop_idle(Conn* conn);
    Conn* take_idle(const std::string& key);

    EventLoop* loop_;
    Resolver* resolver_;
    const size_t max_idle_per_host_;
    const size_t max_response_bytes_;
    RequestId next_id_ = 1;
    uint64_t connections_opened_ = 0;
    std::unordered_map<RequestId, std::shared_ptr<Pending>> pending_;
    std::unordered_map<std::string, std::vector<Conn*>> idle_;
    std::unordered_set<Conn*> conns_;
    std::unordered_set<Conn*> zombies_;  // retired, waiting for the completions of their last io_uring operations
    std::shared_ptr<bool> alive_;  // guards resolver callbacks that outlive the client
};

}  // namespace hola
