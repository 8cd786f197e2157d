// Host name resolution that never blocks an event loop.
//
// The reference resolves its downstream service by name ("aloha", a Kubernetes service name:
// /root/reference/src/main/java/com/redhat/developers/msa/hola/HolaResource.java:95,103) through
// the JVM, inside the Hystrix worker thread, where a slow lookup is cut off by the 600 ms command
// timeout. Here a lookup runs on a small helper thread and its result is posted back to the
// asking loop; the caller's own timer provides the cut-off. Results are cached the way the JVM
// caches them: 30 s for a success, 10 s for a failure.
//
// Numeric IPv4/IPv6 literals and "localhost" are answered inline without touching the helper thread.
#pragma once

#include <condition_variable>
#include <cstdint>
#include <deque>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "net/dns.h"
#include "net/event_loop.h"

namespace hola {

class Resolver {
 public:
    // ok=false means the name did not resolve.
    using Callback = std::function<void(bool ok, const IpAddress& address)>;

    explicit Resolver(int threads = 2, int positive_ttl_s = 30, int negative_ttl_s = 10);
    ~Resolver();
    // Stops answering. Call before the event loops that asked are destroyed. Idempotent.
    void shutdown();

    // The callback runs on `loop`'s thread: immediately (before resolve() returns) on a cache hit
    // or literal, otherwise later via loop->post(). Must be called on `loop`'s thread.
    void resolve(const std::string& host, EventLoop* loop, Callback cb);

    // Pins a name to an address, bypassing hosts file and DNS (--resolve NAME=ADDR).
    void set_static(const std::string& host, const IpAddress& address);

 private:
    struct Entry {
        bool ok = false;
        IpAddress address;
        int64_t expires_ns = 0;
        bool permanent = false;
    };
    struct Waiter {
        EventLoop* loop;
        Callback cb;
    };
    // Shared with the helper threads, which are detached: a lookup stuck inside the system
    // resolver must not hold up process shutdown.
    struct State {
        int64_t positive_ttl_ns = 0;
        int64_t negative_ttl_ns = 0;
        std::mutex mu;
        std::condition_variable cv;
        bool stop = false;
        std::unordered_map<std::string, Entry> cache;
        std::unordered_map<std::string, std::vector<Waiter>> inflight;  // name -> who is waiting
        std::deque<std::string> queue;
    };
    static void thread_main(std::shared_ptr<State> state);

    std::shared_ptr<State> state_;
};

}  // namespace hola
