#include "net/resolver.h"

#include <arpa/inet.h>
#include <netinet/in.h>
#include <sys/socket.h>

#include <cstring>

#include "net/dns.h"

namespace hola {

Resolver::Resolver(int threads, int positive_ttl_s, int negative_ttl_s) : state_(std::make_shared<State>()) {
    state_->positive_ttl_ns = static_cast<int64_t>(positive_ttl_s) * 1000000000ll;
    state_->negative_ttl_ns = static_cast<int64_t>(negative_ttl_s) * 1000000000ll;
    for (int i = 0; i < (threads < 1 ? 1 : threads); ++i) {
        std::thread(thread_main, state_).detach();
    }
}

Resolver::~Resolver() { shutdown(); }

void Resolver::shutdown() {
    {
        std::lock_guard<std::mutex> lock(state_->mu);
        state_->stop = true;
        state_->inflight.clear();  // nobody is left to hear the answers
    }
    state_->cv.notify_all();
}

void Resolver::set_static(const std::string& host, const IpAddress& address) {
    std::lock_guard<std::mutex> lock(state_->mu);
    Entry e;
    e.ok = true;
    e.address = address;
    e.permanent = true;
    state_->cache[host] = e;
}

void Resolver::resolve(const std::string& host, EventLoop* loop, Callback cb) {
    const IpAddress literal = IpAddress::parse(host);
    if (literal.family != 0) {
        cb(true, literal);
        return;
    }
    if (host == "localhost") {
        cb(true, IpAddress::v4(htonl(INADDR_LOOPBACK)));
        return;
    }
    const int64_t now = monotonic_ns();
    {
        std::unique_lock<std::mutex> lock(state_->mu);
        auto it = state_->cache.find(host);
        if (it != state_->cache.end() && (it->second.permanent || it->second.expires_ns > now)) {
            const Entry e = it->second;
            lock.unlock();
            cb(e.ok, e.address);
            return;
        }
        auto& waiters = state_->inflight[host];
        const bool first = waiters.empty();
        waiters.push_back(Waiter{loop, std::move(cb)});
        if (first) {
            state_->queue.push_back(host);
            state_->cv.notify_one();
        }
    }
}

void Resolver::thread_main(std::shared_ptr<State> st) {
    while (true) {
        std::string host;
        {
            std::unique_lock<std::mutex> lock(st->mu);
            st->cv.wait(lock, [&st] { return st->stop || !st->queue.empty(); });
            if (st->stop) return;
            host = std::move(st->queue.front());
            st->queue.pop_front();
        }
        // The system configuration is re-read per lookup (it is tiny, lookups are rare, and a
        // pod's resolv.conf can be rewritten under a running process).
        Entry e;
        e.ok = resolve_address(host, DnsConfig::load_system(), &e.address) == DnsStatus::kOk;
        e.expires_ns = monotonic_ns() + (e.ok ? st->positive_ttl_ns : st->negative_ttl_ns);
        // Answers are posted under the lock so that shutdown() (which takes it) cannot return
        // while a post to a soon-to-be-destroyed loop is still in progress.
        std::lock_guard<std::mutex> lock(st->mu);
        if (st->stop) return;
        auto cached = st->cache.find(host);
        if (cached == st->cache.end() || !cached->second.permanent) st->cache[host] = e;
        auto it = st->inflight.find(host);
        if (it == st->inflight.end()) continue;
        std::vector<Waiter> waiters = std::move(it->second);
        st->inflight.erase(it);
        for (auto& w : waiters) {
            const bool ok = e.ok;
            const IpAddress address = e.address;
            w.loop->post([cb = std::move(w.cb), ok, address] { cb(ok, address); });
        }
    }
}

}  // namespace hola
