// hola-load - HTTP/1.1 load generator for the headline benchmark.
//
// BASELINE.json's metric is "GET /api/hola requests/sec and p50/p99 latency" under "64
// concurrent keep-alive clients for 60 s". The reference ships no load tool and the build box
// has none (no wrk/ab/hey), so this one is part of the project. It is written like the server:
// one epoll loop per thread, connections pinned to a thread, no allocation and no locking per
// request, so that on a shared machine the generator costs no more than the service it measures.
//
// Modes:
//   closed loop (default)  every connection keeps exactly `--pipeline` requests in flight; the
//                          latency of a request runs from the moment it is written to the
//                          moment its last byte is read.
//   open loop (--rate R)   requests are *scheduled* at a fixed rate spread over the connections
//                          and latency runs from the scheduled time, so a stalled server shows
//                          up as growing latency instead of silently fewer samples
//                          (no coordinated omission).
//
// Every response is checked: status code, and optionally the exact body. Time is divided into
// intervals (default 1 s); each interval has its own histogram, the warm-up intervals are
// reported but left out of the summary.
#include <arpa/inet.h>
#include <fcntl.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <pthread.h>
#include <sched.h>
#include <sys/epoll.h>
#include <sys/socket.h>
#include <unistd.h>

#include <atomic>
#include <cerrno>
#include <cmath>
#include <csignal>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "core/histogram.h"
#include "core/http.h"
#include "core/text.h"
#include "net/event_loop.h"  // monotonic_ns

using namespace hola;

namespace {

struct Options {
    std::string host = "127.0.0.1";
    uint16_t port = 8080;
    std::string path = "/api/hola";
    int connections = 64;
    double duration_s = 10;
    double warmup_s = 1;
    double interval_s = 1;
    int threads = 0;
    std::vector<int> cpus;
    double rate = 0;  // 0: closed loop
    int pipeline = 1;
    int expect_status = 200;
    bool check_body = false;
    std::string expect_body;
    std::vector<std::string> headers;
    int timeout_ms = 5000;
    int spin_us = 0;
    bool json = false;
    bool http10 = false;
};

struct Errors {
    uint64_t connect = 0, io = 0, status = 0, body = 0, timeout = 0, protocol = 0;
    uint64_t total() const { return connect + io + status + body + timeout + protocol; }
    void add(const Errors& o) {
        connect += o.connect;
        io += o.io;
        status += o.status;
        body += o.body;
        timeout += o.timeout;
        protocol += o.protocol;
    }
};

struct Interval {
    Histogram hist;
    uint64_t bytes = 0;
};

struct Conn {
    int fd = -1;
    enum State : uint8_t { kIdle, kConnecting, kReady } state = kIdle;
    char rbuf[16384];           // receive buffer; a response head must fit
    size_t rlen = 0;
    size_t sent_off = 0;        // bytes of the pending write already sent
    int to_write = 0;           // requests queued for writing
    int in_flight = 0;          // requests written, response not complete
    int64_t start_ns[16];       // ring of start times, one per in-flight request
    int start_head = 0, start_tail = 0;
    int64_t next_due_ns = 0;    // open loop
    int64_t last_activity_ns = 0;
    // Response being parsed.
    bool head_done = false;
    int64_t body_remaining = 0;
    bool chunked = false;
    bool keep_alive = true;
    ChunkedDecoder decoder;
    std::string body;           // only filled when bodies are checked
    int status = 0;
};

class LoadThread {
 public:
    LoadThread(int id, const Options& opt, sockaddr_in addr, int n_conns, int64_t t0_ns, int n_intervals, double conn_rate)
        : id_(id), opt_(opt), addr_(addr), t0_ns_(t0_ns), conn_period_ns_(conn_rate > 0 ? static_cast<int64_t>(1e9 / conn_rate) : 0) {
        conns_.resize(static_cast<size_t>(n_conns));
        intervals_.resize(static_cast<size_t>(n_intervals));
        interval_ns_ = static_cast<int64_t>(opt.interval_s * 1e9);
        end_ns_ = t0_ns_ + static_cast<int64_t>(n_intervals) * interval_ns_;
        request_ = (opt.http10 ? "GET " + opt.path + " HTTP/1.0\r\nConnection: keep-alive\r\n" : "GET " + opt.path + " HTTP/1.1\r\n");
        request_ += "Host: " + opt.host + ":" + std::to_string(opt.port) + "\r\n";
        for (const auto& h : opt.headers) request_ += h + "\r\n";
        request_ += "\r\n";
        for (int k = 0; k < 16; ++k) request_batch_ += request_;  // 16 = the largest pipeline depth
    }

    void run() {
        if (!opt_.cpus.empty()) {
            cpu_set_t set;
            CPU_ZERO(&set);
            CPU_SET(opt_.cpus[static_cast<size_t>(id_) % opt_.cpus.size()], &set);
            pthread_setaffinity_np(pthread_self(), sizeof set, &set);
        }
        epfd_ = epoll_create1(EPOLL_CLOEXEC);
        int64_t now = monotonic_ns();
        for (size_t i = 0; i < conns_.size(); ++i) {
            // Spread the open-loop schedules evenly inside one period.
            conns_[i].next_due_ns = t0_ns_ + (conn_period_ns_ > 0 ? conn_period_ns_ * static_cast<int64_t>(i) / static_cast<int64_t>(conns_.size()) : 0);
            open_conn(conns_[i], now);
        }
        epoll_event events[256];
        while (true) {
            now = monotonic_ns();
            if (now >= end_ns_) break;
            int64_t wait_ns = std::min<int64_t>(end_ns_ - now, 100000000);
            if (conn_period_ns_ > 0) {
                for (Conn& c : conns_) {
                    if (c.state != Conn::kReady) continue;
                    if (c.next_due_ns <= now) {
                        schedule_open_loop(c, now);
                    }
                    if (c.next_due_ns > now) wait_ns = std::min(wait_ns, c.next_due_ns - now);
                }
            }
            timespec ts{static_cast<time_t>(wait_ns / 1000000000), static_cast<long>(wait_ns % 1000000000)};
            int n = 0;
            if (opt_.spin_us > 0) {
                // Busy-poll while responses are expected; sleep only after spin_us of silence.
                const int64_t spin_until = last_event_ns_ + static_cast<int64_t>(opt_.spin_us) * 1000;
                const int64_t wait_until = now + wait_ns;
                while ((n = epoll_wait(epfd_, events, 256, 0)) == 0) {
                    const int64_t t = monotonic_ns();
                    if (t >= spin_until || t >= wait_until) break;
                    __builtin_ia32_pause();
                }
                if (n == 0) {
                    const int64_t t = monotonic_ns();
                    const int64_t left = wait_until > t ? wait_until - t : 0;
                    timespec ts2{static_cast<time_t>(left / 1000000000), static_cast<long>(left % 1000000000)};
                    n = epoll_pwait2(epfd_, events, 256, &ts2, nullptr);
                }
            } else {
                n = epoll_pwait2(epfd_, events, 256, &ts, nullptr);
            }
            if (n < 0 && errno != EINTR) break;
            if (n > 0) last_event_ns_ = monotonic_ns();
            now = monotonic_ns();
            for (int i = 0; i < n; ++i) {
                Conn& c = *static_cast<Conn*>(events[i].data.ptr);
                if (c.fd < 0) continue;
                handle(c, events[i].events, now);
            }
            if (now - last_sweep_ns_ > 100000000) {
                last_sweep_ns_ = now;
                sweep(now);
            }
        }
        for (Conn& c : conns_) {
            if (c.fd >= 0) close(c.fd);
        }
        close(epfd_);
    }

    const std::vector<Interval>& intervals() const { return intervals_; }
    const Errors& errors() const { return errors_; }
    uint64_t reconnects() const { return reconnects_; }

 private:
    void open_conn(Conn& c, int64_t now) {
        c.fd = socket(AF_INET, SOCK_STREAM | SOCK_NONBLOCK | SOCK_CLOEXEC, 0);
        if (c.fd < 0) {
            ++errors_.connect;
            return;
        }
        const int one = 1;
        setsockopt(c.fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
        c.rlen = 0;
        c.sent_off = 0;
        c.to_write = c.in_flight = 0;
        c.start_head = c.start_tail = 0;
        c.head_done = false;
        c.last_activity_ns = now;
        c.state = Conn::kConnecting;
        const int rc = connect(c.fd, reinterpret_cast<const sockaddr*>(&addr_), sizeof addr_);
        if (rc != 0 && errno != EINPROGRESS) {
            ++errors_.connect;
            close(c.fd);
            c.fd = -1;
            c.state = Conn::kIdle;
            return;
        }
        epoll_event ev{};
        ev.events = EPOLLOUT;
        ev.data.ptr = &c;
        epoll_ctl(epfd_, EPOLL_CTL_ADD, c.fd, &ev);
    }

    void reopen(Conn& c, int64_t now) {
        if (c.fd >= 0) {
            epoll_ctl(epfd_, EPOLL_CTL_DEL, c.fd, nullptr);
            close(c.fd);
            c.fd = -1;
        }
        ++reconnects_;
        open_conn(c, now);
    }

    void set_events(Conn& c, uint32_t events) {
        epoll_event ev{};
        ev.events = events;
        ev.data.ptr = &c;
        epoll_ctl(epfd_, EPOLL_CTL_MOD, c.fd, &ev);
    }

    void enqueue(Conn& c, int64_t start_ns) {
        c.start_ns[c.start_head] = start_ns;
        c.start_head = (c.start_head + 1) & 15;
        ++c.to_write;
    }

    void schedule_open_loop(Conn& c, int64_t now) {
        // Queue every request whose scheduled time has passed (bounded by the pipeline depth;
        // what does not fit stays due and keeps accruing latency).
        while (c.next_due_ns <= now && c.in_flight + c.to_write < std::max(1, opt_.pipeline)) {
            enqueue(c, c.next_due_ns);
            c.next_due_ns += conn_period_ns_;
        }
        flush_writes(c, now);
    }

    void fill_closed_loop(Conn& c, int64_t now) {
        while (c.in_flight + c.to_write < opt_.pipeline) enqueue(c, now);
        flush_writes(c, now);
    }

    void flush_writes(Conn& c, int64_t now) {
        // All queued requests go out in one send(): request_batch_ holds 16 copies back to back,
        // and sent_off is the position inside the first of the still-queued copies.
        while (c.to_write > 0) {
            const size_t len = static_cast<size_t>(c.to_write) * request_.size() - c.sent_off;
            const ssize_t n = send(c.fd, request_batch_.data() + c.sent_off, len, MSG_NOSIGNAL);
            if (n > 0) {
                const size_t progressed = c.sent_off + static_cast<size_t>(n);
                const int completed = static_cast<int>(progressed / request_.size());
                c.sent_off = progressed % request_.size();
                c.to_write -= completed;
                c.in_flight += completed;
                continue;
            }
            if (n < 0 && errno == EINTR) continue;
            if (n < 0 && (errno == EAGAIN || errno == EWOULDBLOCK)) {
                set_events(c, EPOLLIN | EPOLLOUT);
                return;
            }
            ++errors_.io;
            reopen(c, now);
            return;
        }
    }

    void handle(Conn& c, uint32_t events, int64_t now) {
        c.last_activity_ns = now;
        if (c.state == Conn::kConnecting) {
            int err = 0;
            socklen_t len = sizeof err;
            if ((events & (EPOLLERR | EPOLLHUP)) || getsockopt(c.fd, SOL_SOCKET, SO_ERROR, &err, &len) != 0 || err != 0) {
                ++errors_.connect;
                epoll_ctl(epfd_, EPOLL_CTL_DEL, c.fd, nullptr);
                close(c.fd);
                c.fd = -1;
                c.state = Conn::kIdle;
                c.last_activity_ns = now;  // sweep() retries later
                return;
            }
            c.state = Conn::kReady;
            set_events(c, EPOLLIN);
            if (conn_period_ns_ == 0) fill_closed_loop(c, now);
            return;
        }
        if (events & EPOLLOUT) {
            set_events(c, EPOLLIN);
            flush_writes(c, now);
            if (c.fd < 0 || c.state != Conn::kReady) return;
        }
        if (events & (EPOLLIN | EPOLLHUP | EPOLLERR)) read_some(c, now);
    }

    void read_some(Conn& c, int64_t now) {
        while (true) {
            const size_t space = sizeof c.rbuf - c.rlen;
            const ssize_t n = recv(c.fd, c.rbuf + c.rlen, space, 0);
            if (n > 0) {
                c.rlen += static_cast<size_t>(n);
                if (!parse(c, now)) return;
                if (static_cast<size_t>(n) < space) return;
                continue;
            }
            if (n == 0) {
                if (c.in_flight > 0 || c.to_write > 0) ++errors_.io;
                reopen(c, now);
                return;
            }
            if (errno == EINTR) continue;
            if (errno == EAGAIN || errno == EWOULDBLOCK) return;
            ++errors_.io;
            reopen(c, now);
            return;
        }
    }

    // Runs the response parser over the connection's receive buffer. Returns false if the
    // connection was reopened (its buffers are then already reset).
    bool parse(Conn& c, int64_t now) {
        size_t pos = 0;
        while (pos < c.rlen) {
            if (!c.head_done) {
                ResponseHead head;
                const ParseStatus st = parse_response_head(c.rbuf + pos, c.rlen - pos, &head);
                if (st == ParseStatus::kIncomplete) {
                    if (pos == 0 && c.rlen == sizeof c.rbuf) {
                        ++errors_.protocol;
                        reopen(c, now);
                        return false;
                    }
                    break;
                }
                if (st == ParseStatus::kBad || c.in_flight == 0) {
                    ++errors_.protocol;
                    reopen(c, now);
                    return false;
                }
                c.head_done = true;
                c.status = head.status;
                c.keep_alive = head.keep_alive;
                c.chunked = head.chunked;
                c.body.clear();
                c.decoder.reset();
                c.body_remaining = head.chunked ? 0 : (head.content_length > 0 ? head.content_length : 0);
                pos += head.head_len;
                interval_bytes(now) += head.head_len;
                if (!c.chunked && c.body_remaining == 0) {
                    if (!complete(c, now)) return false;
                }
                continue;
            }
            const char* data = c.rbuf + pos;
            const size_t len = c.rlen - pos;
            size_t take = 0;
            bool done = false;
            if (c.chunked) {
                take = c.decoder.feed(data, len, opt_.check_body ? &c.body : nullptr);
                if (c.decoder.failed()) {
                    ++errors_.protocol;
                    reopen(c, now);
                    return false;
                }
                done = c.decoder.done();
            } else {
                take = static_cast<size_t>(std::min<int64_t>(c.body_remaining, static_cast<int64_t>(len)));
                if (opt_.check_body) c.body.append(data, take);
                c.body_remaining -= static_cast<int64_t>(take);
                done = c.body_remaining == 0;
            }
            pos += take;
            interval_bytes(now) += take;
            if (done && !complete(c, now)) return false;
        }
        if (pos >= c.rlen) {
            c.rlen = 0;
        } else if (pos > 0) {
            std::memmove(c.rbuf, c.rbuf + pos, c.rlen - pos);
            c.rlen -= pos;
        }
        // Refill the pipeline / send what has become due. The new request's clock starts at the
        // moment it is written, not at the (earlier) time this batch of events was picked up.
        if (conn_period_ns_ == 0) {
            fill_closed_loop(c, monotonic_ns());
        } else if (c.next_due_ns <= now) {
            schedule_open_loop(c, now);
        }
        return c.fd >= 0 && c.state == Conn::kReady;
    }

    uint64_t& interval_bytes(int64_t now) { return intervals_[interval_index(now)].bytes; }

    size_t interval_index(int64_t now) const {
        int64_t idx = (now - t0_ns_) / interval_ns_;
        if (idx < 0) idx = 0;
        if (idx >= static_cast<int64_t>(intervals_.size())) idx = static_cast<int64_t>(intervals_.size()) - 1;
        return static_cast<size_t>(idx);
    }

    bool complete(Conn& c, int64_t now) {
        c.head_done = false;
        const int64_t started = c.start_ns[c.start_tail];
        c.start_tail = (c.start_tail + 1) & 15;
        --c.in_flight;
        bool ok = true;
        if (c.status != opt_.expect_status) {
            ++errors_.status;
            ok = false;
        } else if (opt_.check_body && c.body != opt_.expect_body) {
            ++errors_.body;
            ok = false;
        }
        if (ok) {
            // One clock read per response: the batch timestamp is too coarse for microsecond
            // latencies when several responses arrive in one epoll round.
            const int64_t t = monotonic_ns();
            if (t >= t0_ns_) intervals_[interval_index(t)].hist.record(static_cast<uint64_t>(t > started ? t - started : 0));
        }
        if (!c.keep_alive) {
            reopen(c, now);
            return false;
        }
        // The replacement request is written by parse() once every response in the receive
        // buffer has been handled, so that a pipelining client refills with one send().
        return true;
    }

    void sweep(int64_t now) {
        const int64_t limit = static_cast<int64_t>(opt_.timeout_ms) * 1000000;
        for (Conn& c : conns_) {
            if (c.state == Conn::kIdle && c.fd < 0) {
                if (now - c.last_activity_ns > 100000000) open_conn(c, now);  // retry a failed connect
                continue;
            }
            if ((c.in_flight > 0 || c.state == Conn::kConnecting) && now - c.last_activity_ns > limit) {
                ++errors_.timeout;
                reopen(c, now);
            }
        }
    }

    const int id_;
    const Options& opt_;
    const sockaddr_in addr_;
    const int64_t t0_ns_;
    const int64_t conn_period_ns_;
    int64_t interval_ns_ = 0;
    int64_t end_ns_ = 0;
    int64_t last_sweep_ns_ = 0;
    int64_t last_event_ns_ = 0;
    int epfd_ = -1;
    std::string request_;
    std::string request_batch_;
    std::vector<Conn> conns_;
    std::vector<Interval> intervals_;
    Errors errors_;
    uint64_t reconnects_ = 0;
};

void usage() {
    std::fprintf(stderr,
                 "usage: hola-load [options] [http://host:port/path]\n"
                 "  -c, --connections N    concurrent keep-alive connections (64)\n"
                 "  -d, --duration S       measured seconds (10)\n"
                 "  -w, --warmup S         warm-up seconds before measuring (1)\n"
                 "  -i, --interval S       reporting interval (1)\n"
                 "  -t, --threads N        generator threads (min(connections, CPUs))\n"
                 "      --cpus LIST        pin thread i to LIST[i] (e.g. 4-7)\n"
                 "      --rate R           open loop at R requests/s in total (default: closed loop)\n"
                 "      --pipeline N       requests in flight per connection (1)\n"
                 "      --expect-status N  (200)\n"
                 "      --expect-body S | --expect-body-file F   verify every body byte-for-byte\n"
                 "  -H, --header 'K: V'    extra request header (repeatable)\n"
                 "      --http10           send HTTP/1.0 keep-alive requests\n"
                 "      --timeout-ms N     per-response timeout (5000)\n"
                 "      --spin-us N        busy-poll for N us after the last event before sleeping (0)\n"
                 "      --json             machine-readable summary on stdout\n");
}

bool parse_url(const std::string& url, Options* o) {
    static const std::string kScheme = "http://";
    if (url.compare(0, kScheme.size(), kScheme) != 0) return false;
    const size_t slash = url.find('/', kScheme.size());
    const std::string authority = url.substr(kScheme.size(), slash == std::string::npos ? std::string::npos : slash - kScheme.size());
    o->path = slash == std::string::npos ? "/" : url.substr(slash);
    const size_t colon = authority.rfind(':');
    if (colon == std::string::npos) {
        o->host = authority;
        o->port = 80;
    } else {
        o->host = authority.substr(0, colon);
        o->port = static_cast<uint16_t>(std::atoi(authority.substr(colon + 1).c_str()));
    }
    return !o->host.empty() && o->port != 0;
}

void json_string(std::string* out, const std::string& s) {
    out->push_back('"');
    for (char c : s) {
        if (c == '"' || c == '\\') {
            out->push_back('\\');
            out->push_back(c);
        } else if (static_cast<unsigned char>(c) < 0x20) {
            char buf[8];
            std::snprintf(buf, sizeof buf, "\\u%04x", c);
            out->append(buf);
        } else {
            out->push_back(c);
        }
    }
    out->push_back('"');
}

double us(uint64_t ns) { return static_cast<double>(ns) / 1000.0; }

}  // namespace

int main(int argc, char** argv) {
    Options opt;
    auto need = [&](int& i) -> const char* {
        if (i + 1 >= argc) {
            usage();
            std::exit(2);
        }
        return argv[++i];
    };
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        if (a == "-c" || a == "--connections") opt.connections = std::atoi(need(i));
        else if (a == "-d" || a == "--duration") opt.duration_s = std::atof(need(i));
        else if (a == "-w" || a == "--warmup") opt.warmup_s = std::atof(need(i));
        else if (a == "-i" || a == "--interval") opt.interval_s = std::atof(need(i));
        else if (a == "-t" || a == "--threads") opt.threads = std::atoi(need(i));
        else if (a == "--cpus") {
            if (!parse_cpu_list(need(i), &opt.cpus)) { std::fprintf(stderr, "hola-load: bad --cpus list\n"); return 2; }
        }
        else if (a == "--rate") opt.rate = std::atof(need(i));
        else if (a == "--pipeline") opt.pipeline = std::atoi(need(i));
        else if (a == "--expect-status") opt.expect_status = std::atoi(need(i));
        else if (a == "--expect-body") { opt.expect_body = need(i); opt.check_body = true; }
        else if (a == "--expect-body-file") {
            FILE* f = std::fopen(need(i), "rb");
            if (!f) { std::perror("hola-load: --expect-body-file"); return 2; }
            char buf[4096]; size_t n;
            while ((n = std::fread(buf, 1, sizeof buf, f)) > 0) opt.expect_body.append(buf, n);
            std::fclose(f);
            opt.check_body = true;
        }
        else if (a == "-H" || a == "--header") opt.headers.push_back(need(i));
        else if (a == "--http10") opt.http10 = true;
        else if (a == "--timeout-ms") opt.timeout_ms = std::atoi(need(i));
        else if (a == "--spin-us") opt.spin_us = std::atoi(need(i));
        else if (a == "--json") opt.json = true;
        else if (a == "-h" || a == "--help") { usage(); return 0; }
        else if (!a.empty() && a[0] != '-') { if (!parse_url(a, &opt)) { std::fprintf(stderr, "hola-load: bad URL %s\n", a.c_str()); return 2; } }
        else { std::fprintf(stderr, "hola-load: unknown option %s\n", a.c_str()); usage(); return 2; }
    }
    if (opt.connections < 1 || opt.pipeline < 1 || opt.pipeline > 16 || opt.interval_s <= 0 || opt.duration_s <= 0) {
        std::fprintf(stderr, "hola-load: connections >= 1, 1 <= pipeline <= 16, interval > 0, duration > 0\n");
        return 2;
    }
    if (opt.threads <= 0) {
        cpu_set_t set;
        CPU_ZERO(&set);
        int n = sched_getaffinity(0, sizeof set, &set) == 0 ? CPU_COUNT(&set) : 1;
        if (!opt.cpus.empty()) n = static_cast<int>(opt.cpus.size());
        opt.threads = std::max(1, std::min(n, opt.connections));
    }
    opt.threads = std::min(opt.threads, opt.connections);

    addrinfo hints{};
    hints.ai_family = AF_INET;
    hints.ai_socktype = SOCK_STREAM;
    addrinfo* res = nullptr;
    if (getaddrinfo(opt.host.c_str(), nullptr, &hints, &res) != 0 || !res) {
        std::fprintf(stderr, "hola-load: cannot resolve %s\n", opt.host.c_str());
        return 1;
    }
    sockaddr_in addr = *reinterpret_cast<sockaddr_in*>(res->ai_addr);
    addr.sin_port = htons(opt.port);
    freeaddrinfo(res);

    signal(SIGPIPE, SIG_IGN);
    const int warm_intervals = static_cast<int>(std::ceil(opt.warmup_s / opt.interval_s - 1e-9));
    const int run_intervals = std::max(1, static_cast<int>(std::ceil(opt.duration_s / opt.interval_s - 1e-9)));
    const int n_intervals = warm_intervals + run_intervals;
    const int64_t t0 = monotonic_ns() + 20000000;  // connections are set up in the first 20 ms

    std::vector<std::unique_ptr<LoadThread>> workers;
    std::vector<std::thread> threads;
    for (int t = 0; t < opt.threads; ++t) {
        const int n = opt.connections / opt.threads + (t < opt.connections % opt.threads ? 1 : 0);
        const double conn_rate = opt.rate > 0 ? opt.rate / opt.connections : 0;
        workers.emplace_back(new LoadThread(t, opt, addr, n, t0, n_intervals, conn_rate));
    }
    for (auto& w : workers) {
        LoadThread* p = w.get();
        threads.emplace_back([p] { p->run(); });
    }
    for (auto& t : threads) t.join();

    // Merge.
    std::vector<Interval> merged(static_cast<size_t>(n_intervals));
    Errors errors;
    uint64_t reconnects = 0;
    for (auto& w : workers) {
        for (int i = 0; i < n_intervals; ++i) {
            merged[static_cast<size_t>(i)].hist.merge(w->intervals()[static_cast<size_t>(i)].hist);
            merged[static_cast<size_t>(i)].bytes += w->intervals()[static_cast<size_t>(i)].bytes;
        }
        errors.add(w->errors());
        reconnects += w->reconnects();
    }
    Histogram total;
    uint64_t bytes = 0;
    for (int i = warm_intervals; i < n_intervals; ++i) {
        total.merge(merged[static_cast<size_t>(i)].hist);
        bytes += merged[static_cast<size_t>(i)].bytes;
    }
    const double measured_s = run_intervals * opt.interval_s;
    const double rps = static_cast<double>(total.count()) / measured_s;

    if (opt.json) {
        std::string o = "{";
        char buf[512];
        o += "\"url\":";
        json_string(&o, "http://" + opt.host + ":" + std::to_string(opt.port) + opt.path);
        std::snprintf(buf, sizeof buf,
                      ",\"mode\":\"%s\",\"connections\":%d,\"threads\":%d,\"pipeline\":%d,\"warmup_s\":%.3f,\"duration_s\":%.3f,"
                      "\"interval_s\":%.3f,\"requests\":%llu,\"rps\":%.1f,\"bytes\":%llu,\"reconnects\":%llu,",
                      opt.rate > 0 ? "open" : "closed", opt.connections, opt.threads, opt.pipeline, warm_intervals * opt.interval_s,
                      measured_s, opt.interval_s, static_cast<unsigned long long>(total.count()), rps,
                      static_cast<unsigned long long>(bytes), static_cast<unsigned long long>(reconnects));
        o += buf;
        if (opt.rate > 0) {
            std::snprintf(buf, sizeof buf, "\"target_rate\":%.1f,", opt.rate);
            o += buf;
        }
        std::snprintf(buf, sizeof buf,
                      "\"errors\":{\"total\":%llu,\"connect\":%llu,\"io\":%llu,\"status\":%llu,\"body\":%llu,\"timeout\":%llu,\"protocol\":%llu},",
                      static_cast<unsigned long long>(errors.total()), static_cast<unsigned long long>(errors.connect),
                      static_cast<unsigned long long>(errors.io), static_cast<unsigned long long>(errors.status),
                      static_cast<unsigned long long>(errors.body), static_cast<unsigned long long>(errors.timeout),
                      static_cast<unsigned long long>(errors.protocol));
        o += buf;
        std::snprintf(buf, sizeof buf,
                      "\"latency_us\":{\"min\":%.1f,\"mean\":%.1f,\"p50\":%.1f,\"p90\":%.1f,\"p99\":%.1f,\"p999\":%.1f,\"max\":%.1f},",
                      us(total.min()), total.mean() / 1000.0, us(total.percentile(50)), us(total.percentile(90)),
                      us(total.percentile(99)), us(total.percentile(99.9)), us(total.max()));
        o += buf;
        o += "\"intervals\":[";
        for (int i = 0; i < n_intervals; ++i) {
            const Histogram& h = merged[static_cast<size_t>(i)].hist;
            std::snprintf(buf, sizeof buf, "%s{\"t\":%.3f,\"warmup\":%s,\"requests\":%llu,\"rps\":%.1f,\"p50_us\":%.1f,\"p99_us\":%.1f}",
                          i ? "," : "", (i + 1) * opt.interval_s, i < warm_intervals ? "true" : "false",
                          static_cast<unsigned long long>(h.count()), static_cast<double>(h.count()) / opt.interval_s,
                          us(h.percentile(50)), us(h.percentile(99)));
            o += buf;
        }
        o += "]}";
        std::puts(o.c_str());
    } else {
        std::printf("%s loop, %d connections, %d threads, %.1f s measured after %.1f s warm-up\n", opt.rate > 0 ? "open" : "closed",
                    opt.connections, opt.threads, measured_s, warm_intervals * opt.interval_s);
        std::printf("  requests  %llu  (%.0f req/s, %.2f MB/s)\n", static_cast<unsigned long long>(total.count()), rps,
                    static_cast<double>(bytes) / measured_s / 1e6);
        std::printf("  latency   min %.1f us  mean %.1f us  p50 %.1f us  p90 %.1f us  p99 %.1f us  p99.9 %.1f us  max %.1f us\n",
                    us(total.min()), total.mean() / 1000.0, us(total.percentile(50)), us(total.percentile(90)),
                    us(total.percentile(99)), us(total.percentile(99.9)), us(total.max()));
        std::printf("  errors    %llu (connect %llu, io %llu, status %llu, body %llu, timeout %llu, protocol %llu), reconnects %llu\n",
                    static_cast<unsigned long long>(errors.total()), static_cast<unsigned long long>(errors.connect),
                    static_cast<unsigned long long>(errors.io), static_cast<unsigned long long>(errors.status),
                    static_cast<unsigned long long>(errors.body), static_cast<unsigned long long>(errors.timeout),
                    static_cast<unsigned long long>(errors.protocol), static_cast<unsigned long long>(reconnects));
    }
    return errors.total() == 0 ? 0 : 3;
}
