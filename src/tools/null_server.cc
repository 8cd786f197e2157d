// hola-null - the ceiling probe: what this machine's TCP stack allows a server with hola's
// threading to do when the server itself does nothing.
//
// Same shape as the real service (src/net/server.cc): one pinned thread per CPU, each with its
// own SO_REUSEPORT listener and its own readiness loop, connections dealt evenly over the threads
// and staying on one thread for life. But no HTTP: whatever arrives is answered with one canned
// response per blank line seen ("\r\n\r\n" ends a request head) - no parsing, no routing, no
// tracing, no date. The canned bytes are what `hola` sends for GET /api/hola, so the load
// generator verifies them as usual and the wire bytes per request are identical.
//
//   hola-null [--port N] [--threads N] [--cpus LIST] [--io-backend epoll|uring]
//
// Run the identical hola-load configuration against `hola` and against `hola-null`: the ratio is
// the fraction of the box's ceiling the service reaches. With
// --io-backend uring the same probe uses io_uring (multishot accept, multishot receive into
// provided buffers, batched sends, one system call per loop iteration) - the upper bound of what
// an io_uring backend could give the service. The reference gets its I/O layer from
// Undertow/XNIO (/root/reference/pom.xml:122-129), which is epoll + read + write per request.
#include <arpa/inet.h>
#include <fcntl.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <pthread.h>
#include <sched.h>
#include <signal.h>
#include <sys/epoll.h>
#include <sys/socket.h>
#include <unistd.h>

#include <atomic>
#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "core/text.h"
#include "net/uring.h"

namespace {

using hola::Uring;

const char kResponse[] =
    "HTTP/1.1 200 OK\r\nConnection: keep-alive\r\nAccess-Control-Allow-Origin: *\r\nContent-Type: text/plain\r\n"
    "Content-Length: 15\r\nDate: Sat, 07 Nov 2026 00:00:00 GMT\r\n\r\nHola de unknown";
constexpr size_t kResponseLen = sizeof kResponse - 1;
constexpr int kMaxBatch = 16;  // canned responses kept back to back for pipelined requests

struct Options {
    uint16_t port = 8080;
    int threads = 1;
    std::vector<int> cpus;
    bool uring = false;
};

struct ThreadCtx {
    int id = 0;
    int listen_fd = -1;
    int pipe_rd = -1, pipe_wr = -1;  // descriptors dealt to this thread by other acceptors
    std::atomic<uint64_t> requests{0};
    std::atomic<uint64_t> connections{0};
};

std::atomic<bool> g_stop{false};
std::atomic<uint32_t> g_next{0};
std::vector<ThreadCtx*> g_threads;
std::string g_batch;  // kMaxBatch responses

int open_listener(uint16_t port) {
    const int fd = socket(AF_INET, SOCK_STREAM | SOCK_NONBLOCK | SOCK_CLOEXEC, 0);
    const int one = 1;
    setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
    setsockopt(fd, SOL_SOCKET, SO_REUSEPORT, &one, sizeof one);
    sockaddr_in a{};
    a.sin_family = AF_INET;
    a.sin_port = htons(port);
    a.sin_addr.s_addr = htonl(INADDR_LOOPBACK);
    if (bind(fd, reinterpret_cast<sockaddr*>(&a), sizeof a) != 0 || listen(fd, 1024) != 0) {
        std::perror("hola-null: bind/listen");
        std::exit(1);
    }
    return fd;
}

size_t count_heads(const char* data, size_t len) {
    size_t n = 0;
    const char* p = data;
    const char* end = data + len;
    while (p + 4 <= end) {
        p = static_cast<const char*>(memmem(p, static_cast<size_t>(end - p), "\r\n\r\n", 4));
        if (!p) break;
        ++n;
        p += 4;
    }
    return n;
}

// A new connection goes to the threads in turn, whoever accepted it (round-robin: the even
// spread the service's balanced placement produces for long-lived connections).
void deal(ThreadCtx& self, int fd, void (*adopt)(ThreadCtx&, int, void*), void* arg) {
    const int one = 1;
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
    ThreadCtx* target = g_threads[g_next.fetch_add(1, std::memory_order_relaxed) % g_threads.size()];
    if (target == &self) {
        adopt(self, fd, arg);
    } else if (write(target->pipe_wr, &fd, sizeof fd) != static_cast<ssize_t>(sizeof fd)) {
        close(fd);
    }
}

// ---- epoll backend: the service's own system-call pattern -------------------------------------

void epoll_adopt(ThreadCtx& t, int fd, void* arg) {
    const int ep = *static_cast<int*>(arg);
    epoll_event ev{};
    ev.events = EPOLLIN | EPOLLRDHUP;
    ev.data.fd = fd;
    if (epoll_ctl(ep, EPOLL_CTL_ADD, fd, &ev) != 0) {
        close(fd);
        return;
    }
    t.connections.fetch_add(1, std::memory_order_relaxed);
}

void run_epoll(ThreadCtx& t) {
    int ep = epoll_create1(EPOLL_CLOEXEC);
    epoll_event ev{};
    ev.events = EPOLLIN;
    ev.data.fd = t.listen_fd;
    epoll_ctl(ep, EPOLL_CTL_ADD, t.listen_fd, &ev);
    ev.data.fd = t.pipe_rd;
    epoll_ctl(ep, EPOLL_CTL_ADD, t.pipe_rd, &ev);
    char buf[4096];
    epoll_event events[256];
    uint64_t served = 0;
    while (!g_stop.load(std::memory_order_relaxed)) {
        const int n = epoll_wait(ep, events, 256, 200);
        for (int i = 0; i < n; ++i) {
            const int fd = events[i].data.fd;
            if (fd == t.listen_fd) {
                while (true) {
                    const int c = accept4(t.listen_fd, nullptr, nullptr, SOCK_NONBLOCK | SOCK_CLOEXEC);
                    if (c < 0) break;
                    deal(t, c, epoll_adopt, &ep);
                }
            } else if (fd == t.pipe_rd) {
                int c;
                while (read(t.pipe_rd, &c, sizeof c) == static_cast<ssize_t>(sizeof c)) epoll_adopt(t, c, &ep);
            } else {
                const ssize_t r = recv(fd, buf, sizeof buf, 0);
                if (r <= 0) {
                    if (r < 0 && (errno == EAGAIN || errno == EINTR)) continue;
                    close(fd);  // also removes it from the epoll set
                    continue;
                }
                size_t k = count_heads(buf, static_cast<size_t>(r));
                served += k;
                while (k > 0) {
                    const size_t chunk = k > kMaxBatch ? kMaxBatch : k;
                    if (send(fd, g_batch.data(), chunk * kResponseLen, MSG_NOSIGNAL) < 0) break;
                    k -= chunk;
                }
            }
        }
        t.requests.store(served, std::memory_order_relaxed);
    }
    close(ep);
}

// ---- io_uring backend --------------------------------------------------------------------------

enum : uint64_t { kAccept = 1, kPipe = 2, kRecv = 3, kSend = 4 };
inline uint64_t tag(uint64_t kind, int fd) { return (kind << 32) | static_cast<uint32_t>(fd); }

struct UringLoop {
    Uring ring;
    int pipe_value = 0;
};

io_uring_sqe* must_sqe(Uring& ring) {
    io_uring_sqe* sqe = ring.get_sqe();
    if (!sqe) {
        ring.submit_and_wait(0);
        sqe = ring.get_sqe();
    }
    return sqe;
}

void arm_recv(Uring& ring, int fd) {
    io_uring_sqe* sqe = must_sqe(ring);
    sqe->opcode = IORING_OP_RECV;
    sqe->fd = fd;
    sqe->ioprio = IORING_RECV_MULTISHOT;
    sqe->flags = IOSQE_BUFFER_SELECT;
    sqe->buf_group = ring.buffer_group();
    sqe->user_data = tag(kRecv, fd);
}

void arm_pipe(UringLoop& loop, ThreadCtx& t) {
    io_uring_sqe* sqe = must_sqe(loop.ring);
    sqe->opcode = IORING_OP_READ;
    sqe->fd = t.pipe_rd;
    sqe->addr = reinterpret_cast<uint64_t>(&loop.pipe_value);
    sqe->len = sizeof loop.pipe_value;
    sqe->user_data = tag(kPipe, t.pipe_rd);
}

void arm_accept(Uring& ring, int listen_fd) {
    io_uring_sqe* sqe = must_sqe(ring);
    sqe->opcode = IORING_OP_ACCEPT;
    sqe->fd = listen_fd;
    sqe->ioprio = IORING_ACCEPT_MULTISHOT;
    sqe->accept_flags = SOCK_CLOEXEC;
    sqe->user_data = tag(kAccept, listen_fd);
}

void uring_adopt(ThreadCtx& t, int fd, void* arg) {
    arm_recv(static_cast<UringLoop*>(arg)->ring, fd);
    t.connections.fetch_add(1, std::memory_order_relaxed);
}

void run_uring(ThreadCtx& t) {
    UringLoop loop;
    int rc = loop.ring.init(4096);
    if (rc == 0) rc = loop.ring.setup_buffers(1, 1024, 2048);
    if (rc != 0) {
        std::fprintf(stderr, "hola-null: io_uring is not usable here: %s\n", std::strerror(-rc));
        std::exit(3);
    }
    Uring& ring = loop.ring;
    arm_accept(ring, t.listen_fd);
    arm_pipe(loop, t);
    uint64_t served = 0;
    while (!g_stop.load(std::memory_order_relaxed)) {
        // One system call per iteration: submits the sends and re-arms queued by the previous
        // batch of completions and waits for the next batch.
        rc = ring.submit_and_wait(1);
        if (rc != 0 && rc != -EINTR && rc != -EAGAIN && rc != -EBUSY) {
            std::fprintf(stderr, "hola-null: io_uring_enter: %s\n", std::strerror(-rc));
            break;
        }
        for (const io_uring_cqe* cqe = ring.peek(); cqe; cqe = ring.peek()) {
            const uint64_t kind = cqe->user_data >> 32;
            const int fd = static_cast<int>(cqe->user_data & 0xFFFFFFFFu);
            const int res = cqe->res;
            const uint32_t flags = cqe->flags;
            ring.seen();
            switch (kind) {
                case kAccept:
                    if (res >= 0) deal(t, res, uring_adopt, &loop);
                    if (!(flags & IORING_CQE_F_MORE)) arm_accept(ring, t.listen_fd);
                    break;
                case kPipe:
                    if (res == static_cast<int>(sizeof loop.pipe_value)) uring_adopt(t, loop.pipe_value, &loop);
                    arm_pipe(loop, t);
                    break;
                case kRecv: {
                    if (flags & IORING_CQE_F_BUFFER) {
                        const uint16_t bid = static_cast<uint16_t>(flags >> IORING_CQE_BUFFER_SHIFT);
                        if (res > 0) {
                            size_t k = count_heads(ring.buffer(bid), static_cast<size_t>(res));
                            served += k;
                            while (k > 0) {
                                const size_t chunk = k > kMaxBatch ? kMaxBatch : k;
                                io_uring_sqe* sqe = must_sqe(ring);
                                sqe->opcode = IORING_OP_SEND;
                                sqe->fd = fd;
                                sqe->addr = reinterpret_cast<uint64_t>(g_batch.data());
                                sqe->len = static_cast<uint32_t>(chunk * kResponseLen);
                                sqe->msg_flags = MSG_NOSIGNAL;
                                sqe->flags = IOSQE_CQE_SKIP_SUCCESS;  // only failed sends report back
                                sqe->user_data = tag(kSend, fd);
                                k -= chunk;
                            }
                        }
                        ring.recycle_buffer(bid, false);
                    }
                    if (res <= 0 && res != -ENOBUFS) {
                        close(fd);  // peer closed or error; the multishot request has ended
                    } else if (!(flags & IORING_CQE_F_MORE)) {
                        arm_recv(ring, fd);
                    }
                    break;
                }
                default:
                    break;  // a failed send: the receive side will notice the dead connection
            }
        }
        ring.publish_buffers();
        t.requests.store(served, std::memory_order_relaxed);
    }
}

void thread_main(ThreadCtx* t, const Options* opt) {
    if (!opt->cpus.empty()) {
        cpu_set_t set;
        CPU_ZERO(&set);
        CPU_SET(opt->cpus[static_cast<size_t>(t->id) % opt->cpus.size()], &set);
        pthread_setaffinity_np(pthread_self(), sizeof set, &set);
    }
    if (opt->uring) {
        run_uring(*t);
    } else {
        run_epoll(*t);
    }
}

}  // namespace

int main(int argc, char** argv) {
    Options opt;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        const char* v = i + 1 < argc ? argv[i + 1] : nullptr;
        if (a == "--port" && v) {
            opt.port = static_cast<uint16_t>(std::atoi(v));
            ++i;
        } else if (a == "--threads" && v) {
            opt.threads = std::atoi(v);
            ++i;
        } else if (a == "--cpus" && v) {
            if (!hola::parse_cpu_list(v, &opt.cpus)) {
                std::fprintf(stderr, "hola-null: bad --cpus list\n");
                return 2;
            }
            ++i;
        } else if (a == "--io-backend" && v) {
            opt.uring = std::string(v) == "uring";
            ++i;
        } else {
            std::fprintf(stderr, "usage: hola-null [--port N] [--threads N] [--cpus LIST] [--io-backend epoll|uring]\n");
            return 2;
        }
    }
    if (opt.threads < 1) opt.threads = 1;
    for (int k = 0; k < kMaxBatch; ++k) g_batch.append(kResponse, kResponseLen);

    sigset_t sigs;
    sigemptyset(&sigs);
    sigaddset(&sigs, SIGINT);
    sigaddset(&sigs, SIGTERM);
    pthread_sigmask(SIG_BLOCK, &sigs, nullptr);
    signal(SIGPIPE, SIG_IGN);

    for (int i = 0; i < opt.threads; ++i) {
        auto* t = new ThreadCtx();
        t->id = i;
        t->listen_fd = open_listener(opt.port);
        int p[2];
        if (pipe2(p, O_NONBLOCK | O_CLOEXEC) != 0) {
            std::perror("hola-null: pipe2");
            return 1;
        }
        t->pipe_rd = p[0];
        t->pipe_wr = p[1];
        g_threads.push_back(t);
    }
    std::vector<std::thread> threads;
    for (ThreadCtx* t : g_threads) threads.emplace_back(thread_main, t, &opt);
    std::fprintf(stderr, "hola-null listening on 127.0.0.1:%u with %d thread(s), backend %s\n", opt.port, opt.threads,
                 opt.uring ? "io_uring" : "epoll");
    int sig = 0;
    sigwait(&sigs, &sig);
    g_stop.store(true);
    // Wake every loop: a connection attempt per listener would do, the pipes are simpler.
    for (ThreadCtx* t : g_threads) {
        const int none = -1;
        ssize_t r = write(t->pipe_wr, &none, sizeof none);
        (void)r;
    }
    for (auto& th : threads) th.join();
    uint64_t total = 0;
    std::string per_thread;
    for (ThreadCtx* t : g_threads) {
        total += t->requests.load();
        per_thread += " " + std::to_string(t->requests.load()) + "/" + std::to_string(t->connections.load());
    }
    std::fprintf(stderr, "hola-null: %llu requests; per thread (requests/connections):%s\n", static_cast<unsigned long long>(total),
                 per_thread.c_str());
    return 0;
}
