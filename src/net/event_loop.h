// One epoll instance + timer heap + cross-thread mailbox, driven by exactly one thread; optionally
// with an io_uring instance in front of it for completion-based socket I/O.
//
// The reference borrows its I/O model from Undertow/XNIO (I/O threads + a worker pool,
// /root/reference/pom.xml:122-129). The rebuild uses one event loop per core instead: a
// connection lives on one loop for its whole life, handlers never block, and slow operations
// (the downstream call, the metrics stream) are expressed as timers and non-blocking sockets on
// the same loop. No locks are taken on the request path.
#pragma once

#include <sys/epoll.h>

#include <atomic>
#include <cstdint>
#include <functional>
#include <memory>
#include <mutex>
#include <queue>
#include <string>
#include <unordered_map>
#include <vector>

struct io_uring_sqe;  // <linux/io_uring.h>

namespace hola {

class EventHandler {
 public:
    virtual ~EventHandler() = default;
    virtual void on_events(uint32_t events) = 0;
};

// Receiver of io_uring completions (uring mode only). `tag` is whatever the submitter put into
// the low 32 bits of the request's user_data; `data` points at the received bytes for a receive
// that used a provided buffer (valid during the call only), else it is null.
class UringHandler {
 public:
    virtual ~UringHandler() = default;
    virtual void on_completion(uint32_t tag, int32_t res, uint32_t cqe_flags, const char* data) = 0;
};

class Uring;

int64_t monotonic_ns();
int64_t realtime_us();

class EventLoop {
 public:
    using TimerId = uint64_t;

    EventLoop();
    ~EventLoop();  // out of line: Uring is an incomplete type here
    EventLoop(const EventLoop&) = delete;
    EventLoop& operator=(const EventLoop&) = delete;

    // fd registration (loop thread only).
    bool add(int fd, uint32_t events, EventHandler* handler);
    bool modify(int fd, uint32_t events, EventHandler* handler);
    void remove(int fd);

    // Timers (loop thread only). The callback runs once, on the loop thread.
    TimerId add_timer(int64_t deadline_ns, std::function<void()> cb);
    void cancel_timer(TimerId id);

    // Runs `fn` on the loop thread soon. Safe from any thread.
    void post(std::function<void()> fn);

    // Clock values sampled once per loop iteration.
    int64_t now_ns() const { return now_ns_; }
    int64_t now_unix_s() const { return now_unix_s_; }
    int64_t now_unix_us() const { return now_unix_us_; }
    void refresh_clock();

    // Busy-poll window: after the last I/O event the loop keeps polling without sleeping for
    // this long before it blocks in epoll_wait. On a pinned, otherwise idle core this trades
    // CPU time for not paying a sleep/wake-up (expensive inside a VM) on every request.
    // 0 (default) always blocks.
    void set_spin_ns(int64_t ns) { spin_ns_ = ns; }

    struct LoopStats {
        uint64_t iterations = 0;  // epoll_wait calls that returned
        uint64_t events = 0;      // fd events dispatched
        uint64_t blocking_waits = 0;
    };
    const LoopStats& stats() const { return stats_; }

    // Runs until stop() is called.
    void run();
    void stop();  // safe from any thread
    bool stopping() const { return stop_.load(std::memory_order_relaxed); }

    // Completion-based I/O. A ring is made in two steps so that a loop never waits for one:
    // prepare_uring() - any thread - creates it (disabled) and registers its buffers, which takes
    // 0.1-0.3 ms and a few memory mappings; a worker should be answering requests meanwhile.
    // adopt_uring() - on the loop's own thread, e.g. posted to it - takes the ring over; the loop
    // switches to it at the end of the current iteration. Sockets registered with epoll before
    // that stay there: in uring mode the epoll set is still served - it is polled through the
    // ring - so listeners, the mailbox and the outbound client work unchanged; sockets that want
    // completions register a UringHandler and submit their own receives and sends:
    //     sqe = loop.uring_sqe(); fill it; sqe->user_data = loop.uring_user_data(slot, tag);
    // Submissions are flushed once per loop iteration, together with the wait for completions.
    // prepare_uring() returns null (reason in *error) when the kernel refuses: io_uring compiled
    // out or switched off (kernel.io_uring_disabled), blocked by a seccomp profile as in many
    // container runtimes, or older than Linux 6.1.
    static std::shared_ptr<Uring> prepare_uring(std::string* error, unsigned entries = 1024, unsigned buffers = 512, unsigned buffer_size = 2048);
    bool adopt_uring(std::shared_ptr<Uring> prepared);
    bool uring_enabled() const { return uring_ != nullptr; }
    int register_uring_handler(UringHandler* h);
    // Only once no operation submitted under the slot is in flight any more (the slot is reused).
    void unregister_uring_handler(int slot);
    static uint64_t uring_user_data(int slot, uint32_t tag) { return (static_cast<uint64_t>(slot + 1) << 32) | tag; }
    ::io_uring_sqe* uring_sqe();  // never null: flushes the submission queue when it is full
    uint16_t uring_buffer_group() const;

    // Deferred deletion: objects whose handlers may still be referenced by the current batch of
    // epoll events are parked here and destroyed at the end of the iteration.
    void defer(std::function<void()> fn) { deferred_.push_back(std::move(fn)); }

 private:
    struct Timer {
        int64_t deadline;
        TimerId id;
        bool operator>(const Timer& o) const { return deadline > o.deadline || (deadline == o.deadline && id > o.id); }
    };

    void drain_mailbox();
    void run_timers();
    void run_deferred();
    void run_uring();
    void dispatch_epoll(int n, epoll_event* events);
    void arm_epoll_poll();
    int next_timeout_ms() const;

    int epfd_ = -1;
    int wake_fd_ = -1;
    std::atomic<bool> stop_{false};
    int64_t now_ns_ = 0;
    int64_t now_unix_s_ = 0;
    int64_t now_unix_us_ = 0;

    std::priority_queue<Timer, std::vector<Timer>, std::greater<Timer>> timers_;
    std::unordered_map<TimerId, std::function<void()>> timer_cbs_;
    TimerId next_timer_id_ = 1;

    std::mutex mailbox_mu_;
    std::vector<std::function<void()>> mailbox_;
    std::atomic<bool> mailbox_signalled_{false};
    std::vector<std::function<void()>> deferred_;
    std::shared_ptr<Uring> uring_;
    std::vector<UringHandler*> uring_handlers_;
    std::vector<int> uring_free_slots_;
    int64_t spin_ns_ = 0;
    int64_t last_event_ns_ = 0;
    LoopStats stats_;
};

}  // namespace hola
