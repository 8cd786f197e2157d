#include "net/event_loop.h"

#include <poll.h>
#include <sys/eventfd.h>
#include <unistd.h>

#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>

#include "net/uring.h"

namespace hola {

namespace {
constexpr uint64_t kEpollPollTag = 1;  // user_data of the multishot poll on the epoll descriptor (handler slots start at 1 << 32)
}

int64_t monotonic_ns() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return static_cast<int64_t>(ts.tv_sec) * 1000000000ll + ts.tv_nsec;
}

int64_t realtime_us() {
    timespec ts;
    clock_gettime(CLOCK_REALTIME, &ts);
    return static_cast<int64_t>(ts.tv_sec) * 1000000ll + ts.tv_nsec / 1000;
}

EventLoop::EventLoop() {
    epfd_ = epoll_create1(EPOLL_CLOEXEC);
    wake_fd_ = eventfd(0, EFD_NONBLOCK | EFD_CLOEXEC);
    if (epfd_ < 0 || wake_fd_ < 0) {
        std::perror("hola: epoll_create1/eventfd");
        std::abort();
    }
    epoll_event ev{};
    ev.events = EPOLLIN;
    ev.data.ptr = nullptr;  // nullptr marks the wake-up descriptor
    epoll_ctl(epfd_, EPOLL_CTL_ADD, wake_fd_, &ev);
    refresh_clock();
}

EventLoop::~EventLoop() {
    for (auto& fn : deferred_) fn();  // pending deletions
    uring_.reset();                   // cancels whatever is still in flight on the ring
    if (wake_fd_ >= 0) close(wake_fd_);
    if (epfd_ >= 0) close(epfd_);
}

void EventLoop::refresh_clock() {
    now_ns_ = monotonic_ns();
    now_unix_us_ = realtime_us();
    now_unix_s_ = now_unix_us_ / 1000000;
}

bool EventLoop::add(int fd, uint32_t events, EventHandler* handler) {
    epoll_event ev{};
    ev.events = events;
    ev.data.ptr = handler;
    return epoll_ctl(epfd_, EPOLL_CTL_ADD, fd, &ev) == 0;
}

bool EventLoop::modify(int fd, uint32_t events, EventHandler* handler) {
    epoll_event ev{};
    ev.events = events;
    ev.data.ptr = handler;
    return epoll_ctl(epfd_, EPOLL_CTL_MOD, fd, &ev) == 0;
}

void EventLoop::remove(int fd) { epoll_ctl(epfd_, EPOLL_CTL_DEL, fd, nullptr); }

EventLoop::TimerId EventLoop::add_timer(int64_t deadline_ns, std::function<void()> cb) {
    const TimerId id = next_timer_id_++;
    timers_.push(Timer{deadline_ns, id});
    timer_cbs_.emplace(id, std::move(cb));
    return id;
}

void EventLoop::cancel_timer(TimerId id) { timer_cbs_.erase(id); }

void EventLoop::post(std::function<void()> fn) {
    {
        std::lock_guard<std::mutex> lock(mailbox_mu_);
        mailbox_.push_back(std::move(fn));
    }
    if (!mailbox_signalled_.exchange(true, std::memory_order_acq_rel)) {
        const uint64_t one = 1;
        ssize_t r = write(wake_fd_, &one, sizeof one);
        (void)r;
    }
}

void EventLoop::stop() {
    stop_.store(true, std::memory_order_release);
    const uint64_t one = 1;
    ssize_t r = write(wake_fd_, &one, sizeof one);
    (void)r;
}

void EventLoop::drain_mailbox() {
    uint64_t counter;
    while (read(wake_fd_, &counter, sizeof counter) > 0) {
    }
    mailbox_signalled_.store(false, std::memory_order_release);
    std::vector<std::function<void()>> batch;
    {
        std::lock_guard<std::mutex> lock(mailbox_mu_);
        batch.swap(mailbox_);
    }
    for (auto& fn : batch) fn();
}

void EventLoop::run_timers() {
    while (!timers_.empty() && timers_.top().deadline <= now_ns_) {
        const TimerId id = timers_.top().id;
        timers_.pop();
        auto it = timer_cbs_.find(id);
        if (it == timer_cbs_.end()) continue;  // cancelled
        std::function<void()> cb = std::move(it->second);
        timer_cbs_.erase(it);
        cb();
    }
}

int EventLoop::next_timeout_ms() const {
    if (!deferred_.empty()) return 0;
    if (timers_.empty()) return 1000;
    const int64_t delta = timers_.top().deadline - now_ns_;
    if (delta <= 0) return 0;
    const int64_t ms = (delta + 999999) / 1000000;
    return ms > 1000 ? 1000 : static_cast<int>(ms);
}

void EventLoop::dispatch_epoll(int n, epoll_event* events) {
    bool woken = false;
    for (int i = 0; i < n; ++i) {
        auto* handler = static_cast<EventHandler*>(events[i].data.ptr);
        if (!handler) {
            woken = true;
            continue;
        }
        handler->on_events(events[i].events);
    }
    if (woken) drain_mailbox();
}

// ---- completion-based mode --------------------------------------------------------------------

std::shared_ptr<Uring> EventLoop::prepare_uring(std::string* error, unsigned entries, unsigned buffers, unsigned buffer_size) {
    std::shared_ptr<Uring> ring(new Uring());
    int rc = ring->init(entries, /*modern_only=*/true, /*disabled=*/true);
    if (rc == 0) rc = ring->setup_buffers(1, buffers, buffer_size);
    if (rc != 0) {
        if (error) *error = std::strerror(-rc);
        return nullptr;
    }
    return ring;
}

bool EventLoop::adopt_uring(std::shared_ptr<Uring> ring) {
    if (uring_ || !ring || ring->enable() != 0) return false;
    uring_ = std::move(ring);
    return true;  // run() notices at the end of this iteration and continues in run_uring()
}

int EventLoop::register_uring_handler(UringHandler* h) {
    if (!uring_free_slots_.empty()) {
        const int slot = uring_free_slots_.back();
        uring_free_slots_.pop_back();
        uring_handlers_[static_cast<size_t>(slot)] = h;
        return slot;
    }
    uring_handlers_.push_back(h);
    return static_cast<int>(uring_handlers_.size()) - 1;
}

void EventLoop::unregister_uring_handler(int slot) {
    if (slot < 0 || static_cast<size_t>(slot) >= uring_handlers_.size()) return;
    uring_handlers_[static_cast<size_t>(slot)] = nullptr;
    uring_free_slots_.push_back(slot);
}

io_uring_sqe* EventLoop::uring_sqe() {
    io_uring_sqe* sqe = uring_->get_sqe();
    // Queue full: hand what is there to the kernel. That can be refused for a moment (-EBUSY while
    // completions that overflowed the completion queue wait to be moved into it; run_uring() frees
    // a slot per completion before dispatching it, so the next attempt makes room).
    for (int attempt = 0; !sqe && attempt < 1000; ++attempt) {
        uring_->submit_and_wait_for(0);
        sqe = uring_->get_sqe();
    }
    if (!sqe) {
        std::fprintf(stderr, "hola: io_uring submission queue stays full\n");
        std::abort();
    }
    return sqe;
}

uint16_t EventLoop::uring_buffer_group() const { return uring_->buffer_group(); }

void EventLoop::arm_epoll_poll() {
    io_uring_sqe* sqe = uring_sqe();
    sqe->opcode = IORING_OP_POLL_ADD;
    sqe->fd = epfd_;
    sqe->poll32_events = POLLIN;
    sqe->len = IORING_POLL_ADD_MULTI;
    sqe->user_data = kEpollPollTag;
}

// One iteration = one io_uring_enter: it submits everything the handlers queued during the
// previous iteration (sends, re-armed receives) and waits for the next completions, bounded by
// the next timer. The epoll set rides along as one multishot poll on the epoll descriptor.
void EventLoop::run_uring() {
    constexpr int kMaxEvents = 256;
    epoll_event events[kMaxEvents];
    arm_epoll_poll();
    bool epoll_backlog = true;  // whatever the epoll loop left pending when it handed over
    while (!stop_.load(std::memory_order_acquire)) {
        const int timeout_ms = epoll_backlog ? 0 : next_timeout_ms();
        int rc;
        if (spin_ns_ > 0 && timeout_ms != 0) {
            // Busy-poll window (set_spin_ns): keep entering without sleeping while traffic is
            // recent; completions are only posted on entry (DEFER_TASKRUN), so every look is a
            // system call - the same price the epoll loop pays for its epoll_wait(0).
            rc = uring_->submit_and_wait_for(0);
            while (rc == 0 && !uring_->peek()) {
                const int64_t t = monotonic_ns();
                if (t - last_event_ns_ >= spin_ns_ || stop_.load(std::memory_order_relaxed)) break;
                if (!timers_.empty() && timers_.top().deadline <= t) break;
                __builtin_ia32_pause();
                rc = uring_->submit_and_wait_for(0);
            }
            if (rc == 0 && !uring_->peek() && (timers_.empty() || timers_.top().deadline > monotonic_ns())) {
                ++stats_.blocking_waits;
                rc = uring_->submit_and_wait_for(static_cast<int64_t>(timeout_ms) * 1000000);
            }
        } else {
            if (timeout_ms != 0) ++stats_.blocking_waits;
            rc = uring_->submit_and_wait_for(static_cast<int64_t>(timeout_ms) * 1000000);
        }
        refresh_clock();
        ++stats_.iterations;
        if (rc != 0 && rc != -ETIME && rc != -EINTR && rc != -EAGAIN && rc != -EBUSY) {
            std::fprintf(stderr, "hola: io_uring_enter: %s\n", std::strerror(-rc));
            break;
        }
        bool epoll_ready = false;
        for (const io_uring_cqe* cqe = uring_->peek(); cqe; cqe = uring_->peek()) {
            const uint64_t user_data = cqe->user_data;
            const int32_t res = cqe->res;
            const uint32_t flags = cqe->flags;
            uring_->seen();
            ++stats_.events;
            last_event_ns_ = now_ns_;
            if (user_data == kEpollPollTag) {
                epoll_ready = true;
                if (!(flags & IORING_CQE_F_MORE)) arm_epoll_poll();
                continue;
            }
            const size_t slot = static_cast<size_t>(user_data >> 32) - 1;
            const char* data = nullptr;
            uint16_t bid = 0;
            if (flags & IORING_CQE_F_BUFFER) {
                bid = static_cast<uint16_t>(flags >> IORING_CQE_BUFFER_SHIFT);
                data = uring_->buffer(bid);
            }
            if (slot < uring_handlers_.size() && uring_handlers_[slot]) uring_handlers_[slot]->on_completion(static_cast<uint32_t>(user_data), res, flags, data);
            if (flags & IORING_CQE_F_BUFFER) uring_->recycle_buffer(bid, true);
        }
        if (epoll_ready || epoll_backlog) {
            // The poll on the epoll descriptor reports *new* readiness only. The epoll set itself
            // is level-triggered and its handlers rely on that (a short read leaves a pending FIN
            // for the next round), so keep asking until nothing is ready; a descriptor that stays
            // ready for several rounds is carried over to the next iteration, which then does not
            // sleep - the same behaviour as the plain epoll loop, at the cost of one more
            // epoll_wait(0) per batch of epoll activity (not on the path of io_uring sockets).
            epoll_backlog = false;
            for (int round = 0; round < 4; ++round) {
                const int n = epoll_wait(epfd_, events, kMaxEvents, 0);
                if (n <= 0) break;
                dispatch_epoll(n, events);
                if (round == 3) epoll_backlog = true;
            }
        }
        run_timers();
        run_deferred();
    }
    run_deferred();
}

void EventLoop::run() {
    if (uring_) {
        run_uring();
        return;
    }
    constexpr int kMaxEvents = 256;
    epoll_event events[kMaxEvents];
    while (!stop_.load(std::memory_order_acquire)) {
        const int timeout_ms = next_timeout_ms();
        int n;
        if (spin_ns_ > 0 && timeout_ms != 0) {
            // Poll without sleeping while traffic is recent; fall back to a blocking wait after
            // spin_ns_ of silence (or at the next timer, whichever is first).
            n = epoll_wait(epfd_, events, kMaxEvents, 0);
            while (n == 0) {
                const int64_t t = monotonic_ns();
                if (t - last_event_ns_ >= spin_ns_ || stop_.load(std::memory_order_relaxed)) break;
                if (!timers_.empty() && timers_.top().deadline <= t) break;
                __builtin_ia32_pause();
                n = epoll_wait(epfd_, events, kMaxEvents, 0);
            }
            if (n == 0 && (timers_.empty() || timers_.top().deadline > monotonic_ns())) {
                ++stats_.blocking_waits;
                n = epoll_wait(epfd_, events, kMaxEvents, timeout_ms);
            }
        } else {
            if (timeout_ms != 0) ++stats_.blocking_waits;
            n = epoll_wait(epfd_, events, kMaxEvents, timeout_ms);
        }
        refresh_clock();
        ++stats_.iterations;
        if (n > 0) {
            stats_.events += static_cast<uint64_t>(n);
            last_event_ns_ = now_ns_;
        }
        if (n < 0) {
            if (errno == EINTR) continue;
            std::perror("hola: epoll_wait");
            break;
        }
        dispatch_epoll(n, events);
        run_timers();
        run_deferred();
        if (uring_) {  // adopt_uring() was called during this iteration
            run_uring();
            return;
        }
    }
    run_deferred();  // left by an early exit from the loop
}

void EventLoop::run_deferred() {
    while (!deferred_.empty()) {
        std::vector<std::function<void()>> batch;
        batch.swap(deferred_);
        for (auto& fn : batch) fn();
    }
}

}  // namespace hola
