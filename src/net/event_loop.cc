#include "net/event_loop.h"

#include <sys/eventfd.h>
#include <unistd.h>

#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <ctime>

namespace hola {

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

void EventLoop::run() {
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
        run_timers();
        run_deferred();
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
