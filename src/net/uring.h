// A minimal io_uring wrapper on the raw system calls (there is no liburing on the build or
// benchmark images): one ring, a provided-buffer ring for multishot receives, submission and
// completion queues accessed with the acquire/release pairs the kernel ABI asks for.
//
// Used by the I/O-ceiling probe (src/tools/null_server.cc) and by the server's optional
// completion-based backend (--io-backend uring, src/net/event_loop.cc and server.cc). The
// question behind both: how much of the per-request kernel time of the epoll + recv + send path
// (/root/reference gets the same three calls from Undertow/XNIO, /root/reference/pom.xml:122-129)
// is system-call and readiness-notification overhead that a completion-based interface removes,
// and how much is TCP/loopback protocol work that stays.
#pragma once

#include <linux/io_uring.h>
#include <sys/mman.h>
#include <sys/syscall.h>
#include <unistd.h>

#include <cerrno>
#include <cstdint>
#include <cstring>

namespace hola {

class Uring {
 public:
    Uring() = default;
    Uring(const Uring&) = delete;
    Uring& operator=(const Uring&) = delete;
    ~Uring() { close_ring(); }

    // Returns 0 or -errno. The ring is meant for one thread: it asks for SINGLE_ISSUER and
    // DEFER_TASKRUN (completions are only posted when this thread enters the kernel to wait).
    // A kernel that refuses those flags is older than 6.1; with `modern_only` that is an error
    // (-ENOSYS) - such a kernel may lack multishot receives and buffer rings too, which cannot
    // be probed for directly - otherwise a plain ring is tried.
    // `disabled`: the ring is created in the disabled state (IORING_SETUP_R_DISABLED), which lets
    // one thread create it and register its buffers and another thread - the one that will use
    // it - take it over with enable().
    int init(unsigned entries, bool modern_only = false, bool disabled = false) {
        io_uring_params p;
        std::memset(&p, 0, sizeof p);
        p.flags = IORING_SETUP_SINGLE_ISSUER | IORING_SETUP_DEFER_TASKRUN | IORING_SETUP_CQSIZE | IORING_SETUP_CLAMP;
        if (disabled) p.flags |= IORING_SETUP_R_DISABLED;
        p.cq_entries = entries * 4;
        int fd = static_cast<int>(syscall(__NR_io_uring_setup, entries, &p));
        if (fd < 0 && modern_only) return errno == EINVAL ? -ENOSYS : -errno;
        if (fd < 0) {
            std::memset(&p, 0, sizeof p);
            p.flags = IORING_SETUP_CQSIZE | IORING_SETUP_CLAMP;
            p.cq_entries = entries * 4;
            fd = static_cast<int>(syscall(__NR_io_uring_setup, entries, &p));
            if (fd < 0) return -errno;
        }
        fd_ = fd;
        sq_len_ = p.sq_off.array + p.sq_entries * sizeof(uint32_t);
        cq_len_ = p.cq_off.cqes + p.cq_entries * sizeof(io_uring_cqe);
        const bool single = (p.features & IORING_FEAT_SINGLE_MMAP) != 0;
        if (single) sq_len_ = cq_len_ = sq_len_ > cq_len_ ? sq_len_ : cq_len_;
        sq_ptr_ = mmap(nullptr, sq_len_, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_POPULATE, fd_, IORING_OFF_SQ_RING);
        if (sq_ptr_ == MAP_FAILED) return fail();
        cq_ptr_ = single ? sq_ptr_ : mmap(nullptr, cq_len_, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_POPULATE, fd_, IORING_OFF_CQ_RING);
        if (cq_ptr_ == MAP_FAILED) return fail();
        sqes_len_ = p.sq_entries * sizeof(io_uring_sqe);
        sqes_ = static_cast<io_uring_sqe*>(mmap(nullptr, sqes_len_, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_POPULATE, fd_, IORING_OFF_SQES));
        if (sqes_ == MAP_FAILED) return fail();
        auto* sq = static_cast<uint8_t*>(sq_ptr_);
        auto* cq = static_cast<uint8_t*>(cq_ptr_);
        sq_head_ = reinterpret_cast<uint32_t*>(sq + p.sq_off.head);
        sq_tail_ = reinterpret_cast<uint32_t*>(sq + p.sq_off.tail);
        sq_mask_ = *reinterpret_cast<uint32_t*>(sq + p.sq_off.ring_mask);
        sq_entries_ = p.sq_entries;
        auto* array = reinterpret_cast<uint32_t*>(sq + p.sq_off.array);
        for (uint32_t i = 0; i < p.sq_entries; ++i) array[i] = i;  // identity: slot i of the ring is SQE i
        cq_head_ = reinterpret_cast<uint32_t*>(cq + p.cq_off.head);
        cq_tail_ = reinterpret_cast<uint32_t*>(cq + p.cq_off.tail);
        cq_mask_ = *reinterpret_cast<uint32_t*>(cq + p.cq_off.ring_mask);
        cqes_ = reinterpret_cast<io_uring_cqe*>(cq + p.cq_off.cqes);
        local_tail_ = *sq_tail_;
        return 0;
    }

    int fd() const { return fd_; }

    // Takes over a ring created with `disabled`: from now on the calling thread is its only
    // submitter. Returns 0 or -errno.
    int enable() { return syscall(__NR_io_uring_register, fd_, IORING_REGISTER_ENABLE_RINGS, nullptr, 0) < 0 ? -errno : 0; }

    // Next free submission entry, zeroed; nullptr when the queue is full (submit first).
    io_uring_sqe* get_sqe() {
        const uint32_t head = __atomic_load_n(sq_head_, __ATOMIC_ACQUIRE);
        if (local_tail_ - head >= sq_entries_) return nullptr;
        io_uring_sqe* sqe = &sqes_[local_tail_ & sq_mask_];
        ++local_tail_;
        std::memset(sqe, 0, sizeof *sqe);
        return sqe;
    }
    unsigned pending() const { return local_tail_ - *sq_tail_; }

    // Publishes the prepared entries and enters the kernel once: submits them and, with
    // wait_for > 0, sleeps until that many completions are there. Returns 0 or -errno.
    int submit_and_wait(unsigned wait_for) {
        const unsigned to_submit = local_tail_ - *sq_tail_;
        __atomic_store_n(sq_tail_, local_tail_, __ATOMIC_RELEASE);
        if (to_submit == 0 && wait_for == 0) return 0;
        const long r = syscall(__NR_io_uring_enter, fd_, to_submit, wait_for, IORING_ENTER_GETEVENTS, nullptr, 0);
        return r < 0 ? -errno : 0;
    }

    // As submit_and_wait(1), but gives up after `timeout_ns` (0: do not wait at all, just submit
    // and let the kernel post what has completed). Returns 0, -ETIME or another -errno.
    int submit_and_wait_for(int64_t timeout_ns) {
        const unsigned to_submit = local_tail_ - *sq_tail_;
        __atomic_store_n(sq_tail_, local_tail_, __ATOMIC_RELEASE);
        if (timeout_ns <= 0) {
            const long r = syscall(__NR_io_uring_enter, fd_, to_submit, 0, IORING_ENTER_GETEVENTS, nullptr, 0);
            return r < 0 ? -errno : 0;
        }
        __kernel_timespec ts;
        ts.tv_sec = timeout_ns / 1000000000ll;
        ts.tv_nsec = timeout_ns % 1000000000ll;
        io_uring_getevents_arg arg;
        std::memset(&arg, 0, sizeof arg);
        arg.ts = reinterpret_cast<uint64_t>(&ts);
        const long r = syscall(__NR_io_uring_enter, fd_, to_submit, 1, IORING_ENTER_GETEVENTS | IORING_ENTER_EXT_ARG, &arg, sizeof arg);
        return r < 0 ? -errno : 0;
    }

    // Completion queue: for (cqe = peek(); cqe; cqe = peek()) { ...; seen(); }
    const io_uring_cqe* peek() const {
        const uint32_t head = *cq_head_;
        if (head == __atomic_load_n(cq_tail_, __ATOMIC_ACQUIRE)) return nullptr;
        return &cqes_[head & cq_mask_];
    }
    void seen() { __atomic_store_n(cq_head_, *cq_head_ + 1, __ATOMIC_RELEASE); }

    // Provided buffers: `count` (a power of two) buffers of `size` bytes under group id `bgid`,
    // handed to the kernel through a shared ring. Returns 0 or -errno.
    int setup_buffers(uint16_t bgid, unsigned count, unsigned size) {
        buf_count_ = count;
        buf_size_ = size;
        bgid_ = bgid;
        const size_t ring_bytes = count * sizeof(io_uring_buf);
        buf_ring_ = static_cast<io_uring_buf_ring*>(mmap(nullptr, ring_bytes, PROT_READ | PROT_WRITE, MAP_ANONYMOUS | MAP_PRIVATE | MAP_POPULATE, -1, 0));
        if (buf_ring_ == MAP_FAILED) return -errno;
        // Not pre-faulted: the pages are touched when the kernel first delivers into them, which
        // keeps start-up short (4 MB of page faults would cost over a millisecond per thread).
        buf_mem_ = static_cast<char*>(mmap(nullptr, static_cast<size_t>(count) * size, PROT_READ | PROT_WRITE, MAP_ANONYMOUS | MAP_PRIVATE, -1, 0));
        if (buf_mem_ == MAP_FAILED) return -errno;
        io_uring_buf_reg reg;
        std::memset(&reg, 0, sizeof reg);
        reg.ring_addr = reinterpret_cast<uint64_t>(buf_ring_);
        reg.ring_entries = count;
        reg.bgid = bgid;
        if (syscall(__NR_io_uring_register, fd_, IORING_REGISTER_PBUF_RING, &reg, 1) < 0) return -errno;
        buf_tail_ = 0;
        for (unsigned i = 0; i < count; ++i) recycle_buffer(static_cast<uint16_t>(i), false);
        publish_buffers();
        return 0;
    }
    char* buffer(uint16_t id) const { return buf_mem_ + static_cast<size_t>(id) * buf_size_; }
    unsigned buffer_size() const { return buf_size_; }
    uint16_t buffer_group() const { return bgid_; }
    // Gives buffer `id` back to the kernel (publish with publish_buffers(), or pass true).
    void recycle_buffer(uint16_t id, bool publish = true) {
        io_uring_buf* b = &buf_ring_->bufs[buf_tail_ & (buf_count_ - 1)];
        b->addr = reinterpret_cast<uint64_t>(buffer(id));
        b->len = buf_size_;
        b->bid = id;
        ++buf_tail_;
        if (publish) publish_buffers();
    }
    void publish_buffers() { __atomic_store_n(&buf_ring_->tail, buf_tail_, __ATOMIC_RELEASE); }

 private:
    int fail() {
        const int e = errno;
        close_ring();
        return -e;
    }
    void close_ring() {
        if (sqes_ && sqes_ != MAP_FAILED) munmap(sqes_, sqes_len_);
        if (cq_ptr_ && cq_ptr_ != MAP_FAILED && cq_ptr_ != sq_ptr_) munmap(cq_ptr_, cq_len_);
        if (sq_ptr_ && sq_ptr_ != MAP_FAILED) munmap(sq_ptr_, sq_len_);
        if (buf_ring_ && buf_ring_ != MAP_FAILED) munmap(buf_ring_, buf_count_ * sizeof(io_uring_buf));
        if (buf_mem_ && buf_mem_ != MAP_FAILED) munmap(buf_mem_, static_cast<size_t>(buf_count_) * buf_size_);
        if (fd_ >= 0) close(fd_);
        sqes_ = nullptr;
        cq_ptr_ = sq_ptr_ = nullptr;
        buf_ring_ = nullptr;
        buf_mem_ = nullptr;
        fd_ = -1;
    }

    int fd_ = -1;
    void* sq_ptr_ = nullptr;
    void* cq_ptr_ = nullptr;
    size_t sq_len_ = 0, cq_len_ = 0, sqes_len_ = 0;
    io_uring_sqe* sqes_ = nullptr;
    uint32_t* sq_head_ = nullptr;
    uint32_t* sq_tail_ = nullptr;
    uint32_t sq_mask_ = 0, sq_entries_ = 0, local_tail_ = 0;
    uint32_t* cq_head_ = nullptr;
    uint32_t* cq_tail_ = nullptr;
    uint32_t cq_mask_ = 0;
    io_uring_cqe* cqes_ = nullptr;

    io_uring_buf_ring* buf_ring_ = nullptr;
    char* buf_mem_ = nullptr;
    unsigned buf_count_ = 0, buf_size_ = 0;
    uint16_t buf_tail_ = 0, bgid_ = 0;
};

}  // namespace hola
