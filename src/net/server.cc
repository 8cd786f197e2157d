#include "net/server.h"

#include <arpa/inet.h>
#include <fcntl.h>
#include <linux/io_uring.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <pthread.h>
#include <sched.h>
#include <sys/socket.h>
#include <unistd.h>

#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>

namespace hola {

namespace {

constexpr size_t kInitialInputCapacity = 4096;

// Canned error responses produced by the server core itself (malformed or oversized heads).
void append_core_error(std::string* out, int status, const char* reason, std::string_view date) {
    out->append("HTTP/1.1 ");
    out->append(std::to_string(status));
    out->push_back(' ');
    out->append(reason);
    out->append("\r\nConnection: close\r\nContent-Length: 0\r\nDate: ");
    out->append(date);
    out->append("\r\n\r\n");
}

// Single-writer counter bump: a plain load/store pair instead of a locked read-modify-write.
inline void bump(std::atomic<uint64_t>& counter) {
    counter.store(counter.load(std::memory_order_relaxed) + 1, std::memory_order_relaxed);
}

}  // namespace

void set_nodelay(int fd) {
    const int one = 1;
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
}

// ---------------------------------------------------------------------------------------------
// Connection

Connection::~Connection() {
    if (fd_ >= 0) close(fd_);
    if (zombie_fd_ >= 0) close(zombie_fd_);
}

void Connection::start(int fd) {
    fd_ = fd;
    if (!in_) {
        in_.reset(new char[kInitialInputCapacity]);
        in_cap_ = kInitialInputCapacity;
    }
    in_len_ = in_pos_ = 0;
    out_.clear();
    out_sent_ = 0;
    want_write_ = close_after_write_ = deferred_ = streaming_ = peer_closed_ = in_process_ = false;
    discard_remaining_ = 0;
    discard_chunked_.reset();
    last_active_ns_ = worker_->loop().now_ns();
    if (worker_->loop().uring_enabled()) {
        uring_ = true;
        if (uring_slot_ < 0) uring_slot_ = worker_->loop().register_uring_handler(this);
        recv_armed_ = recv_cancelling_ = send_inflight_ = zombie_ = false;
        ops_ = 0;
        sending_.clear();
        sending_sent_ = 0;
        uring_arm_recv();
        return;
    }
    registered_events_ = EPOLLIN | EPOLLRDHUP;
    if (!worker_->loop().add(fd_, registered_events_, this)) close_now();
}

void Connection::close_now() {
    if (fd_ < 0) return;
    const ConnRef old = ref();
    const bool notify = deferred_ || streaming_;
    if (uring_) {
        // Operations in flight hold the socket open and will still complete. shutdown() makes
        // them complete now; the descriptor is closed, and the slot given back, by the last one.
        zombie_fd_ = fd_;
        if (ops_ > 0) {
            zombie_ = true;
            shutdown(fd_, SHUT_RDWR);
        }
    } else {
        worker_->loop().remove(fd_);
        close(fd_);
    }
    fd_ = -1;
    ++generation_;
    deferred_ = streaming_ = false;
    if (in_cap_ > kInitialInputCapacity * 4) {  // do not let one large request pin memory
        in_.reset();
        in_cap_ = 0;
    }
    if (out_.capacity() > 65536) std::string().swap(out_);
    out_.clear();
    if (notify) worker_->app_->on_connection_closed(*worker_, old);
    if (zombie_) return;  // uring_reap_zombie() finishes the job
    if (zombie_fd_ >= 0) {
        close(zombie_fd_);
        zombie_fd_ = -1;
    }
    worker_->release(this);
}

// ---- io_uring backend ---------------------------------------------------------------------------

namespace {
enum : uint32_t { kOpRecv = 1, kOpSend = 2, kOpCancel = 3 };
}

void Connection::uring_reap_zombie() {
    if (!zombie_ || ops_ > 0) return;
    zombie_ = false;
    close(zombie_fd_);
    zombie_fd_ = -1;
    if (sending_.capacity() > 65536) std::string().swap(sending_);
    sending_.clear();
    sending_sent_ = 0;
    worker_->release(this);
}

void Connection::uring_arm_recv() {
    EventLoop& loop = worker_->loop();
    io_uring_sqe* sqe = loop.uring_sqe();
    sqe->opcode = IORING_OP_RECV;
    sqe->fd = fd_;
    sqe->ioprio = IORING_RECV_MULTISHOT;
    sqe->flags = IOSQE_BUFFER_SELECT;
    sqe->buf_group = loop.uring_buffer_group();
    sqe->user_data = EventLoop::uring_user_data(uring_slot_, kOpRecv);
    recv_armed_ = true;
    recv_cancelling_ = false;
    ++ops_;
}

void Connection::uring_cancel_recv() {
    io_uring_sqe* sqe = worker_->loop().uring_sqe();
    sqe->opcode = IORING_OP_ASYNC_CANCEL;
    sqe->addr = EventLoop::uring_user_data(uring_slot_, kOpRecv);
    sqe->user_data = EventLoop::uring_user_data(uring_slot_, kOpCancel);
    recv_cancelling_ = true;
    ++ops_;
}

void Connection::uring_submit_send() {
    io_uring_sqe* sqe = worker_->loop().uring_sqe();
    sqe->opcode = IORING_OP_SEND;
    sqe->fd = fd_;
    sqe->addr = reinterpret_cast<uint64_t>(sending_.data() + sending_sent_);
    sqe->len = static_cast<uint32_t>(std::min<size_t>(sending_.size() - sending_sent_, 1u << 30));
    sqe->msg_flags = MSG_NOSIGNAL;
    sqe->user_data = EventLoop::uring_user_data(uring_slot_, kOpSend);
    send_inflight_ = true;
    ++ops_;
}

void Connection::uring_input(const char* data, size_t len) {
    if (streaming_) return;  // a metrics stream has no use for further input
    if (in_pos_ > 0 && in_pos_ == in_len_) in_pos_ = in_len_ = 0;
    if (in_len_ + len > in_cap_) {
        if (in_pos_ > 0) {
            std::memmove(in_.get(), in_.get() + in_pos_, in_len_ - in_pos_);
            in_len_ -= in_pos_;
            in_pos_ = 0;
        }
        if (in_len_ + len > in_cap_) {
            // Unlike a readiness loop, completions cannot be refused: the bytes are here. Input
            // is throttled by cancelling the receive (update_interest), which leaves at most what
            // the socket had buffered; a peer that still gets far beyond the limit is dropped.
            const size_t hard_limit = 4 * (worker_->options().max_header_bytes + kInitialInputCapacity) + (1u << 20);
            if (in_len_ + len > hard_limit) {
                close_now();
                return;
            }
            size_t new_cap = in_cap_ ? in_cap_ : kInitialInputCapacity;
            while (new_cap < in_len_ + len) new_cap *= 2;
            std::unique_ptr<char[]> bigger(new char[new_cap]);
            std::memcpy(bigger.get(), in_.get(), in_len_);
            in_ = std::move(bigger);
            in_cap_ = new_cap;
        }
    }
    std::memcpy(in_.get() + in_len_, data, len);
    in_len_ += len;
    last_active_ns_ = worker_->loop().now_ns();
    process_input();
}

void Connection::on_completion(uint32_t tag, int32_t res, uint32_t cqe_flags, const char* data) {
    switch (tag & 3u) {
        case kOpRecv: {
            const bool more = (cqe_flags & IORING_CQE_F_MORE) != 0;
            if (!more) {
                recv_armed_ = recv_cancelling_ = false;
                --ops_;
            }
            if (zombie_ || fd_ < 0) {
                uring_reap_zombie();
                return;
            }
            if (res > 0) {
                if (data) uring_input(data, static_cast<size_t>(res));
            } else if (res == 0) {
                on_peer_closed();
                if (fd_ < 0) return;
            } else if (res != -ENOBUFS && res != -ECANCELED && res != -EINTR && res != -EAGAIN) {
                close_now();
                return;
            }
            if (fd_ >= 0 && !more) update_interest();  // a multishot that ended is re-armed when reading is still wanted
            return;
        }
        case kOpSend:
            --ops_;
            send_inflight_ = false;
            if (zombie_ || fd_ < 0) {
                uring_reap_zombie();
                return;
            }
            if (res < 0 && res != -EINTR && res != -EAGAIN) {
                close_now();
                return;
            }
            if (res > 0) {
                sending_sent_ += static_cast<size_t>(res);
                if (sending_sent_ < sending_.size()) want_write_ = true;  // short send: the peer is not keeping up
            }
            flush();
            return;
        case kOpCancel:
            --ops_;
            uring_reap_zombie();
            return;
        default:
            return;
    }
}

// ---- both backends --------------------------------------------------------------------------------


bool Connection::input_full() const {
    return in_pos_ == 0 && in_len_ == in_cap_ && in_cap_ >= worker_->options().max_header_bytes + kInitialInputCapacity;
}

void Connection::update_interest() {
    if (fd_ < 0) return;
    if (uring_) {
        // The same decisions as below, expressed as "is a receive in flight".
        bool can_read = !want_write_ || streaming_;
        if (deferred_ && in_len_ - in_pos_ >= worker_->options().max_header_bytes) can_read = false;
        if (peer_closed_) can_read = false;
        if (can_read && !recv_armed_) {
            uring_arm_recv();
        } else if (!can_read && recv_armed_ && !recv_cancelling_) {
            uring_cancel_recv();
        }
        return;
    }
    uint32_t ev = EPOLLRDHUP;
    if (want_write_) ev |= EPOLLOUT;
    bool can_read = !want_write_ || streaming_;
    if (deferred_ && in_len_ >= worker_->options().max_header_bytes) can_read = false;
    if (peer_closed_) can_read = false;
    if (can_read) ev |= EPOLLIN;
    if (peer_closed_) ev &= ~static_cast<uint32_t>(EPOLLRDHUP);
    // Nothing more can be taken in until the pending request is answered: with a level-triggered
    // loop neither "readable" nor "peer half-closed" may stay armed, or the loop spins on an event
    // it cannot consume. EPOLLHUP/EPOLLERR are always reported, so a dead peer is still noticed.
    if (!streaming_ && input_full()) ev &= ~static_cast<uint32_t>(EPOLLIN | EPOLLRDHUP);
    if (ev == registered_events_) return;
    registered_events_ = ev;
    worker_->loop().modify(fd_, ev, this);
}

void Connection::on_events(uint32_t events) {
    if (fd_ < 0) return;  // closed earlier in this batch of events
    if (events & EPOLLERR) {
        close_now();
        return;
    }
    if (events & EPOLLOUT) {
        flush();
        if (fd_ < 0) return;
    }
    if (!(events & (EPOLLIN | EPOLLHUP | EPOLLRDHUP))) return;
    if (peer_closed_) {
        if (events & EPOLLHUP) close_now();
        return;
    }
    last_active_ns_ = worker_->loop().now_ns();
    const size_t max_cap = worker_->options().max_header_bytes + kInitialInputCapacity;
    while (fd_ >= 0) {
        // A metrics stream has no use for further input: whatever the client sends after the
        // request is read and dropped, so that the socket never stays readable.
        if (streaming_) in_len_ = in_pos_ = 0;
        if (in_len_ == in_cap_) {
            if (in_pos_ > 0) {
                std::memmove(in_.get(), in_.get() + in_pos_, in_len_ - in_pos_);
                in_len_ -= in_pos_;
                in_pos_ = 0;
            } else if (in_cap_ < max_cap) {
                const size_t new_cap = std::min(max_cap, in_cap_ * 2);
                std::unique_ptr<char[]> bigger(new char[new_cap]);
                std::memcpy(bigger.get(), in_.get(), in_len_);
                in_ = std::move(bigger);
                in_cap_ = new_cap;
            } else {
                // Full and nothing can be consumed (deferred request with a greedy pipeliner).
                update_interest();
                return;
            }
        }
        const size_t space = in_cap_ - in_len_;
        const ssize_t n = recv(fd_, in_.get() + in_len_, space, 0);
        if (n > 0) {
            if (!streaming_) {
                in_len_ += static_cast<size_t>(n);
                process_input();
            }
            if (static_cast<size_t>(n) < space) break;  // socket drained
            if (want_write_ && !streaming_) break;      // peer is not reading: stop producing for it
            continue;
        }
        if (n == 0) {
            on_peer_closed();
            return;
        }
        if (errno == EINTR) continue;
        if (errno == EAGAIN || errno == EWOULDBLOCK) break;
        close_now();
        return;
    }
}

// The peer has sent its FIN: no further request will come, but everything received so far is
// still owed an answer - including requests that sit unparsed in the buffer behind a deferred
// one (a client may pipeline a batch and half-close at once).
void Connection::on_peer_closed() {
    peer_closed_ = true;
    if (streaming_ || (!deferred_ && !output_pending())) {
        close_now();  // not deferred: every complete request in the buffer has been parsed and answered already
        return;
    }
    if (!deferred_) close_after_write_ = true;  // what is being written is the last answer
    update_interest();
}

bool Connection::discard_body_bytes() {
    const size_t avail = in_len_ - in_pos_;
    if (discard_chunked_) {
        const size_t used = discard_chunked_->feed(in_.get() + in_pos_, avail, nullptr);
        in_pos_ += used;
        if (discard_chunked_->failed()) {
            discard_chunked_.reset();
            close_after_write_ = true;
            return false;
        }
        if (!discard_chunked_->done()) return false;
        discard_chunked_.reset();
        return true;
    }
    const size_t take = static_cast<size_t>(std::min<int64_t>(discard_remaining_, static_cast<int64_t>(avail)));
    in_pos_ += take;
    discard_remaining_ -= static_cast<int64_t>(take);
    return discard_remaining_ == 0;
}

void Connection::process_input() {
    if (in_process_) return;
    in_process_ = true;
    while (fd_ >= 0 && !deferred_ && !streaming_ && !close_after_write_) {
        if ((discard_remaining_ > 0 || discard_chunked_) && !discard_body_bytes()) break;
        if (in_pos_ >= in_len_) break;
        RequestHead head;
        const ParseStatus st = parse_request_head(in_.get() + in_pos_, in_len_ - in_pos_, &head);
        if (st == ParseStatus::kIncomplete) {
            if (in_len_ - in_pos_ >= worker_->options().max_header_bytes) {
                bump(worker_->stats_.bad_requests);
                append_core_error(&out_, 431, "Request Header Fields Too Large", worker_->http_date());
                close_after_write_ = true;
            }
            break;
        }
        if (st == ParseStatus::kBad) {
            bump(worker_->stats_.bad_requests);
            append_core_error(&out_, 400, "Bad Request", worker_->http_date());
            close_after_write_ = true;
            break;
        }
        in_pos_ += head.head_len;
        if (!head.path_is_canonical) {
            // Route by the path the reference's container would route by (core/http.h,
            // canonicalise_path); the path as sent stays available as raw_path and target.
            if (!canonicalise_path(head.raw_path, &worker_->canonical_path_)) {
                bump(worker_->stats_.bad_requests);
                append_core_error(&out_, 400, "Bad Request", worker_->http_date());
                close_after_write_ = true;
                break;
            }
            head.path = worker_->canonical_path_;
            head.path_is_canonical = true;
        }
        if (head.chunked) {
            discard_chunked_.reset(new ChunkedDecoder());
        } else if (head.content_length > 0) {
            discard_remaining_ = head.content_length;
        }
        if (!head.keep_alive || worker_->draining()) close_after_write_ = true;
        if (head.expect_continue && (head.chunked || head.content_length > 0)) close_after_write_ = true;
        bump(worker_->stats_.requests);
        const Disposition d = worker_->app_->handle(*worker_, *this, head);
        if (d == Disposition::kDeferred) {
            deferred_ = true;
        } else if (d == Disposition::kStream) {
            streaming_ = true;
        }
    }
    if (fd_ >= 0) {
        if (in_pos_ >= in_len_) {
            in_pos_ = in_len_ = 0;
        } else if (in_pos_ > 0 && in_pos_ >= in_cap_ / 2) {
            std::memmove(in_.get(), in_.get() + in_pos_, in_len_ - in_pos_);
            in_len_ -= in_pos_;
            in_pos_ = 0;
        }
    }
    in_process_ = false;
    // Everything the peer sent before its FIN has been parsed now (what is left, if anything, is
    // a fragment that will never be completed): the answers being flushed are the last ones.
    if (fd_ >= 0 && peer_closed_ && !deferred_ && !streaming_) close_after_write_ = true;
    flush();
}

void Connection::flush() {
    if (fd_ < 0) return;
    if (uring_) {
        if (send_inflight_) return;  // its completion comes back here
        if (sending_sent_ >= sending_.size()) {
            sending_.clear();
            sending_sent_ = 0;
            if (!out_.empty()) out_.swap(sending_);  // out_ gets the other buffer, with its capacity
        }
        if (sending_sent_ < sending_.size()) {
            if (sending_.size() - sending_sent_ > (1u << 20)) want_write_ = true;
            uring_submit_send();
            if (want_write_) update_interest();
            return;
        }
        want_write_ = false;
        if (close_after_write_ && !deferred_ && !streaming_) {
            close_now();
            return;
        }
        update_interest();
        return;
    }
    while (out_sent_ < out_.size()) {
        const ssize_t n = send(fd_, out_.data() + out_sent_, out_.size() - out_sent_, MSG_NOSIGNAL);
        if (n > 0) {
            out_sent_ += static_cast<size_t>(n);
            continue;
        }
        if (n < 0 && errno == EINTR) continue;
        if (n < 0 && (errno == EAGAIN || errno == EWOULDBLOCK)) {
            want_write_ = true;
            update_interest();
            return;
        }
        close_now();
        return;
    }
    out_.clear();
    out_sent_ = 0;
    want_write_ = false;
    if (close_after_write_ && !deferred_ && !streaming_) {
        close_now();
        return;
    }
    update_interest();
}

// ---------------------------------------------------------------------------------------------
// Worker

Worker::Worker(int id, const ServerOptions& options, Application* app) : id_(id), options_(options), app_(app) {}

Worker::~Worker() {
    if (listen_fd_ >= 0) close(listen_fd_);
}

std::string_view Worker::http_date() {
    const int64_t s = loop_.now_unix_s();
    if (s != date_second_) {
        date_second_ = s;
        format_http_date(s, date_);
    }
    return std::string_view(date_, 29);
}

bool Worker::open_listener(std::string* error) {
    // An address with a colon is IPv6; "::" then means dual stack (IPv4 clients arrive as
    // v4-mapped addresses on the same socket).
    const bool v6 = options_.bind_address.find(':') != std::string::npos;
    sockaddr_storage storage{};
    socklen_t addr_len = 0;
    if (v6) {
        auto* a = reinterpret_cast<sockaddr_in6*>(&storage);
        a->sin6_family = AF_INET6;
        a->sin6_port = htons(options_.port);
        addr_len = sizeof *a;
        if (inet_pton(AF_INET6, options_.bind_address.c_str(), &a->sin6_addr) != 1) {
            *error = "bad bind address: " + options_.bind_address;
            return false;
        }
    } else {
        auto* a = reinterpret_cast<sockaddr_in*>(&storage);
        a->sin_family = AF_INET;
        a->sin_port = htons(options_.port);
        addr_len = sizeof *a;
        if (inet_pton(AF_INET, options_.bind_address.c_str(), &a->sin_addr) != 1) {
            *error = "bad bind address: " + options_.bind_address;
            return false;
        }
    }
    const int fd = socket(v6 ? AF_INET6 : AF_INET, SOCK_STREAM | SOCK_NONBLOCK | SOCK_CLOEXEC, 0);
    if (fd < 0) {
        *error = std::string("socket: ") + std::strerror(errno);
        return false;
    }
    const int one = 1;
    const int zero = 0;
    setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
    if (options_.accept_mode != AcceptMode::kRoundRobin) setsockopt(fd, SOL_SOCKET, SO_REUSEPORT, &one, sizeof one);
    if (v6) setsockopt(fd, IPPROTO_IPV6, IPV6_V6ONLY, &zero, sizeof zero);
    if (bind(fd, reinterpret_cast<sockaddr*>(&storage), addr_len) != 0) {
        *error = "bind " + options_.bind_address + ":" + std::to_string(options_.port) + ": " + std::strerror(errno);
        close(fd);
        return false;
    }
    if (listen(fd, options_.backlog) != 0) {
        *error = std::string("listen: ") + std::strerror(errno);
        close(fd);
        return false;
    }
    listen_fd_ = fd;
    return true;
}

void Worker::on_events(uint32_t /*events*/) {
    // Listening socket readable: accept until the queue is empty.
    while (listen_fd_ >= 0) {
        const int fd = accept4(listen_fd_, nullptr, nullptr, SOCK_NONBLOCK | SOCK_CLOEXEC);
        if (fd < 0) {
            if (errno == EINTR || errno == ECONNABORTED) continue;
            if (errno == EMFILE || errno == ENFILE || errno == ENOBUFS || errno == ENOMEM) {
                // Out of descriptors (or memory): the pending connection stays in the backlog and
                // the listener stays readable, so a level-triggered loop would spin. Stop
                // listening for a moment instead and let open connections finish.
                std::fprintf(stderr, "hola: accept: %s - pausing accepts for 100 ms\n", std::strerror(errno));
                loop_.remove(listen_fd_);
                loop_.add_timer(loop_.now_ns() + 100000000, [this] {
                    if (listen_fd_ >= 0 && !draining_) loop_.add(listen_fd_, EPOLLIN, this);
                });
            }
            return;
        }
        set_nodelay(fd);
            // The connection is counted at the moment of the decision, not when the target gets round
        // to adopting it: several acceptors deciding at the same time must see each other's choices.
        Worker* target = this;
        if (peers_ && peers_->size() > 1 && options_.accept_mode == AcceptMode::kBalanced) {
            target = least_loaded();  // has counted it already
        } else {
            if (peers_ && peers_->size() > 1) target = (*peers_)[rr_next_++ % peers_->size()];
            target->stats_.connections_open.fetch_add(1, std::memory_order_relaxed);
        }
        if (target != this) {
            target->loop().post([target, fd] { target->adopt(fd); });
        } else {
            adopt(fd);
        }
    }
}

// The worker a new connection should go to: this one if no other has fewer open connections,
// else the least loaded (scanning from the next id on, so that ties do not all go to worker 0).
// The connection is booked on the chosen worker with a compare-and-swap against the load that
// made it the choice; if another acceptor got there first the scan is repeated, so two
// simultaneous decisions can never both land on the same "emptiest" worker.
Worker* Worker::least_loaded() {
    const std::vector<Worker*>& all = *peers_;
    const size_t n = all.size();
    while (true) {
        Worker* best = this;
        uint64_t best_load = stats_.connections_open.load(std::memory_order_relaxed);
        for (size_t k = 1; k < n; ++k) {
            Worker* w = all[(static_cast<size_t>(id_) + k) % n];
            const uint64_t load = w->stats_.connections_open.load(std::memory_order_relaxed);
            if (load < best_load) {
                best = w;
                best_load = load;
            }
        }
        if (best->stats_.connections_open.compare_exchange_strong(best_load, best_load + 1, std::memory_order_relaxed)) return best;
    }
}

void Worker::adopt(int fd) {
    Connection* c;
    if (!free_.empty()) {
        c = conns_[free_.back()].get();
        free_.pop_back();
    } else {
        conns_.emplace_back(new Connection(this, static_cast<uint32_t>(conns_.size())));
        c = conns_.back().get();
    }
    ++open_count_;
    stats_.connections_accepted.fetch_add(1, std::memory_order_relaxed);
    c->start(fd);
    if (draining_ && c->fd_ >= 0) {
        // Handed over by a peer that accepted it just before the shutdown began. The client has
        // a connection and may already have written its request: closing now would turn a clean
        // shutdown into a failed request. It gets the same deal as every other connection - its
        // request is answered (with "Connection: close", since draining() is true) - and is
        // closed after the linger period if it stays silent.
        const ConnRef ref = c->ref();
        loop_.add_timer(loop_.now_ns() + static_cast<int64_t>(std::max(1, options_.drain_linger_ms)) * 1000000, [this, ref] {
            Connection* late = lookup(ref);
            if (!late) return;
            if (!late->deferred_ && !late->output_pending()) {
                late->close_now();
            } else {
                late->close_after_write_ = true;
            }
        });
    }
}

void Worker::release(Connection* c) {
    // The slot becomes reusable only at the end of this loop iteration: the batch of epoll events
    // being dispatched may still hold an event for the descriptor that was just closed, and its
    // handler pointer is this object. Reused at once, the object would receive a stale EPOLLERR
    // or EPOLLIN that belongs to the previous connection (HttpClient::retire defers for the same
    // reason).
    parked_.push_back(c->index_);
    if (parked_.size() == 1) {
        loop_.defer([this] {
            free_.insert(free_.end(), parked_.begin(), parked_.end());
            parked_.clear();
        });
    }
    --open_count_;
    stats_.connections_open.fetch_sub(1, std::memory_order_relaxed);
    if (draining_ && open_count_ == 0) loop_.stop();
}

Connection* Worker::lookup(ConnRef ref) {
    if (ref.index >= conns_.size()) return nullptr;
    Connection* c = conns_[ref.index].get();
    if (c->fd_ < 0 || c->generation_ != ref.generation) return nullptr;
    return c;
}

bool Worker::complete(ConnRef ref, std::string_view response, bool close_after) {
    Connection* c = lookup(ref);
    if (!c || !c->deferred_) return false;
    c->deferred_ = false;
    c->out_.append(response);
    if (close_after || draining_) c->close_after_write_ = true;  // a half-closed peer: process_input() decides once the buffer is drained
    c->process_input();  // resumes parsing of pipelined requests, then flushes
    return true;
}

bool Worker::stream_write(ConnRef ref, std::string_view bytes) {
    Connection* c = lookup(ref);
    if (!c || !c->streaming_) return false;
    if (c->out_.size() - c->out_sent_ + c->sending_.size() - c->sending_sent_ > (4u << 20)) {  // reader is not keeping up
        c->close_now();
        return false;
    }
    c->out_.append(bytes);
    c->flush();
    return c->fd_ >= 0;
}

void Worker::sweep_idle() {
    const int64_t limit = static_cast<int64_t>(options_.idle_timeout_ms) * 1000000;
    const int64_t now = loop_.now_ns();
    for (auto& c : conns_) {
        if (c->fd_ < 0 || c->deferred_ || c->streaming_) continue;
        if (c->output_pending()) continue;
        if (now - c->last_active_ns_ > limit) c->close_now();
    }
    if (!draining_) {
        loop_.add_timer(now + limit / 2 + 1000000, [this] { sweep_idle(); });
    }
}

void Worker::begin_drain() {
    if (draining_) return;
    draining_ = true;
    if (listen_fd_ >= 0) {
        loop_.remove(listen_fd_);
        close(listen_fd_);
        listen_fd_ = -1;
    }
    app_->on_worker_stop(*this);
    for (auto& c : conns_) {
        if (c->fd_ >= 0 && c->streaming_) c->close_now();
    }
    if (open_count_ == 0) {
        loop_.stop();
        return;
    }
    // Keep-alive connections are not cut off at once: a client may be writing its next request
    // this very moment, and closing under it turns a clean shutdown into a failed request. For
    // a short linger period every request still gets its answer - marked "Connection: close",
    // because draining() is now true - and only connections that stayed silent are closed after
    // it. A busy client is therefore moved to another instance without seeing an error.
    auto close_idle = [this] {
        for (auto& c : conns_) {
            if (c->fd_ < 0) continue;
            if (!c->deferred_ && !c->output_pending()) {
                c->close_now();
            } else {
                c->close_after_write_ = true;
            }
        }
    };
    if (options_.drain_linger_ms <= 0) {
        close_idle();
        if (open_count_ == 0) return;  // release() has stopped the loop
    } else {
        loop_.add_timer(loop_.now_ns() + static_cast<int64_t>(options_.drain_linger_ms) * 1000000, close_idle);
    }
    loop_.add_timer(loop_.now_ns() + static_cast<int64_t>(options_.shutdown_grace_ms) * 1000000,
                    [this] { loop_.stop(); });
}

void Worker::thread_main() {
    if (!options_.cpu_affinity.empty()) {
        const int cpu = options_.cpu_affinity[static_cast<size_t>(id_) % options_.cpu_affinity.size()];
        cpu_set_t set;
        CPU_ZERO(&set);
        CPU_SET(cpu, &set);
        pthread_setaffinity_np(pthread_self(), sizeof set, &set);
    }
    char name[16];
    std::snprintf(name, sizeof name, "hola-io-%d", id_);
    pthread_setname_np(pthread_self(), name);
    loop_.refresh_clock();
    app_->on_worker_start(*this);
    if (listen_fd_ >= 0) loop_.add(listen_fd_, EPOLLIN, this);
    loop_.set_spin_ns(static_cast<int64_t>(options_.spin_us) * 1000);
    if (options_.idle_timeout_ms > 0) {
        loop_.add_timer(loop_.now_ns() + static_cast<int64_t>(options_.idle_timeout_ms) * 500000,
                        [this] { sweep_idle(); });
    }
    loop_.run();
    // Whatever is still open after the grace period goes down with the loop.
    for (auto& c : conns_) {
        if (c->fd_ >= 0) {
            close(c->fd_);
            c->fd_ = -1;
        }
        if (c->zombie_fd_ >= 0) {
            close(c->zombie_fd_);
            c->zombie_fd_ = -1;
        }
    }
    app_->on_worker_exit(*this);
    timespec ts;
    if (clock_gettime(CLOCK_THREAD_CPUTIME_ID, &ts) == 0)
        cpu_seconds_.store(static_cast<double>(ts.tv_sec) + static_cast<double>(ts.tv_nsec) / 1e9, std::memory_order_relaxed);
}

// ---------------------------------------------------------------------------------------------
// Server

Server::Server(ServerOptions options, Application* app) : options_(std::move(options)), app_(app) {}

Server::~Server() { stop(); }

namespace {

// How many event-loop threads make sense by default: the CPUs this process may run on, capped by
// the container's CPU quota (cgroup v2 cpu.max, or the v1 cfs files). A pod limited to one CPU on
// a 64-core node should get one loop, not 64 taking turns.
int default_thread_count() {
    cpu_set_t set;
    CPU_ZERO(&set);
    int n = sched_getaffinity(0, sizeof set, &set) == 0 ? CPU_COUNT(&set) : 1;
    if (n <= 0) n = 1;
    long quota = -1, period = -1;
    if (FILE* f = std::fopen("/sys/fs/cgroup/cpu.max", "r")) {
        char q[32] = "";
        if (std::fscanf(f, "%31s %ld", q, &period) == 2 && std::strcmp(q, "max") != 0) quota = std::atol(q);
        std::fclose(f);
    } else {
        if (FILE* fq = std::fopen("/sys/fs/cgroup/cpu/cpu.cfs_quota_us", "r")) {
            if (std::fscanf(fq, "%ld", &quota) != 1) quota = -1;
            std::fclose(fq);
        }
        if (FILE* fp = std::fopen("/sys/fs/cgroup/cpu/cpu.cfs_period_us", "r")) {
            if (std::fscanf(fp, "%ld", &period) != 1) period = -1;
            std::fclose(fp);
        }
    }
    if (quota > 0 && period > 0) {
        const int by_quota = static_cast<int>((quota + period - 1) / period);
        if (by_quota >= 1 && by_quota < n) n = by_quota;
    }
    return n;
}

}  // namespace

bool Server::start(std::string* error) {
    int n = options_.threads;
    if (n <= 0) n = default_thread_count();
    options_.threads = n;
    for (int i = 0; i < n; ++i) {
        workers_.emplace_back(new Worker(i, options_, app_));
        worker_ptrs_.push_back(workers_.back().get());
    }
    for (int i = 0; i < n; ++i) {
        Worker& w = *workers_[static_cast<size_t>(i)];
        if (options_.accept_mode == AcceptMode::kRoundRobin && i > 0) break;  // single acceptor
        if (!w.open_listener(error)) {
            workers_.clear();
            worker_ptrs_.clear();
            return false;
        }
        if (i == 0) {
            sockaddr_storage addr{};
            socklen_t len = sizeof addr;
            getsockname(w.listen_fd_, reinterpret_cast<sockaddr*>(&addr), &len);
            bound_port_ = addr.ss_family == AF_INET6 ? ntohs(reinterpret_cast<sockaddr_in6*>(&addr)->sin6_port)
                                                     : ntohs(reinterpret_cast<sockaddr_in*>(&addr)->sin_port);
            options_.port = bound_port_;  // later listeners join the same (possibly ephemeral) port
        }
    }
    if (options_.accept_mode == AcceptMode::kRoundRobin) {
        workers_[0]->peers_ = &worker_ptrs_;
    } else if (options_.accept_mode == AcceptMode::kBalanced) {
        for (auto& w : workers_) w->peers_ = &worker_ptrs_;
    }
    for (auto& w : workers_) {
        Worker* wp = w.get();
        wp->thread_ = std::thread([wp] { wp->thread_main(); });
    }
    started_ = true;
    if (options_.io_backend != IoBackend::kEpoll) {
        ring_thread_ = std::thread([this] { setup_rings(); });
    } else {
        rings_done_.store(true, std::memory_order_release);
    }
    return true;
}

void Server::stop() {
    if (!started_) return;
    started_ = false;
    stopping_rings_.store(true, std::memory_order_relaxed);
    if (ring_thread_.joinable()) ring_thread_.join();
    for (auto& w : workers_) {
        Worker* wp = w.get();
        wp->loop().post([wp] { wp->begin_drain(); });
    }
    for (auto& w : workers_) {
        if (w->thread_.joinable()) w->thread_.join();
    }
}

ServerStats Server::stats() const {
    ServerStats s;
    for (const auto& w : workers_) {
        s.connections_accepted += w->stats_.connections_accepted.load(std::memory_order_relaxed);
        s.connections_open += w->stats_.connections_open.load(std::memory_order_relaxed);
        s.requests += w->stats_.requests.load(std::memory_order_relaxed);
        s.bad_requests += w->stats_.bad_requests.load(std::memory_order_relaxed);
        if (!started_) {
            s.loop_iterations += w->loop_.stats().iterations;
            s.loop_events += w->loop_.stats().events;
            s.loop_blocking_waits += w->loop_.stats().blocking_waits;
        }
    }
    return s;
}

const char* Server::io_backend_name() const {
    if (options_.io_backend == IoBackend::kEpoll) return "epoll";
    size_t active = 0;
    for (const auto& w : workers_) active += w->uring_active_.load(std::memory_order_relaxed) ? 1 : 0;
    if (active == workers_.size()) return "uring";
    if (!rings_done_.load(std::memory_order_acquire)) return "uring (starting)";  // the first few milliseconds
    return active == 0 ? "epoll" : "mixed";
}

// Runs on its own short-lived thread: makes one ring per worker, one after the other, and hands
// each to its worker. The workers serve on epoll meanwhile.
void Server::setup_rings() {
    pthread_setname_np(pthread_self(), "hola-rings");
    for (auto& w : workers_) {
        if (stopping_rings_.load(std::memory_order_relaxed)) break;
        std::string why;
        std::shared_ptr<Uring> ring = EventLoop::prepare_uring(&why);
        if (!ring) {
            if (options_.io_backend == IoBackend::kUring)
                std::fprintf(stderr, "hola: io_uring is not usable here (%s) - staying with epoll\n", why.c_str());
            break;  // what one ring is refused, the next will be too
        }
        Worker* wp = w.get();
        wp->loop().post([wp, ring] {
            if (wp->loop().adopt_uring(ring)) wp->uring_active_.store(true, std::memory_order_relaxed);
        });
    }
    rings_done_.store(true, std::memory_order_release);
}

std::vector<WorkerSnapshot> Server::worker_stats() const {
    std::vector<WorkerSnapshot> out;
    for (const auto& w : workers_) {
        WorkerSnapshot s;
        s.id = w->id_;
        s.connections_accepted = w->stats_.connections_accepted.load(std::memory_order_relaxed);
        s.connections_open = w->stats_.connections_open.load(std::memory_order_relaxed);
        s.requests = w->stats_.requests.load(std::memory_order_relaxed);
        s.cpu_seconds = w->cpu_seconds_.load(std::memory_order_relaxed);
        if (started_ && w->thread_.joinable()) {
            clockid_t clock;
            timespec ts;
            if (pthread_getcpuclockid(w->thread_.native_handle(), &clock) == 0 && clock_gettime(clock, &ts) == 0)
                s.cpu_seconds = static_cast<double>(ts.tv_sec) + static_cast<double>(ts.tv_nsec) / 1e9;
        }
        out.push_back(s);
    }
    return out;
}

}  // namespace hola
