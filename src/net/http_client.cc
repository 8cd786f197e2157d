#include "net/http_client.h"

#include <linux/io_uring.h>
#include <netinet/in.h>
#include <sys/socket.h>
#include <unistd.h>

#include <algorithm>
#include <cerrno>

#include "net/server.h"  // set_nodelay

namespace hola {

const char* ClientResult::describe() const {
    switch (code) {
        case Code::kOk: return "ok";
        case Code::kResolveFailed: return "host name did not resolve";
        case Code::kConnectFailed: return "connect failed";
        case Code::kIoError: return "connection lost";
        case Code::kProtocolError: return "malformed response";
        case Code::kTooLarge: return "response too large";
    }
    return "?";
}

struct HttpClient::Pending {
    RequestId id = 0;
    std::string host;
    uint16_t port = 0;
    std::string key;
    std::string request;
    Callback cb;
    Conn* conn = nullptr;
    int attempts = 0;
    bool cancelled = false;
};

// A connection starts on epoll (connect() completion is a readiness event). Once established,
// and if the loop runs an io_uring, it moves there: a multishot receive stays armed for the life
// of the connection and each request is one queued send - no system call of its own for either,
// where the epoll path costs a send(), the poll round trip and a recv() per downstream call.
// A retired connection with operations in flight lingers (closed for business, descriptor shut
// down) until their completions have arrived, then deletes itself.
class HttpClient::Conn final : public EventHandler, public UringHandler {
 public:
    enum class State : uint8_t { kConnecting, kSending, kReading, kIdle, kClosed };

    Conn(HttpClient* client, int fd, std::string key) : client_(client), fd_(fd), key_(std::move(key)) {}

    void begin(std::shared_ptr<Pending> p, bool reused) {
        req_ = std::move(p);
        req_->conn = this;
        reused_ = reused;
        got_bytes_ = false;
        sent_ = 0;
        in_.clear();
        body_.clear();
        content_type_.clear();
        head_done_ = false;
        status_ = 0;
        body_remaining_ = -1;
        chunked_mode_ = false;
        until_close_ = false;
        keep_alive_ = true;
        chunked_.reset();
        if (reused) {
            state_ = State::kSending;
            write_request();
        }
    }

    void on_events(uint32_t events) override {
        if (state_ == State::kClosed) return;
        if (state_ == State::kIdle) {
            // Anything happening on an idle connection means the peer is done with it.
            client_->drop_idle(this);
            return;
        }
        if (state_ == State::kConnecting) {
            int err = 0;
            socklen_t len = sizeof err;
            if ((events & (EPOLLERR | EPOLLHUP)) || getsockopt(fd_, SOL_SOCKET, SO_ERROR, &err, &len) != 0 || err != 0) {
                client_->retry_or_fail(this, ClientResult::Code::kConnectFailed);
                return;
            }
            state_ = State::kSending;
            if (client_->loop_->uring_enabled()) go_uring();
            write_request();
            return;
        }
        if (state_ == State::kSending && (events & EPOLLOUT)) {
            write_request();
            if (state_ != State::kReading) return;
        }
        if (events & (EPOLLIN | EPOLLHUP | EPOLLRDHUP | EPOLLERR)) read_response();
    }

 private:
    friend class HttpClient;

    // A keep-alive connection spends its life waiting for input; the interest set only changes
    // around connect() and when a write would block, so most requests need no epoll_ctl at all.
    void set_interest(uint32_t events) {
        if (uring_ || events == registered_) return;
        registered_ = events;
        client_->loop_->modify(fd_, events, this);
    }

    enum : uint32_t { kOpRecv = 1, kOpSend = 2 };

    void go_uring() {
        client_->loop_->remove(fd_);
        registered_ = 0;
        uring_ = true;
        slot_ = client_->loop_->register_uring_handler(this);
        arm_recv();
    }

    void arm_recv() {
        EventLoop& loop = *client_->loop_;
        io_uring_sqe* sqe = loop.uring_sqe();
        sqe->opcode = IORING_OP_RECV;
        sqe->fd = fd_;
        sqe->ioprio = IORING_RECV_MULTISHOT;
        sqe->flags = IOSQE_BUFFER_SELECT;
        sqe->buf_group = loop.uring_buffer_group();
        sqe->user_data = EventLoop::uring_user_data(slot_, kOpRecv);
        recv_armed_ = true;
        ++ops_;
    }

    void submit_send() {
        io_uring_sqe* sqe = client_->loop_->uring_sqe();
        sqe->opcode = IORING_OP_SEND;
        sqe->fd = fd_;
        sqe->addr = reinterpret_cast<uint64_t>(send_buf_.data() + sent_);
        sqe->len = static_cast<uint32_t>(send_buf_.size() - sent_);
        sqe->msg_flags = MSG_NOSIGNAL;
        sqe->user_data = EventLoop::uring_user_data(slot_, kOpSend);
        ++ops_;
    }

    // Last completion of a retired connection: now the descriptor and the object can go.
    void reap() {
        if (state_ != State::kClosed || ops_ > 0 || fd_ < 0) return;
        close(fd_);
        fd_ = -1;
        client_->loop_->unregister_uring_handler(slot_);
        HttpClient* client = client_;
        client->zombies_.erase(this);
        client->loop_->defer([this] { delete this; });
    }

    void on_completion(uint32_t tag, int32_t res, uint32_t cqe_flags, const char* data) override {
        if ((tag & 3u) == kOpRecv) {
            const bool more = (cqe_flags & IORING_CQE_F_MORE) != 0;
            if (!more) {
                recv_armed_ = false;
                --ops_;
            }
            if (state_ == State::kClosed) {
                reap();
                return;
            }
            if (res > 0 && data) {
                if (state_ == State::kIdle) {
                    client_->drop_idle(this);  // anything arriving on an idle connection means the peer is done with it
                    return;
                }
                got_bytes_ = true;
                if (state_ == State::kSending) state_ = State::kReading;  // the answer overtook the send's completion
                if (!consume(data, static_cast<size_t>(res))) return;
            } else if (res == 0) {
                if (state_ == State::kIdle) {
                    client_->drop_idle(this);
                } else if (head_done_ && until_close_) {
                    deliver(false);
                } else {
                    client_->retry_or_fail(this, ClientResult::Code::kIoError);
                }
                return;
            } else if (res != -ENOBUFS && res != -EINTR && res != -EAGAIN) {
                if (state_ == State::kIdle) {
                    client_->drop_idle(this);
                } else {
                    client_->retry_or_fail(this, ClientResult::Code::kIoError);
                }
                return;
            }
            if (!more && state_ != State::kClosed) arm_recv();
            return;
        }
        // kOpSend
        --ops_;
        if (state_ == State::kClosed) {
            reap();
            return;
        }
        if (res < 0 && res != -EINTR && res != -EAGAIN) {
            client_->retry_or_fail(this, ClientResult::Code::kIoError);
            return;
        }
        if (res > 0) sent_ += static_cast<size_t>(res);
        if (sent_ < send_buf_.size()) {
            submit_send();
        } else if (state_ == State::kSending) {
            state_ = State::kReading;
        }
    }

    void write_request() {
        if (uring_) {
            send_buf_.assign(req_->request);  // a copy: the request object may be cancelled while the send is in flight
            sent_ = 0;
            submit_send();
            return;
        }
        const std::string& r = req_->request;
        while (sent_ < r.size()) {
            const ssize_t n = ::send(fd_, r.data() + sent_, r.size() - sent_, MSG_NOSIGNAL);
            if (n > 0) {
                sent_ += static_cast<size_t>(n);
                continue;
            }
            if (n < 0 && errno == EINTR) continue;
            if (n < 0 && (errno == EAGAIN || errno == EWOULDBLOCK)) {
                set_interest(EPOLLOUT | EPOLLRDHUP);
                return;
            }
            client_->retry_or_fail(this, ClientResult::Code::kIoError);
            return;
        }
        state_ = State::kReading;
        set_interest(EPOLLIN | EPOLLRDHUP);
    }

    void read_response() {
        char buf[16384];
        while (state_ == State::kReading) {
            const ssize_t n = ::recv(fd_, buf, sizeof buf, 0);
            if (n > 0) {
                got_bytes_ = true;
                if (!consume(buf, static_cast<size_t>(n))) return;
                if (static_cast<size_t>(n) < sizeof buf) return;
                continue;
            }
            if (n == 0) {
                if (head_done_ && until_close_) {
                    deliver(false);
                } else {
                    client_->retry_or_fail(this, ClientResult::Code::kIoError);
                }
                return;
            }
            if (errno == EINTR) continue;
            if (errno == EAGAIN || errno == EWOULDBLOCK) return;
            client_->retry_or_fail(this, ClientResult::Code::kIoError);
            return;
        }
    }

    // Returns false once the connection has been handed back or torn down.
    bool consume(const char* data, size_t len) {
        if (!head_done_) {
            in_.append(data, len);
            while (true) {
                ResponseHead head;
                const ParseStatus st = parse_response_head(in_.data(), in_.size(), &head);
                if (st == ParseStatus::kIncomplete) {
                    if (in_.size() > 65536) {
                        client_->retry_or_fail(this, ClientResult::Code::kProtocolError);
                        return false;
                    }
                    return true;
                }
                if (st == ParseStatus::kBad) {
                    client_->retry_or_fail(this, ClientResult::Code::kProtocolError);
                    return false;
                }
                if (head.status >= 100 && head.status < 200) {  // interim response: skip it
                    in_.erase(0, head.head_len);
                    if (in_.empty()) return true;
                    continue;
                }
                head_done_ = true;
                status_ = head.status;
                content_type_.assign(head.content_type);
                keep_alive_ = head.keep_alive;
                if (response_has_no_body(Method::kGet, head.status)) {
                    body_remaining_ = 0;
                } else if (head.chunked) {
                    chunked_mode_ = true;
                } else if (head.content_length >= 0) {
                    body_remaining_ = head.content_length;
                } else {
                    until_close_ = true;
                }
                const std::string rest = in_.substr(head.head_len);
                in_.clear();
                return consume_body(rest.data(), rest.size());
            }
        }
        return consume_body(data, len);
    }

    bool consume_body(const char* data, size_t len) {
        if (chunked_mode_) {
            chunked_.feed(data, len, &body_);
            if (chunked_.failed()) {
                client_->retry_or_fail(this, ClientResult::Code::kProtocolError);
                return false;
            }
            if (body_.size() > client_->max_response_bytes_) return too_large();
            if (chunked_.done()) {
                deliver(keep_alive_);
                return false;
            }
            return true;
        }
        if (until_close_) {
            body_.append(data, len);
            if (body_.size() > client_->max_response_bytes_) return too_large();
            return true;
        }
        const size_t take = static_cast<size_t>(std::min<int64_t>(body_remaining_, static_cast<int64_t>(len)));
        body_.append(data, take);
        body_remaining_ -= static_cast<int64_t>(take);
        if (body_.size() > client_->max_response_bytes_) return too_large();
        if (body_remaining_ == 0) {
            deliver(keep_alive_ && take == len);  // stray bytes after the body: do not reuse
            return false;
        }
        return true;
    }

    bool too_large() {
        ClientResult r;
        r.code = ClientResult::Code::kTooLarge;
        client_->finish(this, std::move(r), false);
        return false;
    }

    void deliver(bool reusable) {
        ClientResult r;
        r.code = ClientResult::Code::kOk;
        r.status = status_;
        r.content_type = std::move(content_type_);
        r.body = std::move(body_);
        client_->finish(this, std::move(r), reusable);
    }

    HttpClient* client_;
    int fd_;
    std::string key_;
    State state_ = State::kConnecting;
    uint32_t registered_ = 0;
    std::shared_ptr<Pending> req_;
    bool reused_ = false;
    bool got_bytes_ = false;
    size_t sent_ = 0;
    // io_uring mode.
    bool uring_ = false;
    int slot_ = -1;
    int ops_ = 0;
    bool recv_armed_ = false;
    std::string send_buf_;
    std::string in_;
    bool head_done_ = false;
    int status_ = 0;
    std::string content_type_;
    bool keep_alive_ = true;
    bool chunked_mode_ = false;
    bool until_close_ = false;
    int64_t body_remaining_ = -1;
    ChunkedDecoder chunked_;
    std::string body_;
};

HttpClient::HttpClient(EventLoop* loop, Resolver* resolver, size_t max_idle_per_host, size_t max_response_bytes)
    : loop_(loop),
      resolver_(resolver),
      max_idle_per_host_(max_idle_per_host),
      max_response_bytes_(max_response_bytes),
      alive_(std::make_shared<bool>(true)) {}

HttpClient::~HttpClient() {
    *alive_ = false;
    for (Conn* c : conns_) {
        if (c->fd_ >= 0) {
            if (!c->uring_) loop_->remove(c->fd_);
            if (c->uring_) loop_->unregister_uring_handler(c->slot_);  // late completions find an empty slot
            close(c->fd_);
        }
        delete c;
    }
    for (Conn* c : zombies_) {
        loop_->unregister_uring_handler(c->slot_);
        if (c->fd_ >= 0) close(c->fd_);
        delete c;
    }
}

size_t HttpClient::idle_connections() const {
    size_t n = 0;
    for (const auto& kv : idle_) n += kv.second.size();
    return n;
}

HttpClient::Conn* HttpClient::take_idle(const std::string& key) {
    auto it = idle_.find(key);
    if (it == idle_.end() || it->second.empty()) return nullptr;
    Conn* c = it->second.back();
    it->second.pop_back();
    return c;
}

HttpClient::RequestId HttpClient::send(const std::string& host, uint16_t port, std::string request, Callback cb) {
    auto p = std::make_shared<Pending>();
    p->id = next_id_++;
    p->host = host;
    p->port = port;
    p->key = host + ":" + std::to_string(port);
    p->request = std::move(request);
    p->cb = std::move(cb);
    pending_[p->id] = p;
    const RequestId id = p->id;
    dispatch(p);
    return id;
}

void HttpClient::dispatch(const std::shared_ptr<Pending>& p) {
    ++p->attempts;
    if (p->attempts == 1) {
        if (Conn* idle = take_idle(p->key)) {
            idle->begin(p, true);
            return;
        }
    }
    std::shared_ptr<bool> alive = alive_;
    resolver_->resolve(p->host, loop_, [this, alive, p](bool ok, const IpAddress& address) {
        if (!*alive || p->cancelled) return;
        if (!ok) {
            fail(p, ClientResult::Code::kResolveFailed);
            return;
        }
        start_connect(p, address);
    });
}

void HttpClient::start_connect(const std::shared_ptr<Pending>& p, const IpAddress& address) {
    const int fd = socket(address.family == AF_INET6 ? AF_INET6 : AF_INET, SOCK_STREAM | SOCK_NONBLOCK | SOCK_CLOEXEC, 0);
    if (fd < 0) {
        fail(p, ClientResult::Code::kConnectFailed);
        return;
    }
    set_nodelay(fd);
    sockaddr_storage addr;
    const socklen_t addr_len = address.to_sockaddr(p->port, &addr);
    const int rc = connect(fd, reinterpret_cast<sockaddr*>(&addr), addr_len);
    if (rc != 0 && errno != EINPROGRESS) {
        close(fd);
        fail(p, ClientResult::Code::kConnectFailed);
        return;
    }
    ++connections_opened_;
    Conn* c = new Conn(this, fd, p->key);
    conns_.insert(c);
    c->begin(p, false);
    c->state_ = Conn::State::kConnecting;
    c->registered_ = EPOLLOUT | EPOLLRDHUP;
    if (!loop_->add(fd, c->registered_, c)) {
        retry_or_fail(c, ClientResult::Code::kConnectFailed);
    }
}

void HttpClient::fail(const std::shared_ptr<Pending>& p, ClientResult::Code code) {
    pending_.erase(p->id);
    if (p->cancelled) return;
    ClientResult r;
    r.code = code;
    Callback cb = std::move(p->cb);
    cb(std::move(r));
}

void HttpClient::retry_or_fail(Conn* conn, ClientResult::Code code) {
    std::shared_ptr<Pending> p = std::move(conn->req_);
    const bool stale_pooled = conn->reused_ && !conn->got_bytes_;
    retire(conn);
    if (!p) return;
    p->conn = nullptr;
    if (stale_pooled && p->attempts < 2 && !p->cancelled) {
        dispatch(p);  // the pooled connection had been closed under us: once more, fresh
        return;
    }
    fail(p, code);
}

void HttpClient::finish(Conn* conn, ClientResult&& result, bool reusable) {
    std::shared_ptr<Pending> p = std::move(conn->req_);
    if (p) p->conn = nullptr;
    bool pooled = false;
    if (reusable && result.code == ClientResult::Code::kOk) {
        auto& list = idle_[conn->key_];
        if (list.size() < max_idle_per_host_) {
            conn->state_ = Conn::State::kIdle;
            conn->set_interest(EPOLLIN | EPOLLRDHUP);
            list.push_back(conn);
            pooled = true;
        }
    }
    if (!pooled) retire(conn);
    if (!p) return;
    pending_.erase(p->id);
    if (p->cancelled) return;
    Callback cb = std::move(p->cb);
    cb(std::move(result));
}

void HttpClient::drop_idle(Conn* conn) {
    auto it = idle_.find(conn->key_);
    if (it != idle_.end()) {
        auto& list = it->second;
        list.erase(std::remove(list.begin(), list.end(), conn), list.end());
    }
    retire(conn);
}

void HttpClient::retire(Conn* conn) {
    if (conn->state_ == Conn::State::kClosed) return;
    conn->state_ = Conn::State::kClosed;
    conns_.erase(conn);
    if (conn->uring_ && conn->ops_ > 0) {
        // Completions are still to come and they carry this object's address: shut the socket
        // down so that they come now; the last one closes the descriptor and deletes the object.
        shutdown(conn->fd_, SHUT_RDWR);
        zombies_.insert(conn);
        return;
    }
    if (conn->fd_ >= 0) {
        if (conn->uring_) {
            loop_->unregister_uring_handler(conn->slot_);
        } else {
            loop_->remove(conn->fd_);
        }
        close(conn->fd_);
        conn->fd_ = -1;
    }
    // The current batch of epoll events may still point at this object.
    loop_->defer([conn] { delete conn; });
}

void HttpClient::cancel(RequestId id) {
    auto it = pending_.find(id);
    if (it == pending_.end()) return;
    std::shared_ptr<Pending> p = it->second;
    pending_.erase(it);
    p->cancelled = true;
    p->cb = nullptr;
    if (p->conn) {
        Conn* c = p->conn;
        p->conn = nullptr;
        c->req_.reset();
        retire(c);
    }
}

void HttpClient::close_idle() {
    std::vector<Conn*> all;
    for (auto& kv : idle_) {
        for (Conn* c : kv.second) all.push_back(c);
        kv.second.clear();
    }
    for (Conn* c : all) retire(c);
}

}  // namespace hola
