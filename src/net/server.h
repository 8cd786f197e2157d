// HTTP/1.1 server core: one Worker per thread, each with its own event loop, its own listening
// socket (SO_REUSEPORT) or a share of a round-robin acceptor, and its own pool of connections.
//
// Replaces the Undertow listener the reference gets from WildFly Swarm on :8080
// (/root/reference/Readme.md:16, /root/reference/Dockerfile:8, /root/reference/pom.xml:74).
//
// Connection placement (AcceptMode). Undertow's acceptor deals new connections to its I/O threads
// in turn (/root/reference/pom.xml:122-129 pulls it in), so a handful of keep-alive clients ends
// up spread evenly. Plain SO_REUSEPORT does not: the kernel picks the listener by a hash of the
// connection's addresses, and 64 connections over 16 listeners land anywhere between one and eight
// per worker - in a closed loop the worker with eight saturates while the worker with one idles.
// The default mode therefore keeps one listener per worker (accepting stays parallel, a second
// process can still join the port for a rolling restart) but lets the accepting worker pass a new
// connection to the least-loaded worker when it is not itself among the least loaded.
//
// Request life cycle on a worker thread:
//   recv -> parse_request_head (zero copy) -> Application::handle -> bytes appended to the
//   connection's output buffer -> one send() per batch of pipelined requests.
// A handler may instead *defer* (answer later from a timer or another socket's callback on the
// same loop - used for the downstream call) or turn the connection into a *stream* (used for
// the server-sent-events metrics feed). While a request is deferred the connection stops
// parsing, so pipelined responses always go out in request order.
#pragma once

#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <memory>
#include <mutex>
#include <string>
#include <string_view>
#include <thread>
#include <vector>

#include "core/http.h"
#include "net/event_loop.h"

namespace hola {

class Worker;
class Connection;

enum class Disposition : uint8_t {
    kDone,      // a complete response was appended to conn.out()
    kDeferred,  // the handler will call Worker::complete() later
    kStream,    // the handler took over the connection for streaming
};

// Stable handle to a connection that may have gone away in the meantime.
struct ConnRef {
    uint32_t index = 0;
    uint32_t generation = 0;
};

class Application {
 public:
    virtual ~Application() = default;
    // Called once on each worker thread before it starts serving.
    virtual void on_worker_start(Worker& /*worker*/) {}
    // Called once on each worker thread when it is shutting down.
    virtual void on_worker_stop(Worker& /*worker*/) {}
    // Called on each worker thread after its event loop has returned for good; the place to
    // destroy per-worker objects that are registered with that loop.
    virtual void on_worker_exit(Worker& /*worker*/) {}
    virtual Disposition handle(Worker& worker, Connection& conn, const RequestHead& request) = 0;
    // A deferred or streaming connection was closed by the peer (or by shutdown).
    virtual void on_connection_closed(Worker& /*worker*/, ConnRef /*ref*/) {}
};

class Connection final : public EventHandler, public UringHandler {
 public:
    explicit Connection(Worker* worker, uint32_t index) : worker_(worker), index_(index) {}
    ~Connection() override;

    void on_events(uint32_t events) override;                                                      // epoll backend
    void on_completion(uint32_t tag, int32_t res, uint32_t cqe_flags, const char* data) override;  // io_uring backend

    std::string& out() { return out_; }
    // The response being appended is the last one on this connection.
    void close_after_write() { close_after_write_ = true; }
    ConnRef ref() const { return ConnRef{index_, generation_}; }
    bool open() const { return fd_ >= 0; }
    // True once the response being built must be the last one (client asked for close, or the
    // server is shutting down); handlers then send "Connection: close".
    bool closing() const { return close_after_write_; }

 private:
    friend class Worker;
    void start(int fd);
    void close_now();
    void process_input();
    void flush();
    void update_interest();
    bool discard_body_bytes();
    void on_peer_closed();
    bool input_full() const;  // buffer at its limit with nothing consumable (a deferred request and a greedy pipeliner)
    bool output_pending() const {  // bytes owed to the peer that have not been handed to the kernel yet
        return out_sent_ < out_.size() || send_inflight_ || sending_sent_ < sending_.size();
    }
    // io_uring backend (see the comment on IoBackend).
    void uring_arm_recv();
    void uring_cancel_recv();
    void uring_submit_send();
    void uring_input(const char* data, size_t len);
    void uring_reap_zombie();

    Worker* worker_;
    const uint32_t index_;
    uint32_t generation_ = 0;
    int fd_ = -1;

    std::unique_ptr<char[]> in_;
    size_t in_cap_ = 0;
    size_t in_len_ = 0;
    size_t in_pos_ = 0;

    std::string out_;
    size_t out_sent_ = 0;

    bool want_write_ = false;
    bool close_after_write_ = false;
    bool deferred_ = false;
    bool streaming_ = false;
    bool peer_closed_ = false;
    bool in_process_ = false;
    int64_t discard_remaining_ = 0;
    std::unique_ptr<ChunkedDecoder> discard_chunked_;
    uint32_t registered_events_ = 0;
    int64_t last_active_ns_ = 0;

    // io_uring backend state.
    bool uring_ = false;
    int uring_slot_ = -1;           // handler slot in the loop, kept for the life of the object
    bool recv_armed_ = false;       // a multishot receive is in flight
    bool recv_cancelling_ = false;  // ... and has been asked to stop (back-pressure)
    bool send_inflight_ = false;
    bool zombie_ = false;           // closed, but completions of its last operations are still to come
    int ops_ = 0;                   // operations in flight (receive, send, cancel)
    int zombie_fd_ = -1;
    std::string sending_;           // the bytes of the send in flight (must not move until it completes)
    size_t sending_sent_ = 0;
};

enum class AcceptMode : uint8_t {
    kBalanced,    // a listener per worker + hand-off to the least-loaded worker (default)
    kReusePort,   // a listener per worker, the kernel's hash decides (no hand-off)
    kRoundRobin,  // worker 0 accepts and deals connections in turn; no SO_REUSEPORT
};

// How sockets are driven.
//   kEpoll  readiness: epoll_wait, then one recv and one send system call per request - the
//           pattern of the reference's container (Undertow/XNIO, /root/reference/pom.xml:122-129).
//   kUring  completion: one multishot receive per connection delivering into buffers the kernel
//           picks from a shared pool, sends queued as submissions, and a single io_uring_enter per
//           loop iteration that submits the batch and waits for the next completions. Listeners,
//           timers, the mailbox and the outbound client stay on epoll, which is polled through the
//           ring. Falls back to kEpoll when the kernel refuses io_uring. A closed connection keeps
//           its slot until the completions of its last operations have arrived. A worker starts
//           on epoll and switches when its ring arrives (a few ms after start-up; connections
//           accepted before that stay on epoll for their lifetime).
//   kAuto   kUring where the kernel offers it (Linux >= 6.1, not blocked), else kEpoll, silently.
//           The default: on the benchmark box kUring serves 5 % more requests per core and cuts
//           p99 latency by a quarter (profiles/r2_probe4/), and nothing is lost where it is absent.
enum class IoBackend : uint8_t { kAuto, kEpoll, kUring };

struct ServerOptions {
    std::string bind_address = "0.0.0.0";
    uint16_t port = 8080;
    int threads = 0;                  // 0: one per available CPU
    int backlog = 1024;
    size_t max_header_bytes = 65536;  // request line + headers
    AcceptMode accept_mode = AcceptMode::kBalanced;
    IoBackend io_backend = IoBackend::kAuto;
    std::vector<int> cpu_affinity;    // optional: worker i is pinned to cpu_affinity[i % size]
    int shutdown_grace_ms = 5000;     // in-flight work gets this long to finish after SIGTERM
    int drain_linger_ms = 100;        // idle keep-alive connections stay answerable this long (see begin_drain)
    int idle_timeout_ms = 0;          // 0: never close idle keep-alive connections
    int spin_us = 0;                  // busy-poll window after the last event (EventLoop::set_spin_ns)
};

// Counters are kept per worker (single writer, so no cache line is shared between cores on
// the request path) and summed on demand.
struct WorkerStats {
    std::atomic<uint64_t> connections_accepted{0};
    std::atomic<uint64_t> connections_open{0};
    std::atomic<uint64_t> requests{0};
    std::atomic<uint64_t> bad_requests{0};
};

// Snapshot of one worker (Server::worker_stats): what shows whether the load is spread evenly.
struct WorkerSnapshot {
    int id = 0;
    uint64_t connections_accepted = 0;  // connections this worker has served (adopted), not accept() calls
    uint64_t connections_open = 0;
    uint64_t requests = 0;
    double cpu_seconds = 0;             // CPU time of the worker's thread so far
};

struct ServerStats {
    uint64_t connections_accepted = 0;
    uint64_t connections_open = 0;
    uint64_t requests = 0;
    uint64_t bad_requests = 0;
    // Event-loop counters, valid once the workers have stopped (they are plain, unsynchronised).
    uint64_t loop_iterations = 0;
    uint64_t loop_events = 0;
    uint64_t loop_blocking_waits = 0;
};

class Worker final : public EventHandler {
 public:
    Worker(int id, const ServerOptions& options, Application* app);
    ~Worker() override;

    int id() const { return id_; }
    EventLoop& loop() { return loop_; }
    const ServerOptions& options() const { return options_; }
    bool draining() const { return draining_; }

    // 29-byte HTTP date for the current second (cached).
    std::string_view http_date();

    // Completes a deferred request: appends `response` and resumes the connection. Returns
    // false when the connection no longer exists (the response is dropped).
    bool complete(ConnRef ref, std::string_view response, bool close_after);
    // Writes to a streaming connection; false when it is gone (or had to be dropped because the
    // reader is not keeping up).
    bool stream_write(ConnRef ref, std::string_view bytes);
    Connection* lookup(ConnRef ref);

    // Per-worker slot for application state (owned by the application).
    void set_app_state(void* p) { app_state_ = p; }
    void* app_state() const { return app_state_; }

    // Listening-socket events.
    void on_events(uint32_t events) override;

 private:
    friend class Server;
    friend class Connection;
    bool open_listener(std::string* error);
    void adopt(int fd);  // loop thread; the caller has already counted the connection in connections_open
    Worker* least_loaded();
    void release(Connection* c);
    void begin_drain();
    void thread_main();
    void sweep_idle();

    const int id_;
    const ServerOptions& options_;
    Application* app_;
    alignas(64) WorkerStats stats_;
    EventLoop loop_;
    int listen_fd_ = -1;
    bool draining_ = false;
    std::vector<std::unique_ptr<Connection>> conns_;
    std::vector<uint32_t> free_;
    std::vector<uint32_t> parked_;  // closed in this loop iteration; joins free_ when the iteration ends
    size_t open_count_ = 0;
    std::vector<Worker*>* peers_ = nullptr;  // all workers, when this one may hand connections to others
    size_t rr_next_ = 0;
    int64_t date_second_ = -1;
    char date_[29];
    std::string canonical_path_;  // scratch for the request being dispatched
    void* app_state_ = nullptr;
    std::atomic<double> cpu_seconds_{0};  // filled in when the thread exits
    std::atomic<bool> uring_active_{false};
    std::thread thread_;
};

class Server {
 public:
    Server(ServerOptions options, Application* app);
    ~Server();

    // Binds the listeners and starts the worker threads. Returns false with *error set when a
    // socket cannot be opened.
    bool start(std::string* error);
    // Stops accepting, lets in-flight responses finish (bounded by shutdown_grace_ms), joins.
    void stop();
    uint16_t port() const { return bound_port_; }
    int threads() const { return static_cast<int>(workers_.size()); }
    ServerStats stats() const;
    std::vector<WorkerSnapshot> worker_stats() const;
    // "epoll", "uring", "uring (starting)" during the first milliseconds, or "mixed": what the
    // workers actually run on. Rings are made in the background after start() and handed to the
    // workers, which serve on epoll until theirs arrives (and for good when the kernel refuses).
    const char* io_backend_name() const;
    Worker& worker(int i) { return *workers_[static_cast<size_t>(i)]; }

 private:
    ServerOptions options_;
    Application* app_;
    std::vector<std::unique_ptr<Worker>> workers_;
    std::vector<Worker*> worker_ptrs_;
    uint16_t bound_port_ = 0;
    bool started_ = false;
    void setup_rings();
    std::thread ring_thread_;
    std::atomic<bool> stopping_rings_{false};
    std::atomic<bool> rings_done_{false};
};

// Small helper shared with the outbound client.
void set_nodelay(int fd);

}  // namespace hola
