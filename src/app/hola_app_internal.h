// Private to the HolaApp translation units (hola_app.cc, hola_chain.cc, hola_stream.cc): the
// per-worker and per-call state, and the one place where response heads are rendered.
#pragma once

#include <memory>
#include <string>
#include <string_view>
#include <unordered_map>
#include <utility>
#include <vector>

#include "app/hola_app.h"

namespace hola {
namespace detail {

const char* reason_phrase(int status);

// Everything about the shape of a response head lives here, so that it can be corrected in one
// place if a live reference ever shows different Undertow/RESTEasy defaults.
// Returns the offset of the 29-byte date inside *out.
size_t append_response(std::string* out, int status, std::string_view content_type, std::string_view body,
                       bool cors, bool close, bool head_only, std::string_view date,
                       std::string_view extra_headers = {});

// The container's stock error page for non-JAX-RS failures.
std::string undertow_error_page(int status);

constexpr std::string_view kTextPlain = "text/plain";
constexpr std::string_view kTextHtml = "text/html;charset=UTF-8";
constexpr std::string_view kJson = "application/json";
constexpr std::string_view kAllowHeader = "Allow: HEAD, GET, OPTIONS\r\n";

inline uint64_t stream_key(ConnRef ref) { return (static_cast<uint64_t>(ref.index) << 32) | ref.generation; }

}  // namespace detail

struct HolaApp::StreamSub {
    ConnRef ref;
    int delay_ms = 500;
    bool chunked = true;  // false for HTTP/1.0 clients: raw events until the connection closes
    EventLoop::TimerId timer = 0;
};

// Reference to a pooled ChainCall (intrusive count; single-threaded: a call lives on one event
// loop). The last reference hands the object back to its worker's pool, where it keeps the
// capacity of its strings for the next call.
class HolaApp::CallRef {
 public:
    CallRef() = default;
    explicit CallRef(ChainCall* p);
    CallRef(const CallRef& o);
    CallRef(CallRef&& o) noexcept : p_(o.p_) { o.p_ = nullptr; }
    CallRef& operator=(CallRef o) noexcept {
        std::swap(p_, o.p_);
        return *this;
    }
    ~CallRef();
    ChainCall* operator->() const { return p_; }
    ChainCall& operator*() const { return *p_; }

 private:
    ChainCall* p_ = nullptr;
};

struct HolaApp::WorkerState {
    ~WorkerState();
    Worker* worker = nullptr;
    IdGenerator ids;
    SpanRing* ring = nullptr;
    std::unique_ptr<HttpClient> http;

    std::shared_ptr<const GreetingSnapshot> greeting;
    uint64_t greeting_version = 0;

    // Ready-to-send keep-alive responses; only the date is patched from second to second.
    std::string hola_bytes;
    size_t hola_date_at = 0;
    int64_t hola_second = -1;
    uint64_t hola_version = 0;
    std::string health_bytes;
    size_t health_date_at = 0;
    int64_t health_second = -1;

    // Verdicts for the most recently seen Accept header (see HolaApp::handle).
    std::string accept_header;
    bool accept_text_plain = true;
    bool accept_json = true;
    bool accept_yaml = true;

    std::unordered_map<uint64_t, StreamSub> streams;

    // GET /api/hola-chaining.
    std::vector<ChainCall*> call_pool;   // idle ChainCall objects (owned)
    std::string fallback_bytes;          // ready-to-send keep-alive response of the fallback outcome
    size_t fallback_date_at = 0;
    int64_t fallback_second = -1;
    uint64_t fallback_version = 0;
    std::string scratch;                 // response under construction
};

struct HolaApp::ChainCall {
    int refs = 0;
    WorkerState* owner = nullptr;
    std::shared_ptr<const GreetingSnapshot> greeting;  // element 0 of the answer is its text
    ConnRef ref;
    bool head_only = false;
    bool close = false;
    bool deferred = false;  // handle() has returned kDeferred
    bool settled = false;   // the downstream outcome has been decided
    bool finished = false;  // the response has been produced
    std::string response;   // filled when finishing before handle() returns
    TraceContext server_ctx;
    TraceContext client_ctx;
    std::string server_url;
    int64_t server_start_us = 0;
    int64_t client_start_us = 0;
    int64_t exec_start_ns = 0;
    HttpClient::RequestId request = 0;
    EventLoop::TimerId timer = 0;        // the command timeout
    EventLoop::TimerId retry_timer = 0;  // pause before the next attempt
    int attempt = 0;
    bool failed_without_fallback = false;

    // Back to the state of a fresh object, keeping the strings' storage.
    void recycle() {
        greeting.reset();
        ref = ConnRef{};
        head_only = close = deferred = settled = finished = failed_without_fallback = false;
        response.clear();
        server_ctx = TraceContext{};
        client_ctx = TraceContext{};
        server_url.clear();
        server_start_us = client_start_us = exec_start_ns = 0;
        request = 0;
        timer = retry_timer = 0;
        attempt = 0;
    }
};

inline HolaApp::CallRef::CallRef(ChainCall* p) : p_(p) {
    if (p_) ++p_->refs;
}
inline HolaApp::CallRef::CallRef(const CallRef& o) : p_(o.p_) {
    if (p_) ++p_->refs;
}
inline HolaApp::CallRef::~CallRef() {
    if (p_ && --p_->refs == 0) {
        p_->recycle();
        p_->owner->call_pool.push_back(p_);
    }
}
inline HolaApp::WorkerState::~WorkerState() {
    for (ChainCall* c : call_pool) delete c;
}

}  // namespace hola
