#include "app/hola_app.h"

#include <sys/stat.h>

#include <cstdio>
#include <cstring>

#include "app/static_content.h"
#include "core/text.h"

namespace hola {

namespace {

const char* reason_phrase(int status) {
    switch (status) {
        case 200: return "OK";
        case 400: return "Bad Request";
        case 404: return "Not Found";
        case 405: return "Method Not Allowed";
        case 406: return "Not Acceptable";
        case 500: return "Internal Server Error";
        case 503: return "Service Unavailable";
        default: return "Unknown";
    }
}

// Everything about the shape of a response head lives here, so that it can be corrected in one
// place if a live reference ever shows different Undertow/RESTEasy defaults.
// Returns the offset of the 29-byte date inside *out.
size_t append_response(std::string* out, int status, std::string_view content_type, std::string_view body,
                       bool cors, bool close, bool head_only, std::string_view date,
                       std::string_view extra_headers = {}) {
    char num[24];
    out->append("HTTP/1.1 ");
    std::snprintf(num, sizeof num, "%d ", status);
    out->append(num);
    out->append(reason_phrase(status));
    out->append(close ? "\r\nConnection: close\r\n" : "\r\nConnection: keep-alive\r\n");
    out->append(extra_headers);
    if (cors) out->append("Access-Control-Allow-Origin: *\r\n");
    if (!content_type.empty()) {
        out->append("Content-Type: ");
        out->append(content_type);
        out->append("\r\n");
    }
    std::snprintf(num, sizeof num, "%zu", body.size());
    out->append("Content-Length: ");
    out->append(num);
    out->append("\r\nDate: ");
    const size_t date_at = out->size();
    out->append(date);
    out->append("\r\n\r\n");
    if (!head_only) out->append(body);
    return date_at;
}

constexpr std::string_view kTextPlain = "text/plain";
constexpr std::string_view kTextHtml = "text/html;charset=UTF-8";
constexpr std::string_view kJson = "application/json";
constexpr std::string_view kAllowHeader = "Allow: HEAD, GET, OPTIONS\r\n";

std::string undertow_error_page(int status) {
    return std::string("<html><head><title>Error</title></head><body>") + reason_phrase(status) + "</body></html>";
}

uint64_t stream_key(ConnRef ref) { return (static_cast<uint64_t>(ref.index) << 32) | ref.generation; }

// One JAX-RS resource method as far as dispatch is concerned.
enum class ApiRoute { kNone, kHola, kChaining, kHealth, kSwaggerJson, kSwaggerYaml };

ApiRoute match_api(std::string_view sub) {
    if (sub.size() > 1 && sub.back() == '/') sub.remove_suffix(1);  // JAX-RS ignores one trailing slash
    if (sub == "/hola") return ApiRoute::kHola;
    if (sub == "/health") return ApiRoute::kHealth;
    if (sub == "/hola-chaining") return ApiRoute::kChaining;
    if (sub == "/swagger.json") return ApiRoute::kSwaggerJson;
    if (sub == "/swagger.yaml") return ApiRoute::kSwaggerYaml;
    return ApiRoute::kNone;
}

void append_b3_id(std::string* out, const char* name, uint64_t id) {
    char buf[16];
    format_trace_id(id, buf);
    out->append(name);
    out->append(": ");
    out->append(buf, 16);
    out->append("\r\n");
}

}  // namespace

struct HolaApp::StreamSub {
    ConnRef ref;
    int delay_ms = 500;
    EventLoop::TimerId timer = 0;
};

struct HolaApp::WorkerState {
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

    std::unordered_map<uint64_t, StreamSub> streams;
};

struct HolaApp::ChainCall {
    ConnRef ref;
    bool head_only = false;
    bool close = false;
    bool deferred = false;  // handle() has returned kDeferred
    bool settled = false;   // the downstream outcome has been decided
    bool finished = false;  // the response has been produced
    std::string response;   // filled when finishing before handle() returns
    std::string hola_text;
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
};

HolaApp::HolaApp(AppOptions options, GreetingService* greeting, Tracer* tracer, Resolver* resolver)
    : options_(std::move(options)),
      greeting_(greeting),
      tracer_(tracer),
      resolver_(resolver),
      // Key names as HystrixFeign derives them: command = Feign.configKey(type, method),
      // group and thread pool = the target's name, which for a hard-coded target is its URL.
      command_("AlohaService#aloha()", "http://" + options_.aloha_host + ":" + std::to_string(options_.aloha_port) + "/",
               "http://" + options_.aloha_host + ":" + std::to_string(options_.aloha_port) + "/", options_.breaker) {
    const std::string authority = options_.aloha_host + ":" + std::to_string(options_.aloha_port);
    aloha_url_ = "http://" + authority + options_.aloha_path;
    aloha_request_prefix_ = "GET " + options_.aloha_path + " HTTP/1.1\r\nHost: " + authority +
                            "\r\nAccept: */*\r\nConnection: keep-alive\r\nUser-Agent: hola-b200\r\n";
}

HolaApp::~HolaApp() {
    // The workers (and their event loops) are gone by now; see Service::stop().
    for (WorkerState* st : states_) delete st;
}

void HolaApp::on_worker_start(Worker& worker) {
    auto* st = new WorkerState();
    {
        std::lock_guard<std::mutex> lock(states_mu_);
        states_.push_back(st);
    }
    st->worker = &worker;
    if (tracer_->enabled()) st->ring = tracer_->register_producer();
    st->http.reset(new HttpClient(&worker.loop(), resolver_));
    worker.set_app_state(st);
}

void HolaApp::on_worker_stop(Worker& worker) {
    WorkerState& st = state_of(worker);
    st.http->close_idle();
}

void HolaApp::on_worker_exit(Worker& worker) {
    WorkerState& st = state_of(worker);
    for (auto& kv : st.streams) {
        if (kv.second.timer) worker.loop().cancel_timer(kv.second.timer);
    }
    st.streams.clear();
    st.http.reset();  // unregisters its sockets from this worker's loop, which is still alive here
}

const GreetingSnapshot& HolaApp::current_greeting(WorkerState& st) {
    greeting_->poll(st.worker->loop().now_ns());
    const uint64_t v = greeting_->version();
    if (v != st.greeting_version || !st.greeting) {
        st.greeting = greeting_->snapshot();
        st.greeting_version = st.greeting->version;
    }
    return *st.greeting;
}

void HolaApp::record_span(WorkerState& st, const TraceContext& ctx, bool client, int64_t start_us, int64_t end_us,
                          int status, std::string_view url_a, std::string_view url_b, std::string_view url_c) {
    if (!st.ring || !ctx.sampled) return;
    SpanRecord* rec = st.ring->begin_push();
    if (!rec) return;  // collector is not keeping up: drop, never block
    rec->trace_id = ctx.trace_id;
    rec->span_id = ctx.span_id;
    rec->parent_id = ctx.parent_id;
    rec->has_parent = ctx.has_parent;
    rec->client = client;
    rec->shared = ctx.joined;
    rec->start_us = start_us;
    rec->end_us = end_us;
    rec->status = static_cast<uint16_t>(status);
    size_t n = 0;
    for (std::string_view part : {url_a, url_b, url_c}) {
        const size_t take = std::min(part.size(), sizeof rec->url - n);
        if (take == 0) continue;  // an empty view may carry a null pointer
        std::memcpy(rec->url + n, part.data(), take);
        n += take;
    }
    rec->url_len = static_cast<uint16_t>(n);
    st.ring->commit_push();
}

Disposition HolaApp::handle(Worker& worker, Connection& conn, const RequestHead& rq) {
    WorkerState& st = state_of(worker);
    const std::string_view path = rq.path;
    std::string& out = conn.out();

    // The JAX-RS servlet is mapped to /api/* (JaxRsActivator.java:30), which also covers "/api".
    if ((path.size() >= 5 && std::memcmp(path.data(), "/api/", 5) == 0) || path == "/api") {
        const ApiRoute route = match_api(path.substr(4));
        if (route == ApiRoute::kNone) {
            std::string body = "RESTEASY003210: Could not find resource for full path: http://";
            body.append(rq.host.empty() ? std::string_view("localhost") : rq.host);
            body.append(rq.target);
            append_response(&out, 404, kTextHtml, body, true, conn.closing(), rq.method == Method::kHead, worker.http_date());
            return Disposition::kDone;
        }
        if (rq.method == Method::kOptions) {
            append_response(&out, 200, kTextPlain, "HEAD, GET, OPTIONS", true, conn.closing(), false, worker.http_date(),
                            kAllowHeader);
            return Disposition::kDone;
        }
        if (rq.method != Method::kGet && rq.method != Method::kHead) {
            append_response(&out, 405, {}, {}, true, conn.closing(), false, worker.http_date(), kAllowHeader);
            return Disposition::kDone;
        }
        if (rq.has_accept) {
            bool ok = true;
            switch (route) {
                case ApiRoute::kHola:
                case ApiRoute::kHealth:
                    ok = accept_allows(rq.accept, "text", "plain");
                    break;
                case ApiRoute::kChaining:
                case ApiRoute::kSwaggerJson:
                    ok = accept_allows(rq.accept, "application", "json");
                    break;
                case ApiRoute::kSwaggerYaml:
                    ok = accept_allows(rq.accept, "application", "yaml");
                    break;
                case ApiRoute::kNone:
                    break;
            }
            if (!ok) {
                append_response(&out, 406, kTextHtml, "RESTEASY003635: No match for accept header", true, conn.closing(),
                                rq.method == Method::kHead, worker.http_date());
                return Disposition::kDone;
            }
        }
        switch (route) {
            case ApiRoute::kHola:
                return serve_hola(st, conn, rq);
            case ApiRoute::kChaining:
                return serve_chaining(st, conn, rq);
            case ApiRoute::kHealth: {
                if (rq.method == Method::kGet && !conn.closing()) {
                    const int64_t second = worker.loop().now_unix_s();
                    if (st.health_bytes.empty()) {
                        st.health_date_at = append_response(&st.health_bytes, 200, kTextPlain, "I'm ok", true, false, false,
                                                            worker.http_date());
                        st.health_second = second;
                    } else if (second != st.health_second) {
                        std::memcpy(&st.health_bytes[st.health_date_at], worker.http_date().data(), 29);
                        st.health_second = second;
                    }
                    out.append(st.health_bytes);
                } else {
                    append_response(&out, 200, kTextPlain, "I'm ok", true, conn.closing(), rq.method == Method::kHead,
                                    worker.http_date());
                }
                return Disposition::kDone;
            }
            case ApiRoute::kSwaggerJson:
                append_response(&out, 200, kJson, swagger_json(), true, conn.closing(), rq.method == Method::kHead,
                                worker.http_date());
                return Disposition::kDone;
            case ApiRoute::kSwaggerYaml:
                append_response(&out, 200, "application/yaml", swagger_yaml(), true, conn.closing(),
                                rq.method == Method::kHead, worker.http_date());
                return Disposition::kDone;
            case ApiRoute::kNone:
                break;
        }
    }

    if (path == "/hystrix.stream") {
        if (rq.method != Method::kGet) {
            append_response(&out, 405, {}, {}, false, conn.closing(), false, worker.http_date(), "Allow: GET\r\n");
            return Disposition::kDone;
        }
        return serve_stream(st, conn, rq);
    }

    if (path == "/" || path == "/index.html") {
        if (rq.method != Method::kGet && rq.method != Method::kHead) {
            append_response(&out, 405, {}, {}, false, conn.closing(), false, worker.http_date(), "Allow: GET, HEAD\r\n");
            return Disposition::kDone;
        }
        append_response(&out, 200, "text/html", index_html(), false, conn.closing(), rq.method == Method::kHead,
                        worker.http_date());
        return Disposition::kDone;
    }

    if (path.size() > 9 && std::memcmp(path.data(), "/webjars/", 9) == 0 && !options_.webjars_dir.empty() &&
        (rq.method == Method::kGet || rq.method == Method::kHead)) {
        serve_static_file(st, conn, rq);
        return Disposition::kDone;
    }

    append_response(&out, 404, "text/html", undertow_error_page(404), false, conn.closing(), rq.method == Method::kHead,
                    worker.http_date());
    return Disposition::kDone;
}

Disposition HolaApp::serve_hola(WorkerState& st, Connection& conn, const RequestHead& rq) {
    Worker& worker = *st.worker;
    const GreetingSnapshot& g = current_greeting(st);
    // The servlet filter is mapped to the exact URL, so "/api/hola/" is served but not traced.
    const bool traced = st.ring != nullptr && rq.path.size() == 9;
    TraceContext ctx;
    if (traced) ctx = start_server_trace(rq.b3, &st.ids);
    int status = 200;
    if (!g.ok) {
        status = 500;
        if (options_.verbose) std::fprintf(stderr, "hola: GET /api/hola failed: %s\n", g.error.c_str());
        append_response(&conn.out(), 500, "text/html", undertow_error_page(500), false, conn.closing(),
                        rq.method == Method::kHead, worker.http_date());
    } else if (rq.method == Method::kGet && !conn.closing()) {
        const int64_t second = worker.loop().now_unix_s();
        if (st.hola_version != g.version || st.hola_bytes.empty()) {
            st.hola_bytes.clear();
            st.hola_date_at = append_response(&st.hola_bytes, 200, kTextPlain, g.text, true, false, false, worker.http_date());
            st.hola_version = g.version;
            st.hola_second = second;
        } else if (second != st.hola_second) {
            std::memcpy(&st.hola_bytes[st.hola_date_at], worker.http_date().data(), 29);
            st.hola_second = second;
        }
        conn.out().append(st.hola_bytes);
    } else {
        append_response(&conn.out(), 200, kTextPlain, g.text, true, conn.closing(), rq.method == Method::kHead,
                        worker.http_date());
    }
    if (ctx.sampled) {
        record_span(st, ctx, false, worker.loop().now_unix_us(), realtime_us(), status, "http://",
                    rq.host.empty() ? std::string_view("localhost") : rq.host, rq.target);
    }
    return Disposition::kDone;
}

Disposition HolaApp::serve_chaining(WorkerState& st, Connection& conn, const RequestHead& rq) {
    Worker& worker = *st.worker;
    const GreetingSnapshot& g = current_greeting(st);
    const bool traced = st.ring != nullptr && rq.path.size() == 18;  // exactly "/api/hola-chaining"
    auto call = std::make_shared<ChainCall>();
    call->ref = conn.ref();
    call->head_only = rq.method == Method::kHead;
    call->close = conn.closing();
    call->server_start_us = worker.loop().now_unix_us();
    if (traced) {
        call->server_ctx = start_server_trace(rq.b3, &st.ids);
        call->server_url = "http://";
        call->server_url.append(rq.host.empty() ? std::string_view("localhost") : rq.host);
        call->server_url.append(rq.target);
    }
    if (!g.ok) {
        // hola() throws before the downstream call is attempted.
        append_response(&conn.out(), 500, "text/html", undertow_error_page(500), false, conn.closing(), call->head_only,
                        worker.http_date());
        record_span(st, call->server_ctx, false, call->server_start_us, realtime_us(), 500, call->server_url, {}, {});
        return Disposition::kDone;
    }
    call->hola_text = g.text;
    start_chain(st, call);
    if (call->finished) {
        conn.out().append(call->response);
        return Disposition::kDone;
    }
    call->deferred = true;
    return Disposition::kDeferred;
}

// What a failed, timed-out, rejected or short-circuited call contributes to the response: the
// static fallback list (HolaResource.java:112), or nothing when fallbacks are disabled.
JsonStringList HolaApp::fallback_tail(ChainCall& call) const {
    if (options_.breaker.fallback_enabled) return JsonStringList{options_.fallback};
    call.failed_without_fallback = true;
    return {};
}

void HolaApp::start_chain(WorkerState& st, const std::shared_ptr<ChainCall>& call) {
    EventLoop& loop = st.worker->loop();
    const int64_t now_ms = loop.now_ns() / 1000000;
    const HystrixCommand::Admission adm = command_.admit(now_ms);
    if (adm != HystrixCommand::Admission::kRun) {
        if (options_.verbose) {
            std::fprintf(stderr, "hola: aloha call %s -> fallback\n",
                         adm == HystrixCommand::Admission::kShortCircuited ? "short-circuited" : "rejected (pool full)");
        }
        call->settled = true;
        finish_chain(st, call, fallback_tail(*call));
        return;
    }
    call->exec_start_ns = loop.now_ns();

    WorkerState* stp = &st;
    if (options_.breaker.timeout_enabled) {
        call->timer = loop.add_timer(
            loop.now_ns() + static_cast<int64_t>(options_.breaker.timeout_ms) * 1000000, [this, stp, call] {
                if (call->settled) return;
                call->settled = true;
                call->timer = 0;
                EventLoop& lp = stp->worker->loop();
                if (call->retry_timer) {
                    lp.cancel_timer(call->retry_timer);
                    call->retry_timer = 0;
                }
                stp->http->cancel(call->request);
                command_.complete_timeout(lp.now_ns() / 1000000,
                                          static_cast<int>((lp.now_ns() - call->exec_start_ns) / 1000000));
                if (options_.verbose) std::fprintf(stderr, "hola: aloha call timed out -> fallback\n");
                finish_chain(*stp, call, fallback_tail(*call));
            });
    }
    send_chain_attempt(st, call);
}

// One HTTP attempt of the downstream call. Connection-level failures are retried the way Feign's
// default Retryer does (the reference builds its client with HystrixFeign.builder() and does not
// replace the retryer, HolaResource.java:104-112): up to feign_max_attempts tries, sleeping
// 100 ms x 1.5^k (capped at 1 s) between them - all inside the one Hystrix command, so the
// 600 ms command timeout is what usually ends a series of failing attempts. HTTP error statuses
// and undecodable bodies are not retried.
void HolaApp::send_chain_attempt(WorkerState& st, const std::shared_ptr<ChainCall>& call) {
    ++call->attempt;
    call->client_ctx = start_client_trace(call->server_ctx, &st.ids);  // a fresh client span per attempt
    std::string request = aloha_request_prefix_;
    if (st.ring) {
        if (call->client_ctx.sampled) {
            append_b3_id(&request, "X-B3-TraceId", call->client_ctx.trace_id);
            append_b3_id(&request, "X-B3-SpanId", call->client_ctx.span_id);
            append_b3_id(&request, "X-B3-ParentSpanId", call->client_ctx.parent_id);
            request.append("X-B3-Sampled: 1\r\n");
        } else {
            request.append("X-B3-Sampled: 0\r\n");
        }
    }
    request.append("\r\n");
    call->client_start_us = realtime_us();
    WorkerState* stp = &st;
    if (options_.verbose) std::fprintf(stderr, "[AlohaService#aloha] ---> GET %s HTTP/1.1\n", aloha_url_.c_str());
    call->request = st.http->send(
        options_.aloha_host, options_.aloha_port, std::move(request), [this, stp, call](ClientResult&& r) {
            if (call->settled) return;
            EventLoop& lp = stp->worker->loop();
            const int64_t now_ns = monotonic_ns();
            const int exec_ms = static_cast<int>((now_ns - call->exec_start_ns) / 1000000);
            const bool io_failure = r.code == ClientResult::Code::kResolveFailed || r.code == ClientResult::Code::kConnectFailed ||
                                    r.code == ClientResult::Code::kIoError;
            if (io_failure && call->attempt < options_.feign_max_attempts) {
                double interval_ms = options_.feign_retry_period_ms;
                for (int k = 0; k < call->attempt; ++k) interval_ms *= 1.5;
                if (interval_ms > options_.feign_retry_max_period_ms) interval_ms = options_.feign_retry_max_period_ms;
                if (options_.verbose) {
                    std::fprintf(stderr, "[AlohaService#aloha] <--- ERROR %s (%dms)\n[AlohaService#aloha] ---> RETRYING in %.0f ms\n",
                                 r.describe(), exec_ms, interval_ms);
                }
                call->retry_timer = lp.add_timer(now_ns + static_cast<int64_t>(interval_ms * 1e6), [this, stp, call] {
                    call->retry_timer = 0;
                    if (!call->settled) send_chain_attempt(*stp, call);
                });
                return;
            }
            call->settled = true;
            if (call->timer) {
                lp.cancel_timer(call->timer);
                call->timer = 0;
            }
            JsonStringList tail;
            bool ok = false;
            std::string why;
            if (r.code != ClientResult::Code::kOk) {
                why = r.describe();
            } else {
                record_span(*stp, call->client_ctx, true, call->client_start_us, realtime_us(), r.status, aloha_url_, {}, {});
                if (r.status < 200 || r.status > 299) {
                    why = "status " + std::to_string(r.status);
                } else if (!json_decode_string_list(r.body, &tail, &why)) {
                    tail.clear();
                } else {
                    ok = true;
                }
            }
            if (options_.verbose) {
                if (r.code == ClientResult::Code::kOk) {
                    std::fprintf(stderr, "[AlohaService#aloha] <--- HTTP/1.1 %d (%dms)\n", r.status, exec_ms);
                } else {
                    std::fprintf(stderr, "[AlohaService#aloha] <--- ERROR %s (%dms)\n", why.c_str(), exec_ms);
                }
            }
            if (ok) {
                command_.complete_success(now_ns / 1000000, exec_ms);
                finish_chain(*stp, call, std::move(tail));
            } else {
                command_.complete_failure(now_ns / 1000000, exec_ms);
                finish_chain(*stp, call, fallback_tail(*call));
            }
        });
}

void HolaApp::finish_chain(WorkerState& st, const std::shared_ptr<ChainCall>& call, JsonStringList&& tail) {
    if (call->finished) return;
    call->finished = true;
    Worker& worker = *st.worker;
    const bool close = call->close || worker.draining();
    std::string response;
    int status = 200;
    // An empty tail is the marker for "the call failed and fallbacks are switched off"
    // (hystrix.command.default.fallback.enabled=false): Hystrix then throws, the resource method
    // fails and the client sees a 500.
    if (tail.empty() && call->failed_without_fallback) {
        status = 500;
        append_response(&response, 500, "text/html", undertow_error_page(500), false, close, call->head_only, worker.http_date());
    } else {
        JsonStringList all;
        all.reserve(tail.size() + 1);
        all.emplace_back(std::move(call->hola_text));
        for (auto& item : tail) all.emplace_back(std::move(item));
        append_response(&response, 200, kJson, json_encode_string_list(all), true, close, call->head_only, worker.http_date());
    }
    record_span(st, call->server_ctx, false, call->server_start_us, realtime_us(), status, call->server_url, {}, {});
    if (call->deferred) {
        worker.complete(call->ref, response, close);
    } else {
        call->response = std::move(response);
    }
}

Disposition HolaApp::serve_stream(WorkerState& st, Connection& conn, const RequestHead& rq) {
    Worker& worker = *st.worker;
    if (stream_connections_.fetch_add(1, std::memory_order_acq_rel) >= options_.stream_max_connections) {
        stream_connections_.fetch_sub(1, std::memory_order_acq_rel);
        const std::string body = "MaxConcurrentConnections reached: " + std::to_string(options_.stream_max_connections);
        append_response(&conn.out(), 503, "text/html", body, false, conn.closing(), false, worker.http_date());
        return Disposition::kDone;
    }
    int delay_ms = options_.stream_default_delay_ms;
    // ?delay=<ms>, as the Hystrix servlet accepts.
    size_t pos = 0;
    const std::string_view q = rq.query;
    while (pos < q.size()) {
        size_t amp = q.find('&', pos);
        if (amp == std::string_view::npos) amp = q.size();
        const std::string_view kv = q.substr(pos, amp - pos);
        if (kv.size() > 6 && kv.substr(0, 6) == "delay=") {
            int v = 0;
            bool numeric = true;
            for (char c : kv.substr(6)) {
                if (c < '0' || c > '9' || v > 1000000) {
                    numeric = false;
                    break;
                }
                v = v * 10 + (c - '0');
            }
            if (numeric && v > 0) delay_ms = v;
        }
        pos = amp + 1;
    }
    std::string& out = conn.out();
    out.append("HTTP/1.1 200 OK\r\nConnection: keep-alive\r\n"
               "Cache-Control: no-cache, no-store, max-age=0, must-revalidate\r\nPragma: no-cache\r\n"
               "Content-Type: text/event-stream;charset=UTF-8\r\nTransfer-Encoding: chunked\r\nDate: ");
    out.append(worker.http_date());
    out.append("\r\n\r\n");
    StreamSub sub;
    sub.ref = conn.ref();
    sub.delay_ms = delay_ms;
    const uint64_t key = stream_key(sub.ref);
    WorkerState* stp = &st;
    sub.timer = worker.loop().add_timer(worker.loop().now_ns() + static_cast<int64_t>(delay_ms) * 1000000,
                                        [this, stp, key] { stream_tick(*stp, key); });
    st.streams[key] = sub;
    return Disposition::kStream;
}

void HolaApp::stream_tick(WorkerState& st, uint64_t key) {
    auto it = st.streams.find(key);
    if (it == st.streams.end()) return;
    Worker& worker = *st.worker;
    std::string event;
    if (command_.has_executed()) {
        const int64_t mono_ms = worker.loop().now_ns() / 1000000;
        const int64_t wall_ms = worker.loop().now_unix_us() / 1000;
        CommandSnapshot cs = command_.snapshot(mono_ms);
        cs.now_ms = wall_ms;
        PoolSnapshot ps = command_.pool_snapshot(mono_ms);
        ps.now_ms = wall_ms;
        event.append("data: ");
        event.append(hystrix_command_json(command_, cs));
        event.append("\n\ndata: ");
        event.append(hystrix_pool_json(command_, ps));
        event.append("\n\n");
    } else {
        event = "ping: \n\n";
    }
    char size_line[24];
    std::snprintf(size_line, sizeof size_line, "%zx\r\n", event.size());
    std::string chunk = size_line;
    chunk.append(event);
    chunk.append("\r\n");
    const ConnRef ref = it->second.ref;
    const int delay_ms = it->second.delay_ms;
    if (!worker.stream_write(ref, chunk)) return;  // closed; on_connection_closed() cleaned up
    it = st.streams.find(key);
    if (it == st.streams.end()) return;
    WorkerState* stp = &st;
    it->second.timer = worker.loop().add_timer(worker.loop().now_ns() + static_cast<int64_t>(delay_ms) * 1000000,
                                               [this, stp, key] { stream_tick(*stp, key); });
}

void HolaApp::on_connection_closed(Worker& worker, ConnRef ref) {
    WorkerState& st = state_of(worker);
    auto it = st.streams.find(stream_key(ref));
    if (it == st.streams.end()) return;  // a deferred chaining call: it finishes on its own
    if (it->second.timer) worker.loop().cancel_timer(it->second.timer);
    st.streams.erase(it);
    stream_connections_.fetch_sub(1, std::memory_order_acq_rel);
}

void HolaApp::serve_static_file(WorkerState& st, Connection& conn, const RequestHead& rq) {
    Worker& worker = *st.worker;
    const bool head_only = rq.method == Method::kHead;
    std::string rel;
    bool ok = percent_decode(rq.path.substr(8), &rel);  // keeps the leading '/'
    if (ok) {
        // No escaping the directory, no hidden files, no NULs.
        if (rel.find('\0') != std::string::npos || rel.find("/.") != std::string::npos || rel.find('\\') != std::string::npos)
            ok = false;
    }
    std::string full;
    if (ok) {
        full = options_.webjars_dir + rel;
        struct stat sb;
        if (stat(full.c_str(), &sb) != 0) {
            ok = false;
        } else if (S_ISDIR(sb.st_mode)) {
            if (full.back() != '/') full.push_back('/');
            full.append("index.html");
            ok = stat(full.c_str(), &sb) == 0 && S_ISREG(sb.st_mode);
        } else if (!S_ISREG(sb.st_mode)) {
            ok = false;
        }
        if (ok && sb.st_size > (16 << 20)) ok = false;
    }
    std::string body;
    if (ok) {
        FILE* f = std::fopen(full.c_str(), "rb");
        if (!f) {
            ok = false;
        } else {
            char buf[16384];
            size_t got;
            while ((got = std::fread(buf, 1, sizeof buf, f)) > 0) body.append(buf, got);
            std::fclose(f);
        }
    }
    if (!ok) {
        append_response(&conn.out(), 404, "text/html", undertow_error_page(404), false, conn.closing(), head_only,
                        worker.http_date());
        return;
    }
    append_response(&conn.out(), 200, mime_type_for(full), body, false, conn.closing(), head_only, worker.http_date());
}

}  // namespace hola
