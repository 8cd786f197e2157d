// GET /api/hola-chaining: the downstream call with circuit breaker, bulkhead, timeout, Feign-style
// retries and static fallback (/root/reference/src/main/java/com/redhat/developers/msa/hola/HolaResource.java:69-78, 94-113;
// /root/reference/src/main/resources/config.properties:17,19).
#include <cstdio>

#include "app/hola_app_internal.h"

namespace hola {

using namespace detail;

namespace {

void append_b3_id(std::string* out, const char* name, uint64_t id) {
    char buf[16];
    format_trace_id(id, buf);
    out->append(name);
    out->append(": ");
    out->append(buf, 16);
    out->append("\r\n");
}

}  // namespace

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

}  // namespace hola
