// GET /api/hola-chaining: the downstream call with circuit breaker, bulkhead, timeout, Feign-style
// retries and static fallback (/root/reference/src/main/java/com/redhat/developers/msa/hola/HolaResource.java:69-78, 94-113;
// /root/reference/src/main/resources/config.properties:17,19).
//
// Cost model. The reference builds a new HTTP client, a new Feign proxy and a new Hystrix command
// per request (HolaResource.java:94-113). Here the three outcomes cost:
//   * refused by the breaker or the bulkhead (what a dead `aloha` turns every request into once
//     the circuit is open): no allocation, no shared lock - the breaker's counters are sharded per
//     event loop (core/breaker.h) and the fallback answer is a pre-rendered byte string per worker
//     and greeting version, exactly like the /api/hola answer;
//   * the downstream call runs: one pooled ChainCall object per request (recycled with the
//     capacity of its strings), one pooled keep-alive connection, one timer;
//   * fallbacks switched off and the call fails: a 500, rendered on the spot.
#include <cstdio>
#include <cstring>

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

HolaApp::CallRef HolaApp::new_call(WorkerState& st) {
    ChainCall* c;
    if (!st.call_pool.empty()) {
        c = st.call_pool.back();
        st.call_pool.pop_back();
    } else {
        c = new ChainCall();
        c->owner = &st;
    }
    return CallRef(c);
}

// ["<hola text>", ...tail] or, without a tail, ["<hola text>","<fallback>"] - compact, as
// json_encode_string_list writes it, without building the list first.
void HolaApp::append_chain_body(std::string* out, std::string_view hola_text, const JsonStringList* tail) const {
    out->push_back('[');
    json_append_string(out, hola_text);
    if (!tail) {
        out->append(fallback_json_tail_);
        return;
    }
    for (const auto& item : *tail) {
        out->push_back(',');
        if (item) {
            json_append_string(out, *item);
        } else {
            out->append("null");
        }
    }
    out->push_back(']');
}

// The answer of every request that did not get a downstream reply, when fallbacks are on: the
// static fallback list (HolaResource.java:112) after the greeting.
void HolaApp::append_fallback_response(WorkerState& st, Connection& conn, const RequestHead& rq, const GreetingSnapshot& g) {
    Worker& worker = *st.worker;
    if (rq.method == Method::kGet && !conn.closing()) {
        const int64_t second = worker.loop().now_unix_s();
        if (st.fallback_version != g.version || st.fallback_bytes.empty()) {
            st.scratch.clear();
            append_chain_body(&st.scratch, g.text, nullptr);
            st.fallback_bytes.clear();
            st.fallback_date_at = append_response(&st.fallback_bytes, 200, kJson, st.scratch, true, false, false, worker.http_date());
            st.fallback_version = g.version;
            st.fallback_second = second;
        } else if (second != st.fallback_second) {
            std::memcpy(&st.fallback_bytes[st.fallback_date_at], worker.http_date().data(), 29);
            st.fallback_second = second;
        }
        conn.out().append(st.fallback_bytes);
        return;
    }
    st.scratch.clear();
    append_chain_body(&st.scratch, g.text, nullptr);
    append_response(&conn.out(), 200, kJson, st.scratch, true, conn.closing(), rq.method == Method::kHead, worker.http_date());
}

Disposition HolaApp::serve_chaining(WorkerState& st, Connection& conn, const RequestHead& rq) {
    Worker& worker = *st.worker;
    const GreetingSnapshot& g = current_greeting(st);
    const bool traced = st.ring != nullptr && rq.path.size() == 18;  // exactly "/api/hola-chaining"
    const bool head_only = rq.method == Method::kHead;
    const int64_t server_start_us = worker.loop().now_unix_us();
    TraceContext server_ctx;
    if (traced) server_ctx = start_server_trace(rq.b3, &st.ids);
    const std::string_view url_host = rq.host.empty() ? std::string_view("localhost") : rq.host;
    if (!g.ok) {
        // hola() throws before the downstream call is attempted.
        append_response(&conn.out(), 500, "text/html", undertow_error_page(500), false, conn.closing(), head_only, worker.http_date());
        record_span(st, server_ctx, false, server_start_us, realtime_us(), 500, "http://", url_host, rq.target);
        return Disposition::kDone;
    }

    const int64_t now_ms = worker.loop().now_ns() / 1000000;
    const HystrixCommand::Admission adm = command_.admit(now_ms, worker.id());
    if (adm != HystrixCommand::Admission::kRun) {
        // Short-circuited (circuit open) or rejected (bulkhead full): straight to the fallback.
        if (options_.verbose) {
            std::fprintf(stderr, "hola: aloha call %s -> fallback\n",
                         adm == HystrixCommand::Admission::kShortCircuited ? "short-circuited" : "rejected (pool full)");
        }
        int status = 200;
        if (options_.breaker.fallback_enabled) {
            append_fallback_response(st, conn, rq, g);
        } else {
            status = 500;
            append_response(&conn.out(), 500, "text/html", undertow_error_page(500), false, conn.closing(), head_only, worker.http_date());
        }
        if (server_ctx.sampled) record_span(st, server_ctx, false, server_start_us, realtime_us(), status, "http://", url_host, rq.target);
        return Disposition::kDone;
    }

    CallRef call = new_call(st);
    call->greeting = st.greeting;
    call->ref = conn.ref();
    call->head_only = head_only;
    call->close = conn.closing();
    call->server_start_us = server_start_us;
    call->server_ctx = server_ctx;
    if (traced) {
        call->server_url.assign("http://");
        call->server_url.append(url_host);
        call->server_url.append(rq.target);
    }
    start_chain(st, call);
    if (call->finished) {
        conn.out().append(call->response);
        return Disposition::kDone;
    }
    call->deferred = true;
    return Disposition::kDeferred;
}

// The breaker has admitted the call: arm the command timeout and send the first attempt.
void HolaApp::start_chain(WorkerState& st, const CallRef& call) {
    EventLoop& loop = st.worker->loop();
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
                command_.complete_timeout(lp.now_ns() / 1000000, static_cast<int>((lp.now_ns() - call->exec_start_ns) / 1000000),
                                          stp->worker->id());
                if (options_.verbose) std::fprintf(stderr, "hola: aloha call timed out -> fallback\n");
                finish_chain(*stp, call, nullptr);
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
void HolaApp::send_chain_attempt(WorkerState& st, const CallRef& call) {
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
    if (options_.log_feign) std::fprintf(stderr, "[AlohaService#aloha] ---> GET %s HTTP/1.1\n", aloha_url_.c_str());
    call->request = st.http->send(
        options_.aloha_host, options_.aloha_port, std::move(request), [this, stp, call](ClientResult&& r) {
            if (call->settled) return;
            EventLoop& lp = stp->worker->loop();
            const int shard = stp->worker->id();
            const int64_t now_ns = monotonic_ns();
            const int exec_ms = static_cast<int>((now_ns - call->exec_start_ns) / 1000000);
            const bool io_failure = r.code == ClientResult::Code::kResolveFailed || r.code == ClientResult::Code::kConnectFailed ||
                                    r.code == ClientResult::Code::kIoError;
            if (io_failure && call->attempt < options_.feign_max_attempts) {
                double interval_ms = options_.feign_retry_period_ms;
                for (int k = 0; k < call->attempt; ++k) interval_ms *= 1.5;
                if (interval_ms > options_.feign_retry_max_period_ms) interval_ms = options_.feign_retry_max_period_ms;
                if (options_.log_feign) {
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
            if (options_.log_feign) {
                if (r.code == ClientResult::Code::kOk) {
                    std::fprintf(stderr, "[AlohaService#aloha] <--- HTTP/1.1 %d (%dms)\n", r.status, exec_ms);
                } else {
                    std::fprintf(stderr, "[AlohaService#aloha] <--- ERROR %s (%dms)\n", why.c_str(), exec_ms);
                }
            }
            if (ok) {
                command_.complete_success(now_ns / 1000000, exec_ms, shard);
                finish_chain(*stp, call, &tail);
            } else {
                command_.complete_failure(now_ns / 1000000, exec_ms, shard);
                finish_chain(*stp, call, nullptr);
            }
        });
}

void HolaApp::finish_chain(WorkerState& st, const CallRef& call, const JsonStringList* tail) {
    if (call->finished) return;
    call->finished = true;
    Worker& worker = *st.worker;
    const bool close = call->close || worker.draining();
    std::string& response = call->response;
    response.clear();
    int status = 200;
    if (!tail && !options_.breaker.fallback_enabled) {
        // hystrix.command.default.fallback.enabled=false: Hystrix throws, the resource method
        // fails and the client sees a 500.
        call->failed_without_fallback = true;
        status = 500;
        append_response(&response, 500, "text/html", undertow_error_page(500), false, close, call->head_only, worker.http_date());
    } else {
        st.scratch.clear();
        append_chain_body(&st.scratch, call->greeting->text, tail);
        append_response(&response, 200, kJson, st.scratch, true, close, call->head_only, worker.http_date());
    }
    record_span(st, call->server_ctx, false, call->server_start_us, realtime_us(), status, call->server_url, {}, {});
    if (call->deferred) worker.complete(call->ref, response, close);
    // else: serve_chaining() is still on the stack and takes call->response itself
}

}  // namespace hola
