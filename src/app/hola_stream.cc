// GET /hystrix.stream: server-sent events with Hystrix 1.5.1's payloads
// (/root/reference/src/main/java/com/redhat/developers/msa/hola/HystrixServlet.java:12-17).
#include <cstdio>

#include "app/hola_app_internal.h"

namespace hola {

using namespace detail;

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
    // An HTTP/1.0 client cannot read chunked framing: for it the stream is delimited by the
    // connection closing (which, for a stream that never ends, is when the client goes away).
    const bool chunked = rq.version_minor >= 1;
    std::string& out = conn.out();
    out.append(chunked ? "HTTP/1.1 200 OK\r\nConnection: keep-alive\r\n" : "HTTP/1.1 200 OK\r\nConnection: close\r\n");
    out.append("Cache-Control: no-cache, no-store, max-age=0, must-revalidate\r\nPragma: no-cache\r\n"
               "Content-Type: text/event-stream;charset=UTF-8\r\n");
    if (chunked) out.append("Transfer-Encoding: chunked\r\n");
    out.append("Date: ");
    out.append(worker.http_date());
    out.append("\r\n\r\n");
    StreamSub sub;
    sub.ref = conn.ref();
    sub.delay_ms = delay_ms;
    sub.chunked = chunked;
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
    std::string chunk;
    if (it->second.chunked) {
        char size_line[24];
        std::snprintf(size_line, sizeof size_line, "%zx\r\n", event.size());
        chunk = size_line;
        chunk.append(event);
        chunk.append("\r\n");
    } else {
        chunk = std::move(event);
    }
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

}  // namespace hola
