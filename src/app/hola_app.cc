#include "app/hola_app.h"

#include <limits.h>
#include <sys/stat.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "app/hola_app_internal.h"
#include "app/static_content.h"
#include "core/text.h"

namespace hola {

namespace detail {

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

size_t append_response(std::string* out, int status, std::string_view content_type, std::string_view body,
                       bool cors, bool close, bool head_only, std::string_view date,
                       std::string_view extra_headers) {
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

std::string undertow_error_page(int status) {
    return std::string("<html><head><title>Error</title></head><body>") + reason_phrase(status) + "</body></html>";
}

}  // namespace detail

using namespace detail;

namespace {

void append_html_escaped(std::string* out, std::string_view s) {
    for (char c : s) {
        switch (c) {
            case '<': out->append("&lt;"); break;
            case '>': out->append("&gt;"); break;
            case '&': out->append("&amp;"); break;
            case '"': out->append("&quot;"); break;
            case '\'': out->append("&#39;"); break;
            default: out->push_back(c); break;
        }
    }
}

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

}  // namespace

namespace {

int64_t mtime_ns_of(const struct stat& sb) {
    return static_cast<int64_t>(sb.st_mtim.tv_sec) * 1000000000ll + sb.st_mtim.tv_nsec;
}

// host:port as it appears in URLs and Host headers (IPv6 literals go in brackets).
std::string authority_of(const AppOptions& o) {
    const bool v6_literal = o.aloha_host.find(':') != std::string::npos;
    return (v6_literal ? "[" + o.aloha_host + "]" : o.aloha_host) + ":" + std::to_string(o.aloha_port);
}

}  // namespace

HolaApp::HolaApp(AppOptions options, GreetingService* greeting, Tracer* tracer, Resolver* resolver)
    : options_(std::move(options)),
      greeting_(greeting),
      tracer_(tracer),
      resolver_(resolver),
      // Key names as HystrixFeign derives them: command = Feign.configKey(type, method),
      // group and thread pool = the target's name, which for a hard-coded target is its URL.
      command_("AlohaService#aloha()", "http://" + authority_of(options_) + "/", "http://" + authority_of(options_) + "/",
               options_.breaker) {
    const std::string authority = authority_of(options_);
    aloha_stand_in_body_ = json_encode_string_list({std::string("Aloha mai ") + options_.hostname});
    fallback_json_tail_ = ",";
    json_append_string(&fallback_json_tail_, options_.fallback);
    fallback_json_tail_.push_back(']');
    aloha_url_ = "http://" + authority + options_.aloha_path;
    if (!options_.webjars_dir.empty()) {
        char resolved[PATH_MAX];
        webjars_root_ = realpath(options_.webjars_dir.c_str(), resolved) ? std::string(resolved) : options_.webjars_dir;
        while (webjars_root_.size() > 1 && webjars_root_.back() == '/') webjars_root_.pop_back();
    }
    aloha_request_prefix_ = "GET " + options_.aloha_path + " HTTP/1.1\r\nHost: " + authority +
                            "\r\nAccept: */*\r\nConnection: keep-alive\r\nUser-Agent: hola-b200\r\n";
}

// The per-worker states die with the application; the workers (and their event loops) are gone
// by then, see Service::stop() and on_worker_exit().
HolaApp::~HolaApp() = default;

void HolaApp::on_worker_start(Worker& worker) {
    WorkerState* st;
    {
        std::lock_guard<std::mutex> lock(states_mu_);
        states_.emplace_back(new WorkerState());
        st = states_.back().get();
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
        // Stand-in for the next service of the demo chain (--serve-aloha): lets two hola processes
        // be chained for demos and for measuring the downstream path without a JVM or Vert.x.
        if (options_.serve_aloha && path == "/api/aloha-chaining" && (rq.method == Method::kGet || rq.method == Method::kHead)) {
            append_response(&out, 200, kJson, aloha_stand_in_body_, true, conn.closing(), rq.method == Method::kHead,
                            worker.http_date());
            return Disposition::kDone;
        }
        const ApiRoute route = match_api(path.substr(4));
        if (route == ApiRoute::kNone) {
            // The message quotes the request URL and goes out as text/html: escape it, or any
            // link to a non-existent /api path becomes a script-injection vector.
            std::string body = "RESTEASY003210: Could not find resource for full path: http://";
            append_html_escaped(&body, rq.host.empty() ? std::string_view("localhost") : rq.host);
            append_html_escaped(&body, rq.target);
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
            // A client sends the same Accept header on every request; evaluate a header once and
            // remember the verdicts for the three media types this service produces.
            if (rq.accept != st.accept_header) {
                st.accept_header.assign(rq.accept);
                st.accept_text_plain = accept_allows(rq.accept, "text", "plain");
                st.accept_json = accept_allows(rq.accept, "application", "json");
                st.accept_yaml = accept_allows(rq.accept, "application", "yaml");
            }
            bool ok = true;
            switch (route) {
                case ApiRoute::kHola:
                case ApiRoute::kHealth:
                    ok = st.accept_text_plain;
                    break;
                case ApiRoute::kChaining:
                case ApiRoute::kSwaggerJson:
                    ok = st.accept_json;
                    break;
                case ApiRoute::kSwaggerYaml:
                    ok = st.accept_yaml;
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

    if (path.size() > 9 && std::memcmp(path.data(), "/webjars/", 9) == 0 && (rq.method == Method::kGet || rq.method == Method::kHead)) {
        if (!options_.webjars_dir.empty()) {
            serve_static_file(st, conn, rq);
            return Disposition::kDone;
        }
        // No Swagger-UI assets installed: the landing page's iframe still gets a working page.
        if (path == "/webjars/swagger-ui/2.1.4/" || path == "/webjars/swagger-ui/2.1.4/index.html") {
            append_response(&out, 200, "text/html", api_browser_html(), false, conn.closing(), rq.method == Method::kHead,
                            worker.http_date());
            return Disposition::kDone;
        }
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

// Static assets under --webjars-dir (the Swagger-UI webjar of the reference, pom.xml:207-211).
//
// Containment: the decoded path is resolved with realpath() and must stay below the resolved
// root, so neither ".." nor a symbolic link inside the directory can lead out of it.
// Cost: a request normally costs one realpath() and one stat() - both answered from the kernel's
// dentry cache - and the body comes from an in-memory copy keyed by path, size and mtime. Only a
// file that changed (or was never seen) is read, in one go, on the event-loop thread; files above
// kStaticMaxFile are refused rather than allowed to stall the loop.
void HolaApp::serve_static_file(WorkerState& st, Connection& conn, const RequestHead& rq) {
    constexpr off_t kStaticMaxFile = 16 << 20;
    constexpr size_t kStaticCacheBytes = 64u << 20;
    Worker& worker = *st.worker;
    const bool head_only = rq.method == Method::kHead;
    // rq.path is canonical already: decoded (but for %2F/%5C, which then simply name no file),
    // matrix parameters cut, dot segments resolved.
    const std::string rel(rq.path.substr(8));  // keeps the leading '/'
    bool ok = true;
    {
        // No hidden files, no NULs, no backslashes (the resolved-path check below does the rest).
        if (rel.find('\0') != std::string::npos || rel.find("/.") != std::string::npos || rel.find('\\') != std::string::npos)
            ok = false;
    }
    std::string full;
    struct stat sb{};
    if (ok) {
        char resolved[PATH_MAX];
        const std::string candidate = webjars_root_ + rel;
        if (!realpath(candidate.c_str(), resolved)) {
            ok = false;
        } else {
            full = resolved;
            if (stat(full.c_str(), &sb) != 0) {
                ok = false;
            } else if (S_ISDIR(sb.st_mode)) {
                const std::string index = full + "/index.html";
                ok = realpath(index.c_str(), resolved) != nullptr;
                if (ok) {
                    full = resolved;
                    ok = stat(full.c_str(), &sb) == 0;
                }
            }
        }
        if (ok && !S_ISREG(sb.st_mode)) ok = false;
        if (ok && (full.size() <= webjars_root_.size() || full.compare(0, webjars_root_.size(), webjars_root_) != 0 ||
                   full[webjars_root_.size()] != '/'))
            ok = false;  // a link (or the root's own parent) led outside the directory
        if (ok && sb.st_size > kStaticMaxFile) ok = false;
    }
    std::shared_ptr<const StaticFile> file;
    if (ok) {
        {
            std::lock_guard<std::mutex> lock(static_mu_);
            auto it = static_cache_.find(full);
            if (it != static_cache_.end() && it->second->size == sb.st_size && it->second->mtime_ns == mtime_ns_of(sb))
                file = it->second;
        }
        if (!file) {
            auto fresh = std::make_shared<StaticFile>();
            fresh->size = sb.st_size;
            fresh->mtime_ns = mtime_ns_of(sb);
            FILE* f = std::fopen(full.c_str(), "rb");
            if (!f) {
                ok = false;
            } else {
                fresh->body.reserve(static_cast<size_t>(sb.st_size));
                char buf[65536];
                size_t got;
                while ((got = std::fread(buf, 1, sizeof buf, f)) > 0 && fresh->body.size() <= static_cast<size_t>(kStaticMaxFile))
                    fresh->body.append(buf, got);
                std::fclose(f);
                file = fresh;
                std::lock_guard<std::mutex> lock(static_mu_);
                auto old = static_cache_.find(full);
                if (old != static_cache_.end()) {
                    static_cache_bytes_ -= old->second->body.size();
                    static_cache_.erase(old);
                }
                if (static_cache_bytes_ + fresh->body.size() > kStaticCacheBytes) {  // crude but bounded: start over
                    static_cache_.clear();
                    static_cache_bytes_ = 0;
                }
                static_cache_bytes_ += fresh->body.size();
                static_cache_.emplace(full, fresh);
            }
        }
    }
    if (!ok) {
        append_response(&conn.out(), 404, "text/html", undertow_error_page(404), false, conn.closing(), head_only,
                        worker.http_date());
        return;
    }
    append_response(&conn.out(), 200, mime_type_for(full), file->body, false, conn.closing(), head_only, worker.http_date());
}

}  // namespace hola
