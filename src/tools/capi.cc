// Implementation of the C ABI declared in include/hola.h.
#include "hola.h"

#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "app/service.h"
#include "core/breaker.h"
#include "core/config.h"
#include "core/formatter.h"
#include "core/histogram.h"
#include "core/http.h"
#include "core/json.h"
#include "core/properties.h"
#include "core/trace.h"

#ifndef HOLA_VERSION
#define HOLA_VERSION "0.0.1-SNAPSHOT"
#endif

using namespace hola;

namespace {

void give(const std::string& s, char** out, size_t* out_len) {
    char* p = static_cast<char*>(std::malloc(s.size() + 1));
    std::memcpy(p, s.data(), s.size());
    p[s.size()] = '\0';
    *out = p;
    if (out_len) *out_len = s.size();
}

void put_u32(std::string* s, uint32_t v) {
    char b[4] = {static_cast<char>(v & 0xFF), static_cast<char>((v >> 8) & 0xFF), static_cast<char>((v >> 16) & 0xFF),
                 static_cast<char>((v >> 24) & 0xFF)};
    s->append(b, 4);
}

std::string pack(const JsonStringList& items) {
    std::string s;
    put_u32(&s, static_cast<uint32_t>(items.size()));
    for (const auto& it : items) {
        s.push_back(it ? 1 : 0);
        put_u32(&s, it ? static_cast<uint32_t>(it->size()) : 0);
        if (it) s.append(*it);
    }
    return s;
}

bool unpack(const char* data, size_t len, JsonStringList* out) {
    size_t pos = 0;
    auto get_u32 = [&](uint32_t* v) {
        if (pos + 4 > len) return false;
        const auto* p = reinterpret_cast<const unsigned char*>(data + pos);
        *v = static_cast<uint32_t>(p[0]) | (static_cast<uint32_t>(p[1]) << 8) | (static_cast<uint32_t>(p[2]) << 16) |
             (static_cast<uint32_t>(p[3]) << 24);
        pos += 4;
        return true;
    };
    uint32_t count = 0;
    if (!get_u32(&count)) return false;
    for (uint32_t i = 0; i < count; ++i) {
        if (pos + 1 > len) return false;
        const bool present = data[pos++] != 0;
        uint32_t n = 0;
        if (!get_u32(&n) || pos + n > len) return false;
        if (present) {
            out->emplace_back(std::string(data + pos, n));
        } else {
            out->emplace_back(std::nullopt);
        }
        pos += n;
    }
    return true;
}

void json_kv(std::string* o, const char* k, std::string_view v) {
    if (o->back() != '{') o->push_back(',');
    json_append_string(o, k);
    o->push_back(':');
    json_append_string(o, v);
}
void json_kn(std::string* o, const char* k, long long v) {
    if (o->back() != '{') o->push_back(',');
    json_append_string(o, k);
    o->push_back(':');
    o->append(std::to_string(v));
}
void json_kb(std::string* o, const char* k, bool v) {
    if (o->back() != '{') o->push_back(',');
    json_append_string(o, k);
    o->append(v ? ":true" : ":false");
}

}  // namespace

struct hola_config {
    ConfigResolver resolver;
    SystemPropertySource* system = nullptr;
    ConfFileSource* conf = nullptr;
    std::unique_ptr<CachedConfigValue> hello;
};

struct hola_breaker {
    std::unique_ptr<HystrixCommand> cmd;
};

struct hola_histogram {
    Histogram h;
};

struct hola_service {
    std::unique_ptr<Service> service;
};

extern "C" {

const char* hola_version(void) { return HOLA_VERSION; }

void hola_free(void* p) { std::free(p); }

int hola_properties_parse(const char* text, size_t len, char** out, size_t* out_len) {
    PropertyList list;
    std::string error;
    if (!parse_properties(std::string_view(text, len), &list, &error)) {
        give(error, out, out_len);
        return 1;
    }
    JsonStringList flat;
    for (auto& kv : list) {
        flat.emplace_back(std::move(kv.first));
        flat.emplace_back(std::move(kv.second));
    }
    give(pack(flat), out, out_len);
    return 0;
}

int hola_format(const char* pattern, size_t pattern_len, const char* arg, size_t arg_len, char** out, size_t* out_len) {
    FormatResult r = java_format(std::string_view(pattern, pattern_len), std::string_view(arg, arg_len));
    give(r.ok ? r.text : r.error, out, out_len);
    return r.ok ? 0 : 1;
}

int hola_json_encode_list(const char* packed, size_t packed_len, char** out, size_t* out_len) {
    JsonStringList items;
    if (!unpack(packed, packed_len, &items)) {
        give("malformed packed list", out, out_len);
        return 1;
    }
    give(json_encode_string_list(items), out, out_len);
    return 0;
}

int hola_json_decode_list(const char* text, size_t len, char** out, size_t* out_len) {
    JsonStringList items;
    std::string error;
    if (!json_decode_string_list(std::string_view(text, len), &items, &error)) {
        give(error, out, out_len);
        return 1;
    }
    give(pack(items), out, out_len);
    return 0;
}

hola_config* hola_config_new(void) {
    auto* c = new hola_config();
    c->system = new SystemPropertySource();
    c->conf = new ConfFileSource(c->system);
    c->resolver.add_source(std::unique_ptr<ConfigSource>(c->system));
    c->resolver.add_source(std::unique_ptr<ConfigSource>(new EnvironmentSource()));
    c->resolver.add_source(std::unique_ptr<ConfigSource>(c->conf));
    return c;
}

void hola_config_free(hola_config* c) { delete c; }

void hola_config_set_system_property(hola_config* c, const char* key, const char* value) { c->system->set(key, value); }

void hola_config_add_properties_file(hola_config* c, const char* path) {
    c->resolver.add_source(
        std::unique_ptr<ConfigSource>(new StaticPropertiesSource("META-INF/apache-deltaspike.properties", path)));
}

int hola_config_resolve(hola_config* c, const char* key, char** out, size_t* out_len) {
    std::string v;
    if (!c->resolver.resolve(key, &v)) {
        give("", out, out_len);
        return 0;
    }
    give(v, out, out_len);
    return 1;
}

int hola_config_greeting(hola_config* c, const char* hostname, int64_t now_ns, char** out, size_t* out_len, int* refreshed) {
    if (!c->hello) c->hello.reset(new CachedConfigValue(&c->resolver, "hello", "Hola de %s", 5000000000ll, false));
    bool r = false;
    const std::string& pattern = c->hello->get(now_ns, &r);
    if (refreshed) *refreshed = r ? 1 : 0;
    FormatResult f = java_format(pattern, hostname);
    give(f.ok ? f.text : f.error, out, out_len);
    return f.ok ? 0 : 1;
}

uint64_t hola_config_file_reads(const hola_config* c) { return c->conf->reads(); }

hola_breaker* hola_breaker_new(const char* config_text) {
    BreakerConfig cfg = BreakerConfig::reference_defaults();
    if (config_text && *config_text) {
        Properties props;
        if (props.load(config_text)) cfg.apply(props);
    }
    auto* b = new hola_breaker();
    b->cmd.reset(new HystrixCommand("AlohaService#aloha()", "http://aloha:8080/", "http://aloha:8080/", cfg));
    return b;
}

void hola_breaker_free(hola_breaker* b) { delete b; }

int hola_breaker_admit(hola_breaker* b, int64_t now_ms) {
    switch (b->cmd->admit(now_ms)) {
        case HystrixCommand::Admission::kRun: return 0;
        case HystrixCommand::Admission::kShortCircuited: return 1;
        case HystrixCommand::Admission::kPoolRejected: return 2;
    }
    return 0;
}

void hola_breaker_complete(hola_breaker* b, int64_t now_ms, int outcome, int execute_ms) {
    if (outcome == 0) {
        b->cmd->complete_success(now_ms, execute_ms);
    } else if (outcome == 1) {
        b->cmd->complete_failure(now_ms, execute_ms);
    } else {
        b->cmd->complete_timeout(now_ms, execute_ms);
    }
}

int hola_breaker_is_open(hola_breaker* b, int64_t now_ms) { return b->cmd->is_open(now_ms) ? 1 : 0; }

void hola_breaker_health(hola_breaker* b, int64_t now_ms, int64_t* total, int64_t* errors, int* error_pct) {
    const HealthCounts h = b->cmd->health(now_ms);
    if (total) *total = h.total;
    if (errors) *errors = h.errors;
    if (error_pct) *error_pct = h.error_pct;
}

int hola_breaker_command_json(hola_breaker* b, int64_t now_ms, char** out, size_t* out_len) {
    give(hystrix_command_json(*b->cmd, b->cmd->snapshot(now_ms)), out, out_len);
    return 0;
}

int hola_breaker_pool_json(hola_breaker* b, int64_t now_ms, char** out, size_t* out_len) {
    give(hystrix_pool_json(*b->cmd, b->cmd->pool_snapshot(now_ms)), out, out_len);
    return 0;
}

int hola_breaker_timeout_ms(const hola_breaker* b) { return b->cmd->config().timeout_ms; }

int hola_b3_server_decision(const char* trace_id, const char* span_id, const char* parent_span_id, const char* sampled,
                            uint64_t seed, uint64_t ids[3]) {
    B3Headers h;
    if (trace_id) h.trace_id = trace_id;
    if (span_id) h.span_id = span_id;
    if (parent_span_id) h.parent_span_id = parent_span_id;
    if (sampled) {
        h.sampled = sampled;
        h.has_sampled = true;
    }
    IdGenerator gen(seed);
    const TraceContext ctx = start_server_trace(h, &gen);
    ids[0] = ctx.trace_id;
    ids[1] = ctx.span_id;
    ids[2] = ctx.parent_id;
    return (ctx.sampled ? 1 : 0) | (ctx.joined ? 2 : 0) | (ctx.has_parent ? 4 : 0);
}

int hola_span_json(uint64_t trace_id, uint64_t span_id, uint64_t parent_id, int has_parent, int kind, int shared,
                   int64_t start_us, int64_t end_us, int status, const char* url, const char* service_name,
                   uint32_t ipv4_host_order, char** out, size_t* out_len) {
    Span s;
    s.trace_id = trace_id;
    s.id = span_id;
    s.parent_id = parent_id;
    s.has_parent = has_parent != 0;
    s.first_annotation = kind ? "cs" : "sr";
    s.last_annotation = kind ? "cr" : "ss";
    s.first_us = start_us;
    s.last_us = end_us;
    if (!shared) {
        s.timestamp_us = start_us;
        s.duration_us = end_us - start_us;
    }
    s.tags.push_back(BinaryAnnotation{"http.url", url ? url : ""});
    if (status != 0 && (status < 200 || status > 299)) s.tags.push_back(BinaryAnnotation{"http.status_code", std::to_string(status)});
    Endpoint ep;
    ep.service_name = service_name ? service_name : "hola";
    ep.ipv4 = ipv4_host_order;
    std::string json;
    span_append_json(&json, s, ep);
    give(json, out, out_len);
    return 0;
}

int hola_http_parse_request(const char* data, size_t len, char** out, size_t* out_len) {
    RequestHead h;
    const ParseStatus st = parse_request_head(data, len, &h);
    if (st == ParseStatus::kIncomplete) return 1;
    if (st == ParseStatus::kBad) return 2;
    std::string o = "{";
    json_kv(&o, "method", h.method_text);
    json_kn(&o, "method_id", static_cast<int>(h.method));
    json_kv(&o, "target", h.target);
    json_kv(&o, "path", h.path);
    std::string canonical;
    if (canonicalise_path(h.raw_path, &canonical)) {
        json_kv(&o, "canonical_path", canonical);
    } else {
        o.append(",\"canonical_path\":null");
    }
    json_kb(&o, "path_is_canonical", h.path_is_canonical);
    json_kv(&o, "query", h.query);
    json_kn(&o, "version_minor", h.version_minor);
    json_kb(&o, "keep_alive", h.keep_alive);
    json_kb(&o, "chunked", h.chunked);
    json_kb(&o, "expect_continue", h.expect_continue);
    json_kn(&o, "content_length", h.content_length);
    json_kv(&o, "host", h.host);
    json_kb(&o, "has_accept", h.has_accept);
    json_kv(&o, "accept", h.accept);
    json_kv(&o, "b3_trace_id", h.b3.trace_id);
    json_kv(&o, "b3_span_id", h.b3.span_id);
    json_kv(&o, "b3_parent_span_id", h.b3.parent_span_id);
    json_kb(&o, "b3_has_sampled", h.b3.has_sampled);
    json_kv(&o, "b3_sampled", h.b3.sampled);
    json_kn(&o, "head_len", static_cast<long long>(h.head_len));
    o.push_back('}');
    give(o, out, out_len);
    return 0;
}

int hola_http_parse_response(const char* data, size_t len, char** out, size_t* out_len) {
    ResponseHead h;
    const ParseStatus st = parse_response_head(data, len, &h);
    if (st == ParseStatus::kIncomplete) return 1;
    if (st == ParseStatus::kBad) return 2;
    std::string o = "{";
    json_kn(&o, "status", h.status);
    json_kn(&o, "version_minor", h.version_minor);
    json_kb(&o, "keep_alive", h.keep_alive);
    json_kb(&o, "chunked", h.chunked);
    json_kn(&o, "content_length", h.content_length);
    json_kv(&o, "content_type", h.content_type);
    json_kn(&o, "head_len", static_cast<long long>(h.head_len));
    o.push_back('}');
    give(o, out, out_len);
    return 0;
}

int hola_http_dechunk(const char* data, size_t len, char** out, size_t* out_len, size_t* consumed) {
    ChunkedDecoder d;
    std::string body;
    const size_t used = d.feed(data, len, &body);
    if (consumed) *consumed = used;
    give(body, out, out_len);
    if (d.failed()) return 2;
    return d.done() ? 0 : 1;
}

int hola_http_accept_allows(const char* accept, const char* type, const char* subtype) {
    return accept_allows(accept ? accept : "", type, subtype) ? 1 : 0;
}

void hola_http_date(int64_t unix_seconds, char out[30]) {
    format_http_date(unix_seconds, out);
    out[29] = '\0';
}

hola_histogram* hola_histogram_new(void) { return new hola_histogram(); }
void hola_histogram_free(hola_histogram* h) { delete h; }
void hola_histogram_record(hola_histogram* h, uint64_t value) { h->h.record(value); }
uint64_t hola_histogram_percentile(const hola_histogram* h, double pct) { return h->h.percentile(pct); }
uint64_t hola_histogram_count(const hola_histogram* h) { return h->h.count(); }

hola_service* hola_service_start(int argc, const char* const* argv, char** error) {
    std::vector<std::string> args;
    for (int i = 0; i < argc; ++i) args.emplace_back(argv[i]);
    ServiceConfig cfg;
    std::string message;
    if (parse_service_args(args, &cfg, &message) != ParseOutcome::kRun) {
        if (error) give(message, error, nullptr);
        return nullptr;
    }
    auto* s = new hola_service();
    s->service.reset(new Service(std::move(cfg)));
    std::string err;
    if (!s->service->start(&err)) {
        if (error) give(err, error, nullptr);
        delete s;
        return nullptr;
    }
    return s;
}

int hola_service_port(const hola_service* s) { return s->service->port(); }
int hola_service_threads(const hola_service* s) { return s->service->threads(); }

void hola_service_stats(const hola_service* s, uint64_t* requests, uint64_t* connections) {
    const ServerStats st = s->service->server_stats();
    if (requests) *requests = st.requests;
    if (connections) *connections = st.connections_accepted;
}

void hola_service_trace_stats(const hola_service* s, uint64_t* reported, uint64_t* dropped) {
    const TracerStats st = s->service->tracer_stats();
    if (reported) *reported = st.spans_reported;
    if (dropped) *dropped = st.spans_dropped;
}

int hola_service_stats_json(const hola_service* s, char** out, size_t* out_len) {
    give(s->service->stats_json(), out, out_len);
    return 0;
}

void hola_service_stop(hola_service* s) {
    if (!s) return;
    s->service->stop();
    delete s;
}

}  // extern "C"
