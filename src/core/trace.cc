#include "core/trace.h"

#include <arpa/inet.h>
#include <ifaddrs.h>
#include <net/if.h>
#include <netinet/in.h>
#include <sys/random.h>

#include <chrono>
#include <cstring>

#include "core/json.h"
#include "core/text.h"

namespace hola {

bool parse_trace_id(std::string_view hex, uint64_t* out) {
    if (hex.empty() || hex.size() > 32) return false;
    if (hex.size() > 16) hex = hex.substr(hex.size() - 16);
    uint64_t v = 0;
    for (char c : hex) {
        if (!is_hex_digit(c)) return false;
        v = (v << 4) | static_cast<uint64_t>(hex_value(c));
    }
    *out = v;
    return true;
}

void format_trace_id(uint64_t id, char out[16]) {
    static const char kHex[] = "0123456789abcdef";
    for (int i = 15; i >= 0; --i) {
        out[i] = kHex[id & 15];
        id >>= 4;
    }
}

std::string trace_id_string(uint64_t id) {
    char buf[16];
    format_trace_id(id, buf);
    return std::string(buf, 16);
}

namespace {

inline uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }

uint64_t splitmix64(uint64_t* state) {
    uint64_t z = (*state += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

}  // namespace

IdGenerator::IdGenerator() {
    uint64_t seed = 0;
    if (getrandom(&seed, sizeof seed, GRND_NONBLOCK) != static_cast<ssize_t>(sizeof seed)) {
        seed = static_cast<uint64_t>(std::chrono::steady_clock::now().time_since_epoch().count()) ^
               reinterpret_cast<uintptr_t>(this);
    }
    for (auto& s : s_) s = splitmix64(&seed);
}

IdGenerator::IdGenerator(uint64_t seed) {
    for (auto& s : s_) s = splitmix64(&seed);
}

uint64_t IdGenerator::next() {
    while (true) {
        const uint64_t result = rotl(s_[1] * 5, 7) * 9;
        const uint64_t t = s_[1] << 17;
        s_[2] ^= s_[0];
        s_[3] ^= s_[1];
        s_[1] ^= s_[2];
        s_[0] ^= s_[3];
        s_[2] ^= t;
        s_[3] = rotl(s_[3], 45);
        if (result != 0) return result;
    }
}

TraceContext start_server_trace(const B3Headers& in, IdGenerator* ids) {
    TraceContext ctx;
    if (in.has_sampled) {
        if (in.sampled == "0" || ascii_iequals(in.sampled, "false")) return ctx;  // not sampled
        uint64_t trace = 0, span = 0;
        if (parse_trace_id(in.trace_id, &trace) && parse_trace_id(in.span_id, &span)) {
            ctx.sampled = true;
            ctx.joined = true;
            ctx.trace_id = trace;
            ctx.span_id = span;
            uint64_t parent = 0;
            if (parse_trace_id(in.parent_span_id, &parent)) {
                ctx.parent_id = parent;
                ctx.has_parent = true;
            }
            return ctx;
        }
    }
    ctx.sampled = true;  // the reference samples every request
    ctx.span_id = ids->next();
    ctx.trace_id = ctx.span_id;
    return ctx;
}

TraceContext start_client_trace(const TraceContext& server, IdGenerator* ids) {
    TraceContext ctx;
    if (!server.sampled) return ctx;
    ctx.sampled = true;
    ctx.trace_id = server.trace_id;
    ctx.parent_id = server.span_id;
    ctx.has_parent = true;
    ctx.span_id = ids->next();
    return ctx;
}

namespace {

void append_endpoint(std::string* out, const Endpoint& ep) {
    out->append("\"endpoint\":{\"serviceName\":");
    json_append_string(out, ep.service_name);
    if (ep.ipv4 != 0) {
        char buf[INET_ADDRSTRLEN];
        in_addr a;
        a.s_addr = htonl(ep.ipv4);
        inet_ntop(AF_INET, &a, buf, sizeof buf);
        out->append(",\"ipv4\":\"");
        out->append(buf);
        out->push_back('"');
    }
    if (ep.port != 0) {
        out->append(",\"port\":");
        out->append(std::to_string(ep.port));
    }
    out->push_back('}');
}

void append_id(std::string* out, const char* field, uint64_t id) {
    char buf[16];
    format_trace_id(id, buf);
    out->push_back('"');
    out->append(field);
    out->append("\":\"");
    out->append(buf, 16);
    out->push_back('"');
}

void append_annotation(std::string* out, int64_t ts, const char* value, const Endpoint& ep) {
    out->append("{\"timestamp\":");
    out->append(std::to_string(ts));
    out->append(",\"value\":\"");
    out->append(value);
    out->append("\",");
    append_endpoint(out, ep);
    out->push_back('}');
}

}  // namespace

void span_append_json(std::string* out, const Span& span, const Endpoint& endpoint) {
    out->push_back('{');
    append_id(out, "traceId", span.trace_id);
    out->push_back(',');
    append_id(out, "id", span.id);
    out->append(",\"name\":");
    json_append_string(out, span.name);
    if (span.has_parent) {
        out->push_back(',');
        append_id(out, "parentId", span.parent_id);
    }
    if (span.timestamp_us != 0) {
        out->append(",\"timestamp\":");
        out->append(std::to_string(span.timestamp_us));
        out->append(",\"duration\":");
        out->append(std::to_string(span.duration_us > 0 ? span.duration_us : 1));
    }
    out->append(",\"annotations\":[");
    append_annotation(out, span.first_us, span.first_annotation, endpoint);
    out->push_back(',');
    append_annotation(out, span.last_us, span.last_annotation, endpoint);
    out->append("],\"binaryAnnotations\":[");
    bool first = true;
    for (const auto& tag : span.tags) {
        if (!first) out->push_back(',');
        first = false;
        out->append("{\"key\":");
        json_append_string(out, tag.key);
        out->append(",\"value\":");
        json_append_string(out, tag.value);
        out->push_back(',');
        append_endpoint(out, endpoint);
        out->push_back('}');
    }
    out->append("]}");
}

uint32_t local_ipv4() {
    uint32_t result = 0x7F000001u;
    ifaddrs* list = nullptr;
    if (getifaddrs(&list) != 0) return result;
    for (ifaddrs* it = list; it; it = it->ifa_next) {
        if (!it->ifa_addr || it->ifa_addr->sa_family != AF_INET) continue;
        if ((it->ifa_flags & IFF_LOOPBACK) || !(it->ifa_flags & IFF_UP)) continue;
        const auto* sin = reinterpret_cast<const sockaddr_in*>(it->ifa_addr);
        result = ntohl(sin->sin_addr.s_addr);
        break;
    }
    freeifaddrs(list);
    return result;
}

}  // namespace hola
