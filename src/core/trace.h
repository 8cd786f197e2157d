// Zipkin / B3 tracing primitives: header parsing, span model, Zipkin v1 JSON encoding.
//
// The reference traces two URLs with Brave 3.9.1: a servlet filter opens a server span for
// /api/hola and /api/hola-chaining
// (/root/reference/src/main/java/com/redhat/developers/msa/hola/ZipKinFilter.java:38-44), the
// tracer is named "hola" and ships spans over HTTP to $ZIPKIN_SERVER_URL
// (/root/reference/src/main/java/com/redhat/developers/msa/hola/ZipkinResource.java:33-41), and
// the call to `aloha` carries a child client span plus B3 headers
// (/root/reference/src/main/java/com/redhat/developers/msa/hola/HolaResource.java:97-108).
//
// Rules taken from Brave 3.9's HTTP adapters:
//   * `X-B3-Sampled: 0|false` switches tracing off for the request; the downstream call then
//     carries only `X-B3-Sampled: 0`;
//   * inbound ids are honoured only when X-B3-Sampled is present (and true) together with both
//     X-B3-TraceId and X-B3-SpanId; otherwise a new trace starts, trace id == span id;
//   * server spans are annotated sr/ss and tagged http.url (and http.status_code for non-2xx);
//     client spans are annotated cs/cr likewise; span names are the HTTP method.
#pragma once

#include <cstdint>
#include <string>
#include <string_view>
#include <vector>

namespace hola {

// Parses a lower/upper-case hex id. Ids longer than 16 digits keep their low 64 bits.
bool parse_trace_id(std::string_view hex, uint64_t* out);
// Writes exactly 16 lower-case hex digits.
void format_trace_id(uint64_t id, char out[16]);
std::string trace_id_string(uint64_t id);

struct B3Headers {
    std::string_view trace_id;
    std::string_view span_id;
    std::string_view parent_span_id;
    std::string_view sampled;
    bool has_sampled = false;
};

struct TraceContext {
    bool sampled = false;   // false: the request is not traced at all
    bool joined = false;    // ids were supplied by the caller
    uint64_t trace_id = 0;
    uint64_t span_id = 0;
    uint64_t parent_id = 0;
    bool has_parent = false;
};

// Small, fast, per-thread id source (xoshiro256**), seeded from the OS.
class IdGenerator {
 public:
    IdGenerator();
    explicit IdGenerator(uint64_t seed);
    uint64_t next();  // never returns 0

 private:
    uint64_t s_[4];
};

// Server-side decision for an inbound request.
TraceContext start_server_trace(const B3Headers& in, IdGenerator* ids);
// Child context for an outbound call made while handling `server`.
TraceContext start_client_trace(const TraceContext& server, IdGenerator* ids);

struct Endpoint {
    std::string service_name;
    uint32_t ipv4 = 0;  // host byte order; 0 = omit
    uint16_t port = 0;  // 0 = omit
};

struct BinaryAnnotation {
    std::string key;
    std::string value;
};

struct Span {
    uint64_t trace_id = 0;
    uint64_t id = 0;
    uint64_t parent_id = 0;
    bool has_parent = false;
    const char* name = "get";
    int64_t timestamp_us = 0;  // 0 = not reported (span shares its ids with the caller's)
    int64_t duration_us = 0;
    const char* first_annotation = "sr";
    int64_t first_us = 0;
    const char* last_annotation = "ss";
    int64_t last_us = 0;
    std::vector<BinaryAnnotation> tags;
};

// Appends one span as a Zipkin v1 JSON object.
void span_append_json(std::string* out, const Span& span, const Endpoint& endpoint);

// Best-effort primary IPv4 address of this host (first non-loopback interface), host byte
// order; falls back to 127.0.0.1.
uint32_t local_ipv4();

}  // namespace hola
