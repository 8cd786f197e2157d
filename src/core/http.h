// HTTP/1.x wire parsing: request heads (server side), response heads and chunked bodies
// (outbound client and load generator). Zero-copy: all views point into the caller's buffer.
//
// The reference gets its HTTP layer from Undertow inside WildFly Swarm
// (/root/reference/pom.xml:122-129); nothing in the repository itself parses HTTP. This is the
// from-scratch equivalent of that borrowed capability, sized for the service's needs: small
// requests, keep-alive, pipelining, and a handful of headers that change behaviour.
#pragma once

#include <cstddef>
#include <cstdint>
#include <string>
#include <string_view>

#include "core/trace.h"

namespace hola {

enum class Method : uint8_t { kGet, kHead, kPost, kPut, kDelete, kOptions, kPatch, kTrace, kConnect, kOther };

enum class ParseStatus : uint8_t { kComplete, kIncomplete, kBad };

struct RequestHead {
    Method method = Method::kOther;
    std::string_view method_text;
    std::string_view target;  // as sent
    std::string_view path;    // target up to '?', scheme://authority stripped for absolute-form; the server core
                              // replaces it with the canonical form (canonicalise_path) before routing
    std::string_view raw_path;  // always as sent
    bool path_is_canonical = true;  // false: contains '%', ';', '//' or a dot segment - canonicalise before routing
    std::string_view query;   // after '?', may be empty
    int version_minor = 1;
    bool keep_alive = true;
    bool chunked = false;
    bool expect_continue = false;
    bool has_accept = false;
    int64_t content_length = -1;
    std::string_view host;
    std::string_view accept;
    B3Headers b3;
    size_t head_len = 0;  // bytes up to and including the blank line
};

// Parses one request head from data[0..len). kIncomplete means "need more bytes".
ParseStatus parse_request_head(const char* data, size_t len, RequestHead* out);

struct ResponseHead {
    int status = 0;
    int version_minor = 1;
    bool keep_alive = true;
    bool chunked = false;
    int64_t content_length = -1;  // -1: not given (body runs to connection close unless chunked)
    std::string_view content_type;
    size_t head_len = 0;
};

ParseStatus parse_response_head(const char* data, size_t len, ResponseHead* out);

// True when a response to `request_method` with `status` never carries a body.
inline bool response_has_no_body(Method request_method, int status) {
    return request_method == Method::kHead || status == 204 || status == 304 || (status >= 100 && status < 200);
}

// Incremental decoder for Transfer-Encoding: chunked.
class ChunkedDecoder {
 public:
    enum class State : uint8_t { kSize, kSizeExt, kSizeLf, kData, kDataCr, kDataLf, kTrailer, kTrailerLine, kTrailerLf, kDone, kError };
    // Consumes bytes from data[0..len); decoded payload is appended to *body when body is not
    // null. Returns the number of bytes consumed (stops early once the final chunk is done).
    size_t feed(const char* data, size_t len, std::string* body);
    bool done() const { return state_ == State::kDone; }
    bool failed() const { return state_ == State::kError; }
    void reset() { *this = ChunkedDecoder(); }

 private:
    State state_ = State::kSize;
    uint64_t remaining_ = 0;
    int size_digits_ = 0;
    bool trailer_line_empty_ = true;
};

// Does the Accept header admit the given media type ("type/subtype")? An absent or empty header
// admits everything. q-values of 0 exclude.
bool accept_allows(std::string_view accept_header, std::string_view type, std::string_view subtype);

// RFC 7231 IMF-fixdate, 29 bytes: "Sun, 06 Nov 1994 08:49:37 GMT".
void format_http_date(int64_t unix_seconds, char out[29]);

// Percent-decodes a path. Returns false on malformed escapes.
bool percent_decode(std::string_view in, std::string* out);

// The path a request is routed by, from the path as sent. The reference's container does this
// before any resource is matched: Undertow percent-decodes the request path and resolves "." and
// ".." segments and empty segments, and path (matrix) parameters - ";name=value" at the end of a
// segment - are not part of the path RESTEasy matches (@ApplicationPath/@Path,
// /root/reference/src/main/java/com/redhat/developers/msa/hola/JaxRsActivator.java:30,
// HolaResource.java:46,52-53,69-70,80-81). So /api/%68ola, /api/hola;v=1 and /api/../api/hola are
// all GET /api/hola.
//   1. per segment: cut at the first ';'
//   2. percent-decode, except %2F and %5C (an encoded slash stays encoded, as with Undertow's
//      default allow-encoded-slash=false, so it can never introduce a segment boundary)
//   3. drop "." and empty segments, let ".." remove the segment before it (never above the root)
// A trailing slash is kept. Returns false - the request is answered 400 - for a malformed escape
// or a decoded NUL byte.
bool canonicalise_path(std::string_view raw, std::string* out);

}  // namespace hola
