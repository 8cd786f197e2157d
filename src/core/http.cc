#include "core/http.h"

#include <cstring>
#include <ctime>

#include "core/text.h"

namespace hola {
namespace {

// RFC 7230 tchar, as a lookup table: header names are checked byte by byte on every request.
struct TokenTable {
    bool ok[256] = {};
    constexpr TokenTable() {
        for (int c = '0'; c <= '9'; ++c) ok[c] = true;
        for (int c = 'a'; c <= 'z'; ++c) ok[c] = true;
        for (int c = 'A'; c <= 'Z'; ++c) ok[c] = true;
        for (char c : {'!', '#', '$', '%', '&', '\'', '*', '+', '-', '.', '^', '_', '`', '|', '~'}) ok[static_cast<unsigned char>(c)] = true;
    }
};
constexpr TokenTable kToken;

inline bool is_token_char(unsigned char c) { return kToken.ok[c]; }

Method classify_method(std::string_view m) {
    switch (m.size()) {
        case 3:
            if (m == "GET") return Method::kGet;
            if (m == "PUT") return Method::kPut;
            break;
        case 4:
            if (m == "HEAD") return Method::kHead;
            if (m == "POST") return Method::kPost;
            break;
        case 5:
            if (m == "PATCH") return Method::kPatch;
            if (m == "TRACE") return Method::kTrace;
            break;
        case 6:
            if (m == "DELETE") return Method::kDelete;
            break;
        case 7:
            if (m == "OPTIONS") return Method::kOptions;
            if (m == "CONNECT") return Method::kConnect;
            break;
        default:
            break;
    }
    return Method::kOther;
}

bool parse_content_length(std::string_view v, int64_t* out) {
    if (v.empty() || v.size() > 18) return false;
    int64_t n = 0;
    for (char c : v) {
        if (c < '0' || c > '9') return false;
        n = n * 10 + (c - '0');
    }
    *out = n;
    return true;
}

bool contains_token(std::string_view list, std::string_view token) {
    size_t i = 0;
    while (i <= list.size()) {
        size_t comma = list.find(',', i);
        if (comma == std::string_view::npos) comma = list.size();
        if (ascii_iequals(trim_ows(list.substr(i, comma - i)), token)) return true;
        i = comma + 1;
    }
    return false;
}

struct HeaderLine {
    std::string_view name;
    std::string_view value;
};

enum class Line { kHeader, kBlank, kIncomplete, kBad };

// Reads the line starting at data[*pos] (header or the blank line that ends the head) from the
// bytes available so far and advances *pos past its terminator. One pass: the end of the head is
// discovered while the headers are being read, not by a separate scan.
Line next_header(const char* data, size_t len, size_t* pos, HeaderLine* out) {
    const size_t i = *pos;
    if (i >= len) return Line::kIncomplete;
    if (data[i] == '\n') {
        *pos = i + 1;
        return Line::kBlank;
    }
    if (data[i] == '\r') {
        if (i + 1 >= len) return Line::kIncomplete;
        if (data[i + 1] == '\n') {
            *pos = i + 2;
            return Line::kBlank;
        }
        return Line::kBad;  // a header name cannot start with CR
    }
    const char* line = data + i;
    const char* nl = static_cast<const char*>(std::memchr(line, '\n', len - i));
    if (!nl) return Line::kIncomplete;
    size_t line_len = static_cast<size_t>(nl - line);
    *pos = i + line_len + 1;
    if (line_len > 0 && line[line_len - 1] == '\r') --line_len;
    const char* p = line;
    const char* line_end = line + line_len;
    while (p < line_end && is_token_char(static_cast<unsigned char>(*p))) ++p;
    if (p == line || p == line_end || *p != ':') return Line::kBad;
    out->name = std::string_view(line, static_cast<size_t>(p - line));
    out->value = trim_ows(std::string_view(p + 1, static_cast<size_t>(line_end - p - 1)));
    return Line::kHeader;
}

}  // namespace

ParseStatus parse_request_head(const char* data, size_t len, RequestHead* out) {
    // Tolerate stray CRLFs before the request line (RFC 7230 3.5).
    size_t skip = 0;
    while (skip < len && (data[skip] == '\r' || data[skip] == '\n')) ++skip;
    if (skip == len) {
        out->head_len = skip;  // caller may discard these bytes
        return ParseStatus::kIncomplete;
    }
    const char* base = data + skip;
    const size_t avail = len - skip;

    // Request line.
    const char* nl = static_cast<const char*>(std::memchr(base, '\n', avail));
    if (!nl) return ParseStatus::kIncomplete;
    *out = RequestHead();
    size_t line_len = static_cast<size_t>(nl - base);
    size_t pos = line_len + 1;
    if (line_len > 0 && base[line_len - 1] == '\r') --line_len;
    std::string_view line(base, line_len);
    const size_t sp1 = line.find(' ');
    if (sp1 == std::string_view::npos || sp1 == 0) return ParseStatus::kBad;
    const size_t sp2 = line.rfind(' ');
    if (sp2 == sp1) return ParseStatus::kBad;
    out->method_text = line.substr(0, sp1);
    for (char c : out->method_text) {
        if (!is_token_char(static_cast<unsigned char>(c))) return ParseStatus::kBad;
    }
    out->method = classify_method(out->method_text);
    std::string_view target = line.substr(sp1 + 1, sp2 - sp1 - 1);
    if (target.empty() || target.find(' ') != std::string_view::npos) return ParseStatus::kBad;
    const std::string_view version = line.substr(sp2 + 1);
    if (version == "HTTP/1.1") {
        out->version_minor = 1;
    } else if (version == "HTTP/1.0") {
        out->version_minor = 0;
        out->keep_alive = false;
    } else {
        return ParseStatus::kBad;
    }
    out->target = target;
    if (target[0] != '/' && target != "*") {
        // absolute-form: scheme://authority/path
        const size_t scheme = target.find("://");
        if (scheme == std::string_view::npos) return ParseStatus::kBad;
        const size_t slash = target.find('/', scheme + 3);
        target = slash == std::string_view::npos ? std::string_view("/") : target.substr(slash);
    }
    const size_t q = target.find('?');
    if (q == std::string_view::npos) {
        out->path = target;
    } else {
        out->path = target.substr(0, q);
        out->query = target.substr(q + 1);
    }
    out->raw_path = out->path;
    // One pass over a path that is a dozen bytes long: anything that canonicalise_path() would
    // change? (The common request - /api/hola - has nothing of the kind and keeps its view.)
    {
        const std::string_view p = out->path;
        for (size_t i = 0; i < p.size(); ++i) {
            const char c = p[i];
            if (c == '%' || c == ';') {
                out->path_is_canonical = false;
                break;
            }
            if (c == '/' && i + 1 < p.size() && (p[i + 1] == '/' || p[i + 1] == '.')) {
                // "//", or a segment that starts with a dot: "/.", "/..", but also "/.well-known" -
                // the slow path sorts those out.
                out->path_is_canonical = false;
                break;
            }
        }
    }

    // Headers.
    bool saw_content_length = false;
    HeaderLine h;
    while (true) {
        const Line r = next_header(base, avail, &pos, &h);
        if (r == Line::kBlank) break;
        if (r == Line::kIncomplete) return ParseStatus::kIncomplete;
        if (r == Line::kBad) return ParseStatus::kBad;
        switch (h.name.size()) {
            case 4:
                if (ascii_iequals(h.name, "host")) out->host = h.value;
                break;
            case 6:
                if (ascii_iequals(h.name, "accept")) {
                    out->accept = h.value;
                    out->has_accept = true;
                } else if (ascii_iequals(h.name, "expect")) {
                    if (ascii_iequals(h.value, "100-continue")) out->expect_continue = true;
                }
                break;
            case 10:
                if (ascii_iequals(h.name, "connection")) {
                    if (ascii_iequals(h.value, "keep-alive")) {  // by far the usual value
                        out->keep_alive = true;
                    } else if (contains_token(h.value, "close")) {
                        out->keep_alive = false;
                    } else if (contains_token(h.value, "keep-alive")) {
                        out->keep_alive = true;
                    }
                }
                break;
            case 11:
                if (ascii_iequals(h.name, "x-b3-spanid")) out->b3.span_id = h.value;
                break;
            case 12:
                if (ascii_iequals(h.name, "x-b3-traceid")) {
                    out->b3.trace_id = h.value;
                } else if (ascii_iequals(h.name, "x-b3-sampled")) {
                    out->b3.sampled = h.value;
                    out->b3.has_sampled = true;
                }
                break;
            case 14:
                if (ascii_iequals(h.name, "content-length")) {
                    int64_t n = 0;
                    if (!parse_content_length(h.value, &n)) return ParseStatus::kBad;
                    if (saw_content_length && n != out->content_length) return ParseStatus::kBad;
                    saw_content_length = true;
                    out->content_length = n;
                }
                break;
            case 17:
                if (ascii_iequals(h.name, "transfer-encoding")) {
                    if (contains_token(h.value, "chunked")) out->chunked = true;
                } else if (ascii_iequals(h.name, "x-b3-parentspanid")) {
                    out->b3.parent_span_id = h.value;
                }
                break;
            default:
                break;
        }
    }
    if (out->chunked) out->content_length = -1;  // RFC 7230 3.3.3: Transfer-Encoding wins
    out->head_len = skip + pos;
    return ParseStatus::kComplete;
}

ParseStatus parse_response_head(const char* data, size_t len, ResponseHead* out) {
    const char* nl = static_cast<const char*>(std::memchr(data, '\n', len));
    if (!nl) return ParseStatus::kIncomplete;
    *out = ResponseHead();
    size_t line_len = static_cast<size_t>(nl - data);
    size_t pos = line_len + 1;
    if (line_len > 0 && data[line_len - 1] == '\r') --line_len;
    std::string_view line(data, line_len);
    if (line.size() < 12 || line.substr(0, 7) != "HTTP/1.") return ParseStatus::kBad;
    if (line[7] == '1') {
        out->version_minor = 1;
    } else if (line[7] == '0') {
        out->version_minor = 0;
        out->keep_alive = false;
    } else {
        return ParseStatus::kBad;
    }
    if (line[8] != ' ') return ParseStatus::kBad;
    int status = 0;
    for (int i = 9; i < 12; ++i) {
        if (line[static_cast<size_t>(i)] < '0' || line[static_cast<size_t>(i)] > '9') return ParseStatus::kBad;
        status = status * 10 + (line[static_cast<size_t>(i)] - '0');
    }
    if (line.size() > 12 && line[12] != ' ') return ParseStatus::kBad;
    out->status = status;
    HeaderLine h;
    while (true) {
        const Line r = next_header(data, len, &pos, &h);
        if (r == Line::kBlank) break;
        if (r == Line::kIncomplete) return ParseStatus::kIncomplete;
        if (r == Line::kBad) return ParseStatus::kBad;
        switch (h.name.size()) {
            case 10:
                if (ascii_iequals(h.name, "connection")) {
                    if (ascii_iequals(h.value, "keep-alive")) {  // by far the usual value
                        out->keep_alive = true;
                    } else if (contains_token(h.value, "close")) {
                        out->keep_alive = false;
                    } else if (contains_token(h.value, "keep-alive")) {
                        out->keep_alive = true;
                    }
                }
                break;
            case 12:
                if (ascii_iequals(h.name, "content-type")) out->content_type = h.value;
                break;
            case 14:
                if (ascii_iequals(h.name, "content-length")) {
                    int64_t n = 0;
                    if (!parse_content_length(h.value, &n)) return ParseStatus::kBad;
                    out->content_length = n;
                }
                break;
            case 17:
                if (ascii_iequals(h.name, "transfer-encoding") && contains_token(h.value, "chunked")) out->chunked = true;
                break;
            default:
                break;
        }
    }
    if (out->chunked) out->content_length = -1;
    out->head_len = pos;
    return ParseStatus::kComplete;
}

size_t ChunkedDecoder::feed(const char* data, size_t len, std::string* body) {
    size_t i = 0;
    while (i < len && state_ != State::kDone && state_ != State::kError) {
        const char c = data[i];
        switch (state_) {
            case State::kSize:
                if (is_hex_digit(c)) {
                    if (++size_digits_ > 15) {
                        state_ = State::kError;
                        break;
                    }
                    remaining_ = (remaining_ << 4) | static_cast<uint64_t>(hex_value(c));
                    ++i;
                } else if (size_digits_ == 0) {
                    state_ = State::kError;
                } else if (c == '\r') {
                    state_ = State::kSizeLf;
                    ++i;
                } else if (c == '\n') {
                    state_ = remaining_ == 0 ? State::kTrailer : State::kData;
                    ++i;
                } else if (c == ';' || c == ' ' || c == '\t') {
                    state_ = State::kSizeExt;
                    ++i;
                } else {
                    state_ = State::kError;
                }
                break;
            case State::kSizeExt:
                if (c == '\r') {
                    state_ = State::kSizeLf;
                } else if (c == '\n') {
                    state_ = remaining_ == 0 ? State::kTrailer : State::kData;
                }
                ++i;
                break;
            case State::kSizeLf:
                if (c != '\n') {
                    state_ = State::kError;
                    break;
                }
                state_ = remaining_ == 0 ? State::kTrailer : State::kData;
                ++i;
                break;
            case State::kData: {
                const size_t take = static_cast<size_t>(std::min<uint64_t>(remaining_, len - i));
                if (body) body->append(data + i, take);
                i += take;
                remaining_ -= take;
                if (remaining_ == 0) state_ = State::kDataCr;
                break;
            }
            case State::kDataCr:
                if (c == '\r') {
                    state_ = State::kDataLf;
                    ++i;
                } else if (c == '\n') {
                    state_ = State::kSize;
                    size_digits_ = 0;
                    ++i;
                } else {
                    state_ = State::kError;
                }
                break;
            case State::kDataLf:
                if (c != '\n') {
                    state_ = State::kError;
                    break;
                }
                state_ = State::kSize;
                size_digits_ = 0;
                ++i;
                break;
            case State::kTrailer:
                // Start of a trailer line: an empty line ends the message.
                if (c == '\r') {
                    state_ = State::kTrailerLf;
                    trailer_line_empty_ = true;
                    ++i;
                } else if (c == '\n') {
                    state_ = State::kDone;
                    ++i;
                } else {
                    state_ = State::kTrailerLine;
                    ++i;
                }
                break;
            case State::kTrailerLine:
                if (c == '\n') state_ = State::kTrailer;
                ++i;
                break;
            case State::kTrailerLf:
                if (c != '\n') {
                    state_ = State::kError;
                    break;
                }
                state_ = State::kDone;
                ++i;
                break;
            case State::kDone:
            case State::kError:
                break;
        }
    }
    return i;
}

bool accept_allows(std::string_view accept_header, std::string_view type, std::string_view subtype) {
    accept_header = trim_ows(accept_header);
    if (accept_header.empty()) return true;
    size_t i = 0;
    while (i <= accept_header.size()) {
        size_t comma = accept_header.find(',', i);
        if (comma == std::string_view::npos) comma = accept_header.size();
        std::string_view item = trim_ows(accept_header.substr(i, comma - i));
        i = comma + 1;
        if (item.empty()) continue;
        // Split off parameters; look for q=0.
        std::string_view range = item;
        bool excluded = false;
        const size_t semi = item.find(';');
        if (semi != std::string_view::npos) {
            range = trim_ows(item.substr(0, semi));
            std::string_view params = item.substr(semi + 1);
            size_t p = 0;
            while (p <= params.size()) {
                size_t s = params.find(';', p);
                if (s == std::string_view::npos) s = params.size();
                std::string_view param = trim_ows(params.substr(p, s - p));
                p = s + 1;
                if (param.size() >= 2 && (param[0] == 'q' || param[0] == 'Q') && param[1] == '=') {
                    std::string_view q = trim_ows(param.substr(2));
                    bool zero = !q.empty();
                    for (char c : q) {
                        if (c != '0' && c != '.') zero = false;
                    }
                    excluded = zero;
                }
            }
        }
        if (excluded) continue;
        const size_t slash = range.find('/');
        if (slash == std::string_view::npos) {
            if (range == "*") return true;  // what RESTEasy makes of a bare "*"
            continue;
        }
        const std::string_view t = range.substr(0, slash);
        const std::string_view st = range.substr(slash + 1);
        const bool type_ok = t == "*" || ascii_iequals(t, type);
        const bool sub_ok = st == "*" || ascii_iequals(st, subtype);
        if (type_ok && sub_ok) return true;
    }
    return false;
}

void format_http_date(int64_t unix_seconds, char out[29]) {
    static const char kDays[] = "SunMonTueWedThuFriSat";
    static const char kMonths[] = "JanFebMarAprMayJunJulAugSepOctNovDec";
    std::tm tm{};
    const time_t t = static_cast<time_t>(unix_seconds);
    gmtime_r(&t, &tm);
    auto two = [](char* p, int v) {
        p[0] = static_cast<char>('0' + v / 10);
        p[1] = static_cast<char>('0' + v % 10);
    };
    std::memcpy(out, kDays + tm.tm_wday * 3, 3);
    out[3] = ',';
    out[4] = ' ';
    two(out + 5, tm.tm_mday);
    out[7] = ' ';
    std::memcpy(out + 8, kMonths + tm.tm_mon * 3, 3);
    out[11] = ' ';
    const int year = tm.tm_year + 1900;
    two(out + 12, year / 100);
    two(out + 14, year % 100);
    out[16] = ' ';
    two(out + 17, tm.tm_hour);
    out[19] = ':';
    two(out + 20, tm.tm_min);
    out[22] = ':';
    two(out + 23, tm.tm_sec);
    std::memcpy(out + 25, " GMT", 4);
}

bool percent_decode(std::string_view in, std::string* out) {
    out->clear();
    out->reserve(in.size());
    for (size_t i = 0; i < in.size(); ++i) {
        if (in[i] != '%') {
            out->push_back(in[i]);
            continue;
        }
        if (i + 2 >= in.size()) return false;
        if (!is_hex_digit(in[i + 1]) || !is_hex_digit(in[i + 2])) return false;
        out->push_back(static_cast<char>((hex_value(in[i + 1]) << 4) | hex_value(in[i + 2])));
        i += 2;
    }
    return true;
}

bool canonicalise_path(std::string_view raw, std::string* out) {
    out->clear();
    if (raw.empty() || raw[0] != '/') {  // "*" (OPTIONS *) and the like: nothing to canonicalise
        out->assign(raw);
        return true;
    }
    out->reserve(raw.size());
    std::string segment;
    size_t i = 1;
    bool trailing_slash = raw.size() > 1 && raw.back() == '/';
    while (i <= raw.size()) {
        size_t end = raw.find('/', i);
        if (end == std::string_view::npos) end = raw.size();
        std::string_view seg = raw.substr(i, end - i);
        const size_t semi = seg.find(';');
        if (semi != std::string_view::npos) seg = seg.substr(0, semi);
        segment.clear();
        for (size_t k = 0; k < seg.size(); ++k) {
            if (seg[k] != '%') {
                segment.push_back(seg[k]);
                continue;
            }
            if (k + 2 >= seg.size()) return false;
            if (!is_hex_digit(seg[k + 1]) || !is_hex_digit(seg[k + 2])) return false;
            const int v = (hex_value(seg[k + 1]) << 4) | hex_value(seg[k + 2]);
            if (v == 0) return false;
            if (v == '/' || v == '\\') {
                segment.append(seg.substr(k, 3));  // stays encoded
            } else {
                segment.push_back(static_cast<char>(v));
            }
            k += 2;
        }
        if (segment.empty() || segment == ".") {
            if (end == raw.size() && !segment.empty()) trailing_slash = true;  // "/a/." names the directory "/a/"
        } else if (segment == "..") {
            const size_t cut = out->rfind('/');
            if (cut != std::string::npos) out->resize(cut);
            if (end == raw.size()) trailing_slash = true;
        } else {
            out->push_back('/');
            out->append(segment);
        }
        i = end + 1;
    }
    if (out->empty() || (trailing_slash && out->back() != '/')) out->push_back('/');
    return true;
}

}  // namespace hola
