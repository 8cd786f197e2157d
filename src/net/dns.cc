#include "net/dns.h"

#include <arpa/inet.h>
#include <netinet/in.h>
#include <poll.h>
#include <sys/random.h>
#include <sys/socket.h>
#include <unistd.h>

#include <algorithm>
#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <sstream>

#include "core/text.h"
#include "net/event_loop.h"  // monotonic_ns

namespace hola {

namespace {

std::string read_file(const std::string& path) {
    std::string out;
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return out;
    char buf[4096];
    size_t n;
    while ((n = std::fread(buf, 1, sizeof buf, f)) > 0) out.append(buf, n);
    std::fclose(f);
    return out;
}

std::vector<std::string> split_ws(const std::string& line) {
    std::vector<std::string> out;
    std::istringstream is(line);
    std::string tok;
    while (is >> tok) out.push_back(tok);
    return out;
}

std::string lower(std::string s) {
    for (char& c : s) {
        if (c >= 'A' && c <= 'Z') c = static_cast<char>(c + 32);
    }
    return s;
}

}  // namespace

DnsConfig DnsConfig::parse(const std::string& text) {
    DnsConfig cfg;
    std::istringstream is(text);
    std::string line;
    while (std::getline(is, line)) {
        const size_t hash = line.find_first_of("#;");
        if (hash != std::string::npos) line.erase(hash);
        const std::vector<std::string> tok = split_ws(line);
        if (tok.empty()) continue;
        if (tok[0] == "nameserver" && tok.size() >= 2) {
            if (cfg.nameservers.size() < 3) cfg.nameservers.push_back(tok[1]);
        } else if ((tok[0] == "search" || tok[0] == "domain") && tok.size() >= 2) {
            cfg.search.assign(tok.begin() + 1, tok.end());  // the last search/domain line wins
        } else if (tok[0] == "options") {
            for (size_t i = 1; i < tok.size(); ++i) {
                const std::string& o = tok[i];
                if (o.compare(0, 6, "ndots:") == 0) cfg.ndots = std::max(0, std::min(15, std::atoi(o.c_str() + 6)));
                if (o.compare(0, 8, "timeout:") == 0) cfg.timeout_ms = std::max(1, std::min(30, std::atoi(o.c_str() + 8))) * 1000;
                if (o.compare(0, 9, "attempts:") == 0) cfg.attempts = std::max(1, std::min(5, std::atoi(o.c_str() + 9)));
            }
        }
    }
    if (cfg.nameservers.empty()) cfg.nameservers.push_back("127.0.0.1");
    return cfg;
}

DnsConfig DnsConfig::load_system() {
    const char* path = std::getenv("HOLA_RESOLV_CONF");
    DnsConfig cfg = parse(read_file(path && *path ? path : "/etc/resolv.conf"));
    if (const char* hosts = std::getenv("HOLA_HOSTS_FILE")) cfg.hosts_file = hosts;
    if (const char* port = std::getenv("HOLA_DNS_PORT")) {
        const int p = std::atoi(port);
        if (p > 0 && p < 65536) cfg.port = static_cast<uint16_t>(p);
    }
    if (const char* dom = std::getenv("LOCALDOMAIN")) cfg.search = split_ws(dom);
    return cfg;
}

IpAddress IpAddress::v4(uint32_t network_order) {
    IpAddress a;
    a.family = AF_INET;
    std::memcpy(a.bytes, &network_order, 4);
    return a;
}

IpAddress IpAddress::parse(const std::string& text) {
    IpAddress a;
    if (inet_pton(AF_INET, text.c_str(), a.bytes) == 1) {
        a.family = AF_INET;
    } else if (inet_pton(AF_INET6, text.c_str(), a.bytes) == 1) {
        a.family = AF_INET6;
    }
    return a;
}

std::string IpAddress::to_string() const {
    char buf[INET6_ADDRSTRLEN] = "";
    if (family == AF_INET || family == AF_INET6) inet_ntop(family, bytes, buf, sizeof buf);
    return buf;
}

uint32_t IpAddress::v4_network_order() const {
    uint32_t v = 0;
    if (family == AF_INET) std::memcpy(&v, bytes, 4);
    return v;
}

socklen_t IpAddress::to_sockaddr(uint16_t port, sockaddr_storage* out) const {
    std::memset(out, 0, sizeof *out);
    if (family == AF_INET6) {
        auto* a = reinterpret_cast<sockaddr_in6*>(out);
        a->sin6_family = AF_INET6;
        a->sin6_port = htons(port);
        std::memcpy(&a->sin6_addr, bytes, 16);
        return sizeof *a;
    }
    auto* a = reinterpret_cast<sockaddr_in*>(out);
    a->sin_family = AF_INET;
    a->sin_port = htons(port);
    std::memcpy(&a->sin_addr, bytes, 4);
    return sizeof *a;
}

bool IpAddress::operator==(const IpAddress& o) const {
    return family == o.family && std::memcmp(bytes, o.bytes, family == AF_INET6 ? 16 : 4) == 0;
}

bool hosts_file_lookup(const std::string& path, const std::string& host, IpAddress* out) {
    const std::string text = read_file(path);
    const std::string want = lower(host);
    std::istringstream is(text);
    std::string line;
    IpAddress v6_match;
    while (std::getline(is, line)) {
        const size_t hash = line.find('#');
        if (hash != std::string::npos) line.erase(hash);
        const std::vector<std::string> tok = split_ws(line);
        if (tok.size() < 2) continue;
        const IpAddress addr = IpAddress::parse(tok[0]);
        if (addr.family == 0) continue;
        for (size_t i = 1; i < tok.size(); ++i) {
            if (lower(tok[i]) != want) continue;
            if (addr.family == AF_INET) {
                *out = addr;
                return true;
            }
            if (v6_match.family == 0) v6_match = addr;
        }
    }
    if (v6_match.family != 0) {
        *out = v6_match;
        return true;
    }
    return false;
}

std::vector<std::string> dns_candidates(const std::string& host, const DnsConfig& cfg) {
    std::vector<std::string> out;
    if (host.empty()) return out;
    if (host.back() == '.') {  // rooted name: no search list
        out.push_back(host.substr(0, host.size() - 1));
        return out;
    }
    const int dots = static_cast<int>(std::count(host.begin(), host.end(), '.'));
    const bool as_is_first = dots >= cfg.ndots;
    if (as_is_first) out.push_back(host);
    for (const std::string& domain : cfg.search) {
        if (domain.empty() || domain == ".") continue;
        out.push_back(host + "." + (domain.back() == '.' ? domain.substr(0, domain.size() - 1) : domain));
    }
    if (!as_is_first) out.push_back(host);
    return out;
}

bool dns_build_query(const std::string& name, uint16_t id, std::vector<uint8_t>* out, uint16_t qtype) {
    out->clear();
    out->reserve(name.size() + 18);
    auto put16 = [&](uint16_t v) {
        out->push_back(static_cast<uint8_t>(v >> 8));
        out->push_back(static_cast<uint8_t>(v & 0xFF));
    };
    put16(id);
    put16(0x0100);  // standard query, recursion desired
    put16(1);       // one question
    put16(0);
    put16(0);
    put16(0);
    size_t start = 0;
    while (start <= name.size()) {
        size_t dot = name.find('.', start);
        if (dot == std::string::npos) dot = name.size();
        const size_t len = dot - start;
        if (len == 0 || len > 63) return false;
        out->push_back(static_cast<uint8_t>(len));
        out->insert(out->end(), name.begin() + static_cast<long>(start), name.begin() + static_cast<long>(dot));
        start = dot + 1;
        if (dot == name.size()) break;
    }
    if (out->size() - 12 > 254) return false;
    out->push_back(0);
    put16(qtype);
    put16(1);  // QCLASS IN
    return true;
}

namespace {

// Skips a (possibly compressed) name. Returns the offset after it, or 0 on malformed input.
size_t skip_name(const uint8_t* d, size_t len, size_t pos) {
    while (pos < len) {
        const uint8_t c = d[pos];
        if (c == 0) return pos + 1;
        if ((c & 0xC0) == 0xC0) return pos + 2 <= len ? pos + 2 : 0;
        if (c & 0xC0) return 0;
        pos += 1 + c;
    }
    return 0;
}

// Expands the name at `pos` into lower-case dotted text. Returns false on malformed input.
bool read_name(const uint8_t* d, size_t len, size_t pos, std::string* out) {
    out->clear();
    int jumps = 0;
    while (pos < len) {
        const uint8_t c = d[pos];
        if (c == 0) return true;
        if ((c & 0xC0) == 0xC0) {
            if (pos + 1 >= len || ++jumps > 16) return false;
            pos = static_cast<size_t>((c & 0x3F) << 8) | d[pos + 1];
            continue;
        }
        if (c & 0xC0) return false;
        if (pos + 1 + c > len) return false;
        if (!out->empty()) out->push_back('.');
        for (size_t i = 0; i < c; ++i) {
            char ch = static_cast<char>(d[pos + 1 + i]);
            if (ch >= 'A' && ch <= 'Z') ch = static_cast<char>(ch + 32);
            out->push_back(ch);
        }
        pos += 1 + c;
    }
    return false;
}

}  // namespace

DnsStatus dns_parse_response(const uint8_t* d, size_t len, uint16_t id, IpAddress* out, bool* truncated, uint16_t qtype,
                             const std::string* asked_name) {
    if (truncated) *truncated = false;
    if (len < 12) return DnsStatus::kError;
    if (static_cast<uint16_t>((d[0] << 8) | d[1]) != id) return DnsStatus::kError;
    if (!(d[2] & 0x80)) return DnsStatus::kError;  // not a response
    if (asked_name) {
        // The 16-bit id is all that ties a datagram to a query otherwise. The question section of
        // a genuine response repeats what was asked: one question, same name (case-insensitive),
        // same type, class IN. Anything else answers some other question - also when its rcode
        // says NXDOMAIN - and must not be taken for the answer to this one.
        if (((d[4] << 8) | d[5]) != 1) return DnsStatus::kMismatch;
        std::string qname;
        if (!read_name(d, len, 12, &qname)) return DnsStatus::kMismatch;
        const size_t after = skip_name(d, len, 12);
        if (after == 0 || after + 4 > len) return DnsStatus::kMismatch;
        std::string want = *asked_name;
        while (!want.empty() && want.back() == '.') want.pop_back();
        for (char& ch : want) {
            if (ch >= 'A' && ch <= 'Z') ch = static_cast<char>(ch + 32);
        }
        if (qname != want) return DnsStatus::kMismatch;
        if (((d[after] << 8) | d[after + 1]) != qtype || ((d[after + 2] << 8) | d[after + 3]) != 1) return DnsStatus::kMismatch;
    }
    if (d[2] & 0x02) {
        if (truncated) *truncated = true;
    }
    const int rcode = d[3] & 0x0F;
    if (rcode == 3) return DnsStatus::kNotFound;  // NXDOMAIN
    if (rcode != 0) return DnsStatus::kError;     // SERVFAIL, REFUSED, ...
    const int qd = (d[4] << 8) | d[5];
    const int an = (d[6] << 8) | d[7];
    size_t pos = 12;
    std::string target;
    for (int i = 0; i < qd; ++i) {
        if (i == 0 && !read_name(d, len, pos, &target)) return DnsStatus::kError;
        pos = skip_name(d, len, pos);
        if (pos == 0 || pos + 4 > len) return DnsStatus::kError;
        pos += 4;
    }
    // Walk the answers; CNAMEs retarget the name being looked for. A second pass covers servers
    // that list the address record before the alias it belongs to.
    for (int pass = 0; pass < 2; ++pass) {
        size_t p = pos;
        for (int i = 0; i < an; ++i) {
            std::string owner;
            if (!read_name(d, len, p, &owner)) return DnsStatus::kError;
            p = skip_name(d, len, p);
            if (p == 0 || p + 10 > len) return DnsStatus::kError;
            const int type = (d[p] << 8) | d[p + 1];
            const int klass = (d[p + 2] << 8) | d[p + 3];
            const size_t rdlen = static_cast<size_t>((d[p + 8] << 8) | d[p + 9]);
            p += 10;
            if (p + rdlen > len) return DnsStatus::kError;
            if (klass == 1 && owner == target) {
                if (type == qtype && ((qtype == kDnsTypeA && rdlen == 4) || (qtype == kDnsTypeAAAA && rdlen == 16))) {
                    out->family = qtype == kDnsTypeA ? AF_INET : AF_INET6;
                    std::memcpy(out->bytes, d + p, rdlen);
                    return DnsStatus::kOk;
                }
                if (type == 5) {
                    std::string alias;
                    if (!read_name(d, len, p, &alias)) return DnsStatus::kError;
                    target = alias;
                }
            }
            p += rdlen;
        }
    }
    return DnsStatus::kNoData;  // the name exists but has no record of this type
}

namespace {

struct ServerAddr {
    sockaddr_storage ss{};
    socklen_t len = 0;
    int family = AF_INET;
};

bool make_server_addr(const std::string& text, uint16_t port, ServerAddr* out) {
    auto* v4 = reinterpret_cast<sockaddr_in*>(&out->ss);
    if (inet_pton(AF_INET, text.c_str(), &v4->sin_addr) == 1) {
        v4->sin_family = AF_INET;
        v4->sin_port = htons(port);
        out->len = sizeof(sockaddr_in);
        out->family = AF_INET;
        return true;
    }
    auto* v6 = reinterpret_cast<sockaddr_in6*>(&out->ss);
    std::memset(v6, 0, sizeof *v6);
    if (inet_pton(AF_INET6, text.c_str(), &v6->sin6_addr) == 1) {
        v6->sin6_family = AF_INET6;
        v6->sin6_port = htons(port);
        out->len = sizeof(sockaddr_in6);
        out->family = AF_INET6;
        return true;
    }
    return false;
}

int remaining_ms(int64_t deadline_ns, int cap_ms) {
    if (deadline_ns == 0) return cap_ms;
    const int64_t left = (deadline_ns - monotonic_ns()) / 1000000;
    if (left <= 0) return 0;
    return static_cast<int>(std::min<int64_t>(left, cap_ms));
}

// One query/response exchange over UDP (or TCP when `tcp`). Returns kTimeout when nothing usable
// arrived in time.
DnsStatus exchange(const ServerAddr& server, const std::string& name, const std::vector<uint8_t>& query, uint16_t id,
                   uint16_t qtype, bool tcp, int timeout_ms, IpAddress* out, bool* truncated) {
    const int fd = socket(server.family, (tcp ? SOCK_STREAM : SOCK_DGRAM) | SOCK_CLOEXEC | SOCK_NONBLOCK, 0);
    if (fd < 0) return DnsStatus::kError;
    DnsStatus result = DnsStatus::kTimeout;
    const int64_t until = monotonic_ns() + static_cast<int64_t>(timeout_ms) * 1000000;
    auto wait = [&](short events) {
        const int64_t left = (until - monotonic_ns()) / 1000000;
        if (left <= 0) return false;
        pollfd p{fd, events, 0};
        return poll(&p, 1, static_cast<int>(left)) > 0;
    };
    do {
        if (connect(fd, reinterpret_cast<const sockaddr*>(&server.ss), server.len) != 0) {
            if (errno != EINPROGRESS) {
                result = DnsStatus::kError;
                break;
            }
            if (!wait(POLLOUT)) break;
            int err = 0;
            socklen_t l = sizeof err;
            if (getsockopt(fd, SOL_SOCKET, SO_ERROR, &err, &l) != 0 || err != 0) {
                result = DnsStatus::kError;
                break;
            }
        }
        std::vector<uint8_t> wire;
        if (tcp) {
            wire.push_back(static_cast<uint8_t>(query.size() >> 8));
            wire.push_back(static_cast<uint8_t>(query.size() & 0xFF));
        }
        wire.insert(wire.end(), query.begin(), query.end());
        if (send(fd, wire.data(), wire.size(), MSG_NOSIGNAL) != static_cast<ssize_t>(wire.size())) {
            result = DnsStatus::kError;
            break;
        }
        uint8_t buf[4096];
        size_t have = 0;
        while (wait(POLLIN)) {
            const ssize_t n = recv(fd, buf + have, sizeof buf - have, 0);
            if (n < 0 && (errno == EAGAIN || errno == EINTR)) continue;
            if (n <= 0) {
                result = DnsStatus::kError;  // ICMP port unreachable lands here as ECONNREFUSED
                break;
            }
            if (!tcp) {
                const DnsStatus st = dns_parse_response(buf, static_cast<size_t>(n), id, out, truncated, qtype, &name);
                if (st == DnsStatus::kMismatch || (st == DnsStatus::kError && static_cast<size_t>(n) >= 2 &&
                                                   static_cast<uint16_t>((buf[0] << 8) | buf[1]) != id)) {
                    if (truncated) *truncated = false;
                    continue;  // a stray datagram for some other query: keep waiting
                }
                result = st;
                break;
            }
            have += static_cast<size_t>(n);
            if (have < 2) continue;
            const size_t msg_len = static_cast<size_t>((buf[0] << 8) | buf[1]);
            if (msg_len + 2 > sizeof buf) {
                result = DnsStatus::kError;
                break;
            }
            if (have < msg_len + 2) continue;
            result = dns_parse_response(buf + 2, msg_len, id, out, nullptr, qtype, &name);
            if (result == DnsStatus::kMismatch) result = DnsStatus::kError;
            break;
        }
    } while (false);
    close(fd);
    return result;
}

}  // namespace

DnsStatus resolve_address(const std::string& host, const DnsConfig& cfg, IpAddress* out, int deadline_ms) {
    const IpAddress literal = IpAddress::parse(host);
    if (literal.family != 0) {
        *out = literal;
        return DnsStatus::kOk;
    }
    if (ascii_iequals(host, "localhost")) {
        *out = IpAddress::v4(htonl(INADDR_LOOPBACK));
        return DnsStatus::kOk;
    }
    if (hosts_file_lookup(cfg.hosts_file, host, out)) return DnsStatus::kOk;

    const int64_t deadline_ns = deadline_ms > 0 ? monotonic_ns() + static_cast<int64_t>(deadline_ms) * 1000000 : 0;
    std::vector<ServerAddr> servers;
    for (const std::string& ns : cfg.nameservers) {
        ServerAddr a;
        if (make_server_addr(ns, cfg.port, &a)) servers.push_back(a);
    }
    if (servers.empty()) return DnsStatus::kError;

    bool any_timeout = false;
    bool any_error = false;
    for (const std::string& name : dns_candidates(host, cfg)) {
        // IPv4 first; a name that exists only as AAAA (an IPv6-only cluster) is found by the
        // second question.
        for (const uint16_t qtype : {kDnsTypeA, kDnsTypeAAAA}) {
            uint16_t id = 0;
            if (getrandom(&id, sizeof id, GRND_NONBLOCK) != static_cast<ssize_t>(sizeof id))
                id = static_cast<uint16_t>(monotonic_ns() & 0xFFFF);
            std::vector<uint8_t> query;
            if (!dns_build_query(name, id, &query, qtype)) break;
            bool answered = false;  // some server said "no such name" or "no such record"
            bool try_other_family = false;
            for (int attempt = 0; attempt < cfg.attempts && !answered; ++attempt) {
                for (const ServerAddr& server : servers) {
                    const int timeout_ms = remaining_ms(deadline_ns, cfg.timeout_ms);
                    if (timeout_ms <= 0) return DnsStatus::kTimeout;
                    bool truncated = false;
                    DnsStatus st = exchange(server, name, query, id, qtype, false, timeout_ms, out, &truncated);
                    if (truncated && st != DnsStatus::kOk) {
                        const int t2 = remaining_ms(deadline_ns, cfg.timeout_ms);
                        if (t2 > 0) st = exchange(server, name, query, id, qtype, true, t2, out, nullptr);
                    }
                    if (st == DnsStatus::kOk) return DnsStatus::kOk;
                    if (st == DnsStatus::kNotFound || st == DnsStatus::kNoData) {
                        answered = true;  // authoritative "no"
                        try_other_family = st == DnsStatus::kNoData;
                        break;
                    }
                    if (st == DnsStatus::kTimeout) any_timeout = true;
                    if (st == DnsStatus::kError) any_error = true;
                }
            }
            // NXDOMAIN covers every record type, and silence will not turn into an answer by
            // asking for AAAA: in both cases go on to the next candidate name.
            if (!try_other_family) break;
        }
    }
    if (any_timeout) return DnsStatus::kTimeout;
    return any_error ? DnsStatus::kError : DnsStatus::kNotFound;
}

}  // namespace hola
