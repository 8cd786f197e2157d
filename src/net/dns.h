// A small stub resolver: /etc/hosts, then A (and, failing that, AAAA) queries to the name servers
// of /etc/resolv.conf with search-list and ndots handling.
//
// Why not getaddrinfo(): the service is shipped as one fully static executable in a FROM-scratch
// image (the analogue of the reference's self-contained uber-jar, /root/reference/pom.xml:230-241,
// /root/reference/Dockerfile:10). A statically linked glibc still dlopen()s its NSS plug-ins at
// run time, and in an empty image there are none - every lookup of the downstream service name
// "aloha" (/root/reference/src/main/java/com/redhat/developers/msa/hola/HolaResource.java:95,103)
// or of the Zipkin host (/root/reference/Dockerfile:6) would fail. In Kubernetes, "aloha" only
// resolves through the search list in resolv.conf ("<ns>.svc.cluster.local ..."), so that logic
// has to be here too.
//
// The lookup is blocking and meant for helper threads (net/resolver.h, app/tracer.cc); the event
// loops never call it.
#pragma once

#include <sys/socket.h>

#include <cstdint>
#include <string>
#include <vector>

namespace hola {

// An IPv4 or IPv6 address in network byte order.
struct IpAddress {
    int family = 0;  // AF_INET, AF_INET6, or 0 for "none"
    uint8_t bytes[16] = {0};

    static IpAddress v4(uint32_t network_order);
    // Parses a textual address ("10.1.2.3", "fd00::1"). family stays 0 when it is neither.
    static IpAddress parse(const std::string& text);
    std::string to_string() const;
    uint32_t v4_network_order() const;
    // Fills `out` for connect()/bind(); returns the length to pass along.
    socklen_t to_sockaddr(uint16_t port, sockaddr_storage* out) const;
    bool operator==(const IpAddress& o) const;
};

struct DnsConfig {
    std::vector<std::string> nameservers;  // textual IPv4/IPv6 addresses
    std::vector<std::string> search;
    int ndots = 1;
    int timeout_ms = 5000;  // per try (resolv.conf "options timeout:N", seconds)
    int attempts = 2;
    std::string hosts_file = "/etc/hosts";
    uint16_t port = 53;     // overridable for tests (HOLA_DNS_PORT)

    // Parses resolv.conf text. Missing name servers default to 127.0.0.1, as in libc.
    static DnsConfig parse(const std::string& resolv_conf_text);
    // Loads the system configuration. Environment overrides for tests and odd deployments:
    // HOLA_RESOLV_CONF, HOLA_HOSTS_FILE, HOLA_DNS_PORT; LOCALDOMAIN replaces the search list.
    static DnsConfig load_system();
};

// kNotFound: the name does not exist (NXDOMAIN). kNoData: it exists but has no record of the
// type asked for (worth asking again for the other address family).
// kMismatch: the message does not repeat the question that was asked (see dns_parse_response).
enum class DnsStatus { kOk, kNotFound, kNoData, kTimeout, kError, kMismatch };

constexpr uint16_t kDnsTypeA = 1;
constexpr uint16_t kDnsTypeAAAA = 28;

// Looks `host` up in the hosts file; IPv4 entries win over IPv6 ones.
bool hosts_file_lookup(const std::string& path, const std::string& host, IpAddress* out);

// The candidate fully-qualified names for `host`, in the order a stub resolver tries them.
std::vector<std::string> dns_candidates(const std::string& host, const DnsConfig& cfg);

// Wire format helpers (exposed for unit tests).
bool dns_build_query(const std::string& name, uint16_t id, std::vector<uint8_t>* out, uint16_t qtype = kDnsTypeA);
// Parses a response to query `id`. kOk sets *out to the first record of type `qtype` for the
// queried name, following CNAMEs inside the message. *truncated is set when the TC bit is on.
// With `asked_name` the question section must be exactly that question (one entry, same name
// ignoring case, type `qtype`, class IN); otherwise the result is kMismatch and the caller should
// treat the message as a stray datagram.
DnsStatus dns_parse_response(const uint8_t* data, size_t len, uint16_t id, IpAddress* out, bool* truncated,
                             uint16_t qtype = kDnsTypeA, const std::string* asked_name = nullptr);

// Full blocking lookup: literal, "localhost", hosts file, then DNS (A first; AAAA when the name
// has no A record). `deadline_ms` bounds the whole operation (0: only the per-try timeouts apply).
DnsStatus resolve_address(const std::string& host, const DnsConfig& cfg, IpAddress* out, int deadline_ms = 0);

}  // namespace hola
