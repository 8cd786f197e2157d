// Stub resolver: configuration parsing, search-list expansion, wire format.
#include <arpa/inet.h>
#include <unistd.h>

#include <cstdio>
#include <string>
#include <vector>

#include "net/dns.h"
#include "testing.h"

using namespace hola;

namespace {

// Minimal response builder: header + the question + the given raw answer records.
std::vector<uint8_t> response(uint16_t id, const std::string& qname, int rcode, const std::vector<std::vector<uint8_t>>& answers,
                              bool truncated = false) {
    std::vector<uint8_t> q;
    dns_build_query(qname, id, &q);
    std::vector<uint8_t> out(q.begin(), q.end());
    out[2] = static_cast<uint8_t>(0x81 | (truncated ? 0x02 : 0));
    out[3] = static_cast<uint8_t>(0x80 | rcode);
    out[6] = 0;
    out[7] = static_cast<uint8_t>(answers.size());
    for (const auto& a : answers) out.insert(out.end(), a.begin(), a.end());
    return out;
}

std::vector<uint8_t> encode_name(const std::string& name) {
    std::vector<uint8_t> q;
    dns_build_query(name, 0, &q);
    return std::vector<uint8_t>(q.begin() + 12, q.end() - 4);
}

std::vector<uint8_t> record(const std::vector<uint8_t>& owner, uint16_t type, const std::vector<uint8_t>& rdata) {
    std::vector<uint8_t> r = owner;
    const uint8_t fixed[] = {static_cast<uint8_t>(type >> 8), static_cast<uint8_t>(type), 0, 1, 0, 0, 0, 30,
                             static_cast<uint8_t>(rdata.size() >> 8), static_cast<uint8_t>(rdata.size())};
    r.insert(r.end(), fixed, fixed + sizeof fixed);
    r.insert(r.end(), rdata.begin(), rdata.end());
    return r;
}

const std::vector<uint8_t> kPointerToQuestion = {0xC0, 0x0C};

}  // namespace

TEST(dns_resolv_conf_parsing) {
    const DnsConfig c = DnsConfig::parse(
        "# kubernetes pod\nnameserver 10.96.0.10\nsearch msa.svc.cluster.local svc.cluster.local cluster.local\n"
        "options ndots:5 timeout:2 attempts:3\nnameserver 2001:db8::1 ; second\nnameserver 1.1.1.1\nnameserver 8.8.8.8\n");
    CHECK_EQ(c.nameservers.size(), size_t{3});  // libc keeps at most three
    CHECK_EQ(c.nameservers[0], std::string("10.96.0.10"));
    CHECK_EQ(c.nameservers[1], std::string("2001:db8::1"));
    CHECK_EQ(c.search.size(), size_t{3});
    CHECK_EQ(c.ndots, 5);
    CHECK_EQ(c.timeout_ms, 2000);
    CHECK_EQ(c.attempts, 3);
    const DnsConfig d = DnsConfig::parse("");
    CHECK_EQ(d.nameservers.size(), size_t{1});
    CHECK_EQ(d.nameservers[0], std::string("127.0.0.1"));
    CHECK_EQ(d.ndots, 1);
    const DnsConfig e = DnsConfig::parse("domain example.org\nsearch a.test b.test\n");
    CHECK_EQ(e.search.size(), size_t{2});  // the later line replaces the earlier one
}

TEST(dns_candidate_order) {
    DnsConfig c = DnsConfig::parse("search msa.svc.cluster.local svc.cluster.local\noptions ndots:5\n");
    auto v = dns_candidates("aloha", c);
    CHECK_EQ(v.size(), size_t{3});
    CHECK_EQ(v[0], std::string("aloha.msa.svc.cluster.local"));
    CHECK_EQ(v[1], std::string("aloha.svc.cluster.local"));
    CHECK_EQ(v[2], std::string("aloha"));
    v = dns_candidates("zipkin-query.tools.svc.cluster.local.", c);
    CHECK_EQ(v.size(), size_t{1});
    CHECK_EQ(v[0], std::string("zipkin-query.tools.svc.cluster.local"));
    c.ndots = 1;
    v = dns_candidates("api.example.org", c);
    CHECK_EQ(v[0], std::string("api.example.org"));  // enough dots: tried as given first
    CHECK_EQ(v.size(), size_t{3});
    CHECK_EQ(dns_candidates("", c).size(), size_t{0});
}

TEST(dns_query_wire_format) {
    std::vector<uint8_t> q;
    CHECK(dns_build_query("aloha.msa.svc", 0xBEEF, &q));
    const uint8_t expect[] = {0xBE, 0xEF, 0x01, 0x00, 0, 1, 0, 0, 0, 0, 0, 0, 5, 'a', 'l', 'o', 'h', 'a', 3, 'm', 's', 'a',
                              3,    's',  'v',  'c',  0, 0, 1, 0, 1};
    CHECK_EQ(q.size(), sizeof expect);
    CHECK(std::equal(q.begin(), q.end(), expect));
    CHECK(!dns_build_query("bad..name", 1, &q));
    CHECK(!dns_build_query(std::string(64, 'x') + ".com", 1, &q));
    CHECK(!dns_build_query(".", 1, &q));
    CHECK(dns_build_query("a.b", 1, &q, kDnsTypeAAAA));
    CHECK_EQ(static_cast<int>(q[q.size() - 3]), 28);  // QTYPE AAAA
}

TEST(dns_response_parsing) {
    IpAddress ip;
    bool tc = false;
    // Plain A answer with a compressed owner name.
    auto r = response(7, "aloha.test", 0, {record(kPointerToQuestion, 1, {10, 1, 2, 3})});
    CHECK(dns_parse_response(r.data(), r.size(), 7, &ip, &tc) == DnsStatus::kOk);
    CHECK_EQ(ntohl(ip.v4_network_order()), uint32_t{0x0A010203});
    CHECK(!tc);
    // CNAME chain: aloha.test -> real.test -> 192.0.2.9, owner names spelled out and upper-cased.
    r = response(8, "aloha.test", 0,
                 {record(kPointerToQuestion, 5, encode_name("REAL.test")), record(encode_name("real.TEST"), 1, {192, 0, 2, 9})});
    CHECK(dns_parse_response(r.data(), r.size(), 8, &ip, nullptr) == DnsStatus::kOk);
    CHECK_EQ(ntohl(ip.v4_network_order()), uint32_t{0xC0000209});
    // A record listed before the alias that leads to it.
    r = response(8, "aloha.test", 0,
                 {record(encode_name("real.test"), 1, {192, 0, 2, 10}), record(kPointerToQuestion, 5, encode_name("real.test"))});
    CHECK(dns_parse_response(r.data(), r.size(), 8, &ip, nullptr) == DnsStatus::kOk);
    CHECK_EQ(ntohl(ip.v4_network_order()), uint32_t{0xC000020A});
    // An A record for some other name does not count.
    r = response(9, "aloha.test", 0, {record(encode_name("other.test"), 1, {1, 1, 1, 1})});
    CHECK(dns_parse_response(r.data(), r.size(), 9, &ip, nullptr) == DnsStatus::kNoData);
    // NXDOMAIN, SERVFAIL, wrong id, truncation, AAAA only, garbage.
    r = response(10, "aloha.test", 3, {});
    CHECK(dns_parse_response(r.data(), r.size(), 10, &ip, nullptr) == DnsStatus::kNotFound);
    r = response(10, "aloha.test", 2, {});
    CHECK(dns_parse_response(r.data(), r.size(), 10, &ip, nullptr) == DnsStatus::kError);
    CHECK(dns_parse_response(r.data(), r.size(), 11, &ip, nullptr) == DnsStatus::kError);
    r = response(12, "aloha.test", 0, {}, true);
    CHECK(dns_parse_response(r.data(), r.size(), 12, &ip, &tc) == DnsStatus::kNoData);
    CHECK(tc);
    std::vector<uint8_t> v6(16, 0);
    v6[0] = 0xfd;
    v6[15] = 0x01;
    r = response(13, "aloha.test", 0, {record(kPointerToQuestion, 28, v6)});
    CHECK(dns_parse_response(r.data(), r.size(), 13, &ip, nullptr) == DnsStatus::kNoData);  // asked for A, got AAAA only
    CHECK(dns_parse_response(r.data(), r.size(), 13, &ip, nullptr, kDnsTypeAAAA) == DnsStatus::kOk);
    CHECK_EQ(ip.to_string(), std::string("fd00::1"));
    const uint8_t junk[] = {0, 13, 0x81, 0x80, 0, 1, 0, 5, 0, 0, 0, 0, 0xC0};
    CHECK(dns_parse_response(junk, sizeof junk, 13, &ip, nullptr) == DnsStatus::kError);
    // Every truncation of a valid message is handled without reading out of bounds.
    r = response(8, "aloha.test", 0,
                 {record(kPointerToQuestion, 5, encode_name("real.test")), record(encode_name("real.test"), 1, {192, 0, 2, 9})});
    for (size_t n = 0; n < r.size(); ++n) dns_parse_response(r.data(), n, 8, &ip, nullptr);
    // The question must be the one that was asked: same name (any case), same type, exactly one.
    r = response(15, "aloha.test", 0, {record(kPointerToQuestion, 1, {10, 1, 2, 3})});
    const std::string asked = "ALOHA.test.", other = "evil.test";
    CHECK(dns_parse_response(r.data(), r.size(), 15, &ip, nullptr, kDnsTypeA, &asked) == DnsStatus::kOk);
    CHECK(dns_parse_response(r.data(), r.size(), 15, &ip, nullptr, kDnsTypeA, &other) == DnsStatus::kMismatch);
    CHECK(dns_parse_response(r.data(), r.size(), 15, &ip, nullptr, kDnsTypeAAAA, &asked) == DnsStatus::kMismatch);
    r = response(15, "evil.test", 3, {});  // somebody else's NXDOMAIN is not ours
    CHECK(dns_parse_response(r.data(), r.size(), 15, &ip, nullptr, kDnsTypeA, &asked) == DnsStatus::kMismatch);
    r = response(15, "aloha.test", 0, {});
    r[5] = 0;  // QDCOUNT 0: nothing says what this answers
    CHECK(dns_parse_response(r.data(), r.size(), 15, &ip, nullptr, kDnsTypeA, &asked) == DnsStatus::kMismatch);
    // A compression loop terminates.
    std::vector<uint8_t> loop = response(14, "aloha.test", 0, {});
    loop[7] = 1;
    const uint8_t self_ptr[] = {0xC0, static_cast<uint8_t>(loop.size()), 0, 1, 0, 1, 0, 0, 0, 0, 0, 4, 1, 2, 3, 4};
    loop.insert(loop.end(), self_ptr, self_ptr + sizeof self_ptr);
    CHECK(dns_parse_response(loop.data(), loop.size(), 14, &ip, nullptr) == DnsStatus::kError);
}

TEST(dns_hosts_file_and_literals) {
    char tmpl[] = "/tmp/hola_hosts_XXXXXX";
    const int fd = mkstemp(tmpl);
    const std::string text = "127.0.0.1 localhost\n::1 ip6-localhost\n# comment\n10.9.8.7   Aloha aloha.internal # trailing\n";
    CHECK(write(fd, text.data(), text.size()) == static_cast<ssize_t>(text.size()));
    close(fd);
    IpAddress ip;
    CHECK(hosts_file_lookup(tmpl, "aloha", &ip));
    CHECK_EQ(ntohl(ip.v4_network_order()), uint32_t{0x0A090807});
    CHECK(hosts_file_lookup(tmpl, "ALOHA.internal", &ip));
    CHECK(!hosts_file_lookup(tmpl, "trailing", &ip));
    CHECK(hosts_file_lookup(tmpl, "ip6-localhost", &ip));  // IPv6-only entries are usable too
    CHECK_EQ(ip.to_string(), std::string("::1"));
    CHECK(!hosts_file_lookup("/nonexistent/hosts", "aloha", &ip));
    DnsConfig cfg = DnsConfig::parse("nameserver 127.0.0.1\noptions timeout:1 attempts:1\n");
    cfg.hosts_file = tmpl;
    CHECK(resolve_address("192.0.2.55", cfg, &ip) == DnsStatus::kOk);
    CHECK_EQ(ntohl(ip.v4_network_order()), uint32_t{0xC0000237});
    CHECK(resolve_address("fd00::42", cfg, &ip) == DnsStatus::kOk);
    CHECK(ip.family == AF_INET6);
    CHECK(resolve_address("LocalHost", cfg, &ip) == DnsStatus::kOk);
    CHECK_EQ(ntohl(ip.v4_network_order()), uint32_t{0x7F000001});
    CHECK(resolve_address("aloha", cfg, &ip) == DnsStatus::kOk);
    CHECK_EQ(ntohl(ip.v4_network_order()), uint32_t{0x0A090807});
    unlink(tmpl);
}
