// Unit tests for the parsers and codecs in src/core. Golden vectors are written by hand from the
// java.util.Properties / java.util.Formatter / Jackson behaviour the reference depends on
// (there is no JVM on the build box to generate them).
#include <cmath>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <string>
#include <unistd.h>

#include "core/config.h"
#include "core/formatter.h"
#include "core/histogram.h"
#include "core/http.h"
#include "core/json.h"
#include "core/properties.h"
#include "core/text.h"
#include "core/trace.h"
#include "testing.h"

using namespace hola;

namespace {

PropertyList parse_ok(const std::string& text) {
    PropertyList list;
    std::string err;
    const bool ok = parse_properties(text, &list, &err);
    CHECK(ok);
    return list;
}

std::string fmt(const std::string& pattern, const std::string& arg = "host") {
    FormatResult r = java_format(pattern, arg);
    return r.ok ? r.text : "!" + r.error;
}

bool fmt_fails(const std::string& pattern, const std::string& exception) {
    FormatResult r = java_format(pattern, "host");
    return !r.ok && r.error.find(exception) != std::string::npos;
}

}  // namespace

// ---- properties --------------------------------------------------------------------------------

TEST(properties_reference_file_line) {
    // The last line of /root/reference/translation.properties: value ends in backslash + 'n'.
    auto p = parse_ok("# comment\nhello=Hola de %s mediante el archivo de configuraciones\\n");
    CHECK_EQ(p.size(), size_t{1});
    CHECK_EQ(p[0].first, std::string("hello"));
    CHECK_EQ(p[0].second, std::string("Hola de %s mediante el archivo de configuraciones\n"));
}

TEST(properties_separators) {
    auto p = parse_ok("a=1\nb:2\nc 3\nd = 4\ne\t:\t5\nf\n g   =   spaced value  \nh==x\ni=:x\nj : = x\n");
    CHECK_EQ(p.size(), size_t{10});
    CHECK_EQ(p[0].second, std::string("1"));
    CHECK_EQ(p[1].second, std::string("2"));
    CHECK_EQ(p[2].second, std::string("3"));
    CHECK_EQ(p[3].second, std::string("4"));
    CHECK_EQ(p[4].second, std::string("5"));
    CHECK_EQ(p[5].first, std::string("f"));
    CHECK_EQ(p[5].second, std::string(""));
    CHECK_EQ(p[6].first, std::string("g"));
    CHECK_EQ(p[6].second, std::string("spaced value  "));  // trailing blanks are kept
    CHECK_EQ(p[7].second, std::string("=x"));
    CHECK_EQ(p[8].second, std::string(":x"));
    CHECK_EQ(p[9].second, std::string("= x"));  // only one separator is swallowed
}

TEST(properties_comments_and_blank_lines) {
    auto p = parse_ok("\n   \n# c1\n! c2\n   # indented comment\nk=v # not a comment\n#trailing\\\nx=1\n");
    CHECK_EQ(p.size(), size_t{2});
    CHECK_EQ(p[0].second, std::string("v # not a comment"));
    CHECK_EQ(p[1].first, std::string("x"));  // a comment line never continues
}

TEST(properties_line_continuation) {
    auto p = parse_ok("fruits = apple, \\\n          banana, \\\r\n  pear\nodd=a\\\\\nnext=1\nlast=end\\");
    CHECK_EQ(p.size(), size_t{4});
    CHECK_EQ(p[0].second, std::string("apple, banana, pear"));
    CHECK_EQ(p[1].second, std::string("a\\"));  // two backslashes = one literal, no continuation
    CHECK_EQ(p[2].second, std::string("1"));
    CHECK_EQ(p[3].second, std::string("end"));  // dangling backslash at EOF is dropped
}

TEST(properties_escapes) {
    auto p = parse_ok("k\\ e\\=y\\:z = t\\tn\\nr\\rf\\fq\\q\\\\ \\u0041\\u00e9\\u20AC\n");
    CHECK_EQ(p.size(), size_t{1});
    CHECK_EQ(p[0].first, std::string("k e=y:z"));
    CHECK_EQ(p[0].second, std::string("t\tn\nr\rf\fqq\\ A\xC3\xA9\xE2\x82\xAC"));
}

TEST(properties_surrogate_pairs) {
    auto p = parse_ok("smile=\\uD83D\\uDE00!\nlone=\\uD83Dx\n");
    CHECK_EQ(p[0].second, std::string("\xF0\x9F\x98\x80!"));
    CHECK_EQ(p[1].second, std::string("?x"));
}

TEST(properties_malformed_unicode_fails_whole_load) {
    PropertyList list;
    std::string err;
    CHECK(!parse_properties("ok=1\nbad=\\u12G4\n", &list, &err));
    CHECK(err.find("Malformed") != std::string::npos);
    CHECK(!parse_properties("bad=\\u12", &list, &err));
    Properties props;
    CHECK(!props.load("a=1\nb=\\uZZZZ"));
    CHECK_EQ(props.size(), size_t{0});  // nothing survives, like the swallowed exception
}

TEST(properties_duplicates_last_wins_and_crlf) {
    Properties props;
    CHECK(props.load("a=1\r\na=2\r\n\r\nb=3\r"));
    CHECK_EQ(*props.get("a"), std::string("2"));
    CHECK_EQ(*props.get("b"), std::string("3"));
    CHECK(props.get("c") == nullptr);
}

TEST(properties_missing_file) {
    Properties props;
    std::string err;
    CHECK(!props.load_file("/nonexistent/dir/none.properties", &err));
    CHECK(!err.empty());
}

// ---- formatter ---------------------------------------------------------------------------------

TEST(format_reference_patterns) {
    CHECK_EQ(fmt("Hola de %s", "unknown"), std::string("Hola de unknown"));
    CHECK_EQ(fmt("Hola de %s mediante el archivo de configuraciones\n", "pod-1"),
             std::string("Hola de pod-1 mediante el archivo de configuraciones\n"));
    CHECK_EQ(fmt("no directives"), std::string("no directives"));
    CHECK_EQ(fmt(""), std::string(""));
}

TEST(format_percent_and_newline) {
    CHECK_EQ(fmt("100%% %s%n"), std::string("100% host\n"));
    CHECK_EQ(fmt("%5%|"), std::string("%|"));  // JDK 8 ignores the width of %%
    CHECK(fmt_fails("%-%", "MissingFormatWidthException"));
    CHECK(fmt_fails("%5n", "IllegalFormatWidthException"));
    CHECK(fmt_fails("%.2%", "IllegalFormatPrecisionException"));
    CHECK(fmt_fails("%+%", "IllegalFormatFlagsException"));
}

TEST(format_width_precision_case) {
    CHECK_EQ(fmt("[%10s]", "abc"), std::string("[       abc]"));
    CHECK_EQ(fmt("[%-10s]", "abc"), std::string("[abc       ]"));
    CHECK_EQ(fmt("[%.2s]", "abc"), std::string("[ab]"));
    CHECK_EQ(fmt("[%5.1s]", "abc"), std::string("[    a]"));
    CHECK_EQ(fmt("%S", "straße"), std::string("STRASSE"));
    CHECK_EQ(fmt("[%3s]", "\xC3\xA9\xC3\xA9"), std::string("[ \xC3\xA9\xC3\xA9]"));        // width counts chars, not bytes
    CHECK_EQ(fmt("[%3s]", "\xF0\x9F\x98\x80"), std::string("[ \xF0\x9F\x98\x80]"));        // astral = 2 UTF-16 units
    CHECK_EQ(fmt("[%2s]", "abc"), std::string("[abc]"));
}

TEST(format_boolean_and_hash) {
    CHECK_EQ(fmt("%b"), std::string("true"));
    CHECK_EQ(fmt("%B"), std::string("TRUE"));
    CHECK_EQ(fmt("%.2b"), std::string("tr"));
    // "host".hashCode() == 3208616 == 0x30f5a8
    CHECK_EQ(fmt("%h"), std::string("30f5a8"));
    CHECK_EQ(fmt("%H"), std::string("30F5A8"));
    CHECK_EQ(java_string_hash(""), 0);
    CHECK_EQ(java_string_hash("hello"), 99162322);
}

TEST(format_argument_indexing) {
    CHECK_EQ(fmt("%1$s-%1$s"), std::string("host-host"));
    CHECK_EQ(fmt("%s-%<s"), std::string("host-host"));
    CHECK_EQ(fmt("%1$s %s"), std::string("host host"));  // explicit index does not advance the ordinary one
    CHECK(fmt_fails("%s %s", "MissingFormatArgumentException"));
    CHECK(fmt_fails("%2$s", "MissingFormatArgumentException"));
    CHECK(fmt_fails("%<s", "MissingFormatArgumentException"));
}

TEST(format_rejections) {
    CHECK(fmt_fails("%d", "IllegalFormatConversionException"));
    CHECK(fmt_fails("%x", "IllegalFormatConversionException"));
    CHECK(fmt_fails("%f", "IllegalFormatConversionException"));
    CHECK(fmt_fails("%c", "IllegalFormatConversionException"));
    CHECK(fmt_fails("%tY", "IllegalFormatConversionException"));
    CHECK(fmt_fails("%tZZ", "IllegalFormatConversionException"));
    CHECK(fmt_fails("%tq", "UnknownFormatConversionException"));
    CHECK(fmt_fails("%q", "UnknownFormatConversionException"));
    CHECK(fmt_fails("%D", "UnknownFormatConversionException"));
    CHECK(fmt_fails("%N", "UnknownFormatConversionException"));
    CHECK(fmt_fails("100%", "UnknownFormatConversionException"));
    CHECK(fmt_fails("50% off", "IllegalFormatConversionException"));  // "% o" parses as flag ' ' + octal
    CHECK(fmt_fails("50% yes", "UnknownFormatConversionException"));
    CHECK(fmt_fails("%.s", "UnknownFormatConversionException"));
    CHECK(fmt_fails("%-s", "MissingFormatWidthException"));
    CHECK(fmt_fails("%#s", "FormatFlagsConversionMismatchException"));
    CHECK(fmt_fails("%+s", "FormatFlagsConversionMismatchException"));
    CHECK(fmt_fails("%010s", "FormatFlagsConversionMismatchException"));
    CHECK(fmt_fails("%--5s", "DuplicateFormatFlagsException"));
    CHECK(fmt_fails("%s %q", "UnknownFormatConversionException"));  // parse errors win over everything
}

// ---- text --------------------------------------------------------------------------------------

TEST(text_utf8_edges) {
    std::string out;
    append_utf8(&out, 0x41);
    append_utf8(&out, 0xE9);
    append_utf8(&out, 0x20AC);
    append_utf8(&out, 0x1F600);
    CHECK_EQ(out, std::string("A\xC3\xA9\xE2\x82\xAC\xF0\x9F\x98\x80"));
    out.clear();
    append_utf8(&out, 0xD800);    // a lone surrogate has no UTF-8 form: the JVM writes '?'
    append_utf8(&out, 0x110000);  // beyond Unicode
    CHECK_EQ(out, std::string("??"));
    // Malformed input decodes as U+FFFD, one byte at a time, and always advances.
    const std::string bad[] = {"\x80", "\xC3", "\xC3\x41", "\xE2\x82", "\xE2\x41\x41", "\xF0\x9F\x98", "\xF8\x88\x80\x80\x80",
                               "\xC0\x80" /* overlong NUL */, "\xE0\x80\x80" /* overlong */, "\xED\xA0\x80" /* surrogate */, "\xF4\x90\x80\x80" /* > U+10FFFF */};
    for (const std::string& b : bad) {
        size_t pos = 0;
        CHECK_EQ(decode_utf8(b, &pos), uint32_t{0xFFFD});
        CHECK_EQ(pos, size_t{1});
    }
    size_t pos = 0;
    CHECK_EQ(decode_utf8("\xE2\x82\xAC", &pos), uint32_t{0x20AC});
    CHECK_EQ(pos, size_t{3});
    CHECK_EQ(utf16_length("a\xC3\xA9\xF0\x9F\x98\x80"), size_t{4});
    CHECK_EQ(std::string(utf16_prefix("a\xF0\x9F\x98\x80z", 2)), std::string("a"));  // never half of a surrogate pair
    CHECK_EQ(std::string(utf16_prefix("a\xF0\x9F\x98\x80z", 3)), std::string("a\xF0\x9F\x98\x80"));
    // String.hashCode walks UTF-16 units: "\uD83D\uDE00".hashCode() == 0xD83D * 31 + 0xDE00.
    CHECK_EQ(java_string_hash("\xF0\x9F\x98\x80"), static_cast<int32_t>(0xD83D * 31 + 0xDE00));
}

TEST(text_upper_case_like_the_jvm) {
    CHECK_EQ(to_upper_java("hola de mar\xC3\xAD" "a"), std::string("HOLA DE MAR\xC3\x8D" "A"));
    CHECK_EQ(to_upper_java("stra\xC3\x9F" "e"), std::string("STRASSE"));       // one-to-two
    CHECK_EQ(to_upper_java("\xC3\xBF"), std::string("\xC5\xB8"));            // y with diaeresis leaves Latin-1: U+0178
    CHECK_EQ(to_upper_java("\xC2\xB5"), std::string("\xCE\x9C"));            // micro sign -> Greek capital mu
    CHECK_EQ(to_upper_java("\xC3\xB7 \xC3\x97"), std::string("\xC3\xB7 \xC3\x97"));  // division and multiplication signs have no case
    CHECK_EQ(to_upper_java("\xD0\xB6"), std::string("\xD0\xB6"));            // other scripts are left alone (documented)
}

TEST(text_utf16_assembler_and_cpu_lists) {
    std::string out;
    Utf16Assembler a(&out);
    a.push_unit(0xD83D);
    a.push_unit(0xDE00);  // a pair: one astral character
    a.push_unit(0xDE00);  // a low half on its own
    a.push_unit(0xD83D);  // a high half followed by something else
    a.push_unit(0x41);
    a.push_unit(0xD83D);  // ... and one left over at the end
    a.flush();
    CHECK_EQ(out, std::string("\xF0\x9F\x98\x80?" "?A?"));
    std::vector<int> cpus;
    CHECK(parse_cpu_list("0,2,4-7", &cpus));
    CHECK_EQ(cpus.size(), size_t{6});
    CHECK_EQ(cpus[5], 7);
    cpus.clear();
    CHECK(parse_cpu_list("3,,4,", &cpus));  // empty items are skipped
    CHECK_EQ(cpus.size(), size_t{2});
    for (const char* bad : {"", ",", "a", "1-", "-1", "5-2", "1-2-3", "1234567", "0x3"}) {
        cpus.clear();
        CHECK(!parse_cpu_list(bad, &cpus));
    }
}

// ---- json --------------------------------------------------------------------------------------

TEST(json_escapes_and_malformed_documents) {
    std::string out;
    json_append_string(&out, std::string("q\" b\\ /\b\f\n\r\t\x01\x1f z"));
    CHECK_EQ(out, std::string("\"q\\\" b\\\\ /\\b\\f\\n\\r\\t\\u0001\\u001F z\""));
    JsonStringList items;
    std::string err;
    CHECK(json_decode_string_list("[\"\\\"\\\\\\/\\b\\f\\n\\r\\t\\u0041\\uD83D\\uDE00\"]", &items, &err));
    CHECK_EQ(*items[0], std::string("\"\\/\b\f\n\r\tA\xF0\x9F\x98\x80"));
    CHECK(json_decode_string_list(" [ 1 , -2.5e+3 , true , false , null , \"x\" ] ", &items, &err));
    CHECK_EQ(items.size(), size_t{6});
    CHECK_EQ(*items[1], std::string("-2.5e+3"));
    CHECK(!items[4].has_value());
    const char* bad[] = {"", "[", "[\"a\"", "[\"a\",]", "[,]", "{\"a\":1}", "[[1]]", "[{\"a\":1}]", "[\"a\" \"b\"]",
                         "[\"\x01\"]", "[\"\\", "[\"\\u12\"]", "[\"\\u12G4\"]", "[\"\\q\"]", "[-]", "[1.]", "[1e]", "[01x]", "[tru]", "[nul]", "\"a\""};
    CHECK(json_decode_string_list("[\"a\"] trailing", &items, &err));  // Jackson's readValue stops after the first value
    for (const char* doc : bad) {
        err.clear();
        if (json_decode_string_list(doc, &items, &err)) std::fprintf(stderr, "  accepted: %s\n", doc);
        CHECK(!json_decode_string_list(doc, &items, &err));
        CHECK(!err.empty());
    }
}

TEST(json_encode) {
    CHECK_EQ(json_encode_string_list({}), std::string("[]"));
    CHECK_EQ(json_encode_string_list({std::string("Hola de unknown"), std::string("Aloha response (fallback)")}),
             std::string("[\"Hola de unknown\",\"Aloha response (fallback)\"]"));
    CHECK_EQ(json_encode_string_list({std::string("a\"b\\c\n\t\x01"), std::nullopt, std::string("\xC3\xA9")}),
             std::string("[\"a\\\"b\\\\c\\n\\t\\u0001\",null,\"\xC3\xA9\"]"));
}

TEST(json_decode) {
    JsonStringList out;
    std::string err;
    CHECK(json_decode_string_list(" [ \"Aloha mai x\" , \"Ol\\u00e1 de y\", \"Namaste \\uD83D\\uDE00\\n\" ] ", &out, &err));
    CHECK_EQ(out.size(), size_t{3});
    CHECK_EQ(*out[1], std::string("Ol\xC3\xA1 de y"));
    CHECK_EQ(*out[2], std::string("Namaste \xF0\x9F\x98\x80\n"));
    CHECK(json_decode_string_list("[]", &out, &err));
    CHECK_EQ(out.size(), size_t{0});
    CHECK(json_decode_string_list("[1, -2.5e3, true, false, null, \"s\"]", &out, &err));
    CHECK_EQ(out.size(), size_t{6});
    CHECK_EQ(*out[0], std::string("1"));
    CHECK_EQ(*out[1], std::string("-2.5e3"));
    CHECK_EQ(*out[2], std::string("true"));
    CHECK(!out[4].has_value());
}

TEST(json_decode_rejects) {
    JsonStringList out;
    std::string err;
    CHECK(!json_decode_string_list("", &out, &err));
    CHECK(!json_decode_string_list("{\"a\":1}", &out, &err));
    CHECK(!json_decode_string_list("\"just a string\"", &out, &err));
    CHECK(!json_decode_string_list("[\"a\",]", &out, &err));
    CHECK(!json_decode_string_list("[\"a\" \"b\"]", &out, &err));
    CHECK(!json_decode_string_list("[\"unterminated]", &out, &err));
    CHECK(!json_decode_string_list("[[\"nested\"]]", &out, &err));
    CHECK(!json_decode_string_list("[{\"k\":\"v\"}]", &out, &err));
    CHECK(!json_decode_string_list("[\"bad \\x escape\"]", &out, &err));
    CHECK(!json_decode_string_list("[\"a\"", &out, &err));
    CHECK(!json_decode_string_list("<html>502</html>", &out, &err));
}

// ---- config ------------------------------------------------------------------------------------

namespace {
struct TempFile {
    std::string path;
    explicit TempFile(const std::string& content) {
        char tmpl[] = "/tmp/hola_test_XXXXXX";
        const int fd = mkstemp(tmpl);
        path = tmpl;
        write_all(fd, content);
        close(fd);
    }
    ~TempFile() { unlink(path.c_str()); }
    void rewrite(const std::string& content) {
        FILE* f = std::fopen(path.c_str(), "wb");
        std::fwrite(content.data(), 1, content.size(), f);
        std::fclose(f);
    }
    static void write_all(int fd, const std::string& s) {
        size_t off = 0;
        while (off < s.size()) {
            const ssize_t n = write(fd, s.data() + off, s.size() - off);
            if (n <= 0) break;
            off += static_cast<size_t>(n);
        }
    }
};
}  // namespace

TEST(config_layering_and_default) {
    unsetenv("hello");
    unsetenv("hello.Production");
    TempFile file("hello=from file\n");
    ConfigResolver r;
    auto* sys = new SystemPropertySource();
    auto* conf = new ConfFileSource(sys);
    r.add_source(std::unique_ptr<ConfigSource>(conf));  // added first on purpose: ordinal decides
    r.add_source(std::unique_ptr<ConfigSource>(new EnvironmentSource()));
    r.add_source(std::unique_ptr<ConfigSource>(sys));
    std::string v;
    CHECK(!r.resolve("hello", &v));  // no -Dconf: the source yields nothing
    sys->set("conf", file.path);
    CHECK(r.resolve("hello", &v));
    CHECK_EQ(v, std::string("from file"));
    setenv("hello", "from env", 1);
    CHECK(r.resolve("hello", &v));
    CHECK_EQ(v, std::string("from env"));
    sys->set("hello", "from -D");
    CHECK(r.resolve("hello", &v));
    CHECK_EQ(v, std::string("from -D"));
    unsetenv("hello");
}

TEST(config_project_stage_and_variables) {
    unsetenv("hello");
    ConfigResolver r;
    auto* sys = new SystemPropertySource();
    r.add_source(std::unique_ptr<ConfigSource>(sys));
    sys->set("hello", "plain");
    sys->set("hello.Production", "staged ${who}!");
    sys->set("who", "mundo");
    std::string v;
    CHECK(r.resolve("hello", &v));
    CHECK_EQ(v, std::string("staged mundo!"));
    sys->set("org.apache.deltaspike.ProjectStage", "Development");
    CHECK(r.resolve("hello", &v));
    CHECK_EQ(v, std::string("plain"));
    sys->set("hello", "${missing} stays, ${who} goes, ${who} twice");
    CHECK(r.resolve("hello", &v));
    CHECK_EQ(v, std::string("${missing} stays, mundo goes, mundo twice"));
    sys->set("loop", "${loop}");
    CHECK(r.resolve("loop", &v));  // terminates
}

TEST(config_cache_ttl_and_hot_reload) {
    unsetenv("hello");
    TempFile file("hello=uno %s\n");
    ConfigResolver r;
    auto* sys = new SystemPropertySource();
    auto* conf = new ConfFileSource(sys);
    r.add_source(std::unique_ptr<ConfigSource>(sys));
    r.add_source(std::unique_ptr<ConfigSource>(conf));
    sys->set("conf", file.path);
    std::vector<std::string> log;
    CachedConfigValue hello(&r, "hello", "Hola de %s", 5000000000ll, true, [&](const std::string& m) { log.push_back(m); });
    const int64_t s = 1000000000ll;
    CHECK_EQ(hello.get(0), std::string("uno %s"));
    const uint64_t reads_after_first = conf->reads();
    file.rewrite("hello=dos %s\n");
    CHECK_EQ(hello.get(4 * s), std::string("uno %s"));  // still cached
    CHECK_EQ(conf->reads(), reads_after_first);
    CHECK_EQ(hello.get(5 * s), std::string("uno %s"));  // "now <= reloadAfter" keeps the boundary cached
    CHECK_EQ(hello.get(5 * s + 1), std::string("dos %s"));
    file.rewrite("hello=\n");
    CHECK_EQ(hello.get(11 * s), std::string("Hola de %s"));  // empty value -> default
    file.rewrite("hello=\\uXYZW\n");
    CHECK_EQ(hello.get(17 * s), std::string("Hola de %s"));  // unparsable file -> default
    unlink(file.path.c_str());
    CHECK_EQ(hello.get(23 * s), std::string("Hola de %s"));  // missing file -> default
    CHECK_EQ(log.size(), size_t{3});
    CHECK_EQ(log[1], std::string("New value dos %s for key hello."));
}

// ---- http --------------------------------------------------------------------------------------

TEST(http_request_basic) {
    const std::string raw =
        "GET /api/hola?x=1&y=2 HTTP/1.1\r\nHost: localhost:8080\r\nAccept: text/plain, */*;q=0.1\r\n"
        "X-B3-TraceId: 463ac35c9f6413ad\r\nX-B3-SpanId: a2fb4a1d1a96d312\r\nx-b3-parentspanid: 0020000000000001\r\n"
        "X-B3-Sampled: 1\r\n\r\nNEXT";
    RequestHead h;
    CHECK(parse_request_head(raw.data(), raw.size(), &h) == ParseStatus::kComplete);
    CHECK(h.method == Method::kGet);
    CHECK_EQ(std::string(h.path), std::string("/api/hola"));
    CHECK_EQ(std::string(h.query), std::string("x=1&y=2"));
    CHECK_EQ(std::string(h.host), std::string("localhost:8080"));
    CHECK(h.keep_alive);
    CHECK(h.has_accept);
    CHECK_EQ(std::string(h.b3.trace_id), std::string("463ac35c9f6413ad"));
    CHECK_EQ(std::string(h.b3.parent_span_id), std::string("0020000000000001"));
    CHECK(h.b3.has_sampled);
    CHECK_EQ(h.head_len, raw.size() - 4);
    CHECK_EQ(h.content_length, int64_t{-1});
}

TEST(http_request_variants) {
    RequestHead h;
    std::string raw = "GET / HTTP/1.0\r\n\r\n";
    CHECK(parse_request_head(raw.data(), raw.size(), &h) == ParseStatus::kComplete);
    CHECK(!h.keep_alive);
    raw = "GET / HTTP/1.0\r\nConnection: Keep-Alive\r\n\r\n";
    CHECK(parse_request_head(raw.data(), raw.size(), &h) == ParseStatus::kComplete);
    CHECK(h.keep_alive);
    raw = "GET / HTTP/1.1\r\nConnection: foo, Close\r\n\r\n";
    CHECK(parse_request_head(raw.data(), raw.size(), &h) == ParseStatus::kComplete);
    CHECK(!h.keep_alive);
    raw = "POST /api/hola HTTP/1.1\nContent-Length: 12\nExpect: 100-continue\n\n";  // bare LF
    CHECK(parse_request_head(raw.data(), raw.size(), &h) == ParseStatus::kComplete);
    CHECK(h.method == Method::kPost);
    CHECK_EQ(h.content_length, int64_t{12});
    CHECK(h.expect_continue);
    raw = "PUT /x HTTP/1.1\r\nTransfer-Encoding: gzip, chunked\r\nContent-Length: 5\r\n\r\n";
    CHECK(parse_request_head(raw.data(), raw.size(), &h) == ParseStatus::kComplete);
    CHECK(h.chunked);
    CHECK_EQ(h.content_length, int64_t{-1});
    raw = "GET http://example.org:8080/api/health?z HTTP/1.1\r\n\r\n";
    CHECK(parse_request_head(raw.data(), raw.size(), &h) == ParseStatus::kComplete);
    CHECK_EQ(std::string(h.path), std::string("/api/health"));
    raw = "\r\n\r\nGET /late HTTP/1.1\r\n\r\n";
    CHECK(parse_request_head(raw.data(), raw.size(), &h) == ParseStatus::kComplete);
    CHECK_EQ(std::string(h.path), std::string("/late"));
    CHECK_EQ(h.head_len, raw.size());
}

TEST(http_request_incomplete_and_bad) {
    RequestHead h;
    const std::string full = "GET /api/hola HTTP/1.1\r\nHost: a\r\n\r\n";
    for (size_t n = 0; n < full.size(); ++n) CHECK(parse_request_head(full.data(), n, &h) == ParseStatus::kIncomplete);
    const char* bad[] = {
        "GET\r\n\r\n",
        "GET /x\r\n\r\n",
        "GET /x HTTP/2.0\r\n\r\n",
        "GET /x y HTTP/1.1\r\n\r\n",
        "G(T /x HTTP/1.1\r\n\r\n",
        "GET /x HTTP/1.1\r\nNoColonHere\r\n\r\n",
        "GET /x HTTP/1.1\r\nBad Name: v\r\n\r\n",
        "GET /x HTTP/1.1\r\n: empty\r\n\r\n",
        "GET /x HTTP/1.1\r\nContent-Length: 12a\r\n\r\n",
        "GET /x HTTP/1.1\r\nContent-Length: 1\r\nContent-Length: 2\r\n\r\n",
        "GET x HTTP/1.1\r\n\r\n",
    };
    for (const char* b : bad) {
        const ParseStatus st = parse_request_head(b, std::strlen(b), &h);
        if (st != ParseStatus::kBad) std::fprintf(stderr, "  not rejected: %s\n", b);
        CHECK(st == ParseStatus::kBad);
    }
}

TEST(http_request_single_pass_edges) {
    RequestHead h;
    // A malformed line is reported as soon as it is complete, even if the head is not.
    std::string raw = "GET /x HTTP/1.1\r\nBad Name: v\r\nHost: not-yet-termina";
    CHECK(parse_request_head(raw.data(), raw.size(), &h) == ParseStatus::kBad);
    // CR that is not part of a line end.
    raw = "GET /x HTTP/1.1\r\n\rX: y\r\n\r\n";
    CHECK(parse_request_head(raw.data(), raw.size(), &h) == ParseStatus::kBad);
    // Head ends right at the buffer end, with and without CR; a lone trailing CR is "need more".
    raw = "GET /x HTTP/1.1\r\nHost: a\r\n\r";
    CHECK(parse_request_head(raw.data(), raw.size(), &h) == ParseStatus::kIncomplete);
    raw += "\n";
    CHECK(parse_request_head(raw.data(), raw.size(), &h) == ParseStatus::kComplete);
    CHECK_EQ(h.head_len, raw.size());
    raw = "GET /x HTTP/1.1\nHost: a\n\nGET /y HTTP/1.1\n\n";
    CHECK(parse_request_head(raw.data(), raw.size(), &h) == ParseStatus::kComplete);
    CHECK_EQ(h.head_len, size_t{25});
    // Every token character is accepted in a header name, nothing else is.
    raw = "GET /x HTTP/1.1\r\n!#$%&'*+-.^_`|~09azAZ: ok\r\n\r\n";
    CHECK(parse_request_head(raw.data(), raw.size(), &h) == ParseStatus::kComplete);
    for (const char* bad : {"(", ")", "<", ">", "@", ",", ";", "\\", "\"", "/", "[", "]", "?", "=", "{", "}", " ", "\t", "\x7f", "\x80"}) {
        raw = std::string("GET /x HTTP/1.1\r\nX") + bad + "Y: v\r\n\r\n";
        CHECK(parse_request_head(raw.data(), raw.size(), &h) == ParseStatus::kBad);
    }
    // Connection header: usual value, lists, case.
    raw = "GET /x HTTP/1.0\r\nConnection: KEEP-ALIVE\r\n\r\n";
    CHECK(parse_request_head(raw.data(), raw.size(), &h) == ParseStatus::kComplete);
    CHECK(h.keep_alive);
    raw = "GET /x HTTP/1.1\r\nConnection: keep-alive, close\r\n\r\n";
    CHECK(parse_request_head(raw.data(), raw.size(), &h) == ParseStatus::kComplete);
    CHECK(!h.keep_alive);
}

TEST(http_response_and_chunked) {
    ResponseHead r;
    std::string raw = "HTTP/1.1 200 OK\r\nContent-Type: application/json\r\nContent-Length: 9\r\n\r\n[\"aloha\"]";
    CHECK(parse_response_head(raw.data(), raw.size(), &r) == ParseStatus::kComplete);
    CHECK_EQ(r.status, 200);
    CHECK_EQ(r.content_length, int64_t{9});
    CHECK_EQ(std::string(r.content_type), std::string("application/json"));
    raw = "HTTP/1.0 503 Service Unavailable\r\n\r\n";
    CHECK(parse_response_head(raw.data(), raw.size(), &r) == ParseStatus::kComplete);
    CHECK(!r.keep_alive);
    CHECK_EQ(r.content_length, int64_t{-1});
    raw = "HTTP/1.1 20 OK\r\n\r\n";
    CHECK(parse_response_head(raw.data(), raw.size(), &r) == ParseStatus::kBad);

    const std::string chunked = "4\r\nWiki\r\n6;ext=1\r\npedia \r\nE\r\nin \r\n\r\nchunks.\r\n0\r\nTrailer: x\r\n\r\nREST";
    // Feed byte by byte to exercise every state boundary.
    ChunkedDecoder d;
    std::string body;
    size_t used_total = 0;
    for (size_t i = 0; i < chunked.size() && !d.done(); ++i) used_total += d.feed(chunked.data() + i, 1, &body);
    CHECK(d.done());
    CHECK_EQ(body, std::string("Wikipedia in \r\n\r\nchunks."));
    CHECK_EQ(used_total, chunked.size() - 4);
    ChunkedDecoder bad;
    const std::string garbage = "zz\r\n";
    bad.feed(garbage.data(), garbage.size(), nullptr);
    CHECK(bad.failed());
}

TEST(http_accept_matching) {
    CHECK(accept_allows("", "text", "plain"));
    CHECK(accept_allows("*/*", "text", "plain"));
    CHECK(accept_allows("text/*", "text", "plain"));
    CHECK(accept_allows("application/json, text/plain;q=0.5", "text", "plain"));
    CHECK(accept_allows("TEXT/PLAIN", "text", "plain"));
    CHECK(!accept_allows("application/json", "text", "plain"));
    CHECK(!accept_allows("text/plain;q=0", "text", "plain"));
    CHECK(!accept_allows("text/plain;q=0.0, application/xml", "text", "plain"));
    CHECK(accept_allows("text/html, image/gif, image/jpeg, *; q=.2, */*; q=.2", "application", "json"));
}

TEST(http_date_format) {
    char buf[30] = {0};
    format_http_date(784111777, buf);
    CHECK_EQ(std::string(buf, 29), std::string("Sun, 06 Nov 1994 08:49:37 GMT"));
    format_http_date(0, buf);
    CHECK_EQ(std::string(buf, 29), std::string("Thu, 01 Jan 1970 00:00:00 GMT"));
    format_http_date(1793995200, buf);
    CHECK_EQ(std::string(buf, 29), std::string("Fri, 06 Nov 2026 20:00:00 GMT"));
}

TEST(http_percent_decode) {
    std::string out;
    CHECK(percent_decode("/a%20b/%7Ec", &out));
    CHECK_EQ(out, std::string("/a b/~c"));
    CHECK(!percent_decode("/bad%2", &out));
    CHECK(!percent_decode("/bad%zz", &out));
    CHECK(percent_decode("/end%41", &out));
    CHECK_EQ(out, std::string("/endA"));
}

TEST(http_canonical_path) {
    struct Case { const char* raw; const char* want; } cases[] = {
        {"/api/hola", "/api/hola"},
        {"/api/%68ola", "/api/hola"},                      // percent-decoded
        {"/api/hola;a=1", "/api/hola"},                    // matrix parameters are not part of the path
        {"/api;x=1;y/hola;v=2/", "/api/hola/"},            // ... in any segment; trailing slash kept
        {"/api/../api/health", "/api/health"},             // dot segments
        {"/api/./hola", "/api/hola"},
        {"/api//hola", "/api/hola"},                       // empty segments
        {"/../../api/hola", "/api/hola"},                  // never above the root
        {"/api/%2e%2e/api/%2E/hola", "/api/hola"},         // decoded first, then resolved
        {"/webjars/..%2f..%2fsecret.txt", "/webjars/..%2f..%2fsecret.txt"},  // an encoded slash is no segment boundary
        {"/a%5Cb", "/a%5Cb"},
        {"/api/hola%3Bx", "/api/hola;x"},                  // an encoded semicolon is data
        {"/api/hola/.", "/api/hola/"},
        {"/api/hola/..", "/api/"},
        {"/", "/"},
        {"/..", "/"},
        {"/.well-known/x", "/.well-known/x"},
        {"/a%20b/%7Ec", "/a b/~c"},
        {"*", "*"},
    };
    std::string out;
    for (const auto& c : cases) {
        CHECK(canonicalise_path(c.raw, &out));
        CHECK_EQ(out, std::string(c.want));
        // Canonical form is a fixed point (an encoded slash stays encoded, so it is stable too) -
        // except where decoding produced a literal ';', which a second pass would take for a parameter.
        std::string again;
        CHECK(canonicalise_path(out, &again));
        if (out.find(';') == std::string::npos) CHECK_EQ(again, out);
    }
    CHECK(!canonicalise_path("/bad%2", &out));
    CHECK(!canonicalise_path("/bad%zz", &out));
    CHECK(!canonicalise_path("/bad%", &out));
    CHECK(!canonicalise_path("/nul%00byte", &out));
    // The parser flags exactly the paths that need the slow path.
    auto flagged = [](const char* target) {
        const std::string rq = std::string("GET ") + target + " HTTP/1.1\r\n\r\n";
        RequestHead h;
        CHECK(parse_request_head(rq.data(), rq.size(), &h) == ParseStatus::kComplete);
        CHECK_EQ(std::string(h.raw_path), std::string(h.path));
        return !h.path_is_canonical;
    };
    CHECK(!flagged("/api/hola"));
    CHECK(!flagged("/api/hola?x=%41;y=/../"));  // the query is not the path
    CHECK(!flagged("/webjars/swagger-ui/2.1.4/swagger-ui.js"));
    CHECK(flagged("/api/%68ola"));
    CHECK(flagged("/api/hola;a=1"));
    CHECK(flagged("/api/../api/health"));
    CHECK(flagged("/api//hola"));
    CHECK(flagged("/api/hola/."));
}

// ---- trace -------------------------------------------------------------------------------------

TEST(trace_ids_and_decisions) {
    uint64_t v = 0;
    CHECK(parse_trace_id("463ac35c9f6413ad", &v));
    CHECK_EQ(v, uint64_t{0x463ac35c9f6413adull});
    CHECK(parse_trace_id("1", &v));
    CHECK_EQ(v, uint64_t{1});
    CHECK(parse_trace_id("0000000000000001463ac35c9f6413ad", &v));  // 128-bit: keep the low half
    CHECK_EQ(v, uint64_t{0x463ac35c9f6413adull});
    CHECK(!parse_trace_id("", &v));
    CHECK(!parse_trace_id("xyz", &v));
    CHECK_EQ(trace_id_string(0xabc), std::string("0000000000000abc"));

    IdGenerator ids(42);
    B3Headers h;
    TraceContext c = start_server_trace(h, &ids);
    CHECK(c.sampled && !c.joined && !c.has_parent);
    CHECK_EQ(c.trace_id, c.span_id);
    CHECK(c.span_id != 0);

    h.trace_id = "463ac35c9f6413ad";
    h.span_id = "a2fb4a1d1a96d312";
    c = start_server_trace(h, &ids);  // ids without X-B3-Sampled are ignored (Brave 3.9)
    CHECK(c.sampled && !c.joined);
    h.sampled = "1";
    h.has_sampled = true;
    c = start_server_trace(h, &ids);
    CHECK(c.sampled && c.joined && !c.has_parent);
    CHECK_EQ(c.trace_id, uint64_t{0x463ac35c9f6413adull});
    h.parent_span_id = "0020000000000001";
    c = start_server_trace(h, &ids);
    CHECK(c.has_parent);
    h.sampled = "0";
    c = start_server_trace(h, &ids);
    CHECK(!c.sampled);
    h.sampled = "FALSE";
    CHECK(!start_server_trace(h, &ids).sampled);

    h.sampled = "true";
    const TraceContext server = start_server_trace(h, &ids);
    const TraceContext client = start_client_trace(server, &ids);
    CHECK(client.sampled && client.has_parent);
    CHECK_EQ(client.trace_id, server.trace_id);
    CHECK_EQ(client.parent_id, server.span_id);
    CHECK(client.span_id != server.span_id);
    TraceContext off;
    CHECK(!start_client_trace(off, &ids).sampled);
}

TEST(trace_span_json) {
    Span s;
    s.trace_id = 0x1;
    s.id = 0x2;
    s.parent_id = 0x3;
    s.has_parent = true;
    s.timestamp_us = 1000;
    s.duration_us = 250;
    s.first_us = 1000;
    s.last_us = 1250;
    s.tags.push_back(BinaryAnnotation{"http.url", "http://localhost:8080/api/hola?q=\"x\""});
    Endpoint ep;
    ep.service_name = "hola";
    ep.ipv4 = 0x0A000005;
    std::string json;
    span_append_json(&json, s, ep);
    CHECK_EQ(json,
             std::string("{\"traceId\":\"0000000000000001\",\"id\":\"0000000000000002\",\"name\":\"get\","
                         "\"parentId\":\"0000000000000003\",\"timestamp\":1000,\"duration\":250,\"annotations\":["
                         "{\"timestamp\":1000,\"value\":\"sr\",\"endpoint\":{\"serviceName\":\"hola\",\"ipv4\":\"10.0.0.5\"}},"
                         "{\"timestamp\":1250,\"value\":\"ss\",\"endpoint\":{\"serviceName\":\"hola\",\"ipv4\":\"10.0.0.5\"}}],"
                         "\"binaryAnnotations\":[{\"key\":\"http.url\",\"value\":\"http://localhost:8080/api/hola?q=\\\"x\\\"\","
                         "\"endpoint\":{\"serviceName\":\"hola\",\"ipv4\":\"10.0.0.5\"}}]}"));
}

// ---- histogram ---------------------------------------------------------------------------------

TEST(histogram_accuracy) {
    Histogram h;
    for (uint64_t v = 1; v <= 100000; ++v) h.record(v * 1000);  // 1 us .. 100 ms
    CHECK_EQ(h.count(), uint64_t{100000});
    auto within = [](uint64_t got, double want) { return std::abs(static_cast<double>(got) - want) <= want * 0.01; };
    CHECK(within(h.percentile(50), 50000e3));
    CHECK(within(h.percentile(99), 99000e3));
    CHECK(within(h.percentile(99.9), 99900e3));
    CHECK_EQ(h.percentile(100), uint64_t{100000000});
    CHECK_EQ(h.min(), uint64_t{1000});
    for (uint64_t v : {0ull, 1ull, 127ull, 128ull, 129ull, 255ull, 256ull, 1000000ull, 123456789012ull, ~0ull}) {
        const size_t idx = Histogram::index_of(v);
        CHECK(idx < Histogram::kBuckets);
        CHECK(Histogram::upper_edge(idx) >= v);
        if (idx > 0) CHECK(Histogram::upper_edge(idx - 1) < v);
    }
    Histogram a, b;
    a.record(10);
    b.record(1000);
    a.merge(b);
    CHECK_EQ(a.count(), uint64_t{2});
    CHECK_EQ(a.max(), uint64_t{1000});
}
