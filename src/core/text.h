// UTF-8 / UTF-16 helpers shared by the properties parser, the formatter and the JSON codec.
//
// The reference runs on the JVM, where strings are UTF-16 sequences: widths and precisions in
// String.format count UTF-16 units, String.hashCode walks UTF-16 units, and \uXXXX escapes in
// property files and JSON name UTF-16 units (surrogate pairs for astral characters). This
// service keeps everything as UTF-8 bytes and only converts at those few points.
#pragma once

#include <cstdint>
#include <string>
#include <string_view>
#include <vector>

namespace hola {

// Appends the UTF-8 encoding of code point `cp`. Lone surrogates and out-of-range values are
// written as '?', which is what the JVM produces when it encodes such a string to UTF-8.
void append_utf8(std::string* out, uint32_t cp);

// Decodes one code point starting at s[*pos] and advances *pos. Malformed bytes decode as
// U+FFFD and advance by one byte so that the caller always makes progress.
uint32_t decode_utf8(std::string_view s, size_t* pos);

// Number of UTF-16 code units the string would occupy on the JVM.
size_t utf16_length(std::string_view s);

// Longest prefix of `s` that is at most `units` UTF-16 units long (never splits a code point).
std::string_view utf16_prefix(std::string_view s, size_t units);

// java.lang.String.hashCode() of the string's UTF-16 form.
int32_t java_string_hash(std::string_view s);

// Upper-cases ASCII and Latin-1 letters the way String.toUpperCase() does in the ROOT/English
// locale (including the one-to-two mapping of U+00DF to "SS"). Other scripts are left alone.
std::string to_upper_java(std::string_view s);

// Incremental assembler for sequences of \uXXXX escapes: pairs a high surrogate with the low
// surrogate that follows it and flushes unpaired halves as '?'.
class Utf16Assembler {
 public:
    explicit Utf16Assembler(std::string* out) : out_(out) {}
    void push_unit(uint16_t unit);
    // Call before appending anything that is not a \uXXXX unit, and once at the end.
    void flush();

 private:
    std::string* out_;
    uint16_t pending_high_ = 0;
};

// Parses a CPU list such as "0,2,4-7" into {0,2,4,5,6,7}. Returns false on anything else.
bool parse_cpu_list(std::string_view text, std::vector<int>* out);

inline bool is_hex_digit(char c) {
    return (c >= '0' && c <= '9') || (c >= 'a' && c <= 'f') || (c >= 'A' && c <= 'F');
}
inline int hex_value(char c) {
    if (c >= '0' && c <= '9') return c - '0';
    if (c >= 'a' && c <= 'f') return c - 'a' + 10;
    return c - 'A' + 10;
}

inline bool ascii_iequals(std::string_view a, std::string_view b) {
    if (a.size() != b.size()) return false;
    for (size_t i = 0; i < a.size(); ++i) {
        char x = a[i], y = b[i];
        if (x >= 'A' && x <= 'Z') x = static_cast<char>(x + 32);
        if (y >= 'A' && y <= 'Z') y = static_cast<char>(y + 32);
        if (x != y) return false;
    }
    return true;
}

inline std::string_view trim_ows(std::string_view s) {
    while (!s.empty() && (s.front() == ' ' || s.front() == '\t')) s.remove_prefix(1);
    while (!s.empty() && (s.back() == ' ' || s.back() == '\t')) s.remove_suffix(1);
    return s;
}

}  // namespace hola
