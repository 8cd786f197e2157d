#include "core/properties.h"

#include <cerrno>
#include <cstdio>
#include <cstring>

#include "core/text.h"

namespace hola {
namespace {

inline bool is_space(char c) { return c == ' ' || c == '\t' || c == '\f'; }
inline bool is_eol(char c) { return c == '\n' || c == '\r'; }

// Splits the text into logical lines: comments and blank lines removed, continuations joined.
// Escapes are left in place for the second pass, which needs them to find the key's end.
void split_logical_lines(std::string_view text, std::vector<std::string>* lines) {
    size_t i = 0;
    const size_t n = text.size();
    while (i < n) {
        // Start of a logical line: drop leading blanks and empty natural lines.
        while (i < n && (is_space(text[i]) || is_eol(text[i]))) ++i;
        if (i >= n) break;
        if (text[i] == '#' || text[i] == '!') {
            while (i < n && !is_eol(text[i])) ++i;
            continue;
        }
        std::string line;
        bool backslash = false;  // true when an odd run of backslashes precedes the cursor
        while (i < n) {
            const char c = text[i];
            if (!is_eol(c)) {
                line.push_back(c);
                backslash = (c == '\\') ? !backslash : false;
                ++i;
                continue;
            }
            // Line terminator: consume \r\n as one.
            ++i;
            if (c == '\r' && i < n && text[i] == '\n') ++i;
            if (!backslash) break;
            line.pop_back();  // the continuation backslash itself
            backslash = false;
            while (i < n && is_space(text[i])) ++i;
        }
        if (backslash) line.pop_back();  // dangling backslash at end of input
        lines->push_back(std::move(line));
    }
}

bool unescape(std::string_view in, std::string* out, std::string* error) {
    out->clear();
    out->reserve(in.size());
    Utf16Assembler utf16(out);
    size_t i = 0;
    while (i < in.size()) {
        const char c = in[i++];
        if (c != '\\') {
            utf16.flush();
            out->push_back(c);
            continue;
        }
        if (i >= in.size()) {  // cannot happen for well-formed logical lines; keep it harmless
            utf16.flush();
            break;
        }
        const char e = in[i++];
        if (e == 'u') {
            if (i + 4 > in.size()) {
                if (error) *error = "Malformed \\uxxxx encoding.";
                return false;
            }
            unsigned value = 0;
            for (int k = 0; k < 4; ++k) {
                const char h = in[i + k];
                if (!is_hex_digit(h)) {
                    if (error) *error = "Malformed \\uxxxx encoding.";
                    return false;
                }
                value = (value << 4) | static_cast<unsigned>(hex_value(h));
            }
            i += 4;
            utf16.push_unit(static_cast<uint16_t>(value));
            continue;
        }
        utf16.flush();
        switch (e) {
            case 't': out->push_back('\t'); break;
            case 'n': out->push_back('\n'); break;
            case 'r': out->push_back('\r'); break;
            case 'f': out->push_back('\f'); break;
            default: out->push_back(e); break;
        }
    }
    utf16.flush();
    return true;
}

}  // namespace

bool parse_properties(std::string_view text, PropertyList* out, std::string* error) {
    std::vector<std::string> lines;
    split_logical_lines(text, &lines);
    for (const std::string& line : lines) {
        const size_t limit = line.size();
        size_t key_len = 0;
        size_t value_start = limit;
        bool has_sep = false;
        bool backslash = false;
        while (key_len < limit) {
            const char c = line[key_len];
            if ((c == '=' || c == ':') && !backslash) {
                value_start = key_len + 1;
                has_sep = true;
                break;
            }
            if (is_space(c) && !backslash) {
                value_start = key_len + 1;
                break;
            }
            backslash = (c == '\\') ? !backslash : false;
            ++key_len;
        }
        while (value_start < limit) {
            const char c = line[value_start];
            if (!is_space(c)) {
                if (!has_sep && (c == '=' || c == ':')) {
                    has_sep = true;
                } else {
                    break;
                }
            }
            ++value_start;
        }
        std::string key, value;
        if (!unescape(std::string_view(line).substr(0, key_len), &key, error)) return false;
        if (!unescape(std::string_view(line).substr(value_start), &value, error)) return false;
        out->emplace_back(std::move(key), std::move(value));
    }
    return true;
}

bool Properties::load(std::string_view text, std::string* error) {
    map_.clear();
    PropertyList list;
    if (!parse_properties(text, &list, error)) return false;
    for (auto& kv : list) map_[std::move(kv.first)] = std::move(kv.second);
    return true;
}

bool Properties::load_file(const std::string& path, std::string* error) {
    map_.clear();
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) {
        if (error) *error = path + ": " + std::strerror(errno);
        return false;
    }
    std::string text;
    char buf[4096];
    size_t got;
    while ((got = std::fread(buf, 1, sizeof buf, f)) > 0) text.append(buf, got);
    const bool failed = std::ferror(f) != 0;
    std::fclose(f);
    if (failed) {
        if (error) *error = path + ": read error";
        return false;
    }
    return load(text, error);
}

const std::string* Properties::get(std::string_view key) const {
    auto it = map_.find(std::string(key));
    return it == map_.end() ? nullptr : &it->second;
}

}  // namespace hola
