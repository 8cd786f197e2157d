#include "core/json.h"

#include "core/text.h"

namespace hola {

void json_append_string(std::string* out, std::string_view s) {
    static const char kHex[] = "0123456789ABCDEF";
    out->push_back('"');
    size_t run = 0;
    for (size_t i = 0; i < s.size(); ++i) {
        const unsigned char c = static_cast<unsigned char>(s[i]);
        if (c >= 0x20 && c != '"' && c != '\\') continue;
        out->append(s.substr(run, i - run));
        run = i + 1;
        switch (c) {
            case '"': out->append("\\\""); break;
            case '\\': out->append("\\\\"); break;
            case '\b': out->append("\\b"); break;
            case '\f': out->append("\\f"); break;
            case '\n': out->append("\\n"); break;
            case '\r': out->append("\\r"); break;
            case '\t': out->append("\\t"); break;
            default:
                out->append("\\u00");
                out->push_back(kHex[c >> 4]);
                out->push_back(kHex[c & 15]);
                break;
        }
    }
    out->append(s.substr(run));
    out->push_back('"');
}

std::string json_encode_string_list(const JsonStringList& items) {
    std::string out;
    out.push_back('[');
    bool first = true;
    for (const auto& item : items) {
        if (!first) out.push_back(',');
        first = false;
        if (item) {
            json_append_string(&out, *item);
        } else {
            out.append("null");
        }
    }
    out.push_back(']');
    return out;
}

namespace {

struct Cursor {
    std::string_view s;
    size_t i = 0;
    void skip_ws() {
        while (i < s.size() && (s[i] == ' ' || s[i] == '\t' || s[i] == '\n' || s[i] == '\r')) ++i;
    }
    bool at_end() const { return i >= s.size(); }
    char peek() const { return s[i]; }
};

bool parse_string(Cursor* c, std::string* out, std::string* error) {
    // c->peek() == '"'
    ++c->i;
    out->clear();
    Utf16Assembler utf16(out);
    while (true) {
        if (c->at_end()) {
            *error = "unterminated string";
            return false;
        }
        const char ch = c->s[c->i++];
        if (ch == '"') break;
        if (static_cast<unsigned char>(ch) < 0x20) {
            *error = "unescaped control character in string";
            return false;
        }
        if (ch != '\\') {
            utf16.flush();
            out->push_back(ch);
            continue;
        }
        if (c->at_end()) {
            *error = "unterminated escape";
            return false;
        }
        const char e = c->s[c->i++];
        if (e == 'u') {
            if (c->i + 4 > c->s.size()) {
                *error = "truncated \\u escape";
                return false;
            }
            unsigned v = 0;
            for (int k = 0; k < 4; ++k) {
                const char h = c->s[c->i + k];
                if (!is_hex_digit(h)) {
                    *error = "bad \\u escape";
                    return false;
                }
                v = (v << 4) | static_cast<unsigned>(hex_value(h));
            }
            c->i += 4;
            utf16.push_unit(static_cast<uint16_t>(v));
            continue;
        }
        utf16.flush();
        switch (e) {
            case '"': out->push_back('"'); break;
            case '\\': out->push_back('\\'); break;
            case '/': out->push_back('/'); break;
            case 'b': out->push_back('\b'); break;
            case 'f': out->push_back('\f'); break;
            case 'n': out->push_back('\n'); break;
            case 'r': out->push_back('\r'); break;
            case 't': out->push_back('\t'); break;
            default:
                *error = "unknown escape";
                return false;
        }
    }
    utf16.flush();
    return true;
}

bool parse_number_text(Cursor* c, std::string* out, std::string* error) {
    const size_t start = c->i;
    auto digits = [&]() {
        size_t n = 0;
        while (!c->at_end() && c->peek() >= '0' && c->peek() <= '9') {
            ++c->i;
            ++n;
        }
        return n;
    };
    if (!c->at_end() && c->peek() == '-') ++c->i;
    if (c->at_end()) {
        *error = "bad number";
        return false;
    }
    if (c->peek() == '0') {
        ++c->i;
    } else if (digits() == 0) {
        *error = "bad number";
        return false;
    }
    if (!c->at_end() && c->peek() == '.') {
        ++c->i;
        if (digits() == 0) {
            *error = "bad number";
            return false;
        }
    }
    if (!c->at_end() && (c->peek() == 'e' || c->peek() == 'E')) {
        ++c->i;
        if (!c->at_end() && (c->peek() == '+' || c->peek() == '-')) ++c->i;
        if (digits() == 0) {
            *error = "bad number";
            return false;
        }
    }
    out->assign(c->s.substr(start, c->i - start));
    return true;
}

bool match_literal(Cursor* c, std::string_view word) {
    if (c->s.substr(c->i, word.size()) != word) return false;
    c->i += word.size();
    return true;
}

}  // namespace

bool json_decode_string_list(std::string_view text, JsonStringList* out, std::string* error) {
    std::string local_error;
    if (!error) error = &local_error;
    out->clear();
    Cursor c{text, 0};
    c.skip_ws();
    if (c.at_end() || c.peek() != '[') {
        *error = "expected a JSON array";
        return false;
    }
    ++c.i;
    c.skip_ws();
    if (!c.at_end() && c.peek() == ']') {
        ++c.i;
    } else {
        while (true) {
            c.skip_ws();
            if (c.at_end()) {
                *error = "unterminated array";
                return false;
            }
            const char ch = c.peek();
            if (ch == '"') {
                std::string s;
                if (!parse_string(&c, &s, error)) return false;
                out->emplace_back(std::move(s));
            } else if (ch == '-' || (ch >= '0' && ch <= '9')) {
                std::string s;
                if (!parse_number_text(&c, &s, error)) return false;
                out->emplace_back(std::move(s));
            } else if (match_literal(&c, "true")) {
                out->emplace_back("true");
            } else if (match_literal(&c, "false")) {
                out->emplace_back("false");
            } else if (match_literal(&c, "null")) {
                out->emplace_back(std::nullopt);
            } else {
                *error = (ch == '[' || ch == '{') ? "nested value where a string was expected" : "unexpected character";
                return false;
            }
            c.skip_ws();
            if (c.at_end()) {
                *error = "unterminated array";
                return false;
            }
            if (c.peek() == ',') {
                ++c.i;
                continue;
            }
            if (c.peek() == ']') {
                ++c.i;
                break;
            }
            *error = "expected ',' or ']'";
            return false;
        }
    }
    // Jackson's readValue stops after the first value and does not complain about trailing
    // content by default; be equally lenient.
    return true;
}

}  // namespace hola
