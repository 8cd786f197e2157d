#include "core/text.h"

namespace hola {

void append_utf8(std::string* out, uint32_t cp) {
    if (cp < 0x80) {
        out->push_back(static_cast<char>(cp));
    } else if (cp < 0x800) {
        out->push_back(static_cast<char>(0xC0 | (cp >> 6)));
        out->push_back(static_cast<char>(0x80 | (cp & 0x3F)));
    } else if (cp >= 0xD800 && cp <= 0xDFFF) {
        out->push_back('?');
    } else if (cp < 0x10000) {
        out->push_back(static_cast<char>(0xE0 | (cp >> 12)));
        out->push_back(static_cast<char>(0x80 | ((cp >> 6) & 0x3F)));
        out->push_back(static_cast<char>(0x80 | (cp & 0x3F)));
    } else if (cp <= 0x10FFFF) {
        out->push_back(static_cast<char>(0xF0 | (cp >> 18)));
        out->push_back(static_cast<char>(0x80 | ((cp >> 12) & 0x3F)));
        out->push_back(static_cast<char>(0x80 | ((cp >> 6) & 0x3F)));
        out->push_back(static_cast<char>(0x80 | (cp & 0x3F)));
    } else {
        out->push_back('?');
    }
}

uint32_t decode_utf8(std::string_view s, size_t* pos) {
    const size_t i = *pos;
    const unsigned char b0 = static_cast<unsigned char>(s[i]);
    if (b0 < 0x80) {
        *pos = i + 1;
        return b0;
    }
    int need = 0;
    uint32_t cp = 0;
    uint32_t min = 0;
    if ((b0 & 0xE0) == 0xC0) {
        need = 1; cp = b0 & 0x1F; min = 0x80;
    } else if ((b0 & 0xF0) == 0xE0) {
        need = 2; cp = b0 & 0x0F; min = 0x800;
    } else if ((b0 & 0xF8) == 0xF0) {
        need = 3; cp = b0 & 0x07; min = 0x10000;
    } else {
        *pos = i + 1;
        return 0xFFFD;
    }
    if (i + static_cast<size_t>(need) >= s.size()) {  // truncated sequence
        *pos = i + 1;
        return 0xFFFD;
    }
    for (int k = 1; k <= need; ++k) {
        const unsigned char b = static_cast<unsigned char>(s[i + k]);
        if ((b & 0xC0) != 0x80) {
            *pos = i + 1;
            return 0xFFFD;
        }
        cp = (cp << 6) | (b & 0x3F);
    }
    if (cp < min || cp > 0x10FFFF || (cp >= 0xD800 && cp <= 0xDFFF)) {
        *pos = i + 1;
        return 0xFFFD;
    }
    *pos = i + 1 + need;
    return cp;
}

size_t utf16_length(std::string_view s) {
    size_t units = 0;
    size_t pos = 0;
    while (pos < s.size()) {
        const uint32_t cp = decode_utf8(s, &pos);
        units += cp >= 0x10000 ? 2 : 1;
    }
    return units;
}

std::string_view utf16_prefix(std::string_view s, size_t units) {
    size_t used = 0;
    size_t pos = 0;
    while (pos < s.size()) {
        size_t next = pos;
        const uint32_t cp = decode_utf8(s, &next);
        const size_t w = cp >= 0x10000 ? 2 : 1;
        if (used + w > units) break;
        used += w;
        pos = next;
    }
    return s.substr(0, pos);
}

int32_t java_string_hash(std::string_view s) {
    uint32_t h = 0;
    size_t pos = 0;
    while (pos < s.size()) {
        const uint32_t cp = decode_utf8(s, &pos);
        if (cp >= 0x10000) {
            const uint32_t v = cp - 0x10000;
            h = 31 * h + (0xD800 + (v >> 10));
            h = 31 * h + (0xDC00 + (v & 0x3FF));
        } else {
            h = 31 * h + cp;
        }
    }
    return static_cast<int32_t>(h);
}

std::string to_upper_java(std::string_view s) {
    std::string out;
    out.reserve(s.size());
    size_t pos = 0;
    while (pos < s.size()) {
        const size_t start = pos;
        const uint32_t cp = decode_utf8(s, &pos);
        if (cp >= 'a' && cp <= 'z') {
            out.push_back(static_cast<char>(cp - 32));
        } else if (cp == 0xDF) {
            out.append("SS");
        } else if (cp == 0xFF) {
            append_utf8(&out, 0x178);
        } else if (cp == 0xB5) {
            append_utf8(&out, 0x39C);
        } else if (cp >= 0xE0 && cp <= 0xFE && cp != 0xF7) {
            append_utf8(&out, cp - 32);
        } else {
            out.append(s.substr(start, pos - start));
        }
    }
    return out;
}

bool parse_cpu_list(std::string_view text, std::vector<int>* out) {
    auto number = [](std::string_view s, int* v) {
        if (s.empty() || s.size() > 6) return false;
        int n = 0;
        for (char c : s) {
            if (c < '0' || c > '9') return false;
            n = n * 10 + (c - '0');
        }
        *v = n;
        return true;
    };
    size_t i = 0;
    while (i <= text.size()) {
        size_t comma = text.find(',', i);
        if (comma == std::string_view::npos) comma = text.size();
        const std::string_view part = text.substr(i, comma - i);
        i = comma + 1;
        if (part.empty()) continue;
        const size_t dash = part.find('-');
        int a = 0, b = 0;
        if (dash == std::string_view::npos) {
            if (!number(part, &a)) return false;
            out->push_back(a);
        } else {
            if (!number(part.substr(0, dash), &a) || !number(part.substr(dash + 1), &b) || b < a) return false;
            for (int v = a; v <= b; ++v) out->push_back(v);
        }
    }
    return !out->empty();
}

void Utf16Assembler::push_unit(uint16_t unit) {
    if (pending_high_ != 0) {
        if (unit >= 0xDC00 && unit <= 0xDFFF) {
            const uint32_t cp = 0x10000 + ((static_cast<uint32_t>(pending_high_) - 0xD800) << 10) +
                                (unit - 0xDC00);
            pending_high_ = 0;
            append_utf8(out_, cp);
            return;
        }
        out_->push_back('?');
        pending_high_ = 0;
    }
    if (unit >= 0xD800 && unit <= 0xDBFF) {
        pending_high_ = unit;
    } else {
        append_utf8(out_, unit);  // lone low surrogates come out as '?'
    }
}

void Utf16Assembler::flush() {
    if (pending_high_ != 0) {
        out_->push_back('?');
        pending_high_ = 0;
    }
}

}  // namespace hola
