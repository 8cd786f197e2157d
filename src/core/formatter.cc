#include "core/formatter.h"

#include <cstdint>
#include <cstdio>
#include <vector>

#include "core/text.h"

namespace hola {
namespace {

enum Flag : unsigned {
    kLeftJustify = 1u << 0,   // '-'
    kUpperCase = 1u << 1,     // set by an upper-case conversion letter
    kAlternate = 1u << 2,     // '#'
    kPlus = 1u << 3,          // '+'
    kLeadingSpace = 1u << 4,  // ' '
    kZeroPad = 1u << 5,       // '0'
    kGroup = 1u << 6,         // ','
    kParentheses = 1u << 7,   // '('
    kPrevious = 1u << 8,      // '<'
};

unsigned flag_of(char c) {
    switch (c) {
        case '-': return kLeftJustify;
        case '#': return kAlternate;
        case '+': return kPlus;
        case ' ': return kLeadingSpace;
        case '0': return kZeroPad;
        case ',': return kGroup;
        case '(': return kParentheses;
        case '<': return kPrevious;
        default: return 0;
    }
}

std::string flags_to_string(unsigned f) {
    std::string s;
    if (f & kLeftJustify) s.push_back('-');
    if (f & kUpperCase) s.push_back('^');
    if (f & kAlternate) s.push_back('#');
    if (f & kPlus) s.push_back('+');
    if (f & kLeadingSpace) s.push_back(' ');
    if (f & kZeroPad) s.push_back('0');
    if (f & kGroup) s.push_back(',');
    if (f & kParentheses) s.push_back('(');
    if (f & kPrevious) s.push_back('<');
    return s;
}

inline bool is_digit(char c) { return c >= '0' && c <= '9'; }
inline bool is_letter(char c) { return (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z'); }

// Argument index classes used by java.util.Formatter.
constexpr int kIndexText = -2;      // consumes no argument (%n, %%)
constexpr int kIndexPrevious = -1;  // '<' flag
constexpr int kIndexOrdinary = 0;

struct Spec {
    bool literal = false;
    std::string_view text;  // literal text, or the raw specifier for error messages
    int index = kIndexOrdinary;
    unsigned flags = 0;
    long width = -1;
    long precision = -1;
    bool date_time = false;
    char conv = 0;  // lower-cased conversion letter
};

bool parse_number(std::string_view digits, long* out) {
    long v = 0;
    for (char c : digits) {
        v = v * 10 + (c - '0');
        if (v > INT32_MAX) return false;
    }
    *out = v;
    return true;
}

// Tries to match one format specifier at pattern[pos] == '%', following the JVM's grammar
//   %[index$][flags][width][.precision][t|T]conversion
// Returns the length matched, or 0 when the text at pos is not a specifier.
size_t match_specifier(std::string_view p, size_t pos, std::string_view* index, std::string_view* flags,
                       std::string_view* width, std::string_view* precision, bool* dt, bool* dt_upper,
                       char* conv) {
    const size_t n = p.size();
    size_t j = pos + 1;
    size_t k = j;
    while (k < n && is_digit(p[k])) ++k;
    if (k > j && k < n && p[k] == '$') {
        *index = p.substr(j, k - j);
        j = k + 1;
    } else {
        *index = {};
    }
    size_t f = j;
    while (j < n && flag_of(p[j]) != 0) ++j;
    *flags = p.substr(f, j - f);
    size_t w = j;
    while (j < n && is_digit(p[j])) ++j;
    *width = p.substr(w, j - w);
    *precision = {};
    if (j < n && p[j] == '.') {
        size_t q = j + 1;
        while (q < n && is_digit(p[q])) ++q;
        if (q == j + 1) return 0;  // '.' must be followed by digits
        *precision = p.substr(j + 1, q - j - 1);
        j = q;
    }
    *dt = false;
    *dt_upper = false;
    if (j + 1 < n && (p[j] == 't' || p[j] == 'T') && (is_letter(p[j + 1]) || p[j + 1] == '%')) {
        *dt = true;
        *dt_upper = p[j] == 'T';
        ++j;
    }
    if (j >= n || !(is_letter(p[j]) || p[j] == '%')) return 0;
    *conv = p[j];
    return j + 1 - pos;
}

bool is_general(char c) { return c == 'b' || c == 's' || c == 'h'; }
bool is_character(char c) { return c == 'c'; }
bool is_integer(char c) { return c == 'd' || c == 'o' || c == 'x'; }
bool is_float(char c) { return c == 'e' || c == 'g' || c == 'f' || c == 'a'; }
bool is_text(char c) { return c == 'n' || c == '%'; }

bool is_valid_conversion(char raw) {
    switch (raw) {
        case 'b': case 'B': case 's': case 'S': case 'h': case 'H':
        case 'c': case 'C':
        case 'd': case 'o': case 'x': case 'X':
        case 'e': case 'E': case 'g': case 'G': case 'f': case 'a': case 'A':
        case 'n': case '%':
        case 't':
            return true;
        default:
            return false;
    }
}

bool is_valid_date_time(char c) {
    static const std::string_view k = "HIklMNLSQpszZABbhaCYyjmdeRTrDFc";
    return k.find(c) != std::string_view::npos;
}

std::string fail(const char* exception, std::string detail) {
    std::string s = "java.util.";
    s += exception;
    if (!detail.empty()) {
        s += ": ";
        s += detail;
    }
    return s;
}

// Checks a parsed specifier the way FormatSpecifier's constructor does. Returns an empty string
// when the specifier is acceptable.
std::string check_spec(const Spec& s) {
    const std::string raw(s.text);
    const unsigned bad_general = kPlus | kLeadingSpace | kZeroPad | kGroup | kParentheses;
    if (s.date_time) {
        if (s.precision != -1) return fail("IllegalFormatPrecisionException", std::to_string(s.precision));
        if (!is_valid_date_time(s.conv)) return fail("UnknownFormatConversionException", std::string("t") + s.conv);
        if (s.flags & (kAlternate | bad_general))
            return fail("FormatFlagsConversionMismatchException", flags_to_string(s.flags));
        if (s.width == -1 && (s.flags & kLeftJustify)) return fail("MissingFormatWidthException", raw);
        return {};
    }
    if (is_general(s.conv)) {
        if ((s.conv == 'b' || s.conv == 'h') && (s.flags & kAlternate))
            return fail("FormatFlagsConversionMismatchException", "Conversion = " + std::string(1, s.conv) + ", Flags = #");
        if (s.width == -1 && (s.flags & kLeftJustify)) return fail("MissingFormatWidthException", raw);
        if (s.flags & bad_general)
            return fail("FormatFlagsConversionMismatchException",
                        "Conversion = " + std::string(1, s.conv) + ", Flags = " + flags_to_string(s.flags & bad_general));
        return {};
    }
    if (is_character(s.conv)) {
        if (s.precision != -1) return fail("IllegalFormatPrecisionException", std::to_string(s.precision));
        if (s.flags & (kAlternate | bad_general))
            return fail("FormatFlagsConversionMismatchException", flags_to_string(s.flags));
        if (s.width == -1 && (s.flags & kLeftJustify)) return fail("MissingFormatWidthException", raw);
        return {};
    }
    if (is_integer(s.conv) || is_float(s.conv)) {
        if (is_integer(s.conv) && s.precision != -1)
            return fail("IllegalFormatPrecisionException", std::to_string(s.precision));
        if (s.width == -1 && (s.flags & (kLeftJustify | kZeroPad))) return fail("MissingFormatWidthException", raw);
        if (((s.flags & kPlus) && (s.flags & kLeadingSpace)) || ((s.flags & kLeftJustify) && (s.flags & kZeroPad)))
            return fail("IllegalFormatFlagsException", flags_to_string(s.flags));
        return {};
    }
    if (is_text(s.conv)) {
        if (s.precision != -1) return fail("IllegalFormatPrecisionException", std::to_string(s.precision));
        if (s.conv == '%') {
            if (s.flags != kLeftJustify && s.flags != 0)
                return fail("IllegalFormatFlagsException", flags_to_string(s.flags));
            if (s.width == -1 && (s.flags & kLeftJustify)) return fail("MissingFormatWidthException", raw);
        } else {
            if (s.width != -1) return fail("IllegalFormatWidthException", std::to_string(s.width));
            if (s.flags != 0) return fail("IllegalFormatFlagsException", flags_to_string(s.flags));
        }
        return {};
    }
    return fail("UnknownFormatConversionException", std::string(1, s.conv));
}

void append_justified(std::string* out, std::string_view s, const Spec& spec) {
    if (spec.width == -1) {
        out->append(s);
        return;
    }
    const size_t len = utf16_length(s);
    const size_t width = static_cast<size_t>(spec.width);
    const size_t pad = width > len ? width - len : 0;
    if (spec.flags & kLeftJustify) {
        out->append(s);
        out->append(pad, ' ');
    } else {
        out->append(pad, ' ');
        out->append(s);
    }
}

void append_string_conversion(std::string* out, std::string_view value, const Spec& spec) {
    std::string_view v = value;
    if (spec.precision != -1 && static_cast<size_t>(spec.precision) < utf16_length(v))
        v = utf16_prefix(v, static_cast<size_t>(spec.precision));
    if (spec.flags & kUpperCase) {
        const std::string upper = to_upper_java(v);
        append_justified(out, upper, spec);
    } else {
        append_justified(out, v, spec);
    }
}

}  // namespace

FormatResult java_format(std::string_view pattern, std::string_view arg) {
    FormatResult result;
    std::vector<Spec> specs;

    // Phase 1: split into literals and specifiers. The JVM parses the whole pattern before it
    // formats anything, so a malformed tail fails the call even if the head was fine.
    size_t i = 0;
    const size_t n = pattern.size();
    size_t literal_start = 0;
    while (i < n) {
        if (pattern[i] != '%') {
            ++i;
            continue;
        }
        std::string_view index, flags, width, precision;
        bool dt = false, dt_upper = false;
        char conv = 0;
        const size_t len = match_specifier(pattern, i, &index, &flags, &width, &precision, &dt, &dt_upper, &conv);
        if (len == 0) {
            const char c = (i + 1 < n) ? pattern[i + 1] : '%';
            result.error = fail("UnknownFormatConversionException", std::string("Conversion = '") + c + "'");
            return result;
        }
        if (i > literal_start) {
            Spec lit;
            lit.literal = true;
            lit.text = pattern.substr(literal_start, i - literal_start);
            specs.push_back(lit);
        }
        Spec s;
        s.text = pattern.substr(i, len);
        if (!index.empty()) {
            long v = 0;
            s.index = parse_number(index, &v) ? static_cast<int>(v) : kIndexOrdinary;
        }
        for (char c : flags) {
            const unsigned f = flag_of(c);
            if (s.flags & f) {
                result.error = fail("DuplicateFormatFlagsException", "Flags = '" + std::string(1, c) + "'");
                return result;
            }
            s.flags |= f;
        }
        if (s.flags & kPrevious) s.index = kIndexPrevious;
        if (!width.empty() && !parse_number(width, &s.width)) s.width = -1;
        if (!precision.empty() && !parse_number(precision, &s.precision)) s.precision = -1;
        s.date_time = dt;
        if (dt) {
            if (dt_upper) s.flags |= kUpperCase;
            s.conv = conv;
        } else {
            if (!is_valid_conversion(conv)) {
                result.error = fail("UnknownFormatConversionException", std::string("Conversion = '") + conv + "'");
                return result;
            }
            if (conv >= 'A' && conv <= 'Z') {
                s.flags |= kUpperCase;
                conv = static_cast<char>(conv + 32);
            }
            s.conv = conv;
            if (is_text(conv)) s.index = kIndexText;
        }
        std::string err = check_spec(s);
        if (!err.empty()) {
            result.error = std::move(err);
            return result;
        }
        specs.push_back(s);
        i += len;
        literal_start = i;
    }
    if (literal_start < n) {
        Spec lit;
        lit.literal = true;
        lit.text = pattern.substr(literal_start);
        specs.push_back(lit);
    }

    // Phase 2: format. There is exactly one argument (index 0).
    constexpr int kArgCount = 1;
    int last = -1;
    int last_ordinary = -1;
    std::string out;
    out.reserve(pattern.size() + arg.size());
    for (const Spec& s : specs) {
        if (s.literal) {
            out.append(s.text);
            continue;
        }
        if (s.index == kIndexText) {
            if (s.conv == 'n') {
                out.push_back('\n');
            } else {
                out.push_back('%');  // JDK 8 ignores the width of %% (the reference image is openjdk8)
            }
            continue;
        }
        int arg_index;
        if (s.index == kIndexPrevious) {
            if (last < 0 || last > kArgCount - 1) {
                result.error = fail("MissingFormatArgumentException", "Format specifier '" + std::string(s.text) + "'");
                return result;
            }
            arg_index = last;
        } else if (s.index == kIndexOrdinary) {
            ++last_ordinary;
            last = last_ordinary;
            arg_index = last_ordinary;
        } else {
            last = s.index - 1;
            arg_index = last;
        }
        if (arg_index > kArgCount - 1) {
            result.error = fail("MissingFormatArgumentException", "Format specifier '" + std::string(s.text) + "'");
            return result;
        }
        if (s.date_time || is_character(s.conv) || is_integer(s.conv) || is_float(s.conv)) {
            result.error = fail("IllegalFormatConversionException",
                                std::string(1, s.conv) + " != java.lang.String");
            return result;
        }
        switch (s.conv) {
            case 's':
                if (s.flags & kAlternate) {
                    result.error = fail("FormatFlagsConversionMismatchException", "Conversion = s, Flags = #");
                    return result;
                }
                append_string_conversion(&out, arg, s);
                break;
            case 'b':
                append_string_conversion(&out, "true", s);
                break;
            case 'h': {
                char buf[16];
                std::snprintf(buf, sizeof buf, "%x", static_cast<uint32_t>(java_string_hash(arg)));
                append_string_conversion(&out, buf, s);
                break;
            }
            default:
                result.error = fail("UnknownFormatConversionException", std::string(1, s.conv));
                return result;
        }
    }
    result.ok = true;
    result.text = std::move(out);
    return result;
}

}  // namespace hola
