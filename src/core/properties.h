// Parser for the java.util.Properties text format.
//
// The reference reads its external greeting file with `new Properties().load(new FileReader(..))`
// (/root/reference/src/main/java/com/redhat/developers/msa/hola/Configuration.java:26-33), so the
// bytes of `translation.properties` only mean what that loader says they mean. In particular the
// shipped file ends in the two characters backslash + 'n' (/root/reference/translation.properties:15),
// which the loader turns into one line feed at the end of the greeting.
//
// Format rules implemented here (from the java.util.Properties#load(Reader) specification):
//   * natural lines end at \n, \r or \r\n; a logical line may span several natural lines when the
//     line terminator is preceded by an odd number of backslashes, in which case the backslash,
//     the terminator and the next line's leading white space are dropped;
//   * white space is space, \t and \f; blank lines and lines whose first non-blank character is
//     '#' or '!' are skipped, and a comment line never continues onto the next line;
//   * the key ends at the first unescaped '=', ':' or white space; white space around a single
//     '=' or ':' separator is dropped; the rest of the logical line is the value;
//   * escapes: \t \n \r \f, \uXXXX (exactly four hex digits, otherwise the whole load fails) and
//     backslash + any other character, which stands for that character;
//   * when a key occurs twice the later entry wins.
#pragma once

#include <string>
#include <string_view>
#include <unordered_map>
#include <utility>
#include <vector>

namespace hola {

using PropertyList = std::vector<std::pair<std::string, std::string>>;

// Parses `text`; entries are appended to `out` in file order (duplicates included). Returns false
// and sets `error` when the loader would have thrown (malformed \uXXXX).
bool parse_properties(std::string_view text, PropertyList* out, std::string* error);

class Properties {
 public:
    // Replaces the content with the entries of `text`. On failure the map is left empty, which is
    // what the reference's swallow-everything catch block amounts to (Configuration.java:34-36).
    bool load(std::string_view text, std::string* error = nullptr);
    // Reads and parses a file. A missing or unreadable file counts as failure.
    bool load_file(const std::string& path, std::string* error = nullptr);

    const std::string* get(std::string_view key) const;
    size_t size() const { return map_.size(); }
    const std::unordered_map<std::string, std::string>& entries() const { return map_; }

 private:
    std::unordered_map<std::string, std::string> map_;
};

}  // namespace hola
