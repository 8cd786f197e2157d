// Minimal JSON support: exactly what the service needs and nothing else.
//
//  * encoding a list of strings, which is the body of GET /api/hola-chaining
//    (/root/reference/src/main/java/com/redhat/developers/msa/hola/HolaResource.java:73-78);
//  * decoding the downstream reply into a list of strings the way the reference's JacksonDecoder
//    does for `List<String>` (HolaResource.java:110, AlohaService.java:25-26): the document must
//    be an array; string elements are taken as they are, numbers and booleans are coerced to
//    their text, null stays null, nested arrays/objects make the decode fail (and a failed
//    decode sends the caller to the circuit-breaker fallback);
//  * string escaping for the span reporter, the metrics stream and the API description.
#pragma once

#include <optional>
#include <string>
#include <string_view>
#include <vector>

namespace hola {

// Appends `s` as a quoted JSON string: `"` and `\` escaped, control characters as \b \f \n \r \t
// or \u00XX (upper-case hex, as Jackson writes them), everything else copied through as UTF-8.
void json_append_string(std::string* out, std::string_view s);

using JsonStringList = std::vector<std::optional<std::string>>;

// Compact array, no white space: ["a","b",null]
std::string json_encode_string_list(const JsonStringList& items);

// Returns false (with a reason in *error) unless `text` is a single JSON array of scalars
// surrounded by optional white space.
bool json_decode_string_list(std::string_view text, JsonStringList* out, std::string* error);

}  // namespace hola
