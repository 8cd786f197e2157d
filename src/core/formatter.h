// The slice of java.util.Formatter that `String.format(translation, hostname)` can reach.
//
// The reference formats the greeting with exactly one String argument
// (/root/reference/src/main/java/com/redhat/developers/msa/hola/HolaResource.java:65). The pattern
// is operator-editable (it comes from the hot-reloaded properties file), so more than "%s" has to
// behave: a pattern the JVM would reject turns into an HTTP 500 there, and does here too.
//
// Supported, with JVM semantics:
//   %s %S      the argument, optional '-' flag, width and .precision (counted in UTF-16 units)
//   %b %B      "true" (the argument is a non-null, non-Boolean object)
//   %h %H      Integer.toHexString(arg.hashCode())
//   %n         line separator ("\n" on Linux, which is where the reference image runs)
//   %%         a literal percent sign (accepts '-' and a width)
//   %1$s       explicit argument index; %<s re-uses the previous argument
// Rejected, as on the JVM: numeric/char/date conversions applied to a String
// (IllegalFormatConversion), a second ordinary argument or an index above 1
// (MissingFormatArgument), unknown conversions or a dangling '%' (UnknownFormatConversion),
// flags that the conversion does not take (FormatFlagsConversionMismatch / IllegalFormatFlags),
// '-' without a width (MissingFormatWidth), precision where none is allowed
// (IllegalFormatPrecision).
#pragma once

#include <string>
#include <string_view>

namespace hola {

struct FormatResult {
    bool ok = false;
    std::string text;   // formatted output when ok
    std::string error;  // JVM-style exception name and detail when !ok
};

// Formats `pattern` with the single string argument `arg`.
FormatResult java_format(std::string_view pattern, std::string_view arg);

}  // namespace hola
