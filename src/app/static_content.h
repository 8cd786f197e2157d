// Fixed documents served by the application: the landing page and the API description.
//
//  * GET /  -> the page at /root/reference/src/main/webapp/index.html (485 bytes, no trailing
//    newline): a full-window iframe pointed at the bundled Swagger-UI with this service's
//    swagger.json pre-selected.
//  * GET /api/swagger.json, /api/swagger.yaml -> what swagger-jaxrs 1.5.8 generates from the
//    BeanConfig in /root/reference/src/main/java/com/redhat/developers/msa/hola/SwaggerListener.java:29-41
//    and the three @ApiOperation annotations in HolaResource.java:55,72,83 (pretty-printed,
//    because setPrettyPrint(true)). The text is reconstructed from the library's known output
//    format; it could not be diffed against a live reference (no JVM on the build box).
#pragma once

#include <string_view>

namespace hola {

std::string_view index_html();
std::string_view swagger_json();
std::string_view swagger_yaml();
// Self-contained page served where Swagger-UI's index would be when its assets are not installed.
std::string_view api_browser_html();

// Content type for a static file name, by extension (used for the optional webjars directory).
std::string_view mime_type_for(std::string_view path);

}  // namespace hola
