#include "app/static_content.h"

#include "core/text.h"

namespace hola {

// Byte-for-byte the page the reference serves at "/": tabs for indentation, LF line ends, no
// newline after </html>. Its bytes are part of the observable contract.
std::string_view index_html() {
    static const char kPage[] =
        "<!DOCTYPE html>\n"
        "<html>\n"
        "<head>\n"
        "<meta charset=\"UTF-8\">\n"
        "<title>Microservices endpoint</title>\n"
        "</head>\n"
        "<body style=\"margin: 0px; padding: 0px; overflow: hidden\">\n"
        "\t<div>\n"
        "\t\t<iframe id=\"iframe\" style=\"position: absolute;\" height=\"100%\" width=\"100%\"></iframe>\n"
        "\t</div>\n"
        "\t<script type=\"text/javascript\">\n"
        "\t\tvar loc = window.location.href;\n"
        "\t\tvar dest = 'webjars/swagger-ui/2.1.4/?url=' + loc + 'api/swagger.json#!/default';\n"
        "\t\tdocument.getElementById('iframe').src = dest;\n"
        "\t</script>\n"
        "</body>\n"
        "</html>";
    return std::string_view(kPage, sizeof kPage - 1);
}

std::string_view swagger_json() {
    static const char kDoc[] = R"JSON({
  "swagger" : "2.0",
  "info" : {
    "description" : "Operations that can be invoked in the hola microservices",
    "version" : "1.0.0",
    "title" : "Hola microservices REST API",
    "contact" : {
      "name" : "developer@redhat.com"
    },
    "license" : {
      "name" : "Apache 2.0",
      "url" : "http://www.apache.org/licenses/LICENSE-2.0.html"
    }
  },
  "basePath" : "/api",
  "schemes" : [ "http" ],
  "paths" : {
    "/health" : {
      "get" : {
        "summary" : "Used to verify the health of the service",
        "description" : "",
        "operationId" : "health",
        "produces" : [ "text/plain" ],
        "parameters" : [ ],
        "responses" : {
          "200" : {
            "description" : "successful operation",
            "schema" : {
              "type" : "string"
            }
          }
        }
      }
    },
    "/hola" : {
      "get" : {
        "summary" : "Returns the greeting in Spanish",
        "description" : "",
        "operationId" : "hola",
        "produces" : [ "text/plain" ],
        "parameters" : [ ],
        "responses" : {
          "200" : {
            "description" : "successful operation",
            "schema" : {
              "type" : "string"
            }
          }
        }
      }
    },
    "/hola-chaining" : {
      "get" : {
        "summary" : "Returns the greeting plus the next service in the chain",
        "description" : "",
        "operationId" : "holaChaining",
        "produces" : [ "application/json" ],
        "parameters" : [ ],
        "responses" : {
          "200" : {
            "description" : "successful operation",
            "schema" : {
              "type" : "array",
              "items" : {
                "type" : "string"
              }
            }
          }
        }
      }
    }
  }
})JSON";
    return std::string_view(kDoc, sizeof kDoc - 1);
}

std::string_view swagger_yaml() {
    static const char kDoc[] = R"YAML(---
swagger: "2.0"
info:
  description: "Operations that can be invoked in the hola microservices"
  version: "1.0.0"
  title: "Hola microservices REST API"
  contact:
    name: "developer@redhat.com"
  license:
    name: "Apache 2.0"
    url: "http://www.apache.org/licenses/LICENSE-2.0.html"
basePath: "/api"
schemes:
- "http"
paths:
  /health:
    get:
      summary: "Used to verify the health of the service"
      description: ""
      operationId: "health"
      produces:
      - "text/plain"
      parameters: []
      responses:
        200:
          description: "successful operation"
          schema:
            type: "string"
  /hola:
    get:
      summary: "Returns the greeting in Spanish"
      description: ""
      operationId: "hola"
      produces:
      - "text/plain"
      parameters: []
      responses:
        200:
          description: "successful operation"
          schema:
            type: "string"
  /hola-chaining:
    get:
      summary: "Returns the greeting plus the next service in the chain"
      description: ""
      operationId: "holaChaining"
      produces:
      - "application/json"
      parameters: []
      responses:
        200:
          description: "successful operation"
          schema:
            type: "array"
            items:
              type: "string"
)YAML";
    return std::string_view(kDoc, sizeof kDoc - 1);
}

std::string_view mime_type_for(std::string_view path) {
    const size_t dot = path.rfind('.');
    const std::string_view ext = dot == std::string_view::npos ? std::string_view() : path.substr(dot + 1);
    struct Entry {
        const char* ext;
        const char* type;
    };
    static const Entry kTable[] = {
        {"html", "text/html"},        {"htm", "text/html"},         {"css", "text/css"},
        {"js", "application/javascript"}, {"json", "application/json"}, {"png", "image/png"},
        {"gif", "image/gif"},         {"jpg", "image/jpeg"},        {"jpeg", "image/jpeg"},
        {"svg", "image/svg+xml"},     {"ico", "image/x-icon"},      {"txt", "text/plain"},
        {"map", "application/json"},  {"woff", "application/font-woff"}, {"woff2", "font/woff2"},
        {"ttf", "application/x-font-ttf"}, {"eot", "application/vnd.ms-fontobject"}, {"yaml", "application/yaml"},
        {"xml", "application/xml"},
    };
    for (const Entry& e : kTable) {
        if (ascii_iequals(ext, e.ext)) return e.type;
    }
    return "application/octet-stream";
}

}  // namespace hola
