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

// Stand-in for the Swagger-UI page the landing page frames. The real Swagger-UI 2.1.4 bundle
// comes from a webjar (/root/reference/pom.xml:207-211) that is not available offline; this
// single self-contained page gives the same capability - read the API description named by
// ?url=..., list its operations, invoke them from the browser - without any third-party asset.
std::string_view api_browser_html() {
    static const char kPage[] = R"HTML(<!DOCTYPE html>
<html>
<head>
<meta charset="UTF-8">
<title>API browser</title>
<style>
body { font: 14px/1.4 sans-serif; margin: 1.5em; color: #222; }
h1 { font-size: 1.4em; margin-bottom: 0; }
.desc { color: #555; margin-top: .2em; }
.op { border: 1px solid #c3d9ec; border-radius: 4px; margin: .8em 0; background: #ebf3f9; }
.op .head { padding: .5em .7em; cursor: pointer; }
.method { display: inline-block; min-width: 3.5em; text-align: center; background: #0f6ab4; color: #fff; border-radius: 3px; padding: .1em .4em; font-weight: bold; }
.path { font-family: monospace; font-size: 1.1em; margin: 0 .7em; }
.summary { color: #555; }
.body { display: none; padding: .6em .9em; border-top: 1px solid #c3d9ec; background: #fff; }
.body.open { display: block; }
pre { background: #f6f6f6; border: 1px solid #ddd; padding: .5em; overflow: auto; max-height: 20em; }
button { padding: .3em .9em; }
.note { color: #777; font-size: .9em; margin-top: 2em; }
</style>
</head>
<body>
<h1 id="title">API</h1>
<div class="desc" id="desc"></div>
<div id="meta" class="desc"></div>
<div id="ops"></div>
<div class="note">Built-in API browser of the hola service. To use the original Swagger-UI instead, unpack its
distribution into a directory as <code>swagger-ui/2.1.4/</code> and start the service with <code>--webjars-dir</code>.</div>
<script type="text/javascript">
(function () {
  function text(s) { return document.createTextNode(s); }
  function el(tag, cls, content) {
    var e = document.createElement(tag);
    if (cls) e.className = cls;
    if (content !== undefined) e.appendChild(text(content));
    return e;
  }
  var m = /[?&]url=([^&#]*)/.exec(window.location.search);
  var specUrl = m ? decodeURIComponent(m[1]) : '/api/swagger.json';
  fetch(specUrl).then(function (r) { return r.json(); }).then(function (spec) {
    var info = spec.info || {};
    document.getElementById('title').textContent = info.title || 'API';
    document.getElementById('desc').textContent = info.description || '';
    var scheme = (spec.schemes && spec.schemes[0]) || window.location.protocol.replace(':', '');
    var host = spec.host || window.location.host;
    var base = scheme + '://' + host + (spec.basePath || '');
    document.getElementById('meta').textContent = 'version ' + (info.version || '?') + ' - base URL ' + base +
        (info.license ? ' - ' + info.license.name : '');
    var ops = document.getElementById('ops');
    Object.keys(spec.paths || {}).forEach(function (path) {
      Object.keys(spec.paths[path]).forEach(function (method) {
        var op = spec.paths[path][method];
        var box = el('div', 'op');
        var head = el('div', 'head');
        head.appendChild(el('span', 'method', method.toUpperCase()));
        head.appendChild(el('span', 'path', path));
        head.appendChild(el('span', 'summary', op.summary || ''));
        var body = el('div', 'body');
        body.appendChild(el('div', '', 'Produces: ' + (op.produces || []).join(', ')));
        var go = el('button', '', 'Try it out');
        var out = el('pre', '', '');
        out.style.display = 'none';
        go.onclick = function () {
          var t0 = Date.now();
          fetch(base + path, { method: method.toUpperCase(), headers: { 'Accept': (op.produces || ['*/*']).join(', ') } })
            .then(function (r) { return r.text().then(function (t) { return [r, t]; }); })
            .then(function (rt) {
              var r = rt[0], lines = ['HTTP ' + r.status + ' ' + r.statusText + '   (' + (Date.now() - t0) + ' ms)'];
              r.headers.forEach(function (v, k) { lines.push(k + ': ' + v); });
              out.textContent = lines.join('\n') + '\n\n' + rt[1];
              out.style.display = 'block';
            })
            .catch(function (e) { out.textContent = 'request failed: ' + e; out.style.display = 'block'; });
        };
        body.appendChild(go);
        body.appendChild(out);
        head.onclick = function () { body.className = body.className === 'body' ? 'body open' : 'body'; };
        box.appendChild(head);
        box.appendChild(body);
        ops.appendChild(box);
      });
    });
  }).catch(function (e) {
    document.getElementById('desc').textContent = 'Could not load ' + specUrl + ': ' + e;
  });
})();
</script>
</body>
</html>
)HTML";
    return std::string_view(kPage, sizeof kPage - 1);
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
