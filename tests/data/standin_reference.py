"""A stand-in for the Java reference, used only to prove that `bench.py --impl reference` drives
*some other program* through the same procedure and prints the same schema. It is deliberately
not this repository's server: a few lines of Python's http.server answering GET /api/hola the way
the reference does without -Dconf and without HOSTNAME exported (HolaResource.java:57-65)."""
import sys
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer


class Handler(BaseHTTPRequestHandler):
    protocol_version = "HTTP/1.1"

    def log_message(self, *a):
        pass

    def do_GET(self):
        body = b"Hola de unknown" if self.path == "/api/hola" else b"I'm ok"
        self.send_response(200 if self.path in ("/api/hola", "/api/health") else 404)
        self.send_header("Content-Type", "text/plain")
        self.send_header("Access-Control-Allow-Origin", "*")
        self.send_header("Content-Length", str(len(body)))
        self.end_headers()
        self.wfile.write(body)


class Server(ThreadingHTTPServer):
    request_queue_size = 128
    daemon_threads = True


if __name__ == "__main__":
    Server(("127.0.0.1", int(sys.argv[1])), Handler).serve_forever()
