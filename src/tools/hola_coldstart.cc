// hola-coldstart - time from exec() to the first HTTP 200.
//
// BASELINE.json config 3: "standalone uber-jar (target/hola-swarm.jar equivalent) cold start to
// first HTTP 200". Start-up is a headline property of a microservice (pod scale-out, readiness
// probes: /root/reference/pom.xml:91-94 allows the JVM service 5 s before the first probe), and
// for a process that is up in about a millisecond the measuring tool must not be the slow part,
// hence native code: fork+exec the command, then hammer connect()+GET until a 200 arrives.
//
//   hola-coldstart [--port N] [--path /api/hola] [--runs K] [--timeout-ms T] -- command [args...]
//
// Prints one JSON line with the per-run times in microseconds. The child gets SIGTERM after each
// run and is waited for, so runs do not overlap.
#include <arpa/inet.h>
#include <fcntl.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <signal.h>
#include <sys/socket.h>
#include <sys/wait.h>
#include <unistd.h>

#include <algorithm>
#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "net/event_loop.h"  // monotonic_ns

namespace {

// One attempt: connect, send the request, read until the status line is in. Returns the status
// code, 0 when the server is not there yet.
int try_once(uint16_t port, const std::string& request) {
    const int fd = socket(AF_INET, SOCK_STREAM | SOCK_CLOEXEC, 0);
    if (fd < 0) return 0;
    sockaddr_in addr{};
    addr.sin_family = AF_INET;
    addr.sin_port = htons(port);
    addr.sin_addr.s_addr = htonl(INADDR_LOOPBACK);
    int status = 0;
    if (connect(fd, reinterpret_cast<sockaddr*>(&addr), sizeof addr) == 0) {
        const int one = 1;
        setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
        timeval tv{1, 0};
        setsockopt(fd, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof tv);
        if (send(fd, request.data(), request.size(), MSG_NOSIGNAL) == static_cast<ssize_t>(request.size())) {
            char buf[256];
            size_t have = 0;
            while (have < 12) {
                const ssize_t n = recv(fd, buf + have, sizeof buf - have, 0);
                if (n <= 0) break;
                have += static_cast<size_t>(n);
            }
            if (have >= 12 && std::memcmp(buf, "HTTP/1.", 7) == 0) status = std::atoi(buf + 9);
        }
    }
    close(fd);
    return status;
}

}  // namespace

int main(int argc, char** argv) {
    uint16_t port = 8080;
    std::string path = "/api/hola";
    int runs = 5;
    int timeout_ms = 30000;
    int i = 1;
    for (; i < argc; ++i) {
        const std::string a = argv[i];
        if (a == "--") {
            ++i;
            break;
        }
        if (i + 1 >= argc) break;
        if (a == "--port") port = static_cast<uint16_t>(std::atoi(argv[++i]));
        else if (a == "--path") path = argv[++i];
        else if (a == "--runs") runs = std::atoi(argv[++i]);
        else if (a == "--timeout-ms") timeout_ms = std::atoi(argv[++i]);
        else break;
    }
    if (i >= argc || runs < 1) {
        std::fprintf(stderr, "usage: hola-coldstart [--port N] [--path P] [--runs K] [--timeout-ms T] -- command [args...]\n");
        return 2;
    }
    const std::string request = "GET " + path + " HTTP/1.1\r\nHost: localhost\r\nConnection: close\r\n\r\n";
    std::vector<double> times_us;
    for (int r = 0; r < runs; ++r) {
        const int64_t t0 = hola::monotonic_ns();
        const pid_t pid = fork();
        if (pid < 0) {
            std::perror("fork");
            return 1;
        }
        if (pid == 0) {
            // Child: silence it and become the server.
            const int devnull = open("/dev/null", O_WRONLY);
            if (devnull >= 0) {
                dup2(devnull, 1);
                dup2(devnull, 2);
            }
            execvp(argv[i], argv + i);
            _exit(127);
        }
        int status = 0;
        int64_t t1 = 0;
        const int64_t give_up = t0 + static_cast<int64_t>(timeout_ms) * 1000000;
        while (true) {
            status = try_once(port, request);
            t1 = hola::monotonic_ns();
            if (status == 200 || t1 > give_up) break;
            int wstatus = 0;
            if (waitpid(pid, &wstatus, WNOHANG) == pid) {
                std::fprintf(stderr, "hola-coldstart: the command exited before serving (status %d)\n", wstatus);
                return 1;
            }
            if (status == 0) usleep(50);  // connection refused: the listener is not up yet
        }
        kill(pid, SIGTERM);
        int wstatus = 0;
        waitpid(pid, &wstatus, 0);
        if (status != 200) {
            std::fprintf(stderr, "hola-coldstart: no HTTP 200 within %d ms (last status %d)\n", timeout_ms, status);
            return 1;
        }
        times_us.push_back(static_cast<double>(t1 - t0) / 1000.0);
    }
    std::vector<double> sorted = times_us;
    std::sort(sorted.begin(), sorted.end());
    std::string out = "{\"metric\":\"exec to first HTTP 200\",\"unit\":\"us\",\"runs\":[";
    char buf[64];
    for (size_t k = 0; k < times_us.size(); ++k) {
        std::snprintf(buf, sizeof buf, "%s%.1f", k ? "," : "", times_us[k]);
        out += buf;
    }
    std::snprintf(buf, sizeof buf, "],\"min\":%.1f,", sorted.front());
    out += buf;
    std::snprintf(buf, sizeof buf, "\"median\":%.1f,", sorted[sorted.size() / 2]);
    out += buf;
    std::snprintf(buf, sizeof buf, "\"max\":%.1f}", sorted.back());
    out += buf;
    std::puts(out.c_str());
    return 0;
}
