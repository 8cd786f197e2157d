#include "app/service.h"

#include "core/text.h"

#include <arpa/inet.h>
#include <fcntl.h>
#include <sys/eventfd.h>
#include <sys/resource.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <ctime>

#ifndef HOLA_VERSION
#define HOLA_VERSION "0.0.1-SNAPSHOT"
#endif

namespace hola {

std::string usage_text() {
    return "usage: hola [-Dkey=value ...] [options]\n"
           "\n"
           "JVM-style definitions (same meaning as for the Java service):\n"
           "  -Dconf=FILE                 properties file holding the `hello` greeting pattern (hot-reloaded, 5 s)\n"
           "  -Dhello=PATTERN             greeting pattern; beats the environment and the file\n"
           "  -Dswarm.http.port=N         listen port (default 8080)\n"
           "  -Dswarm.bind.address=ADDR   listen address (default 0.0.0.0)\n"
           "  -Dhystrix.command.default.KEY=V, -Dhystrix.threadpool.default.KEY=V   breaker settings\n"
           "\n"
           "Options:\n"
           "  --conf FILE                 same as -Dconf=FILE\n"
           "  --port N, --bind ADDR       same as the swarm.* definitions (port 0: pick a free one; --bind :: listens\n"
           "                              on IPv6 and IPv4)\n"
           "  --threads N                 event-loop threads (default: one per available CPU, capped by the cgroup CPU quota)\n"
           "  --cpus LIST                 pin thread i to the i-th CPU of LIST (e.g. 0,1,2,3 or 0-3)\n"
           "  --accept-mode MODE          balanced (default: a listener per thread, new connections go to the least-loaded\n"
           "                              thread) | reuseport (a listener per thread, the kernel's hash decides) |\n"
           "                              rr (one acceptor deals connections in turn; no SO_REUSEPORT)\n"
           "  --io-backend BACKEND        epoll (readiness: one recv and one send system call per request) | uring (io_uring:\n"
           "                              multishot receives into kernel-picked buffers, sends batched into one system call\n"
           "                              per loop iteration; needs Linux 6.1+, says so and uses epoll where refused) |\n"
           "                              auto (default: uring where available, else epoll, silently)\n"
           "  --aloha-url URL             downstream base URL (default http://aloha:8080/)\n"
           "  --feign-max-attempts N      tries per downstream call on connection failures (default 5, as Feign's\n"
           "                              default Retryer: pauses of 150, 225, 337, 506 ms; 1 disables retrying)\n"
           "  --resolve NAME=ADDRESS      answer lookups of NAME with that IPv4/IPv6 address instead of hosts file and DNS\n"
           "  --zipkin-url URL            overrides env ZIPKIN_SERVER_URL; unset/empty disables tracing\n"
           "  --zipkin-flush-ms N         span reporter period (default 1000)\n"
           "  --zipkin-ring N             spans queued per thread before new ones are dropped (default 1024)\n"
           "  --zipkin-max-batch N        spans per POST to the collector (default 1000, the size of the reference's queue)\n"
           "  --hystrix-config FILE       properties file with hystrix.* keys (default: the reference's config.properties values)\n"
           "  --deltaspike-properties FILE  extra config source at ordinal 100 (META-INF/apache-deltaspike.properties)\n"
           "  --webjars-dir DIR           directory served under /webjars/ (Swagger-UI assets)\n"
           "  --idle-timeout MS           close keep-alive connections idle for MS (default: never)\n"
           "  --drain-linger-ms N         on SIGTERM keep answering on open connections for N ms (each answer says\n"
           "                              \"Connection: close\") before closing the idle ones (default 100)\n"
           "  --spin-us N                 keep polling for N us after the last event before sleeping (default 0;\n"
           "                              for pinned, dedicated cores: trades idle CPU for lower latency)\n"
           "  --max-header-bytes N        request head limit (default 65536)\n"
           "  --log-config-reads          print the conf path on every config-file read, as the reference does\n"
           "  --log-feign                 log every downstream call to stderr in Feign's BASIC format, as the reference does\n"
           "  --reference-logging         both of the above: the log output of the Java service, for log scrapers\n"
           "  --serve-aloha               also answer GET /api/aloha-chaining with [\"Aloha mai <HOSTNAME>\"]: a stand-in\n"
           "                              for the next service of the demo chain (point another hola at it)\n"
           "  --print-greeting            print what GET /api/hola would answer with this configuration and exit\n"
           "                              (exit 1 if the pattern is one String.format would reject)\n"
           "  --verbose                   log downstream calls and request failures to stderr\n"
           "  --quiet                     no start-up/shutdown log lines\n"
           "  --version, --help\n";
}

namespace {

bool parse_uint(const std::string& s, long max, long* out) {
    if (s.empty() || s.size() > 10) return false;
    long v = 0;
    for (char c : s) {
        if (c < '0' || c > '9') return false;
        v = v * 10 + (c - '0');
    }
    if (v > max) return false;
    *out = v;
    return true;
}

}  // namespace

ParseOutcome parse_service_args(const std::vector<std::string>& args, ServiceConfig* cfg, std::string* message) {
    std::string hystrix_config;
    auto fail = [&](const std::string& m) {
        *message = "hola: " + m + "\n";
        return ParseOutcome::kExitUsage;
    };
    auto set_prop = [&](const std::string& k, const std::string& v) { cfg->system_properties.emplace_back(k, v); };
    for (size_t i = 0; i < args.size(); ++i) {
        const std::string& a = args[i];
        auto value = [&](std::string* out) {
            if (i + 1 >= args.size()) return false;
            *out = args[++i];
            return true;
        };
        std::string v;
        long n = 0;
        if (a.size() > 2 && a[0] == '-' && a[1] == 'D') {
            const size_t eq = a.find('=');
            if (eq == std::string::npos) {
                set_prop(a.substr(2), "");
            } else {
                set_prop(a.substr(2, eq - 2), a.substr(eq + 1));
            }
        } else if (a == "--conf") {
            if (!value(&v)) return fail("--conf needs a value");
            set_prop("conf", v);
        } else if (a == "--port") {
            if (!value(&v)) return fail("--port needs a value");
            set_prop("swarm.http.port", v);
        } else if (a == "--bind") {
            if (!value(&v)) return fail("--bind needs a value");
            set_prop("swarm.bind.address", v);
        } else if (a == "--threads") {
            if (!value(&v) || !parse_uint(v, 4096, &n)) return fail("--threads needs a number");
            cfg->server.threads = static_cast<int>(n);
        } else if (a == "--cpus") {
            if (!value(&v) || !parse_cpu_list(v, &cfg->server.cpu_affinity)) return fail("bad --cpus list");
        } else if (a == "--accept-mode") {
            if (!value(&v) || (v != "balanced" && v != "reuseport" && v != "rr")) return fail("--accept-mode is balanced, reuseport or rr");
            cfg->server.accept_mode = v == "balanced" ? AcceptMode::kBalanced : v == "reuseport" ? AcceptMode::kReusePort : AcceptMode::kRoundRobin;
        } else if (a == "--io-backend") {
            if (!value(&v) || (v != "auto" && v != "epoll" && v != "uring")) return fail("--io-backend is auto, epoll or uring");
            cfg->server.io_backend = v == "uring" ? IoBackend::kUring : v == "epoll" ? IoBackend::kEpoll : IoBackend::kAuto;
        } else if (a == "--aloha-url") {
            std::string host, path;
            uint16_t port = 80;
            if (!value(&v) || !parse_http_url(v, &host, &port, &path)) return fail("--aloha-url must look like http://host:port/");
            cfg->app.aloha_host = host;
            cfg->app.aloha_port = port;
            cfg->app.aloha_path = path + "/api/aloha-chaining";
        } else if (a == "--feign-max-attempts") {
            if (!value(&v) || !parse_uint(v, 100, &n) || n == 0) return fail("--feign-max-attempts needs a number >= 1");
            cfg->app.feign_max_attempts = static_cast<int>(n);
        } else if (a == "--resolve") {
            if (!value(&v)) return fail("--resolve needs NAME=ADDRESS");
            const size_t eq = v.find('=');
            if (eq == std::string::npos || eq == 0 || IpAddress::parse(v.substr(eq + 1)).family == 0)
                return fail("--resolve needs NAME=ADDRESS (an IPv4 or IPv6 address)");
            cfg->static_hosts.emplace_back(v.substr(0, eq), v.substr(eq + 1));
        } else if (a == "--zipkin-url") {
            if (!value(&v)) return fail("--zipkin-url needs a value");
            cfg->zipkin_url = v;
            cfg->zipkin_url_given = true;
        } else if (a == "--zipkin-flush-ms") {
            if (!value(&v) || !parse_uint(v, 3600000, &n) || n == 0) return fail("--zipkin-flush-ms needs a positive number");
            cfg->zipkin_flush_ms = static_cast<int>(n);
        } else if (a == "--zipkin-ring") {
            if (!value(&v) || !parse_uint(v, 1 << 22, &n) || n < 2) return fail("--zipkin-ring needs a number of spans >= 2");
            cfg->zipkin_ring = static_cast<int>(n);
        } else if (a == "--zipkin-max-batch") {
            if (!value(&v) || !parse_uint(v, 1 << 22, &n) || n == 0) return fail("--zipkin-max-batch needs a positive number");
            cfg->zipkin_max_batch = static_cast<int>(n);
        } else if (a == "--hystrix-config") {
            if (!value(&hystrix_config)) return fail("--hystrix-config needs a value");
        } else if (a == "--deltaspike-properties") {
            if (!value(&cfg->deltaspike_properties)) return fail("--deltaspike-properties needs a value");
        } else if (a == "--webjars-dir") {
            if (!value(&v)) return fail("--webjars-dir needs a value");
            while (v.size() > 1 && v.back() == '/') v.pop_back();
            cfg->app.webjars_dir = v;
        } else if (a == "--idle-timeout") {
            if (!value(&v) || !parse_uint(v, 86400000, &n)) return fail("--idle-timeout needs milliseconds");
            cfg->server.idle_timeout_ms = static_cast<int>(n);
        } else if (a == "--drain-linger-ms") {
            if (!value(&v) || !parse_uint(v, 600000, &n)) return fail("--drain-linger-ms needs milliseconds");
            cfg->server.drain_linger_ms = static_cast<int>(n);
        } else if (a == "--spin-us") {
            if (!value(&v) || !parse_uint(v, 1000000, &n)) return fail("--spin-us needs microseconds");
            cfg->server.spin_us = static_cast<int>(n);
        } else if (a == "--max-header-bytes") {
            if (!value(&v) || !parse_uint(v, 1 << 30, &n) || n < 256) return fail("--max-header-bytes needs a number >= 256");
            cfg->server.max_header_bytes = static_cast<size_t>(n);
        } else if (a == "--log-config-reads") {
            cfg->log_config_reads = true;
        } else if (a == "--verbose") {
            cfg->app.verbose = true;
            cfg->app.log_feign = true;
        } else if (a == "--log-feign") {
            cfg->app.log_feign = true;
        } else if (a == "--reference-logging") {
            cfg->app.log_feign = true;
            cfg->log_config_reads = true;
        } else if (a == "--quiet") {
            cfg->quiet = true;
        } else if (a == "--serve-aloha") {
            cfg->app.serve_aloha = true;
        } else if (a == "--print-greeting") {
            cfg->print_greeting = true;
        } else if (a == "--version") {
            *message = std::string("hola ") + HOLA_VERSION + "\n";
            return ParseOutcome::kExitOk;
        } else if (a == "--help" || a == "-h") {
            *message = usage_text();
            return ParseOutcome::kExitOk;
        } else {
            *message = "hola: unknown argument: " + a + "\n" + usage_text();
            return ParseOutcome::kExitUsage;
        }
    }

    // Listener settings from the swarm.* definitions (later definitions win, as on the JVM).
    for (const auto& kv : cfg->system_properties) {
        long n = 0;
        if (kv.first == "swarm.http.port") {
            if (!parse_uint(kv.second, 65535, &n)) return fail("bad port: " + kv.second);
            cfg->server.port = static_cast<uint16_t>(n);
        } else if (kv.first == "swarm.bind.address") {
            cfg->server.bind_address = kv.second;
        }
    }

    // Breaker settings: the reference's config.properties values, then an optional file, then
    // -D definitions (Archaius gives system properties the last word).
    if (!hystrix_config.empty()) {
        Properties props;
        std::string err;
        if (!props.load_file(hystrix_config, &err)) return fail("--hystrix-config: " + err);
        cfg->app.breaker.apply(props);
    }
    for (const auto& kv : cfg->system_properties) {
        if (kv.first.compare(0, 8, "hystrix.") == 0) cfg->app.breaker.apply_key(kv.first, kv.second);
    }
    return ParseOutcome::kRun;
}

std::unique_ptr<ConfigResolver> build_config_resolver(const ServiceConfig& config) {
    // Configuration layers, highest ordinal first (core/config.h).
    std::unique_ptr<ConfigResolver> resolver(new ConfigResolver());
    auto* system_props = new SystemPropertySource();
    for (const auto& kv : config.system_properties) system_props->set(kv.first, kv.second);
    auto* conf_source = new ConfFileSource(system_props);
    if (config.log_config_reads) {
        conf_source->set_on_read([](const std::string* path) {
            std::printf("%s\n", path ? path->c_str() : "null");  // Configuration.java:28
            std::fflush(stdout);
        });
    }
    resolver->add_source(std::unique_ptr<ConfigSource>(system_props));
    resolver->add_source(std::unique_ptr<ConfigSource>(new EnvironmentSource()));
    if (!config.deltaspike_properties.empty()) {
        resolver->add_source(std::unique_ptr<ConfigSource>(
            new StaticPropertiesSource("META-INF/apache-deltaspike.properties", config.deltaspike_properties)));
    }
    resolver->add_source(std::unique_ptr<ConfigSource>(conf_source));
    return resolver;
}

bool resolve_greeting_once(const ServiceConfig& config, std::string* body, std::string* error) {
    std::unique_ptr<ConfigResolver> resolver = build_config_resolver(config);
    const char* h = std::getenv("HOSTNAME");
    GreetingService greeting(resolver.get(), h ? h : "unknown", nullptr);
    greeting.poll(monotonic_ns());
    const std::shared_ptr<const GreetingSnapshot> snap = greeting.snapshot();
    if (!snap->ok) {
        *error = "pattern \"" + snap->pattern + "\" cannot be formatted: " + snap->error;
        return false;
    }
    *body = snap->text;
    return true;
}

Service::Service(ServiceConfig config) : config_(std::move(config)) {}

Service::~Service() { stop(); }

void Service::log(const char* level, const std::string& message) const {
    if (config_.quiet && level[0] == 'I') return;
    char stamp[32];
    const time_t t = time(nullptr);
    tm tmv;
    localtime_r(&t, &tmv);
    strftime(stamp, sizeof stamp, "%Y-%m-%d %H:%M:%S", &tmv);
    std::fprintf(stderr, "%s %s [hola] %s\n", stamp, level, message.c_str());
}

bool Service::start(std::string* error) {
    if (started_) return true;
    const int64_t t_start = monotonic_ns();

    // Every connection is a descriptor: use what the system allows rather than the usual soft
    // default of 1024.
    rlimit nofile{};
    if (getrlimit(RLIMIT_NOFILE, &nofile) == 0 && nofile.rlim_cur < nofile.rlim_max) {
        nofile.rlim_cur = nofile.rlim_max;
        setrlimit(RLIMIT_NOFILE, &nofile);
    }

    // Grow the descriptor table now, while the process is still single-threaded. The kernel
    // starts a process with room for 64 descriptors and doubles the table on demand; once other
    // threads share it, every doubling waits for an RCU grace period - 20-40 ms on the benchmark
    // boxes - in whichever thread happens to open the 65th, 129th, ... descriptor. With one epoll
    // set, one eventfd, one listener and one ring per worker that was the start-up path on any
    // machine with more than a dozen CPUs, and later the accept path (measured:
    // profiles/r2_coldstart/). One dup to a high number sizes the table for 16 K descriptors
    // (128 KB); beyond that the doubling happens again, once per doubling.
    {
        const int probe = eventfd(0, EFD_CLOEXEC);
        if (probe >= 0) {
            rlimit lim{};
            long want = 16384;
            if (getrlimit(RLIMIT_NOFILE, &lim) == 0 && lim.rlim_cur != RLIM_INFINITY && static_cast<long>(lim.rlim_cur) <= want)
                want = static_cast<long>(lim.rlim_cur) - 1;
            if (want > probe) {
                const int high = fcntl(probe, F_DUPFD_CLOEXEC, static_cast<int>(want));
                if (high >= 0) close(high);
            }
            close(probe);
        }
    }

    if (const char* h = std::getenv("HOSTNAME")) config_.app.hostname = h;
    std::string zipkin_url = config_.zipkin_url;
    if (!config_.zipkin_url_given) {
        if (const char* z = std::getenv("ZIPKIN_SERVER_URL")) zipkin_url = z;
    }

    resolver_ = build_config_resolver(config_);

    greeting_.reset(new GreetingService(resolver_.get(), config_.app.hostname,
                                        [this](const std::string& m) { log("INFO", m); }));
    tracer_.reset(new Tracer(zipkin_url, "hola", config_.zipkin_flush_ms, static_cast<size_t>(config_.zipkin_ring),
                             static_cast<size_t>(config_.zipkin_max_batch)));
    if (!tracer_->error().empty()) log("WARN", tracer_->error() + " - tracing disabled");
    dns_.reset(new Resolver());
    for (const auto& kv : config_.static_hosts) dns_->set_static(kv.first, IpAddress::parse(kv.second));
    app_.reset(new HolaApp(config_.app, greeting_.get(), tracer_.get(), dns_.get()));
    server_.reset(new Server(config_.server, app_.get()));
    if (!server_->start(error)) {
        server_.reset();
        return false;
    }
    tracer_->start();
    started_ = true;
    char msg[256];
    std::snprintf(msg, sizeof msg, "hola %s listening on %s:%u with %d thread(s), started in %.1f ms, %s%s", HOLA_VERSION,
                  config_.server.bind_address.c_str(), server_->port(), server_->threads(),
                  static_cast<double>(monotonic_ns() - t_start) / 1e6, server_->io_backend_name(), tracer_->enabled() ? ", tracing on" : "");
    log("INFO", msg);
    return true;
}

std::string Service::stats_json() const {
    if (!server_) return "{}";
    rusage ru{};
    getrusage(RUSAGE_SELF, &ru);
    const ServerStats st = server_->stats();
    const TracerStats ts = tracer_->stats();
    const char* mode = config_.server.accept_mode == AcceptMode::kBalanced    ? "balanced"
                       : config_.server.accept_mode == AcceptMode::kReusePort ? "reuseport"
                                                                              : "rr";
    char buf[512];
    std::snprintf(buf, sizeof buf,
                  "{\"requests\":%llu,\"bad_requests\":%llu,\"connections_accepted\":%llu,\"connections_open\":%llu,"
                  "\"cpu_user_s\":%.6f,\"cpu_sys_s\":%.6f,\"accept_mode\":\"%s\",\"io_backend\":\"%s\",\"threads\":%d,"
                  "\"spans\":{\"reported\":%llu,\"dropped\":%llu,\"posts_ok\":%llu,\"posts_failed\":%llu},\"workers\":[",
                  static_cast<unsigned long long>(st.requests), static_cast<unsigned long long>(st.bad_requests),
                  static_cast<unsigned long long>(st.connections_accepted), static_cast<unsigned long long>(st.connections_open),
                  static_cast<double>(ru.ru_utime.tv_sec) + static_cast<double>(ru.ru_utime.tv_usec) / 1e6,
                  static_cast<double>(ru.ru_stime.tv_sec) + static_cast<double>(ru.ru_stime.tv_usec) / 1e6, mode, server_->io_backend_name(),
                  server_->threads(),
                  static_cast<unsigned long long>(ts.spans_reported), static_cast<unsigned long long>(ts.spans_dropped),
                  static_cast<unsigned long long>(ts.posts_ok), static_cast<unsigned long long>(ts.posts_failed));
    std::string out = buf;
    bool first = true;
    for (const WorkerSnapshot& w : server_->worker_stats()) {
        std::snprintf(buf, sizeof buf, "%s{\"id\":%d,\"requests\":%llu,\"connections\":%llu,\"connections_open\":%llu,\"cpu_s\":%.6f}",
                      first ? "" : ",", w.id, static_cast<unsigned long long>(w.requests),
                      static_cast<unsigned long long>(w.connections_accepted), static_cast<unsigned long long>(w.connections_open),
                      w.cpu_seconds);
        out.append(buf);
        first = false;
    }
    out.append("]}");
    return out;
}

void Service::stop() {
    if (!started_) return;
    started_ = false;
    server_->stop();
    dns_->shutdown();
    tracer_->stop();
    if (config_.app.verbose) {
        // A cheap profile for the record: how the process spent its CPU, how well the event
        // loops batched their work, and how evenly the workers shared it.
        rusage ru{};
        getrusage(RUSAGE_SELF, &ru);
        const ServerStats st = server_->stats();
        const TracerStats ts = tracer_->stats();
        const double user = static_cast<double>(ru.ru_utime.tv_sec) + static_cast<double>(ru.ru_utime.tv_usec) / 1e6;
        const double sys = static_cast<double>(ru.ru_stime.tv_sec) + static_cast<double>(ru.ru_stime.tv_usec) / 1e6;
        char msg[512];
        std::snprintf(msg, sizeof msg,
                      "stats: requests=%llu connections=%llu cpu_user_s=%.3f cpu_sys_s=%.3f us_per_request(user/sys)=%.2f/%.2f "
                      "loop_iterations=%llu events=%llu blocking_waits=%llu spans(reported/dropped)=%llu/%llu ctx_switches(vol/invol)=%ld/%ld",
                      static_cast<unsigned long long>(st.requests), static_cast<unsigned long long>(st.connections_accepted), user, sys,
                      st.requests ? user * 1e6 / static_cast<double>(st.requests) : 0.0,
                      st.requests ? sys * 1e6 / static_cast<double>(st.requests) : 0.0,
                      static_cast<unsigned long long>(st.loop_iterations), static_cast<unsigned long long>(st.loop_events),
                      static_cast<unsigned long long>(st.loop_blocking_waits), static_cast<unsigned long long>(ts.spans_reported),
                      static_cast<unsigned long long>(ts.spans_dropped), ru.ru_nvcsw, ru.ru_nivcsw);
        log("INFO", msg);
        std::string per_worker = "workers (requests/connections/cpu_s):";
        for (const WorkerSnapshot& w : server_->worker_stats()) {
            char item[96];
            std::snprintf(item, sizeof item, " %d=%llu/%llu/%.2f", w.id, static_cast<unsigned long long>(w.requests),
                          static_cast<unsigned long long>(w.connections_accepted), w.cpu_seconds);
            per_worker.append(item);
        }
        log("INFO", per_worker);
    }
}

}  // namespace hola
