// The whole service as one object: command-line parsing, configuration layers, greeting cache,
// tracer, resolver, application and server, wired together. `hola` (app/main.cc) is a thin
// wrapper around it, and libhola.so exposes it through the C ABI (include/hola.h) so that tests
// and host programs can run the service in-process.
//
// The accepted arguments are those of the executable and mirror how the reference is launched:
// `java -jar hola-swarm.jar [-Dconf=…] [-Dhello=…] [-Dswarm.http.port=…]`
// (/root/reference/Readme.md:13-16, /root/reference/src/main/java/com/redhat/developers/msa/hola/Configuration.java:27).
#pragma once

#include <memory>
#include <string>
#include <vector>

#include "app/greeting.h"
#include "app/hola_app.h"
#include "app/tracer.h"
#include "core/config.h"
#include "net/resolver.h"
#include "net/server.h"

namespace hola {

struct ServiceConfig {
    ServerOptions server;
    AppOptions app;
    std::vector<std::pair<std::string, std::string>> system_properties;  // -Dkey=value, in order
    std::string zipkin_url;
    bool zipkin_url_given = false;  // false: fall back to env ZIPKIN_SERVER_URL
    std::string deltaspike_properties;
    bool log_config_reads = false;
    bool quiet = false;
    bool print_greeting = false;  // resolve and format the greeting once, print it, exit (see resolve_greeting_once)
    // Test hooks.
    std::vector<std::pair<std::string, std::string>> static_hosts;  // --resolve name=address (IPv4 or IPv6)
    int zipkin_flush_ms = 1000;
    int zipkin_ring = 1024;        // spans queued per worker thread before new ones are dropped
    int zipkin_max_batch = 1000;   // spans per POST (the reference's queue, and so its batch, holds 1000)
};

enum class ParseOutcome { kRun, kExitOk, kExitUsage };

// Parses `args` (without argv[0]). Diagnostics and --help/--version output go to *message.
ParseOutcome parse_service_args(const std::vector<std::string>& args, ServiceConfig* config, std::string* message);

std::string usage_text();

// The service's configuration stack (system properties > environment > optional static file >
// the hot-reloaded `conf` file) for a parsed command line.
std::unique_ptr<ConfigResolver> build_config_resolver(const ServiceConfig& config);

// `hola --print-greeting`: what GET /api/hola would answer right now with this command line and
// environment, without starting a server. False (with the JVM-style reason in *error) when the
// configured pattern is one that would turn every request into a 500 - worth checking before a
// properties file goes to production.
bool resolve_greeting_once(const ServiceConfig& config, std::string* body, std::string* error);

class Service {
 public:
    explicit Service(ServiceConfig config);
    ~Service();

    bool start(std::string* error);
    void stop();

    uint16_t port() const { return server_ ? server_->port() : 0; }
    int threads() const { return server_ ? server_->threads() : 0; }
    bool tracing() const { return tracer_ && tracer_->enabled(); }
    ServerStats server_stats() const { return server_->stats(); }
    TracerStats tracer_stats() const { return tracer_->stats(); }
    // Counters as one JSON object: totals, process CPU time, span counts and one entry per worker
    // (requests, connections, CPU seconds). Printed on SIGUSR1 by the executable (`hola-stats {...}`
    // on stderr) and available through the C ABI; safe to call while the service runs.
    std::string stats_json() const;
    HolaApp& app() { return *app_; }
    const ServiceConfig& config() const { return config_; }

 private:
    void log(const char* level, const std::string& message) const;

    ServiceConfig config_;
    std::unique_ptr<ConfigResolver> resolver_;
    std::unique_ptr<GreetingService> greeting_;
    std::unique_ptr<Tracer> tracer_;
    std::unique_ptr<Resolver> dns_;
    std::unique_ptr<HolaApp> app_;
    std::unique_ptr<Server> server_;
    bool started_ = false;
};

}  // namespace hola
