// The hola service itself: routing and the handlers behind every URL the reference answers.
//
//   GET /api/hola            HolaResource.hola()          HolaResource.java:52-67
//   GET /api/hola-chaining   HolaResource.holaChaining()  HolaResource.java:69-78, 94-113
//   GET /api/health          HolaResource.health()        HolaResource.java:80-86
//   GET /api/swagger.{json,yaml}                          SwaggerListener.java:28-42
//   GET /hystrix.stream                                   HystrixServlet.java:12-17
//   GET /  and /index.html                                src/main/webapp/index.html
//   GET /webjars/**          (Swagger-UI webjar, pom.xml:207-211; served from --webjars-dir)
// plus the cross-cutting pieces: "/api" mount point (JaxRsActivator.java:30), the CORS header on
// every JAX-RS response (CorsFilter.java:26-32) and Zipkin tracing on exactly /api/hola and
// /api/hola-chaining (ZipKinFilter.java:38-44).
// (All paths relative to /root/reference/src/main/java/com/redhat/developers/msa/hola/ unless
// they start with src/ or pom.xml.)
#pragma once

#include <atomic>
#include <memory>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "app/greeting.h"
#include "app/tracer.h"
#include "core/breaker.h"
#include "core/json.h"
#include "net/http_client.h"
#include "net/resolver.h"
#include "net/server.h"

namespace hola {

struct AppOptions {
    std::string hostname = "unknown";             // env HOSTNAME (HolaResource.java:57)
    std::string aloha_host = "aloha";             // HolaResource.java:95
    uint16_t aloha_port = 8080;                   // HolaResource.java:103
    std::string aloha_path = "/api/aloha-chaining";  // AlohaService.java:25
    std::string fallback = "Aloha response (fallback)";  // HolaResource.java:112
    std::string webjars_dir;                      // empty: /webjars/** answers 404
    BreakerConfig breaker = BreakerConfig::reference_defaults();
    // Feign's default Retryer (period 100 ms, max period 1 s, 5 attempts); 1 attempt = no retry.
    int feign_max_attempts = 5;
    int feign_retry_period_ms = 100;
    int feign_retry_max_period_ms = 1000;
    int stream_max_connections = 5;               // hystrix.stream.maxConcurrentConnections default
    int stream_default_delay_ms = 500;
    bool verbose = false;
    // Feign's Logger.Level.BASIC lines on stderr for every downstream call, as the reference
    // configures them (HolaResource.java:109). Off by default: see DESIGN.md section 5.
    bool log_feign = false;
    bool serve_aloha = false;  // also answer GET /api/aloha-chaining like the next service of the demo would
};

class HolaApp final : public Application {
 public:
    HolaApp(AppOptions options, GreetingService* greeting, Tracer* tracer, Resolver* resolver);
    ~HolaApp() override;

    void on_worker_start(Worker& worker) override;
    void on_worker_stop(Worker& worker) override;
    void on_worker_exit(Worker& worker) override;
    Disposition handle(Worker& worker, Connection& conn, const RequestHead& request) override;
    void on_connection_closed(Worker& worker, ConnRef ref) override;

    HystrixCommand& command() { return command_; }
    const AppOptions& options() const { return options_; }

 private:
    struct WorkerState;
    struct ChainCall;
    class CallRef;
    struct StreamSub;

    static WorkerState& state_of(Worker& w) { return *static_cast<WorkerState*>(w.app_state()); }

    // Handlers.
    Disposition serve_hola(WorkerState& st, Connection& conn, const RequestHead& rq);
    Disposition serve_chaining(WorkerState& st, Connection& conn, const RequestHead& rq);
    Disposition serve_stream(WorkerState& st, Connection& conn, const RequestHead& rq);
    void serve_static_file(WorkerState& st, Connection& conn, const RequestHead& rq);

    // Greeting text for this worker (refreshing its cached snapshot when the version moved).
    const GreetingSnapshot& current_greeting(WorkerState& st);

    // Downstream call with breaker, timeout and fallback.
    CallRef new_call(WorkerState& st);
    void start_chain(WorkerState& st, const CallRef& call);
    void send_chain_attempt(WorkerState& st, const CallRef& call);
    // `tail` null: the call failed, timed out or was refused - the fallback (or, with fallbacks
    // switched off, a 500) is the answer.
    void finish_chain(WorkerState& st, const CallRef& call, const JsonStringList* tail);
    void append_chain_body(std::string* out, std::string_view hola_text, const JsonStringList* tail) const;
    void append_fallback_response(WorkerState& st, Connection& conn, const RequestHead& rq, const GreetingSnapshot& g);

    // Tracing.
    void record_span(WorkerState& st, const TraceContext& ctx, bool client, int64_t start_us, int64_t end_us,
                     int status, std::string_view url_a, std::string_view url_b, std::string_view url_c);

    // Metrics stream.
    void stream_tick(WorkerState& st, uint64_t key);

    AppOptions options_;
    GreetingService* greeting_;
    Tracer* tracer_;
    Resolver* resolver_;
    HystrixCommand command_;
    std::mutex states_mu_;
    std::vector<std::unique_ptr<WorkerState>> states_;  // one per worker thread
    std::atomic<int> stream_connections_{0};
    std::string aloha_request_prefix_;  // request line + fixed headers, B3 headers are appended
    std::string aloha_url_;             // for the client span's http.url tag
    std::string aloha_stand_in_body_;   // what --serve-aloha answers on /api/aloha-chaining
    std::string fallback_json_tail_;    // `,"<fallback>"]` - the end of the body when the downstream call did not succeed

    // /webjars/**: resolved root and an in-memory copy of the files served so far.
    struct StaticFile {
        int64_t size = 0;
        int64_t mtime_ns = 0;
        std::string body;
    };
    std::string webjars_root_;
    std::mutex static_mu_;
    std::unordered_map<std::string, std::shared_ptr<const StaticFile>> static_cache_;
    size_t static_cache_bytes_ = 0;
};

}  // namespace hola
