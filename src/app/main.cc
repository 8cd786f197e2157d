// hola - the single self-contained executable, the analogue of `java -jar hola-swarm.jar`
// (/root/reference/Readme.md:13-16, /root/reference/Dockerfile:3).
//
// JVM-style definitions are accepted where the reference takes them, so an existing launch line
// carries over: `hola -Dconf=/path/translation.properties` (Configuration.java:27),
// `-Dhello=...` (DeltaSpike system-property source), `-Dswarm.http.port=N`,
// `-Dhystrix.command.default....=V` (Archaius). Environment: HOSTNAME (HolaResource.java:57),
// ZIPKIN_SERVER_URL (ZipkinResource.java:37). See app/service.h for the rest.
#include <pthread.h>
#include <signal.h>

#include <cstdio>
#include <string>
#include <vector>

#include "app/service.h"

int main(int argc, char** argv) {
    using namespace hola;
    std::vector<std::string> args(argv + 1, argv + argc);
    ServiceConfig config;
    std::string message;
    switch (parse_service_args(args, &config, &message)) {
        case ParseOutcome::kExitOk:
            std::fputs(message.c_str(), stdout);
            return 0;
        case ParseOutcome::kExitUsage:
            std::fputs(message.c_str(), stderr);
            return 2;
        case ParseOutcome::kRun:
            break;
    }

    if (config.print_greeting) {
        std::string body, error;
        if (!resolve_greeting_once(config, &body, &error)) {
            std::fprintf(stderr, "hola: %s\n", error.c_str());
            return 1;
        }
        std::fwrite(body.data(), 1, body.size(), stdout);
        return 0;
    }

    // The worker threads must not receive the termination signals; this thread waits for them.
    sigset_t sigs;
    sigemptyset(&sigs);
    sigaddset(&sigs, SIGINT);
    sigaddset(&sigs, SIGTERM);
    sigaddset(&sigs, SIGUSR1);  // dump counters (Service::stats_json) and carry on
    pthread_sigmask(SIG_BLOCK, &sigs, nullptr);
    signal(SIGPIPE, SIG_IGN);

    Service service(std::move(config));
    std::string error;
    if (!service.start(&error)) {
        std::fprintf(stderr, "hola: %s\n", error.c_str());
        return 1;
    }
    int sig = 0;
    while (sigwait(&sigs, &sig) == 0 && sig == SIGUSR1) {
        std::fprintf(stderr, "hola-stats %s\n", service.stats_json().c_str());
        std::fflush(stderr);
    }
    if (!service.config().quiet) std::fprintf(stderr, "hola: signal %d, shutting down\n", sig);
    service.stop();
    return 0;
}
