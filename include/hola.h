/* hola.h - C ABI of libhola.so.
 *
 * Two audiences: (1) the Python test-suite, which drives the building blocks of the service
 * through ctypes and checks them against hand-written golden vectors and property-based tests;
 * (2) host programs that want to run the service in-process instead of spawning the `hola`
 * executable.
 *
 * Conventions: functions that produce text return it through `char** out` + `size_t* out_len`
 * as a malloc'ed, NUL-terminated buffer that the caller releases with hola_free(). Lists of
 * strings travel in a simple packed form:
 *     u32 count, then per item: u8 present (0 = null), u32 length, bytes
 * (little endian), also released with hola_free(). Return codes are 0 for success unless
 * stated otherwise. All functions are thread-safe unless stated otherwise.
 */
#ifndef HOLA_H_
#define HOLA_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* hola_version(void);
void hola_free(void* p);

/* ---- java.util.Properties text format ------------------------------------------------------
 * Reference: Configuration.java:26-33 loads the greeting file with Properties.load(Reader).
 * On success returns 0 and the entries, in file order, as a packed list k0,v0,k1,v1,...
 * On a malformed \uXXXX escape returns 1 and an error message in *out. */
int hola_properties_parse(const char* text, size_t len, char** out, size_t* out_len);

/* ---- String.format(pattern, oneStringArgument) ---------------------------------------------
 * Reference: HolaResource.java:65. Returns 0 and the formatted text, or 1 and the name of the
 * exception the JVM would have thrown (which the service turns into HTTP 500). */
int hola_format(const char* pattern, size_t pattern_len, const char* arg, size_t arg_len, char** out, size_t* out_len);

/* ---- JSON list-of-strings codec -------------------------------------------------------------
 * Reference: HolaResource.java:73-78 (response of /api/hola-chaining) and :110 (JacksonDecoder
 * for the downstream reply). encode: packed list -> JSON text. decode: JSON text -> packed list
 * (return 0), or 1 and an error message when the text is not an array of scalars. */
int hola_json_encode_list(const char* packed, size_t packed_len, char** out, size_t* out_len);
int hola_json_decode_list(const char* text, size_t len, char** out, size_t* out_len);

/* ---- layered configuration ------------------------------------------------------------------
 * Reference: HolaResource.java:58-64 + Configuration.java. A config object has the service's
 * source stack: system properties (400) > environment (300) > optional static properties file
 * (100) > the file named by system property `conf`, re-read on every lookup (0). Not
 * thread-safe per object. */
typedef struct hola_config hola_config;
hola_config* hola_config_new(void);
void hola_config_free(hola_config* c);
void hola_config_set_system_property(hola_config* c, const char* key, const char* value);
void hola_config_add_properties_file(hola_config* c, const char* path);
/* Raw project-stage-aware lookup with ${var} expansion. Returns 1 when found, else 0. */
int hola_config_resolve(hola_config* c, const char* key, char** out, size_t* out_len);
/* The /api/hola body at time now_ns (any monotonic clock): `hello` looked up with default
 * "Hola de %s" and a 5 s cache, formatted with `hostname`. Returns 0 and the body, or 1 and the
 * formatter's exception text. *refreshed (optional) is set to 1 when the sources were consulted. */
int hola_config_greeting(hola_config* c, const char* hostname, int64_t now_ns, char** out, size_t* out_len, int* refreshed);
/* Number of times the `conf` file has been opened so far. */
uint64_t hola_config_file_reads(const hola_config* c);

/* ---- circuit breaker ------------------------------------------------------------------------
 * Reference: HolaResource.java:104-112 + config.properties:17,19 (Hystrix 1.5.1 semantics).
 * `config_text` is Archaius-style properties text (hystrix.command.default.…); NULL or empty
 * selects the reference's settings. All times are caller-supplied milliseconds. */
typedef struct hola_breaker hola_breaker;
hola_breaker* hola_breaker_new(const char* config_text);
void hola_breaker_free(hola_breaker* b);
/* 0 = run (then call hola_breaker_complete exactly once), 1 = short-circuited, 2 = rejected
 * because the bulkhead is full. */
int hola_breaker_admit(hola_breaker* b, int64_t now_ms);
/* outcome: 0 success, 1 failure, 2 timeout. */
void hola_breaker_complete(hola_breaker* b, int64_t now_ms, int outcome, int execute_ms);
int hola_breaker_is_open(hola_breaker* b, int64_t now_ms);
/* Health counts as the breaker sees them: total requests, errors, error percentage. */
void hola_breaker_health(hola_breaker* b, int64_t now_ms, int64_t* total, int64_t* errors, int* error_pct);
/* The two `data:` payloads of /hystrix.stream. */
int hola_breaker_command_json(hola_breaker* b, int64_t now_ms, char** out, size_t* out_len);
int hola_breaker_pool_json(hola_breaker* b, int64_t now_ms, char** out, size_t* out_len);
int hola_breaker_timeout_ms(const hola_breaker* b);

/* ---- tracing ----------------------------------------------------------------------------------
 * Reference: ZipKinFilter.java:38-44, ZipkinResource.java:33-41, HolaResource.java:97-108.
 * Decides how an inbound request is traced from its B3 headers (NULL = header absent).
 * ids[0..2] = trace id, span id, parent id. Returns a bit set: 1 sampled, 2 joined (ids taken
 * from the caller), 4 has parent. `seed` seeds the id generator (deterministic for tests). */
int hola_b3_server_decision(const char* trace_id, const char* span_id, const char* parent_span_id, const char* sampled,
                            uint64_t seed, uint64_t ids[3]);
/* One span as Zipkin v1 JSON. kind: 0 server (sr/ss), 1 client (cs/cr). shared != 0 omits
 * timestamp/duration. status 0 or 2xx adds no http.status_code tag. */
int hola_span_json(uint64_t trace_id, uint64_t span_id, uint64_t parent_id, int has_parent, int kind, int shared,
                   int64_t start_us, int64_t end_us, int status, const char* url, const char* service_name,
                   uint32_t ipv4_host_order, char** out, size_t* out_len);

/* ---- HTTP wire helpers ------------------------------------------------------------------------
 * Parses one request head. Returns 0 complete, 1 incomplete, 2 malformed. On 0, *out is a JSON
 * object describing what the server understood (method, path, query, version, keep_alive,
 * content_length, chunked, host, accept, b3 headers, head_len). */
int hola_http_parse_request(const char* data, size_t len, char** out, size_t* out_len);
/* Same for a response head (status, keep_alive, content_length, chunked, content_type, head_len). */
int hola_http_parse_response(const char* data, size_t len, char** out, size_t* out_len);
/* Decodes a complete chunked body. Returns 0 done, 1 incomplete, 2 malformed; *consumed gets
 * the number of input bytes used. */
int hola_http_dechunk(const char* data, size_t len, char** out, size_t* out_len, size_t* consumed);
/* 1 when the Accept header admits type/subtype. */
int hola_http_accept_allows(const char* accept, const char* type, const char* subtype);
/* Writes the 29-character IMF-fixdate plus NUL. */
void hola_http_date(int64_t unix_seconds, char out[30]);

/* ---- latency histogram (the load generator's) -------------------------------------------------*/
typedef struct hola_histogram hola_histogram;
hola_histogram* hola_histogram_new(void);
void hola_histogram_free(hola_histogram* h);
void hola_histogram_record(hola_histogram* h, uint64_t value);
uint64_t hola_histogram_percentile(const hola_histogram* h, double pct);
uint64_t hola_histogram_count(const hola_histogram* h);

/* ---- the service, in-process ------------------------------------------------------------------
 * argv takes the same arguments as the `hola` executable (without the program name). Returns
 * NULL and an error message in *error (hola_free it) when the arguments are bad or the port
 * cannot be bound. The server runs on its own threads until hola_service_stop(). */
typedef struct hola_service hola_service;
hola_service* hola_service_start(int argc, const char* const* argv, char** error);
int hola_service_port(const hola_service* s);
int hola_service_threads(const hola_service* s);
/* requests served / connections accepted so far. */
void hola_service_stats(const hola_service* s, uint64_t* requests, uint64_t* connections);
/* spans accepted by the collector / spans dropped so far. */
void hola_service_trace_stats(const hola_service* s, uint64_t* reported, uint64_t* dropped);
/* All counters as one JSON object: totals, process CPU seconds, span counts, the accept mode and
 * one entry per worker thread {id, requests, connections, connections_open, cpu_s} - the place to
 * see whether connections are spread evenly over the threads (the reference relies on Undertow's
 * acceptor dealing them in turn, /root/reference/pom.xml:122-129). The `hola` executable prints
 * the same object as "hola-stats {...}" on stderr when it receives SIGUSR1. */
int hola_service_stats_json(const hola_service* s, char** out, size_t* out_len);
void hola_service_stop(hola_service* s);

#ifdef __cplusplus
}
#endif

#endif /* HOLA_H_ */
