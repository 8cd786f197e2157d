// Layered configuration lookup with the semantics the reference gets from Apache DeltaSpike.
//
// The greeting pattern is looked up as
//   ConfigResolver.resolve("hello").withDefault("Hola de %s").logChanges(true)
//                 .cacheFor(SECONDS, 5).getValue()
// (/root/reference/src/main/java/com/redhat/developers/msa/hola/HolaResource.java:58-64). The
// sources DeltaSpike consults, highest ordinal first:
//   400  JVM system properties      -> here: the -Dkey=value arguments of the executable
//   300  process environment
//   100  META-INF/apache-deltaspike.properties (optional; --deltaspike-properties <file> here)
//     0  "ConfigFile": the properties file named by system property `conf`, re-read on every
//        lookup, any failure meaning "no entries"
//        (/root/reference/src/main/java/com/redhat/developers/msa/hola/Configuration.java:23-48)
// A typed lookup is project-stage aware (tries "<key>.<ProjectStage>" first; the stage defaults
// to Production), expands ${other.key} references, and falls back to the default when the
// result is missing or empty.
//
// None of this is on the request path: the server formats the greeting once per refresh and the
// workers serve precomputed bytes (see app/greeting.h).
#pragma once

#include <cstdint>
#include <functional>
#include <map>
#include <memory>
#include <string>
#include <string_view>
#include <vector>

#include "core/properties.h"

namespace hola {

class ConfigSource {
 public:
    virtual ~ConfigSource() = default;
    virtual int ordinal() const = 0;
    virtual std::string name() const = 0;
    // Returns true and sets *value when this source defines `key`.
    virtual bool lookup(const std::string& key, std::string* value) = 0;
};

// Ordinal 400. Holds the -D definitions given on the command line.
class SystemPropertySource : public ConfigSource {
 public:
    void set(std::string key, std::string value) { props_[std::move(key)] = std::move(value); }
    const std::string* get(const std::string& key) const {
        auto it = props_.find(key);
        return it == props_.end() ? nullptr : &it->second;
    }
    int ordinal() const override { return 400; }
    std::string name() const override { return "system-properties"; }
    bool lookup(const std::string& key, std::string* value) override;

 private:
    std::map<std::string, std::string> props_;
};

// Ordinal 300. Reads the live process environment.
class EnvironmentSource : public ConfigSource {
 public:
    int ordinal() const override { return 300; }
    std::string name() const override { return "environment-properties"; }
    bool lookup(const std::string& key, std::string* value) override;
};

// Ordinal 100. A properties file loaded once at start-up.
class StaticPropertiesSource : public ConfigSource {
 public:
    StaticPropertiesSource(std::string label, const std::string& path);
    int ordinal() const override { return 100; }
    std::string name() const override { return label_; }
    bool lookup(const std::string& key, std::string* value) override;

 private:
    std::string label_;
    Properties props_;
};

// Ordinal 0. The reference's own source: opens and parses the file named by system property
// `conf` on every lookup. `on_read` (optional) is told the path each time, which is where the
// reference prints it to stdout (Configuration.java:28).
class ConfFileSource : public ConfigSource {
 public:
    explicit ConfFileSource(const SystemPropertySource* system) : system_(system) {}
    void set_on_read(std::function<void(const std::string* path)> cb) { on_read_ = std::move(cb); }
    int ordinal() const override { return 0; }
    std::string name() const override { return "ConfigFile"; }
    bool lookup(const std::string& key, std::string* value) override;
    uint64_t reads() const { return reads_; }

 private:
    const SystemPropertySource* system_;
    std::function<void(const std::string* path)> on_read_;
    uint64_t reads_ = 0;
};

class ConfigResolver {
 public:
    // Takes ownership. Sources are consulted in descending ordinal order; among equal ordinals
    // the one added first wins.
    void add_source(std::unique_ptr<ConfigSource> source);

    // Raw lookup across the sources (DeltaSpike's ConfigResolver.getPropertyValue).
    bool property_value(const std::string& key, std::string* value);

    // Typed-resolver lookup: project-stage aware, then ${var} expansion.
    bool resolve(const std::string& key, std::string* value);

    // The active project stage: `org.apache.deltaspike.ProjectStage`, then
    // `javax.faces.PROJECT_STAGE`, `faces.PROJECT_STAGE`, else "Production".
    std::string project_stage();

    std::string expand_variables(std::string value);

 private:
    std::vector<std::unique_ptr<ConfigSource>> sources_;
    int depth_ = 0;
};

// The equivalent of one TypedResolver held for the life of the process: key, default, time-based
// cache and change logging. Not thread-safe; the caller serialises access.
class CachedConfigValue {
 public:
    using Logger = std::function<void(const std::string& message)>;

    CachedConfigValue(ConfigResolver* resolver, std::string key, std::string default_value,
                      int64_t cache_ns, bool log_changes, Logger logger = nullptr);

    // Returns the value, re-resolving when the cache period has elapsed. `now_ns` is any
    // monotonic clock. `*refreshed` (optional) is set when a re-resolve took place.
    const std::string& get(int64_t now_ns, bool* refreshed = nullptr);
    int64_t reload_after_ns() const { return reload_after_; }
    uint64_t version() const { return version_; }

 private:
    ConfigResolver* resolver_;
    std::string key_;
    std::string default_;
    int64_t cache_ns_;
    bool log_changes_;
    Logger logger_;
    bool has_value_ = false;
    std::string value_;
    int64_t reload_after_ = INT64_MIN;
    uint64_t version_ = 0;
};

}  // namespace hola
