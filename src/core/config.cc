#include "core/config.h"

#include <algorithm>
#include <cstdlib>

namespace hola {

bool SystemPropertySource::lookup(const std::string& key, std::string* value) {
    auto it = props_.find(key);
    if (it == props_.end()) return false;
    *value = it->second;
    return true;
}

bool EnvironmentSource::lookup(const std::string& key, std::string* value) {
    if (key.empty() || key.find('=') != std::string::npos) return false;
    const char* v = std::getenv(key.c_str());
    if (!v) return false;
    *value = v;
    return true;
}

StaticPropertiesSource::StaticPropertiesSource(std::string label, const std::string& path)
    : label_(std::move(label)) {
    props_.load_file(path);
}

bool StaticPropertiesSource::lookup(const std::string& key, std::string* value) {
    const std::string* v = props_.get(key);
    if (!v) return false;
    *value = *v;
    return true;
}

bool ConfFileSource::lookup(const std::string& key, std::string* value) {
    const std::string* path = system_->get("conf");
    if (on_read_) on_read_(path);
    if (!path) return false;  // the reference hits a NullPointerException here and swallows it
    ++reads_;
    Properties props;
    if (!props.load_file(*path)) return false;
    const std::string* v = props.get(key);
    if (!v) return false;
    *value = *v;
    return true;
}

void ConfigResolver::add_source(std::unique_ptr<ConfigSource> source) {
    sources_.push_back(std::move(source));
    std::stable_sort(sources_.begin(), sources_.end(),
                     [](const auto& a, const auto& b) { return a->ordinal() > b->ordinal(); });
}

bool ConfigResolver::property_value(const std::string& key, std::string* value) {
    for (auto& source : sources_) {
        if (source->lookup(key, value)) return true;
    }
    return false;
}

std::string ConfigResolver::project_stage() {
    static const char* const kKeys[] = {"org.apache.deltaspike.ProjectStage", "javax.faces.PROJECT_STAGE",
                                        "faces.PROJECT_STAGE"};
    std::string v;
    for (const char* key : kKeys) {
        if (property_value(key, &v) && !v.empty()) return v;
    }
    return "Production";
}

bool ConfigResolver::resolve(const std::string& key, std::string* value) {
    // A value may refer to other keys, which may refer to further keys; a cycle must not take
    // the process down.
    struct DepthGuard {
        int& d;
        explicit DepthGuard(int& depth) : d(depth) { ++d; }
        ~DepthGuard() { --d; }
    } guard(depth_);
    if (depth_ > 16) return false;
    const std::string stage = project_stage();
    bool found = false;
    if (!stage.empty()) found = property_value(key + "." + stage, value);
    if (!found) found = property_value(key, value);
    if (!found) return false;
    *value = expand_variables(std::move(*value));
    return true;
}

std::string ConfigResolver::expand_variables(std::string value) {
    size_t start = 0;
    int guard = 0;  // a self-referencing value must not spin forever
    while ((start = value.find("${", start)) != std::string::npos && guard++ < 64) {
        const size_t end = value.find('}', start);
        if (end == std::string::npos) break;
        const std::string name = value.substr(start + 2, end - start - 2);
        if (name.empty()) break;
        std::string replacement;
        if (resolve(name, &replacement)) {
            const std::string token = "${" + name + "}";
            size_t at = 0;
            while ((at = value.find(token, at)) != std::string::npos) {
                value.replace(at, token.size(), replacement);
                at += replacement.size();
            }
        }
        ++start;
    }
    return value;
}

CachedConfigValue::CachedConfigValue(ConfigResolver* resolver, std::string key, std::string default_value,
                                     int64_t cache_ns, bool log_changes, Logger logger)
    : resolver_(resolver),
      key_(std::move(key)),
      default_(std::move(default_value)),
      cache_ns_(cache_ns),
      log_changes_(log_changes),
      logger_(std::move(logger)) {}

const std::string& CachedConfigValue::get(int64_t now_ns, bool* refreshed) {
    if (refreshed) *refreshed = false;
    if (has_value_ && cache_ns_ > 0 && now_ns <= reload_after_) return value_;
    std::string resolved;
    const bool found = resolver_->resolve(key_, &resolved);
    std::string next = (found && !resolved.empty()) ? std::move(resolved) : default_;
    if (!has_value_ || next != value_) {
        if (log_changes_ && logger_) logger_("New value " + next + " for key " + key_ + ".");
        ++version_;
    }
    value_ = std::move(next);
    has_value_ = true;
    reload_after_ = now_ns + cache_ns_;
    if (refreshed) *refreshed = true;
    return value_;
}

}  // namespace hola
