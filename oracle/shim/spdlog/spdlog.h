// Minimal stand-in for the spdlog 1.14.1 API surface that the reference sources use
// (the reference's submodules/spdlog is missing from the mount; it is logging only, no
// arithmetic). TEST INFRASTRUCTURE: used solely to compile the unmodified reference
// into oracle/_ref/. Written from scratch; `{}` formatting goes through std::vformat.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <ctime>
#include <format>
#include <fstream>
#include <functional>
#include <iostream>
#include <memory>
#include <mutex>
#include <string>
#include <string_view>
#include <vector>

namespace spdlog {
namespace level {
enum level_enum { trace = 0, debug = 1, info = 2, warn = 3, err = 4, critical = 5, off = 6 };
inline const char* name(level_enum l) {
    static const char* n[] = {"trace", "debug", "info", "warning", "error", "critical", "off"};
    return n[l];
}
} // namespace level

namespace sinks {
class sink {
public:
    virtual ~sink() = default;
    virtual void write(const std::string& line) = 0;
    virtual void flush()                        = 0;
    void set_level(level::level_enum l) { lvl_ = l; }
    bool should_log(level::level_enum l) const { return l >= lvl_; }

private:
    level::level_enum lvl_ = level::trace;
};
} // namespace sinks
using sink_ptr = std::shared_ptr<sinks::sink>;

class logger {
public:
    template <class It>
    logger(std::string name, It b, It e) : name_(std::move(name)), sinks_(b, e) {}
    explicit logger(std::string name) : name_(std::move(name)) {}
    void set_level(level::level_enum l) { lvl_ = l; }
    void flush_on(level::level_enum l) { flush_lvl_ = l; }
    void flush() {
        for (auto& s : sinks_) s->flush();
    }
    void log(level::level_enum l, const std::string& msg) {
        if (l < lvl_) return;
        std::time_t t = std::time(nullptr);
        char ts[32];
        std::strftime(ts, sizeof ts, "%Y-%m-%d %H:%M:%S", std::localtime(&t));
        std::string line = std::string("[") + ts + "] [" + level::name(l) + "] " + msg + "\n";
        std::lock_guard<std::mutex> g(mu_);
        if (sinks_.empty()) {
            if (l >= level::info) std::fputs(line.c_str(), stdout);
        }
        for (auto& s : sinks_)
            if (s->should_log(l)) s->write(line);
        if (l >= flush_lvl_) flush();
    }
    template <class... A>
    void error(std::string_view f, A&&... a) { log(level::err, std::vformat(f, std::make_format_args(a...))); }
    template <class... A>
    void info(std::string_view f, A&&... a) { log(level::info, std::vformat(f, std::make_format_args(a...))); }
    template <class... A>
    void debug(std::string_view f, A&&... a) { log(level::debug, std::vformat(f, std::make_format_args(a...))); }

private:
    std::string name_;
    std::vector<sink_ptr> sinks_;
    level::level_enum lvl_ = level::info, flush_lvl_ = level::off;
    std::mutex mu_;
};

inline std::shared_ptr<logger>& default_logger_ref() {
    static std::shared_ptr<logger> l = std::make_shared<logger>("");
    return l;
}
inline std::shared_ptr<logger> default_logger() { return default_logger_ref(); }
inline void set_default_logger(std::shared_ptr<logger> l) { default_logger_ref() = std::move(l); }
inline void set_pattern(const std::string&) {}
inline void set_level(level::level_enum l) { default_logger()->set_level(l); }
inline void flush_on(level::level_enum l) { default_logger()->flush_on(l); }
inline void set_error_handler(std::function<void(const std::string&)>) {}

template <class... A>
inline void error(std::string_view f, A&&... a) { default_logger()->log(level::err, std::vformat(f, std::make_format_args(a...))); }
template <class... A>
inline void info(std::string_view f, A&&... a) { default_logger()->log(level::info, std::vformat(f, std::make_format_args(a...))); }
template <class... A>
inline void debug(std::string_view f, A&&... a) { default_logger()->log(level::debug, std::vformat(f, std::make_format_args(a...))); }
template <class... A>
inline void warn(std::string_view f, A&&... a) { default_logger()->log(level::warn, std::vformat(f, std::make_format_args(a...))); }
} // namespace spdlog
