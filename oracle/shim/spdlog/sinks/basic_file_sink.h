// spdlog shim (see ../spdlog.h): file sink.
#pragma once
#include "../spdlog.h"
namespace spdlog::sinks {
class basic_file_sink_mt : public sink {
public:
    explicit basic_file_sink_mt(const std::string& path, bool truncate = false)
        : f_(path, truncate ? std::ios::trunc : std::ios::app) {}
    void write(const std::string& line) override { f_ << line; }
    void flush() override { f_.flush(); }

private:
    std::ofstream f_;
};
} // namespace spdlog::sinks
