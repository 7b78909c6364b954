// spdlog shim (see ../spdlog.h): stdout sink (no colours).
#pragma once
#include "../spdlog.h"
namespace spdlog::sinks {
class stdout_color_sink_mt : public sink {
public:
    void write(const std::string& line) override { std::fputs(line.c_str(), stdout); }
    void flush() override { std::fflush(stdout); }
};
} // namespace spdlog::sinks
