// log.hpp — the logging the reference does through spdlog (reference src/simulation.cpp:79-110):
// info and above to the console, debug and above to <result>/execution.log, lines formatted
// "[%Y-%m-%d %H:%M:%S] [level] message", error flushed immediately. Own code, no spdlog.
#pragma once

#include <cstdio>
#include <ctime>
#include <fstream>
#include <iostream>
#include <mutex>
#include <string>

namespace mpslog {
enum Level { Debug = 0, Info = 1, Error = 2 };

class Logger {
public:
    static Logger& get() {
        static Logger L;
        return L;
    }
    void openFile(const std::string& path) {
        std::lock_guard<std::mutex> g(mu_);
        if (file_.is_open()) file_.close();
        file_.open(path, std::ios::trunc);
    }
    void closeFile() {
        std::lock_guard<std::mutex> g(mu_);
        if (file_.is_open()) file_.close();
    }
    void setConsole(bool on) { console_ = on; }
    void log(Level l, const std::string& msg) {
        static const char* names[] = {"debug", "info", "error"};
        std::time_t t = std::time(nullptr);
        char ts[32];
        std::tm tmv;
        localtime_r(&t, &tmv);
        std::strftime(ts, sizeof ts, "%Y-%m-%d %H:%M:%S", &tmv);
        std::string line = std::string("[") + ts + "] [" + names[l] + "] " + msg + "\n";
        std::lock_guard<std::mutex> g(mu_);
        if (console_ && l >= Info) {
            std::cout << line;
            if (l >= Error) std::cout.flush();
        }
        if (file_.is_open()) {
            file_ << line;
            if (l >= Error) file_.flush();
        }
    }

private:
    std::mutex mu_;
    std::ofstream file_;
    bool console_ = true;
};

inline void debug(const std::string& m) { Logger::get().log(Debug, m); }
inline void info(const std::string& m) { Logger::get().log(Info, m); }
inline void error(const std::string& m) { Logger::get().log(Error, m); }
} // namespace mpslog
