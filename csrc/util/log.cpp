#include "log.hpp"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <mutex>

namespace pccl {

static LogLevel parse_level() {
    const char *env = std::getenv("PCCL_LOG_LEVEL");
    if (!env) return ERR;
    static const struct {
        const char *name;
        LogLevel lvl;
    } table[] = {{"TRACE", TRACE}, {"DEBUG", DEBUG}, {"INFO", INFO}, {"WARN", WARN},
                 {"ERR", ERR},     {"ERROR", ERR},   {"FATAL", FATAL}, {"NONE", NONE}};
    for (const auto &e : table)
        if (strcasecmp(env, e.name) == 0) return e.lvl;
    return ERR;
}

LogLevel log_threshold() {
    static const LogLevel lvl = parse_level();
    return lvl;
}

void log_emit(LogLevel level, const char *file, int line, const std::string &msg) {
    static std::mutex mu;
    static const char *names[] = {"TRACE", "DEBUG", "INFO", "WARN", "ERROR", "FATAL", "BUG", "NONE"};
    if (level >= log_threshold() && log_threshold() != NONE) {
        const auto now = std::chrono::system_clock::now();
        const auto t = std::chrono::system_clock::to_time_t(now);
        const auto ms = std::chrono::duration_cast<std::chrono::milliseconds>(now.time_since_epoch()).count() % 1000;
        struct tm tmv {};
        localtime_r(&t, &tmv);
        char stamp[32];
        strftime(stamp, sizeof stamp, "%H:%M:%S", &tmv);
        const char *base = strrchr(file, '/');
        base = base ? base + 1 : file;
        std::lock_guard<std::mutex> lock(mu);
        fprintf(stdout, "[%s.%03d] [%s] %s:%d %s\n", stamp, static_cast<int>(ms), names[level], base, line,
                msg.c_str());
        fflush(stdout);
    }
    if (level == FATAL || level == BUG) {
        if (level == BUG) fprintf(stderr, "pccl: internal invariant violated at %s:%d: %s\n", file, line, msg.c_str());
        std::_Exit(1);
    }
}

} // namespace pccl
