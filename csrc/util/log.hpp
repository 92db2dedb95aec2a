// Minimal levelled logger.  Level comes from PCCL_LOG_LEVEL (default ERR), lines go to stdout with
// a wall-clock stamp.  FATAL and BUG terminate the process (behaviour parity with reference:
// log/src/pccl_log.cpp:28-75).
#pragma once
#include <sstream>

namespace pccl {

enum LogLevel { TRACE = 0, DEBUG = 1, INFO = 2, WARN = 3, ERR = 4, FATAL = 5, BUG = 6, NONE = 7 };

LogLevel log_threshold();
void log_emit(LogLevel level, const char *file, int line, const std::string &msg);

class LogLine {
public:
    LogLine(LogLevel level, const char *file, int line) : level_(level), file_(file), line_(line) {}
    ~LogLine() { log_emit(level_, file_, line_, os_.str()); }
    template <typename T> LogLine &operator<<(const T &v) {
        os_ << v;
        return *this;
    }

private:
    LogLevel level_;
    const char *file_;
    int line_;
    std::ostringstream os_;
};

} // namespace pccl

namespace pccl {
struct LogVoidify {
    void operator&(const LogLine &) const {}
};
} // namespace pccl

#define PLOG(level)                                                                                                    \
    (::pccl::level < ::pccl::log_threshold() && ::pccl::level < ::pccl::FATAL)                                         \
        ? (void) 0                                                                                                     \
        : ::pccl::LogVoidify() & ::pccl::LogLine(::pccl::level, __FILE__, __LINE__)
