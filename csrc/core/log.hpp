// Levelled logger with elapsed timestamps, stall detector, small utilities.
// Parity: /root/reference/srcs/go/log/logger.go:14-150,
// /root/reference/srcs/go/utils/stalldetector.go:15-46, /root/reference/srcs/go/utils/*.go.
#pragma once
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <string>
#include <thread>

namespace kf {

enum LogLevel { LOG_DEBUG = 0, LOG_INFO = 1, LOG_WARN = 2, LOG_ERROR = 3 };

void log_set_level(LogLevel l);
LogLevel log_level();
void log_set_level_from_env();  // KUNGFU_CONFIG_LOG_LEVEL
void log_set_file(const std::string &path);
void log_msg(LogLevel l, const char *fmt, ...) __attribute__((format(printf, 2, 3)));

#define KF_DEBUG(...) do { if (kf::log_level() <= kf::LOG_DEBUG) kf::log_msg(kf::LOG_DEBUG, __VA_ARGS__); } while (0)
#define KF_INFO(...) kf::log_msg(kf::LOG_INFO, __VA_ARGS__)
#define KF_WARN(...) kf::log_msg(kf::LOG_WARN, __VA_ARGS__)
#define KF_ERROR(...) kf::log_msg(kf::LOG_ERROR, __VA_ARGS__)

double now_seconds();             // monotonic
double seconds_since_start();     // since process start (or KUNGFU_PROC_START_TIMESTAMP)
std::string show_rate(double bytes_per_sec);
std::string show_size(double bytes);

// Prints "<name> stalled for Ns" every `period` while alive and
// "<name> recovered after Ns" if it had stalled. Opt-in via
// KUNGFU_CONFIG_ENABLE_STALL_DETECTION.
class StallDetector
{
    std::string name_;
    double t0_;
    std::mutex mu_;
    std::condition_variable cv_;
    bool done_ = false;
    bool stalled_ = false;
    std::thread th_;

  public:
    static bool enabled();
    StallDetector(const std::string &name, double period_s = 3.0);
    ~StallDetector();
};

// Repeatedly call f until it returns true or `timeout_s` elapsed.
bool poll_until(const std::function<bool()> &f, double interval_s, double timeout_s);

std::string getenv_str(const char *name, const std::string &def = "");
bool getenv_bool(const char *name, bool def = false);
double parse_duration_seconds(const std::string &s, double def);  // "10ms", "3s", "1m"

}  // namespace kf
