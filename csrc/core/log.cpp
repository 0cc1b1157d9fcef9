#include "core/log.hpp"

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace kf {

static std::atomic<int> g_level{LOG_INFO};
static FILE *g_file = nullptr;
static std::mutex g_log_mu;
static const double g_t0 = now_seconds();

void log_set_level(LogLevel l) { g_level = l; }
LogLevel log_level() { return (LogLevel)g_level.load(); }

void log_set_level_from_env()
{
    const std::string v = getenv_str("KUNGFU_CONFIG_LOG_LEVEL");
    if (v == "DEBUG") g_level = LOG_DEBUG;
    else if (v == "INFO") g_level = LOG_INFO;
    else if (v == "WARN") g_level = LOG_WARN;
    else if (v == "ERROR") g_level = LOG_ERROR;
}

void log_set_file(const std::string &path)
{
    std::lock_guard<std::mutex> lk(g_log_mu);
    if (g_file) fclose(g_file);
    g_file = path.empty() ? nullptr : fopen(path.c_str(), "ae");  // close-on-exec: a log file is not for the workers to inherit
}

void log_msg(LogLevel l, const char *fmt, ...)
{
    if ((int)l < g_level.load()) return;
    static const char *tags[] = {"[D]", "[I]", "[W]", "[E]"};
    char buf[2048];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    std::lock_guard<std::mutex> lk(g_log_mu);
    FILE *out = g_file ? g_file : (l >= LOG_WARN ? stderr : stdout);
    fprintf(out, "%s [%9.3fs] %s\n", tags[l], now_seconds() - g_t0, buf);
    fflush(out);
}

double now_seconds()
{
    using namespace std::chrono;
    return duration_cast<duration<double>>(steady_clock::now().time_since_epoch()).count();
}

double seconds_since_start() { return now_seconds() - g_t0; }

std::string show_size(double b)
{
    static const char *u[] = {"B", "KiB", "MiB", "GiB", "TiB"};
    int i = 0;
    while (b >= 1024 && i < 4) {
        b /= 1024;
        ++i;
    }
    char buf[64];
    snprintf(buf, sizeof(buf), "%.2f%s", b, u[i]);
    return buf;
}

std::string show_rate(double r) { return show_size(r) + "/s"; }

bool StallDetector::enabled()
{
    static const bool on = getenv_bool("KUNGFU_CONFIG_ENABLE_STALL_DETECTION");
    return on;
}

StallDetector::StallDetector(const std::string &name, double period_s) : name_(name), t0_(now_seconds())
{
    if (!enabled()) return;
    th_ = std::thread([this, period_s] {
        std::unique_lock<std::mutex> lk(mu_);
        while (!cv_.wait_for(lk, std::chrono::duration<double>(period_s), [this] { return done_; })) {
            stalled_ = true;
            KF_WARN("%s stalled for %.1fs", name_.c_str(), now_seconds() - t0_);
        }
    });
}

StallDetector::~StallDetector()
{
    if (!th_.joinable()) return;
    {
        std::lock_guard<std::mutex> lk(mu_);
        done_ = true;
    }
    cv_.notify_all();
    th_.join();
    if (stalled_) KF_WARN("%s recovered after %.1fs", name_.c_str(), now_seconds() - t0_);
}

bool poll_until(const std::function<bool()> &f, double interval_s, double timeout_s)
{
    const double t0 = now_seconds();
    for (;;) {
        if (f()) return true;
        if (now_seconds() - t0 >= timeout_s) return false;
        std::this_thread::sleep_for(std::chrono::duration<double>(interval_s));
    }
}

std::string getenv_str(const char *name, const std::string &def)
{
    const char *v = getenv(name);
    return v ? std::string(v) : def;
}

bool getenv_bool(const char *name, bool def)
{
    const char *v = getenv(name);
    if (!v) return def;
    return !strcmp(v, "1") || !strcmp(v, "true") || !strcmp(v, "TRUE") || !strcmp(v, "True");
}

double parse_duration_seconds(const std::string &s, double def)
{
    if (s.empty()) return def;
    char *end = nullptr;
    double v  = strtod(s.c_str(), &end);
    if (end == s.c_str()) return def;
    std::string u(end);
    if (u == "ms") return v / 1e3;
    if (u == "us") return v / 1e6;
    if (u == "s" || u.empty()) return v;
    if (u == "m") return v * 60;
    if (u == "h") return v * 3600;
    return def;
}

}  // namespace kf
