// Small header-only helpers for C++ users of the runtime (agents, schedulers
// built on kungfu::Peer). Counterparts of the reference's utils directory
// (/root/reference/srcs/cpp/include/kungfu/utils/{channel,waiter,ema,
// handler_manager,error_checker,trace}.hpp), re-designed: channels can be
// closed and polled, the latch counts and times out, handles complete through
// a condition variable instead of a spin-sleep loop.
#pragma once
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <deque>
#include <map>
#include <mutex>
#include <optional>
#include <string>

namespace kungfu
{
// Unbounded many-producer / single-consumer queue. get() blocks until an item
// arrives or the channel is closed and drained (then returns nullopt).
template <typename T> class Channel
{
    std::mutex mu_;
    std::condition_variable cv_;
    std::deque<T> items_;
    bool closed_ = false;

  public:
    bool put(T x)
    {
        {
            std::lock_guard<std::mutex> lk(mu_);
            if (closed_) return false;
            items_.push_back(std::move(x));
        }
        cv_.notify_one();
        return true;
    }

    std::optional<T> get()
    {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [&] { return !items_.empty() || closed_; });
        return pop_locked();
    }

    template <typename Rep, typename Period> std::optional<T> get_for(std::chrono::duration<Rep, Period> d)
    {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait_for(lk, d, [&] { return !items_.empty() || closed_; });
        return pop_locked();
    }

    std::optional<T> try_get()
    {
        std::lock_guard<std::mutex> lk(mu_);
        return pop_locked();
    }

    void close()
    {
        {
            std::lock_guard<std::mutex> lk(mu_);
            closed_ = true;
        }
        cv_.notify_all();
    }

    size_t size()
    {
        std::lock_guard<std::mutex> lk(mu_);
        return items_.size();
    }

  private:
    std::optional<T> pop_locked()
    {
        if (items_.empty()) return std::nullopt;
        std::optional<T> x(std::move(items_.front()));
        items_.pop_front();
        return x;
    }
};

// Count-down latch: wait() returns once done() has been called `n` times.
class Latch
{
    std::mutex mu_;
    std::condition_variable cv_;
    int pending_;

  public:
    explicit Latch(int n = 1) : pending_(n) {}

    void done()
    {
        std::lock_guard<std::mutex> lk(mu_);
        if (pending_ > 0 && --pending_ == 0) cv_.notify_all();
    }

    void wait()
    {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [&] { return pending_ == 0; });
    }

    template <typename Rep, typename Period> bool wait_for(std::chrono::duration<Rep, Period> d)
    {
        std::unique_lock<std::mutex> lk(mu_);
        return cv_.wait_for(lk, d, [&] { return pending_ == 0; });
    }
};

// v <- alpha v + (1 - alpha) x; the first sample passes through.
template <typename T> class ExponentialMovingAverage
{
    T alpha_;
    T value_{};
    bool primed_ = false;

  public:
    explicit ExponentialMovingAverage(T alpha) : alpha_(alpha) {}

    T update(T x)
    {
        value_  = primed_ ? alpha_ * value_ + (T(1) - alpha_) * x : x;
        primed_ = true;
        return value_;
    }

    T get() const { return value_; }
    bool primed() const { return primed_; }
};

// Integer handles for asynchronous operations: create() before launching,
// done(h) from the completion callback, wait(h) / wait_all() from the caller.
class HandleTable
{
    std::mutex mu_;
    std::condition_variable cv_;
    std::map<int, bool> done_;
    int next_ = 1;

  public:
    int create()
    {
        std::lock_guard<std::mutex> lk(mu_);
        done_[next_] = false;
        return next_++;
    }

    void done(int h)
    {
        {
            std::lock_guard<std::mutex> lk(mu_);
            auto it = done_.find(h);
            if (it != done_.end()) it->second = true;
        }
        cv_.notify_all();
    }

    // false if the handle is unknown (never created or already waited for)
    bool wait(int h)
    {
        std::unique_lock<std::mutex> lk(mu_);
        auto it = done_.find(h);
        if (it == done_.end()) return false;
        cv_.wait(lk, [&] { return done_[h]; });
        done_.erase(h);
        return true;
    }

    void wait_all()
    {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [&] {
            for (auto &kv : done_)
                if (!kv.second) return false;
            return true;
        });
        done_.clear();
    }

    size_t pending()
    {
        std::lock_guard<std::mutex> lk(mu_);
        size_t n = 0;
        for (auto &kv : done_) n += !kv.second;
        return n;
    }
};

// Wall-clock scope timer; prints when KUNGFU_ENABLE_TRACE is set in the environment.
class TraceScope
{
    const char *name_;
    std::chrono::steady_clock::time_point t0_;

  public:
    explicit TraceScope(const char *name) : name_(name), t0_(std::chrono::steady_clock::now()) {}
    ~TraceScope()
    {
        static const bool on = std::getenv("KUNGFU_ENABLE_TRACE") != nullptr;
        if (on) std::fprintf(stderr, "[trace] %s took %.3f ms\n", name_, elapsed_ms());
    }
    double elapsed_ms() const
    {
        return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0_).count();
    }
};
}  // namespace kungfu

#define KUNGFU_TRACE_SCOPE(name) ::kungfu::TraceScope kungfu_trace_scope_##__LINE__(name)

// Abort with file:line when a runtime call (kf_* / kfb_* returning non-zero on failure) fails.
#define KUNGFU_CHECK(expr)                                                                                           \
    do {                                                                                                             \
        const int kungfu_check_rc = (expr);                                                                          \
        if (kungfu_check_rc != 0) {                                                                                  \
            std::fprintf(stderr, "%s:%d: %s failed with %d\n", __FILE__, __LINE__, #expr, kungfu_check_rc);          \
            std::exit(1);                                                                                            \
        }                                                                                                            \
    } while (0)
