// kungfu::Peer — C++ convenience wrapper over the C ABI of libkungfu_host.so.
//
// Same shape as the reference's class (/root/reference/srcs/cpp/include/kungfu/
// peer.hpp:13-161): every collective has a blocking overload and an
// asynchronous one taking a DoneCallback; AllGatherTransform = gather to rank 0,
// apply f there, broadcast the result (/root/reference/srcs/cpp/src/session.cpp:162-181).
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <stdexcept>
#include <string>
#include <vector>

#include "../kungfu_host.h"

namespace kungfu
{
using DoneCallback = std::function<void()>;

// dtype codes (KungFu_Datatype compatible) and ops
enum DType : int { U8 = 0x000108, I32 = 0x010408, I64 = 0x010808, F16 = 0x020208, F32 = 0x020408, F64 = 0x020808, BF16 = 0x040208 };
enum Op : int { SUM = 0, MIN = 1, MAX = 2, PROD = 3 };

class Peer
{
    kf_peer *p_;

    static void check(int rc, const char *what)
    {
        if (rc != 0) throw std::runtime_error(std::string(what) + ": " + kf_last_error());
    }

    // Callback overloads run on the runtime's own worker threads (kf_submit): the op's return code is
    // checked there, and ~Peer() (kf_peer_close) runs what is still queued and joins those threads, so
    // no operation can outlive the object it captured. A failed asynchronous operation is handed to the
    // error handler; the default prints the runtime's message and exits, which is what the reference
    // did for any failed op (/root/reference/srcs/go/libkungfu-comm/main.go:163-187).
    struct Job {
        std::function<int()> op;
        DoneCallback done;
        const char *what;
        Peer *self;
    };
    std::function<void(int, const char *)> on_error_;

    static int run_job(void *arg) { return static_cast<Job *>(arg)->op(); }
    static void end_job(void *arg, int rc)
    {
        Job *j = static_cast<Job *>(arg);
        if (rc != 0) j->self->fail_async(rc, j->what);
        else if (j->done) j->done();
        delete j;
    }
    void fail_async(int rc, const char *what)
    {
        if (on_error_) {
            on_error_(rc, what);
            return;
        }
        std::fprintf(stderr, "[kungfu] asynchronous %s failed (%d): %s\n", what, rc, kf_last_error());
        std::exit(1);
    }
    template <typename F> int async(const char *what, F f, DoneCallback done)
    {
        Job *j = new Job{std::function<int()>(f), std::move(done), what, this};
        const int rc = kf_submit(p_, &Peer::run_job, j, &Peer::end_job);
        if (rc != 0) delete j;
        return rc;
    }

  public:
    Peer() : p_(kf_peer_create_from_env())
    {
        if (!p_) throw std::runtime_error(std::string("kungfu::Peer: ") + kf_last_error());
        check(kf_peer_start(p_), "start");
    }
    ~Peer() { kf_peer_close(p_); }
    Peer(const Peer &) = delete;

    int Rank() const { return kf_rank(p_); }
    int Size() const { return kf_size(p_); }
    int LocalRank() const { return kf_local_rank(p_); }
    int LocalSize() const { return kf_local_size(p_); }
    int HostCount() const { return kf_host_count(p_); }
    uint64_t Uid() const { return kf_uid(p_); }
    bool Detached() const { return kf_detached(p_) != 0; }

    int Barrier() { return kf_barrier(p_); }
    int Barrier(DoneCallback done)
    {
        return async("Barrier", [this] { return kf_barrier(p_); }, done);
    }
    // what to do when a callback-style operation fails (default: print and exit(1))
    void SetAsyncErrorHandler(std::function<void(int, const char *)> h) { on_error_ = std::move(h); }

    int AllReduce(const void *send, void *recv, int64_t count, int dtype, int op, const char *name)
    {
        return kf_all_reduce(p_, send, recv, count, dtype, op, name);
    }
    int AllReduce(const void *send, void *recv, int64_t count, int dtype, int op, const char *name, DoneCallback done)
    {
        std::string n(name);
        return async("AllReduce", [=] { return kf_all_reduce(p_, send, recv, count, dtype, op, n.c_str()); }, done);
    }
    // all-reduce along `tree` (father array of Size() entries, nullptr = active strategy) that also feeds the
    // per-strategy throughput statistics CheckInterference() votes on
    int MonitoredAllReduce(const void *send, void *recv, int64_t count, int dtype, int op, const int32_t *tree,
                           const char *name)
    {
        return kf_monitored_all_reduce(p_, send, recv, count, dtype, op, tree, tree ? Size() : 0, name);
    }
    int MonitoredAllReduce(const void *send, void *recv, int64_t count, int dtype, int op, const int32_t *tree,
                           const char *name, DoneCallback done)
    {
        std::string n(name);
        std::vector<int32_t> t(tree ? tree : nullptr, tree ? tree + Size() : nullptr);
        return async("MonitoredAllReduce", [=] {
            return kf_monitored_all_reduce(p_, send, recv, count, dtype, op, t.empty() ? nullptr : t.data(), (int)t.size(), n.c_str());
        }, done);
    }
    // runs `done` on the runtime's callback path without doing any communication (ordering / wake-up hook)
    int Noop(DoneCallback done)
    {
        return async("Noop", [] { return 0; }, done);
    }
    int CrossAllReduce(const void *send, void *recv, int64_t count, int dtype, int op, const char *name)
    {
        return kf_cross_all_reduce(p_, send, recv, count, dtype, op, name);
    }
    int Reduce(const void *send, void *recv, int64_t count, int dtype, int op, const char *name)
    {
        return kf_reduce(p_, send, recv, count, dtype, op, name);
    }
    int Broadcast(const void *send, void *recv, int64_t count, int dtype, const char *name)
    {
        return kf_broadcast(p_, send, recv, count, dtype, name);
    }
    int Broadcast(const void *send, void *recv, int64_t count, int dtype, const char *name, DoneCallback done)
    {
        std::string n(name);
        return async("Broadcast", [=] { return kf_broadcast(p_, send, recv, count, dtype, n.c_str()); }, done);
    }
    int LocalReduce(const void *send, void *recv, int64_t count, int dtype, int op, const char *name)
    {
        return kf_local_reduce(p_, send, recv, count, dtype, op, name);
    }
    int LocalBroadcast(const void *send, void *recv, int64_t count, int dtype, const char *name)
    {
        return kf_local_broadcast(p_, send, recv, count, dtype, name);
    }
    int Gather(const void *send, void *recv, int64_t count, int dtype, const char *name)
    {
        return kf_gather(p_, send, recv, count, dtype, name);
    }
    int AllGather(const void *send, void *recv, int64_t count, int dtype, const char *name)
    {
        return kf_all_gather(p_, send, recv, count, dtype, name);
    }
    bool Consensus(const void *data, int64_t nbytes, const char *name)
    {
        int agree = 0;
        check(kf_consensus(p_, data, nbytes, name, &agree), "consensus");
        return agree != 0;
    }

    // Gather -> f on rank 0 -> Broadcast. T must be trivially copyable.
    template <typename T, typename F>
    int AllGatherTransform(const T *input, int64_t n, T *output, int64_t m, const char *name, F f)
    {
        const int np = Size();
        std::vector<T> all((size_t)n * np);
        int rc = kf_gather(p_, input, all.data(), n * (int64_t)sizeof(T), U8, (std::string(name) + "::gather").c_str());
        if (rc) return rc;
        if (Rank() == 0) f(all.data(), np, n, output, m);
        return kf_broadcast(p_, output, output, m * (int64_t)sizeof(T), U8, (std::string(name) + "::bcast").c_str());
    }

    // p2p model store
    int Save(const char *name, const void *buf, int64_t nbytes) { return kf_save(p_, nullptr, name, buf, nbytes); }
    int Save(const char *version, const char *name, const void *buf, int64_t nbytes)
    {
        return kf_save(p_, version, name, buf, nbytes);
    }
    int Request(int rank, const char *name, void *buf, int64_t nbytes) { return kf_request(p_, rank, nullptr, name, buf, nbytes); }
    int Request(int rank, const char *version, const char *name, void *buf, int64_t nbytes)
    {
        return kf_request(p_, rank, version, name, buf, nbytes);
    }
    int Request(int rank, const char *name, void *buf, int64_t nbytes, DoneCallback done)
    {
        std::string n(name);
        return async("Request", [=] { return kf_request(p_, rank, nullptr, n.c_str(), buf, nbytes); }, done);
    }

    // control / adaptation / monitoring
    int ResizeCluster(int new_size, bool *changed, bool *detached)
    {
        int c = 0, d = 0;
        int rc = kf_resize_cluster(p_, new_size, &c, &d);
        *changed = c, *detached = d;
        return rc;
    }
    int ResizeClusterFromURL(bool *changed, bool *detached)
    {
        int c = 0, d = 0;
        int rc = kf_resize_cluster_from_url(p_, &c, &d);
        *changed = c, *detached = d;
        return rc;
    }
    int ProposeNewSize(int n) { return kf_propose_new_size(p_, n); }
    int SetTree(const int32_t *tree) { return kf_set_tree(p_, tree, Size()); }
    int GetPeerLatencies(float *out, int n) { return kf_get_peer_latencies(p_, out, n); }
    int GetEgressRates(float *out) { return kf_egress_rates(p_, out, Size()); }
    bool CheckInterference() { return kf_check_interference(p_) != 0; }
    void CalcStats() { kf_calc_stats(p_); }
    void LogStats() { kf_log_stats(p_); }
    void PrintStategyStats() { kf_print_strategy_stats(p_); }
    kf_peer *raw() { return p_; }
};
}  // namespace kungfu
