// Session: the CPU collective engine for one cluster version.
//
// A strategy is a (reduce graph, broadcast graph) pair; a buffer is cut into
// 1 MiB chunks, each chunk is hashed onto one strategy of the active list and
// the chunks run concurrently. Parity:
// /root/reference/srcs/go/kungfu/session/{session.go,allreduce.go,allgather.go,
// strategy.go,shard.go,adaptation.go,adaptiveStrategies.go,monitoring.go}.
//
// On a B200 node the tensors themselves are reduced by the NVLink kernels; this
// engine serves CPU tensors, the control protocol (barrier, consensus, handle
// exchange) and the cross-host leg of hierarchical all-reduce.
#pragma once
#include <cstdint>
#include <mutex>
#include <string>
#include <vector>

#include "core/base.hpp"
#include "core/graph.hpp"
#include "ctrl/transport.hpp"

namespace kf {

struct StrategyStat {
    double throughput = 0;  // bytes/s over the last window
    uint64_t acc_size = 0;
    double first_begin = -1, last_end = 0;
    double reff_throughput = 0;
    std::mutex mu;
    void update(double begin, double end, uint64_t size);
    void reset();
};

class Session
{
    PeerList peers_;
    PeerID self_;
    int rank_, local_rank_, local_size_, host_count_;
    Strategy strategy_;
    StrategyList local_, global_, cross_;
    std::vector<std::unique_ptr<StrategyStat>> global_stats_;
    std::vector<double> stat_log_;
    Client *client_;
    Server *server_;
    double timeout_s_;
    bool hash_by_name_;
    std::string last_error_;
    std::mutex err_mu_;

    bool run_graphs(const Workspace &w, const std::vector<const Graph *> &graphs);
    bool run_strategies(const Workspace &w, const StrategyList &sl, bool monitored);
    bool send_to(int rank, const std::string &name, const void *data, size_t n);
    bool recv_from(int rank, const std::string &name, void *buf, size_t n);
    void fail(const std::string &msg);
    void set_global(const StrategyList &sl);

  public:
    Session(const PeerList &peers, const PeerID &self, Strategy strategy, Client *client, Server *server);

    int rank() const { return rank_; }
    const StrategyList &global_strategies() const { return global_; }
    int size() const { return (int)peers_.size(); }
    int local_rank() const { return local_rank_; }
    int local_size() const { return local_size_; }
    int host_count() const { return host_count_; }
    const PeerList &peers() const { return peers_; }
    std::string last_error();

    bool barrier();
    // true iff every rank passed identical bytes
    bool bytes_consensus(const void *data, size_t n, const std::string &name, bool *agree);
    bool all_reduce(const Workspace &w);
    bool monitored_all_reduce(const Workspace &w);
    bool all_reduce_with(const std::vector<int32_t> &forest, const Workspace &w);
    bool cross_all_reduce(const Workspace &w);
    bool reduce(const Workspace &w);     // to rank 0
    bool broadcast(const Workspace &w);  // from rank 0
    bool local_reduce(const Workspace &w);
    bool local_broadcast(const Workspace &w);
    bool gather(const Workspace &w);      // recv holds size()*count on rank 0
    bool all_gather(const Workspace &w);  // recv holds size()*count everywhere

    // adaptation
    bool set_global_strategy(const StrategyList &sl);  // with digest consensus
    bool set_tree(const std::vector<int32_t> &forest);
    bool set_strategy(Strategy s);
    int num_strategies() const { return (int)global_.size(); }
    // text form of the active global strategies, one "reduce=... bcast=..." line each
    std::string strategies_text() const;
    Strategy strategy() const { return strategy_; }
    void calc_stats();
    void log_stats();
    void print_strategy_stats();
    bool check_interference();  // majority vote: throughput < 0.8 * reference
    std::vector<float> peer_latencies();
    std::vector<float> egress_rates(NetMonitor *m);
};

}  // namespace kf
