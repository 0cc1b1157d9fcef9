// Directed graphs over ranks and the topology generators behind every
// all-reduce strategy; Prim MST over measured latencies.
//
// A strategy is a pair (reduce graph, broadcast graph): data flows along the
// reduce graph's edges being combined at every node that has a self loop, then
// along the broadcast graph. Parity: /root/reference/srcs/go/plan/graph/graph.go,
// /root/reference/srcs/go/plan/topology.go:17-160,
// /root/reference/srcs/go/plan/subgraph/subgraph.go:5-31,
// /root/reference/srcs/cpp/include/kungfu/mst.hpp:9-59.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "core/base.hpp"
#include "core/plan.hpp"

namespace kf {

class Graph
{
    struct Node {
        bool self_loop = false;
        std::vector<int> prevs, nexts;
    };
    std::vector<Node> nodes_;

  public:
    Graph() {}
    explicit Graph(int n) : nodes_(n) {}
    int size() const { return (int)nodes_.size(); }
    void add_edge(int i, int j);  // i == j marks a self loop
    bool is_self_loop(int i) const { return nodes_[i].self_loop; }
    const std::vector<int> &prevs(int i) const { return nodes_[i].prevs; }
    const std::vector<int> &nexts(int i) const { return nodes_[i].nexts; }
    bool is_isolated(int i) const { return nodes_[i].prevs.empty() && nodes_[i].nexts.empty(); }
    Graph reverse() const;
    // canonical bytes: used to check that every rank built the same graph
    std::vector<uint8_t> digest_bytes() const;
    std::string debug_string() const;
    // father array -> tree(s); root(s) satisfy father[i] == i. Returns false if
    // an entry is out of range. *roots = number of roots.
    static bool from_forest_array(const std::vector<int32_t> &father, Graph *out, int *roots);
    // true iff every node is reachable from `root` along nexts exactly once
    bool is_valid_tree_with_root(int root) const;
};

// reduce graph derived from a broadcast tree: reversed edges + self loops
Graph gen_default_reduce_graph(const Graph &bcast);
Graph gen_tree(const PeerList &peers);  // star per host + star over masters
Graph gen_binary_tree(int k);
Graph gen_star_bcast_graph(int k, int root);
Graph gen_binary_tree_star(const PeerList &peers, int offset = 0);
Graph gen_multi_star_one(const PeerList &peers, int root_master_idx);
std::vector<Graph> gen_multi_binary_tree_star(const PeerList &peers);
std::vector<Graph> gen_multi_star(const PeerList &peers);
// ring starting after r: (reduce chain ending at r, broadcast chain from r)
void gen_circular_graph_pair(int k, int r, Graph *reduce, Graph *bcast);
// same shapes restricted to a subset of ranks (the host masters)
void gen_sub_circular_graph_pair(int n, const std::vector<int> &ranks, int r, Graph *reduce, Graph *bcast);
Graph gen_sub_binary_tree(int n, const std::vector<int> &ranks);

struct StrategyGraphs {
    Graph reduce, bcast;
};
using StrategyList = std::vector<StrategyGraphs>;

Strategy auto_select(const PeerList &peers);
StrategyList gen_global_strategies(const PeerList &peers, Strategy s);
StrategyList gen_local_strategies(const PeerList &peers);
StrategyList gen_cross_strategies(const PeerList &peers, Strategy s);
StrategyList strategies_from_tree(const Graph &bcast);
std::vector<uint8_t> digest_bytes(const StrategyList &sl);

// One rank's view of a strategy list, compiled for the device executor (kfb_graph_allreduce in
// include/kungfu_b200.h): the buffer is cut into one chunk per strategy; in round r the rank pulls
// chunk edge_plan[e] from rank edge_src[e] for e in [round_begin[r], round_begin[r+1]), reducing
// onto its own copy while r < n_reduce_rounds and overwriting it afterwards. A node's round is its
// depth (longest path from a source) in the reduce graph, then in the broadcast graph, so every
// source of an edge is final when the edge runs. (The host walks the same graphs edge by edge,
// session.cpp:run_graphs; reference: srcs/go/kungfu/session/session.go:218-286.)
struct DevicePlan {
    int n_plans = 0, n_rounds = 0, n_reduce_rounds = 0;
    std::vector<int32_t> round_begin, edge_plan, edge_src;
};
bool compile_device_plan(const StrategyList &sl, int rank, DevicePlan *out, std::string *err);

// Prim on the symmetrised n x n weight matrix (row-major). Returns n-1 edges
// (u, v) flattened; edges[2k] is already in the tree when edges[2k+1] joins.
std::vector<int32_t> minimum_spanning_tree(const std::vector<float> &w, int n);
// father array (rooted at 0) from MST edges
std::vector<int32_t> mst_to_forest(const std::vector<int32_t> &edges, int n);

}  // namespace kf
