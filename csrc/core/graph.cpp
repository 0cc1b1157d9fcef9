#include "core/graph.hpp"

#include <algorithm>
#include <limits>
#include <map>
#include <sstream>

namespace kf {

void Graph::add_edge(int i, int j)
{
    if (i == j) {
        nodes_[i].self_loop = true;
        return;
    }
    nodes_[i].nexts.push_back(j);
    nodes_[j].prevs.push_back(i);
}

Graph Graph::reverse() const
{
    Graph r(size());
    for (int i = 0; i < size(); ++i) {
        if (nodes_[i].self_loop) r.add_edge(i, i);
        for (int j : nodes_[i].nexts) r.add_edge(j, i);
    }
    return r;
}

static void put32(std::vector<uint8_t> *b, uint32_t v)
{
    for (int k = 0; k < 4; ++k) b->push_back((v >> (8 * k)) & 255);
}

std::vector<uint8_t> Graph::digest_bytes() const
{
    std::vector<uint8_t> b;
    put32(&b, (uint32_t)size());
    for (int i = 0; i < size(); ++i) {
        put32(&b, nodes_[i].self_loop ? 1u : 0u);
        std::vector<int> nx = nodes_[i].nexts;
        std::sort(nx.begin(), nx.end());
        put32(&b, (uint32_t)nx.size());
        for (int j : nx) put32(&b, (uint32_t)j);
    }
    return b;
}

std::string Graph::debug_string() const
{
    std::ostringstream os;
    os << "[" << size() << "]{";
    for (int i = 0; i < size(); ++i) {
        if (nodes_[i].self_loop) os << "(" << i << ")";
        for (int j : nodes_[i].nexts) os << "(" << i << "->" << j << ")";
    }
    os << "}";
    return os.str();
}

bool Graph::from_forest_array(const std::vector<int32_t> &father, Graph *out, int *roots)
{
    const int n = (int)father.size();
    Graph g(n);
    int m = 0;
    for (int i = 0; i < n; ++i) {
        const int f = father[i];
        if (f < 0 || f >= n) return false;
        if (f == i) {
            ++m;
        } else {
            g.add_edge(f, i);
        }
    }
    *out = g;
    if (roots) *roots = m;
    return true;
}

bool Graph::is_valid_tree_with_root(int root) const
{
    const int n = size();
    if (root < 0 || root >= n) return false;
    std::vector<int> seen(n, 0), stack{root};
    int visited = 0;
    while (!stack.empty()) {
        int u = stack.back();
        stack.pop_back();
        if (seen[u]++) return false;
        ++visited;
        for (int v : nodes_[u].nexts) stack.push_back(v);
    }
    if (visited != n) return false;
    if (!nodes_[root].prevs.empty()) return false;
    for (int i = 0; i < n; ++i)
        if (i != root && nodes_[i].prevs.size() != 1) return false;
    return true;
}

Graph gen_default_reduce_graph(const Graph &bcast)
{
    Graph r = bcast.reverse();
    for (int i = 0; i < bcast.size(); ++i) r.add_edge(i, i);
    return r;
}

static void local_masters(const PeerList &peers, std::vector<int> *masters, std::vector<int> *parent)
{
    peers.partition_by_host(masters, parent);
}

static Graph local_stars(const PeerList &peers, std::vector<int> *masters)
{
    Graph g((int)peers.size());
    std::vector<int> parent;
    local_masters(peers, masters, &parent);
    for (int r = 0; r < (int)peers.size(); ++r)
        if (parent[r] != r) g.add_edge(parent[r], r);
    return g;
}

Graph gen_tree(const PeerList &peers)
{
    std::vector<int> masters;
    Graph g = local_stars(peers, &masters);
    for (size_t i = 1; i < masters.size(); ++i) g.add_edge(masters[0], masters[i]);
    return g;
}

Graph gen_binary_tree(int k)
{
    Graph g(k);
    for (int i = 0; i < k; ++i) {
        if (2 * i + 1 < k) g.add_edge(i, 2 * i + 1);
        if (2 * i + 2 < k) g.add_edge(i, 2 * i + 2);
    }
    return g;
}

Graph gen_star_bcast_graph(int k, int root)
{
    Graph g(k);
    for (int i = 0; i < k; ++i)
        if (i != root) g.add_edge(root, i);
    return g;
}

Graph gen_binary_tree_star(const PeerList &peers, int offset)
{
    std::vector<int> masters;
    Graph g     = local_stars(peers, &masters);
    const int k = (int)masters.size();
    if (k > 1) {
        auto idx = [&](int i) { return masters[(i + offset) % k]; };
        for (int i = 0; i < k; ++i) {
            if (2 * i + 1 < k) g.add_edge(idx(i), idx(2 * i + 1));
            if (2 * i + 2 < k) g.add_edge(idx(i), idx(2 * i + 2));
        }
    }
    return g;
}

Graph gen_multi_star_one(const PeerList &peers, int root)
{
    std::vector<int> masters;
    Graph g     = local_stars(peers, &masters);
    const int k = (int)masters.size();
    for (int i = 0; i < k; ++i)
        if (k > 1 && i != root) g.add_edge(masters[root], masters[i]);
    return g;
}

std::vector<Graph> gen_multi_binary_tree_star(const PeerList &peers)
{
    std::vector<int> masters, parent;
    local_masters(peers, &masters, &parent);
    std::vector<Graph> gs;
    for (size_t i = 0; i < masters.size(); ++i) gs.push_back(gen_binary_tree_star(peers, (int)i));
    return gs;
}

std::vector<Graph> gen_multi_star(const PeerList &peers)
{
    std::vector<int> masters, parent;
    local_masters(peers, &masters, &parent);
    std::vector<Graph> gs;
    for (size_t i = 0; i < masters.size(); ++i) gs.push_back(gen_multi_star_one(peers, (int)i));
    return gs;
}

void gen_circular_graph_pair(int k, int r, Graph *reduce, Graph *bcast)
{
    std::vector<int> all(k);
    for (int i = 0; i < k; ++i) all[i] = i;
    gen_sub_circular_graph_pair(k, all, r, reduce, bcast);
}

void gen_sub_circular_graph_pair(int n, const std::vector<int> &ranks, int r, Graph *reduce, Graph *bcast)
{
    const int k = (int)ranks.size();
    Graph g(n), b(n);
    for (int i = 0; i < k; ++i) g.add_edge(ranks[i], ranks[i]);
    for (int i = 1; i < k; ++i) {
        g.add_edge(ranks[(r + i) % k], ranks[(r + i + 1) % k]);
        b.add_edge(ranks[(r + i - 1) % k], ranks[(r + i) % k]);
    }
    *reduce = g;
    *bcast  = b;
}

Graph gen_sub_binary_tree(int n, const std::vector<int> &ranks)
{
    Graph g(n);
    const int k = (int)ranks.size();
    for (int i = 0; i < k; ++i) {
        if (2 * i + 1 < k) g.add_edge(ranks[i], ranks[2 * i + 1]);
        if (2 * i + 2 < k) g.add_edge(ranks[i], ranks[2 * i + 2]);
    }
    return g;
}

Strategy auto_select(const PeerList &peers)
{
    return peers.host_count() <= 1 ? Star : BinaryTreeStar;
}

StrategyList strategies_from_tree(const Graph &bcast)
{
    return StrategyList{StrategyGraphs{gen_default_reduce_graph(bcast), bcast}};
}

StrategyList gen_global_strategies(const PeerList &peers, Strategy s)
{
    const int k = (int)peers.size();
    if (s == Auto) s = auto_select(peers);
    StrategyList sl;
    switch (s) {
    case Star: return strategies_from_tree(gen_star_bcast_graph(k, 0));
    case Tree: return strategies_from_tree(gen_tree(peers));
    case BinaryTree: return strategies_from_tree(gen_binary_tree(k));
    case BinaryTreeStar: return strategies_from_tree(gen_binary_tree_star(peers));
    case MultiStar:
        for (auto &g : gen_multi_star(peers)) sl.push_back({gen_default_reduce_graph(g), g});
        return sl;
    case MultiBinaryTreeStar:
        for (auto &g : gen_multi_binary_tree_star(peers)) sl.push_back({gen_default_reduce_graph(g), g});
        return sl;
    case Clique:
        for (int r = 0; r < k; ++r) {
            Graph g = gen_star_bcast_graph(k, r);
            sl.push_back({gen_default_reduce_graph(g), g});
        }
        return sl;
    case Ring:
        for (int r = 0; r < k; ++r) {
            StrategyGraphs sg;
            gen_circular_graph_pair(k, r, &sg.reduce, &sg.bcast);
            sl.push_back(sg);
        }
        return sl;
    default: break;
    }
    return strategies_from_tree(gen_star_bcast_graph(k, 0));
}

StrategyList gen_local_strategies(const PeerList &peers)
{
    std::vector<int> masters, parents;
    peers.partition_by_host(&masters, &parents);
    std::vector<int32_t> father(parents.begin(), parents.end());
    Graph g;
    int roots = 0;
    Graph::from_forest_array(father, &g, &roots);
    return strategies_from_tree(g);
}

StrategyList gen_cross_strategies(const PeerList &peers, Strategy s)
{
    std::vector<int> masters, parents;
    peers.partition_by_host(&masters, &parents);
    const int n = (int)peers.size();
    StrategyList sl;
    if (s == Ring) {
        for (size_t r = 0; r < masters.size(); ++r) {
            StrategyGraphs sg;
            gen_sub_circular_graph_pair(n, masters, (int)r, &sg.reduce, &sg.bcast);
            sl.push_back(sg);
        }
        return sl;
    }
    // reduce graph of a sub tree only loops on the participating ranks
    Graph b = gen_sub_binary_tree(n, masters);
    Graph r = b.reverse();
    for (int m : masters) r.add_edge(m, m);
    sl.push_back({r, b});
    return sl;
}

std::vector<uint8_t> digest_bytes(const StrategyList &sl)
{
    std::vector<uint8_t> b;
    for (auto &s : sl) {
        auto x = s.reduce.digest_bytes();
        auto y = s.bcast.digest_bytes();
        b.insert(b.end(), x.begin(), x.end());
        b.insert(b.end(), y.begin(), y.end());
    }
    return b;
}

std::vector<int32_t> minimum_spanning_tree(const std::vector<float> &w, int n)
{
    std::vector<int32_t> edges;
    if (n <= 1) return edges;
    auto weight = [&](int i, int j) { return 0.5f * (w[i * n + j] + w[j * n + i]); };
    std::vector<bool> in(n, false);
    std::vector<float> best(n, std::numeric_limits<float>::infinity());
    std::vector<int> from(n, 0);
    in[0] = true;
    for (int j = 1; j < n; ++j) best[j] = weight(0, j);
    for (int step = 1; step < n; ++step) {
        int v = -1;
        for (int j = 0; j < n; ++j)
            if (!in[j] && (v < 0 || best[j] < best[v])) v = j;
        in[v] = true;
        edges.push_back(from[v]);
        edges.push_back(v);
        for (int j = 0; j < n; ++j) {
            if (!in[j] && weight(v, j) < best[j]) {
                best[j] = weight(v, j);
                from[j] = v;
            }
        }
    }
    return edges;
}

std::vector<int32_t> mst_to_forest(const std::vector<int32_t> &edges, int n)
{
    std::vector<int32_t> father(n);
    for (int i = 0; i < n; ++i) father[i] = i;
    for (size_t k = 0; k + 1 < edges.size(); k += 2) father[edges[k + 1]] = edges[k];
    return father;
}

// ---- device plan -----------------------------------------------------------------------
static bool graph_depths(const Graph &g, std::vector<int> *depth, std::string *err)
{
    const int n = g.size();
    depth->assign(n, -1);
    std::vector<int> state(n, 0);  // 0 new, 1 on the stack, 2 done
    // iterative longest-path-from-a-source; a cycle is an error (self loops are not edges here)
    for (int root = 0; root < n; ++root) {
        if (state[root] == 2) continue;
        std::vector<std::pair<int, size_t>> stack{{root, 0}};
        state[root] = 1;
        while (!stack.empty()) {
            auto &[v, i]   = stack.back();
            const auto &ps = g.prevs(v);
            if (i < ps.size()) {
                const int u = ps[i++];
                if (state[u] == 1) {
                    if (err) *err = "cycle in strategy graph";
                    return false;
                }
                if (state[u] == 0) {
                    state[u] = 1;
                    stack.push_back({u, 0});
                }
            } else {
                int d = 0;
                for (int u : ps) d = std::max(d, (*depth)[u] + 1);
                (*depth)[v] = d;
                state[v]    = 2;
                stack.pop_back();
            }
        }
    }
    return true;
}

bool compile_device_plan(const StrategyList &sl, int rank, DevicePlan *out, std::string *err)
{
    DevicePlan p;
    p.n_plans = (int)sl.size();
    if (sl.empty() || rank < 0 || rank >= sl[0].reduce.size()) {
        if (err) *err = "compile_device_plan: bad rank or empty strategy list";
        return false;
    }
    std::vector<std::vector<int>> rdepth(sl.size()), bdepth(sl.size());
    int max_r = 0, max_b = 0;
    for (size_t i = 0; i < sl.size(); ++i) {
        if (!graph_depths(sl[i].reduce, &rdepth[i], err) || !graph_depths(sl[i].bcast, &bdepth[i], err)) return false;
        for (int d : rdepth[i]) max_r = std::max(max_r, d);
        for (int d : bdepth[i]) max_b = std::max(max_b, d);
    }
    p.n_reduce_rounds = max_r;
    p.n_rounds        = max_r + max_b;
    p.round_begin.push_back(0);
    for (int rnd = 1; rnd <= max_r; ++rnd) {
        for (size_t i = 0; i < sl.size(); ++i) {
            // only nodes with a self loop combine what they receive (reduce graph semantics)
            if (rdepth[i][rank] != rnd || !sl[i].reduce.is_self_loop(rank)) continue;
            for (int src : sl[i].reduce.prevs(rank)) {
                p.edge_plan.push_back((int32_t)i);
                p.edge_src.push_back(src);
            }
        }
        p.round_begin.push_back((int32_t)p.edge_plan.size());
    }
    for (int rnd = 1; rnd <= max_b; ++rnd) {
        for (size_t i = 0; i < sl.size(); ++i) {
            if (bdepth[i][rank] != rnd) continue;
            for (int src : sl[i].bcast.prevs(rank)) {
                p.edge_plan.push_back((int32_t)i);
                p.edge_src.push_back(src);
            }
        }
        p.round_begin.push_back((int32_t)p.edge_plan.size());
    }
    *out = p;
    return true;
}

}  // namespace kf
