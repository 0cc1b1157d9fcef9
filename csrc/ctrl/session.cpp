#include "ctrl/session.hpp"

#include <atomic>
#include <cstring>
#include <thread>

#include "core/log.hpp"

namespace kf {

static constexpr size_t kChunkBytes           = 1 << 20;
static constexpr double kInterferenceThreshold = 0.8;

void StrategyStat::update(double begin, double end, uint64_t size)
{
    std::lock_guard<std::mutex> lk(mu);
    if (first_begin < 0) first_begin = begin;
    if (end > last_end) last_end = end;
    acc_size += size;
}

void StrategyStat::reset()
{
    acc_size    = 0;
    first_begin = -1;
    last_end    = 0;
}

Session::Session(const PeerList &peers, const PeerID &self, Strategy strategy, Client *client, Server *server)
    : peers_(peers), self_(self), strategy_(strategy), client_(client), server_(server)
{
    rank_       = peers_.rank_of(self);
    local_rank_ = peers_.local_rank_of(self);
    local_size_ = peers_.local_size_of(self);
    host_count_ = peers_.host_count();
    if (strategy_ == Auto) strategy_ = auto_select(peers_);
    local_ = gen_local_strategies(peers_);
    set_global(gen_global_strategies(peers_, strategy_));
    cross_        = gen_cross_strategies(peers_, strategy_);
    timeout_s_    = parse_duration_seconds(getenv_str("KUNGFU_CONFIG_OP_TIMEOUT"), 600);
    hash_by_name_ = getenv_str("KUNGFU_CONFIG_STRATEGY_HASH_METHOD", "NAME") != "SIMPLE";
}

void Session::set_global(const StrategyList &sl)
{
    global_ = sl;
    global_stats_.clear();
    for (size_t i = 0; i < sl.size(); ++i) global_stats_.emplace_back(new StrategyStat());
}

void Session::fail(const std::string &msg)
{
    std::lock_guard<std::mutex> lk(err_mu_);
    last_error_ = msg;
    KF_ERROR("session: %s", msg.c_str());
}

std::string Session::last_error()
{
    std::lock_guard<std::mutex> lk(err_mu_);
    return last_error_;
}

bool Session::send_to(int rank, const std::string &name, const void *data, size_t n)
{
    std::string err;
    if (!client_->send(peers_[rank], CONN_COLLECTIVE, name, FLAG_NONE, data, n, &err)) {
        fail(err);
        return false;
    }
    return true;
}

bool Session::recv_from(int rank, const std::string &name, void *buf, size_t n)
{
    const long got = server_->mailbox.recv_into(peers_[rank], name, buf, n, timeout_s_);
    if (got < 0) {
        fail("timed out waiting for '" + name + "' from rank " + std::to_string(rank));
        return false;
    }
    if ((size_t)got != n) {
        fail("'" + name + "' from rank " + std::to_string(rank) + " has " + std::to_string(got) +
             " bytes, expected " + std::to_string(n));
        return false;
    }
    return true;
}

// Walk the graphs in order for this rank (reference: session.go:218-286).
bool Session::run_graphs(const Workspace &w, const std::vector<const Graph *> &graphs)
{
    if (size() == 1) {
        w.forward();
        return true;
    }
    const size_t n = w.bytes();
    std::vector<uint8_t> tmp;
    int recv_count        = 0;
    const void *effective = w.send;
    for (const Graph *g : graphs) {
        const auto &prevs = g->prevs(rank_);
        const auto &nexts = g->nexts(rank_);
        if (g->is_self_loop(rank_)) {
            // Fan-in: the receives of all predecessors are posted together (the reference receives
            // them in parallel, session.go:242-251), the reduction then runs in predecessor order so
            // that the sum is formed in the same order on every run.
            const size_t np = prevs.size();
            if (tmp.size() < n * np) tmp.resize(n * np);
            std::vector<char> got(np, 1);
            if (np > 1) {
                std::vector<std::thread> rx;
                for (size_t i = 1; i < np; ++i)
                    rx.emplace_back([&, i] { got[i] = recv_from(prevs[i], w.name, tmp.data() + i * n, n) ? 1 : 0; });
                got[0] = recv_from(prevs[0], w.name, tmp.data(), n) ? 1 : 0;
                for (auto &t : rx) t.join();
            } else if (np == 1) {
                got[0] = recv_from(prevs[0], w.name, tmp.data(), n) ? 1 : 0;
            }
            for (size_t i = 0; i < np; ++i) {
                if (!got[i]) return false;
                transform2(effective, tmp.data() + i * n, w.recv, w.count, w.dtype, w.op);
                effective = w.recv;
                ++recv_count;
            }
            for (int q : nexts)
                if (!send_to(q, w.name, effective, n)) return false;
        } else {
            if (prevs.empty()) {
                if (recv_count == 0) w.forward();
            } else {
                for (int p : prevs) {
                    if (!recv_from(p, w.name, w.recv, n)) return false;
                    ++recv_count;
                }
            }
            for (int q : nexts)
                if (!send_to(q, w.name, w.recv, n)) return false;
        }
    }
    // a node that never received anything (e.g. the only worker of its host in
    // a local reduce) still owes the caller a result in recv
    if (recv_count == 0) w.forward();
    return true;
}

static size_t name_hash(const std::string &s)
{
    size_t h = 0;
    for (unsigned char c : s) h += (size_t)c * c;
    return h;
}

bool Session::run_strategies(const Workspace &w, const StrategyList &sl, bool monitored)
{
    const size_t bytes = w.bytes();
    const size_t k     = std::max<size_t>(1, (bytes + kChunkBytes - 1) / kChunkBytes);
    auto parts         = even_partition(w.count, k);
    auto run_chunk     = [&](size_t i) -> bool {
        const Workspace c = (k == 1) ? w : w.slice(parts[i].begin, parts[i].end);
        const size_t si   = (hash_by_name_ ? name_hash(c.name) : i) % sl.size();
        const double t0   = monitored ? now_seconds() : 0;
        const bool ok     = run_graphs(c, {&sl[si].reduce, &sl[si].bcast});
        if (monitored && &sl == &global_ && si < global_stats_.size())
            global_stats_[si]->update(t0, now_seconds(), c.bytes());
        return ok;
    };
    if (k == 1) return run_chunk(0);
    std::atomic<size_t> next{0};
    std::atomic<bool> ok{true};
    const size_t nthreads = std::min<size_t>(k, 8);
    std::vector<std::thread> ths;
    for (size_t t = 0; t < nthreads; ++t) {
        ths.emplace_back([&] {
            for (;;) {
                const size_t i = next.fetch_add(1);
                if (i >= k) return;
                if (!run_chunk(i)) ok = false;
            }
        });
    }
    for (auto &t : ths) t.join();
    return ok;
}

bool Session::all_reduce(const Workspace &w) { return run_strategies(w, global_, false); }
bool Session::monitored_all_reduce(const Workspace &w) { return run_strategies(w, global_, true); }
bool Session::cross_all_reduce(const Workspace &w) { return run_strategies(w, cross_, false); }

bool Session::all_reduce_with(const std::vector<int32_t> &forest, const Workspace &w)
{
    if (forest.empty()) return all_reduce(w);
    Graph g;
    int roots = 0;
    if ((int)forest.size() != size() || !Graph::from_forest_array(forest, &g, &roots) || roots != 1) {
        fail("all_reduce_with: the tree must be a father array of size np with one root");
        return false;
    }
    return run_strategies(w, strategies_from_tree(g), false);
}

bool Session::reduce(const Workspace &w)
{
    const auto &s = global_[0];
    return run_graphs(w, {&s.reduce});
}

bool Session::broadcast(const Workspace &w)
{
    const auto &s = global_[0];
    return run_graphs(w, {&s.bcast});
}

bool Session::local_reduce(const Workspace &w) { return run_graphs(w, {&local_[0].reduce}); }
bool Session::local_broadcast(const Workspace &w) { return run_graphs(w, {&local_[0].bcast}); }

bool Session::barrier()
{
    const int n = size();
    std::vector<uint8_t> a(n, 1), b(n, 0);
    Workspace w;
    w.send  = a.data();
    w.recv  = b.data();
    w.count = (size_t)n;
    w.dtype = U8;
    w.op    = SUM;
    w.name  = "kungfu::barrier";
    StallDetector sd("barrier");
    return all_reduce(w);
}

bool Session::bytes_consensus(const void *data, size_t n, const std::string &name, bool *agree)
{
    *agree = false;
    {
        int32_t len = (int32_t)n, lo = 0, hi = 0;
        Workspace w;
        w.send  = &len;
        w.count = 1;
        w.dtype = I32;
        w.recv  = &lo;
        w.op    = MIN;
        w.name  = "consensus::len::min::" + name;
        if (!all_reduce(w)) return false;
        w.recv = &hi;
        w.op   = MAX;
        w.name = "consensus::len::max::" + name;
        if (!all_reduce(w)) return false;
        if (lo != hi) return true;
    }
    if (n == 0) {
        *agree = true;
        return true;
    }
    std::vector<uint8_t> lo(n), hi(n);
    Workspace w;
    w.send  = data;
    w.count = n;
    w.dtype = U8;
    w.recv  = lo.data();
    w.op    = MIN;
    w.name  = "consensus::min::" + name;
    if (!all_reduce(w)) return false;
    w.recv = hi.data();
    w.op   = MAX;
    w.name = "consensus::max::" + name;
    if (!all_reduce(w)) return false;
    *agree = (memcmp(lo.data(), hi.data(), n) == 0);
    return true;
}

bool Session::gather(const Workspace &w)
{
    const size_t n = w.bytes();
    if (rank_ != 0) return send_to(0, w.name, w.send, n);
    char *out = static_cast<char *>(w.recv);
    for (int r = 0; r < size(); ++r) {
        if (r == rank_) {
            memcpy(out + (size_t)r * n, w.send, n);
        } else if (!recv_from(r, w.name, out + (size_t)r * n, n)) {
            return false;
        }
    }
    return true;
}

bool Session::all_gather(const Workspace &w)
{
    const size_t n = w.bytes();
    char *out      = static_cast<char *>(w.recv);
    std::atomic<bool> ok{true};
    std::thread sender([&] {
        for (int k = 1; k < size(); ++k)
            if (!send_to((rank_ + k) % size(), w.name, w.send, n)) ok = false;
    });
    memcpy(out + (size_t)rank_ * n, w.send, n);
    for (int k = 1; k < size(); ++k) {
        const int r = (rank_ - k + size()) % size();
        if (!recv_from(r, w.name, out + (size_t)r * n, n)) ok = false;
    }
    sender.join();
    return ok;
}

bool Session::set_global_strategy(const StrategyList &sl)
{
    const auto digest = digest_bytes(sl);
    bool agree        = false;
    if (!bytes_consensus(digest.data(), digest.size(), "kungfu::set-strategy", &agree)) return false;
    if (!agree) {
        fail("set_global_strategy: ranks proposed different graphs");
        return false;
    }
    set_global(sl);
    return true;
}

bool Session::set_tree(const std::vector<int32_t> &forest)
{
    Graph g;
    int roots = 0;
    if ((int)forest.size() != size() || !Graph::from_forest_array(forest, &g, &roots) || roots != 1) {
        fail("set_tree: need a father array of size np with exactly one root");
        return false;
    }
    int root = 0;
    for (int i = 0; i < size(); ++i)
        if (forest[i] == i) root = i;
    if (!g.is_valid_tree_with_root(root)) {
        fail("set_tree: the father array contains a cycle");
        return false;
    }
    return set_global_strategy(strategies_from_tree(g));
}

bool Session::set_strategy(Strategy s)
{
    if (!set_global_strategy(gen_global_strategies(peers_, s))) return false;
    strategy_ = s == Auto ? auto_select(peers_) : s;
    return true;
}

void Session::calc_stats()
{
    if (global_.size() != 1) {
        KF_ERROR("calc_stats needs exactly one active strategy (have %zu)", global_.size());
        return;
    }
    StrategyStat &st = *global_stats_[0];
    std::lock_guard<std::mutex> lk(st.mu);
    if (st.acc_size == 0) return;
    const double dt = st.last_end - st.first_begin;
    st.throughput   = dt > 0 ? (double)st.acc_size / dt : 0;
    st.reset();
}

void Session::log_stats()
{
    if (global_.size() != 1) {
        KF_ERROR("log_stats needs exactly one active strategy");
        return;
    }
    stat_log_.push_back(global_stats_[0]->throughput);
}

void Session::print_strategy_stats()
{
    KF_INFO("strategy stats: %zu active strateg%s", global_.size(), global_.size() == 1 ? "y" : "ies");
    for (size_t i = 0; i < stat_log_.size(); ++i)
        KF_INFO("global step #%zu throughput=%s", i, show_rate(stat_log_[i]).c_str());
}

bool Session::check_interference()
{
    if (global_.size() != 1) {
        KF_ERROR("check_interference needs exactly one active strategy");
        return false;
    }
    StrategyStat &st = *global_stats_[0];
    if (st.reff_throughput == 0) {
        st.reff_throughput = st.throughput;
        if (rank_ == 0) KF_INFO("adaptation: reference throughput = %s", show_rate(st.reff_throughput).c_str());
        return false;
    }
    int32_t vote = st.throughput < kInterferenceThreshold * st.reff_throughput ? 1 : 0, total = 0;
    Workspace w;
    w.send  = &vote;
    w.recv  = &total;
    w.count = 1;
    w.dtype = I32;
    w.op    = SUM;
    w.name  = "kungfu::strategy-monitor";
    if (!all_reduce(w)) return false;
    if (rank_ == 0)
        KF_INFO("adaptation: throughput %s vs reference %s, %d/%d votes to switch", show_rate(st.throughput).c_str(),
                show_rate(st.reff_throughput).c_str(), total, size());
    return total > size() / 2;
}

std::string Session::strategies_text() const
{
    std::string text;
    for (auto &sg : global_) text += "reduce=" + sg.reduce.debug_string() + " bcast=" + sg.bcast.debug_string() + "\n";
    return text;
}

std::vector<float> Session::peer_latencies()
{
    std::vector<float> out(size(), 0.f);
    for (int r = 0; r < size(); ++r) {
        if (r == rank_) continue;
        out[r] = (float)client_->ping(peers_[r]);
    }
    return out;
}

std::vector<float> Session::egress_rates(NetMonitor *m)
{
    std::vector<float> out(size(), 0.f);
    if (!m) return out;
    m->sample();
    for (int r = 0; r < size(); ++r) out[r] = (float)m->egress_rate(peers_[r]);
    return out;
}

}  // namespace kf
