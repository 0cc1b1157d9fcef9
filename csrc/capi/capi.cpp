// extern "C" surface of the host runtime (see include/kungfu_host.h).
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/kungfu_host.h"
#include "core/base.hpp"
#include "core/graph.hpp"
#include "core/log.hpp"
#include "core/plan.hpp"
#include "ctrl/peer.hpp"
#include "ctrl/transport.hpp"
#include "elastic/runner.hpp"

using namespace kf;

struct kf_peer {
    std::unique_ptr<Peer> peer;
};

static thread_local std::string g_err;

static int fail(const std::string &e)
{
    g_err = e;
    return -1;
}

static int done(kf_peer *p, bool ok)
{
    if (ok) return 0;
    g_err = p->peer->last_error();
    return -1;
}

static Workspace ws(const void *send, void *recv, int64_t count, int dtype, int op, const char *name)
{
    Workspace w;
    w.send  = send;
    w.recv  = recv;
    w.count = (size_t)count;
    w.dtype = (uint32_t)dtype;
    w.op    = (Op)op;
    w.name  = name ? name : "";
    return w;
}

static int copy_out(const std::string &s, char *out, int cap)
{
    if ((int)s.size() + 1 > cap) return fail("output buffer too small");
    memcpy(out, s.c_str(), s.size() + 1);
    return (int)s.size();
}

extern "C" {

const char *kf_last_error(void) { return g_err.c_str(); }

kf_peer *kf_peer_create_from_env(void)
{
    PeerConfig cfg;
    std::string err;
    if (!PeerConfig::from_env(&cfg, &err)) {
        g_err = err;
        return nullptr;
    }
    kf_peer *p = new kf_peer();
    p->peer.reset(new Peer(cfg));
    return p;
}

int kf_peer_start(kf_peer *p) { return done(p, p->peer->start()); }

void kf_peer_close(kf_peer *p)
{
    if (!p) return;
    p->peer->close();
    delete p;
}

int kf_rank(kf_peer *p) { return p->peer->session()->rank(); }
int kf_size(kf_peer *p) { return p->peer->session()->size(); }
int kf_local_rank(kf_peer *p) { return p->peer->session()->local_rank(); }
int kf_local_size(kf_peer *p) { return p->peer->session()->local_size(); }
int kf_host_count(kf_peer *p) { return p->peer->session()->host_count(); }
int kf_detached(kf_peer *p) { return p->peer->detached() ? 1 : 0; }
int kf_cluster_version(kf_peer *p) { return p->peer->cluster_version(); }
uint64_t kf_uid(kf_peer *p) { return p->peer->uid(); }

int kf_peer_addr(kf_peer *p, int rank, char *buf, int cap)
{
    Session *s = p->peer->session();
    if (rank < 0 || rank >= s->size()) return fail("bad rank");
    return copy_out(s->peers()[rank].str(), buf, cap) < 0 ? -1 : 0;
}

int kf_barrier(kf_peer *p) { return done(p, p->peer->session()->barrier()); }

#define KF_CHECK_DTYPE(dt)                                                     \
    if (!dtype_valid((uint32_t)(dt))) return fail("unsupported dtype code " + std::to_string(dt))

int kf_all_reduce(kf_peer *p, const void *send, void *recv, int64_t count, int dtype, int op, const char *name)
{
    KF_CHECK_DTYPE(dtype);
    StallDetector sd(std::string("all_reduce ") + (name ? name : ""));
    return done(p, p->peer->session()->all_reduce(ws(send, recv, count, dtype, op, name)));
}

int kf_monitored_all_reduce(kf_peer *p, const void *send, void *recv, int64_t count, int dtype, int op,
                            const int32_t *tree, int tree_len, const char *name)
{
    KF_CHECK_DTYPE(dtype);
    Session *s = p->peer->session();
    if (tree && tree_len > 0) {
        std::vector<int32_t> forest(tree, tree + tree_len);
        if (!s->set_tree(forest)) return done(p, false);
    }
    return done(p, s->monitored_all_reduce(ws(send, recv, count, dtype, op, name)));
}

int kf_all_reduce_with(kf_peer *p, const int32_t *tree, int tree_len, const void *send, void *recv, int64_t count,
                       int dtype, int op, const char *name)
{
    KF_CHECK_DTYPE(dtype);
    std::vector<int32_t> forest;
    if (tree && tree_len > 0) forest.assign(tree, tree + tree_len);
    return done(p, p->peer->session()->all_reduce_with(forest, ws(send, recv, count, dtype, op, name)));
}

int kf_cross_all_reduce(kf_peer *p, const void *send, void *recv, int64_t count, int dtype, int op, const char *name)
{
    KF_CHECK_DTYPE(dtype);
    return done(p, p->peer->session()->cross_all_reduce(ws(send, recv, count, dtype, op, name)));
}

int kf_reduce(kf_peer *p, const void *send, void *recv, int64_t count, int dtype, int op, const char *name)
{
    KF_CHECK_DTYPE(dtype);
    return done(p, p->peer->session()->reduce(ws(send, recv, count, dtype, op, name)));
}

int kf_broadcast(kf_peer *p, const void *send, void *recv, int64_t count, int dtype, const char *name)
{
    KF_CHECK_DTYPE(dtype);
    return done(p, p->peer->session()->broadcast(ws(send, recv, count, dtype, SUM, name)));
}

int kf_local_reduce(kf_peer *p, const void *send, void *recv, int64_t count, int dtype, int op, const char *name)
{
    KF_CHECK_DTYPE(dtype);
    return done(p, p->peer->session()->local_reduce(ws(send, recv, count, dtype, op, name)));
}

int kf_local_broadcast(kf_peer *p, const void *send, void *recv, int64_t count, int dtype, const char *name)
{
    KF_CHECK_DTYPE(dtype);
    return done(p, p->peer->session()->local_broadcast(ws(send, recv, count, dtype, SUM, name)));
}

int kf_all_gather(kf_peer *p, const void *send, void *recv, int64_t count, int dtype, const char *name)
{
    KF_CHECK_DTYPE(dtype);
    return done(p, p->peer->session()->all_gather(ws(send, recv, count, dtype, SUM, name)));
}

int kf_gather(kf_peer *p, const void *send, void *recv, int64_t count, int dtype, const char *name)
{
    KF_CHECK_DTYPE(dtype);
    return done(p, p->peer->session()->gather(ws(send, recv, count, dtype, SUM, name)));
}

int kf_consensus(kf_peer *p, const void *data, int64_t nbytes, const char *name, int *agree)
{
    bool a        = false;
    const bool ok = p->peer->session()->bytes_consensus(data, (size_t)nbytes, name ? name : "", &a);
    if (agree) *agree = a ? 1 : 0;
    return done(p, ok);
}

int64_t kf_all_reduce_async(kf_peer *p, const void *send, void *recv, int64_t count, int dtype, int op, const char *name)
{
    if (!dtype_valid((uint32_t)dtype)) return fail("unsupported dtype");
    Workspace w = ws(send, recv, count, dtype, op, name);
    Peer *peer  = p->peer.get();
    return peer->submit([peer, w] { return peer->session()->all_reduce(w); });
}

int64_t kf_broadcast_async(kf_peer *p, const void *send, void *recv, int64_t count, int dtype, const char *name)
{
    if (!dtype_valid((uint32_t)dtype)) return fail("unsupported dtype");
    Workspace w = ws(send, recv, count, dtype, SUM, name);
    Peer *peer  = p->peer.get();
    return peer->submit([peer, w] { return peer->session()->broadcast(w); });
}

int kf_wait_handle(kf_peer *p, int64_t handle) { return done(p, p->peer->wait(handle)); }

int kf_submit(kf_peer *p, int (*fn)(void *), void *arg, void (*on_done)(void *, int))
{
    if (!p || !fn) return fail("kf_submit: null argument");
    p->peer->submit_detached([fn, arg, on_done] {
        const int rc = fn(arg);
        if (on_done) on_done(arg, rc);
    });
    return 0;
}

int kf_save(kf_peer *p, const char *version, const char *name, const void *data, int64_t nbytes)
{
    p->peer->save(version ? version : "", name, data, (size_t)nbytes);
    return 0;
}

int kf_request(kf_peer *p, int target_rank, const char *version, const char *name, void *buf, int64_t nbytes)
{
    return done(p, p->peer->request(target_rank, version ? version : "", name, buf, (size_t)nbytes));
}

int kf_resize_cluster(kf_peer *p, int new_size, int *changed, int *detached)
{
    bool c = false, d = false;
    const bool ok = p->peer->resize_cluster(new_size, &c, &d);
    *changed  = c;
    *detached = d;
    return done(p, ok);
}

int kf_peek_cluster_change(kf_peer *p, int *differs)
{
    bool d        = false;
    const bool ok = p->peer->peek_cluster_change(&d);
    if (differs) *differs = d ? 1 : 0;
    return done(p, ok);
}

int kf_resize_cluster_from_url(kf_peer *p, int *changed, int *detached)
{
    bool c = false, d = false;
    const bool ok = p->peer->resize_cluster_from_url(&c, &d);
    *changed  = c;
    *detached = d;
    return done(p, ok);
}

int kf_propose_new_size(kf_peer *p, int new_size) { return done(p, p->peer->propose_new_size(new_size)); }

int kf_set_tree(kf_peer *p, const int32_t *forest, int n)
{
    return done(p, p->peer->session()->set_tree(std::vector<int32_t>(forest, forest + n)));
}

int kf_set_strategy(kf_peer *p, const char *name)
{
    Strategy s;
    if (!parse_strategy(name, &s)) return fail(std::string("unknown strategy ") + name);
    return done(p, p->peer->session()->set_strategy(s));
}

int kf_calc_stats(kf_peer *p)
{
    p->peer->session()->calc_stats();
    return 0;
}
int kf_log_stats(kf_peer *p)
{
    p->peer->session()->log_stats();
    return 0;
}
int kf_print_strategy_stats(kf_peer *p)
{
    p->peer->session()->print_strategy_stats();
    return 0;
}
int kf_check_interference(kf_peer *p) { return p->peer->session()->check_interference() ? 1 : 0; }

int kf_get_peer_latencies(kf_peer *p, float *out, int n)
{
    auto v = p->peer->session()->peer_latencies();
    for (int i = 0; i < n && i < (int)v.size(); ++i) out[i] = v[i];
    return 0;
}

int kf_egress_rates(kf_peer *p, float *out, int n)
{
    auto v = p->peer->egress_rates();
    for (int i = 0; i < n && i < (int)v.size(); ++i) out[i] = v[i];
    return 0;
}

int64_t kf_request_async(kf_peer *p, int target_rank, const char *version, const char *name, void *buf, int64_t nbytes)
{
    if (!p || !name || !buf || nbytes < 0) return fail("kf_request_async: bad argument");
    Peer *peer = p->peer.get();
    const std::string v = version ? version : "", n = name;
    return peer->submit([peer, target_rank, v, n, buf, nbytes] { return peer->request(target_rank, v, n, buf, (size_t)nbytes); });
}

struct kf_prefetch_state {
    std::mutex mu;
    std::condition_variable cv;
    size_t nbytes = 0;  // of either buffer; fixed (the vectors themselves are swapped under mu)
    std::vector<uint8_t> model, incoming;
    bool have = false, requesting = false;
    int64_t fresh = 0;
};

struct kf_prefetch {
    kf_peer *peer;
    std::string name;
    std::shared_ptr<kf_prefetch_state> st;
};

kf_prefetch *kf_prefetch_new(kf_peer *p, const char *name, int64_t nbytes)
{
    if (!p || !name || nbytes <= 0) {
        fail("kf_prefetch_new: bad argument");
        return nullptr;
    }
    auto *pf = new kf_prefetch{p, name, std::make_shared<kf_prefetch_state>()};
    pf->st->nbytes = (size_t)nbytes;
    pf->st->model.resize((size_t)nbytes);
    pf->st->incoming.resize((size_t)nbytes);
    return pf;
}

// First call: blocking pull. Every call: start the pull for the next one unless one is in flight.
static bool prefetch_acquire(kf_prefetch *pf, int target_rank)
{
    auto st    = pf->st;
    Peer *peer = pf->peer->peer.get();
    const size_t n = st->nbytes;
    if (!st->have) {
        if (!peer->request(target_rank, "", pf->name, st->model.data(), n)) return false;
        st->have = true;
    }
    bool start = false;
    {
        std::lock_guard<std::mutex> lk(st->mu);
        if (!st->requesting) st->requesting = start = true;
    }
    if (start) {
        const std::string name = pf->name;
        // the job owns a reference to the buffers: kf_prefetch_free may run before it does
        peer->submit_detached([peer, st, name, target_rank, n] {
            const bool ok = peer->request(target_rank, "", name, st->incoming.data(), n);
            std::lock_guard<std::mutex> lk(st->mu);
            if (ok) {
                st->model.swap(st->incoming);
                ++st->fresh;
            }
            st->requesting = false;
            st->cv.notify_all();
        });
    }
    return true;
}

int kf_prefetch_next(kf_prefetch *pf, int target_rank, void *out, int64_t *fresh)
{
    if (!pf || !out) return fail("kf_prefetch_next: null argument");
    if (!prefetch_acquire(pf, target_rank)) return done(pf->peer, false);
    auto st = pf->st;
    std::lock_guard<std::mutex> lk(st->mu);
    memcpy(out, st->model.data(), st->nbytes);
    if (fresh) *fresh = st->fresh;
    return 0;
}

int kf_save_parts(kf_peer *p, const char *version, const char *name, const void *const *parts, const int64_t *nbytes, int n)
{
    if (!p || !name || n < 0 || (n > 0 && (!parts || !nbytes))) return fail("kf_save_parts: bad argument");
    size_t total = 0;
    for (int i = 0; i < n; ++i) {
        if (nbytes[i] < 0 || (nbytes[i] > 0 && !parts[i])) return fail("kf_save_parts: bad part");
        total += (size_t)nbytes[i];
    }
    auto blob = new_blob(total);  // recycled storage when a previous version of this size has been released
    size_t off = 0;
    for (int i = 0; i < n; ++i) {
        if (nbytes[i]) memcpy(blob->data() + off, parts[i], (size_t)nbytes[i]);
        off += (size_t)nbytes[i];
    }
    p->peer->save_blob(version ? version : "", name, std::move(blob));
    return 0;
}

__attribute__((optimize("O3", "tree-vectorize"))) static void blend_half(float *t, const float *o, int64_t n)
{
#pragma GCC ivdep
    for (int64_t i = 0; i < n; ++i) t[i] = 0.5f * (t[i] + o[i]);
}

void kf_blend_f32(float *const *tensors, const int64_t *counts, int n, const float *other)
{
    int64_t off = 0;
    for (int i = 0; i < n; ++i) {
        blend_half(tensors[i], other + off, counts[i]);
        off += counts[i];
    }
}

int kf_prefetch_blend_f32(kf_prefetch *pf, int target_rank, float *const *tensors, const int64_t *counts, int n, int64_t *fresh)
{
    if (!pf || n < 0 || (n > 0 && (!tensors || !counts))) return fail("kf_prefetch_blend_f32: bad argument");
    int64_t total = 0;
    for (int i = 0; i < n; ++i) total += counts[i];
    if ((size_t)total * sizeof(float) != pf->st->nbytes) return fail("kf_prefetch_blend_f32: tensors do not add up to the model size");
    if (!prefetch_acquire(pf, target_rank)) return done(pf->peer, false);
    auto st = pf->st;
    std::lock_guard<std::mutex> lk(st->mu);  // the background pull swaps buffers under this lock
    kf_blend_f32(tensors, counts, n, (const float *)st->model.data());
    if (fresh) *fresh = st->fresh;
    return 0;
}

void kf_prefetch_free(kf_prefetch *pf)
{
    if (!pf) return;
    {
        std::unique_lock<std::mutex> lk(pf->st->mu);
        pf->st->cv.wait_for(lk, std::chrono::seconds(30), [&] { return !pf->st->requesting; });
    }
    delete pf;
}

int kf_send_fd(kf_peer *p, int target_rank, const char *tag, int fd)
{
    return done(p, p->peer->send_fd(target_rank, tag, fd));
}

int kf_recv_fd(kf_peer *p, int src_rank, const char *tag, int *fd) { return done(p, p->peer->recv_fd(src_rank, tag, fd)); }

int kf_current_strategies(kf_peer *p, char *out, int cap)
{
    return copy_out(p->peer->session()->strategies_text(), out, cap);
}

int kf_current_strategy_name(kf_peer *p, char *out, int cap)
{
    return copy_out(strategy_name(p->peer->session()->strategy()), out, cap);
}

// ---- planner utilities ---------------------------------------------------------------
int kf_gen_strategy_graphs(const char *peers_csv, const char *strategy, char *out, int cap)
{
    PeerList peers;
    Strategy s;
    if (!parse_peer_list(peers_csv, &peers)) return fail("bad peer list");
    if (!parse_strategy(strategy, &s)) return fail("unknown strategy");
    std::string text;
    for (auto &sg : gen_global_strategies(peers, s))
        text += "reduce=" + sg.reduce.debug_string() + " bcast=" + sg.bcast.debug_string() + "\n";
    return copy_out(text, out, cap);
}

static int plan_out(const StrategyList &sl, int rank, int32_t *out, int cap)
{
    DevicePlan p;
    std::string err;
    if (!compile_device_plan(sl, rank, &p, &err)) return fail(err);
    const size_t need = 4 + p.round_begin.size() + 2 * p.edge_plan.size();
    if ((size_t)cap < need) return fail("device plan does not fit the caller's buffer");
    size_t k = 0;
    out[k++] = p.n_plans;
    out[k++] = p.n_rounds;
    out[k++] = p.n_reduce_rounds;
    out[k++] = (int32_t)p.edge_plan.size();
    for (int32_t v : p.round_begin) out[k++] = v;
    for (int32_t v : p.edge_plan) out[k++] = v;
    for (int32_t v : p.edge_src) out[k++] = v;
    return (int)k;
}

int kf_device_plan(kf_peer *p, int32_t *out, int cap)
{
    return plan_out(p->peer->session()->global_strategies(), p->peer->session()->rank(), out, cap);
}

int kf_device_plan_for(const char *peers_csv, const char *strategy, int rank, int32_t *out, int cap)
{
    PeerList peers;
    Strategy s;
    if (!parse_peer_list(peers_csv, &peers)) return fail("bad peer list");
    if (!parse_strategy(strategy, &s)) return fail("unknown strategy");
    return plan_out(gen_global_strategies(peers, s), rank, out, cap);
}

int kf_minimum_spanning_tree(const float *weights, int n, int32_t *edges_out)
{
    std::vector<float> w(weights, weights + (size_t)n * n);
    auto e = minimum_spanning_tree(w, n);
    for (size_t i = 0; i < e.size(); ++i) edges_out[i] = e[i];
    return (int)e.size() / 2;
}

void kf_transform2(const void *x, const void *y, void *z, int64_t n, int dtype, int op)
{
    transform2(x, y, z, (size_t)n, (uint32_t)dtype, (Op)op);
}

int kf_cluster_resize(const char *cluster_json, int new_size, char *out, int cap)
{
    Cluster c, next;
    std::string err;
    if (!parse_cluster_json(cluster_json, &c, &err)) return fail(err);
    if (!c.validate(&err)) return fail(err);
    if (!c.resize(new_size, &next, &err)) return fail(err);
    return copy_out(next.json(), out, cap);
}

int kf_parse_hostfile(const char *text, char *out, int cap)
{
    HostList hl;
    std::string err;
    if (!parse_hostfile(text, &hl, &err)) return fail(err);
    return copy_out(hl.str(), out, cap);
}

int kf_gen_peer_list(const char *hostlist, int np, const char *port_range, char *out, int cap)
{
    HostList hl;
    PortRange pr;
    PeerList pl;
    std::string err;
    if (!parse_host_list(hostlist, &hl)) return fail("bad host list");
    if (port_range && *port_range && !parse_port_range(port_range, &pr)) return fail("bad port range");
    if (!hl.gen_peer_list(np, pr, &pl, &err)) return fail(err);
    return copy_out(pl.str(), out, cap);
}

int kf_run_main(int argc, char **argv) { return kungfu_run_main(argc, argv); }
int kf_config_server_main(int argc, char **argv) { return config_server_main(argc, argv); }

}  // extern "C"
