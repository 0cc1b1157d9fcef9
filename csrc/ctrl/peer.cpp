#include "ctrl/peer.hpp"

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <unistd.h>

#include "core/log.hpp"
#include "core/platforms.hpp"

namespace kf {

bool PeerConfig::from_env(PeerConfig *out, std::string *err)
{
    PeerConfig c;
    log_set_level_from_env();
    c.monitoring        = getenv_bool("KUNGFU_CONFIG_ENABLE_MONITORING");
    c.monitoring_period = parse_duration_seconds(getenv_str("KUNGFU_CONFIG_MONITORING_PERIOD"), 1.0);
    c.config_server     = getenv_str(env::kConfigServer);
    const std::string strategy = getenv_str(env::kStrategy, "AUTO");
    if (!parse_strategy(strategy, &c.strategy)) {
        if (err) *err = "unknown strategy '" + strategy + "'";
        return false;
    }
    const std::string self = getenv_str(env::kSelfSpec);
    if (self.empty()) {
        // a hosting platform may describe the cluster instead of kungfu-run
        PeerList platform_peers;
        int task        = 0;
        std::string perr;
        const int found = modelarts_cluster(&platform_peers, &task, &perr);
        if (found < 0) {
            if (err) *err = "ModelArts environment: " + perr;
            return false;
        }
        if (found > 0 && platform_peers.size() > 1) {
            c.single     = false;
            c.init_peers = platform_peers;
            c.self       = platform_peers[task];
            *out         = c;
            return true;
        }
        // not launched by kungfu-run: one process, rank 0 of 1
        c.single = true;
        parse_peer_id("127.0.0.1:0", &c.self);
        c.init_peers.push_back(c.self);
        *out = c;
        return true;
    }
    c.single = false;
    if (!parse_peer_id(self, &c.self)) {
        if (err) *err = std::string("bad ") + env::kSelfSpec + " '" + self + "'";
        return false;
    }
    if (!parse_peer_list(getenv_str(env::kInitPeers), &c.init_peers)) {
        if (err) *err = std::string("bad ") + env::kInitPeers;
        return false;
    }
    if (!parse_peer_list(getenv_str(env::kInitRunners), &c.init_runners)) {
        if (err) *err = std::string("bad ") + env::kInitRunners;
        return false;
    }
    const std::string parent = getenv_str(env::kParentID);
    if (!parent.empty() && !parse_peer_id(parent, &c.parent)) {
        if (err) *err = std::string("bad ") + env::kParentID;
        return false;
    }
    c.init_version = atoi(getenv_str(env::kInitVersion, "0").c_str());
    if (!c.init_peers.contains(c.self)) {
        if (err) *err = "self " + c.self.str() + " is not in the initial peer list";
        return false;
    }
    *out = c;
    return true;
}

Peer::Peer(const PeerConfig &cfg) : cfg_(cfg), tcfg_(TransportConfig::from_env())
{
    cluster_.runners = cfg_.init_runners;
    cluster_.workers = cfg_.init_peers;
    version_         = cfg_.init_version;
}

Peer::~Peer() { close(); }

void Peer::set_error(const std::string &e)
{
    {
        std::lock_guard<std::mutex> lk(err_mu_);
        last_error_ = e;
    }
    KF_ERROR("%s", e.c_str());
}

std::string Peer::last_error()
{
    std::string e;
    {
        std::lock_guard<std::mutex> lk(err_mu_);
        e = last_error_;
    }
    if (session_) {
        const std::string s = session_->last_error();
        if (!s.empty()) e += (e.empty() ? "" : "; ") + s;
    }
    return e;
}

bool Peer::start()
{
    if (started_) return true;
    client_.reset(new Client(cfg_.self, tcfg_));
    server_.reset(new Server(cfg_.self, tcfg_.use_unix_sock));
    client_->monitor = cfg_.monitoring ? &monitor_ : nullptr;
    server_->monitor = cfg_.monitoring ? &monitor_ : nullptr;
    server_->on_control = [](const PeerID &, const std::string &name, const std::string &) {
        if (name == "exit") {
            KF_WARN("exit requested over the control channel");
            _exit(0);
        }
    };
    if (!cfg_.single) {
        std::string err;
        if (!server_->start(&err)) {
            set_error("cannot start server on " + cfg_.self.str() + ": " + err);
            return false;
        }
        if (cfg_.monitoring) {
            std::string merr;
            const uint16_t mport = (uint16_t)(cfg_.self.port + 10000);
            if (metrics_.start(
                    mport,
                    [this](const HttpRequest &rq, HttpResponse *rs) {
                        if (rq.path == "/metrics") rs->body = monitor_.metrics_text();
                        else rs->status = 404;
                    },
                    &merr)) {
                monitor_thread_ = std::thread([this] {
                    while (!closing_) {
                        monitor_.sample();
                        // sleep in slices so close() never waits a whole period
                        for (double left = cfg_.monitoring_period; left > 0 && !closing_; left -= 0.05)
                            std::this_thread::sleep_for(std::chrono::duration<double>(std::min(left, 0.05)));
                    }
                });
            } else {
                KF_WARN("metrics endpoint disabled: %s", merr.c_str());
            }
        }
    }
    for (int i = 0; i < 4; ++i) async_workers_.emplace_back([this] { async_loop(); });
    started_ = true;
    return update_to(cluster_.workers);
}

void Peer::close()
{
    if (closing_.exchange(true)) return;
    {
        std::lock_guard<std::mutex> lk(async_mu_);
        async_stop_ = true;
    }
    async_cv_.notify_all();
    for (auto &t : async_workers_)
        if (t.joinable()) t.join();
    if (monitor_thread_.joinable()) monitor_thread_.join();
    metrics_.stop();
    if (client_) client_->reset();
    if (server_) server_->stop();
}

Session *Peer::session()
{
    std::lock_guard<std::mutex> lk(mu_);
    return session_.get();
}

Cluster Peer::cluster()
{
    std::lock_guard<std::mutex> lk(mu_);
    return cluster_;
}

uint64_t Peer::uid() const
{
    return ((uint64_t)cfg_.self.ipv4 << 32) | ((uint64_t)cfg_.self.port << 16) | (uint64_t)(cfg_.init_version & 0xffff);
}

// New membership: new token (= version), fresh connections, fresh session,
// then a barrier that returns once every member (including just-spawned ones)
// has reached the same point (/root/reference/srcs/go/kungfu/peer/peer.go:144-166).
bool Peer::update_to(const PeerList &workers)
{
    StallDetector sd("peer update");
    // Order matters: drop what belongs to the previous epoch while the server still answers with the OLD
    // token (members already in the new epoch are refused and retry), and only then open the new epoch.
    // The other way round a fast peer's first message of the new epoch could land between the two steps
    // and be discarded, leaving everybody in the barrier below forever (seen once in ~5 runs under TSAN).
    client_->reset();
    server_->mailbox.reset();
    client_->set_token((uint32_t)version_);
    server_->set_token((uint32_t)version_);
    {
        std::lock_guard<std::mutex> lk(mu_);
        session_.reset(new Session(workers, cfg_.self, cfg_.strategy, client_.get(), server_.get()));
    }
    if (!session_->barrier()) {
        set_error("barrier after update failed");
        return false;
    }
    return true;
}

bool Peer::notify_runners(const Stage &stage)
{
    const std::string body = stage.json();
    for (auto &r : stage.cluster.runners) {
        if (!client_->wait(r, tcfg_.wait_runner_timeout)) {
            set_error("runner " + r.str() + " is unreachable");
            return false;
        }
        std::string err;
        if (!client_->send(r, CONN_CONTROL, "update", FLAG_NONE, body.data(), body.size(), &err)) {
            set_error("cannot notify runner " + r.str() + ": " + err);
            return false;
        }
    }
    return true;
}

bool Peer::propose(const Cluster &next, bool *changed, bool *detached)
{
    StallDetector sd("propose");
    *changed  = false;
    *detached = false;
    Cluster cur = cluster();
    if (cur == next) return true;
    if (cur.workers.disjoint(next.workers))
        KF_ERROR("full update detected: %s -> %s, state will be lost", cur.workers.str().c_str(), next.workers.str().c_str());
    else if (!next.workers.empty() && !cur.workers.contains(next.workers[0]))
        KF_ERROR("new root can't be a new worker: state will be lost");
    Stage stage;
    stage.version = version_ + 1;
    stage.cluster = next;
    if (!notify_runners(stage)) return false;
    {
        std::lock_guard<std::mutex> lk(mu_);
        cluster_ = next;
        version_ = stage.version;
    }
    *changed  = true;
    *detached = !next.workers.contains(cfg_.self);
    detached_ = *detached;
    return true;
}

bool Peer::fetch_agreed_cluster(Cluster *out)
{
    Cluster next;
    for (int attempt = 0;; ++attempt) {
        next = cluster();
        if (!cfg_.config_server.empty()) {
            std::string body, err;
            if (open_url(cfg_.config_server, &body, &err)) {
                Cluster c;
                std::string perr;
                if (parse_cluster_json(body, &c, &perr)) next = c;
                else KF_WARN("ignoring malformed cluster config: %s", perr.c_str());
            } else {
                KF_WARN("config server unreachable (%s); using the current cluster", err.c_str());
            }
        }
        const auto digest = next.bytes();
        bool agree        = false;
        if (!session()->bytes_consensus(digest.data(), digest.size(), "kungfu::resize", &agree)) {
            set_error("consensus on the new cluster failed");
            return false;
        }
        if (agree) break;
        if (attempt % 20 == 0) KF_WARN("diverged cluster proposals, retrying");
        std::this_thread::sleep_for(std::chrono::milliseconds(50));
    }
    *out = next;
    return true;
}

bool Peer::resize_cluster_from_url(bool *changed, bool *detached)
{
    *changed = *detached = false;
    if (cfg_.single) return true;
    Cluster next;
    if (!fetch_agreed_cluster(&next)) return false;
    if (!propose(next, changed, detached)) return false;
    if (*changed && !*detached) return update_to(next.workers);
    return true;
}

bool Peer::peek_cluster_change(bool *differs)
{
    *differs = false;
    if (cfg_.single) return true;
    Cluster next;
    if (!fetch_agreed_cluster(&next)) return false;
    *differs = next.bytes() != cluster().bytes();
    return true;
}

bool Peer::propose_new_size(int new_size)
{
    if (cfg_.config_server.empty()) {
        set_error("propose_new_size needs KUNGFU_CONFIG_SERVER (kungfu-run -config-server / -builtin-config-port)");
        return false;
    }
    Cluster next;
    std::string err;
    if (!cluster().resize(new_size, &next, &err)) {
        set_error("cannot resize to " + std::to_string(new_size) + ": " + err);
        return false;
    }
    int status = 0;
    std::string resp;
    if (!http_do("PUT", cfg_.config_server, next.json(), &status, &resp, &err) || status != 200) {
        set_error("PUT " + cfg_.config_server + " failed: " + err + " status " + std::to_string(status));
        return false;
    }
    return true;
}

bool Peer::resize_cluster(int new_size, bool *changed, bool *detached)
{
    *changed = *detached = false;
    if (cfg_.single) return true;
    if (!cfg_.config_server.empty()) {
        // Rank 0 publishes the new cluster; the others fetch it only after that PUT has landed (one tiny MIN
        // all-reduce of rank 0's outcome), so the first fetch already agrees: no "diverged proposal" round with
        // its 50 ms back-off, and a failed PUT fails the call on every rank instead of leaving the others polling.
        int32_t ok = 1, all_ok = 0;
        if (session()->rank() == 0 && !propose_new_size(new_size)) ok = 0;
        Workspace w;
        w.send  = &ok;
        w.recv  = &all_ok;
        w.count = 1;
        w.dtype = I32;
        w.op    = MIN;
        w.name  = "kungfu::resize::proposed";
        if (!session()->all_reduce(w)) return false;
        if (!all_ok) {
            if (session()->rank() != 0) set_error("rank 0 could not publish the resized cluster");
            return false;
        }
        return resize_cluster_from_url(changed, detached);
    }
    // No config server: every rank derives the same next cluster locally.
    Cluster next;
    std::string err;
    if (!cluster().resize(new_size, &next, &err)) {
        set_error("cannot resize to " + std::to_string(new_size) + ": " + err);
        return false;
    }
    const auto digest = next.bytes();
    bool agree        = false;
    if (!session()->bytes_consensus(digest.data(), digest.size(), "kungfu::resize", &agree) || !agree) {
        set_error("ranks disagree on the resized cluster");
        return false;
    }
    if (!propose(next, changed, detached)) return false;
    if (*changed && !*detached) return update_to(next.workers);
    return true;
}

void Peer::save(const std::string &version, const std::string &name, const void *data, size_t n)
{
    if (version.empty()) server_->store.save(name, data, n);
    else server_->vstore.save(version, name, data, n);
}

void Peer::save_blob(const std::string &version, const std::string &name, std::shared_ptr<std::vector<uint8_t>> blob)
{
    if (version.empty()) server_->store.adopt(name, std::move(blob));
    else server_->vstore.adopt(version, name, std::move(blob));
}

bool Peer::request(int target_rank, const std::string &version, const std::string &name, void *buf, size_t n)
{
    Session *s = session();
    if (target_rank < 0 || target_rank >= s->size()) {
        set_error("request: bad target rank");
        return false;
    }
    if (target_rank == s->rank()) {
        auto blob = version.empty() ? server_->store.get(name) : server_->vstore.get(version, name);
        if (!blob || blob->size() != n) return false;
        memcpy(buf, blob->data(), n);
        return true;
    }
    std::string err;
    if (!client_->request(s->peers()[target_rank], version, name, buf, n, &err)) {
        std::lock_guard<std::mutex> lk(err_mu_);
        last_error_ = err;  // a missing blob is an expected outcome: recorded, not logged
        return false;
    }
    return true;
}

bool Peer::send_fd(int target_rank, const std::string &tag, int fd)
{
    Session *s = session();
    std::string err;
    if (!client_->pass_fd(s->peers()[target_rank], tag, fd, &err)) {
        set_error(err);
        return false;
    }
    return true;
}

bool Peer::recv_fd(int src_rank, const std::string &tag, int *fd)
{
    Session *s = session();
    const int got = server_->mailbox.recv_fd(s->peers()[src_rank], tag, 120.0);
    if (got < 0) {
        set_error("timed out waiting for descriptor '" + tag + "' from rank " + std::to_string(src_rank));
        return false;
    }
    *fd = got;
    return true;
}

void Peer::async_loop()
{
    for (;;) {
        std::pair<int64_t, std::function<bool()>> job;
        {
            std::unique_lock<std::mutex> lk(async_mu_);
            async_cv_.wait(lk, [this] { return async_stop_ || !async_q_.empty(); });
            if (async_q_.empty()) return;
            job = std::move(async_q_.front());
            async_q_.pop_front();
        }
        const bool ok = job.second();
        if (job.first == 0) continue;  // fire-and-forget job: its own completion callback reports
        {
            std::lock_guard<std::mutex> lk(async_mu_);
            async_done_[job.first] = ok;
        }
        async_cv_.notify_all();
    }
}

int64_t Peer::submit(std::function<bool()> op)
{
    std::lock_guard<std::mutex> lk(async_mu_);
    const int64_t h = next_handle_++;
    async_q_.emplace_back(h, std::move(op));
    async_cv_.notify_all();
    return h;
}

void Peer::submit_detached(std::function<void()> op)
{
    std::lock_guard<std::mutex> lk(async_mu_);
    async_q_.emplace_back(0, [op] {
        op();
        return true;
    });
    async_cv_.notify_all();
}

bool Peer::wait(int64_t handle)
{
    std::unique_lock<std::mutex> lk(async_mu_);
    async_cv_.wait(lk, [&] { return async_done_.count(handle) > 0; });
    const bool ok = async_done_[handle];
    async_done_.erase(handle);
    return ok;
}

std::vector<float> Peer::egress_rates() { return session()->egress_rates(cfg_.monitoring ? &monitor_ : nullptr); }

}  // namespace kf
