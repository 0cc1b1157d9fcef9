// Peer: the process-level runtime object (one per worker).
//
// Owns the server, the client, the current Session, the p2p store and the
// elastic state (cluster + version). Parity:
// /root/reference/srcs/go/kungfu/peer/{peer.go,p2p.go,legacy.go,router.go},
// /root/reference/srcs/go/kungfu/env/{config.go,envs.go},
// /root/reference/srcs/cpp/include/kungfu/peer.hpp:13-161 (sync + async API).
#pragma once
#include <condition_variable>
#include <deque>
#include <functional>
#include <future>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>

#include "ctrl/http.hpp"
#include "ctrl/session.hpp"
#include "ctrl/transport.hpp"

namespace kf {

namespace env {
constexpr const char *kSelfSpec        = "KUNGFU_SELF_SPEC";
constexpr const char *kInitPeers       = "KUNGFU_INIT_PEERS";
constexpr const char *kInitRunners     = "KUNGFU_INIT_RUNNERS";
constexpr const char *kParentID        = "KUNGFU_PARENT_ID";
constexpr const char *kInitVersion     = "KUNGFU_INIT_CLUSTER_VERSION";
constexpr const char *kStrategy        = "KUNGFU_ALLREDUCE_STRATEGY";
constexpr const char *kConfigServer    = "KUNGFU_CONFIG_SERVER";
constexpr const char *kAllowNvlink     = "KUNGFU_ALLOW_NVLINK";
constexpr const char *kJobStart        = "KUNGFU_JOB_START_TIMESTAMP";
constexpr const char *kProcStart       = "KUNGFU_PROC_START_TIMESTAMP";
constexpr const char *kCudaVisible     = "KUNGFU_CUDA_VISIBLE_DEVICES";
}  // namespace env

struct PeerConfig {
    bool single = true;
    PeerID self, parent;
    PeerList init_peers, init_runners;
    int init_version  = 0;
    Strategy strategy = Auto;
    std::string config_server;
    bool monitoring          = false;
    double monitoring_period = 1.0;
    static bool from_env(PeerConfig *out, std::string *err);
};

class Peer
{
    PeerConfig cfg_;
    TransportConfig tcfg_;
    std::unique_ptr<Server> server_;
    std::unique_ptr<Client> client_;
    NetMonitor monitor_;
    HttpServer metrics_;
    std::thread monitor_thread_;
    std::atomic<bool> closing_{false};

    std::mutex mu_;
    std::unique_ptr<Session> session_;
    Cluster cluster_;
    int version_   = 0;
    bool detached_ = false;
    bool started_  = false;
    std::mutex err_mu_;  // async ops fail on worker threads
    std::string last_error_;

    // async ops
    std::mutex async_mu_;
    std::condition_variable async_cv_;
    std::deque<std::pair<int64_t, std::function<bool()>>> async_q_;
    std::map<int64_t, bool> async_done_;
    std::vector<std::thread> async_workers_;
    int64_t next_handle_ = 1;
    bool async_stop_     = false;
    void async_loop();
    bool fetch_agreed_cluster(Cluster *out);

    bool update_to(const PeerList &workers);
    bool propose(const Cluster &next, bool *changed, bool *detached);
    bool notify_runners(const Stage &stage);
    void set_error(const std::string &e);

  public:
    explicit Peer(const PeerConfig &cfg);
    ~Peer();
    bool start();
    void close();
    std::string last_error();

    Session *session();  // current session (never null after start)
    const PeerConfig &config() const { return cfg_; }
    bool single() const { return cfg_.single; }
    bool detached() const { return detached_; }
    int cluster_version() const { return version_; }
    uint64_t uid() const;
    Cluster cluster();

    // elastic
    bool resize_cluster(int new_size, bool *changed, bool *detached);
    bool resize_cluster_from_url(bool *changed, bool *detached);
    // Does the config server publish a cluster other than the current one? Same fetch + byte
    // consensus as resize_cluster_from_url (every rank gets the same answer), nothing is changed:
    // lets a caller prepare (e.g. make sharded optimizer state whole) only when a resize is coming.
    bool peek_cluster_change(bool *differs);
    bool propose_new_size(int new_size);

    // p2p model store
    void save(const std::string &version, const std::string &name, const void *data, size_t n);
    // publish a blob that was assembled by the caller (one copy from the tensors into the store, not two)
    void save_blob(const std::string &version, const std::string &name, std::shared_ptr<std::vector<uint8_t>> blob);
    bool request(int target_rank, const std::string &version, const std::string &name, void *buf, size_t n);

    // descriptor passing for the device layer
    bool send_fd(int target_rank, const std::string &tag, int fd);
    bool recv_fd(int src_rank, const std::string &tag, int *fd);

    // async: returns a handle; wait() reports the op's success
    int64_t submit(std::function<bool()> op);
    bool wait(int64_t handle);
    // runs op on the same worker pool; nothing is recorded (the op reports its own completion).
    // Jobs queued before close() are still run: the destructor drains the queue, then joins.
    void submit_detached(std::function<void()> op);

    std::vector<float> egress_rates();
};

}  // namespace kf
