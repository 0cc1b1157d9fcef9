// kungfu-run (static and watch mode), the REST config server, process control.
//
// Parity: /root/reference/srcs/go/kungfu/runner/{flags.go,simple.go,watch.go,
// handler.go,discovery.go}, /root/reference/srcs/go/kungfu/job/{job.go,
// gpu_resource.go,cuda_visible_device.go}, /root/reference/srcs/go/proc/proc.go,
// /root/reference/srcs/go/utils/runner/local/local.go,
// /root/reference/srcs/go/kungfu/elastic/configserver/configserver.go.
#pragma once
#include <atomic>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "core/base.hpp"
#include "core/plan.hpp"
#include "ctrl/http.hpp"

namespace kf {

// ---- config server -------------------------------------------------------------------------
// GET  <endpoint>        200 cluster json | 404
// PUT  <endpoint>        validate, store, version++ (403 once cleared to 0 workers)
// POST <endpoint>        reset (forget the cluster)
// DELETE <endpoint>      same as POST, logged as a warning
// GET  /stop             stop the server
class ConfigServer
{
    HttpServer http_;
    std::mutex mu_;
    bool has_ = false;
    Cluster cluster_;
    int version_ = 0;
    std::string endpoint_;
    std::atomic<bool> stop_requested_{false};
    void handle(const HttpRequest &rq, HttpResponse *rs);

  public:
    bool start(uint16_t port, const std::string &endpoint, const Cluster *init, std::string *err);
    void stop() { http_.stop(); }
    bool stop_requested() const { return stop_requested_; }
    int version();
};

int config_server_main(int argc, char **argv);

// ---- kungfu-run -----------------------------------------------------------------------------
struct RunFlags {
    int np = 1;
    std::string host_list = "127.0.0.1:8";  // default: this machine
    std::string hostfile;
    std::string peer_list;
    std::string user;
    PortRange port_range;
    std::string self;
    double timeout = 0;
    bool verbose   = true;
    std::string nic;
    bool allow_nvlink = false;
    Strategy strategy = Auto;
    int port       = kDefaultRunnerPort;
    int debug_port = 0;
    bool watch     = false;
    bool keep      = false;
    int init_version = 0;
    std::string config_server;
    long t0 = 0;
    std::string logfile, logdir;
    bool quiet = false;
    double delay = 0;
    int builtin_config_port = 0;
    std::string prog;
    std::vector<std::string> args;
    HostList hosts;
};
bool parse_run_flags(int argc, char **argv, RunFlags *f, std::string *err);
std::string run_usage();

// IPv4 of this host: -self, else the address of -nic, else 127.0.0.1
bool infer_self_ipv4(const RunFlags &f, uint32_t *out, std::string *err);

// Pool of local GPU ids handed to workers (KUNGFU_CUDA_VISIBLE_DEVICES).
class GpuPool
{
    std::mutex mu_;
    std::vector<bool> used_;

  public:
    explicit GpuPool(int n) : used_(n, false) {}
    int get();  // -1 when exhausted
    void put(int id);
    int cap() const { return (int)used_.size(); }
};

// Device for pool slot `slot` given the runner's CUDA_VISIBLE_DEVICES (`visible`, nullptr if unset):
// the value of KUNGFU_CUDA_VISIBLE_DEVICES, -1 if there is none; *token = what CUDA_VISIBLE_DEVICES of
// the worker is narrowed to. /root/reference/srcs/go/kungfu/job/cuda_visible_device.go:17-58
int map_cuda_index(int slot, const char *visible, std::string *token);

struct ProcSpec {
    std::string name;  // shown as the log prefix
    std::string prog;
    std::vector<std::string> args;
    std::map<std::string, std::string> env;  // merged over the runner's env
    std::string logdir;
    std::string log_tag;
    PeerID self;
    int gpu = -1;
};

// Environment contract for one worker (job.go:31-72).
ProcSpec make_worker_proc(const RunFlags &f, const PeerID &self, const PeerID &parent, const Cluster &cluster,
                          int version, int gpu_id);

int kungfu_run_main(int argc, char **argv);

}  // namespace kf
