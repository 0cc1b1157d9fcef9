#include "elastic/runner.hpp"

#include <arpa/inet.h>
#include <fcntl.h>
#include <ifaddrs.h>
#include <net/if.h>
#include <netdb.h>
#include <signal.h>
#include <sys/prctl.h>
#include <sys/stat.h>
#include <sys/wait.h>
#include <unistd.h>

#include <algorithm>
#include <cctype>
#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <fstream>
#include <sstream>
#include <thread>

#include "core/log.hpp"
#include "ctrl/peer.hpp"
#include "ctrl/transport.hpp"

extern char **environ;

namespace kf {

// =============================== config server ===============================
void ConfigServer::handle(const HttpRequest &rq, HttpResponse *rs)
{
    if (rq.path == "/stop") {
        stop_requested_ = true;
        rs->body        = "stopping\n";
        return;
    }
    if (rq.path != endpoint_) {
        rs->status = 404;
        rs->body   = "not found\n";
        return;
    }
    std::lock_guard<std::mutex> lk(mu_);
    if (rq.method == "GET") {
        if (!has_) {
            rs->status = 404;
            rs->body   = "No Config Found.\n";
            return;
        }
        rs->content_type = "application/json";
        rs->body         = cluster_.json() + "\n";
    } else if (rq.method == "PUT") {
        Cluster c;
        std::string err;
        if (!parse_cluster_json(rq.body, &c, &err) || !c.validate(&err)) {
            KF_ERROR("config server: rejected PUT: %s", err.c_str());
            rs->status = 400;
            rs->body   = err + "\n";
            return;
        }
        if (!has_) {
            KF_INFO("config server: first config with %zu workers", c.workers.size());
            has_     = true;
            version_ = 1;
            cluster_ = c;
        } else if (!cluster_.workers.empty()) {
            ++version_;
            cluster_ = c;
            KF_INFO("config server: v%d -> %zu workers: %s", version_, c.workers.size(), c.workers.str().c_str());
        } else {
            KF_INFO("config server: config was cleared, update rejected");
            rs->status = 403;
            rs->body   = "config was cleared\n";
        }
    } else if (rq.method == "POST" || rq.method == "DELETE") {
        has_ = false;
        if (rq.method == "DELETE") KF_WARN("config server: config deleted");
        else KF_INFO("config server: config reset");
    } else {
        rs->status = 405;
        rs->body   = "unsupported method\n";
    }
}

bool ConfigServer::start(uint16_t port, const std::string &endpoint, const Cluster *init, std::string *err)
{
    endpoint_ = endpoint.empty() ? "/config" : endpoint;
    if (init) {
        has_     = true;
        cluster_ = *init;
        version_ = 1;
    }
    return http_.start(port, [this](const HttpRequest &rq, HttpResponse *rs) { handle(rq, rs); }, err);
}

int ConfigServer::version()
{
    std::lock_guard<std::mutex> lk(mu_);
    return version_;
}

static std::atomic<bool> g_signalled{false};
static void on_signal(int) { g_signalled = true; }

int config_server_main(int argc, char **argv)
{
    int port = 9100;
    std::string init_file, endpoint = "/config";
    double ttl = 0;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        auto next = [&]() -> std::string { return i + 1 < argc ? argv[++i] : ""; };
        if (a == "-port") port = atoi(next().c_str());
        else if (a == "-init") init_file = next();
        else if (a == "-endpoint") endpoint = next();
        else if (a == "-ttl") ttl = parse_duration_seconds(next(), 0);
        else if (a == "-h" || a == "-help" || a == "--help") {
            printf("usage: kungfu-config-server [-port 9100] [-init cluster.json] [-endpoint /config] [-ttl 0s]\n");
            return 0;
        } else {
            fprintf(stderr, "unknown flag %s\n", a.c_str());
            return 2;
        }
    }
    Cluster init;
    bool has_init = false;
    if (!init_file.empty()) {
        std::string text, err;
        // the same checks a PUT goes through: a cluster that workers would refuse must not be published
        if (!open_url(init_file, &text, &err) || !parse_cluster_json(text, &init, &err) || !init.validate(&err)) {
            fprintf(stderr, "bad -init %s: %s\n", init_file.c_str(), err.c_str());
            return 1;
        }
        has_init = true;
    }
    ConfigServer srv;
    std::string err;
    if (!srv.start((uint16_t)port, endpoint, has_init ? &init : nullptr, &err)) {
        fprintf(stderr, "config server: %s\n", err.c_str());
        return 1;
    }
    signal(SIGINT, on_signal);
    signal(SIGTERM, on_signal);
    KF_INFO("config server: http://127.0.0.1:%d%s", port, endpoint.c_str());
    const double t0 = now_seconds();
    while (!g_signalled && !srv.stop_requested() && (ttl <= 0 || now_seconds() - t0 < ttl))
        std::this_thread::sleep_for(std::chrono::milliseconds(50));
    srv.stop();
    return 0;
}

// =============================== flags ======================================
std::string run_usage()
{
    return "usage: kungfu-run [flags] <program> [args...]\n"
           "  -np N                 number of workers (default 1)\n"
           "  -H ip:slots[:public],...   hosts (default 127.0.0.1:8)\n"
           "  -hostfile PATH        OpenMPI-style hostfile, overrides -H\n"
           "  -P ip:port,...        explicit worker list\n"
           "  -u USER               ssh user (kungfu-rrun)\n"
           "  -port-range B-E       worker ports (default 10000-11000)\n"
           "  -self IPV4 | -nic IF  this host's address\n"
           "  -timeout DUR          kill the job after DUR (0 = none)\n"
           "  -v=BOOL               echo worker output (default true)\n"
           "  -allow-nvlink         keep every GPU visible to every worker (needed for NVLink peer mapping)\n"
           "  -strategy NAME        STAR|RING|CLIQUE|TREE|BINARY_TREE|BINARY_TREE_STAR|MULTI_STAR|MULTI_BINARY_TREE_STAR|AUTO\n"
           "  -port N               runner control port (default 38080)\n"
           "  -debug-port N         HTTP dump of received stages\n"
           "  -w                    watch mode (elastic): obey cluster updates\n"
           "  -k                    stay alive after all workers finished\n"
           "  -init-version N       initial cluster version\n"
           "  -config-server URL    cluster config REST endpoint\n"
           "  -builtin-config-port N  also run a config server on this port\n"
           "  -t0 TS -logfile F -logdir D -q -delay DUR\n";
}

// "-H node1:4,node2:4" may name hosts: each name is resolved to its IPv4 and
// kept as the public address (/root/reference/srcs/go/kungfu/runner/discovery.go:157-215).
static bool resolve_host_names(const std::string &list, std::string *out, std::string *err)
{
    out->clear();
    size_t pos = 0;
    while (pos <= list.size()) {
        size_t comma = list.find(',', pos);
        if (comma == std::string::npos) comma = list.size();
        std::string spec = list.substr(pos, comma - pos);
        pos              = comma + 1;
        if (spec.empty()) continue;
        const size_t colon     = spec.find(':');
        const std::string host = spec.substr(0, colon);
        uint32_t ip;
        if (!parse_ipv4(host, &ip)) {
            addrinfo hints, *res = nullptr;
            memset(&hints, 0, sizeof(hints));
            hints.ai_family = AF_INET;
            if (getaddrinfo(host.c_str(), nullptr, &hints, &res) != 0 || !res) {
                *err = "cannot resolve host '" + host + "' in -H";
                return false;
            }
            ip = ntohl(((sockaddr_in *)res->ai_addr)->sin_addr.s_addr);
            freeaddrinfo(res);
            std::string rest = colon == std::string::npos ? ":1" : spec.substr(colon);
            // ip:slots:public (public defaults to the name the user gave)
            if (std::count(rest.begin(), rest.end(), ':') < 2) rest += ":" + host;
            spec = format_ipv4(ip) + rest;
        }
        if (!out->empty()) *out += ",";
        *out += spec;
    }
    return true;
}

static bool parse_bool(const std::string &v) { return v == "1" || v == "true" || v == "True" || v == "TRUE" || v.empty(); }

bool parse_run_flags(int argc, char **argv, RunFlags *f, std::string *err)
{
    f->t0 = (long)time(nullptr);
    int i = 1;
    for (; i < argc; ++i) {
        std::string a = argv[i];
        if (a.empty() || a[0] != '-') break;
        if (a == "--") {
            ++i;
            break;
        }
        while (a.size() > 1 && a[0] == '-') a.erase(0, 1);
        std::string val;
        bool has_val = false;
        auto eq      = a.find('=');
        if (eq != std::string::npos) {
            val     = a.substr(eq + 1);
            a       = a.substr(0, eq);
            has_val = true;
        }
        auto need = [&]() -> bool {
            if (has_val) return true;
            if (i + 1 >= argc) {
                *err = "flag -" + a + " needs a value";
                return false;
            }
            val = argv[++i];
            return true;
        };
        if (a == "np") { if (!need()) return false; f->np = atoi(val.c_str()); }
        else if (a == "H") { if (!need()) return false; f->host_list = val; }
        else if (a == "hostfile") { if (!need()) return false; f->hostfile = val; }
        else if (a == "P") { if (!need()) return false; f->peer_list = val; }
        else if (a == "u") { if (!need()) return false; f->user = val; }
        else if (a == "port-range") {
            if (!need()) return false;
            if (!parse_port_range(val, &f->port_range)) { *err = "bad -port-range " + val; return false; }
        }
        else if (a == "self") { if (!need()) return false; f->self = val; }
        else if (a == "timeout") { if (!need()) return false; f->timeout = parse_duration_seconds(val, 0); }
        else if (a == "v") { f->verbose = has_val ? parse_bool(val) : true; }
        else if (a == "nic") { if (!need()) return false; f->nic = val; }
        else if (a == "allow-nvlink") { f->allow_nvlink = has_val ? parse_bool(val) : true; }
        else if (a == "strategy") {
            if (!need()) return false;
            if (!parse_strategy(val, &f->strategy)) { *err = "unknown -strategy " + val; return false; }
        }
        else if (a == "port") { if (!need()) return false; f->port = atoi(val.c_str()); }
        else if (a == "debug-port") { if (!need()) return false; f->debug_port = atoi(val.c_str()); }
        else if (a == "w") { f->watch = has_val ? parse_bool(val) : true; }
        else if (a == "k") { f->keep = has_val ? parse_bool(val) : true; }
        else if (a == "init-version") { if (!need()) return false; f->init_version = atoi(val.c_str()); }
        else if (a == "config-server") { if (!need()) return false; f->config_server = val; }
        else if (a == "t0") { if (!need()) return false; f->t0 = atol(val.c_str()); }
        else if (a == "logfile") { if (!need()) return false; f->logfile = val; }
        else if (a == "logdir") { if (!need()) return false; f->logdir = val; }
        else if (a == "q") { f->quiet = has_val ? parse_bool(val) : true; }
        else if (a == "delay") { if (!need()) return false; f->delay = parse_duration_seconds(val, 0); }
        else if (a == "builtin-config-port") { if (!need()) return false; f->builtin_config_port = atoi(val.c_str()); }
        else if (a == "h" || a == "help") { *err = run_usage(); return false; }
        else { *err = "unknown flag -" + a + "\n" + run_usage(); return false; }
    }
    if (i >= argc) {
        *err = "missing program name\n" + run_usage();
        return false;
    }
    f->prog = argv[i++];
    for (; i < argc; ++i) f->args.push_back(argv[i]);
    if (!f->hostfile.empty()) {
        std::string text;
        if (!open_url(f->hostfile, &text, err)) return false;
        if (!parse_hostfile(text, &f->hosts, err)) return false;
    } else {
        std::string resolved;
        if (!resolve_host_names(f->host_list, &resolved, err)) return false;
        if (!parse_host_list(resolved, &f->hosts)) {
            *err = "bad -H " + f->host_list;
            return false;
        }
    }
    return true;
}

bool infer_self_ipv4(const RunFlags &f, uint32_t *out, std::string *err)
{
    if (!f.self.empty()) {
        if (parse_ipv4(f.self, out)) return true;
        *err = "bad -self " + f.self;
        return false;
    }
    if (!f.nic.empty()) {
        ifaddrs *ifs = nullptr;
        if (getifaddrs(&ifs) == 0) {
            for (ifaddrs *p = ifs; p; p = p->ifa_next) {
                if (p->ifa_addr && p->ifa_addr->sa_family == AF_INET && f.nic == p->ifa_name) {
                    *out = ntohl(((sockaddr_in *)p->ifa_addr)->sin_addr.s_addr);
                    freeifaddrs(ifs);
                    return true;
                }
            }
            freeifaddrs(ifs);
        }
        *err = "no IPv4 address on -nic " + f.nic;
        return false;
    }
    // single host in the list: that's us; otherwise loopback
    if (f.hosts.size() == 1) {
        *out = f.hosts[0].ipv4;
        return true;
    }
    parse_ipv4("127.0.0.1", out);
    return true;
}

int GpuPool::get()
{
    std::lock_guard<std::mutex> lk(mu_);
    for (size_t i = 0; i < used_.size(); ++i) {
        if (!used_[i]) {
            used_[i] = true;
            return (int)i;
        }
    }
    return -1;
}

void GpuPool::put(int id)
{
    std::lock_guard<std::mutex> lk(mu_);
    if (id >= 0 && id < (int)used_.size()) used_[id] = false;
}

// The runner's own CUDA_VISIBLE_DEVICES narrows and orders the GPUs its workers may use: slot k
// of the pool is the k-th entry of that list. A list of integers gives the device number (negative
// entries are skipped, a repeated one makes the list invalid); any other form (GPU-<uuid>, MIG-...)
// is handed on token by token and the slot stays the index into it.
int map_cuda_index(int slot, const char *visible, std::string *token)
{
    if (token) token->clear();
    if (slot < 0) return -1;
    if (visible == nullptr) {
        if (token) *token = std::to_string(slot);
        return slot;
    }
    std::vector<std::string> toks;
    {
        std::string cur;
        for (const char *c = visible;; ++c) {
            if (*c == ',' || *c == 0) {
                size_t b = cur.find_first_not_of(" \t"), e = cur.find_last_not_of(" \t");
                if (b != std::string::npos) toks.push_back(cur.substr(b, e - b + 1));
                cur.clear();
                if (*c == 0) break;
            } else {
                cur += *c;
            }
        }
    }
    if (toks.empty()) return -1;  // CUDA_VISIBLE_DEVICES="": a CPU-only job by configuration, nothing to warn about
    bool numeric = true;
    for (auto &t : toks) {
        size_t i = (t[0] == '-' || t[0] == '+') ? 1 : 0;
        if (i >= t.size() || t.size() > 9) numeric = false;
        for (; i < t.size(); ++i) numeric = numeric && isdigit((unsigned char)t[i]);
    }
    if (!numeric) {
        if (slot >= (int)toks.size()) {
            KF_WARN("CUDA_VISIBLE_DEVICES=%s has no entry for local slot %d", visible, slot);
            return -1;
        }
        if (token) *token = toks[slot];
        return slot;
    }
    std::vector<int> ids;
    for (auto &t : toks) {
        const int n = atoi(t.c_str());
        if (n < 0) continue;
        if (std::find(ids.begin(), ids.end(), n) != ids.end()) {
            KF_WARN("invalid CUDA_VISIBLE_DEVICES=%s: device %d is listed twice", visible, n);
            return -1;
        }
        ids.push_back(n);
    }
    if (slot >= (int)ids.size()) {
        KF_WARN("CUDA_VISIBLE_DEVICES=%s has no entry for local slot %d", visible, slot);
        return -1;
    }
    if (token) *token = std::to_string(ids[slot]);
    return ids[slot];
}

ProcSpec make_worker_proc(const RunFlags &f, const PeerID &self, const PeerID &parent, const Cluster &cluster,
                          int version, int gpu_id)
{
    ProcSpec p;
    p.name    = self.host_str() + "." + std::to_string(self.port);
    p.prog    = f.prog;
    p.args    = f.args;
    p.logdir  = f.logdir;
    p.log_tag = p.name + "@" + std::to_string(version);
    p.self    = self;
    p.gpu     = gpu_id;
    auto &e   = p.env;
    e[env::kSelfSpec]    = self.str();
    e[env::kInitPeers]   = cluster.workers.str();
    e[env::kInitRunners] = cluster.runners.str();
    e[env::kParentID]    = parent.str();
    e[env::kInitVersion] = std::to_string(version);
    e[env::kStrategy]    = strategy_name(f.strategy);
    e[env::kConfigServer] = f.config_server;
    e[env::kAllowNvlink] = f.allow_nvlink ? "true" : "false";
    e[env::kJobStart]    = std::to_string(f.t0);
    e[env::kProcStart]   = std::to_string((long)time(nullptr));
    // slot of the GPU pool -> device, through the runner's own CUDA_VISIBLE_DEVICES
    // (/root/reference/srcs/go/kungfu/job/cuda_visible_device.go:17-58, job.go:47-53)
    std::string token;
    const int cuda_idx   = map_cuda_index(gpu_id, getenv("CUDA_VISIBLE_DEVICES"), &token);
    e[env::kCudaVisible] = std::to_string(cuda_idx);
    // Without -allow-nvlink each worker only sees its own GPU (the reference's
    // default). Peer-memory collectives need every GPU of the node mapped, so
    // GPU jobs should pass -allow-nvlink; the worker then selects its device
    // from KUNGFU_CUDA_VISIBLE_DEVICES (looked up in the inherited CUDA_VISIBLE_DEVICES).
    if (!f.allow_nvlink && cuda_idx >= 0) e["CUDA_VISIBLE_DEVICES"] = token;
    e["PYTHONUNBUFFERED"] = "1";
    return p;
}

// =============================== processes ===================================
namespace {

const char *kColors[] = {"\x1b[1;32m", "\x1b[1;34m", "\x1b[1;33m", "\x1b[1;35m", "\x1b[1;36m", "\x1b[1;31m"};

struct RunningProc {
    pid_t pid = -1;
    ProcSpec spec;
    std::thread out_th, err_th;
    int exit_code = -1;
    bool done     = false;
    int attempt   = 0;
    std::atomic<bool> loader_bug{false};  // stderr carried the dynamic-linker assertion (see kLoaderBug)
};

// glibc's dynamic linker occasionally dies at start-up with this assertion when many processes load the
// same CUDA / NCCL libraries at once; the process has done nothing yet, so starting it again is safe.
// (/root/reference/srcs/go/utils/runner/local/hack.go:14-35, srcs/go/nccl/bug.go:10)
constexpr const char *kLoaderBug = "Inconsistency detected by ld.so";
constexpr int kLoaderBugRetries  = 3;

void pump(int fd, const std::string &prefix, const char *color, bool echo, FILE *copy, bool to_stderr,
          std::atomic<bool> *loader_bug = nullptr)
{
    FILE *in = fdopen(fd, "r");
    if (!in) return;
    char *line = nullptr;
    size_t cap = 0;
    ssize_t n;
    while ((n = getline(&line, &cap, in)) > 0) {
        if (loader_bug && strstr(line, kLoaderBug)) *loader_bug = true;
        if (copy) {
            fwrite(line, 1, (size_t)n, copy);
            fflush(copy);
        }
        if (echo) {
            FILE *out = to_stderr ? stderr : stdout;
            fprintf(out, "%s[%s]\x1b[m %s", color, prefix.c_str(), line);
            if (line[n - 1] != '\n') fputc('\n', out);
            fflush(out);
        }
    }
    free(line);
    fclose(in);
}

class ProcTable
{
    std::mutex mu_;
    std::condition_variable cv_;
    std::map<PeerID, std::shared_ptr<RunningProc>> procs_;
    std::deque<std::pair<PeerID, int>> exits_;
    bool verbose_;
    int color_ = 0;

  public:
    explicit ProcTable(bool verbose) : verbose_(verbose) {}

    bool spawn(const ProcSpec &spec, std::string *err, int attempt = 0)
    {
        int po[2], pe[2];
        if (pipe2(po, O_CLOEXEC) != 0 || pipe2(pe, O_CLOEXEC) != 0) {
            *err = "pipe failed";
            return false;
        }
        std::vector<std::string> envs;
        std::map<std::string, std::string> merged;
        for (char **e = environ; *e; ++e) {
            std::string kv = *e;
            auto eq        = kv.find('=');
            if (eq != std::string::npos) merged[kv.substr(0, eq)] = kv.substr(eq + 1);
        }
        for (auto &kv : spec.env) merged[kv.first] = kv.second;
        for (auto &kv : merged) envs.push_back(kv.first + "=" + kv.second);
        std::vector<char *> envp, argv;
        for (auto &s : envs) envp.push_back(const_cast<char *>(s.c_str()));
        envp.push_back(nullptr);
        argv.push_back(const_cast<char *>(spec.prog.c_str()));
        for (auto &a : spec.args) argv.push_back(const_cast<char *>(a.c_str()));
        argv.push_back(nullptr);
        pid_t pid = fork();
        if (pid < 0) {
            *err = "fork failed";
            return false;
        }
        if (pid == 0) {
            setpgid(0, 0);
            // never outlive the runner: an orphaned worker would keep its port and GPU
            prctl(PR_SET_PDEATHSIG, SIGKILL);
            if (getppid() == 1) _exit(1);  // runner died between fork and prctl
            dup2(po[1], 1);
            dup2(pe[1], 2);
            execvpe(spec.prog.c_str(), argv.data(), envp.data());
            fprintf(stderr, "exec %s failed: %s\n", spec.prog.c_str(), strerror(errno));
            _exit(127);
        }
        ::close(po[1]);
        ::close(pe[1]);
        auto rp  = std::make_shared<RunningProc>();
        rp->pid  = pid;
        rp->spec = spec;
        rp->attempt = attempt;
        FILE *lo = nullptr, *le = nullptr;
        if (!spec.logdir.empty()) {
            mkdir(spec.logdir.c_str(), 0755);
            lo = fopen((spec.logdir + "/" + spec.log_tag + ".stdout.log").c_str(), "we");
            le = fopen((spec.logdir + "/" + spec.log_tag + ".stderr.log").c_str(), "we");
        }
        const char *color = kColors[color_++ % 6];
        rp->out_th = std::thread([=] { pump(po[0], spec.name + "::stdout", color, verbose_, lo, false); if (lo) fclose(lo); });
        rp->err_th = std::thread([=] { pump(pe[0], spec.name + "::stderr", color, verbose_, le, true, &rp->loader_bug); if (le) fclose(le); });
        {
            std::lock_guard<std::mutex> lk(mu_);
            procs_[spec.self] = rp;
        }
        std::thread([this, rp] {
            int status = 0;
            while (waitpid(rp->pid, &status, 0) < 0 && errno == EINTR) {}
            rp->out_th.join();
            rp->err_th.join();
            const int code = WIFEXITED(status) ? WEXITSTATUS(status) : 128 + (WIFSIGNALED(status) ? WTERMSIG(status) : 0);
            if (code != 0 && rp->loader_bug && rp->attempt < kLoaderBugRetries) {
                KF_WARN("runner: %s hit the dynamic-linker start-up assertion (exit %d): starting it again (%d/%d)",
                        rp->spec.name.c_str(), code, rp->attempt + 1, kLoaderBugRetries);
                std::string e;
                if (spawn(rp->spec, &e, rp->attempt + 1)) {
                    // PR_SET_PDEATHSIG follows the forking THREAD: this thread has to outlive the child it just
                    // started, or the kernel kills the child the moment we return
                    std::unique_lock<std::mutex> lk(mu_);
                    std::shared_ptr<RunningProc> child = procs_[rp->spec.self];
                    cv_.wait(lk, [&] { return child->done; });
                    return;
                }
                KF_ERROR("runner: restart of %s failed: %s", rp->spec.name.c_str(), e.c_str());
            }
            std::lock_guard<std::mutex> lk(mu_);
            rp->exit_code = code;
            rp->done      = true;
            exits_.push_back({rp->spec.self, code});
            cv_.notify_all();
        }).detach();
        return true;
    }

    bool running(const PeerID &id)
    {
        std::lock_guard<std::mutex> lk(mu_);
        auto it = procs_.find(id);
        return it != procs_.end() && !it->second->done;
    }

    int live()
    {
        std::lock_guard<std::mutex> lk(mu_);
        int n = 0;
        for (auto &kv : procs_)
            if (!kv.second->done) ++n;
        return n;
    }

    // wait for one exit event; false on timeout
    bool next_exit(std::pair<PeerID, int> *out, double timeout_s)
    {
        std::unique_lock<std::mutex> lk(mu_);
        if (!cv_.wait_for(lk, std::chrono::duration<double>(timeout_s), [this] { return !exits_.empty(); })) return false;
        *out = exits_.front();
        exits_.pop_front();
        return true;
    }

    void wait_for(const PeerID &id, double timeout_s)
    {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait_for(lk, std::chrono::duration<double>(timeout_s), [&] {
            auto it = procs_.find(id);
            return it == procs_.end() || it->second->done;
        });
    }

    void forget(const PeerID &id)
    {
        std::lock_guard<std::mutex> lk(mu_);
        auto it = procs_.find(id);
        if (it != procs_.end() && it->second->done) procs_.erase(it);
    }

    void kill_all(int sig)
    {
        std::lock_guard<std::mutex> lk(mu_);
        for (auto &kv : procs_)
            if (!kv.second->done) ::kill(-kv.second->pid, sig);
    }

    void wait_all(double timeout_s)
    {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait_for(lk, std::chrono::duration<double>(timeout_s), [this] {
            for (auto &kv : procs_)
                if (!kv.second->done) return false;
            return true;
        });
    }
};

void dump_startup_info(const RunFlags &f, int argc, char **argv)
{
    if (f.quiet) return;
    std::string cmd;
    for (int i = 0; i < argc; ++i) cmd += std::string(i ? " " : "") + argv[i];
    KF_INFO("kungfu-run: %s", cmd.c_str());
    for (char **e = environ; *e; ++e) {
        if (!strncmp(*e, "KUNGFU_", 7) || !strncmp(*e, "CUDA_", 5) || !strncmp(*e, "NCCL_", 5)) KF_INFO("env %s", *e);
    }
    ifaddrs *ifs = nullptr;
    if (getifaddrs(&ifs) == 0) {
        for (ifaddrs *p = ifs; p; p = p->ifa_next) {
            if (p->ifa_addr && p->ifa_addr->sa_family == AF_INET) {
                char buf[64];
                inet_ntop(AF_INET, &((sockaddr_in *)p->ifa_addr)->sin_addr, buf, sizeof(buf));
                KF_INFO("nic %s %s", p->ifa_name, buf);
            }
        }
        freeifaddrs(ifs);
    }
}

// The reference retried workers that died of a specific dynamic-linker crash
// (/root/reference/srcs/go/utils/runner/local/hack.go:14-35). That failure
// mode belongs to its NCCL/TF stack; here a failing worker fails the job.

int simple_run(const RunFlags &f, const PeerID &parent, const Cluster &cluster)
{
    ProcTable table(f.verbose);
    GpuPool gpus(std::max(1, (int)cluster.workers.on_host(parent.ipv4).size()));
    int started = 0;
    for (auto &w : cluster.workers.on_host(parent.ipv4)) {
        std::string err;
        if (!table.spawn(make_worker_proc(f, w, parent, cluster, f.init_version, gpus.get()), &err)) {
            KF_ERROR("cannot start worker %s: %s", w.str().c_str(), err.c_str());
            table.kill_all(SIGTERM);
            return 1;
        }
        ++started;
    }
    if (started == 0) {
        KF_WARN("no worker of this job runs on %s", parent.host_str().c_str());
        return 0;
    }
    const double t0 = now_seconds();
    int rc          = 0;
    while (table.live() > 0 || rc == 0) {
        std::pair<PeerID, int> ev;
        if (table.next_exit(&ev, 0.2)) {
            if (ev.second != 0) {
                KF_ERROR("worker %s exited with %d; stopping the others", ev.first.str().c_str(), ev.second);
                rc = ev.second;
                table.kill_all(SIGTERM);
                table.wait_all(5);
                table.kill_all(SIGKILL);
                break;
            }
            if (table.live() == 0) break;
        }
        if (g_signalled || (f.timeout > 0 && now_seconds() - t0 > f.timeout)) {
            KF_ERROR(g_signalled ? "interrupted" : "timeout");
            rc = g_signalled ? 130 : 124;
            table.kill_all(SIGTERM);
            table.wait_all(5);
            table.kill_all(SIGKILL);
            break;
        }
    }
    table.wait_all(5);
    return rc;
}

int watch_run(const RunFlags &f, const PeerID &parent, const Cluster &init_cluster)
{
    std::mutex mu;
    std::deque<Stage> pending;
    std::map<int, std::string> seen;  // version -> cluster json
    std::vector<std::string> history;
    std::atomic<bool> exit_requested{false}, inconsistent{false};

    Server server(parent, true);
    server.set_token((uint32_t)f.init_version);
    server.on_control = [&](const PeerID &src, const std::string &name, const std::string &body) {
        if (name == "exit") {
            exit_requested = true;
            return;
        }
        if (name != "update") return;
        Stage s;
        std::string err;
        if (!parse_stage_json(body, &s, &err)) {
            KF_ERROR("runner: bad update from %s: %s", src.str().c_str(), err.c_str());
            return;
        }
        std::lock_guard<std::mutex> lk(mu);
        auto it = seen.find(s.version);
        if (it != seen.end()) {
            if (it->second != s.cluster.json()) {
                KF_ERROR("runner: inconsistent update for version %d", s.version);
                inconsistent = true;
            }
            return;  // every worker sends the same stage; the first one wins
        }
        seen[s.version] = s.cluster.json();
        history.push_back(s.json());
        pending.push_back(s);
    };
    std::string err;
    if (!server.start(&err)) {
        KF_ERROR("runner: cannot listen on %s: %s", parent.str().c_str(), err.c_str());
        return 1;
    }
    HttpServer debug;
    if (f.debug_port > 0) {
        debug.start((uint16_t)f.debug_port,
                    [&](const HttpRequest &, HttpResponse *rs) {
                        std::lock_guard<std::mutex> lk(mu);
                        rs->content_type = "application/json";
                        rs->body         = "[";
                        for (size_t i = 0; i < history.size(); ++i) rs->body += (i ? "," : "") + history[i];
                        rs->body += "]\n";
                    },
                    &err);
    }

    ProcTable table(f.verbose);
    int slots = 0;
    for (auto &h : f.hosts)
        if (h.ipv4 == parent.ipv4) slots = h.slots;
    GpuPool gpus(std::max(slots, 1));
    std::map<PeerID, int> gpu_of;
    Cluster current;  // nothing runs yet
    if (f.init_version < 0) {
        // started empty; the first stage arrives over the control channel (kungfu-notify-start,
        // a cluster manager). /root/reference/srcs/go/cmd/kungfu-run/app/kungfu-run.go:99-101
        KF_INFO("runner %s: waiting to be initialized", parent.str().c_str());
    } else {
        Stage s0;
        s0.version = f.init_version;
        s0.cluster = init_cluster;
        std::lock_guard<std::mutex> lk(mu);
        seen[s0.version] = s0.cluster.json();
        history.push_back(s0.json());
        pending.push_back(s0);
    }
    KF_INFO("runner %s: watching for cluster updates", parent.str().c_str());
    const double t0 = now_seconds();
    int rc          = 0;
    bool ever_ran   = false;
    for (;;) {
        Stage s;
        bool have = false;
        {
            std::lock_guard<std::mutex> lk(mu);
            if (!pending.empty()) {
                s = pending.front();
                pending.pop_front();
                have = true;
            }
        }
        if (have) {
            server.set_token((uint32_t)s.version);
            PeerList removed, added;
            current.workers.diff(s.cluster.workers, &removed, &added);
            PeerList del = removed.on_host(parent.ipv4), add = added.on_host(parent.ipv4);
            KF_INFO("runner: v%d np=%zu local +%zu/-%zu global +%zu/-%zu", s.version, s.cluster.workers.size(),
                    add.size(), del.size(), added.size(), removed.size());
            for (auto &id : del) {
                table.wait_for(id, 120);  // detached workers leave by themselves
                if (table.running(id)) KF_WARN("runner: %s did not stop after being removed", id.str().c_str());
                gpus.put(gpu_of[id]);
                gpu_of.erase(id);
            }
            for (auto &id : add) {
                const int gpu = gpus.get();
                if (gpu < 0) KF_WARN("runner: no free GPU slot for %s", id.str().c_str());
                gpu_of[id] = gpu;
                table.forget(id);
                std::string e;
                if (!table.spawn(make_worker_proc(f, id, parent, s.cluster, s.version, gpu), &e)) {
                    KF_ERROR("runner: cannot start %s: %s", id.str().c_str(), e.c_str());
                    rc = 1;
                    break;
                }
                ever_ran = true;
            }
            current = s.cluster;
            if (rc) break;
            continue;
        }
        std::pair<PeerID, int> ev;
        if (table.next_exit(&ev, 0.1)) {
            if (ev.second != 0) {
                KF_ERROR("runner: worker %s exited with %d", ev.first.str().c_str(), ev.second);
                rc = ev.second;
                break;
            }
            bool more;
            {
                std::lock_guard<std::mutex> lk(mu);
                more = !pending.empty();
            }
            if (table.live() == 0 && !more && ever_ran && !f.keep) break;
        }
        if (exit_requested || inconsistent || g_signalled) {
            rc = inconsistent ? 1 : (g_signalled ? 130 : 0);
            break;
        }
        if (f.timeout > 0 && now_seconds() - t0 > f.timeout) {
            KF_ERROR("runner: timeout");
            rc = 124;
            break;
        }
    }
    table.kill_all(SIGTERM);
    table.wait_all(5);
    table.kill_all(SIGKILL);
    debug.stop();
    server.stop();
    KF_INFO("runner %s: stop watching", parent.str().c_str());
    return rc;
}

}  // namespace

int kungfu_run_main(int argc, char **argv)
{
    RunFlags f;
    std::string err;
    if (!parse_run_flags(argc, argv, &f, &err)) {
        fprintf(stderr, "%s\n", err.c_str());
        return 2;
    }
    log_set_level_from_env();
    if (!f.logfile.empty()) log_set_file(f.logfile);
    dump_startup_info(f, argc, argv);
    signal(SIGINT, on_signal);
    signal(SIGTERM, on_signal);
    if (f.delay > 0) std::this_thread::sleep_for(std::chrono::duration<double>(f.delay));

    uint32_t self_ip;
    if (!infer_self_ipv4(f, &self_ip, &err)) {
        fprintf(stderr, "%s\n", err.c_str());
        return 2;
    }
    const PeerID parent{self_ip, (uint16_t)f.port};
    Cluster cluster;
    cluster.runners = f.hosts.gen_runner_list((uint16_t)f.port);
    if (!f.peer_list.empty()) {
        if (!parse_peer_list(f.peer_list, &cluster.workers)) {
            fprintf(stderr, "bad -P %s\n", f.peer_list.c_str());
            return 2;
        }
    } else if (!f.hosts.gen_peer_list(f.np, f.port_range, &cluster.workers, &err)) {
        fprintf(stderr, "%s\n", err.c_str());
        return 2;
    }
    if (!cluster.validate(&err)) {
        fprintf(stderr, "invalid cluster: %s\n", err.c_str());
        return 2;
    }

    std::unique_ptr<ConfigServer> builtin;
    if (f.builtin_config_port > 0) {
        builtin.reset(new ConfigServer());
        if (!builtin->start((uint16_t)f.builtin_config_port, "/config", &cluster, &err)) {
            fprintf(stderr, "builtin config server: %s\n", err.c_str());
            return 1;
        }
        if (f.config_server.empty())
            f.config_server = "http://127.0.0.1:" + std::to_string(f.builtin_config_port) + "/config";
        KF_INFO("builtin config server: %s", f.config_server.c_str());
    }
    int rc = f.watch ? watch_run(f, parent, cluster) : simple_run(f, parent, cluster);
    if (builtin) builtin->stop();
    return rc;
}

}  // namespace kf
