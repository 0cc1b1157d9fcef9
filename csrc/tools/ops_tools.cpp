// Operator tools for elastic jobs (one source, dispatched on the program name):
//
//   kungfu-test-util   -kill ip:port              ask a worker to exit over its control channel
//                      -propose cluster.json [-server URL]   PUT a cluster to the config server
//   kungfu-notify-start -config-server URL        read the cluster and send stage 0 to every runner
//                                                 (starts the workers of `kungfu-run -w -np 0` runners)
//   kungfu-config-client -server URL [-H ip:slots,...] [-port-range a-b] [-period 1s] [-ttl 10s]
//                      [-reset | -clear | -example]           random-size cluster every period
//                      [-shuffle] [-seed N]    -shuffle: a random SUBSET of the slots in random order (ranks move,
//                                              hosts empty out, sometimes no worker survives), as the reference's
//                                              client does; without it the first n slots
//
// Reference: /root/reference/tests/go/cmd/kungfu-test-util/kungfu-test-util.go:20-86,
// kungfu-notify-start/kungfu-notify-start.go:25-70, kungfu-config-client-example/
// kungfu-config-client-example.go:22-156.
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <random>
#include <sstream>
#include <string>
#include <thread>

#include "core/log.hpp"
#include "core/plan.hpp"
#include "ctrl/http.hpp"
#include "ctrl/transport.hpp"

using namespace kf;

namespace
{
struct Flags {
    std::map<std::string, std::string> kv;
    bool has(const std::string &k) const { return kv.count(k) != 0; }
    std::string get(const std::string &k, const std::string &def) const
    {
        auto it = kv.find(k);
        return it == kv.end() ? def : it->second;
    }
};

// -flag value | -boolflag
Flags parse(int argc, char **argv, const char *const *bool_flags)
{
    Flags f;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        while (!a.empty() && a[0] == '-') a.erase(0, 1);
        bool is_bool = false;
        for (const char *const *b = bool_flags; *b; ++b) is_bool |= a == *b;
        if (is_bool || i + 1 >= argc)
            f.kv[a] = "1";
        else
            f.kv[a] = argv[++i];
    }
    return f;
}

bool put_cluster(const std::string &server, const Cluster &c)
{
    int status = 0;
    std::string resp, err;
    if (!http_do("PUT", server, c.json(), &status, &resp, &err)) {
        fprintf(stderr, "PUT %s failed: %s\n", server.c_str(), err.c_str());
        return false;
    }
    printf("PUT %s -> %d (%zu workers)\n", server.c_str(), status, c.workers.size());
    return status == 200;
}

int test_util(const Flags &f)
{
    if (f.has("kill")) {
        PeerID target;
        if (!parse_peer_id(f.get("kill", ""), &target)) {
            fprintf(stderr, "bad peer id %s (want ip:port)\n", f.get("kill", "").c_str());
            return 2;
        }
        TransportConfig cfg = TransportConfig::from_env();
        cfg.conn_retry_count = 3;
        Client client(PeerID(), cfg);
        std::string err;
        if (!client.send(target, CONN_CONTROL, "exit", FLAG_NONE, nullptr, 0, &err)) {
            fprintf(stderr, "cannot send the exit signal to %s: %s\n", target.str().c_str(), err.c_str());
            return 1;
        }
        printf("exit signal sent to %s\n", target.str().c_str());
        return 0;
    }
    if (f.has("propose")) {
        std::ifstream in(f.get("propose", ""));
        std::stringstream ss;
        ss << in.rdbuf();
        Cluster c;
        std::string err;
        if (!in || !parse_cluster_json(ss.str(), &c, &err)) {
            fprintf(stderr, "cannot read a cluster from %s: %s\n", f.get("propose", "").c_str(), err.c_str());
            return 1;
        }
        return put_cluster(f.get("server", "http://127.0.0.1:9100/config"), c) ? 0 : 1;
    }
    fprintf(stderr, "usage: kungfu-test-util -kill ip:port | -propose cluster.json [-server URL]\n");
    return 2;
}

int notify_start(const Flags &f)
{
    const std::string url = f.get("config-server", "");
    std::string body, err;
    Cluster c;
    if (url.empty() || !open_url(url, &body, &err) || !parse_cluster_json(body, &c, &err)) {
        fprintf(stderr, "kungfu-notify-start: cannot get a cluster from '%s': %s\n", url.c_str(), err.c_str());
        return 1;
    }
    Stage stage;
    stage.version = atoi(f.get("version", "0").c_str());
    stage.cluster = c;
    const std::string msg = stage.json();
    TransportConfig cfg   = TransportConfig::from_env();
    Client client(PeerID(), cfg);
    int failed = 0;
    for (auto &r : c.runners) {
        if (!client.wait(r, parse_duration_seconds(f.get("timeout", "30s"), 30))) {
            fprintf(stderr, "runner %s is unreachable\n", r.str().c_str());
            ++failed;
            continue;
        }
        if (!client.send(r, CONN_CONTROL, "update", FLAG_NONE, msg.data(), msg.size(), &err)) {
            fprintf(stderr, "cannot notify %s: %s\n", r.str().c_str(), err.c_str());
            ++failed;
            continue;
        }
        printf("stage %d (%zu workers) sent to runner %s\n", stage.version, c.workers.size(), r.str().c_str());
    }
    return failed ? 1 : 0;
}

int config_client(const Flags &f)
{
    const std::string server = f.get("server", "http://127.0.0.1:9100/config");
    int status = 0;
    std::string resp, err;
    if (f.has("reset") || f.has("clear")) {
        const char *method = f.has("reset") ? "POST" : "DELETE";
        if (!http_do(method, server, "", &status, &resp, &err)) {
            fprintf(stderr, "%s %s failed: %s\n", method, server.c_str(), err.c_str());
            return 1;
        }
        printf("%s %s -> %d\n", method, server.c_str(), status);
        return status == 200 ? 0 : 1;
    }
    HostList hl;
    PortRange pr;
    if (!parse_host_list(f.get("H", "127.0.0.1:4"), &hl) || !parse_port_range(f.get("port-range", "10000-11000"), &pr)) {
        fprintf(stderr, "bad -H or -port-range\n");
        return 2;
    }
    const int cap = hl.cap() > 0 ? hl.cap() : 1;
    std::mt19937 rng(f.has("seed") ? (unsigned)atoi(f.get("seed", "0").c_str()) : (unsigned)getpid());
    const bool shuffle = f.has("shuffle");
    auto make = [&](int n, Cluster *c) {
        c->runners = hl.gen_runner_list((uint16_t)atoi(f.get("port", "38080").c_str()));
        if (!shuffle) return hl.gen_peer_list(n, pr, &c->workers, &err);
        PeerList all;
        if (!hl.gen_peer_list(cap, pr, &all, &err)) return false;
        std::shuffle(all.begin(), all.end(), rng);
        all.resize((size_t)n);
        c->workers = all;
        return true;
    };
    if (f.has("example")) {
        Cluster c;
        if (!make(cap, &c)) return 1;
        printf("%s\n", c.json().c_str());
        return 0;
    }
    const double period = parse_duration_seconds(f.get("period", "1s"), 1.0);
    const double ttl    = parse_duration_seconds(f.get("ttl", "0s"), 0.0);
    const auto t0 = std::chrono::steady_clock::now();
    for (int i = 0;; ++i) {
        Cluster c;
        const int n = 1 + (int)(rng() % (unsigned)cap);
        if (make(n, &c)) put_cluster(server, c);
        std::this_thread::sleep_for(std::chrono::duration<double>(period));
        const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        if (ttl > 0 && dt >= ttl) break;
    }
    return 0;
}
}  // namespace

int main(int argc, char **argv)
{
    static const char *const bools[] = {"reset", "clear", "example", "shuffle", nullptr};
    const Flags f = parse(argc, argv, bools);
    if (strstr(argv[0], "notify-start")) return notify_start(f);
    if (strstr(argv[0], "config-client")) return config_client(f);
    return test_util(f);
}
