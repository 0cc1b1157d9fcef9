// Cluster description from a hosting platform's environment instead of
// kungfu-run. Huawei ModelArts batch jobs start one container per task and
// export the task index / count and one "host:port" per task; a worker that
// finds those variables (and no KUNGFU_SELF_SPEC) derives its peer list from
// them, see PeerConfig::from_env.
// Behaviour: /root/reference/srcs/go/platforms/modelarts/modelarts.go:14-115
// (a single task is 127.0.0.1:38888; host names must resolve to one address).
#include "core/platforms.hpp"

#include <arpa/inet.h>
#include <netdb.h>
#include <sys/socket.h>

#include <cstdlib>
#include <cstring>
#include <string>

namespace kf {

static bool resolve_ipv4(const std::string &host, uint32_t *ip, std::string *err)
{
    if (parse_ipv4(host, ip)) return true;
    addrinfo hints;
    memset(&hints, 0, sizeof(hints));
    hints.ai_family   = AF_INET;
    hints.ai_socktype = SOCK_STREAM;
    addrinfo *res     = nullptr;
    if (getaddrinfo(host.c_str(), nullptr, &hints, &res) != 0 || !res) {
        if (err) *err = "cannot resolve '" + host + "'";
        return false;
    }
    // exactly one address is expected, as in the reference
    uint32_t first = 0;
    int distinct   = 0;
    for (addrinfo *a = res; a; a = a->ai_next) {
        const uint32_t v = ntohl(reinterpret_cast<sockaddr_in *>(a->ai_addr)->sin_addr.s_addr);
        if (distinct == 0 || v != first) {
            if (distinct == 0) first = v;
            if (v != first || distinct == 0) ++distinct;
        }
    }
    freeaddrinfo(res);
    if (distinct != 1) {
        if (err) *err = "'" + host + "' resolves to several addresses; exactly 1 is expected";
        return false;
    }
    *ip = first;
    return true;
}

int modelarts_cluster(PeerList *workers, int *self_task, std::string *err)
{
    const char *idx = getenv("DLS_TASK_INDEX");
    const char *num = getenv("DLS_TASK_NUMBER");
    if (!idx || !num || !*idx || !*num) return 0;
    char *end   = nullptr;
    const int n = (int)strtol(num, &end, 10);
    if (*end || n < 1) {
        if (err) *err = std::string("DLS_TASK_NUMBER='") + num + "' is not a positive integer";
        return -1;
    }
    int me = (int)strtol(idx, &end, 10);
    if (*end) {
        if (err) *err = std::string("DLS_TASK_INDEX='") + idx + "' is not an integer";
        return -1;
    }
    if (n == 1 && me == 1) me = 0;  // the platform numbers a lone task 1
    if (me < 0 || me >= n) {
        if (err) *err = "DLS_TASK_INDEX must satisfy 0 <= index < DLS_TASK_NUMBER";
        return -1;
    }
    workers->clear();
    if (n == 1) {
        PeerID p;
        parse_peer_id("127.0.0.1:38888", &p);
        workers->push_back(p);
    } else {
        for (int t = 0; t < n; ++t) {
            const std::string key = "BATCH_CUSTOM" + std::to_string(t) + "_HOSTS";
            const char *hp        = getenv(key.c_str());
            if (!hp || !*hp) {
                if (err) *err = key + " not set";
                return -1;
            }
            const std::string v(hp);
            const auto colon = v.rfind(':');
            if (colon == std::string::npos) {
                if (err) *err = key + "='" + v + "' is not host:port";
                return -1;
            }
            const int port = atoi(v.c_str() + colon + 1);
            uint32_t ip    = 0;
            if (port <= 0 || port > 65535) {
                if (err) *err = key + "='" + v + "' has no valid port";
                return -1;
            }
            if (!resolve_ipv4(v.substr(0, colon), &ip, err)) return -1;
            workers->push_back(PeerID{ip, (uint16_t)port});
        }
    }
    if (self_task) *self_task = me;
    return 1;
}

}  // namespace kf
