// Cluster description from a hosting platform's environment instead of
// kungfu-run: Huawei ModelArts batch jobs export the task index/count and a
// per-task host name pattern.
// Parity: /root/reference/srcs/go/platforms/modelarts/modelarts.go:25-50.
#include <cstdlib>
#include <string>

#include "core/plan.hpp"

namespace kf {

// Returns false when the process is not running under ModelArts.
// DLS_TASK_INDEX, DLS_TASK_NUMBER, BATCH_CUSTOMx_HOSTS ("host:port" per task).
bool modelarts_cluster(int slots_per_host, const PortRange &pr, PeerList *workers, int *self_task)
{
    const char *idx = getenv("DLS_TASK_INDEX");
    const char *num = getenv("DLS_TASK_NUMBER");
    if (!idx || !num) return false;
    const int n = atoi(num);
    workers->clear();
    for (int t = 0; t < n; ++t) {
        const std::string key = "BATCH_CUSTOM" + std::to_string(t) + "_HOSTS";
        const char *hp        = getenv(key.c_str());
        if (!hp) return false;
        std::string host(hp);
        auto colon = host.find(':');
        if (colon != std::string::npos) host = host.substr(0, colon);
        uint32_t ip;
        if (!parse_ipv4(host, &ip)) return false;  // the launcher resolves names before exporting
        for (int s = 0; s < slots_per_host; ++s) workers->push_back(PeerID{ip, (uint16_t)(pr.begin + s)});
    }
    if (self_task) *self_task = atoi(idx);
    return true;
}

}  // namespace kf
