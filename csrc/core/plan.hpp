// Addressing and cluster description: PeerID, PeerList, HostSpec, Cluster.
//
// Parity: /root/reference/srcs/go/plan/{addr.go,id.go,peerlist.go,hostspec.go,
// cluster.go,interval.go}, /root/reference/srcs/go/plan/hostfile/hostfile.go.
// The JSON form of Cluster ({"Runners":[..],"Workers":[..]}, ids as "ip:port")
// is what the config server stores and what runners receive in an "update".
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <vector>

namespace kf {

struct PeerID {
    uint32_t ipv4 = 0;  // host byte order, a.b.c.d = a<<24|b<<16|c<<8|d
    uint16_t port = 0;

    bool operator==(const PeerID &o) const { return ipv4 == o.ipv4 && port == o.port; }
    bool operator!=(const PeerID &o) const { return !(*this == o); }
    bool operator<(const PeerID &o) const { return ipv4 != o.ipv4 ? ipv4 < o.ipv4 : port < o.port; }
    std::string str() const;       // "a.b.c.d:port"
    std::string host_str() const;  // "a.b.c.d"
    std::string sock_file() const; // unix socket path for colocated peers
    bool colocated_with(const PeerID &o) const { return ipv4 == o.ipv4; }
};

bool parse_ipv4(const std::string &s, uint32_t *out);
std::string format_ipv4(uint32_t ip);
bool parse_peer_id(const std::string &s, PeerID *out);

class PeerList : public std::vector<PeerID>
{
  public:
    using std::vector<PeerID>::vector;
    int rank_of(const PeerID &p) const;  // -1 if absent
    int local_rank_of(const PeerID &p) const;
    int local_size_of(const PeerID &p) const;
    int host_count() const;
    bool contains(const PeerID &p) const { return rank_of(p) >= 0; }
    bool disjoint(const PeerList &o) const;
    // (this \ o, o \ this)
    void diff(const PeerList &o, PeerList *removed, PeerList *added) const;
    PeerList on_host(uint32_t ipv4) const;
    // masters = first rank of each host; parents[i] = master rank of i's host
    void partition_by_host(std::vector<int> *masters, std::vector<int> *parents) const;
    std::string str() const;  // comma separated
    std::vector<uint8_t> bytes() const;
};
bool parse_peer_list(const std::string &s, PeerList *out);

struct PortRange {
    uint16_t begin = 10000, end = 11000;
    uint16_t cap() const { return end - begin; }
};
bool parse_port_range(const std::string &s, PortRange *out);

struct HostSpec {
    uint32_t ipv4 = 0;
    int slots     = 0;
    std::string public_addr;
    std::string str() const;
};
bool parse_host_spec(const std::string &s, HostSpec *out);  // ip[:slots[:public]]

class HostList : public std::vector<HostSpec>
{
  public:
    int cap() const;
    // np workers packed host by host onto consecutive ports of `pr`
    bool gen_peer_list(int np, const PortRange &pr, PeerList *out, std::string *err) const;
    PeerList gen_runner_list(uint16_t port) const;
    std::string str() const;
};
bool parse_host_list(const std::string &s, HostList *out);  // comma separated HostSpecs
// OpenMPI-style hostfile: "<host> slots=<n> [public_addr=<x>]", '#' comments
bool parse_hostfile(const std::string &text, HostList *out, std::string *err);

constexpr uint16_t kDefaultRunnerPort = 38080;

struct Cluster {
    PeerList runners;
    PeerList workers;

    bool validate(std::string *err) const;
    // shrink = drop the tail; grow = append to the least-loaded runner's host
    // on the next free port (/root/reference/srcs/go/plan/cluster.go:72-113)
    bool resize(int new_size, Cluster *out, std::string *err) const;
    std::string json() const;
    std::vector<uint8_t> bytes() const;  // canonical bytes for consensus
    bool operator==(const Cluster &o) const { return runners == o.runners && workers == o.workers; }
};
bool parse_cluster_json(const std::string &s, Cluster *out, std::string *err);

// One version of the cluster as sent to runners.
struct Stage {
    int version = 0;
    Cluster cluster;
    std::string json() const;
};
bool parse_stage_json(const std::string &s, Stage *out, std::string *err);

}  // namespace kf
