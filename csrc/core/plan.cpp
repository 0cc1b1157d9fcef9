#include "core/plan.hpp"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <set>
#include <sstream>

#include "core/json.hpp"

namespace kf {

static std::vector<std::string> split(const std::string &s, char sep)
{
    std::vector<std::string> out;
    std::string cur;
    for (char c : s) {
        if (c == sep) {
            out.push_back(cur);
            cur.clear();
        } else {
            cur += c;
        }
    }
    out.push_back(cur);
    return out;
}

static bool parse_uint(const std::string &s, unsigned long max, unsigned long *out)
{
    if (s.empty() || s.size() > 10) return false;
    for (char c : s)
        if (c < '0' || c > '9') return false;
    unsigned long v = strtoul(s.c_str(), nullptr, 10);
    if (v > max) return false;
    *out = v;
    return true;
}

bool parse_ipv4(const std::string &s, uint32_t *out)
{
    auto parts = split(s, '.');
    if (parts.size() != 4) return false;
    uint32_t ip = 0;
    for (auto &p : parts) {
        unsigned long v;
        if (!parse_uint(p, 255, &v)) return false;
        ip = (ip << 8) | (uint32_t)v;
    }
    *out = ip;
    return true;
}

std::string format_ipv4(uint32_t ip)
{
    char buf[32];
    snprintf(buf, sizeof(buf), "%u.%u.%u.%u", ip >> 24, (ip >> 16) & 255, (ip >> 8) & 255, ip & 255);
    return buf;
}

std::string PeerID::str() const { return format_ipv4(ipv4) + ":" + std::to_string(port); }
std::string PeerID::host_str() const { return format_ipv4(ipv4); }
std::string PeerID::sock_file() const
{
    // the address is part of the name so that several "hosts" emulated on
    // loopback aliases (127.0.0.x) of one machine do not collide
    return "/tmp/kungfu-b200-" + format_ipv4(ipv4) + "-" + std::to_string(port) + ".sock";
}

bool parse_peer_id(const std::string &s, PeerID *out)
{
    auto parts = split(s, ':');
    if (parts.size() != 2) return false;
    unsigned long port;
    if (!parse_ipv4(parts[0], &out->ipv4) || !parse_uint(parts[1], 65535, &port)) return false;
    out->port = (uint16_t)port;
    return true;
}

int PeerList::rank_of(const PeerID &p) const
{
    for (size_t i = 0; i < size(); ++i)
        if ((*this)[i] == p) return (int)i;
    return -1;
}

int PeerList::local_rank_of(const PeerID &p) const
{
    int k = 0;
    for (auto &q : *this) {
        if (q == p) return k;
        if (q.ipv4 == p.ipv4) ++k;
    }
    return -1;
}

int PeerList::local_size_of(const PeerID &p) const
{
    int k = 0;
    for (auto &q : *this)
        if (q.ipv4 == p.ipv4) ++k;
    return k;
}

int PeerList::host_count() const
{
    std::set<uint32_t> hosts;
    for (auto &q : *this) hosts.insert(q.ipv4);
    return (int)hosts.size();
}

bool PeerList::disjoint(const PeerList &o) const
{
    for (auto &q : *this)
        if (o.contains(q)) return false;
    return true;
}

void PeerList::diff(const PeerList &o, PeerList *removed, PeerList *added) const
{
    for (auto &q : *this)
        if (!o.contains(q)) removed->push_back(q);
    for (auto &q : o)
        if (!contains(q)) added->push_back(q);
}

PeerList PeerList::on_host(uint32_t ipv4) const
{
    PeerList out;
    for (auto &q : *this)
        if (q.ipv4 == ipv4) out.push_back(q);
    return out;
}

void PeerList::partition_by_host(std::vector<int> *masters, std::vector<int> *parents) const
{
    std::map<uint32_t, int> host_master;
    parents->assign(size(), 0);
    for (size_t r = 0; r < size(); ++r) {
        auto it = host_master.find((*this)[r].ipv4);
        if (it == host_master.end()) {
            host_master[(*this)[r].ipv4] = (int)r;
            masters->push_back((int)r);
            (*parents)[r] = (int)r;
        } else {
            (*parents)[r] = it->second;
        }
    }
}

std::string PeerList::str() const
{
    std::string s;
    for (size_t i = 0; i < size(); ++i) {
        if (i) s += ",";
        s += (*this)[i].str();
    }
    return s;
}

std::vector<uint8_t> PeerList::bytes() const
{
    std::vector<uint8_t> b;
    for (auto &q : *this) {
        for (int k = 3; k >= 0; --k) b.push_back((q.ipv4 >> (8 * k)) & 255);
        b.push_back(q.port >> 8);
        b.push_back(q.port & 255);
    }
    return b;
}

bool parse_peer_list(const std::string &s, PeerList *out)
{
    out->clear();
    if (s.empty()) return true;
    for (auto &part : split(s, ',')) {
        PeerID p;
        if (!parse_peer_id(part, &p)) return false;
        out->push_back(p);
    }
    return true;
}

bool parse_port_range(const std::string &s, PortRange *out)
{
    auto parts = split(s, '-');
    unsigned long b, e;
    if (parts.size() != 2 || !parse_uint(parts[0], 65535, &b) || !parse_uint(parts[1], 65535, &e) || e < b)
        return false;
    out->begin = (uint16_t)b;
    out->end   = (uint16_t)e;
    return true;
}

std::string HostSpec::str() const
{
    return format_ipv4(ipv4) + ":" + std::to_string(slots) + ":" + public_addr;
}

bool parse_host_spec(const std::string &s, HostSpec *out)
{
    auto parts = split(s, ':');
    if (parts.empty() || parts.size() > 3) return false;
    if (!parse_ipv4(parts[0], &out->ipv4)) return false;
    out->slots       = 1;
    out->public_addr = parts[0];
    if (parts.size() >= 2) {
        unsigned long n;
        if (!parse_uint(parts[1], 1 << 20, &n)) return false;
        out->slots = (int)n;
    }
    if (parts.size() == 3) out->public_addr = parts[2];
    return true;
}

int HostList::cap() const
{
    int c = 0;
    for (auto &h : *this) c += h.slots;
    return c;
}

bool HostList::gen_peer_list(int np, const PortRange &pr, PeerList *out, std::string *err) const
{
    out->clear();
    if (cap() < np) {
        if (err) *err = "cluster capacity " + std::to_string(cap()) + " < np " + std::to_string(np);
        return false;
    }
    for (auto &h : *this) {
        for (int j = 0; j < h.slots; ++j) {
            if ((int)out->size() >= np) return true;
            if (j >= pr.cap()) {
                if (err) *err = "port range too small for host " + format_ipv4(h.ipv4);
                return false;
            }
            out->push_back(PeerID{h.ipv4, (uint16_t)(pr.begin + j)});
        }
    }
    return true;
}

PeerList HostList::gen_runner_list(uint16_t port) const
{
    PeerList out;
    for (auto &h : *this) out.push_back(PeerID{h.ipv4, port});
    return out;
}

std::string HostList::str() const
{
    std::string s;
    for (size_t i = 0; i < size(); ++i) {
        if (i) s += ",";
        s += (*this)[i].str();
    }
    return s;
}

bool parse_host_list(const std::string &s, HostList *out)
{
    out->clear();
    if (s.empty()) return true;  // no hosts: valid and empty, like the peer list
    for (auto &part : split(s, ',')) {
        HostSpec h;
        if (!parse_host_spec(part, &h)) return false;
        out->push_back(h);
    }
    return true;
}

bool parse_hostfile(const std::string &text, HostList *out, std::string *err)
{
    out->clear();
    std::istringstream in(text);
    std::string line;
    int lineno = 0;
    while (std::getline(in, line)) {
        ++lineno;
        auto hash = line.find('#');
        if (hash != std::string::npos) line = line.substr(0, hash);
        std::istringstream ls(line);
        std::vector<std::string> tok;
        std::string t;
        while (ls >> t) tok.push_back(t);
        if (tok.empty()) continue;
        HostSpec h;
        if (!parse_ipv4(tok[0], &h.ipv4)) {
            if (err) *err = "hostfile line " + std::to_string(lineno) + ": bad host '" + tok[0] + "'";
            return false;
        }
        h.slots       = 1;
        h.public_addr = tok[0];
        for (size_t i = 1; i < tok.size(); ++i) {
            auto eq = tok[i].find('=');
            if (eq == std::string::npos) {
                if (err) *err = "hostfile line " + std::to_string(lineno) + ": expected key=value";
                return false;
            }
            const std::string k = tok[i].substr(0, eq), v = tok[i].substr(eq + 1);
            if (k == "slots") {
                unsigned long n;
                if (!parse_uint(v, 1 << 20, &n)) {
                    if (err) *err = "hostfile line " + std::to_string(lineno) + ": bad slots";
                    return false;
                }
                h.slots = (int)n;
            } else if (k == "public_addr") {
                h.public_addr = v;
            }
        }
        out->push_back(h);
    }
    return true;
}

bool Cluster::validate(std::string *err) const
{
    std::set<PeerID> seen;
    for (auto &w : workers) {
        if (!seen.insert(w).second) {
            if (err) *err = "duplicated worker " + w.str();
            return false;
        }
    }
    std::set<uint32_t> hosts;
    for (auto &r : runners) {
        if (!hosts.insert(r.ipv4).second) {
            if (err) *err = "duplicated runner host " + r.host_str();
            return false;
        }
        if (seen.count(r)) {
            if (err) *err = "runner " + r.str() + " collides with a worker";
            return false;
        }
    }
    for (auto &w : workers) {
        if (!hosts.count(w.ipv4)) {
            if (err) *err = "worker " + w.str() + " has no runner on its host";
            return false;
        }
    }
    return true;
}

bool Cluster::resize(int new_size, Cluster *out, std::string *err) const
{
    if (new_size < 0) {
        if (err) *err = "negative size";
        return false;
    }
    *out = *this;
    if ((int)out->workers.size() > new_size) out->workers.resize(new_size);
    while ((int)out->workers.size() < new_size) {
        if (out->runners.empty()) {
            if (err) *err = "no runner to grow on";
            return false;
        }
        std::map<uint32_t, int> used;
        for (auto &r : out->runners) used[r.ipv4] = 0;
        for (auto &w : out->workers) used[w.ipv4]++;
        uint32_t host = out->runners[0].ipv4;
        for (auto &r : out->runners)
            if (used[r.ipv4] < used[host]) host = r.ipv4;
        uint16_t port = 0;
        for (auto &w : out->workers)
            if (w.ipv4 == host && port <= w.port) port = w.port + 1;
        if (port == 0) {
            // a host without workers: start where the job's other hosts start (a job launched with -port-range keeps
            // its ports on new hosts too); the default range only for an empty cluster, as in the reference
            for (auto &w : out->workers)
                if (port == 0 || w.port < port) port = w.port;
            if (port == 0) port = PortRange().begin;
        }
        out->workers.push_back(PeerID{host, port});
    }
    return true;
}

static std::string id_array_json(const PeerList &l)
{
    std::string s = "[";
    for (size_t i = 0; i < l.size(); ++i) {
        if (i) s += ",";
        s += json_quote(l[i].str());
    }
    return s + "]";
}

std::string Cluster::json() const
{
    return "{\"Runners\":" + id_array_json(runners) + ",\"Workers\":" + id_array_json(workers) + "}";
}

std::vector<uint8_t> Cluster::bytes() const
{
    std::vector<uint8_t> b = runners.bytes();
    b.push_back(0xff);
    auto w = workers.bytes();
    b.insert(b.end(), w.begin(), w.end());
    return b;
}

static bool id_array_from(const Json *j, PeerList *out, std::string *err)
{
    out->clear();
    if (!j || j->kind == Json::Null) return true;
    if (j->kind != Json::Arr) {
        if (err) *err = "expected an array of \"ip:port\"";
        return false;
    }
    for (auto &e : j->arr) {
        PeerID p;
        if (e.kind != Json::Str || !parse_peer_id(e.str, &p)) {
            if (err) *err = "bad peer id in cluster json";
            return false;
        }
        out->push_back(p);
    }
    return true;
}

static bool cluster_from(const Json &j, Cluster *out, std::string *err)
{
    if (j.kind != Json::Obj) {
        if (err) *err = "cluster json must be an object";
        return false;
    }
    return id_array_from(j.get("Runners"), &out->runners, err) &&
           id_array_from(j.get("Workers"), &out->workers, err);
}

bool parse_cluster_json(const std::string &s, Cluster *out, std::string *err)
{
    Json j;
    if (!JsonParser(s).parse(&j)) {
        if (err) *err = "malformed json";
        return false;
    }
    return cluster_from(j, out, err);
}

std::string Stage::json() const
{
    return "{\"Version\":" + std::to_string(version) + ",\"Cluster\":" + cluster.json() + "}";
}

bool parse_stage_json(const std::string &s, Stage *out, std::string *err)
{
    Json j;
    if (!JsonParser(s).parse(&j) || j.kind != Json::Obj) {
        if (err) *err = "malformed stage json";
        return false;
    }
    const Json *v = j.get("Version");
    const Json *c = j.get("Cluster");
    if (!v || v->kind != Json::Num || !c) {
        if (err) *err = "stage json needs Version and Cluster";
        return false;
    }
    out->version = (int)v->num;
    return cluster_from(*c, &out->cluster, err);
}

}  // namespace kf
