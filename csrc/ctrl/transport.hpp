// Control-plane transport: framed, named messages over TCP (remote) or unix
// sockets (colocated), with a token handshake that keeps cluster versions from
// mixing during a resize, fd passing for CUDA VMM handles, a p2p blob RPC, and
// per-peer egress accounting.
//
// Parity: /root/reference/srcs/go/rchannel/connection/{connection.go,message.go},
// /root/reference/srcs/go/rchannel/client/{client.go,connection_pool.go},
// /root/reference/srcs/go/rchannel/server/{server.go,composed.go},
// /root/reference/srcs/go/rchannel/handler/{collective.go,p2p.go,control.go,ping.go}.
// Differences by design: bulk tensor data never travels here on the GPU path
// (kernels move it over NVLink); this plane carries bootstrap, consensus,
// small CPU collectives, the elastic protocol and memory handles.
#pragma once
#include <atomic>
#include <condition_variable>
#include <deque>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <string>
#include <thread>
#include <vector>

#include "core/plan.hpp"

namespace kf {

enum ConnType : uint16_t {
    CONN_PING       = 0,
    CONN_CONTROL    = 1,
    CONN_COLLECTIVE = 2,
    CONN_P2P        = 3,
    CONN_FD         = 4,  // SCM_RIGHTS descriptor passing (unix sockets only)
};

enum MsgFlags : uint32_t {
    FLAG_NONE           = 0,
    FLAG_IS_RESPONSE    = 2,
    FLAG_REQUEST_FAILED = 4,
};

struct TransportConfig {
    int conn_retry_count      = 500;
    double conn_retry_period  = 0.2;   // seconds
    double wait_runner_timeout = 300;  // seconds
    bool use_unix_sock        = true;
    static TransportConfig from_env();
};

// ---- low level socket helpers ---------------------------------------------------
// bind_ipv4 == 0 listens on every interface
int tcp_listen(uint16_t port, std::string *err, uint32_t bind_ipv4 = 0);
int unix_listen(const std::string &path, std::string *err);
int tcp_connect(uint32_t ipv4, uint16_t port);
int unix_connect(const std::string &path);
bool read_full(int fd, void *buf, size_t n);
bool write_full(int fd, const void *buf, size_t n);
bool send_fd(int sock, int fd);
int recv_fd(int sock);

struct MsgHeader {
    std::string name;
    uint32_t flags = 0;
    uint64_t len   = 0;
};
bool write_message(int fd, const std::string &name, uint32_t flags, const void *data, uint64_t len);
bool read_message_header(int fd, MsgHeader *h);

// ---- per-peer byte counters (the "monitor") ------------------------------------------
class NetMonitor
{
    mutable std::mutex mu_;
    std::map<PeerID, uint64_t> egress_, ingress_;
    std::map<PeerID, uint64_t> last_egress_;
    std::map<PeerID, double> rate_;
    double last_t_ = 0;

  public:
    void egress(const PeerID &p, uint64_t n);
    void ingress(const PeerID &p, uint64_t n);
    void sample();  // call periodically: turns deltas into rates (bytes/s)
    double egress_rate(const PeerID &p) const;
    uint64_t egress_total(const PeerID &p) const;
    std::string metrics_text() const;  // Prometheus style
};

// ---- rendezvous of named messages ----------------------------------------------------------
// Message bodies that arrive before their receiver is parked are held in pooled buffers:
// power-of-two size classes from 512 B, a few buffers kept per class, so the steady stream of
// 1 MiB chunks of a graph walk does not malloc / zero-fill / free once per message.
// Parity: /root/reference/srcs/go/rchannel/connection/byte_slice_pool.go:7-60 (size-keyed sync.Pool).
struct Body {
    std::unique_ptr<uint8_t[]> data;
    size_t cap = 0, len = 0;
};

class BodyPool
{
    static constexpr size_t kMinClass = 512, kKeepPerClass = 8, kMaxPooled = size_t(64) << 20;
    std::mutex mu_;
    std::map<size_t, std::vector<std::unique_ptr<uint8_t[]>>> free_;
    uint64_t hits_ = 0, misses_ = 0;
    static size_t size_class(size_t n);

  public:
    Body acquire(size_t len);
    void release(Body &&b);
    uint64_t hits() const { return hits_; }
    uint64_t misses() const { return misses_; }
    static BodyPool &instance();
};

class Mailbox
{
    struct Waiter {
        void *buf  = nullptr;
        size_t cap = 0;
        bool done  = false;
        bool busy  = false;  // a reader thread is filling buf right now
        size_t got = 0;
    };
    struct Box {
        std::deque<Body> ready;
        Waiter *waiter = nullptr;
    };
    std::mutex mu_;
    std::condition_variable cv_;
    std::map<std::string, Box> boxes_;
    std::map<std::string, std::deque<int>> fds_;
    bool closed_ = false;

    static std::string key(const PeerID &src, const std::string &name);

  public:
    // called by server reader threads: body of `len` bytes is still in `fd`
    bool deliver(const PeerID &src, const std::string &name, int fd, uint64_t len);
    // blocks until a message named `name` from `src` is available; returns its
    // size (<= cap) or -1 on timeout/close
    long recv_into(const PeerID &src, const std::string &name, void *buf, size_t cap, double timeout_s);
    void deliver_fd(const PeerID &src, const std::string &tag, int fd);
    int recv_fd(const PeerID &src, const std::string &tag, double timeout_s);
    void close();
    void reset();  // drop everything buffered (new session epoch)
};

// ---- versioned in-memory blob store (the model store PairAveraging pulls from) -------------
// Parity: /root/reference/srcs/go/store/{store.go,versionedstore.go,blob.go}
// Storage for published blobs. A model is published every step and each version lives until its last reader is
// done with it; a fresh 100 MB vector per step costs more in page faults than the copy into it. new_blob(n) hands
// out a vector of n bytes (contents unspecified) whose storage goes back to a small free list when the last
// shared_ptr to it dies, in whichever thread that happens.
std::shared_ptr<std::vector<uint8_t>> new_blob(size_t n);

class BlobStore
{
    std::mutex mu_;
    std::map<std::string, std::shared_ptr<std::vector<uint8_t>>> blobs_;

  public:
    void save(const std::string &name, const void *data, size_t n);
    void adopt(const std::string &name, std::shared_ptr<std::vector<uint8_t>> blob);  // no copy
    std::shared_ptr<std::vector<uint8_t>> get(const std::string &name);
};

class VersionedStore
{
    std::mutex mu_;
    size_t window_;
    std::deque<std::pair<std::string, std::shared_ptr<BlobStore>>> versions_;

  public:
    explicit VersionedStore(size_t window = 3) : window_(window) {}
    void save(const std::string &version, const std::string &name, const void *data, size_t n);
    void adopt(const std::string &version, const std::string &name, std::shared_ptr<std::vector<uint8_t>> blob);
    std::shared_ptr<std::vector<uint8_t>> get(const std::string &version, const std::string &name);
    size_t versions() const { return versions_.size(); }
};

// ---- server ---------------------------------------------------------------------------------------
using ControlHandler = std::function<void(const PeerID &src, const std::string &name, const std::string &body)>;

class Server
{
    PeerID self_;
    bool use_unix_;
    int tcp_fd_ = -1, unix_fd_ = -1;
    // Until the owner joins an epoch (set_token) collective connections are refused: a faster peer must not
    // be able to park messages in the mailbox that the owner's first reset() would then throw away.
    std::atomic<uint32_t> token_{0xFFFFFFFFu};
    std::atomic<bool> stopping_{false};
    std::thread tcp_thread_, unix_thread_;
    std::mutex conn_mu_;
    std::condition_variable conn_cv_;
    std::set<int> conns_;
    int active_ = 0;

    void accept_loop(int lfd);
    void serve(int fd);

  public:
    Mailbox mailbox;
    BlobStore store;
    VersionedStore vstore;
    NetMonitor *monitor = nullptr;
    ControlHandler on_control;

    Server(const PeerID &self, bool use_unix) : self_(self), use_unix_(use_unix) {}
    ~Server() { stop(); }
    bool start(std::string *err);
    void stop();
    void set_token(uint32_t t) { token_ = t; }
    uint32_t token() const { return token_; }
    const PeerID &self() const { return self_; }
};

// ---- client ---------------------------------------------------------------------------------------
class Client
{
    struct Conn {
        int fd = -1;
        std::mutex mu;
    };
    PeerID self_;
    TransportConfig cfg_;
    std::atomic<uint32_t> token_{0};
    std::mutex mu_;
    std::map<std::pair<PeerID, uint16_t>, std::shared_ptr<Conn>> pool_;

    int dial(const PeerID &dst, ConnType type, bool retry, std::string *err);
    std::shared_ptr<Conn> get(const PeerID &dst, ConnType type, std::string *err);
    void drop(const PeerID &dst, ConnType type);

  public:
    NetMonitor *monitor = nullptr;
    Client(const PeerID &self, const TransportConfig &cfg) : self_(self), cfg_(cfg) {}
    ~Client() { reset(); }
    void set_token(uint32_t t) { token_ = t; }
    bool send(const PeerID &dst, ConnType type, const std::string &name, uint32_t flags, const void *data,
              uint64_t len, std::string *err);
    // p2p RPC: fetch blob `name` (optionally of `version`) from dst's store
    bool request(const PeerID &dst, const std::string &version, const std::string &name, void *buf,
                 uint64_t len, std::string *err);
    bool pass_fd(const PeerID &dst, const std::string &tag, int fd, std::string *err);
    // round trip in seconds, <0 on failure
    double ping(const PeerID &dst);
    // ping until reachable
    bool wait(const PeerID &dst, double timeout_s);
    void reset();  // close every pooled connection
};

}  // namespace kf
