#include "ctrl/transport.hpp"

#include <arpa/inet.h>
#include <errno.h>
#include <fcntl.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <signal.h>
#include <sys/socket.h>
#include <sys/stat.h>
#include <sys/un.h>
#include <unistd.h>

#include <chrono>
#include <cstring>
#include <new>
#include <sstream>

#include "core/log.hpp"

namespace kf {

static constexpr uint32_t kMagic = 0x4b46b200u;  // "KF" b200
// Bodies of the non-collective connection types are read into fresh zero-filled strings, so their announced
// length must be bounded BEFORE the allocation: 600 hostile "1 GiB ping" headers used to cost 60 GB of RSS.
static constexpr uint64_t kMaxSmallBody   = 64 << 10;
static constexpr uint64_t kMaxControlBody = 16 << 20;

TransportConfig TransportConfig::from_env()
{
    TransportConfig c;
    const std::string n = getenv_str("KUNGFU_CONFIG_CONN_RETRY_COUNT");
    if (!n.empty()) c.conn_retry_count = atoi(n.c_str());
    c.conn_retry_period   = parse_duration_seconds(getenv_str("KUNGFU_CONFIG_CONN_RETRY_PERIOD"), c.conn_retry_period);
    c.wait_runner_timeout = parse_duration_seconds(getenv_str("KUNGFU_CONFIG_WAIT_RUNNER_TIMEOUT"), c.wait_runner_timeout);
    c.use_unix_sock       = getenv_bool("KUNGFU_CONFIG_USE_UNIX_SOCK", true);
    return c;
}

// ---- sockets ----------------------------------------------------------------------------
static void ignore_sigpipe()
{
    static std::once_flag once;
    std::call_once(once, [] { signal(SIGPIPE, SIG_IGN); });
}

int tcp_listen(uint16_t port, std::string *err, uint32_t bind_ipv4)
{
    ignore_sigpipe();
    int fd = socket(AF_INET, SOCK_STREAM | SOCK_CLOEXEC, 0);
    if (fd < 0) {
        if (err) *err = std::string("socket: ") + strerror(errno);
        return -1;
    }
    int one = 1;
    setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
    sockaddr_in a;
    memset(&a, 0, sizeof(a));
    a.sin_family      = AF_INET;
    a.sin_addr.s_addr = htonl(bind_ipv4 ? bind_ipv4 : INADDR_ANY);
    a.sin_port        = htons(port);
    if (bind(fd, (sockaddr *)&a, sizeof(a)) != 0 || listen(fd, 512) != 0) {
        if (err) *err = "bind/listen :" + std::to_string(port) + ": " + strerror(errno);
        ::close(fd);
        return -1;
    }
    return fd;
}

int unix_listen(const std::string &path, std::string *err)
{
    ignore_sigpipe();
    int fd = socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
    if (fd < 0) {
        if (err) *err = std::string("socket: ") + strerror(errno);
        return -1;
    }
    unlink(path.c_str());
    sockaddr_un a;
    memset(&a, 0, sizeof(a));
    a.sun_family = AF_UNIX;
    strncpy(a.sun_path, path.c_str(), sizeof(a.sun_path) - 1);
    if (bind(fd, (sockaddr *)&a, sizeof(a)) != 0 || listen(fd, 512) != 0) {
        if (err) *err = "bind/listen " + path + ": " + strerror(errno);
        ::close(fd);
        return -1;
    }
    return fd;
}

int tcp_connect(uint32_t ipv4, uint16_t port)
{
    ignore_sigpipe();
    int fd = socket(AF_INET, SOCK_STREAM | SOCK_CLOEXEC, 0);
    if (fd < 0) return -1;
    sockaddr_in a;
    memset(&a, 0, sizeof(a));
    a.sin_family      = AF_INET;
    a.sin_addr.s_addr = htonl(ipv4);
    a.sin_port        = htons(port);
    if (connect(fd, (sockaddr *)&a, sizeof(a)) != 0) {
        ::close(fd);
        return -1;
    }
    int one = 1;
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
    return fd;
}

int unix_connect(const std::string &path)
{
    ignore_sigpipe();
    int fd = socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
    if (fd < 0) return -1;
    sockaddr_un a;
    memset(&a, 0, sizeof(a));
    a.sun_family = AF_UNIX;
    strncpy(a.sun_path, path.c_str(), sizeof(a.sun_path) - 1);
    if (connect(fd, (sockaddr *)&a, sizeof(a)) != 0) {
        ::close(fd);
        return -1;
    }
    return fd;
}

bool read_full(int fd, void *buf, size_t n)
{
    char *p = static_cast<char *>(buf);
    while (n > 0) {
        ssize_t r = ::read(fd, p, n);
        if (r == 0) return false;
        if (r < 0) {
            if (errno == EINTR) continue;
            return false;
        }
        p += r;
        n -= (size_t)r;
    }
    return true;
}

bool write_full(int fd, const void *buf, size_t n)
{
    const char *p = static_cast<const char *>(buf);
    while (n > 0) {
        ssize_t r = ::send(fd, p, n, MSG_NOSIGNAL);
        if (r < 0) {
            if (errno == EINTR) continue;
            return false;
        }
        p += r;
        n -= (size_t)r;
    }
    return true;
}

bool send_fd(int sock, int fd)
{
    char data = 'F';
    iovec iov{&data, 1};
    char ctrl[CMSG_SPACE(sizeof(int))];
    memset(ctrl, 0, sizeof(ctrl));
    msghdr msg;
    memset(&msg, 0, sizeof(msg));
    msg.msg_iov        = &iov;
    msg.msg_iovlen     = 1;
    msg.msg_control    = ctrl;
    msg.msg_controllen = sizeof(ctrl);
    cmsghdr *c         = CMSG_FIRSTHDR(&msg);
    c->cmsg_level      = SOL_SOCKET;
    c->cmsg_type       = SCM_RIGHTS;
    c->cmsg_len        = CMSG_LEN(sizeof(int));
    memcpy(CMSG_DATA(c), &fd, sizeof(int));
    for (;;) {
        ssize_t r = sendmsg(sock, &msg, MSG_NOSIGNAL);
        if (r < 0 && errno == EINTR) continue;
        return r == 1;
    }
}

int recv_fd(int sock)
{
    char data = 0;
    iovec iov{&data, 1};
    char ctrl[CMSG_SPACE(sizeof(int))];
    memset(ctrl, 0, sizeof(ctrl));
    msghdr msg;
    memset(&msg, 0, sizeof(msg));
    msg.msg_iov        = &iov;
    msg.msg_iovlen     = 1;
    msg.msg_control    = ctrl;
    msg.msg_controllen = sizeof(ctrl);
    for (;;) {
        ssize_t r = recvmsg(sock, &msg, MSG_CMSG_CLOEXEC);
        if (r < 0 && errno == EINTR) continue;
        if (r != 1) return -1;
        break;
    }
    for (cmsghdr *c = CMSG_FIRSTHDR(&msg); c; c = CMSG_NXTHDR(&msg, c)) {
        if (c->cmsg_level == SOL_SOCKET && c->cmsg_type == SCM_RIGHTS) {
            int fd;
            memcpy(&fd, CMSG_DATA(c), sizeof(int));
            return fd;
        }
    }
    return -1;
}

bool write_message(int fd, const std::string &name, uint32_t flags, const void *data, uint64_t len)
{
    // header and small bodies go out in one write
    std::vector<char> h(4 + name.size() + 4 + 8);
    uint32_t nl = (uint32_t)name.size();
    memcpy(h.data(), &nl, 4);
    memcpy(h.data() + 4, name.data(), name.size());
    memcpy(h.data() + 4 + name.size(), &flags, 4);
    memcpy(h.data() + 8 + name.size(), &len, 8);
    if (len > 0 && len <= 4096) {
        const size_t off = h.size();
        h.resize(off + len);
        memcpy(h.data() + off, data, len);
        return write_full(fd, h.data(), h.size());
    }
    if (!write_full(fd, h.data(), h.size())) return false;
    return len == 0 || write_full(fd, data, len);
}

bool read_message_header(int fd, MsgHeader *h)
{
    uint32_t nl;
    if (!read_full(fd, &nl, 4) || nl > (1u << 20)) return false;
    h->name.resize(nl);
    if (nl && !read_full(fd, &h->name[0], nl)) return false;
    if (!read_full(fd, &h->flags, 4)) return false;
    // no tensor of a model is anywhere near 64 GiB: a larger length is a corrupt or hostile header
    return read_full(fd, &h->len, 8) && h->len <= (uint64_t(1) << 36);
}

// ---- monitor ----------------------------------------------------------------------------
void NetMonitor::egress(const PeerID &p, uint64_t n)
{
    std::lock_guard<std::mutex> lk(mu_);
    egress_[p] += n;
}

void NetMonitor::ingress(const PeerID &p, uint64_t n)
{
    std::lock_guard<std::mutex> lk(mu_);
    ingress_[p] += n;
}

void NetMonitor::sample()
{
    std::lock_guard<std::mutex> lk(mu_);
    const double t = now_seconds();
    if (last_t_ > 0 && t > last_t_) {
        for (auto &kv : egress_) rate_[kv.first] = (double)(kv.second - last_egress_[kv.first]) / (t - last_t_);
    }
    last_egress_ = egress_;
    last_t_      = t;
}

double NetMonitor::egress_rate(const PeerID &p) const
{
    std::lock_guard<std::mutex> lk(mu_);
    auto it = rate_.find(p);
    return it == rate_.end() ? 0.0 : it->second;
}

uint64_t NetMonitor::egress_total(const PeerID &p) const
{
    std::lock_guard<std::mutex> lk(mu_);
    auto it = egress_.find(p);
    return it == egress_.end() ? 0 : it->second;
}

std::string NetMonitor::metrics_text() const
{
    std::lock_guard<std::mutex> lk(mu_);
    std::ostringstream os;
    for (auto &kv : egress_) os << "egress_total_bytes{peer=\"" << kv.first.str() << "\"} " << kv.second << "\n";
    for (auto &kv : ingress_) os << "ingress_total_bytes{peer=\"" << kv.first.str() << "\"} " << kv.second << "\n";
    for (auto &kv : rate_) os << "egress_rate_bytes_per_sec{peer=\"" << kv.first.str() << "\"} " << kv.second << "\n";
    return os.str();
}

// ---- mailbox ----------------------------------------------------------------------------
std::string Mailbox::key(const PeerID &src, const std::string &name) { return src.str() + "/" + name; }

size_t BodyPool::size_class(size_t n)
{
    size_t c = kMinClass;
    while (c < n) c <<= 1;
    return c;
}

BodyPool &BodyPool::instance()
{
    static BodyPool pool;
    return pool;
}

Body BodyPool::acquire(size_t len)
{
    Body b;
    b.len = len;
    if (len == 0) return b;
    b.cap = size_class(len);
    if (b.cap <= kMaxPooled) {
        std::lock_guard<std::mutex> lk(mu_);
        auto it = free_.find(b.cap);
        if (it != free_.end() && !it->second.empty()) {
            b.data = std::move(it->second.back());
            it->second.pop_back();
            ++hits_;
            return b;
        }
        ++misses_;
    }
    b.data.reset(new (std::nothrow) uint8_t[b.cap]);  // not value-initialised: the socket read fills it
    if (!b.data) b.cap = b.len = 0;                    // out of memory: the caller drops the connection
    return b;
}

void BodyPool::release(Body &&b)
{
    if (!b.data || b.cap > kMaxPooled) return;
    std::lock_guard<std::mutex> lk(mu_);
    auto &v = free_[b.cap];
    if (v.size() < kKeepPerClass) v.push_back(std::move(b.data));
}

bool Mailbox::deliver(const PeerID &src, const std::string &name, int fd, uint64_t len)
{
    const std::string k = key(src, name);
    Waiter *w           = nullptr;
    {
        std::unique_lock<std::mutex> lk(mu_);
        Box &b = boxes_[k];
        if (b.waiter && !b.waiter->busy && !b.waiter->done && b.ready.empty() && b.waiter->cap >= len) {
            w       = b.waiter;
            w->busy = true;
        }
    }
    if (w) {
        // zero-copy path: the receiver is already parked with its buffer
        const bool ok = len == 0 || read_full(fd, w->buf, len);
        std::unique_lock<std::mutex> lk(mu_);
        w->got  = len;
        w->busy = false;
        w->done = true;
        boxes_[k].waiter = nullptr;
        cv_.notify_all();
        return ok;
    }
    Body body = BodyPool::instance().acquire(len);
    if (len && !body.data) return false;
    if (len && !read_full(fd, body.data.get(), len)) {
        BodyPool::instance().release(std::move(body));
        return false;
    }
    std::unique_lock<std::mutex> lk(mu_);
    boxes_[k].ready.push_back(std::move(body));
    cv_.notify_all();
    return true;
}

long Mailbox::recv_into(const PeerID &src, const std::string &name, void *buf, size_t cap, double timeout_s)
{
    const std::string k = key(src, name);
    std::unique_lock<std::mutex> lk(mu_);
    Box &b = boxes_[k];
    Waiter w;
    w.buf = buf;
    w.cap = cap;
    const auto deadline = std::chrono::steady_clock::now() + std::chrono::duration<double>(timeout_s);
    for (;;) {
        if (w.done) return (long)w.got;
        if (!w.busy && !b.ready.empty()) {
            if (b.waiter == &w) b.waiter = nullptr;
            Body body = std::move(b.ready.front());
            b.ready.pop_front();
            const size_t n = std::min(cap, body.len), total = body.len;
            lk.unlock();
            if (n) memcpy(buf, body.data.get(), n);
            BodyPool::instance().release(std::move(body));
            return (long)total;
        }
        if (closed_) {
            if (b.waiter == &w && !w.busy) b.waiter = nullptr;
            if (!w.busy) return -1;
        }
        if (!b.waiter) b.waiter = &w;
        if (timeout_s <= 0) {
            cv_.wait(lk);
        } else if (cv_.wait_until(lk, deadline) == std::cv_status::timeout && !w.busy && !w.done) {
            if (b.waiter == &w) b.waiter = nullptr;
            if (!b.ready.empty()) continue;
            return -1;
        }
    }
}

void Mailbox::deliver_fd(const PeerID &src, const std::string &tag, int fd)
{
    std::lock_guard<std::mutex> lk(mu_);
    fds_[key(src, tag)].push_back(fd);
    cv_.notify_all();
}

int Mailbox::recv_fd(const PeerID &src, const std::string &tag, double timeout_s)
{
    const std::string k = key(src, tag);
    std::unique_lock<std::mutex> lk(mu_);
    const auto deadline = std::chrono::steady_clock::now() + std::chrono::duration<double>(timeout_s);
    for (;;) {
        auto it = fds_.find(k);
        if (it != fds_.end() && !it->second.empty()) {
            int fd = it->second.front();
            it->second.pop_front();
            return fd;
        }
        if (closed_) return -1;
        if (cv_.wait_until(lk, deadline) == std::cv_status::timeout) return -1;
    }
}

void Mailbox::close()
{
    std::lock_guard<std::mutex> lk(mu_);
    closed_ = true;
    cv_.notify_all();
}

void Mailbox::reset()
{
    std::lock_guard<std::mutex> lk(mu_);
    for (auto it = boxes_.begin(); it != boxes_.end();) {
        if (it->second.waiter) {
            it->second.ready.clear();
            ++it;
        } else {
            it = boxes_.erase(it);
        }
    }
    for (auto &kv : fds_)
        for (int fd : kv.second) ::close(fd);
    fds_.clear();
}

// ---- stores -----------------------------------------------------------------------------
namespace {
class BlobPool
{
    static constexpr size_t kMinPooled = size_t(1) << 20, kKeep = 4;
    std::mutex mu_;
    std::vector<std::vector<uint8_t> *> free_;

  public:
    std::vector<uint8_t> *take(size_t n)
    {
        if (n >= kMinPooled) {
            std::lock_guard<std::mutex> lk(mu_);
            for (size_t i = 0; i < free_.size(); ++i) {
                if (free_[i]->capacity() >= n && free_[i]->capacity() <= n + n / 8) {
                    std::vector<uint8_t> *v = free_[i];
                    free_.erase(free_.begin() + i);
                    v->resize(n);  // within capacity: no reallocation, no page faults on the recycled part
                    return v;
                }
            }
        }
        return new std::vector<uint8_t>(n);
    }
    void give(std::vector<uint8_t> *v)
    {
        if (v->capacity() >= kMinPooled) {
            std::lock_guard<std::mutex> lk(mu_);
            if (free_.size() < kKeep) {
                free_.push_back(v);
                return;
            }
        }
        delete v;
    }
};
BlobPool &blob_pool()
{
    static BlobPool *p = new BlobPool();  // never destroyed: blobs may outlive static destructors
    return *p;
}
}  // namespace

std::shared_ptr<std::vector<uint8_t>> new_blob(size_t n)
{
    return std::shared_ptr<std::vector<uint8_t>>(blob_pool().take(n), [](std::vector<uint8_t> *v) { blob_pool().give(v); });
}

void BlobStore::save(const std::string &name, const void *data, size_t n)
{
    // A vector of its own per save (new, or storage that no reader holds any more): readers
    // holding the old shared_ptr never see a torn blob (the reference copied in place without
    // a lock, /root/reference/srcs/go/store/blob.go:20-26).
    auto v = new_blob(n);
    if (n) memcpy(v->data(), data, n);
    std::lock_guard<std::mutex> lk(mu_);
    blobs_[name] = v;
}

void BlobStore::adopt(const std::string &name, std::shared_ptr<std::vector<uint8_t>> blob)
{
    std::shared_ptr<std::vector<uint8_t>> old;  // dies after the lock is released: its deleter takes the pool's lock
    {
        std::lock_guard<std::mutex> lk(mu_);
        auto &slot = blobs_[name];
        old.swap(slot);
        slot = std::move(blob);
    }
}

std::shared_ptr<std::vector<uint8_t>> BlobStore::get(const std::string &name)
{
    std::lock_guard<std::mutex> lk(mu_);
    auto it = blobs_.find(name);
    return it == blobs_.end() ? nullptr : it->second;
}

void VersionedStore::adopt(const std::string &version, const std::string &name, std::shared_ptr<std::vector<uint8_t>> blob)
{
    std::shared_ptr<BlobStore> s;
    {
        std::lock_guard<std::mutex> lk(mu_);
        for (auto &kv : versions_)
            if (kv.first == version) s = kv.second;
        if (!s) {
            s = std::make_shared<BlobStore>();
            versions_.push_back({version, s});
            while (versions_.size() > window_) versions_.pop_front();
        }
    }
    s->adopt(name, std::move(blob));
}

void VersionedStore::save(const std::string &version, const std::string &name, const void *data, size_t n)
{
    auto v = new_blob(n);
    if (n) memcpy(v->data(), data, n);
    adopt(version, name, std::move(v));
}

std::shared_ptr<std::vector<uint8_t>> VersionedStore::get(const std::string &version, const std::string &name)
{
    std::shared_ptr<BlobStore> s;
    {
        std::lock_guard<std::mutex> lk(mu_);
        for (auto &kv : versions_)
            if (kv.first == version) s = kv.second;
    }
    return s ? s->get(name) : nullptr;
}

// ---- server -----------------------------------------------------------------------------
struct WireHello {
    uint32_t magic;
    uint16_t type;
    uint16_t src_port;
    uint32_t src_ipv4;
    uint32_t token;
};

bool Server::start(std::string *err)
{
    // loopback aliases (multi-host emulation on one machine) bind their own address
    tcp_fd_ = tcp_listen(self_.port, err, (self_.ipv4 >> 24) == 127 ? self_.ipv4 : 0);
    if (tcp_fd_ < 0) return false;
    if (use_unix_) {
        unix_fd_ = unix_listen(self_.sock_file(), err);
        if (unix_fd_ < 0) {
            ::close(tcp_fd_);
            tcp_fd_ = -1;
            return false;
        }
    }
    tcp_thread_ = std::thread([this] { accept_loop(tcp_fd_); });
    if (unix_fd_ >= 0) unix_thread_ = std::thread([this] { accept_loop(unix_fd_); });
    return true;
}

void Server::stop()
{
    if (stopping_.exchange(true)) return;
    mailbox.close();
    if (tcp_fd_ >= 0) shutdown(tcp_fd_, SHUT_RDWR), ::close(tcp_fd_);
    if (unix_fd_ >= 0) {
        shutdown(unix_fd_, SHUT_RDWR), ::close(unix_fd_);
        unlink(self_.sock_file().c_str());
    }
    if (tcp_thread_.joinable()) tcp_thread_.join();
    if (unix_thread_.joinable()) unix_thread_.join();
    std::unique_lock<std::mutex> lk(conn_mu_);
    for (int fd : conns_) shutdown(fd, SHUT_RDWR);
    conn_cv_.wait_for(lk, std::chrono::seconds(5), [this] { return active_ == 0; });
}

void Server::accept_loop(int lfd)
{
    while (!stopping_) {
        int fd = accept4(lfd, nullptr, nullptr, SOCK_CLOEXEC);
        if (fd < 0) {
            if (errno == EINTR) continue;
            break;
        }
        {
            std::lock_guard<std::mutex> lk(conn_mu_);
            conns_.insert(fd);
            ++active_;
        }
        std::thread([this, fd] {
            serve(fd);
            ::close(fd);
            std::lock_guard<std::mutex> lk(conn_mu_);
            conns_.erase(fd);
            --active_;
            conn_cv_.notify_all();
        }).detach();
    }
}

void Server::serve(int fd)
{
    WireHello hello;
    if (!read_full(fd, &hello, sizeof(hello)) || hello.magic != kMagic) return;
    const uint32_t my_token = token_;
    if (!write_full(fd, &my_token, 4)) return;
    const ConnType type = (ConnType)hello.type;
    const PeerID src{hello.src_ipv4, hello.src_port};
    if ((type == CONN_COLLECTIVE || type == CONN_FD) && hello.token != my_token) return;  // stale epoch
    for (;;) {
        MsgHeader h;
        if (!read_message_header(fd, &h)) return;
        if (monitor) monitor->ingress(src, h.len);
        switch (type) {
        case CONN_COLLECTIVE:
            if (!mailbox.deliver(src, h.name, fd, h.len)) return;
            break;
        case CONN_FD: {
            int got = recv_fd(fd);
            if (got < 0) return;
            mailbox.deliver_fd(src, h.name, got);
            break;
        }
        case CONN_P2P: {
            if (h.len > kMaxSmallBody) return;  // a version label, not a payload
            std::string version(h.len, '\0');
            if (h.len && !read_full(fd, &version[0], h.len)) return;
            auto blob = version.empty() ? store.get(h.name) : vstore.get(version, h.name);
            bool ok;
            if (blob) {
                ok = write_message(fd, h.name, FLAG_IS_RESPONSE, blob->data(), blob->size());
                if (monitor) monitor->egress(src, blob->size());
            } else {
                ok = write_message(fd, h.name, FLAG_IS_RESPONSE | FLAG_REQUEST_FAILED, nullptr, 0);
            }
            if (!ok) return;
            break;
        }
        case CONN_CONTROL: {
            if (h.len > kMaxControlBody) return;  // a stage document (cluster JSON)
            std::string body(h.len, '\0');
            if (h.len && !read_full(fd, &body[0], h.len)) return;
            if (on_control) on_control(src, h.name, body);
            break;
        }
        case CONN_PING: {
            if (h.len > kMaxSmallBody) return;
            std::string body(h.len, '\0');
            if (h.len && !read_full(fd, &body[0], h.len)) return;
            if (!write_message(fd, h.name, FLAG_IS_RESPONSE, body.data(), body.size())) return;
            break;
        }
        default: return;
        }
    }
}

// ---- client -----------------------------------------------------------------------------
int Client::dial(const PeerID &dst, ConnType type, bool retry, std::string *err)
{
    // Total patience = retry count x retry period (500 x 200 ms by default, like the reference). A peer that
    // answers with another epoch's token is alive and merely has not reached the same membership yet (it
    // is still starting, or still finishing the previous epoch): poll it quickly instead of sleeping a
    // whole retry period, so start-up and resizes do not pay 200 ms for losing a race by a microsecond.
    const double budget = retry ? std::max(1, cfg_.conn_retry_count) * cfg_.conn_retry_period : 0.0;
    const double t0     = now_seconds();
    for (;;) {
        int fd        = -1;
        bool mismatch = false;
        if (cfg_.use_unix_sock && dst.colocated_with(self_)) fd = unix_connect(dst.sock_file());
        if (fd < 0 && type != CONN_FD) fd = tcp_connect(dst.ipv4, dst.port);
        if (fd >= 0) {
            WireHello hello{kMagic, (uint16_t)type, self_.port, self_.ipv4, token_.load()};
            uint32_t remote = 0;
            if (write_full(fd, &hello, sizeof(hello)) && read_full(fd, &remote, 4)) {
                if ((type != CONN_COLLECTIVE && type != CONN_FD) || remote == hello.token) return fd;
                mismatch = true;
                KF_DEBUG("token mismatch dialing %s: mine %u theirs %u", dst.str().c_str(), hello.token, remote);
            }
            ::close(fd);
        }
        const double left = budget - (now_seconds() - t0);
        if (left <= 0) break;
        std::this_thread::sleep_for(
            std::chrono::duration<double>(std::min(left, mismatch ? std::min(0.005, cfg_.conn_retry_period) : cfg_.conn_retry_period)));
    }
    if (err) *err = "cannot connect to " + dst.str();
    return -1;
}

std::shared_ptr<Client::Conn> Client::get(const PeerID &dst, ConnType type, std::string *err)
{
    std::shared_ptr<Conn> c;
    {
        std::lock_guard<std::mutex> lk(mu_);
        auto &slot = pool_[{dst, (uint16_t)type}];
        if (!slot) slot = std::make_shared<Conn>();
        c = slot;
    }
    std::lock_guard<std::mutex> lk(c->mu);
    if (c->fd < 0) {
        c->fd = dial(dst, type, type != CONN_PING, err);
        if (c->fd < 0) return nullptr;
    }
    return c;
}

void Client::drop(const PeerID &dst, ConnType type)
{
    std::lock_guard<std::mutex> lk(mu_);
    auto it = pool_.find({dst, (uint16_t)type});
    if (it != pool_.end()) {
        std::lock_guard<std::mutex> lk2(it->second->mu);
        if (it->second->fd >= 0) ::close(it->second->fd);
        it->second->fd = -1;
    }
}

bool Client::send(const PeerID &dst, ConnType type, const std::string &name, uint32_t flags, const void *data,
                  uint64_t len, std::string *err)
{
    for (int attempt = 0; attempt < 2; ++attempt) {
        auto c = get(dst, type, err);
        if (!c) return false;
        bool ok;
        {
            std::lock_guard<std::mutex> lk(c->mu);
            ok = c->fd >= 0 && write_message(c->fd, name, flags, data, len);
        }
        if (ok) {
            if (monitor) monitor->egress(dst, len);
            return true;
        }
        drop(dst, type);  // stale pooled connection (peer restarted): redial once
    }
    if (err) *err = "send to " + dst.str() + " failed";
    return false;
}

bool Client::request(const PeerID &dst, const std::string &version, const std::string &name, void *buf,
                     uint64_t len, std::string *err)
{
    auto c = get(dst, CONN_P2P, err);
    if (!c) return false;
    std::lock_guard<std::mutex> lk(c->mu);
    MsgHeader h;
    if (c->fd < 0 || !write_message(c->fd, name, FLAG_NONE, version.data(), version.size()) ||
        !read_message_header(c->fd, &h)) {
        if (c->fd >= 0) ::close(c->fd);
        c->fd = -1;
        if (err) *err = "p2p request to " + dst.str() + " failed (connection)";
        return false;
    }
    if (h.flags & FLAG_REQUEST_FAILED) {
        if (err) *err = "peer " + dst.str() + " has no blob named '" + name + "'";
        return false;
    }
    if (h.len != len) {
        // wrong size: the stream can no longer be trusted to be aligned with what we expect - drop the
        // connection (the next request dials again) instead of allocating a sink of the announced size
        ::close(c->fd);
        c->fd = -1;
        if (err) *err = "blob '" + name + "' has " + std::to_string(h.len) + " bytes, expected " + std::to_string(len);
        return false;
    }
    if (len && !read_full(c->fd, buf, len)) {
        ::close(c->fd);
        c->fd = -1;
        if (err) *err = "p2p response truncated";
        return false;
    }
    return true;
}

bool Client::pass_fd(const PeerID &dst, const std::string &tag, int fd, std::string *err)
{
    if (!dst.colocated_with(self_)) {
        if (err) *err = "descriptors can only be passed to a colocated peer";
        return false;
    }
    auto c = get(dst, CONN_FD, err);
    if (!c) return false;
    std::lock_guard<std::mutex> lk(c->mu);
    if (c->fd < 0 || !write_message(c->fd, tag, FLAG_NONE, nullptr, 0) || !send_fd(c->fd, fd)) {
        if (err) *err = "passing fd to " + dst.str() + " failed";
        return false;
    }
    return true;
}

double Client::ping(const PeerID &dst)
{
    std::string err;
    int fd = dial(dst, CONN_PING, false, &err);
    if (fd < 0) return -1;
    const double t0 = now_seconds();
    MsgHeader h;
    char payload[8] = "kfping";
    bool ok         = write_message(fd, "ping", FLAG_NONE, payload, sizeof(payload)) && read_message_header(fd, &h);
    if (ok && h.len > kMaxSmallBody) ok = false;  // an echo of our 8 bytes, nothing bigger
    if (ok && h.len) {
        std::vector<char> b(h.len);
        ok = read_full(fd, b.data(), h.len);
    }
    ::close(fd);
    return ok ? now_seconds() - t0 : -1;
}

bool Client::wait(const PeerID &dst, double timeout_s)
{
    return poll_until([&] { return ping(dst) >= 0; }, 0.1, timeout_s);
}

void Client::reset()
{
    std::lock_guard<std::mutex> lk(mu_);
    for (auto &kv : pool_) {
        std::lock_guard<std::mutex> lk2(kv.second->mu);
        if (kv.second->fd >= 0) ::close(kv.second->fd);
        kv.second->fd = -1;
    }
    pool_.clear();
}

}  // namespace kf
