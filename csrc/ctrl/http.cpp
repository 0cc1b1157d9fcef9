#include "ctrl/http.hpp"

#include <arpa/inet.h>
#include <netdb.h>
#include <sys/socket.h>
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <cstring>
#include <fstream>
#include <sstream>

#include "core/plan.hpp"
#include "ctrl/transport.hpp"

namespace kf {

static std::string lower(std::string s)
{
    std::transform(s.begin(), s.end(), s.begin(), ::tolower);
    return s;
}

// Reads "head\r\n\r\n" + Content-Length body. Returns false on malformed input.
static bool read_http(int fd, std::string *start_line, std::map<std::string, std::string> *headers,
                      std::string *body)
{
    std::string buf;
    char tmp[4096];
    size_t head_end = std::string::npos;
    while (head_end == std::string::npos) {
        ssize_t r = ::read(fd, tmp, sizeof(tmp));
        if (r <= 0) return false;
        buf.append(tmp, (size_t)r);
        head_end = buf.find("\r\n\r\n");
        if (buf.size() > (1u << 20)) return false;
    }
    std::istringstream head(buf.substr(0, head_end));
    std::getline(head, *start_line);
    if (!start_line->empty() && start_line->back() == '\r') start_line->pop_back();
    std::string line;
    while (std::getline(head, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        auto c = line.find(':');
        if (c == std::string::npos) continue;
        std::string v = line.substr(c + 1);
        while (!v.empty() && v[0] == ' ') v.erase(0, 1);
        (*headers)[lower(line.substr(0, c))] = v;
    }
    *body      = buf.substr(head_end + 4);
    size_t len = 0;
    bool has_len = false;
    auto it = headers->find("content-length");
    if (it != headers->end()) {
        if (it->second.empty() || it->second[0] == '-') return false;
        len     = (size_t)strtoull(it->second.c_str(), nullptr, 10);
        has_len = true;
        if (len > (size_t(64) << 20)) return false;  // a cluster document is a few KiB: refuse absurd bodies
    }
    if (has_len) {
        while (body->size() < len) {
            ssize_t r = ::read(fd, tmp, sizeof(tmp));
            if (r <= 0) return false;
            body->append(tmp, (size_t)r);
        }
        body->resize(len);
    }
    return true;
}

bool HttpServer::start(uint16_t port, HttpHandler h, std::string *err)
{
    handler_ = h;
    fd_      = tcp_listen(port, err);
    if (fd_ < 0) return false;
    port_ = port;
    th_   = std::thread([this] { loop(); });
    return true;
}

void HttpServer::stop()
{
    if (stopping_.exchange(true)) return;
    if (fd_ >= 0) {
        shutdown(fd_, SHUT_RDWR);
        ::close(fd_);
    }
    if (th_.joinable()) th_.join();
    for (int i = 0; i < 200 && inflight_ > 0; ++i) std::this_thread::sleep_for(std::chrono::milliseconds(10));
}

void HttpServer::loop()
{
    while (!stopping_) {
        int c = accept4(fd_, nullptr, nullptr, SOCK_CLOEXEC);
        if (c < 0) {
            if (errno == EINTR) continue;
            break;
        }
        // a request that stalls (half a header, a body that never comes) must not pin a thread forever
        timeval tv{10, 0};
        setsockopt(c, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
        setsockopt(c, SOL_SOCKET, SO_SNDTIMEO, &tv, sizeof(tv));
        ++inflight_;
        std::thread([this, c] {
            serve(c);
            ::close(c);
            --inflight_;
        }).detach();
    }
}

static const char *status_text(int s)
{
    switch (s) {
    case 200: return "OK";
    case 400: return "Bad Request";
    case 403: return "Forbidden";
    case 404: return "Not Found";
    case 405: return "Method Not Allowed";
    }
    return "Error";
}

void HttpServer::serve(int fd)
{
    HttpRequest req;
    std::string start;
    if (!read_http(fd, &start, &req.headers, &req.body)) return;
    std::istringstream ss(start);
    ss >> req.method >> req.path;
    HttpResponse resp;
    handler_(req, &resp);
    std::ostringstream os;
    os << "HTTP/1.1 " << resp.status << " " << status_text(resp.status) << "\r\n"
       << "Content-Type: " << resp.content_type << "\r\n"
       << "Content-Length: " << resp.body.size() << "\r\n"
       << "Connection: close\r\n\r\n"
       << resp.body;
    const std::string out = os.str();
    write_full(fd, out.data(), out.size());
}

bool parse_url(const std::string &s, Url *out)
{
    auto p = s.find("://");
    if (p == std::string::npos) return false;
    out->scheme      = s.substr(0, p);
    std::string rest = s.substr(p + 3);
    auto slash       = rest.find('/');
    std::string hp   = slash == std::string::npos ? rest : rest.substr(0, slash);
    out->path        = slash == std::string::npos ? "/" : rest.substr(slash);
    auto colon       = hp.find(':');
    out->host        = colon == std::string::npos ? hp : hp.substr(0, colon);
    out->port        = colon == std::string::npos ? 80 : (uint16_t)atoi(hp.substr(colon + 1).c_str());
    return !out->host.empty();
}

static int connect_host(const std::string &host, uint16_t port)
{
    uint32_t ip;
    if (parse_ipv4(host, &ip)) return tcp_connect(ip, port);
    addrinfo hints, *res = nullptr;
    memset(&hints, 0, sizeof(hints));
    hints.ai_family   = AF_INET;
    hints.ai_socktype = SOCK_STREAM;
    if (getaddrinfo(host.c_str(), nullptr, &hints, &res) != 0 || !res) return -1;
    ip = ntohl(((sockaddr_in *)res->ai_addr)->sin_addr.s_addr);
    freeaddrinfo(res);
    return tcp_connect(ip, port);
}

bool http_do(const std::string &method, const std::string &url, const std::string &body, int *status,
             std::string *response, std::string *err)
{
    Url u;
    if (!parse_url(url, &u) || u.scheme != "http") {
        if (err) *err = "unsupported url " + url;
        return false;
    }
    int fd = connect_host(u.host, u.port);
    if (fd < 0) {
        if (err) *err = "cannot connect to " + u.host + ":" + std::to_string(u.port);
        return false;
    }
    std::ostringstream os;
    os << method << " " << u.path << " HTTP/1.1\r\nHost: " << u.host << "\r\nUser-Agent: kungfu-b200\r\n"
       << "Content-Type: application/json\r\nContent-Length: " << body.size() << "\r\nConnection: close\r\n\r\n"
       << body;
    const std::string out = os.str();
    std::string start, rbody;
    std::map<std::string, std::string> headers;
    bool ok = write_full(fd, out.data(), out.size());
    if (ok) {
        // responses without Content-Length end at EOF
        ok = read_http(fd, &start, &headers, &rbody);
        if (ok && !headers.count("content-length")) {
            char tmp[4096];
            ssize_t r;
            while ((r = ::read(fd, tmp, sizeof(tmp))) > 0) rbody.append(tmp, (size_t)r);
        }
    }
    ::close(fd);
    if (!ok) {
        if (err) *err = "http exchange with " + url + " failed";
        return false;
    }
    std::istringstream ss(start);
    std::string ver;
    int st = 0;
    ss >> ver >> st;
    if (status) *status = st;
    if (response) *response = rbody;
    return true;
}

bool open_url(const std::string &url, std::string *content, std::string *err)
{
    std::string path;
    if (url.compare(0, 7, "file://") == 0) {
        path = url.substr(7);
    } else if (url.find("://") == std::string::npos) {
        path = url;
    }
    if (!path.empty()) {
        std::ifstream in(path);
        if (!in) {
            if (err) *err = "cannot open " + path;
            return false;
        }
        std::ostringstream os;
        os << in.rdbuf();
        *content = os.str();
        return true;
    }
    int status = 0;
    if (!http_do("GET", url, "", &status, content, err)) return false;
    if (status != 200) {
        if (err) *err = "GET " + url + " -> " + std::to_string(status);
        return false;
    }
    return true;
}

}  // namespace kf
