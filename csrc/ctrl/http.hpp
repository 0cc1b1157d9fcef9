// Tiny HTTP/1.1 server and client (Connection: close, Content-Length bodies):
// config server REST API, /metrics endpoint, runner debug endpoint, and the
// workers' GET/PUT of the cluster config.
// Parity: /root/reference/srcs/go/kungfu/elastic/configserver/configserver.go,
// /root/reference/srcs/go/monitor/server.go, /root/reference/srcs/go/utils/http.go.
#pragma once
#include <atomic>
#include <functional>
#include <map>
#include <string>
#include <thread>

namespace kf {

struct HttpRequest {
    std::string method, path, body;
    std::map<std::string, std::string> headers;
};

struct HttpResponse {
    int status = 200;
    std::string content_type = "text/plain";
    std::string body;
};

using HttpHandler = std::function<void(const HttpRequest &, HttpResponse *)>;

class HttpServer
{
    int fd_ = -1;
    uint16_t port_ = 0;
    std::atomic<bool> stopping_{false};
    std::atomic<int> inflight_{0};
    std::thread th_;
    HttpHandler handler_;
    void loop();
    void serve(int fd);

  public:
    ~HttpServer() { stop(); }
    bool start(uint16_t port, HttpHandler h, std::string *err);
    void stop();
    uint16_t port() const { return port_; }
};

struct Url {
    std::string scheme, host, path;
    uint16_t port = 80;
};
bool parse_url(const std::string &s, Url *out);

// Returns false on transport errors; *status carries the HTTP status.
bool http_do(const std::string &method, const std::string &url, const std::string &body, int *status,
             std::string *response, std::string *err);
// http(s)://, file:// and plain paths (/root/reference/srcs/go/utils/http.go OpenURL)
bool open_url(const std::string &url, std::string *content, std::string *err);

}  // namespace kf
