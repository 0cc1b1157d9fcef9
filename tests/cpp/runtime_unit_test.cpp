// Unit tests of the host runtime's internals, one block per unit test file of the reference's Go
// runtime (the subjects, not the code):
//   plan/hostspec_test.go, plan/cluster_test.go, plan/hostfile/hostfile_test.go, plan/graph/graph_test.go,
//   kungfu/runner/flags_test.go, kungfu/runner/peerspec_test.go, kungfu/job/cuda_visible_device_test.go,
//   store/versionedstore_test.go, rchannel/connection/message_test.go, monitor/counters_test.go.
// Links against libkungfu_host.so and includes the runtime's own headers (csrc/).
#include <sys/socket.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "core/graph.hpp"
#include "core/plan.hpp"
#include "ctrl/transport.hpp"
#include "elastic/runner.hpp"

using namespace kf;

static int g_failed = 0;
#define CHECK(cond)                                                        \
    do {                                                                   \
        if (!(cond)) {                                                     \
            fprintf(stderr, "%s:%d: CHECK(%s) failed\n", __FILE__, __LINE__, #cond); \
            ++g_failed;                                                    \
        }                                                                  \
    } while (0)

static void test_host_specs_and_peer_lists()
{
    HostList hl;
    CHECK(parse_host_list("", &hl) && hl.empty());  // an empty list is valid and empty
    CHECK(parse_host_list("192.168.1.11:4,192.168.1.12:4:pub.example", &hl) && hl.size() == 2);
    CHECK(hl.cap() == 8 && hl[0].public_addr == "192.168.1.11" && hl[1].public_addr == "pub.example");
    PortRange pr;
    CHECK(parse_port_range("10000-11000", &pr));
    PeerList pl;
    std::string err;
    CHECK(hl.gen_peer_list(0, pr, &pl, &err) && pl.empty());  // zero peers is fine
    CHECK(hl.gen_peer_list(5, pr, &pl, &err) && pl.size() == 5);
    CHECK(pl[3].port == 10003 && pl[4].port == 10000 && pl[4].ipv4 == hl[1].ipv4);  // host by host, ports from the range
    CHECK(!hl.gen_peer_list(9, pr, &pl, &err) && !err.empty());                     // over capacity
    PeerList rl = hl.gen_runner_list(38080);
    CHECK(rl.size() == 2 && rl[1].str() == "192.168.1.12:38080");

    // peer lists: "ip:port,ip:port", empty list allowed, round trip through str()
    PeerList a;
    CHECK(parse_peer_list("", &a) && a.empty());
    CHECK(parse_peer_list("127.0.0.1:10000,127.0.0.1:10001,127.0.0.1:10008", &a) && a.size() == 3);
    PeerList b;
    CHECK(parse_peer_list(a.str(), &b) && a == b);
    CHECK(!parse_peer_list("127.0.0.1", &b) && !parse_peer_list("127.0.0.1:x", &b) && !parse_peer_list("1.2.3:80", &b));
    CHECK(a.rank_of(a[2]) == 2 && a.local_rank_of(a[2]) == 2 && a.local_size_of(a[0]) == 3 && a.host_count() == 1);

    // diff / disjoint (what a runner computes on every update)
    PeerList c;
    CHECK(parse_peer_list("127.0.0.1:10001,127.0.0.2:10000", &c));
    PeerList removed, added;
    a.diff(c, &removed, &added);
    CHECK(removed.size() == 2 && added.size() == 1 && added[0].str() == "127.0.0.2:10000");
    CHECK(!a.disjoint(c));
    PeerList d;
    CHECK(parse_peer_list("127.0.0.3:1", &d) && a.disjoint(d));
    uint32_t ip2;
    CHECK(parse_ipv4("127.0.0.2", &ip2) && c.on_host(ip2).size() == 1);
}

static void test_cluster_resize()
{
    // two runners, three workers: the fourth worker goes to the less loaded host on its next free port
    Cluster c;
    c.runners = {PeerID{1, 31300}, PeerID{2, 31200}};
    c.workers = {PeerID{1, 100}, PeerID{1, 101}, PeerID{2, 100}};
    Cluster d;
    std::string err;
    CHECK(c.resize(4, &d, &err) && d.workers.size() == 4 && d.workers[3] == (PeerID{2, 101}));
    CHECK(c.resize(1, &d, &err) && d.workers.size() == 1 && d.workers[0] == c.workers[0]);  // shrink keeps the head
    CHECK(c.resize(3, &d, &err) && d == c);
    // a host that has no worker yet starts at the lowest port the job uses elsewhere; an empty cluster at the default
    Cluster e2;
    e2.runners = {PeerID{1, 38080}, PeerID{2, 38080}};
    e2.workers = {PeerID{1, 31700}, PeerID{1, 31701}};
    CHECK(e2.resize(3, &d, &err) && d.workers[2] == (PeerID{2, 31700}));
    e2.workers.clear();
    CHECK(e2.resize(1, &d, &err) && d.workers[0] == (PeerID{1, 10000}));
    // JSON round trip (what the config server stores)
    Cluster e;
    CHECK(parse_cluster_json(c.json(), &e, &err) && e == c);
    Stage s, s2;
    s.version = 7;
    s.cluster = c;
    CHECK(parse_stage_json(s.json(), &s2, &err) && s2.version == 7 && s2.cluster == c);
    // a worker on a host without a runner is rejected
    Cluster bad = c;
    bad.workers.push_back(PeerID{9, 100});
    CHECK(!bad.validate(&err));
}

static void test_hostfile()
{
    const char *text =
        "\n\t# ...\n\t192.168.0.3 slots=4 # ...\n\t# ...\n   \t127.0.0.1 slots=8 public_addr=x.y.z# ...\n\t";
    HostList hl;
    std::string err;
    CHECK(parse_hostfile(text, &hl, &err));
    CHECK(hl.size() == 2);
    uint32_t ip;
    CHECK(parse_ipv4("192.168.0.3", &ip) && hl[0].ipv4 == ip && hl[0].slots == 4 && hl[0].public_addr == "192.168.0.3");
    CHECK(parse_ipv4("127.0.0.1", &ip) && hl[1].ipv4 == ip && hl[1].slots == 8 && hl[1].public_addr == "x.y.z");
}

static void test_forest_array()
{
    Graph g;
    int roots = -1;
    CHECK(Graph::from_forest_array({1, 1, 1}, &g, &roots) && roots == 1 && g.size() == 3);
    CHECK(g.is_valid_tree_with_root(1) && !g.is_valid_tree_with_root(0));
    CHECK(Graph::from_forest_array({0, 1, 2}, &g, &roots) && roots == 3);
    CHECK(!Graph::from_forest_array({1, 2, 3}, &g, &roots));  // father out of range
}

static void test_run_flags()
{
    const char *argv[] = {"kungfu-run", "-port-range", "8080-8088", "-np", "3", "-H", "127.0.0.1:2,127.0.0.2:2",
                          "-strategy", "RING", "-w", "-timeout", "90s", "prog", "-np", "9"};
    RunFlags f;
    std::string err;
    CHECK(parse_run_flags((int)(sizeof(argv) / sizeof(argv[0])), const_cast<char **>(argv), &f, &err));
    CHECK(f.port_range.begin == 8080 && f.port_range.end == 8088);
    CHECK(f.np == 3 && f.watch && f.strategy == Ring && f.timeout == 90.0);
    CHECK(f.prog == "prog" && f.args.size() == 2 && f.args[1] == "9");  // flags after the program belong to the program
    CHECK(f.hosts.size() == 2 && f.hosts.cap() == 4);
    const char *bad[] = {"kungfu-run", "-port-range", "8088-8080", "prog"};
    RunFlags g;
    CHECK(!parse_run_flags(4, const_cast<char **>(bad), &g, &err));
}

static void test_cuda_visible_devices()
{
    std::string tok;
    CHECK(map_cuda_index(1, nullptr, &tok) == 1 && tok == "1");  // unset: the slot is the device
    CHECK(map_cuda_index(1, "2,3", &tok) == 3 && tok == "3");
    CHECK(map_cuda_index(0, "2,3", &tok) == 2 && tok == "2");
    CHECK(map_cuda_index(0, "", &tok) == -1 && tok.empty());     // set but empty: no GPU for anybody
    CHECK(map_cuda_index(2, "2,3", &tok) == -1);                 // not enough devices
    CHECK(map_cuda_index(1, " 5 , -1, 7 ", &tok) == 7);           // blanks trimmed, negative entries skipped
    CHECK(map_cuda_index(0, "1,1", &tok) == -1);                 // listed twice
    CHECK(map_cuda_index(-1, "0,1", &tok) == -1);                // pool exhausted upstream
    CHECK(map_cuda_index(1, "GPU-aaaa,GPU-bbbb", &tok) == 1 && tok == "GPU-bbbb");  // uuid form: passed on by position
    CHECK(map_cuda_index(2, "GPU-aaaa,GPU-bbbb", &tok) == -1);

    GpuPool pool(2);
    const int a = pool.get(), b = pool.get();
    CHECK(a == 0 && b == 1 && pool.get() == -1);
    pool.put(a);
    CHECK(pool.get() == 0 && pool.cap() == 2);

    // the environment contract of one worker
    RunFlags f;
    f.prog = "python";
    f.args = {"train.py"};
    Cluster c;
    c.runners = {PeerID{0x7f000001, 38080}};
    c.workers = {PeerID{0x7f000001, 10000}, PeerID{0x7f000001, 10001}};
    setenv("CUDA_VISIBLE_DEVICES", "4,6", 1);
    ProcSpec p = make_worker_proc(f, c.workers[1], c.runners[0], c, 3, 1);
    CHECK(p.env["KUNGFU_CUDA_VISIBLE_DEVICES"] == "6" && p.env["CUDA_VISIBLE_DEVICES"] == "6");
    CHECK(p.env["KUNGFU_SELF_SPEC"] == "127.0.0.1:10001" && p.env["KUNGFU_INIT_PEERS"] == c.workers.str());
    CHECK(p.env["KUNGFU_INIT_CLUSTER_VERSION"] == "3" && p.env["KUNGFU_PARENT_ID"] == "127.0.0.1:38080");
    CHECK(p.name == "127.0.0.1.10001" && p.log_tag == "127.0.0.1.10001@3");
    f.allow_nvlink = true;  // the whole node stays visible; the worker finds device 6 in the inherited list
    p = make_worker_proc(f, c.workers[1], c.runners[0], c, 3, 1);
    CHECK(p.env["KUNGFU_CUDA_VISIBLE_DEVICES"] == "6" && p.env.count("CUDA_VISIBLE_DEVICES") == 0);
    CHECK(p.env["KUNGFU_ALLOW_NVLINK"] == "true");
    unsetenv("CUDA_VISIBLE_DEVICES");
    p = make_worker_proc(f, c.workers[0], c.runners[0], c, 0, 0);
    CHECK(p.env["KUNGFU_CUDA_VISIBLE_DEVICES"] == "0");
}

static void test_versioned_store()
{
    VersionedStore vs(1);  // window of one version
    const uint8_t one = 1, two = 2;
    vs.save("0xff", "a.idx", &one, 1);
    CHECK(vs.get("0x00", "a.idx") == nullptr);  // unknown version
    CHECK(vs.get("0xff", "b.idx") == nullptr);  // unknown name
    auto b = vs.get("0xff", "a.idx");
    CHECK(b && b->size() == 1 && (*b)[0] == 1);
    vs.save("0x100", "a.idx", &two, 1);         // a new version pushes the old one out of the window
    CHECK(vs.get("0xff", "a.idx") == nullptr && vs.versions() == 1);
    auto c = vs.get("0x100", "a.idx");
    CHECK(c && (*c)[0] == 2);
    CHECK(b && (*b)[0] == 1);  // a blob handed out stays valid for its holder

    VersionedStore w3(3);
    for (int v = 0; v < 5; ++v) w3.save(std::to_string(v), "m", &one, 1);
    CHECK(w3.versions() == 3 && w3.get("1", "m") == nullptr && w3.get("2", "m") != nullptr && w3.get("4", "m") != nullptr);

    BlobStore bs;
    bs.save("x", "abc", 3);
    bs.save("x", "de", 2);  // overwrite
    auto x = bs.get("x");
    CHECK(x && x->size() == 2 && (*x)[0] == 'd' && bs.get("y") == nullptr);
}

static void test_message_framing()
{
    int sv[2];
    CHECK(socketpair(AF_UNIX, SOCK_STREAM, 0, sv) == 0);
    // short message
    CHECK(write_message(sv[0], "123456", FLAG_NONE, "abcdef", 6));
    MsgHeader h;
    CHECK(read_message_header(sv[1], &h) && h.name == "123456" && h.len == 6 && h.flags == FLAG_NONE);
    char buf[8] = {0};
    CHECK(read_full(sv[1], buf, 6) && std::string(buf, 6) == "abcdef");
    // empty body, flags carried
    CHECK(write_message(sv[0], "n", FLAG_IS_RESPONSE | FLAG_REQUEST_FAILED, nullptr, 0));
    CHECK(read_message_header(sv[1], &h) && h.name == "n" && h.len == 0 && h.flags == (FLAG_IS_RESPONSE | FLAG_REQUEST_FAILED));
    // 16 MiB body written from another thread (a socket buffer is far smaller)
    const size_t big = size_t(16) << 20;
    std::vector<char> payload(big), got(big);
    for (size_t i = 0; i < big; ++i) payload[i] = (char)('0' + i % 8);
    std::thread wr([&] { CHECK(write_message(sv[0], std::string(1024, 'k'), FLAG_NONE, payload.data(), big)); });
    CHECK(read_message_header(sv[1], &h) && h.name.size() == 1024 && h.len == big);
    CHECK(read_full(sv[1], got.data(), big) && memcmp(got.data(), payload.data(), big) == 0);
    wr.join();
    // hostile headers are refused: a 2 MiB name, a 128 GiB body
    uint32_t nl = 2u << 20;
    CHECK(write_full(sv[0], &nl, 4) && !read_message_header(sv[1], &h));
    nl = 1;
    const uint32_t fl = 0;
    const uint64_t len = uint64_t(1) << 37;
    CHECK(write_full(sv[0], &nl, 4) && write_full(sv[0], "x", 1) && write_full(sv[0], &fl, 4) && write_full(sv[0], &len, 8));
    CHECK(!read_message_header(sv[1], &h));
    // a peer that goes away mid-header
    close(sv[0]);
    CHECK(!read_message_header(sv[1], &h));
    close(sv[1]);

    // rendezvous: a body that arrives before its receiver is parked, and the other way round
    Mailbox mb;
    const PeerID src{0x7f000001, 10000};
    CHECK(socketpair(AF_UNIX, SOCK_STREAM, 0, sv) == 0);
    CHECK(write_full(sv[0], "hello", 5) && mb.deliver(src, "early", sv[1], 5));
    char r[16] = {0};
    CHECK(mb.recv_into(src, "early", r, sizeof(r), 1.0) == 5 && std::string(r, 5) == "hello");
    std::thread late([&] {
        std::this_thread::sleep_for(std::chrono::milliseconds(50));
        write_full(sv[0], "world!", 6);
        mb.deliver(src, "late", sv[1], 6);
    });
    CHECK(mb.recv_into(src, "late", r, sizeof(r), 5.0) == 6 && std::string(r, 6) == "world!");
    late.join();
    CHECK(mb.recv_into(src, "never", r, sizeof(r), 0.05) == -1);  // timeout
    close(sv[0]);
    close(sv[1]);
}

static void test_rate_accumulators()
{
    NetMonitor nm;
    const PeerID a{0x7f000001, 1}, b{0x7f000001, 2};
    nm.egress(a, 3);
    nm.ingress(a, 2);
    nm.sample();
    CHECK(nm.egress_total(a) == 3 && nm.egress_total(b) == 0 && nm.egress_rate(a) == 0.0);
    std::this_thread::sleep_for(std::chrono::milliseconds(20));
    nm.egress(a, 1000);
    nm.egress(b, 10);
    nm.sample();
    CHECK(nm.egress_total(a) == 1003 && nm.egress_rate(a) > 0 && nm.egress_rate(a) < 1000 / 0.015);
    CHECK(nm.egress_rate(b) > 0 && nm.egress_rate(a) > nm.egress_rate(b));
    const std::string text = nm.metrics_text();
    CHECK(text.find("egress") != std::string::npos && text.find("1003") != std::string::npos);
    std::this_thread::sleep_for(std::chrono::milliseconds(5));
    nm.sample();  // nothing sent in this window
    CHECK(nm.egress_rate(a) == 0.0);
}

int main()
{
    test_host_specs_and_peer_lists();
    test_cluster_resize();
    test_hostfile();
    test_forest_array();
    test_run_flags();
    test_cuda_visible_devices();
    test_versioned_store();
    test_message_framing();
    test_rate_accumulators();
    if (g_failed) {
        fprintf(stderr, "%d checks failed\n", g_failed);
        return 1;
    }
    printf("OK runtime-unit\n");
    return 0;
}
