// kungfu-bench-allreduce / kungfu-bench-p2p: micro-benchmarks of the HOST data
// plane (TCP / unix-socket graph walks, p2p blob store), run under kungfu-run.
// (The device collectives have their own benchmark, benchmarks/allreduce_bench.py.)
//
//   kungfu-run -np 4 kungfu-bench-allreduce -model resnet50 -mode par -epochs 5
//   kungfu-run -np 4 kungfu-bench-p2p -model vgg16 -epochs 3
//   kungfu-run -np 4 kungfu-bench-allreduce -model resnet50 -fuse -batch 32     (fake trainer: Img/sec lines,
//       /root/reference/tests/cpp/integration/fake_trainer.hpp:169-176)
//
// Metrics follow the reference's tools so numbers are comparable:
//   all-reduce "rate" = 4 (np-1) bytes / time   (/root/reference/tests/go/cmd/kungfu-bench-allreduce/kungfu-bench-allreduce.go:74-75)
//   p2p "rate"        = 2 np bytes / time       (/root/reference/tests/go/cmd/kungfu-bench-p2p/kungfu-bench-p2p.go:89-90)
// The gradient-size lists are generated from the architectures (ResNet-50:
// 161 tensors with BN statistics excluded / VGG16: 32 / BERT-base: 199+).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/kungfu/peer.hpp"
#include "../../include/kungfu/utils.hpp"

namespace
{
typedef std::vector<int64_t> Sizes;

void conv_bn(Sizes &s, int64_t cin, int64_t cout, int k)
{
    s.push_back(cout * cin * k * k);
    s.push_back(cout);  // bn weight
    s.push_back(cout);  // bn bias
}

Sizes resnet50_sizes()
{
    Sizes s;
    conv_bn(s, 3, 64, 7);
    const int blocks[4] = {3, 4, 6, 3};
    int64_t inplanes    = 64;
    for (int st = 0; st < 4; ++st) {
        const int64_t planes = 64LL << st;
        for (int b = 0; b < blocks[st]; ++b) {
            conv_bn(s, inplanes, planes, 1);
            conv_bn(s, planes, planes, 3);
            conv_bn(s, planes, planes * 4, 1);
            if (b == 0) conv_bn(s, inplanes, planes * 4, 1);  // downsample
            inplanes = planes * 4;
        }
    }
    s.push_back(2048 * 1000);
    s.push_back(1000);
    return s;
}

Sizes vgg16_sizes()
{
    Sizes s;
    const int cfg[] = {64, 64, 0, 128, 128, 0, 256, 256, 256, 0, 512, 512, 512, 0, 512, 512, 512, 0};
    int64_t cin     = 3;
    for (int c : cfg) {
        if (c == 0) continue;
        s.push_back((int64_t)c * cin * 9);
        s.push_back(c);
        cin = c;
    }
    const int64_t fc[3][2] = {{512 * 7 * 7, 4096}, {4096, 4096}, {4096, 1000}};
    for (auto &f : fc) {
        s.push_back(f[0] * f[1]);
        s.push_back(f[1]);
    }
    return s;
}

Sizes bert_base_sizes()
{
    Sizes s;
    const int64_t H = 768, I = 3072, V = 30522;
    s.push_back(V * H), s.push_back(512 * H), s.push_back(2 * H), s.push_back(H), s.push_back(H);
    for (int l = 0; l < 12; ++l) {
        s.push_back(3 * H * H), s.push_back(3 * H);  // qkv
        s.push_back(H * H), s.push_back(H);          // proj
        s.push_back(H), s.push_back(H);              // ln1
        s.push_back(I * H), s.push_back(I);          // fc1
        s.push_back(H * I), s.push_back(H);          // fc2
        s.push_back(H), s.push_back(H);              // ln2
    }
    s.push_back(H * H), s.push_back(H);  // pooler
    s.push_back(H * H), s.push_back(H), s.push_back(H), s.push_back(H), s.push_back(V);  // mlm head
    s.push_back(2 * H), s.push_back(2);                                                  // nsp
    return s;
}

// "resnet50" | "vgg16" | "bert" | "<count>x<elems>"
bool parse_model(const std::string &m, Sizes *out)
{
    if (m == "resnet50") return *out = resnet50_sizes(), true;
    if (m == "vgg16") return *out = vgg16_sizes(), true;
    if (m == "bert") return *out = bert_base_sizes(), true;
    const size_t x = m.find('x');
    if (x == std::string::npos) return false;
    const long n = atol(m.substr(0, x).c_str()), e = atol(m.substr(x + 1).c_str());
    if (n <= 0 || e <= 0) return false;
    out->assign((size_t)n, (int64_t)e);
    return true;
}

double now_s()
{
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

struct Args {
    std::string model = "resnet50", mode = "par";
    int epochs = 5, warmup = 2, batch = 0;
    bool fuse = false;
};

bool parse_args(int argc, char **argv, Args *a)
{
    for (int i = 1; i < argc; ++i) {
        const std::string f = argv[i];
        auto val = [&](const char *name) -> const char * {
            if (i + 1 >= argc) {
                fprintf(stderr, "%s needs a value\n", name);
                exit(2);
            }
            return argv[++i];
        };
        if (f == "-model") a->model = val("-model");
        else if (f == "-mode") a->mode = val("-mode");
        else if (f == "-epochs") a->epochs = atoi(val("-epochs"));
        else if (f == "-warmup") a->warmup = atoi(val("-warmup"));
        else if (f == "-fuse") a->fuse = true;
        else if (f == "-batch") a->batch = atoi(val("-batch"));
        else return false;
    }
    return a->mode == "par" || a->mode == "seq";
}

int bench_allreduce(const Args &a, const Sizes &sizes_in)
{
    kungfu::Peer peer;
    Sizes sizes = sizes_in;
    if (a.fuse) {
        int64_t tot = 0;
        for (auto s : sizes) tot += s;
        sizes.assign(1, tot);
    }
    int64_t total = 0;
    std::vector<std::vector<float>> send(sizes.size()), recv(sizes.size());
    std::vector<std::string> names(sizes.size());
    for (size_t i = 0; i < sizes.size(); ++i) {
        send[i].assign((size_t)sizes[i], 1.0f);
        recv[i].assign((size_t)sizes[i], 0.0f);
        names[i] = "bench::grad::" + std::to_string(i);
        total += sizes[i];
    }
    const int np     = peer.Size();
    const double mib = total * 4.0 / (1 << 20);
    for (int ep = 0; ep < a.warmup + a.epochs; ++ep) {
        peer.Barrier();
        const double t0 = now_s();
        if (a.mode == "seq") {
            for (size_t i = 0; i < sizes.size(); ++i)
                KUNGFU_CHECK(peer.AllReduce(send[i].data(), recv[i].data(), sizes[i], kungfu::F32, kungfu::SUM,
                                            names[i].c_str()));
        } else {
            kungfu::Latch all((int)sizes.size());
            for (size_t i = 0; i < sizes.size(); ++i)
                peer.AllReduce(send[i].data(), recv[i].data(), sizes[i], kungfu::F32, kungfu::SUM, names[i].c_str(),
                               [&all] { all.done(); });
            all.wait();
        }
        const double dt = now_s() - t0;
        for (size_t i = 0; i < sizes.size(); i += sizes.size() / 4 + 1)
            if (recv[i][0] != (float)np || recv[i].back() != (float)np) {
                fprintf(stderr, "wrong result in tensor %zu: %f (np=%d)\n", i, recv[i][0], np);
                return 1;
            }
        if (peer.Rank() == 0 && a.batch > 0 && ep >= a.warmup)
            // fake-trainer view (no compute, communication only): one epoch = one training step of `batch` samples
            printf("Img/sec %.1f per worker, %.1f per cluster (communication bound)\n", a.batch / dt, a.batch * np / dt);
        if (peer.Rank() == 0) {
            const double rate = 4.0 * (np - 1) * total * 4.0 / dt / (1 << 30);
            printf("%s epoch %d: %zu tensors, %.1f MiB, %.2f ms, rate %.3f GiB/s (4(np-1)bytes/t), algbw %.3f GiB/s\n",
                   ep < a.warmup ? "warmup" : "timed ", ep, sizes.size(), mib, dt * 1e3, rate,
                   total * 4.0 / dt / (1 << 30));
        }
    }
    peer.Barrier();
    if (peer.Rank() == 0) printf("OK kungfu-bench-allreduce np=%d model=%s mode=%s\n", np, a.model.c_str(), a.mode.c_str());
    return 0;
}

int bench_p2p(const Args &a, const Sizes &sizes)
{
    kungfu::Peer peer;
    int64_t total = 0;
    for (auto s : sizes) total += s;
    std::vector<float> model((size_t)total, (float)(peer.Rank() + 1)), other((size_t)total, 0.0f);
    const int np = peer.Size(), target = (peer.Rank() + 1) % np;
    KUNGFU_CHECK(peer.Save("bench::model", model.data(), total * 4));
    peer.Barrier();
    for (int ep = 0; ep < a.warmup + a.epochs; ++ep) {
        peer.Barrier();
        const double t0 = now_s();
        KUNGFU_CHECK(peer.Request(target, "bench::model", other.data(), total * 4));
        peer.Barrier();
        const double dt = now_s() - t0;
        if (other[0] != (float)(target + 1) || other.back() != (float)(target + 1)) {
            fprintf(stderr, "wrong model from rank %d: %f\n", target, other[0]);
            return 1;
        }
        if (peer.Rank() == 0)
            printf("%s epoch %d: %.1f MiB per request, %.2f ms, rate %.3f GiB/s (2 np bytes/t)\n",
                   ep < a.warmup ? "warmup" : "timed ", ep, total * 4.0 / (1 << 20), dt * 1e3,
                   2.0 * np * total * 4.0 / dt / (1 << 30));
    }
    if (peer.Rank() == 0) printf("OK kungfu-bench-p2p np=%d model=%s\n", np, a.model.c_str());
    return 0;
}
}  // namespace

int main(int argc, char **argv)
{
    const bool p2p = strstr(argv[0], "p2p") != nullptr;
    Args a;
    Sizes sizes;
    if (!parse_args(argc, argv, &a) || !parse_model(a.model, &sizes)) {
        fprintf(stderr,
                "usage: kungfu-run -np N %s [-model resnet50|vgg16|bert|<count>x<elems>] [-mode par|seq] [-fuse] [-batch N] "
                "[-epochs 5] [-warmup 2]\n",
                argv[0]);
        return 2;
    }
    try {
        return p2p ? bench_p2p(a, sizes) : bench_allreduce(a, sizes);
    } catch (const std::exception &e) {
        fprintf(stderr, "%s: %s\n", argv[0], e.what());
        return 1;
    }
}
