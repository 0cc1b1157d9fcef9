// C++ integration test of the kungfu::Peer API, run under kungfu-run for
// np = 1..4 and every strategy (the reference's fake-agent,
// /root/reference/tests/cpp/integration/fake_agent.cpp:15-205).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <numeric>
#include <vector>

#include <kungfu/peer.hpp>
#include <kungfu/utils.hpp>

#define CHECK(cond)                                                            \
    do {                                                                       \
        if (!(cond)) {                                                         \
            fprintf(stderr, "CHECK failed %s:%d: %s\n", __FILE__, __LINE__, #cond); \
            exit(1);                                                           \
        }                                                                      \
    } while (0)

int main()
{
    kungfu::Peer peer;
    const int np = peer.Size(), rank = peer.Rank();
    // AllReduce: y[i] = i * np
    {
        const int n = 1 << 18;
        std::vector<int32_t> x(n), y(n);
        std::iota(x.begin(), x.end(), 0);
        CHECK(peer.AllReduce(x.data(), y.data(), n, kungfu::I32, kungfu::SUM, "ar-int") == 0);
        for (int i = 0; i < n; ++i) CHECK(y[i] == i * np);
    }
    // Reduce / AllReduce x100 with changing names
    for (int k = 0; k < 100; ++k) {
        float x = (float)(rank + k), y = 0;
        char name[32];
        snprintf(name, sizeof(name), "loop%d", k);
        CHECK(peer.AllReduce(&x, &y, 1, kungfu::F32, kungfu::SUM, name) == 0);
        CHECK(std::fabs(y - (np * (np - 1) / 2.0f + (float)k * np)) < 1e-3);
        float r = 0;
        CHECK(peer.Reduce(&x, &r, 1, kungfu::F32, kungfu::MAX, name) == 0);
        if (rank == 0) CHECK(r == (float)(np - 1 + k));
    }
    // async AllReduce with callback
    {
        std::vector<float> x(1000, 1.0f), y(1000, 0.0f);
        kungfu::Latch done;  // a real happens-before edge from the callback to this thread (TSAN-clean)
        peer.AllReduce(x.data(), y.data(), 1000, kungfu::F32, kungfu::SUM, "async", [&] { done.done(); });
        done.wait();
        CHECK(y[999] == (float)np);
    }
    // Gather / AllGather
    {
        std::vector<int32_t> x(3, rank), g(3 * np, -1), a(3 * np, -1);
        CHECK(peer.Gather(x.data(), g.data(), 3, kungfu::I32, "gather") == 0);
        CHECK(peer.AllGather(x.data(), a.data(), 3, kungfu::I32, "allgather") == 0);
        for (int r = 0; r < np; ++r) {
            CHECK(a[3 * r] == r);
            if (rank == 0) CHECK(g[3 * r + 2] == r);
        }
    }
    // AllGatherTransform: sum of every rank's value computed on rank 0, delivered to all
    {
        float in = (float)(rank + 1), out = 0;
        CHECK(peer.AllGatherTransform(&in, 1, &out, 1, "agt", [](const float *all, int n, int, float *o, int) {
                  *o = 0;
                  for (int i = 0; i < n; ++i) *o += all[i];
              }) == 0);
        CHECK(out == np * (np + 1) / 2.0f);
    }
    // Broadcast, Consensus, p2p store (success and expected failure)
    {
        int32_t v = rank == 0 ? 42 : -1;
        CHECK(peer.Broadcast(&v, &v, 1, kungfu::I32, "bcast") == 0);
        CHECK(v == 42);
        CHECK(peer.Consensus(&v, sizeof(v), "same"));
        int32_t mine = rank;
        CHECK(peer.Consensus(&mine, sizeof(mine), "diff") == (np == 1));
        float model[4] = {(float)rank, 1, 2, 3};
        CHECK(peer.Save("m", model, sizeof(model)) == 0);
        CHECK(peer.Barrier() == 0);
        float other[4] = {-1, -1, -1, -1};
        const int target = (rank + 1) % np;
        CHECK(peer.Request(target, "m", other, sizeof(other)) == 0);
        CHECK(other[0] == (float)target);
        CHECK(peer.Request(target, "m!", other, sizeof(other)) != 0);
        CHECK(peer.Barrier() == 0);
    }
    // double-buffered model pull (kf_prefetch_*) and the async request handle: the pull for the next
    // call runs on the runtime's workers while this thread keeps publishing new models
    {
        const int target = (rank + 1) % np;
        std::vector<float> m(4096, (float)rank);
        CHECK(peer.Save("pf", m.data(), m.size() * 4) == 0);
        CHECK(peer.Barrier() == 0);
        kf_prefetch *pf = kf_prefetch_new(peer.raw(), "pf", (int64_t)m.size() * 4);
        CHECK(pf != nullptr);
        std::vector<float> got(m.size(), -1.f);
        int64_t fresh = 0;
        for (int it = 0; it < 40; ++it) {
            std::fill(m.begin(), m.end(), (float)(rank + 1000 * (it + 1)));
            if (it % 2) {
                CHECK(peer.Save("pf", m.data(), m.size() * 4) == 0);  // overwrite while peers pull
            } else {  // the same from two parts, gathered by the runtime into a recycled blob
                const void *parts[2]    = {m.data(), m.data() + 1000};
                const int64_t nbytes[2] = {1000 * 4, (int64_t)(m.size() - 1000) * 4};
                CHECK(kf_save_parts(peer.raw(), nullptr, "pf", parts, nbytes, 2) == 0);
            }
            CHECK(kf_prefetch_next(pf, target, got.data(), &fresh) == 0);
            CHECK(got.front() == got.back());  // one snapshot, never a mixture of two
            CHECK(((int)got.front() - target) % 1000 == 0);
        }
        // blend straight out of the double buffer: t <- 0.5 (t + newest model), two tensors
        std::vector<float> t0(1000, 2.f), t1(m.size() - 1000, 4.f);
        float *tensors[2]      = {t0.data(), t1.data()};
        const int64_t cnts[2]  = {(int64_t)t0.size(), (int64_t)t1.size()};
        CHECK(kf_prefetch_blend_f32(pf, target, tensors, cnts, 2, &fresh) == 0);
        CHECK(t0.front() - 1.f == t1.back() - 2.f && t0.front() == t0.back());  // both halves of ONE snapshot v: 1 + v/2, 2 + v/2
        const int64_t bad_cnts[2] = {1, 1};
        CHECK(kf_prefetch_blend_f32(pf, target, tensors, bad_cnts, 2, &fresh) != 0);  // sizes must add up to the model
        kf_prefetch_free(pf);  // waits for the pull in flight
        std::vector<float> a(m.size(), -1.f);
        const int64_t h = kf_request_async(peer.raw(), target, nullptr, "pf", a.data(), (int64_t)a.size() * 4);
        CHECK(h > 0 && kf_wait_handle(peer.raw(), h) == 0 && a.front() == a.back());
        const int64_t h2 = kf_request_async(peer.raw(), target, nullptr, "pf", a.data(), 8);  // wrong size
        CHECK(h2 > 0 && kf_wait_handle(peer.raw(), h2) != 0);
        CHECK(peer.Barrier() == 0);
    }
    // latencies + set_tree
    {
        std::vector<float> lat(np);
        CHECK(peer.GetPeerLatencies(lat.data(), np) == 0);
        std::vector<int32_t> tree(np, 0);
        CHECK(peer.SetTree(tree.data()) == 0);
        int32_t one = 1, sum = 0;
        CHECK(peer.AllReduce(&one, &sum, 1, kungfu::I32, kungfu::SUM, "after-set-tree") == 0);
        CHECK(sum == np);
    }
    // MonitoredAllReduce along an explicit star rooted at the last rank (sync and callback forms), Noop
    {
        std::vector<int32_t> star(np, np - 1);
        std::vector<int64_t> x(1000, rank + 1), y(1000, 0), z(1000, 0);
        CHECK(peer.MonitoredAllReduce(x.data(), y.data(), 1000, kungfu::I64, kungfu::SUM, star.data(), "monitored") == 0);
        CHECK(y[0] == (int64_t)np * (np + 1) / 2 && y[999] == y[0]);
        kungfu::Latch done(2);
        peer.MonitoredAllReduce(x.data(), z.data(), 1000, kungfu::I64, kungfu::SUM, nullptr, "monitored-async",
                                [&] { done.done(); });
        peer.Noop([&] { done.done(); });
        done.wait();
        CHECK(z[500] == y[0]);
        peer.CalcStats();
    }
    // a callback-style operation that fails reports through the error handler (not silently, not by a
    // dangling thread): ask for a blob nobody saved
    if (np > 1) {
        kungfu::Latch failed(1);
        int seen_rc = 0;
        peer.SetAsyncErrorHandler([&](int rc, const char *) {
            seen_rc = rc;
            failed.done();
        });
        float buf4[4];
        bool called = false;
        CHECK(peer.Request((rank + 1) % np, "nobody-saved-this", buf4, sizeof(buf4), [&] { called = true; }) == 0);
        failed.wait();
        CHECK(seen_rc != 0 && !called);
        peer.SetAsyncErrorHandler(nullptr);
    }
    // operations still queued when the Peer goes away are run and joined by its destructor
    {
        static int32_t late_one = 1, late_sum = 0;
        peer.Barrier();
        peer.AllReduce(&late_one, &late_sum, 1, kungfu::I32, kungfu::SUM, "late-async", [] {});
    }
    peer.Barrier();
    printf("OK fake-agent rank %d/%d\n", rank, np);
    return 0;
}
