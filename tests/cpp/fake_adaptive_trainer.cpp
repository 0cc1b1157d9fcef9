// A trainer with no model: all-reduces a fake gradient every step and follows a
// "steps:size,steps:size" schedule through kungfu::Peer::ResizeCluster, re-syncing
// the step counter after every membership change. Run under `kungfu-run -w`.
// Reference: /root/reference/tests/go/cmd/kungfu-fake-adaptive-trainer/.
#include <kungfu/peer.hpp>
#include <kungfu/utils.hpp>

#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

struct Phase {
    int begin, end, size;
};

static std::vector<Phase> parse_schedule(const std::string &s, int *total)
{
    std::vector<Phase> out;
    int t = 0;
    size_t i = 0;
    while (i < s.size()) {
        size_t j = s.find(',', i);
        if (j == std::string::npos) j = s.size();
        const std::string part = s.substr(i, j - i);
        const size_t c = part.find(':');
        const int n = atoi(part.substr(0, c).c_str()), size = atoi(part.substr(c + 1).c_str());
        out.push_back({t, t + n, size});
        t += n;
        i = j + 1;
    }
    *total = t;
    return out;
}

static int64_t sync_step(kungfu::Peer &peer, int64_t step)
{
    int64_t out = 0;
    KUNGFU_CHECK(peer.AllReduce(&step, &out, 1, kungfu::I64, kungfu::MAX, "fake-trainer::step"));
    return out;
}

int main(int argc, char **argv)
{
    const std::string schedule = argc > 1 ? argv[1] : "3:2,3:4,3:1";
    const int64_t count        = argc > 2 ? atol(argv[2]) : 100000;
    int total = 0;
    const std::vector<Phase> phases = parse_schedule(schedule, &total);
    kungfu::Peer peer;
    std::vector<float> grad((size_t)count, 1.0f), sum((size_t)count);
    int64_t step = sync_step(peer, 0);
    while (step < total) {
        const int np = peer.Size();
        KUNGFU_CHECK(peer.AllReduce(grad.data(), sum.data(), count, kungfu::F32, kungfu::SUM, "fake-trainer::grad"));
        if (sum[0] != (float)np || sum.back() != (float)np) {
            fprintf(stderr, "step %ld: wrong sum %f with np=%d\n", (long)step, sum[0], np);
            return 1;
        }
        ++step;
        int want = np;
        for (auto &p : phases)
            if (p.begin <= step && step < p.end) want = p.size;
        if (want != np) {
            bool changed = false, detached = false;
            kungfu::TraceScope t("resize");
            KUNGFU_CHECK(peer.ResizeCluster(want, &changed, &detached));
            if (detached) {
                printf("detached at step %ld\n", (long)step);
                return 0;
            }
            if (changed) {
                step = sync_step(peer, step);
                if (peer.Rank() == 0)
                    printf("step %ld: resized %d -> %d in %.1f ms\n", (long)step, np, peer.Size(), t.elapsed_ms());
            }
        }
    }
    printf("OK fake-adaptive-trainer rank %d finished %ld steps with np=%d\n", peer.Rank(), (long)step, peer.Size());
    return 0;
}
