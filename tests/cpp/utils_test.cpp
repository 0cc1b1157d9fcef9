// Unit test of include/kungfu/utils.hpp (no runtime needed).
#include <kungfu/utils.hpp>

#include <cassert>
#include <cmath>
#include <thread>
#include <vector>

int main()
{
    using namespace std::chrono_literals;
    {
        kungfu::Channel<int> ch;
        std::vector<std::thread> producers;
        for (int p = 0; p < 4; ++p)
            producers.emplace_back([&ch, p] {
                for (int i = 0; i < 100; ++i) ch.put(p * 1000 + i);
            });
        long sum = 0;
        for (int i = 0; i < 400; ++i) sum += *ch.get();
        for (auto &t : producers) t.join();
        long want = 0;
        for (int p = 0; p < 4; ++p)
            for (int i = 0; i < 100; ++i) want += p * 1000 + i;
        assert(sum == want);
        assert(!ch.try_get().has_value());
        assert(!ch.get_for(5ms).has_value());
        ch.put(7);
        ch.close();
        assert(!ch.put(8));
        assert(*ch.get() == 7);
        assert(!ch.get().has_value());  // closed and drained: does not block
    }
    {
        kungfu::Latch l(3);
        assert(!l.wait_for(5ms));
        std::thread t([&l] {
            for (int i = 0; i < 3; ++i) l.done();
        });
        l.wait();
        t.join();
        assert(l.wait_for(0ms));
    }
    {
        kungfu::ExponentialMovingAverage<double> ema(0.9);
        assert(!ema.primed());
        assert(ema.update(10.0) == 10.0);
        assert(std::fabs(ema.update(0.0) - 9.0) < 1e-12);
        assert(std::fabs(ema.get() - 9.0) < 1e-12);
    }
    {
        kungfu::HandleTable h;
        const int a = h.create(), b = h.create();
        assert(a != b && h.pending() == 2);
        std::thread t([&h, a, b] {
            std::this_thread::sleep_for(5ms);
            h.done(a);
            h.done(b);
        });
        assert(h.wait(a));
        h.wait_all();
        t.join();
        assert(!h.wait(a));  // consumed
        assert(h.pending() == 0);
    }
    {
        KUNGFU_TRACE_SCOPE("utils_test");
        KUNGFU_CHECK(0);
    }
    std::printf("OK utils\n");
    return 0;
}
