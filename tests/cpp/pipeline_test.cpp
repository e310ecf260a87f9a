// CPU test of the prepare / device / finish pipeline (sviper_b200/host/pipeline.hpp, fiber.hpp) with a SIMULATED device: a
// collector thread that completes parked jobs after a short delay, in shuffled order, the way Gateway::collect does. Checks
// what the reference's polish_variant relies on (per-thread BAM / FAI handles, include/sviper/sviper.h:133-135): a body always
// continues on the thread that started it, and a thread never runs two bodies at once; plus: every body finishes exactly
// once, the number of parked bodies per thread respects the cap, stack data survives the suspension, an exception in one
// body does not lose the others and is rethrown at the end.
#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <stdexcept>

#include "sviper_b200/host/pipeline.hpp"

using namespace sviper_b200;

struct FakeDevice
{
    std::mutex m;
    std::condition_variable cv;
    std::vector<Completion *> q;
    bool stop = false;
    std::thread th;
    std::atomic<long> batches{0}, jobs{0};

    FakeDevice() : th([this] { run(); }) {}
    ~FakeDevice()
    {
        { std::lock_guard<std::mutex> lk(m); stop = true; }
        cv.notify_all();
        th.join();
    }
    // what Gateway::polish does around svp_submit_polish, for a fiber
    void submit_and_park()
    {
        Completion c;
        Fiber * f = Fiber::current();
        if (!f) throw std::logic_error("not in a fiber");
        c.fiber = f; c.home = static_cast<FiberHome *>(f->user);
        { std::lock_guard<std::mutex> lk(m); q.push_back(&c); }
        cv.notify_one();
        Fiber::yield();
    }
    void run()
    {
        std::mt19937 rng(7);
        for (;;)
        {
            std::vector<Completion *> batch;
            {
                std::unique_lock<std::mutex> lk(m);
                cv.wait(lk, [&] { return stop || !q.empty(); });
                if (q.empty()) return;
                batch.swap(q);
            }
            std::shuffle(batch.begin(), batch.end(), rng);
            if (rng() % 4 == 0) std::this_thread::sleep_for(std::chrono::microseconds(200));
            batches++; jobs += (long)batch.size();
            for (Completion * c : batch) c->home->pipe->make_ready(c->home->owner, c->fiber);
        }
    }
};

#define CHECK(x) do { if (!(x)) { std::fprintf(stderr, "FAILED: %s (line %d)\n", #x, __LINE__); return 1; } } while (0)

int main(int argc, char ** argv)
{
    const int threads = argc > 1 ? std::atoi(argv[1]) : 4;
    const size_t n = argc > 2 ? (size_t)std::atol(argv[2]) : 2000;
    const int cap = argc > 3 ? std::atoi(argv[3]) : 8;
    // --- 1. every body parks three times (MSA-like, polish, a second seam) and continues on its own thread
    {
        FakeDevice dev;
        Pipeline pipe(threads, cap);
        std::vector<int> done(n, 0), started_on(n, -1);
        std::vector<std::atomic<int>> running(threads);
        for (auto & r : running) r = 0;
        std::atomic<int> wrong_thread{0}, overlapped{0}, stack_broken{0};
        pipe.run(n, [&](size_t i) {
            const int me = omp_get_thread_num();
            started_on[i] = me;
            volatile unsigned long canary[64];
            for (int k = 0; k < 64; ++k) canary[k] = i * 1315423911ul + k;
            for (int seam = 0; seam < 3; ++seam)
            {
                if (running[me].fetch_add(1) != 0) overlapped++;      // this thread is inside another body right now
                for (volatile int spin = 0; spin < 200; ++spin) {}
                running[me].fetch_sub(1);
                dev.submit_and_park();
                if (omp_get_thread_num() != me) wrong_thread++;
            }
            for (int k = 0; k < 64; ++k) if (canary[k] != i * 1315423911ul + k) stack_broken++;
            done[i]++;
        });
        CHECK(wrong_thread == 0);
        CHECK(overlapped == 0);
        CHECK(stack_broken == 0);
        for (size_t i = 0; i < n; ++i) CHECK(done[i] == 1 && started_on[i] >= 0 && started_on[i] < threads);
        CHECK(pipe.peak_inflight() <= (size_t)cap);
        CHECK(dev.jobs == (long)(3 * n));
        std::printf("parked/resumed %ld jobs in %ld simulated batches (mean %.1f), peak parked per thread %zu of %d\n", dev.jobs.load(),
                    dev.batches.load(), (double)dev.jobs / (double)dev.batches, pipe.peak_inflight(), cap);
        CHECK(threads == 1 || n < 100 || pipe.peak_inflight() >= 2);       // bodies really overlap on a thread
    }
    // --- 2. bodies that never reach the device (SKIP / FAIL records) and an exception in one body
    {
        FakeDevice dev;
        Pipeline pipe(threads, cap);
        std::atomic<size_t> finished{0};
        bool thrown = false;
        try
        {
            pipe.run(300, [&](size_t i) {
                if (i % 3 == 0) { finished++; return; }                      // filtered before the device seam
                dev.submit_and_park();
                if (i == 101) throw std::runtime_error("variant 101");
                finished++;
            });
        }
        catch (std::runtime_error const & e) { thrown = std::string(e.what()) == "variant 101"; }
        CHECK(thrown);
        CHECK(finished == 299);
    }
    // --- 3. the idle hook runs before a thread blocks (the Gateway flushes the device queue there)
    {
        FakeDevice dev;
        Pipeline pipe(2, 4);
        std::atomic<int> idle_calls{0};
        pipe.set_on_idle([&] { idle_calls++; });
        pipe.run(40, [&](size_t) { dev.submit_and_park(); });
        CHECK(idle_calls > 0);
    }
    std::puts("ok");
    return 0;
}
