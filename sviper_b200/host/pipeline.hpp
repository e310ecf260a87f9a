// sviper_b200/host/pipeline.hpp -- the device batch scheduler that replaces the loop at src/sviper.cpp:43-47.
//
//   reference:  #pragma omp parallel for schedule(guided)  ->  polish_variant(variants[i], info)   (one variant per thread,
//               the thread sits in polish_to_perfection for ~14 s of CPU per variant)
//   here:       Pipeline::run(n, body): T host threads, each running MANY variants as fibers (fiber.hpp). body(i) is the
//               reference's unchanged polish_variant. Where it reaches the device seam (Gateway::polish, called by the shim
//               of polish_to_perfection) the fiber is suspended and the thread starts preparing the next variant; the
//               variant joins whatever device batch a lane forms next (asynchronous C ABI: svp_submit_polish /
//               svp_wait_any); when its result is back, the thread that owns the fiber resumes it and the finish half
//               (VCF fields, log) runs. prepare (sviper.h:130-426) | device (polishing + final alignment) | finish (:441-507).
//
// A fiber is always resumed by the thread that started it: the reference binds one BAM/FAI handle set per thread at the top
// of polish_variant (sviper.h:133-135, auxiliary.h:34-84), and a thread runs one fiber at a time, so the handles are never
// used concurrently. Variants are dealt dynamically (shared counter), results land in the caller's objects in place -- the
// reference writes its VCF in input order from the same vector (io.h:89-90). No data moves between GPUs.
#pragma once
#include <omp.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <deque>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "fiber.hpp"
#include "polisher.hpp"

namespace sviper_b200
{
class Pipeline;

// what a suspended fiber belongs to
struct FiberHome
{
    Pipeline * pipe = nullptr;
    int owner = -1;          // thread (omp_get_thread_num) that started the fiber
    void * scope = nullptr;  // per-variant data of the shims (VariantScope)
};

// one device job a caller waits for: an OS thread blocks on the condition variable, a fiber is parked and re-queued
struct Completion
{
    std::mutex m;
    std::condition_variable cv;
    bool done = false;
    Fiber * fiber = nullptr;
    FiberHome * home = nullptr;
};

class Pipeline
{
public:
    // max_inflight: suspended variants per thread (each holds its reads and a fiber stack)
    explicit Pipeline(int threads, int max_inflight = 64) : T_(threads < 1 ? 1 : threads), cap_(max_inflight < 1 ? 1 : max_inflight), slots_(T_) {}

    // Runs body(0..n-1), each call inside a fiber, on T OpenMP threads (so omp_get_thread_num() inside body is the
    // running thread, as the reference expects). Returns when every body has returned. Exceptions from body propagate
    // (the first one) after all other fibers have finished.
    void run(size_t n, std::function<void(size_t)> const & body)
    {
        next_.store(0);
        std::exception_ptr first_error;
        std::mutex err_m;
#pragma omp parallel num_threads(T_)
        {
            const int me = omp_get_thread_num();
            Slot & s = slots_[me];
            size_t inflight = 0;
            auto step = [&](Fiber * f, FiberHome * h) {
                try { f->resume(); }
                catch (...)
                {
                    std::lock_guard<std::mutex> lk(err_m);
                    if (!first_error) first_error = std::current_exception();
                }
                if (f->done()) { delete f; delete h; return true; }
                return false;
            };
            for (;;)
            {
                // 1. finish halves first: they release memory and turn into output
                Fiber * f = nullptr;
                {
                    std::lock_guard<std::mutex> lk(s.m);
                    if (!s.ready.empty()) { f = s.ready.front(); s.ready.pop_front(); }
                }
                if (f)
                {
                    if (step(f, static_cast<FiberHome *>(f->user))) --inflight;
                    continue;
                }
                // 2. start the prepare half of the next variant
                if (inflight < (size_t)cap_)
                {
                    const size_t i = next_.fetch_add(1);
                    if (i < n)
                    {
                        FiberHome * h = new FiberHome{this, me, nullptr};
                        Fiber * nf = new Fiber([&body, i] { body(i); }, stack_bytes());
                        nf->user = h;
                        started_.fetch_add(1);
                        if (!step(nf, h)) { ++inflight; peak(inflight); }
                        continue;
                    }
                }
                if (inflight == 0) break;
                // 3. nothing to prepare (or the in-flight cap is reached): whatever is queued on the devices must run now,
                // then wait for one of this thread's variants
                if (on_idle_) on_idle_();
                std::unique_lock<std::mutex> lk(s.m);
                waits_.fetch_add(1);
                s.cv.wait(lk, [&] { return !s.ready.empty(); });
            }
        }
        if (first_error) std::rethrow_exception(first_error);
    }

    // called by the collector thread of a device when the job of a parked fiber is done
    void make_ready(int owner, Fiber * f)
    {
        Slot & s = slots_[(size_t)owner];
        {
            std::lock_guard<std::mutex> lk(s.m);
            s.ready.push_back(f);
        }
        s.cv.notify_one();
    }

    // called by a thread right before it blocks for lack of work (the Gateway hooks svp_flush in here)
    void set_on_idle(std::function<void()> f) { on_idle_ = std::move(f); }
    int threads() const { return T_; }
    size_t peak_inflight() const { return peak_.load(); }
    size_t thread_waits() const { return waits_.load(); }

private:
    struct Slot
    {
        std::mutex m;
        std::condition_variable cv;
        std::deque<Fiber *> ready;
    };
    static size_t stack_bytes()
    {
        char const * e = std::getenv("SVIPER_B200_FIBER_STACK_MB");
        return (size_t)(e && std::atoi(e) > 0 ? std::atoi(e) : 8) << 20;
    }
    void peak(size_t mine)
    {
        // (per-thread figure; the sum over threads is bounded by T * cap)
        size_t p = peak_.load();
        while (mine > p && !peak_.compare_exchange_weak(p, mine)) {}
    }

    int T_, cap_;
    std::function<void()> on_idle_;
    std::vector<Slot> slots_;
    std::atomic<size_t> next_{0}, started_{0}, peak_{0}, waits_{0};
};

// Process-wide door to the GPUs: one asynchronous context (svp_submit_polish / svp_wait_any) per device, one collector
// thread per device that hands finished jobs back to whoever waits for them. GPUs: SVIPER_B200_DEVICES="0,1,2,3" (default
// "0"); a job goes to the device with the fewest outstanding jobs. There is no CPU fallback: without a GPU this throws.
class Gateway
{
public:
    static Gateway & instance()
    {
        static Gateway g;
        return g;
    }

    // polish_to_perfection (+ final alignment when pv carries a reference part) for one variant: returns when the result
    // is in rb. Inside a Pipeline fiber the fiber is suspended meanwhile, otherwise the calling thread blocks.
    void polish(PackedVariant & pv, ResultBuffers & rb)
    {
        Completion c;
        Fiber * f = Fiber::current();
        FiberHome * home = f ? static_cast<FiberHome *>(f->user) : nullptr;
        if (home && home->pipe) { c.fiber = f; c.home = home; }
        size_t d = 0;
        for (size_t i = 1; i < dev_.size(); ++i)
            if (dev_[i]->outstanding.load() < dev_[d]->outstanding.load()) d = i;
        Device & dev = *dev_[d];
        dev.outstanding.fetch_add(1);
        dev.submitted.fetch_add(1);
        if (svp_submit_polish(dev.ctx, (uint64_t)(uintptr_t)&c, &pv.win, &rb.res) != SVP_OK)
        {
            dev.outstanding.fetch_sub(1);
            throw std::runtime_error(std::string("svp_submit_polish failed: ") + svp_last_error(dev.ctx));
        }
        {
            std::lock_guard<std::mutex> lk(dev.m);
            dev.cv.notify_one();    // the collector sleeps while nothing is outstanding
        }
        if (c.fiber)
        {
            Fiber::yield();         // resumed by the owning thread after Pipeline::make_ready
            return;
        }
        std::unique_lock<std::mutex> lk(c.m);
        c.cv.wait(lk, [&] { return c.done; });
    }

    // the final alignment alone (sviper.h:457-466), for callers whose polish job did not carry the reference part
    std::vector<CigarElement> final_alignment(std::string const & ref_part, std::string const & polished, double * score)
    {
        std::lock_guard<std::mutex> lk(final_m_);
        if (!final_) final_.reset(new Polisher(dev_[0]->id));
        return final_->final_alignment(ref_part, polished, score);
    }

    struct Stats { uint64_t batches = 0, variants = 0; };
    Stats stats()
    {
        Stats s;
        for (auto & d : dev_)
        {
            uint64_t b = 0, v = 0;
            svp_engine_stats(d->ctx, &b, &v);
            s.batches += b; s.variants += v;
        }
        return s;
    }
    size_t n_gpus() const { return dev_.size(); }
    // start on whatever is queued now (a lane otherwise waits for SVP_BATCH_TARGET variants or for the linger time)
    void flush() { for (auto & d : dev_) svp_flush(d->ctx); }

    ~Gateway()
    {
        for (auto & d : dev_)
        {
            {
                std::lock_guard<std::mutex> lk(d->m);
                d->stop = true;
            }
            d->cv.notify_all();
            if (d->collector.joinable()) d->collector.join();
        }
        final_.reset();
        for (auto & d : dev_) svp_destroy(d->ctx);
    }

private:
    struct Device
    {
        int id = 0;
        svp_ctx * ctx = nullptr;
        std::atomic<long> outstanding{0};
        std::atomic<uint64_t> submitted{0};
        std::mutex m;
        std::condition_variable cv;
        bool stop = false;
        std::thread collector;
    };

    Gateway()
    {
        char const * e = std::getenv("SVIPER_B200_DEVICES");
        std::string s = e ? e : "0";
        std::vector<int> ids;
        for (size_t p = 0; p < s.size();)
        {
            size_t q = s.find(',', p);
            if (q == std::string::npos) q = s.size();
            if (q > p) ids.push_back(std::atoi(s.substr(p, q - p).c_str()));
            p = q + 1;
        }
        if (ids.empty()) ids.push_back(0);
        for (int id : ids)
        {
            std::unique_ptr<Device> d(new Device);
            d->id = id;
            d->ctx = svp_create(id, nullptr);
            if (!d->ctx) throw std::runtime_error("sviper_b200: no usable CUDA device " + std::to_string(id) + " (there is no CPU fallback)");
            // work-conserving: a lane starts on whatever is queued as soon as it is free; while every lane is busy the
            // queue deepens and the next device batch is larger (SVP_BATCH_TARGET holds lanes back until that many wait)
            if (!std::getenv("SVP_BATCH_TARGET")) svp_engine_config(d->ctx, 1, 0);
            dev_.push_back(std::move(d));
        }
        for (auto & d : dev_) d->collector = std::thread([this, dp = d.get()] { collect(*dp); });
    }

    void collect(Device & d)
    {
        for (;;)
        {
            {
                std::unique_lock<std::mutex> lk(d.m);
                d.cv.wait(lk, [&] { return d.stop || d.outstanding.load() > 0; });
                if (d.stop && d.outstanding.load() <= 0) return;
            }
            uint64_t tag = 0;
            svp_result * r = nullptr;
            const int rc = svp_wait_any(d.ctx, &tag, &r);
            if (rc != SVP_OK) continue;     // 1: nothing outstanding (raced with the submitter's counter); errors are in r->status
            d.outstanding.fetch_sub(1);
            Completion * c = reinterpret_cast<Completion *>((uintptr_t)tag);
            if (c->fiber) c->home->pipe->make_ready(c->home->owner, c->fiber);
            else
            {
                std::lock_guard<std::mutex> lk(c->m);
                c->done = true;
                c->cv.notify_one();
            }
        }
    }

    std::vector<std::unique_ptr<Device>> dev_;
    std::mutex final_m_;
    std::unique_ptr<Polisher> final_;
};
}  // namespace sviper_b200
