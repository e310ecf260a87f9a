// sviper_b200/host/fiber.hpp -- stackful coroutines (POSIX ucontext) for the prepare / device / finish pipeline.
//
// The reference polishes one variant per call of sviper::polish_variant (include/sviper/sviper.h:130-510): file I/O and
// the long-read consensus, then polish_to_perfection (:431), then the final alignment (:457) and the VCF fields. The
// pipeline runs that UNCHANGED function inside a fiber: where it reaches the device seams the fiber is suspended, the
// OS thread goes on preparing the next variant, and the fiber is resumed on the same thread once the device batch
// that holds its variant has finished. That is the re-slicing of polish_variant into prepare / device / finish without
// rewriting it.
#pragma once
#include <sys/mman.h>
#include <ucontext.h>

#include <cstddef>
#include <exception>
#include <functional>
#include <stdexcept>
#include <utility>

namespace sviper_b200
{
class Fiber
{
public:
    explicit Fiber(std::function<void()> fn, size_t stack_bytes = (size_t)8 << 20) : fn_(std::move(fn)), stack_bytes_(stack_bytes)
    {
        // lazily committed stack with a guard page below it
        void * p = mmap(nullptr, stack_bytes_ + 4096, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_STACK | MAP_NORESERVE, -1, 0);
        if (p == MAP_FAILED) throw std::runtime_error("sviper_b200: cannot map a fiber stack");
        stack_ = static_cast<char *>(p);
        mprotect(stack_, 4096, PROT_NONE);
        getcontext(&ctx_);
        ctx_.uc_stack.ss_sp = stack_ + 4096;
        ctx_.uc_stack.ss_size = stack_bytes_;
        ctx_.uc_link = &caller_;
        makecontext(&ctx_, reinterpret_cast<void (*)()>(&Fiber::trampoline), 2, (unsigned)((uintptr_t)this >> 32),
                    (unsigned)((uintptr_t)this & 0xffffffffu));
    }
    ~Fiber() { if (stack_) munmap(stack_, stack_bytes_ + 4096); }
    Fiber(Fiber const &) = delete;
    Fiber & operator=(Fiber const &) = delete;

    // Runs the fiber on the calling thread until it yields or finishes. An exception that escaped the fiber's function
    // is rethrown here.
    void resume()
    {
        Fiber * prev = current_ref();
        current_ref() = this;
        swapcontext(&caller_, &ctx_);
        current_ref() = prev;
        if (ex_) { std::exception_ptr e = ex_; ex_ = nullptr; std::rethrow_exception(e); }
    }
    bool done() const { return done_; }

    // the fiber running on this thread (nullptr outside any fiber)
    static Fiber * current() { return current_ref(); }
    // suspends the running fiber; control returns from the resume() that started it
    static void yield()
    {
        Fiber * f = current_ref();
        if (!f) throw std::logic_error("sviper_b200: Fiber::yield outside a fiber");
        swapcontext(&f->ctx_, &f->caller_);
    }

    void * user = nullptr;   // owner's per-fiber data (the pipeline's variant scope)

private:
    static Fiber *& current_ref()
    {
        static thread_local Fiber * cur = nullptr;
        return cur;
    }
    static void trampoline(unsigned hi, unsigned lo)
    {
        Fiber * f = reinterpret_cast<Fiber *>(((uintptr_t)hi << 32) | (uintptr_t)lo);
        try { f->fn_(); }
        catch (...) { f->ex_ = std::current_exception(); }
        f->done_ = true;
        // returning switches to uc_link == caller_
    }

    std::function<void()> fn_;
    size_t stack_bytes_;
    char * stack_ = nullptr;
    ucontext_t ctx_, caller_;
    bool done_ = false;
    std::exception_ptr ex_;
};
}  // namespace sviper_b200
