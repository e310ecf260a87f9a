// sviper_b200/host/polisher.hpp -- C++14 host side above the C ABI (include/sviper_b200.h).
//
// Mirrors the reference's own interface for the hot path so that a call site in SViper changes by one line:
//
//   reference                                                        here
//   ---------------------------------------------------------------  -----------------------------------------
//   compute_baseQ_stats(config, reads1, reads2)   polishing.h:197    sviper_b200::compute_baseQ_stats(cfg, r1, r2)
//   polish_to_perfection(reads1, reads2, ref, config)  :319-346      Polisher::polish_to_perfection(r1, r2, ref, cfg)
//   globalAlignment(..., ConvexGaps()) + getCigarString  sviper.h:457-466   Polisher::final_alignment(ref_part, polished)
//   assign_quality(record, var, config)  evaluate_final_mapping.h:17  sviper_b200::assign_quality(n_cigar, k)
//   assign_quality + refine_variant + FAIL5  sviper.h:468-478         sviper_b200::evaluate_final_mapping(var, cigar, polished, beginPos, k)
//   #pragma omp parallel for over polish_variant  src/sviper.cpp:43   BatchScheduler (prepare threads -> device batches)
//
// No SeqAn types cross this header: sequences are std::string over "ACGTN", qualities are Phred+33 strings
// (what seqan::Dna5QString prints). Everything is computed by the CUDA library; a failed call throws.
#pragma once
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <deque>
#include <functional>
#include <future>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "../../include/sviper_b200.h"

namespace sviper_b200
{
// the SViperConfig fields the path reads and writes (include/sviper/config.h:52-144)
struct Config
{
    // inputs
    double baseQ_mean = 0.0;                 // compute_baseQ_stats (polishing.h:197-234)
    double mean_insert_size = 280.054;       // CmdOptions defaults (config.h:16-17)
    double stdev_insert_size = 145.162;
    int ref_flank_length = 500;              // sviper.h:363
    double min_coverage = 4.0;               // config.h:79
    // side effects of polish_to_perfection (config.h:72-90; logged at sviper.h:433-443)
    unsigned rounds = 0, substituted_bases = 0, inserted_bases = 0, deleted_bases = 0;
    double mappQ_mean = 0.0;
    std::vector<unsigned> cov_profile;
};

struct QRead
{
    std::string seq;    // ACGTN
    std::string qual;   // Phred+33, same length as seq; an empty read is the dummy mate (basics.h:317,398)
};

struct CigarElement
{
    char operation;
    uint32_t count;
};

inline uint8_t ord_value(char c)   // seqan::Dna5 ordValue
{
    switch (c) { case 'A': case 'a': return 0; case 'C': case 'c': return 1; case 'G': case 'g': return 2;
                 case 'T': case 't': return 3; default: return 4; }
}

inline std::vector<uint8_t> encode(std::string const & s)
{
    std::vector<uint8_t> v(s.size());
    for (size_t i = 0; i < s.size(); ++i) v[i] = ord_value(s[i]);
    return v;
}

inline std::string decode(uint8_t const * p, size_t n)
{
    std::string s(n, 'N');
    for (size_t i = 0; i < n; ++i) s[i] = "ACGTN"[p[i] > 4 ? 4 : p[i]];
    return s;
}

// polishing.h:197-234. Like the reference it divides by n * length(reads[0]) of each set, and N carries quality 0.
inline void compute_baseQ_stats(Config & cfg, std::vector<QRead> const & reads1, std::vector<QRead> const & reads2)
{
    double sum = 0;
    for (auto const * set : {&reads1, &reads2})
        for (QRead const & r : *set)
            for (size_t i = 0; i < r.seq.size(); ++i)
            {
                int q = (int)r.qual[i] - 33;
                if (q < 0) q = 0;
                if (q > 62) q = 62;
                if (ord_value(r.seq[i]) == 4) q = 0;
                sum += q;
            }
    if (reads1.empty()) return;   // polishing.h:201-202: nothing to average, baseQ_mean untouched
    double den = 0;
    if (!reads1.empty()) den += (double)reads1.size() * (double)reads1[0].seq.size();
    if (!reads2.empty()) den += (double)reads2.size() * (double)reads2[0].seq.size();
    cfg.baseQ_mean = sum / den;
}

inline double assign_quality(int n_cigar, int flanking_region) { return svp_assign_quality(n_cigar, flanking_region); }

// the Variant members the tail of polish_variant reads and writes (include/sviper/variant.h:119-131)
struct VariantFields
{
    int sv_type = SVP_SV_DEL;   // SV_TYPE::DEL = 1, SV_TYPE::INS = 2 (variant.h:13-21)
    int sv_length = -1;
    int ref_pos = -1;
    int ref_pos_end = -1;
    double quality = 0;
    std::string filter;
    std::string info;
    // QUAL as Variant::write prints it (variant.h:142)
    std::string quality_string() const
    {
        char buf[64];
        svp_format_quality(quality, buf, (int32_t)sizeof buf);
        return buf;
    }
};

// sviper.h:468-478: assign_quality, then refine_variant; FILTER becomes FAIL5 when the polished sequence no longer
// holds the variant (var otherwise unchanged). begin_pos = final_record.beginPos (sviper.h:461), polished over "ACGTN".
// Returns what refine_variant returns.
inline bool evaluate_final_mapping(VariantFields & var, std::vector<CigarElement> const & cigar,
                                   std::string const & polished, int begin_pos, int flanking_region)
{
    var.quality = svp_assign_quality((int32_t)cigar.size(), flanking_region);
    std::string ops(cigar.size(), 'M');
    std::vector<uint32_t> cnt(cigar.size());
    for (size_t i = 0; i < cigar.size(); ++i) { ops[i] = cigar[i].operation; cnt[i] = cigar[i].count; }
    svp_variant_pos vp = {var.sv_type, var.ref_pos, var.ref_pos_end, var.sv_length};
    int32_t ins_begin = 0;
    int rc = svp_refine_variant(begin_pos, ops.data(), cnt.data(), (int32_t)cigar.size(), &vp, &ins_begin);
    if (rc < 0) throw std::logic_error("UNKNOWN CIGAR CHARACTER.");   // basics.h:169
    if (rc == 0) { var.filter = "FAIL5"; return false; }
    std::vector<uint8_t> pol = encode(polished);
    int need = svp_refine_info(var.info.c_str(), &vp, pol.data(), (int32_t)pol.size(), ins_begin, nullptr, 0);
    if (need < 0) throw std::out_of_range("inserted sequence lies outside the polished sequence");
    std::vector<char> buf((size_t)need + 1);
    svp_refine_info(var.info.c_str(), &vp, pol.data(), (int32_t)pol.size(), ins_begin, buf.data(), need + 1);
    var.info.assign(buf.data(), (size_t)need);
    var.ref_pos = vp.ref_pos; var.ref_pos_end = vp.ref_pos_end; var.sv_length = vp.sv_length;
    return true;
}

// One variant's inputs in the C ABI's layout (owns the flattened buffers the svp_window_in points into)
struct PackedVariant
{
    std::vector<uint8_t> cons, ref_part, b1, q1, b2, q2;
    std::vector<int32_t> o1, o2;
    svp_window_in win;

    PackedVariant(std::vector<QRead> const & reads1, std::vector<QRead> const & reads2, std::string const & ref,
                  Config const & cfg, std::string const * ref_part_seq = nullptr)
    {
        if (reads1.size() != reads2.size()) throw std::invalid_argument("reads1 and reads2 must have the same size");
        cons = encode(ref);
        if (ref_part_seq) ref_part = encode(*ref_part_seq);
        flatten(reads1, b1, q1, o1);
        flatten(reads2, b2, q2, o2);
        std::memset(&win, 0, sizeof win);
        win.consensus = cons.data(); win.consensus_len = (int32_t)cons.size();
        win.ref_flank_length = cfg.ref_flank_length;
        win.n_pairs = (int32_t)reads1.size();
        win.reads1_bases = b1.data(); win.reads1_quals = q1.data(); win.reads1_off = o1.data();
        win.reads2_bases = b2.data(); win.reads2_quals = q2.data(); win.reads2_off = o2.data();
        win.baseQ_mean = cfg.baseQ_mean; win.insert_mean = cfg.mean_insert_size; win.insert_sd = cfg.stdev_insert_size;
        win.min_coverage = cfg.min_coverage;
        win.ref_part = ref_part_seq ? ref_part.data() : nullptr;
        win.ref_part_len = ref_part_seq ? (int32_t)ref_part.size() : 0;
    }
    PackedVariant(PackedVariant const &) = delete;
    PackedVariant & operator=(PackedVariant const &) = delete;

private:
    static void flatten(std::vector<QRead> const & reads, std::vector<uint8_t> & b, std::vector<uint8_t> & q,
                        std::vector<int32_t> & off)
    {
        off.assign(1, 0);
        for (QRead const & r : reads)
        {
            if (r.qual.size() != r.seq.size()) throw std::invalid_argument("quality string length differs from read length");
            for (size_t i = 0; i < r.seq.size(); ++i)
            {
                b.push_back(ord_value(r.seq[i]));
                int v = (int)r.qual[i] - 33;
                q.push_back((uint8_t)(v < 0 ? 0 : v));
            }
            off.push_back((int32_t)b.size());
        }
        b.push_back(0); q.push_back(0);   // keep data() non-null for empty sets
    }
};

// result buffers of one variant (owns what the svp_result points into)
struct ResultBuffers
{
    std::vector<uint8_t> polished;
    std::vector<uint32_t> cov, cig_cnt;
    std::vector<char> cig_op;
    svp_result res;

    explicit ResultBuffers(size_t consensus_len, int cigar_cap = 4096)
        : polished((size_t)svp_window_capacity((int32_t)consensus_len)),
          cov(3 * (size_t)svp_window_capacity((int32_t)consensus_len) + 1024),
          cig_cnt(cigar_cap), cig_op(cigar_cap)
    {
        std::memset(&res, 0, sizeof res);
        res.polished = polished.data(); res.polished_cap = (int32_t)polished.size();
        res.cov_profile = cov.data(); res.cov_cap = (int32_t)cov.size();
        res.cigar_op = cig_op.data(); res.cigar_cnt = cig_cnt.data(); res.cigar_cap = cigar_cap;
    }
    ResultBuffers(ResultBuffers const &) = delete;
    ResultBuffers & operator=(ResultBuffers const &) = delete;

    std::string sequence() const { return decode(polished.data(), (size_t)res.polished_len); }
    std::vector<CigarElement> cigar() const
    {
        std::vector<CigarElement> c((size_t)res.n_cigar);
        for (int i = 0; i < res.n_cigar; ++i) c[i] = {cig_op[i], cig_cnt[i]};
        return c;
    }
    void to_config(Config & cfg) const
    {
        cfg.rounds = res.rounds; cfg.substituted_bases = res.substituted_bases; cfg.inserted_bases = res.inserted_bases;
        cfg.deleted_bases = res.deleted_bases; cfg.mappQ_mean = res.mappQ_mean;
        cfg.cov_profile.assign(cov.begin(), cov.begin() + res.cov_len);
    }
};

// One context on one GPU. Not thread-safe: use one Polisher per submitting thread, or the BatchScheduler.
class Polisher
{
public:
    explicit Polisher(int cuda_device = 0, void * cuda_stream = nullptr) : ctx_(svp_create(cuda_device, cuda_stream))
    {
        if (!ctx_) throw std::runtime_error("sviper_b200: no usable CUDA device (there is no CPU fallback)");
    }
    ~Polisher() { svp_destroy(ctx_); }
    Polisher(Polisher const &) = delete;
    Polisher & operator=(Polisher const &) = delete;
    svp_ctx * ctx() const { return ctx_; }
    // number of concurrent parts a batch is split into on the GPU (default 4); results do not depend on it
    void set_lanes(int n) { check(svp_set_lanes(ctx_, n), "svp_set_lanes"); }

    // include/sviper/polishing.h:319-346 (same argument order, same side effects on config)
    std::string polish_to_perfection(std::vector<QRead> const & reads1, std::vector<QRead> const & reads2,
                                     std::string ref, Config & config)
    {
        PackedVariant pv(reads1, reads2, ref, config);
        ResultBuffers rb(ref.size());
        check(svp_polish_batch(ctx_, 1, &pv.win, &rb.res), "svp_polish_batch");
        check(rb.res.status, "polish_to_perfection");
        rb.to_config(config);
        return rb.sequence();
    }

    // include/sviper/sviper.h:452-466: convex-gap global alignment of `polished` (rows) against `ref_part`
    // (columns) and its CIGAR; final_score (discarded by the reference) is returned through *score.
    std::vector<CigarElement> final_alignment(std::string const & ref_part, std::string const & polished,
                                              double * score = nullptr)
    {
        std::vector<uint8_t> a = encode(ref_part), b = encode(polished);
        svp_final_in in = {a.data(), (int32_t)a.size(), b.data(), (int32_t)b.size()};
        ResultBuffers rb(16, (int)(a.size() + b.size() + 16));
        check(svp_final_align_batch(ctx_, 1, &in, &rb.res), "svp_final_align_batch");
        check(rb.res.status, "final_alignment");
        if (score) *score = rb.res.final_score;
        return rb.cigar();
    }

    // a whole batch in one device pass (what the scheduler calls)
    void polish_batch(std::vector<PackedVariant *> const & in, std::vector<ResultBuffers *> const & out)
    {
        std::vector<svp_window_in> w(in.size());
        std::vector<svp_result> r(in.size());
        for (size_t i = 0; i < in.size(); ++i) { w[i] = in[i]->win; r[i] = out[i]->res; }
        check(svp_polish_batch(ctx_, (int32_t)in.size(), w.data(), r.data()), "svp_polish_batch");
        for (size_t i = 0; i < in.size(); ++i) out[i]->res = r[i];
    }

private:
    void check(int rc, char const * what)
    {
        if (rc != SVP_OK) throw std::runtime_error(std::string(what) + " failed: " + svp_last_error(ctx_));
    }
    svp_ctx * ctx_;
};

// Device batch scheduler: replaces the body of the OpenMP loop at src/sviper.cpp:43-47.
// Host "prepare" threads (polish_variant's I/O half, sviper.h:130-426) submit packed variants; one worker
// thread per GPU drains the shared queue in batches (dynamic balancing across GPUs: a long insertion window
// on one GPU does not hold up the others), runs svp_polish_batch and fulfils the futures; the host "finish"
// half (assign_quality / refine_variant / log, sviper.h:441-507) runs on whoever waits on the future.
// No data moves between GPUs, so there is no collective; results land in the caller's own objects.
class BatchScheduler
{
public:
    struct Job
    {
        std::unique_ptr<PackedVariant> in;
        std::unique_ptr<ResultBuffers> out;
        std::promise<std::unique_ptr<ResultBuffers>> done;
    };

    // linger_us: after the first submission arrives a worker waits this long for more before it forms the batch
    // (submitters that block on their future, like the OpenMP threads of the GPU build, arrive one by one)
    BatchScheduler(std::vector<int> const & cuda_devices, size_t max_batch = 256, unsigned linger_us = 0)
        : max_batch_(max_batch ? max_batch : 1), linger_us_(linger_us)
    {
        for (int dev : cuda_devices) polishers_.emplace_back(new Polisher(dev));
        for (size_t g = 0; g < polishers_.size(); ++g) workers_.emplace_back([this, g] { run(g); });
    }
    ~BatchScheduler()
    {
        {
            std::lock_guard<std::mutex> lk(m_);
            stop_ = true;
        }
        cv_.notify_all();
        for (std::thread & t : workers_) t.join();
    }
    BatchScheduler(BatchScheduler const &) = delete;
    BatchScheduler & operator=(BatchScheduler const &) = delete;

    std::future<std::unique_ptr<ResultBuffers>> submit(std::unique_ptr<PackedVariant> pv)
    {
        std::unique_ptr<Job> j(new Job);
        j->out.reset(new ResultBuffers(pv->cons.size()));
        j->in = std::move(pv);
        std::future<std::unique_ptr<ResultBuffers>> f = j->done.get_future();
        {
            std::lock_guard<std::mutex> lk(m_);
            queue_.push_back(std::move(j));
        }
        cv_.notify_all();   // a lingering worker must not swallow the wake-up of an idle one
        return f;
    }
    size_t n_gpus() const { return polishers_.size(); }

private:
    void run(size_t g)
    {
        for (;;)
        {
            std::vector<std::unique_ptr<Job>> batch;
            {
                std::unique_lock<std::mutex> lk(m_);
                cv_.wait(lk, [this] { return stop_ || !queue_.empty(); });
                if (queue_.empty()) return;   // stop requested and nothing left
                if (linger_us_ && !stop_)
                    cv_.wait_for(lk, std::chrono::microseconds(linger_us_), [this] { return stop_ || queue_.size() >= max_batch_; });
                // leave work for the other GPUs: take at most an even share of what is queued
                size_t take = (queue_.size() + polishers_.size() - 1) / polishers_.size();
                if (take > max_batch_) take = max_batch_;
                for (size_t i = 0; i < take; ++i) { batch.push_back(std::move(queue_.front())); queue_.pop_front(); }
                if (!queue_.empty()) cv_.notify_all();
            }
            std::vector<PackedVariant *> in;
            std::vector<ResultBuffers *> out;
            for (auto & j : batch) { in.push_back(j->in.get()); out.push_back(j->out.get()); }
            std::exception_ptr failed;
            try { polishers_[g]->polish_batch(in, out); }
            catch (...) { failed = std::current_exception(); }
            // every promise is settled exactly once; a throwing set_value must not take the worker down
            for (auto & j : batch)
            {
                try
                {
                    if (failed) j->done.set_exception(failed);
                    else j->done.set_value(std::move(j->out));
                }
                catch (std::future_error const &) {}
            }
        }
    }

    size_t max_batch_;
    unsigned linger_us_;
    std::vector<std::unique_ptr<Polisher>> polishers_;
    std::vector<std::thread> workers_;
    std::deque<std::unique_ptr<Job>> queue_;
    std::mutex m_;
    std::condition_variable cv_;
    bool stop_ = false;
};
}  // namespace sviper_b200
