// sviper_b200 -- SViper's command line program with the per-variant OpenMP loop (src/sviper.cpp:43-47) replaced by the
// prepare / device / finish pipeline (pipeline.hpp). Everything else is the reference's own code, called in the order
// its main() calls it (src/sviper.cpp:5-59): parseCommandLine (sviper.h:26-128), print_log_header / read_vcf / write_vcf
// (io.h), prep_file_handles (auxiliary.h:34-84) and, per variant, the UNCHANGED sviper::polish_variant (sviper.h:130-510),
// compiled with the shim directory in front of the include path so that its two hot calls go to the GPU.
//
// What changes is who waits: the reference keeps one variant per thread and the thread computes polish_to_perfection
// itself; here a thread runs the I/O and long-read consensus of a variant, parks it at the device seam and prepares the
// next one, so the device sees whatever all threads have prepared (not one variant per thread) and no thread idles while
// the GPU works. Same options, same VCF (input order), same per-variant log text (blocks in completion order, as in the
// reference). Extra knobs are environment variables, so the reference's command line stays as it is:
//   SVIPER_B200_DEVICES=0,1,..   GPUs to use (default 0)
//   SVIPER_B200_INFLIGHT=n       parked variants per thread (default 64)
//   SVIPER_B200_STATS=1          one line with pipeline / device-batch statistics on stderr at the end
#include <chrono>
#include <cstdlib>
#include <iostream>

#include <sviper/io.h>
#include <sviper/sviper.h>

using namespace sviper;

int main(int argc, char const ** argv)
{
    input_output_information info{};
    seqan::ArgumentParser::ParseResult res = parseCommandLine(info.cmd_options, argc, argv);
    if (res != seqan::ArgumentParser::PARSE_OK)
        return res == seqan::ArgumentParser::PARSE_ERROR;

    if (!open_file_success(info.long_read_bai, (info.cmd_options.long_read_file_name + ".bai").c_str()) ||
        !open_file_success(info.short_read_bai, (info.cmd_options.short_read_file_name + ".bai").c_str()) ||
        !open_file_success(info.log_file, (info.cmd_options.output_prefix + ".log").c_str()))
        return 1;
    print_log_header(info.cmd_options, info.log_file);

    std::vector<std::string> vcf_header;
    std::vector<Variant> variants;
    if (!read_vcf(variants, vcf_header, info))
        return 1;

    // one BAM / BAM / FAI handle set per host thread, as in the reference (also sets the OpenMP thread count)
    if (!prep_file_handles(info))
        return 1;

    info.log_file << "======================================================================" << std::endl
                  << "START polishing variants in of file " << info.cmd_options.candidate_file_name << std::endl
                  << "======================================================================" << std::endl;

    auto const t0 = std::chrono::steady_clock::now();
    {
        char const * e = std::getenv("SVIPER_B200_INFLIGHT");
        sviper_b200::Gateway::instance();   // fail before any work if there is no usable GPU
        sviper_b200::Pipeline pipe((int)info.cmd_options.threads, e && std::atoi(e) > 0 ? std::atoi(e) : 64);
        pipe.set_on_idle([] { sviper_b200::Gateway::instance().flush(); });
        pipe.run(variants.size(), [&](size_t vidx) { polish_variant(variants[vidx], info); });
        if (std::getenv("SVIPER_B200_STATS"))
        {
            sviper_b200::Gateway::Stats st = sviper_b200::Gateway::instance().stats();
            double const s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            std::cerr << "[sviper_b200] variants=" << variants.size() << " seconds=" << s << " threads=" << pipe.threads()
                      << " gpus=" << sviper_b200::Gateway::instance().n_gpus() << " device_batches=" << st.batches
                      << " device_variants=" << st.variants
                      << " mean_device_batch=" << (st.batches ? (double)st.variants / (double)st.batches : 0.0)
                      << " peak_parked_per_thread=" << pipe.peak_inflight() << " thread_waits=" << pipe.thread_waits() << std::endl;
        }
    }

    if (!write_vcf(variants, vcf_header, info))
        return 1;

    info.log_file << "======================================================================" << std::endl
                  << "                                 DONE" << std::endl
                  << "======================================================================" << std::endl;
    return 0;
}
