// sched.cu -- multi-GPU block scheduler (north_star item 5).  LD blocks are independent
// units: one worker thread per GPU pulls blocks from a queue sorted by descending size (longest
// processing time first), with no collective between GPUs.  The reference has no scheduler; users
// launch one `ci-gwas.py cusk <block-index>` process per block (README.md:62).
// Resume: a block whose <stem>.mdim already exists is skipped (the de-facto checkpoint of the
// reference's workflow; merge-block-outputs tolerates missing blocks, merge_blocks.py:364-386).
#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <mutex>
#include <thread>

#include "../csrc/common.cuh"
#include "host.hpp"

namespace cuskhost {

std::vector<int> devices_from_env()
{
    std::vector<int> out;
    const char *e = getenv("CUSK_DEVICES");
    if (e && *e) {
        const char *p = e;
        while (*p) {
            char *end = nullptr;
            const long v = strtol(p, &end, 10);
            if (end == p) break;
            out.push_back((int)v);
            p = (*end == ',') ? end + 1 : end;
        }
    }
    if (out.empty()) {
        const int n = cusk_device_count();
        for (int i = 0; i < n; i++) out.push_back(i);
    }
    return out;
}

// Longest-processing-time-first order: cost model m^2 (level 0 and the correlation Gram are
// quadratic in the block size; the deeper levels depend on the data).
std::vector<size_t> lpt_order(const std::vector<MarkerBlock> &blocks)
{
    std::vector<size_t> order(blocks.size());
    for (size_t i = 0; i < order.size(); i++) order[i] = i;
    std::stable_sort(order.begin(), order.end(),
                     [&](size_t a, size_t b) { return blocks[a].size() > blocks[b].size(); });
    return order;
}

namespace {

struct Tally {
    std::mutex mu;
    size_t done = 0, skipped_existing = 0, skipped_nosig = 0, failed = 0;
    int64_t tests = 0;
    double load_s = 0, device_s = 0, write_s = 0;  // summed over blocks (host threads overlap them)
    std::vector<double> device_s_of;               // per device: seconds its blocks spent in the device phase
    std::vector<size_t> blocks_of;
    std::string first_error;
};

// Host threads per GPU (CUSK_SLOTS_PER_GPU), each with its own context (stream, arenas): while one block is in its
// device phase the other slots of the GPU read, assemble, upload or write the next ones, and the kernels of several small
// blocks share the GPU (overlap file IO / H2D with compute) -- the level loop of a small block is a chain of
// short, latency-bound launches that leaves most of the device idle.  Default: the host's hardware threads per GPU,
// between 2 and 4 (measured on the 1 700-block genome, one GPU: 0.67 / 0.54 / 0.47 s of processing with 2 / 3 / 4 slots).
int slots_per_gpu(size_t ndevices)
{
    const char *e = getenv("CUSK_SLOTS_PER_GPU");
    const int v = e ? atoi(e) : 0;
    if (v > 0) return v > 8 ? 8 : v;
    const unsigned hw = std::thread::hardware_concurrency();
    const int per = (int)(hw / (unsigned)std::max<size_t>(ndevices, 1));
    return std::min(4, std::max(2, per));
}

int64_t tests_of(const RunReport &r)
{
    int64_t t = 0;
    for (int s = 0; s < r.stages; s++)
        for (int l = 0; l <= CUSK_ML; l++) t += r.stage[s].tests_executed[l];
    return t;
}

// load_shared() runs on the calling thread WHILE the workers create their CUDA contexts (seconds on an 8-GPU box);
// it fills `blocks` (and whatever run_one reads); the workers start pulling blocks once it has returned.
template <class LoadShared, class RunOne>
int run_all(std::vector<MarkerBlock> &blocks, const std::string &outdir, const std::vector<int> &devices, bool resume,
            LoadShared load_shared, RunOne run_one)
{
    if (devices.empty()) {
        std::cerr << "no CUDA device available" << std::endl;
        return 1;
    }
    std::vector<size_t> order;
    std::mutex ready_mu;
    std::condition_variable ready_cv;
    int ready = 0;  // 1: shared inputs loaded, -1: loading failed
    std::atomic<int> ctx_ready{0};
    std::atomic<long long> last_ctx_ready_us{0};  // when the last worker had its CUDA context (since t0)
    std::atomic<size_t> next{0};
    Tally tally;
    tally.device_s_of.assign(devices.size(), 0.0);
    tally.blocks_of.assign(devices.size(), 0);
    const int slots = slots_per_gpu(devices.size());
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> workers;
    for (size_t di = 0; di < devices.size() * (size_t)slots; di++) {
        const int dev = devices[di / (size_t)slots];
        const size_t dslot = di / (size_t)slots;
        workers.emplace_back([&, dev, dslot]() {
            cusk_ctx *ctx = nullptr;
            if (cusk_ctx_create(dev, &ctx) != CUSK_OK) {
                std::lock_guard<std::mutex> lk(tally.mu);
                tally.failed++;
                if (tally.first_error.empty()) tally.first_error = cusk_last_error();
                return;
            }
            {
                const long long us = std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - t0).count();
                long long prev = last_ctx_ready_us.load();
                while (us > prev && !last_ctx_ready_us.compare_exchange_weak(prev, us)) {
                }
                ctx_ready.fetch_add(1);
                std::unique_lock<std::mutex> lk(ready_mu);
                ready_cv.wait(lk, [&] { return ready != 0; });
                if (ready < 0) {
                    cusk_ctx_destroy(ctx);
                    return;
                }
            }
            for (;;) {
                const size_t k = next.fetch_add(1);
                if (k >= order.size()) break;
                const size_t bid = order[k];
                const std::string stem = make_path(outdir, blocks[bid].stem(), "");
                if (resume && path_exists(stem + ".mdim")) {
                    std::lock_guard<std::mutex> lk(tally.mu);
                    tally.skipped_existing++;
                    continue;
                }
                RunReport rep;
                int rc;
                std::string err;
                try {
                    rc = run_one(ctx, bid, &rep);
                    if (rc) err = cusk_last_error();
                } catch (const HostError &e) {
                    rc = 1;
                    err = e.msg;
                }
                std::lock_guard<std::mutex> lk(tally.mu);
                if (rc) {
                    tally.failed++;
                    if (tally.first_error.empty()) tally.first_error = "block " + blocks[bid].stem() + ": " + err;
                } else if (rep.skipped) {
                    tally.skipped_nosig++;
                } else {
                    tally.done++;
                    tally.tests += tests_of(rep);
                }
                tally.load_s += rep.load_s;
                tally.device_s += rep.device_s;
                tally.write_s += rep.write_s;
                tally.device_s_of[dslot] += rep.device_s;
                tally.blocks_of[dslot]++;
                printf("{\"block\": \"%s\", \"device\": %d, \"rc\": %d, \"skipped\": %s, \"n_var\": %zu, \"n_out\": %zu, "
                       "\"load_s\": %.4f, \"device_s\": %.4f, \"write_s\": %.4f, \"tests\": %lld}\n",
                       blocks[bid].stem().c_str(), dev, rc, rep.skipped ? "true" : "false", rep.num_var_in,
                       rep.num_var_out, rep.load_s, rep.device_s, rep.write_s, (long long)tests_of(rep));
            }
            cusk_ctx_destroy(ctx);
        });
    }
    std::string load_error;
    int load_code = 0;
    try {
        load_shared();
        order = lpt_order(blocks);
    } catch (const HostError &e) {
        load_error = e.msg;
        load_code = e.code ? e.code : 1;
    }
    const double load_shared_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    {
        std::lock_guard<std::mutex> lk(ready_mu);
        ready = load_error.empty() ? 1 : -1;
    }
    ready_cv.notify_all();
    for (auto &w : workers) w.join();
    if (!load_error.empty()) fail(load_error, load_code);
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    // startup (CUDA contexts of every worker, shared tables) against the time spent on the blocks themselves
    const double contexts_ready_s = 1e-6 * (double)last_ctx_ready_us.load();
    std::string per_gpu = "[";
    for (size_t i = 0; i < devices.size(); i++) {
        char buf[160];
        snprintf(buf, sizeof(buf), "%s{\"device\": %d, \"blocks\": %zu, \"device_phase_s\": %.4f, \"busy\": %.3f}", i ? ", " : "",
                 devices[i], tally.blocks_of[i], tally.device_s_of[i], secs > 0 ? tally.device_s_of[i] / secs : 0.0);
        per_gpu += buf;
    }
    per_gpu += "]";
    printf("{\"summary\": true, \"blocks\": %zu, \"done\": %zu, \"skipped_existing\": %zu, \"skipped_no_signal\": %zu, "
           "\"failed\": %zu, \"gpus\": %zu, \"slots_per_gpu\": %d, \"seconds\": %.3f, \"load_shared_s\": %.3f, "
           "\"contexts_ready_s\": %.3f, \"processing_s\": %.3f, \"ci_tests\": %lld, "
           "\"ci_tests_per_s\": %.4g, \"sum_load_s\": %.3f, \"sum_device_s\": %.3f, \"sum_write_s\": %.3f, \"per_gpu\": %s}\n",
           blocks.size(), tally.done, tally.skipped_existing, tally.skipped_nosig, tally.failed, devices.size(), slots, secs,
           load_shared_s, contexts_ready_s, secs - std::max(load_shared_s, contexts_ready_s), (long long)tally.tests, secs > 0 ? (double)tally.tests / secs : 0.0, tally.load_s, tally.device_s, tally.write_s,
           per_gpu.c_str());
    if (tally.failed) {
        std::cerr << tally.first_error << std::endl;
        return 1;
    }
    return 0;
}

}  // namespace

int run_cusk_all(const CuskArgs &a, const std::vector<int> &devices, bool resume)
{
    CuskShared shared;
    return run_all(shared.blocks, a.outdir, devices, resume, [&] { load_cusk_shared(a, &shared, true); },
                   [&](cusk_ctx *ctx, size_t bid, RunReport *rep) { return run_cusk_block(ctx, a, shared, bid, false, rep); });
}

// mxm_pattern: path of the per-block mxm file with "{}" standing for the block stem <chr>_<first>_<last>
int run_cuskss_all(const CuskssArgs &a, const std::string &mxm_pattern, const std::vector<int> &devices, bool resume)
{
    CuskssShared shared;
    return run_all(shared.blocks, a.outdir, devices, resume, [&] { load_cuskss_shared(a, &shared); },
                   [&](cusk_ctx *ctx, size_t bid, RunReport *rep) {
        CuskssArgs b = a;
        b.block_ix = (int)bid;
        b.merged = false;
        b.trait_only = false;
        std::string path = mxm_pattern;
        const size_t pos = path.find("{}");
        if (pos != std::string::npos) path.replace(pos, 2, shared.blocks[bid].stem());
        b.mxm_path = path;
        check_path(path);
        return run_cuskss(ctx, b, false, rep, &shared);
    });
}

}  // namespace cuskhost
