// tbf_pipeline.cpp -- initToneGenerator for MANY organs at once (BASELINE config 4: one tuning per
// organ): builder threads turn specs into table sets on the host cores (tbf_tables.cpp) while the
// calling thread uploads them in order and creates the instances; with device_waves the 256
// wavetables of every organ are synthesised on the GPU (tbf_gpu_tonegen_tables_create_synth).
// Host C++ only, on top of the C-ABI.
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "tbf_gpu.h"
#include "tbf_tables.h"

namespace {

struct Built {
    tbf_tonegen_tables tt;
    tbf_wave_synth ws;
    int rc = TBF_OK;
    bool ready = false;
};

} // namespace

extern "C" int tbf_gpu_organs_create(tbf_ctx *ctx, const tbf_tonegen_spec *specs, uint32_t n, int whirl_tables,
                                     int n_threads, int device_waves, int *tonegen_handles,
                                     uint32_t *instance_ids, double *stats)
{
    if (!ctx || !specs || !n || !instance_ids)
        return TBF_ERR_ARG;
    if (n_threads < 1)
        n_threads = (int)std::max(1u, std::thread::hardware_concurrency());
    n_threads = (int)std::min<uint32_t>((uint32_t)n_threads, n);
    const uint32_t window = (uint32_t)n_threads * 4; // table sets built ahead of the uploader
    std::vector<Built> slot(window);
    std::mutex mu;
    std::condition_variable cv;
    std::atomic<uint32_t> next{0};
    uint32_t consumed = 0; // guarded by mu
    bool stop = false;     // guarded by mu

    auto builder = [&] {
        for (;;) {
            const uint32_t i = next.fetch_add(1);
            if (i >= n)
                return;
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return stop || i < consumed + window; });
                if (stop)
                    return;
            }
            Built &b = slot[i % window];
            const int rc = device_waves ? tbf_build_tonegen_tables_synth(&specs[i], &b.tt, &b.ws)
                                        : tbf_build_tonegen_tables(&specs[i], &b.tt);
            std::lock_guard<std::mutex> lk(mu);
            b.rc = rc;
            b.ready = true;
            cv.notify_all();
        }
    };
    std::vector<std::thread> threads;
    for (int t = 0; t < n_threads; ++t)
        threads.emplace_back(builder);

    const auto t_begin = std::chrono::steady_clock::now();
    double wait_s = 0.0, upload_s = 0.0;
    int result = TBF_OK;
    for (uint32_t i = 0; i < n && result == TBF_OK; ++i) {
        Built &b = slot[i % window];
        const auto t0 = std::chrono::steady_clock::now();
        {
            std::unique_lock<std::mutex> lk(mu);
            cv.wait(lk, [&] { return b.ready; });
        }
        const auto t1 = std::chrono::steady_clock::now();
        result = b.rc;
        if (result == TBF_OK) {
            int h = -1;
            result = device_waves ? tbf_gpu_tonegen_tables_create_synth(ctx, &b.tt, &b.ws, &h)
                                  : tbf_gpu_tonegen_tables_create(ctx, &b.tt, &h);
            if (result == TBF_OK) {
                if (tonegen_handles)
                    tonegen_handles[i] = h;
                result = tbf_gpu_instance_create(ctx, h, whirl_tables, 1, &instance_ids[i]);
            }
            tbf_free_tonegen_tables(&b.tt);
        }
        const auto t2 = std::chrono::steady_clock::now();
        wait_s += std::chrono::duration<double>(t1 - t0).count();
        upload_s += std::chrono::duration<double>(t2 - t1).count();
        std::lock_guard<std::mutex> lk(mu);
        b.ready = false;
        consumed = i + 1;
        cv.notify_all();
    }
    {
        std::lock_guard<std::mutex> lk(mu);
        stop = true;
        cv.notify_all();
    }
    for (auto &t : threads)
        t.join();
    // table sets a builder finished after the uploader gave up
    for (Built &b : slot)
        if (b.ready && b.rc == TBF_OK)
            tbf_free_tonegen_tables(&b.tt);
    if (stats) {
        stats[0] = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
        stats[1] = wait_s;
        stats[2] = upload_s;
    }
    return result;
}
