// tbf_multi.cpp -- one host process, several GPUs: a context and a worker thread per device,
// static partition of the organ list sorted by table key, no collective (include/tbf_multi.h,
// SURVEY 8(e)).  Host C++ only; all device work goes through the C-ABI of tbf_gpu.h.
#include "tbf_multi.h"

#include <algorithm>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <numeric>
#include <string>
#include <thread>
#include <vector>

namespace {

// one device's worker: runs the jobs posted to it, in order, on its own host thread
struct Worker {
    tbf_ctx *ctx = nullptr;
    std::thread thread;
    std::mutex mu;
    std::condition_variable cv;
    std::function<void()> job;
    bool has_job = false, busy = false, stop = false;

    void loop()
    {
        std::unique_lock<std::mutex> lk(mu);
        for (;;) {
            cv.wait(lk, [&] { return has_job || stop; });
            if (stop)
                return;
            std::function<void()> j = std::move(job);
            has_job = false;
            busy = true;
            lk.unlock();
            j();
            lk.lock();
            busy = false;
            cv.notify_all();
        }
    }
    void post(std::function<void()> j)
    {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return !has_job && !busy; });
        job = std::move(j);
        has_job = true;
        cv.notify_all();
    }
    void wait_idle()
    {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return !has_job && !busy; });
    }
};

// consecutive slots of a batch that live on one device
struct Run {
    uint32_t first, count;
};

} // namespace

struct tbf_multi {
    std::vector<Worker *> workers;
    std::string err;
    std::mutex err_mu;
};

extern "C" {

int tbf_multi_create(tbf_multi **out, int n_devices, const int *devices)
{
    if (!out || n_devices < 1)
        return TBF_ERR_ARG;
    *out = nullptr;
    tbf_multi *m = new tbf_multi();
    for (int i = 0; i < n_devices; ++i) {
        Worker *w = new Worker();
        const int rc = tbf_gpu_create(&w->ctx, devices ? devices[i] : i);
        if (rc != TBF_OK) {
            delete w;
            tbf_multi_destroy(m);
            return rc;
        }
        w->thread = std::thread([w] { w->loop(); });
        m->workers.push_back(w);
    }
    *out = m;
    return TBF_OK;
}

void tbf_multi_destroy(tbf_multi *m)
{
    if (!m)
        return;
    for (Worker *w : m->workers) {
        {
            std::unique_lock<std::mutex> lk(w->mu);
            w->cv.wait(lk, [&] { return !w->has_job && !w->busy; });
            w->stop = true;
            w->cv.notify_all();
        }
        if (w->thread.joinable())
            w->thread.join();
        tbf_gpu_destroy(w->ctx);
        delete w;
    }
    delete m;
}

int tbf_multi_device_count(const tbf_multi *m) { return m ? (int)m->workers.size() : TBF_ERR_ARG; }

tbf_ctx *tbf_multi_context(tbf_multi *m, int index)
{
    return (m && index >= 0 && (size_t)index < m->workers.size()) ? m->workers[index]->ctx : nullptr;
}

const char *tbf_multi_last_error(const tbf_multi *m) { return m ? m->err.c_str() : "null tbf_multi"; }

int tbf_multi_plan(int n_devices, uint32_t n, const uint64_t *table_key, uint32_t *device_of, uint32_t *order)
{
    if (n_devices < 1 || !n || !table_key || !device_of)
        return TBF_ERR_ARG;
    std::vector<uint32_t> ord(n);
    std::iota(ord.begin(), ord.end(), 0u);
    std::stable_sort(ord.begin(), ord.end(),
                     [&](uint32_t a, uint32_t b) { return table_key[a] < table_key[b]; });
    const uint32_t per = (n + (uint32_t)n_devices - 1) / (uint32_t)n_devices;
    for (uint32_t k = 0; k < n; ++k) {
        device_of[ord[k]] = k / per;
        if (order)
            order[k] = ord[k];
    }
    return TBF_OK;
}

int tbf_multi_render_batch(tbf_multi *m, const tbf_multi_batch *b, float *device_ms)
{
    if (!m || !b || !b->n_slots || !b->device_of || !b->instance_ids)
        return TBF_ERR_ARG;
    const size_t nd = m->workers.size();
    if (b->flags & TBF_RENDER_OUT_CHUNKED) {
        m->err = "tbf_multi_render_batch: OUT_CHUNKED is a single-device layout";
        return TBF_ERR_ARG;
    }
    const bool dev_out = (b->flags & TBF_RENDER_OUT_DEVICE) != 0;
    if (dev_out && (!b->dev_out_l || !b->dev_out_r)) {
        m->err = "tbf_multi_render_batch: OUT_DEVICE needs dev_out_l / dev_out_r";
        return TBF_ERR_ARG;
    }
    // runs of consecutive slots per device; events are sorted by slot, so a run's events are one
    // contiguous range
    std::vector<std::vector<Run>> runs(nd);
    for (uint32_t i = 0; i < b->n_slots;) {
        const uint32_t d = b->device_of[i];
        if (d >= nd) {
            m->err = "tbf_multi_render_batch: device index out of range";
            return TBF_ERR_ARG;
        }
        uint32_t j = i + 1;
        while (j < b->n_slots && b->device_of[j] == d)
            ++j;
        runs[d].push_back(Run{i, j - i});
        i = j;
    }
    if (b->n_events && !b->events) {
        m->err = "tbf_multi_render_batch: events missing";
        return TBF_ERR_ARG;
    }
    std::vector<uint64_t> ev_first((size_t)b->n_slots + 1, 0);
    for (uint64_t e = 0; e < b->n_events; ++e) {
        const uint32_t s = b->events[e].instance;
        if (s >= b->n_slots || (e && s < b->events[e - 1].instance)) {
            m->err = "tbf_multi_render_batch: events not sorted by slot, or slot out of range";
            return TBF_ERR_ARG;
        }
        ev_first[s + 1] += 1;
    }
    for (uint32_t s = 0; s < b->n_slots; ++s)
        ev_first[s + 1] += ev_first[s];
    std::vector<int> rc(nd, TBF_OK);
    std::vector<float> ms(nd, 0.f);
    for (size_t d = 0; d < nd; ++d) {
        if (runs[d].empty())
            continue;
        Worker *w = m->workers[d];
        w->post([=, &runs, &ev_first, &rc, &ms] {
            uint32_t dev_row = 0; // rows already written to this device's own output buffers
            for (const Run &r : runs[d]) {
                const uint64_t e0 = ev_first[r.first], e1 = ev_first[r.first + r.count];
                std::vector<tbf_event> ev(b->events + e0, b->events + e1);
                for (tbf_event &e : ev)
                    e.instance -= r.first; // slot inside this call
                tbf_batch_desc desc = {};
                desc.n_instances = r.count;
                desc.instance_ids = b->instance_ids + r.first;
                desc.n_fragments = b->n_fragments;
                desc.events = ev.data();
                desc.n_events = ev.size();
                desc.out_stride = b->out_stride;
                desc.flags = b->flags;
                if (dev_out) {
                    desc.out_l = b->dev_out_l[d] + (size_t)dev_row * b->out_stride;
                    desc.out_r = b->dev_out_r[d] + (size_t)dev_row * b->out_stride;
                } else if (!(b->flags & TBF_RENDER_NO_OUTPUT)) {
                    desc.out_l = b->out_l + (size_t)r.first * b->out_stride;
                    desc.out_r = b->out_r + (size_t)r.first * b->out_stride;
                    for (int j = 0; j < 4; ++j)
                        desc.stems[j] = b->stems[j] ? b->stems[j] + (size_t)r.first * b->out_stride : nullptr;
                }
                desc.checksums = b->checksums ? b->checksums + r.first : nullptr;
                if (b->in) {
                    desc.in = b->in + (size_t)r.first * b->in_stride;
                    desc.in_stride = b->in_stride;
                }
                const int one = tbf_gpu_render_batch(w->ctx, &desc);
                if (one != TBF_OK) {
                    rc[d] = one;
                    std::lock_guard<std::mutex> g(m->err_mu);
                    m->err = "device entry " + std::to_string(d) + ": " + tbf_gpu_last_error(w->ctx);
                    return;
                }
                float t = 0.f;
                if (tbf_gpu_last_launch_ms(w->ctx, &t, nullptr) == TBF_OK)
                    ms[d] += t;
                dev_row += r.count;
            }
        });
    }
    for (size_t d = 0; d < nd; ++d)
        m->workers[d]->wait_idle();
    for (size_t d = 0; d < nd; ++d) {
        if (device_ms)
            device_ms[d] = ms[d];
    }
    for (size_t d = 0; d < nd; ++d)
        if (rc[d] != TBF_OK)
            return rc[d];
    return TBF_OK;
}

} // extern "C"
