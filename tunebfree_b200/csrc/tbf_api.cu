/* tbf_api.cu -- host side of the C-ABI declared in include/tbf_gpu.h.
 *
 * One context per GPU.  Tables are uploaded once and shared by any number of
 * instances; per-instance state is a flat blob array in HBM (tbf_device.h).
 * Nothing here computes audio on the host: if the CUDA device or a kernel
 * launch is unavailable every entry point fails with TBF_ERR_CUDA.
 */
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <chrono>
#include <map>
#include <string>
#include <thread>
#include <vector>

#include "tbf_device.h"
#include "tbf_tables.h"

/* Device memory of the tone-generator table sets: slabs from cudaMalloc, blocks bumped off them
 * and recycled by size class.  A sweep uploads thousands of sets and a re-tuning host frees and
 * creates one per change (src/clap.cpp:129-160): neither should pay a driver allocation (or a
 * pool growth) per set -- that was the largest part of round 1's 1.4 ms per set. */
struct DevArena {
    struct Slab {
        unsigned char *p;
        size_t cap, used;
    };
    std::vector<Slab> slabs;
    std::multimap<size_t, unsigned char *> free_blocks; /* size class -> block */
    size_t next_slab = (size_t)32 << 20;

    static size_t size_class(size_t bytes)
    {
        const size_t g = bytes >= ((size_t)4 << 20) ? ((size_t)1 << 20) : ((size_t)64 << 10);
        return (bytes + g - 1) / g * g;
    }
    unsigned char *alloc(size_t bytes)
    {
        const size_t sz = size_class(bytes);
        auto it = free_blocks.find(sz);
        if (it != free_blocks.end()) {
            unsigned char *p = it->second;
            free_blocks.erase(it);
            return p;
        }
        if (slabs.empty() || slabs.back().cap - slabs.back().used < sz) {
            const size_t cap = std::max(next_slab, sz);
            unsigned char *p = nullptr;
            if (cudaMalloc((void **)&p, cap) != cudaSuccess) {
                cudaGetLastError();
                return nullptr;
            }
            slabs.push_back(Slab{p, cap, 0});
            next_slab = std::min(next_slab * 2, (size_t)1 << 30);
        }
        Slab &sl = slabs.back();
        unsigned char *p = sl.p + sl.used;
        sl.used += sz;
        return p;
    }
    void release(unsigned char *p, size_t bytes) { free_blocks.insert({size_class(bytes), p}); }
    void destroy()
    {
        for (auto &sl : slabs)
            cudaFree(sl.p);
        slabs.clear();
        free_blocks.clear();
    }
};

struct TgSet {
    size_t blob_bytes = 0;   /* of the arena block `wave` points to */
    tbf_tonegen_tables host; /* scalars only; pointers nulled */
    float *wave = nullptr;
    uint16_t *cw = nullptr, *cb = nullptr;
    float *cl = nullptr;
    uint16_t *rw = nullptr, *rn = nullptr; /* per-key wheel runs: wheel, entries */
    /* lifetime: instances using the set; released by the caller; device memory gone */
    int refs = 0;
    bool destroyed = false, freed = false;
};
struct WhSet {
    tbf_whirl_tables host;
    float *tables = nullptr; /* 4 x 16384 displacement + 2 x 16384 x 5 FIR */
    int ring_size = TBF_WHIRL_BUF, gather_ok = 0, window = 0;
    int refs = 0;
    bool destroyed = false, freed = false;
};

/* Streams between the stage kernels for one chunk of fragments, [slot][chunk samples]. */
struct StageScratch {
    FragCtl *ctl = nullptr;
    float *f32 = nullptr;   /* u, php, dx0, hx0, line[4]: 8 planes */
    double *f64 = nullptr;  /* hang, dang: 2 planes */
    size_t cap_ctl = 0, cap_f32 = 0, cap_f64 = 0;
    /* tensor maps over f32 / f64 for the 2-D TMA tiles of stages R and D, and what they describe */
    TileMaps maps;
    bool maps_ok = false;
    const void *map_f32 = nullptr, *map_f64 = nullptr;
    size_t map_pitch = 0, map_rows32 = 0, map_rows64 = 0;
};

/* cuTensorMapEncodeTiled through the runtime (no link against libcuda) */
typedef CUresult (*tbf_encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                        const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                        CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                        CUtensorMapFloatOOBfill);
static tbf_encode_tiled_fn tbf_encode_tiled()
{
    static tbf_encode_tiled_fn fn = [] {
        void *f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            f = nullptr;
        cudaGetLastError();
        return (tbf_encode_tiled_fn)f;
    }();
    return fn;
}
/* rows x row_floats floats at `base` (row pitch = row_floats * 4 bytes) as boxes of 32 rows x 32
 * floats (128 bytes, the 128-byte swizzle's span) */
static bool tile_map_make(CUtensorMap *m, void *base, size_t row_floats, size_t rows)
{
    tbf_encode_tiled_fn enc = tbf_encode_tiled();
    if (!enc || ((uintptr_t)base & 127u) || (row_floats % 32) || !rows || rows >> 31 || row_floats >> 31)
        return false;
    const cuuint64_t dims[2] = {(cuuint64_t)row_floats, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)row_floats * 4};
    const cuuint32_t box[2] = {32, 32}, estr[2] = {1, 1};
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

#define TBF_MAX_GROUPS 4
#define TBF_MAX_STAGE_BUFS 4
#define TBF_SCAT_BLOCK 32 /* samples the kernels scatter before they read (tbf_kernels.cu BLK) */

struct tbf_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    std::string err;
    std::vector<TgSet> tg;
    std::vector<WhSet> wh;
    DevTonegenTables *d_tg = nullptr;
    DevWhirlTables *d_wh = nullptr;
    size_t d_tg_cap = 0, d_wh_cap = 0;
    InstState *d_states = nullptr;
    size_t n_inst = 0, cap_inst = 0;
    std::vector<int> inst_tg, inst_wh;
    /* lifetime: slots of destroyed instances and handles of freed table sets are recycled */
    std::vector<uint8_t> inst_alive;
    size_t n_live = 0;
    std::vector<int> free_tg, free_wh;
    DevArena arena;
    double upload_us[4] = {0, 0, 0, 0};
    uint64_t upload_n = 0;
    InstState *d_tmpl = nullptr;
    /* table upload: page-locked staging and the device copy of the raw waves + their index,
     * grown on demand (tbf_gpu_tonegen_tables_create) */
    /* two staging slots in rotation: the copies and kernels of one table set run while the host
     * fills the other slot; a slot is reused once its event has fired (no per-call sync) */
    struct StageSlot {
        unsigned char *h = nullptr, *d = nullptr;
        size_t cap_h = 0, cap_d = 0;
        cudaEvent_t done = nullptr;
        bool used = false;
        /* device-side wave synthesis: the samples the kernel flagged as rounding-ambiguous are
         * checked against the host's libm AFTER the call that made them has returned (the next
         * use of the slot, or the next entry point that reads tables): no synchronisation per
         * table set.  [0] = count, [4..] = uint4 records (wheel, sample, float bits, -) */
        uint32_t *d_flags = nullptr, *h_flags = nullptr;
        bool pending = false;
        int pending_handle = -1;
        tbf_wave_synth pending_synth;
    } stage[2];
    int stage_next = 0;
    /* device-side wave synthesis (tbf_gpu_tonegen_tables_create_synth): the dither decisions
     * of each rand() seed, host and device copies, and the list of flagged samples */
    struct DitherStream {
        uint32_t seed = 0;
        uint64_t n = 0; /* draws covered */
        std::vector<uint32_t> host;
        uint32_t *dev = nullptr;
    };
    std::vector<DitherStream> dither;
    uint64_t synth_flagged = 0, synth_differed = 0;
    /* per-render scratch, grown on demand */
    tbf_event *d_events = nullptr;
    size_t cap_events = 0;
    uint64_t *d_ev_start = nullptr;
    uint32_t *d_ids = nullptr;
    unsigned long long *d_csum = nullptr;
    size_t cap_slots = 0;
    float *d_out = nullptr;
    size_t cap_out = 0;
    /* pending host copies of the launch in flight */
    tbf_batch_desc prepared{};
    bool has_prepared = false, prepared_has_ids = false;
    uint32_t prepared_chunk = 0;
    /* fragments per device -> host staging chunk.  0 = not set by the caller: 32 (one stage launch)
     * for the chunk-major layout, whose chunks are single contiguous copies -- the first one leaves
     * after one stage launch instead of four -- and 125 (1/3 s) for the row layout, whose strided
     * copies want long rows (config 3 through host buffers, same run: chunk-major 136.8 k organ-s/s
     * with 32, 132.3 k with 125; rows 117.5 k with 32, 128.3 k with 125) */
    uint32_t host_chunk_fragments = 0;
    uint32_t host_chunk(bool chunked) const { return host_chunk_fragments ? host_chunk_fragments : chunked ? 32u : 125u; }
    int prepared_ring = 512; /* stage W's shared-memory ring: the largest the batch's whirl sets need */
    uint32_t prepared_tg_sets = 0; /* distinct tone-generator table sets among the batch's organs */
    cudaStream_t copy_stream = nullptr, copy_stream2 = nullptr; /* left / right channel */
    cudaEvent_t render_done[2] = {nullptr, nullptr}, copy_done[2] = {nullptr, nullptr},
                copy_done2[2] = {nullptr, nullptr};
    cudaEvent_t t0 = nullptr, t1 = nullptr;
    uint64_t *d_ev_cursor = nullptr; /* [3][cap_slots]: stage T, stage R, the rotor kernel */
    StageScratch scratch;
    /* instance groups: each renders its share of a batch on its own stream, so the
     * latency-bound lane = instance stages of one group overlap the others' stages */
    /* ... and every stage of a group has its own stream: stage s of chunk c overlaps stage
     * s-1 of chunk c+1 (the stage streams between them are double-buffered by chunk parity) */
    cudaStream_t sstream[TBF_MAX_GROUPS][4] = {};
    cudaStream_t rot_stream[TBF_MAX_GROUPS] = {}; /* the rotor kernel, ahead of stage R's audio chain */
    cudaEvent_t rot_done[TBF_MAX_GROUPS][TBF_MAX_STAGE_BUFS] = {};
    cudaEvent_t sdone[TBF_MAX_GROUPS][4][TBF_MAX_STAGE_BUFS] = {}; /* [group][stage][chunk % n_bufs] */
    int n_bufs = 3; /* copies of the stage streams in flight, 2..TBF_MAX_STAGE_BUFS (TBF_STAGE_BUFS) */
    cudaEvent_t fork = nullptr;
    size_t scratch_budget = (size_t)8 << 30; /* bytes of HBM the stage streams may take */
    uint32_t max_stage_chunk = 32;           /* fragments per stage-kernel launch       */
    int n_groups_max = 1; /* measured best from 512 to 8192 organs once stage T and W of
                             consecutive chunks overlap on the stage streams; TBF_GROUPS raises it */
    bool in_flight = false;
    int launches = 0;
    /* stage profiling: every stage kernel serialised on the main stream between two events */
    bool profile_stages = false;
    bool timeline = false;                        /* TBF_TIMELINE=1: see tbf_gpu_synchronize() */
    std::vector<cudaEvent_t> prof_events;        /* pool, pairs */
    std::vector<int> prof_stage;                  /* stage of pair i of the last launch */
    float stage_ms[4] = {0, 0, 0, 0};
    int stage_launches[4] = {0, 0, 0, 0};
    std::vector<uint64_t> ev_start_host;
};

#define CK(call)                                                                            \
    do {                                                                                    \
        cudaError_t e_ = (call);                                                            \
        if (e_ != cudaSuccess) {                                                            \
            ctx->err = std::string(#call) + ": " + cudaGetErrorString(e_);                  \
            return TBF_ERR_CUDA;                                                            \
        }                                                                                   \
    } while (0)

/* 1 - l with l = exp(-1.0 / (SR / n * tau)), the per-call smoothing factor of a rotor ramp
 * (src/whirl.cpp:1255-1257, :1306-1308); glibc's exp, as the reference's host computes it */
static double ramp_one_minus_l(double sample_rate, float tau)
{
    const double l = exp(-1.0 / (sample_rate / (double)(size_t)TBF_FRAGMENT * tau));
    return 1 - l;
}

static int fail(tbf_ctx *ctx, int code, const char *msg)
{
    if (ctx)
        ctx->err = msg;
    return code;
}

extern "C" {

int tbf_gpu_create(tbf_ctx **out, int device)
{
    if (!out)
        return TBF_ERR_ARG;
    *out = nullptr;
    /* 2 groups x 4 stage streams + copy + control: keep them on separate hardware queues
     * (only effective if set before the process creates its CUDA context) */
    setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n)
        return TBF_ERR_CUDA;
    tbf_ctx *ctx = new tbf_ctx();
    ctx->device = device;
    if (cudaSetDevice(device) != cudaSuccess ||
        cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&ctx->copy_stream2, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->copy_done2[0], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->copy_done2[1], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreate(&ctx->ev0) != cudaSuccess || cudaEventCreate(&ctx->ev1) != cudaSuccess ||
        cudaEventCreate(&ctx->t0) != cudaSuccess || cudaEventCreate(&ctx->t1) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->render_done[0], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->render_done[1], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->copy_done[0], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->copy_done[1], cudaEventDisableTiming) != cudaSuccess ||
        cudaMalloc(&ctx->d_tmpl, sizeof(InstState)) != cudaSuccess) {
        delete ctx;
        return TBF_ERR_CUDA;
    }
    bool ok = cudaEventCreateWithFlags(&ctx->fork, cudaEventDisableTiming) == cudaSuccess;
    {
        for (auto &sl : ctx->stage)
            ok = ok && cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming) == cudaSuccess;
    }
    /* the lane = instance stages (R, D) are latency-bound and small: their CTAs go first
     * whenever an SM has room, so they run beside the throughput stages instead of after */
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
    /* bit s: stage s on a high-priority stream.  Measured on the v5 kernels (config 3, 4096 | 16 384
     * organs, k organ-s/s): 0: 205.7 | 217.1, 2 (stage R): 203.4 | 217.8, 8: 201.3 | 215.3, 10: 199.6 |
     * 215.8, 12: 197.5 | 212.5, 14: 197.6 | 208.4, 6: 196.9 | 209.6 */
    const int prio_mask = getenv("TBF_PRIO_MASK") ? atoi(getenv("TBF_PRIO_MASK")) : 2;
    for (int g = 0; g < TBF_MAX_GROUPS && ok; ++g)
        for (int st = 0; st < 4 && ok; ++st) {
            ok = cudaStreamCreateWithPriority(&ctx->sstream[g][st], cudaStreamNonBlocking,
                                              ((prio_mask >> st) & 1) ? prio_hi : prio_lo) == cudaSuccess;
            for (int b = 0; ok && b < TBF_MAX_STAGE_BUFS; ++b)
                ok = cudaEventCreateWithFlags(&ctx->sdone[g][st][b], cudaEventDisableTiming) == cudaSuccess;
        }
    for (int g = 0; g < TBF_MAX_GROUPS && ok; ++g) {
        ok = cudaStreamCreateWithPriority(&ctx->rot_stream[g], cudaStreamNonBlocking, prio_hi) == cudaSuccess;
        for (int b = 0; ok && b < TBF_MAX_STAGE_BUFS; ++b)
            ok = cudaEventCreateWithFlags(&ctx->rot_done[g][b], cudaEventDisableTiming) == cudaSuccess;
    }
    if (!ok) {
        delete ctx;
        return TBF_ERR_CUDA;
    }
    /* tuning knobs (bench sweeps): instance groups and fragments per stage launch */
    if (const char *e = getenv("TBF_GROUPS"))
        ctx->n_groups_max = std::max(1, std::min(TBF_MAX_GROUPS, atoi(e)));
    if (const char *e = getenv("TBF_STAGE_BUFS"))
        ctx->n_bufs = std::max(2, std::min(TBF_MAX_STAGE_BUFS, atoi(e)));
    if (const char *e = getenv("TBF_STAGE_CHUNK"))
        ctx->max_stage_chunk = (uint32_t)std::max(1, std::min(256, atoi(e)));
    *out = ctx;
    return TBF_OK;
}

void tbf_gpu_destroy(tbf_ctx *ctx)
{
    if (!ctx)
        return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    ctx->arena.destroy(); /* every table set's block */
    if (ctx->upload_n && getenv("TBF_UPLOAD_PROFILE"))
        fprintf(stderr, "tbf upload profile: %llu sets; per set: validate+runs %.0f us, slot wait %.0f us, "
                        "staging fill %.0f us, descriptor+CUDA calls %.0f us\n",
                (unsigned long long)ctx->upload_n, ctx->upload_us[0] / ctx->upload_n, ctx->upload_us[1] / ctx->upload_n,
                ctx->upload_us[2] / ctx->upload_n, ctx->upload_us[3] / ctx->upload_n);
    for (auto &w : ctx->wh)
        if (!w.freed)
            cudaFree(w.tables);
    cudaFree(ctx->d_tg);
    cudaFree(ctx->d_wh);
    cudaFree(ctx->d_states);
    cudaFree(ctx->d_tmpl);
    for (auto &sl : ctx->stage) {
        cudaFreeHost(sl.h);
        cudaFree(sl.d);
        cudaFree(sl.d_flags);
        cudaFreeHost(sl.h_flags);
        if (sl.done)
            cudaEventDestroy(sl.done);
    }
    for (auto &ds : ctx->dither)
        cudaFree(ds.dev);
    cudaFree(ctx->d_events);
    cudaFree(ctx->d_ev_start);
    cudaFree(ctx->d_ev_cursor);
    cudaFree(ctx->d_ids);
    cudaFree(ctx->d_csum);
    cudaFree(ctx->d_out);
    cudaFree(ctx->scratch.ctl);
    cudaFree(ctx->scratch.f32);
    cudaFree(ctx->scratch.f64);
    cudaEventDestroy(ctx->ev0);
    cudaEventDestroy(ctx->ev1);
    for (int i = 0; i < 2; ++i) {
        cudaEventDestroy(ctx->render_done[i]);
        cudaEventDestroy(ctx->copy_done[i]);
        cudaEventDestroy(ctx->copy_done2[i]);
    }
    cudaEventDestroy(ctx->t0);
    cudaEventDestroy(ctx->t1);
    for (int g = 0; g < TBF_MAX_GROUPS; ++g)
        for (int st = 0; st < 4; ++st) {
            if (ctx->sstream[g][st])
                cudaStreamDestroy(ctx->sstream[g][st]);
            if (st == 0 && ctx->rot_stream[g])
                cudaStreamDestroy(ctx->rot_stream[g]);
            for (int par = 0; st == 0 && par < TBF_MAX_STAGE_BUFS; ++par)
                if (ctx->rot_done[g][par])
                    cudaEventDestroy(ctx->rot_done[g][par]);
            for (int par = 0; par < TBF_MAX_STAGE_BUFS; ++par)
                if (ctx->sdone[g][st][par])
                    cudaEventDestroy(ctx->sdone[g][st][par]);
        }
    if (ctx->fork)
        cudaEventDestroy(ctx->fork);
    for (cudaEvent_t e : ctx->prof_events)
        cudaEventDestroy(e);
    cudaStreamDestroy(ctx->copy_stream);
    cudaStreamDestroy(ctx->copy_stream2);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char *tbf_gpu_last_error(const tbf_ctx *ctx) { return ctx ? ctx->err.c_str() : "null context"; }

} /* extern "C" */

extern "C" {
static bool tg_usable(const tbf_ctx *ctx, int h);
}

/* the dither decisions of the first `n` rand() draws after srand(seed), on both sides */
static int dither_stream(tbf_ctx *ctx, uint32_t seed, uint64_t n, const uint32_t **host, const uint32_t **dev)
{
    tbf_ctx::DitherStream *ds = nullptr;
    for (auto &d : ctx->dither)
        if (d.seed == seed)
            ds = &d;
    if (!ds) {
        ctx->dither.emplace_back();
        ds = &ctx->dither.back();
        ds->seed = seed;
    }
    if (ds->n < n) {
        const uint64_t want = std::max<uint64_t>(n, (uint64_t)1 << 20); /* 1 Mi draws cover 48 kHz sets */
        std::vector<uint32_t> bits((want + 31) / 32);
        tbf_dither_bits(seed, want, bits.data());
        uint32_t *dev_new = nullptr;
        CK(cudaMalloc(&dev_new, sizeof(uint32_t) * bits.size()));
        CK(cudaMemcpy(dev_new, bits.data(), sizeof(uint32_t) * bits.size(), cudaMemcpyHostToDevice));
        if (ds->dev) {
            CK(cudaStreamSynchronize(ctx->stream)); /* a synthesis in flight may still read the old copy */
            cudaFree(ds->dev);
        }
        ds->dev = dev_new;
        ds->host.swap(bits);
        ds->n = want;
    }
    *host = ds->host.data();
    *dev = ds->dev;
    return TBF_OK;
}

#define TBF_SYNTH_FLAG_CAP 4096u
#define TBF_SYNTH_FLAG_EAGER 255u /* records that travel with the count (4 KB) */

/* The flagged samples of the table set a staging slot synthesised last: recompute each with the
 * host's libm (tbf_wave_sample) and, where the device's float differs, patch the set's wave
 * store -- every image of the sample in the four rotated, wrap-extended copies. */
static int finish_pending_synth(tbf_ctx *ctx, tbf_ctx::StageSlot &sl)
{
    if (!sl.pending)
        return TBF_OK;
    sl.pending = false;
    CK(cudaSetDevice(ctx->device));
    CK(cudaEventSynchronize(sl.done));
    const uint32_t n_flag = sl.h_flags[0];
    const int h = sl.pending_handle;
    if (!n_flag || h < 0 || (size_t)h >= ctx->tg.size() || ctx->tg[h].freed)
        return TBF_OK;
    if (n_flag > TBF_SYNTH_FLAG_CAP) {
        ctx->err = "wave synthesis: too many ambiguous samples in table set " + std::to_string(h) + " (" +
                   std::to_string(n_flag) + ", degenerate spec)";
        return TBF_ERR_ARG;
    }
    if (n_flag > TBF_SYNTH_FLAG_EAGER)
        CK(cudaMemcpy(sl.h_flags + 4, sl.d_flags + 4, sizeof(uint32_t) * 4 * n_flag, cudaMemcpyDeviceToHost));
    const tbf_tonegen_tables &t = ctx->tg[h].host;
    const uint32_t *bits_h = nullptr, *bits_d = nullptr;
    if (int rc = dither_stream(ctx, sl.pending_synth.seed, t.wave_total, &bits_h, &bits_d))
        return rc;
    size_t off4[TBF_NOF_WHEELS + 1] = {0}, acc = 0;
    for (int w = 1; w <= TBF_NOF_WHEELS; ++w) {
        off4[w] = acc;
        if (t.wave_length[w])
            acc += (size_t)4 * tbf_wave_copy_len(t.wave_length[w]);
    }
    for (uint32_t k = 0; k < n_flag; ++k) {
        const uint32_t *f = sl.h_flags + 4 + 4 * k;
        const uint32_t w = f[0], i = f[1];
        if (w < 1 || w > TBF_NOF_WHEELS || i >= t.wave_length[w])
            continue;
        const uint32_t pos = t.wave_offset[w] + i;
        const float v = tbf_wave_sample(&sl.pending_synth, (int)w, i, (bits_h[pos >> 5] >> (pos & 31u)) & 1u);
        uint32_t vb;
        memcpy(&vb, &v, 4);
        ctx->synth_flagged += 1;
        if (vb == f[2])
            continue;
        ctx->synth_differed += 1;
        const uint32_t len = t.wave_length[w], cl = tbf_wave_copy_len(len);
        for (uint32_t c = 0; c < 4; ++c) /* copy c holds wave[(j + c) % len] at j < cl */
            for (uint32_t j = (i + len - (c % len)) % len; j < cl; j += len)
                CK(cudaMemcpy(ctx->tg[h].wave + off4[w] + (size_t)c * cl + j, &v, sizeof(float),
                              cudaMemcpyHostToDevice));
    }
    return TBF_OK;
}

static int finish_all_pending_synth(tbf_ctx *ctx)
{
    for (auto &sl : ctx->stage)
        if (int rc = finish_pending_synth(ctx, sl))
            return rc;
    return TBF_OK;
}

/* upload one tone-generator table set; with `synth` the wave samples are made on the device */
static int create_tonegen_tables(tbf_ctx *ctx, const tbf_tonegen_tables *t, const tbf_wave_synth *synth,
                                 int *handle)
{
    if (!ctx || !t || !handle || (!synth && !t->wave_data))
        return fail(ctx, TBF_ERR_ARG, "tonegen tables: null argument");
    /* TBF_UPLOAD_PROFILE=1: where the host time of a table-set upload goes (printed at destroy) */
    static const bool prof = getenv("TBF_UPLOAD_PROFILE") != nullptr;
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto lap = [&](int k, std::chrono::steady_clock::time_point &t0) {
        if (prof) {
            const auto t1 = now();
            ctx->upload_us[k] += std::chrono::duration<double, std::micro>(t1 - t0).count();
            t0 = t1;
        }
    };
    auto tp = now();
    const uint32_t nc = t->contrib_start[TBF_MAX_KEYS];
    for (int w = 1; w <= TBF_NOF_WHEELS; ++w) {
        if ((uint64_t)t->wave_offset[w] + t->wave_length[w] > t->wave_total)
            return fail(ctx, TBF_ERR_ARG, "tonegen tables: wave slice out of range");
        if (t->wave_length[w] && t->wave_length[w] < TBF_FRAGMENT)
            return fail(ctx, TBF_ERR_ARG, "tonegen tables: wave shorter than one fragment");
        if (t->wave_length[w] > 0xFFFFu)
            return fail(ctx, TBF_ERR_ARG, "tonegen tables: wave longer than 65535 samples");
    }
    for (int k = 0; k < TBF_MAX_KEYS; ++k)
        if (t->contrib_start[k] > t->contrib_start[k + 1])
            return fail(ctx, TBF_ERR_ARG, "tonegen tables: contrib_start not monotone");
    for (uint32_t e = 0; e < nc; ++e)
        if (t->contrib_wheel[e] < 1 || t->contrib_wheel[e] > TBF_NOF_WHEELS ||
            t->contrib_bus[e] >= TBF_NOF_BUSES || t->wave_length[t->contrib_wheel[e]] == 0)
            return fail(ctx, TBF_ERR_ARG, "tonegen tables: contribution out of range");
    /* the kernels keep a TBF_VRING-slot window of the scanner's delay line in shared memory:
     * every write must land inside it (offsets are (1 + A + A sin) * 65536 with A <= 12
     * for everything scanner.modulation.* accepts, src/vibrato.cpp:249-255, :343-352) */
    for (int k = 0; k < 3; ++k)
        for (int i = 0; i < TBF_VIB_TBL; ++i)
            if ((t->vib_offset_table[k][i] >> 16) + TBF_SCAT_BLOCK + 2 >= TBF_VRING)
                return fail(ctx, TBF_ERR_GEOMETRY, "tonegen tables: scanner offset beyond the shared-memory window");
    /* wheel runs of every key's contribution list; a warp applies a key event with
     * lane = contribution when the list is wheel-sorted with unique (wheel, bus) pairs
     * (compilePlayMatrix builds it that way, src/tonegen.cpp:1183-1201) */
    std::vector<uint16_t> run_wheel, run_count;
    std::vector<uint32_t> run_start(TBF_MAX_KEYS + 1, 0);
    bool coop = true;
    {
        std::vector<uint8_t> wheel_seen(TBF_NOF_WHEELS + 1);
        for (int k = 0; k < TBF_MAX_KEYS; ++k) {
            run_start[k] = (uint32_t)run_wheel.size();
            std::fill(wheel_seen.begin(), wheel_seen.end(), 0);
            for (uint32_t e = t->contrib_start[k]; e < t->contrib_start[k + 1];) {
                const uint16_t w = t->contrib_wheel[e];
                if (wheel_seen[w])
                    coop = false; /* the wheel comes back after another one */
                wheel_seen[w] = 1;
                uint32_t e2 = e;
                uint32_t bus_seen = 0;
                while (e2 < t->contrib_start[k + 1] && t->contrib_wheel[e2] == w) {
                    if (bus_seen & (1u << t->contrib_bus[e2]))
                        coop = false;
                    bus_seen |= 1u << t->contrib_bus[e2];
                    ++e2;
                }
                run_wheel.push_back(w);
                run_count.push_back((uint16_t)(e2 - e));
                e = e2;
            }
        }
    }
    run_start[TBF_MAX_KEYS] = (uint32_t)run_wheel.size();
    const size_t nr = run_wheel.size();
    CK(cudaSetDevice(ctx->device));
    lap(0, tp); /* validation + wheel runs */

    TgSet s;
    s.host = *t;
    /* The device keeps four copies of every wave, copy c rotated left by c samples and extended
     * by two fragments of wrap-around: the reads of two consecutive fragments from any position
     * are then one contiguous, 16-byte aligned kilobyte in copy (pos & 3) that never wraps, which
     * one TMA bulk copy fetches (tbf_kernels.cu, wave_request).  The copies are made on the
     * device from the raw waves (tbf_wave_store_kernel): a quarter of the bytes over the link,
     * none at all when the raw waves are synthesised there. */
    uint32_t wave4_offset[TBF_NOF_WHEELS + 1] = {0};
    size_t n_wave4 = 0;
    for (int w = 1; w <= TBF_NOF_WHEELS; ++w) {
        const uint32_t len = t->wave_length[w];
        if (!len)
            continue;
        const uint32_t cl = tbf_wave_copy_len(len);
        if (n_wave4 + (size_t)4 * cl > 0xFFFFFFFFull)
            return fail(ctx, TBF_ERR_ARG, "tonegen tables: wave store too large");
        wave4_offset[w] = (uint32_t)n_wave4;
        n_wave4 += (size_t)4 * cl;
    }
    /* one allocation per table set (config 4 uploads thousands of them): the wave store first,
     * then the contribution and run arrays, each 16-byte aligned */
    const size_t n_wave = std::max<size_t>(4, n_wave4);
    auto up16 = [](size_t x) { return (x + 15) & ~(size_t)15; };
    const size_t o_cl = up16(sizeof(float) * n_wave);
    const size_t o_cw = up16(o_cl + sizeof(float) * nc);
    const size_t o_cb = up16(o_cw + sizeof(uint16_t) * nc);
    const size_t o_rw = up16(o_cb + sizeof(uint16_t) * nc);
    const size_t o_rn = up16(o_rw + sizeof(uint16_t) * nr);
    const size_t blob_bytes = up16(o_rn + sizeof(uint16_t) * nr) + 16;
    const size_t tail_bytes = blob_bytes - o_cl;
    /* staging, host and device sides alike:
     *   [raw waves][wave_offset, wave_length, wave4_offset][f1, attenuation (synthesis)]
     * host side only, behind it: [DevTonegenTables][the arrays behind the wave store] */
    const size_t n_meta = 3 * (TBF_NOF_WHEELS + 1);
    const size_t s_meta = up16(sizeof(float) * (size_t)t->wave_total);
    const size_t s_synth = up16(s_meta + sizeof(uint32_t) * n_meta);
    const size_t s_desc = up16(s_synth + sizeof(double) * 2 * (TBF_NOF_WHEELS + 1));
    const size_t s_tail = up16(s_desc + sizeof(DevTonegenTables));
    const size_t stage_bytes = s_tail + tail_bytes;

    tbf_ctx::StageSlot &sl = ctx->stage[ctx->stage_next];
    ctx->stage_next ^= 1;
    if (sl.pending) {
        if (int rc = finish_pending_synth(ctx, sl)) /* waits for the slot's last table set, too */
            return rc;
    } else if (sl.used) {
        CK(cudaEventSynchronize(sl.done)); /* the slot's last table set has left it */
    }
    lap(1, tp); /* waiting for the staging slot (device-bound when this dominates) */
    if (stage_bytes > sl.cap_h) {
        cudaFreeHost(sl.h);
        sl.h = nullptr;
        sl.cap_h = 0;
        CK(cudaHostAlloc((void **)&sl.h, stage_bytes + (stage_bytes >> 2), cudaHostAllocDefault));
        sl.cap_h = stage_bytes + (stage_bytes >> 2);
    }
    if (s_desc > sl.cap_d) {
        cudaFree(sl.d);
        sl.d = nullptr;
        sl.cap_d = 0;
        CK(cudaMalloc((void **)&sl.d, s_desc + (s_desc >> 2)));
        sl.cap_d = s_desc + (s_desc >> 2);
    }
    unsigned char *hs = sl.h;
    if (!synth)
        memcpy(hs, t->wave_data, sizeof(float) * (size_t)t->wave_total);
    uint32_t *meta = reinterpret_cast<uint32_t *>(hs + s_meta);
    memcpy(meta, t->wave_offset, sizeof(t->wave_offset));
    memcpy(meta + (TBF_NOF_WHEELS + 1), t->wave_length, sizeof(t->wave_length));
    memcpy(meta + 2 * (TBF_NOF_WHEELS + 1), wave4_offset, sizeof(wave4_offset));
    if (synth) {
        memcpy(hs + s_synth, synth->f1, sizeof(double) * (TBF_NOF_WHEELS + 1));
        memcpy(hs + s_synth + sizeof(double) * (TBF_NOF_WHEELS + 1), synth->attenuation,
               sizeof(double) * (TBF_NOF_WHEELS + 1));
    }
    unsigned char *tail = hs + s_tail - o_cl; /* tail + o_x = the staged copy of array x */
    memset(hs + s_tail, 0, tail_bytes);
    if (nc) {
        memcpy(tail + o_cl, t->contrib_level, sizeof(float) * nc);
        memcpy(tail + o_cw, t->contrib_wheel, sizeof(uint16_t) * nc);
        memcpy(tail + o_cb, t->contrib_bus, sizeof(uint16_t) * nc);
    }
    if (nr) {
        memcpy(tail + o_rw, run_wheel.data(), sizeof(uint16_t) * nr);
        memcpy(tail + o_rn, run_count.data(), sizeof(uint16_t) * nr);
    }

    lap(2, tp); /* filling the staging slot */
    /* the set's slot in the device array of descriptors */
    const bool reuse_tg = !ctx->free_tg.empty();
    const size_t idx = reuse_tg ? (size_t)ctx->free_tg.back() : ctx->tg.size();
    if (idx + 1 > ctx->d_tg_cap) {
        const size_t ncap = std::max<size_t>(64, ctx->d_tg_cap * 2);
        DevTonegenTables *nd = nullptr;
        cudaError_t e = cudaMalloc(&nd, sizeof(DevTonegenTables) * ncap);
        if (e != cudaSuccess) {
            ctx->err = cudaGetErrorString(e);
            return TBF_ERR_NOMEM;
        }
        if (ctx->tg.size()) {
            CK(cudaStreamSynchronize(ctx->stream)); /* descriptors still on their way into the old array */
            cudaMemcpy(nd, ctx->d_tg, sizeof(DevTonegenTables) * ctx->tg.size(), cudaMemcpyDeviceToDevice);
        }
        cudaFree(ctx->d_tg);
        ctx->d_tg = nd;
        ctx->d_tg_cap = ncap;
    }

    unsigned char *dev = ctx->arena.alloc(blob_bytes);
    if (!dev)
        return fail(ctx, TBF_ERR_NOMEM, "tonegen tables: out of device memory");
    float *d_raw = reinterpret_cast<float *>(sl.d);
    const uint32_t *d_meta = reinterpret_cast<const uint32_t *>(sl.d + s_meta);
    cudaError_t ue = cudaSuccess;
    if (!synth) {
        ue = cudaMemcpyAsync(sl.d, hs, s_synth, cudaMemcpyHostToDevice, ctx->stream);
    } else {
        /* ---- the waves are made here (tbf_wave_synth_kernel) ---- */
        const uint32_t *bits_h = nullptr, *bits_d = nullptr;
        int rc = dither_stream(ctx, synth->seed, t->wave_total, &bits_h, &bits_d);
        if (rc != TBF_OK) {
            ctx->arena.release(dev, blob_bytes);
            return rc;
        }
        const size_t flag_bytes = sizeof(uint32_t) * 4 * (1 + TBF_SYNTH_FLAG_CAP);
        if (!sl.d_flags) {
            CK(cudaMalloc((void **)&sl.d_flags, flag_bytes));
            CK(cudaHostAlloc((void **)&sl.h_flags, flag_bytes, cudaHostAllocDefault));
        }
        WaveSynthParams sp;
        sp.rate = synth->sample_rate;
        memcpy(sp.harmonics, synth->harmonics, sizeof(sp.harmonics));
        ue = cudaMemcpyAsync(sl.d + s_meta, hs + s_meta, s_desc - s_meta, cudaMemcpyHostToDevice, ctx->stream);
        if (ue == cudaSuccess)
            ue = cudaMemsetAsync(sl.d_flags, 0, 16, ctx->stream);
        if (ue == cudaSuccess)
            ue = (cudaError_t)tbf_launch_wave_synth(
                d_raw, d_meta, reinterpret_cast<const double *>(sl.d + s_synth),
                reinterpret_cast<const double *>(sl.d + s_synth) + (TBF_NOF_WHEELS + 1), sp, bits_d,
                sl.d_flags, sl.d_flags + 4, TBF_SYNTH_FLAG_CAP, ctx->stream);
        /* the count and the first records come back with the rest of the set's work; they are
         * looked at in finish_pending_synth(), not here: no wait per table set */
        if (ue == cudaSuccess)
            ue = cudaMemcpyAsync(sl.h_flags, sl.d_flags, 16 * (1 + TBF_SYNTH_FLAG_EAGER), cudaMemcpyDeviceToHost,
                                 ctx->stream);
        sl.pending_synth = *synth;
    }
    if (ue == cudaSuccess)
        ue = cudaMemcpyAsync(dev + o_cl, hs + s_tail, tail_bytes, cudaMemcpyHostToDevice, ctx->stream);
    if (ue == cudaSuccess && n_wave4 < n_wave)
        ue = cudaMemsetAsync(dev, 0, sizeof(float) * n_wave, ctx->stream);
    if (ue == cudaSuccess && n_wave4)
        ue = (cudaError_t)tbf_launch_wave_store(reinterpret_cast<float *>(dev), d_raw, d_meta, ctx->stream);

    DevTonegenTables *d = reinterpret_cast<DevTonegenTables *>(hs + s_desc);
    memset(d, 0, sizeof(*d));
    d->wave4 = reinterpret_cast<float *>(dev);
    d->contrib_wheel = reinterpret_cast<const uint16_t *>(dev + o_cw);
    d->contrib_bus = reinterpret_cast<const uint16_t *>(dev + o_cb);
    d->contrib_level = reinterpret_cast<const float *>(dev + o_cl);
    d->run_wheel = reinterpret_cast<const uint16_t *>(dev + o_rw);
    d->run_count = reinterpret_cast<const uint16_t *>(dev + o_rn);
    memcpy(d->run_start, run_start.data(), sizeof(d->run_start));
    d->coop_events = coop ? 1 : 0;
    memcpy(d->wave4_offset, wave4_offset, sizeof(d->wave4_offset));
    memcpy(d->wave_length, t->wave_length, sizeof(d->wave_length));
    memcpy(d->contrib_start, t->contrib_start, sizeof(d->contrib_start));
    for (int r = 0; r < 8; ++r) {
        memcpy(d->env[r], t->attack_env[r], sizeof(float) * TBF_FRAGMENT);
        memcpy(d->env[8 + r], t->release_env[r], sizeof(float) * TBF_FRAGMENT);
    }
    memcpy(d->key_comp_table, t->key_comp_table, sizeof(d->key_comp_table));
    memcpy(d->draw_bar_level, t->draw_bar_level, sizeof(d->draw_bar_level));
    d->perc_env_gain_reset_norm = t->perc_env_gain_reset_norm;
    d->perc_env_gain_reset_soft = t->perc_env_gain_reset_soft;
    d->perc_env_scaling = t->perc_env_scaling;
    d->perc_decay_fast_norm = t->perc_decay_fast_norm;
    d->perc_decay_fast_soft = t->perc_decay_fast_soft;
    d->perc_decay_slow_norm = t->perc_decay_slow_norm;
    d->perc_decay_slow_soft = t->perc_decay_slow_soft;
    d->perc_drawbar_normal_gain = t->perc_drawbar_normal_gain;
    d->perc_drawbar_soft_gain = t->perc_drawbar_soft_gain;
    d->perc_send_bus_a = t->perc_send_bus_a;
    d->perc_send_bus_b = t->perc_send_bus_b;
    d->perc_trigger_bus = t->perc_trigger_bus;
    d->output_level_trim = t->output_level_trim;
    memcpy(d->vib_offset_table, t->vib_offset_table, sizeof(d->vib_offset_table));
    d->vib_stator_increment = t->vib_stator_increment;
    if (ue == cudaSuccess)
        ue = cudaMemcpyAsync(ctx->d_tg + idx, d, sizeof(DevTonegenTables), cudaMemcpyHostToDevice, ctx->stream);
    if (ue == cudaSuccess)
        ue = cudaEventRecord(sl.done, ctx->stream);
    sl.used = true;
    lap(3, tp); /* descriptor + CUDA calls */
    ctx->upload_n += 1;
    if (ue != cudaSuccess) {
        cudaStreamSynchronize(ctx->stream);
        ctx->arena.release(dev, blob_bytes);
        ctx->err = std::string("tonegen tables upload: ") + cudaGetErrorString(ue);
        cudaGetLastError();
        return TBF_ERR_CUDA;
    }
    s.wave = reinterpret_cast<float *>(dev); /* owns the block; the rest point into it */
    s.blob_bytes = blob_bytes;
    s.cl = nullptr;
    s.cw = s.cb = s.rw = s.rn = nullptr;
    s.host.wave_data = nullptr;
    s.host.contrib_wheel = s.host.contrib_bus = nullptr;
    s.host.contrib_level = nullptr;
    if (reuse_tg) {
        ctx->free_tg.pop_back();
        ctx->tg[idx] = s;
    } else {
        ctx->tg.push_back(s);
    }
    if (synth) {
        sl.pending = true;
        sl.pending_handle = (int)idx;
    }
    *handle = (int)idx;
    return TBF_OK;
}

extern "C" {

int tbf_gpu_tonegen_tables_create(tbf_ctx *ctx, const tbf_tonegen_tables *t, int *handle)
{
    return create_tonegen_tables(ctx, t, nullptr, handle);
}

int tbf_gpu_tonegen_tables_create_synth(tbf_ctx *ctx, const tbf_tonegen_tables *t, const tbf_wave_synth *synth,
                                        int *handle)
{
    if (!synth)
        return fail(ctx, TBF_ERR_ARG, "tonegen tables: null synthesis description");
    return create_tonegen_tables(ctx, t, synth, handle);
}

int tbf_gpu_synth_stats(const tbf_ctx *ctx, uint64_t *flagged, uint64_t *differed)
{
    if (!ctx)
        return TBF_ERR_ARG;
    if (finish_all_pending_synth(const_cast<tbf_ctx *>(ctx)) != TBF_OK)
        return TBF_ERR_CUDA;
    if (flagged)
        *flagged = ctx->synth_flagged;
    if (differed)
        *differed = ctx->synth_differed;
    return TBF_OK;
}

/* the raw waves of a set (copy 0 of every wheel, wave_offset layout): parity tests of the
 * device-side synthesis */
int tbf_gpu_tonegen_tables_read_waves(tbf_ctx *ctx, int handle, float *out, uint64_t n_floats)
{
    if (!ctx || !out || !tg_usable(ctx, handle))
        return fail(ctx, TBF_ERR_HANDLE, "read_waves: unknown handle");
    CK(cudaSetDevice(ctx->device));
    if (int rc = finish_all_pending_synth(ctx))
        return rc;
    CK(cudaStreamSynchronize(ctx->stream));
    const TgSet &s = ctx->tg[handle];
    size_t off4 = 0;
    for (int w = 1; w <= TBF_NOF_WHEELS; ++w) {
        const uint32_t len = s.host.wave_length[w];
        if (!len)
            continue;
        if ((uint64_t)s.host.wave_offset[w] + len > n_floats)
            return fail(ctx, TBF_ERR_ARG, "read_waves: buffer too small");
        CK(cudaMemcpy(out + s.host.wave_offset[w], s.wave + off4, sizeof(float) * len, cudaMemcpyDeviceToHost));
        off4 += (size_t)4 * tbf_wave_copy_len(len);
    }
    return TBF_OK;
}

int tbf_gpu_whirl_tables_create(tbf_ctx *ctx, const tbf_whirl_tables *w, int *handle)
{
    if (!ctx || !w || !handle || !w->hn_fwd_displ || !w->hn_bwd_displ || !w->dr_fwd_displ ||
        !w->dr_bwd_displ || !w->bfw || !w->bbw)
        return fail(ctx, TBF_ERR_ARG, "whirl tables: null argument");
    if (!(w->sample_rate > 0))
        return fail(ctx, TBF_ERR_ARG, "whirl tables: bad sample rate");
    CK(cudaSetDevice(ctx->device));
    WhSet s;
    s.host = *w;
    const size_t N = TBF_WHIRL_DISPLC;
    /* device layout: 4 displacement tables as pairs (2N floats each), then the two FIR banks
     * as five planes of N floats (coefficient c of position k at [c][k]): the 32 lanes of a warp
     * read positions a few steps apart, so a plane read touches ~3 cache lines where an
     * array-of-positions layout touches one 32-byte sector per lane */
    {
        std::vector<float> t(N * 24, 0.0f);
        const float *src[4] = {w->hn_fwd_displ, w->hn_bwd_displ, w->dr_fwd_displ, w->dr_bwd_displ};
        for (int i = 0; i < 4; ++i)
            for (size_t k = 0; k < N; ++k) {
                t[2 * N * i + 2 * k] = src[i][k];
                t[2 * N * i + 2 * k + 1] = src[i][(k + 1) & (N - 1)];
            }
        const float *fir[2] = {w->bfw, w->bbw};
        for (int i = 0; i < 2; ++i)
            for (size_t k = 0; k < N; ++k)
                for (int c = 0; c < 5; ++c)
                    t[8 * N + 8 * N * i + (size_t)c * N + k] = fir[i][5 * k + c];
        CK(cudaMalloc(&s.tables, sizeof(float) * t.size()));
        CK(cudaMemcpy(s.tables, t.data(), sizeof(float) * t.size(), cudaMemcpyHostToDevice));
    }
    s.host.hn_fwd_displ = s.host.hn_bwd_displ = s.host.dr_fwd_displ = s.host.dr_bwd_displ = nullptr;
    s.host.bfw = s.host.bbw = nullptr;

    DevWhirlTables d;
    memset(&d, 0, sizeof(d));
    d.hn_fwd_displ = reinterpret_cast<const float2 *>(s.tables);
    d.hn_bwd_displ = reinterpret_cast<const float2 *>(s.tables + 2 * N);
    d.dr_fwd_displ = reinterpret_cast<const float2 *>(s.tables + 4 * N);
    d.dr_bwd_displ = reinterpret_cast<const float2 *>(s.tables + 6 * N);
    d.bfw = s.tables + 8 * N;
    d.bbw = s.tables + 16 * N;
    d.sample_rate = w->sample_rate;
    d.fw_ang = w->mic_angle * .25;            /* src/whirl.cpp:1379 */
    d.bw_ang = 1. + w->mic_angle * -.25;      /* src/whirl.cpp:1380 */
    const float tau[4] = {w->horn_acc, w->horn_dec, w->drum_acc, w->drum_dec};
    for (int i = 0; i < 4; ++i) {
        d.one_minus_l[i] = ramp_one_minus_l(w->sample_rate, tau[i]);
        for (int v = 0; v < 128; ++v) {
            /* setHornAcceleration ... setDrumDeceleration, src/whirl.cpp:895-917: the
             * sum is a double, the field a float */
            const float t = (float)(.01 + (double)v / (i < 2 ? 80.0 : 14.0));
            d.ramp_one_minus_l[i][v] = ramp_one_minus_l(w->sample_rate, t);
        }
    }
    for (int i = 0; i < 9; ++i) {
        d.rev_horn_target[i] = w->rev_horn_target[i];
        d.rev_drum_target[i] = w->rev_drum_target[i];
    }
    d.hn_brake_pos = w->hn_brake_pos;
    d.dr_brake_pos = w->dr_brake_pos;
    for (int i = 0; i < 6; ++i) {
        d.horn_phase[i] = w->horn_phase[i];
        d.drum_phase[i] = w->drum_phase[i];
        d.horn_spacing[i] = w->horn_spacing[i];
        d.drum_spacing[i] = w->drum_spacing[i];
    }
    memcpy(d.iir[0], w->hafw, sizeof(float) * 5);
    memcpy(d.iir[1], w->hbfw, sizeof(float) * 5);
    memcpy(d.iir[2], w->drfL, sizeof(float) * 5);
    memcpy(d.iir[3], w->drfR, sizeof(float) * 5);
    for (int i = 0; i < 3; ++i) {
        if (w->revselects[i] < 0) {
            cudaFree(s.tables);
            return fail(ctx, TBF_ERR_ARG, "whirl tables: negative revselect");
        }
        d.revselects[i] = w->revselects[i];
    }
    d.horn_level = w->horn_level;
    d.leakage = w->leakage;
    d.mic_hll = w->mic_hll;
    d.mic_hlr = w->mic_hlr;
    d.mic_hrl = w->mic_hrl;
    d.mic_hrr = w->mic_hrr;
    d.mic_dll = w->mic_dll;
    d.mic_dlr = w->mic_dlr;
    d.mic_drl = w->mic_drl;
    d.mic_drr = w->mic_drr;
    d.bypass = w->bypass;

    /* Delay-line geometry: how far ahead of the reader each tap can write
     * (t = spacing + displacement + outpos, slot = floor(t), src/whirl.cpp:1432-1469).
     * One slot of margin on each side covers the float rounding of the sum; the
     * secondary slot (n + 1) is included in dmax. */
    {
        const float *disp[12];
        for (int p = 0; p < 6; ++p) {
            const bool fwd = (p == 0 || p == 3 || p == 4);
            disp[p] = fwd ? w->hn_fwd_displ : w->hn_bwd_displ;
            disp[6 + p] = fwd ? w->dr_fwd_displ : w->dr_bwd_displ;
        }
        int wmax = 0;
        bool ok = true;
        for (int p = 0; p < 12; ++p) {
            float mn = disp[p][0], mx = disp[p][0];
            for (size_t i = 1; i < N; ++i) {
                mn = std::min(mn, disp[p][i]);
                mx = std::max(mx, disp[p][i]);
            }
            const float sp = p < 6 ? w->horn_spacing[p] : w->drum_spacing[p - 6];
            if (!(mn == mn) || !(mx == mx) || !(sp == sp)) {
                cudaFree(s.tables);
                return fail(ctx, TBF_ERR_GEOMETRY, "whirl tables: NaN in displacement tables");
            }
            /* the float sum spacing + displacement + outpos (< 4096) is within 1e-3 of
             * the exact value; +1 is the secondary slot */
            d.tap_dmin[p] = (int)floor((double)sp + (double)mn - 0.01);
            d.tap_dmax[p] = (int)floor((double)sp + (double)mx + 0.01) + 1;
            if (d.tap_dmin[p] < 0)
                ok = false;
            wmax = std::max(wmax, d.tap_dmax[p]);
        }
        /* the kernels scatter 32 samples, then read them: after a block's scatter the
         * live slots reach wmax + 32 ahead of the block's first read */
        d.window = wmax + 1 + TBF_SCAT_BLOCK;
        if (d.window >= TBF_WHIRL_BUF - TBF_SCAT_BLOCK) {
            cudaFree(s.tables);
            return fail(ctx, TBF_ERR_GEOMETRY,
                        "whirl tables: tap delays do not fit the 2048-sample delay lines");
        }
        for (int line = 0; line < 4; ++line) {
            int t[3];
            const int base = (line < 2) ? (line & 1) : 6 + (line & 1);
            for (int k = 0; k < 3; ++k)
                t[k] = base + 2 * k;
            std::sort(t, t + 3, [&](int a, int b) { return d.tap_dmin[a] > d.tap_dmin[b]; });
            for (int k = 0; k < 3; ++k)
                d.tap_order[line][k] = t[k];
            for (int k = 0; k + 1 < 3; ++k) /* farther tap strictly beyond the nearer one */
                if (!(d.tap_dmin[t[k]] > d.tap_dmax[t[k + 1]]))
                    ok = false;
        }
        /* shared-memory ring per delay line: the smallest power of two that holds the
         * live window; the serial path keeps the whole 2048-slot line */
        d.gather_ok = ok ? 1 : 0;
        d.ring_size = TBF_WHIRL_BUF;
        if (ok)
            for (int r = 512; r <= TBF_WHIRL_BUF; r *= 2)
                if (d.window <= r) {
                    d.ring_size = r;
                    break;
                }
        /* conversion-free index math of stage W (tap_position<true>): phases small and non-
         * negative, microphone angle offsets within [0, 64), delays non-negative */
        bool small = ok && d.fw_ang >= 0.0 && d.fw_ang < 64.0 && d.bw_ang >= 0.0 && d.bw_ang < 64.0;
        for (int i = 0; i < 6; ++i)
            small = small && d.horn_phase[i] >= 0 && d.horn_phase[i] < (1 << 21) && d.drum_phase[i] >= 0 &&
                    d.drum_phase[i] < (1 << 21);
        d.small_index = small ? 1 : 0;
        d.fast_max_angle = 1.0;
        /* Can a tap's destination slot ever fail the parallel scatter's per-block checks (slots
         * monotone in the sample index, at most two samples per slot, inside the window)?  Not if
         * its delay changes by less than half a sample per sample.  The delay is piecewise linear
         * in the rotor angle (16384 steps per turn), the angle moves |incr| * 16384 steps per
         * sample, and |incr| never exceeds the largest speed the set knows: the ramp
         * incr += (1 - l)(target - incr), 0 < 1 - l < 1, approaches its target from one side
         * (src/whirl.cpp:1255-1257), the brake moves at most 60 / 100 rpm (:1333-1374).  The float
         * rounding of the table index (1/128 of a step at most) rides on the step count. */
        {
            double vmax[2] = {fabs(w->init_horn_incr), fabs(w->init_drum_incr)};
            for (int i = 0; i < 9; ++i) {
                vmax[0] = std::max(vmax[0], fabs(w->rev_horn_target[i]));
                vmax[1] = std::max(vmax[1], fabs(w->rev_drum_target[i]));
            }
            vmax[0] = std::max(vmax[0], 60.0 / (60.0 * w->sample_rate));
            vmax[1] = std::max(vmax[1], 100.0 / (60.0 * w->sample_rate));
            bool cert = ok;
            for (int p = 0; p < 12 && cert; ++p) {
                double slope = 0.0;
                for (size_t i = 0; i < N; ++i)
                    slope = std::max(slope, fabs((double)disp[p][(i + 1) & (N - 1)] - (double)disp[p][i]));
                const double steps = vmax[p < 6 ? 0 : 1] * (double)N + 0.02;
                cert = slope * steps <= 0.49 && vmax[p < 6 ? 0 : 1] < 0.25;
            }
            d.taps_certified = cert ? 1 : 0;
        }
        s.ring_size = d.ring_size;
        s.gather_ok = d.gather_ok;
        s.window = d.window;
    }

    const bool reuse_wh = !ctx->free_wh.empty();
    const size_t idx = reuse_wh ? (size_t)ctx->free_wh.back() : ctx->wh.size();
    if (idx + 1 > ctx->d_wh_cap) {
        const size_t ncap = std::max<size_t>(8, ctx->d_wh_cap * 2);
        DevWhirlTables *nd = nullptr;
        CK(cudaMalloc(&nd, sizeof(DevWhirlTables) * ncap));
        if (ctx->wh.size())
            cudaMemcpy(nd, ctx->d_wh, sizeof(DevWhirlTables) * ctx->wh.size(), cudaMemcpyDeviceToDevice);
        cudaFree(ctx->d_wh);
        ctx->d_wh = nd;
        ctx->d_wh_cap = ncap;
    }
    CK(cudaMemcpy(ctx->d_wh + idx, &d, sizeof(d), cudaMemcpyHostToDevice));
    if (reuse_wh) {
        ctx->free_wh.pop_back();
        ctx->wh[idx] = s;
    } else {
        ctx->wh.push_back(s);
    }
    *handle = (int)idx;
    return TBF_OK;
}

/* the state allocTonegen+initToneGenerator+init_vibrato and allocWhirl+initWhirl
 * leave behind (src/tonegen.cpp:196-300, :2905-3022; src/vibrato.cpp:288-317;
 * src/whirl.cpp:40-131, :626-662): everything not set here starts as zero */
static void fill_init(InstInit *s, const tbf_tonegen_tables &t, int tgh, const tbf_whirl_tables &w, int whh)
{
    memset(s, 0, sizeof(*s));
    s->tg.newRouting = s->tg.oldRouting = t.init_routing;
    s->tg.percSendBus = t.init_perc_send_bus;
    s->tg.swellPedalGain = t.init_swell_pedal_gain;
    s->tg.drawBarChange = t.init_draw_bar_change;
    s->tg.percEnabled = t.init_perc_enabled;
    s->tg.percTrigRestore = t.init_perc_trig_restore;
    s->tg.percIsSoft = t.init_perc_is_soft;
    s->tg.percIsFast = t.init_perc_is_fast;
    s->rp.percEnvGain = 0.0f;
    s->tg.percEnvGainReset = t.init_perc_env_gain_reset;
    s->tg.percEnvGainDecay = t.init_perc_env_gain_decay;
    s->tg.percDrawbarGain = t.init_perc_drawbar_gain;
    s->tg.outputGain = 1.0f;
    s->tg.pz = 0.0f;
    s->rp.keyCompLevel = 1.0f; /* KEYCOMP_ZERO_LEVEL */
    memcpy(s->tg.drawBarGain, t.init_draw_bar_gain, sizeof(s->tg.drawBarGain));
    s->tg.vibStator = 0;
    s->tg.vibOutPos = 0x3FF / 2; /* BUF_MASK_SAMPLES / 2, src/vibrato.cpp:294 */
    s->tg.vibTable = t.init_vib_table;
    s->tg.vibMixed = t.init_vib_mixed;
    s->rp.hornAngle = w.init_horn_angle;
    s->rp.drumAngle = w.init_drum_angle;
    s->rp.hornIncr = w.init_horn_incr;
    s->rp.drumIncr = w.init_drum_incr;
    s->rp.hornTarget = w.init_horn_target;
    s->rp.drumTarget = w.init_drum_target;
    s->rp.hornAcDc = w.init_horn_acdc;
    s->rp.drumAcDc = w.init_drum_acdc;
    s->rp.revSelect = w.init_rev_select;
    memcpy(s->rp.iir[0], w.hafw, sizeof(float) * 5);
    memcpy(s->rp.iir[1], w.hbfw, sizeof(float) * 5);
    s->rp.brakePos[0] = w.hn_brake_pos;
    s->rp.brakePos[1] = w.dr_brake_pos;
    const float tau[4] = {w.horn_acc, w.horn_dec, w.drum_acc, w.drum_dec};
    for (int i = 0; i < 4; ++i)
        s->rp.oml[i] = ramp_one_minus_l(w.sample_rate, tau[i]);
    s->tg.tgTables = tgh;
    s->tg.whTables = whh;
}

/* stream-ordered: the initialising kernel takes its few hundred bytes as a kernel argument,
 * so creating or resetting instances costs a launch, no upload and no synchronisation */
static int replicate(tbf_ctx *ctx, int tgh, int whh, uint32_t first, uint32_t count)
{
    InstInit init;
    fill_init(&init, ctx->tg[tgh].host, tgh, ctx->wh[whh].host, whh);
    const cudaError_t e = (cudaError_t)tbf_launch_init_instances(ctx->d_states, first, count, init, ctx->stream);
    if (e != cudaSuccess) {
        ctx->err = std::string("instance init: ") + cudaGetErrorString(e);
        return TBF_ERR_CUDA;
    }
    return TBF_OK;
}

static bool tg_usable(const tbf_ctx *ctx, int h)
{
    return h >= 0 && (size_t)h < ctx->tg.size() && !ctx->tg[h].destroyed && !ctx->tg[h].freed;
}
static bool wh_usable(const tbf_ctx *ctx, int h)
{
    return h >= 0 && (size_t)h < ctx->wh.size() && !ctx->wh[h].destroyed && !ctx->wh[h].freed;
}

/* device memory of a table set nobody can reach any more; its handle is recycled */
static void tg_release_if_unused(tbf_ctx *ctx, int h)
{
    TgSet &t = ctx->tg[h];
    if (t.freed || !t.destroyed || t.refs > 0)
        return;
    /* back to the arena: its next user writes it with work ordered on the library's stream,
     * which every render has joined by then */
    ctx->arena.release(reinterpret_cast<unsigned char *>(t.wave), t.blob_bytes);
    t.wave = nullptr;
    t.freed = true;
    ctx->free_tg.push_back(h);
}
static void wh_release_if_unused(tbf_ctx *ctx, int h)
{
    WhSet &w = ctx->wh[h];
    if (w.freed || !w.destroyed || w.refs > 0)
        return;
    cudaFree(w.tables);
    w.tables = nullptr;
    w.freed = true;
    ctx->free_wh.push_back(h);
}

int tbf_gpu_instance_create(tbf_ctx *ctx, int tgh, int whh, uint32_t count, uint32_t *first_id)
{
    if (!ctx || !count)
        return fail(ctx, TBF_ERR_ARG, "instance_create: bad argument");
    if (!tg_usable(ctx, tgh) || !wh_usable(ctx, whh))
        return fail(ctx, TBF_ERR_HANDLE, "instance_create: unknown table handle");
    if (ctx->in_flight)
        return fail(ctx, TBF_ERR_ARG, "instance_create: a launch is in flight");
    CK(cudaSetDevice(ctx->device));
    /* first fit among the slots of destroyed instances, else append */
    size_t first = ctx->n_inst;
    if (ctx->n_inst - ctx->n_live >= count) {
        size_t run = 0;
        for (size_t i = 0; i < ctx->n_inst; ++i) {
            run = ctx->inst_alive[i] ? 0 : run + 1;
            if (run == count) {
                first = i + 1 - count;
                break;
            }
        }
    }
    const size_t need = std::max(ctx->n_inst, first + count);
    if (need > ctx->cap_inst) {
        size_t ncap = std::max<size_t>(need, ctx->cap_inst * 2);
        InstState *nd = nullptr;
        cudaError_t e = cudaMalloc(&nd, sizeof(InstState) * ncap);
        if (e != cudaSuccess) {
            ctx->err = std::string("instance_create: ") + cudaGetErrorString(e);
            cudaGetLastError();
            return TBF_ERR_NOMEM;
        }
        if (ctx->n_inst) {
            /* instance init kernels are stream-ordered on the library's (non-blocking) stream:
             * they must have written the old array before it is copied */
            CK(cudaStreamSynchronize(ctx->stream));
            CK(cudaMemcpy(nd, ctx->d_states, sizeof(InstState) * ctx->n_inst, cudaMemcpyDeviceToDevice));
        }
        cudaFree(ctx->d_states);
        ctx->d_states = nd;
        ctx->cap_inst = ncap;
    }
    int rc = replicate(ctx, tgh, whh, (uint32_t)first, count);
    if (rc != TBF_OK)
        return rc;
    if (need > ctx->n_inst) {
        ctx->n_inst = need;
        ctx->inst_tg.resize(need, -1);
        ctx->inst_wh.resize(need, -1);
        ctx->inst_alive.resize(need, 0);
    }
    for (size_t i = first; i < first + count; ++i) {
        ctx->inst_tg[i] = tgh;
        ctx->inst_wh[i] = whh;
        ctx->inst_alive[i] = 1;
    }
    ctx->n_live += count;
    ctx->tg[tgh].refs += (int)count;
    ctx->wh[whh].refs += (int)count;
    if (first_id)
        *first_id = (uint32_t)first;
    return TBF_OK;
}

int tbf_gpu_instance_destroy(tbf_ctx *ctx, uint32_t first_id, uint32_t count)
{
    if (!ctx || (size_t)first_id + count > ctx->n_inst)
        return fail(ctx, TBF_ERR_HANDLE, "instance_destroy: range out of bounds");
    if (ctx->in_flight)
        return fail(ctx, TBF_ERR_ARG, "instance_destroy: a launch is in flight");
    for (size_t i = first_id; i < (size_t)first_id + count; ++i)
        if (!ctx->inst_alive[i])
            return fail(ctx, TBF_ERR_HANDLE, "instance_destroy: instance already destroyed");
    CK(cudaSetDevice(ctx->device));
    if (int rc = finish_all_pending_synth(ctx))
        return rc;
    ctx->has_prepared = false; /* a prepared batch may list the instance */
    for (size_t i = first_id; i < (size_t)first_id + count; ++i) {
        const int tgh = ctx->inst_tg[i], whh = ctx->inst_wh[i];
        ctx->inst_alive[i] = 0;
        ctx->tg[tgh].refs -= 1;
        ctx->wh[whh].refs -= 1;
        tg_release_if_unused(ctx, tgh);
        wh_release_if_unused(ctx, whh);
    }
    ctx->n_live -= count;
    /* trailing free slots are given back, so ids stay small on a host that re-creates organs */
    while (ctx->n_inst && !ctx->inst_alive[ctx->n_inst - 1])
        ctx->n_inst -= 1;
    ctx->inst_tg.resize(ctx->n_inst);
    ctx->inst_wh.resize(ctx->n_inst);
    ctx->inst_alive.resize(ctx->n_inst);
    return TBF_OK;
}

int tbf_gpu_tonegen_tables_destroy(tbf_ctx *ctx, int handle)
{
    if (!ctx || !tg_usable(ctx, handle))
        return fail(ctx, TBF_ERR_HANDLE, "tonegen_tables_destroy: unknown handle");
    if (ctx->in_flight)
        return fail(ctx, TBF_ERR_ARG, "tonegen_tables_destroy: a launch is in flight");
    CK(cudaSetDevice(ctx->device));
    if (int rc = finish_all_pending_synth(ctx))
        return rc;
    ctx->tg[handle].destroyed = true;
    tg_release_if_unused(ctx, handle);
    return TBF_OK;
}

int tbf_gpu_whirl_tables_destroy(tbf_ctx *ctx, int handle)
{
    if (!ctx || !wh_usable(ctx, handle))
        return fail(ctx, TBF_ERR_HANDLE, "whirl_tables_destroy: unknown handle");
    if (ctx->in_flight)
        return fail(ctx, TBF_ERR_ARG, "whirl_tables_destroy: a launch is in flight");
    CK(cudaSetDevice(ctx->device));
    ctx->wh[handle].destroyed = true;
    wh_release_if_unused(ctx, handle);
    return TBF_OK;
}

int tbf_gpu_resource_counts(const tbf_ctx *ctx, uint32_t *live_instances, uint32_t *instance_slots,
                            uint32_t *live_tonegen_sets, uint32_t *live_whirl_sets)
{
    if (!ctx)
        return TBF_ERR_ARG;
    if (live_instances)
        *live_instances = (uint32_t)ctx->n_live;
    if (instance_slots)
        *instance_slots = (uint32_t)ctx->n_inst;
    uint32_t nt = 0, nw = 0;
    for (auto &t : ctx->tg)
        nt += t.freed ? 0 : 1;
    for (auto &w : ctx->wh)
        nw += w.freed ? 0 : 1;
    if (live_tonegen_sets)
        *live_tonegen_sets = nt;
    if (live_whirl_sets)
        *live_whirl_sets = nw;
    return TBF_OK;
}

int tbf_gpu_instance_stats(tbf_ctx *ctx, const uint32_t *ids, uint32_t n, uint64_t *fragments,
                           uint64_t *active_wheels)
{
    if (!ctx || !n || (!ids && n > ctx->n_inst))
        return fail(ctx, TBF_ERR_ARG, "instance_stats: bad argument");
    if (ctx->in_flight)
        return fail(ctx, TBF_ERR_ARG, "instance_stats: synchronize first");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    /* the two 64-bit counters of every instance in one strided copy */
    std::vector<unsigned long long> st(2 * ctx->n_inst);
    CK(cudaMemcpy2D(st.data(), 16, &ctx->d_states[0].tg.statFragments, sizeof(InstState), 16, ctx->n_inst,
                    cudaMemcpyDeviceToHost));
    uint64_t f = 0, w = 0;
    for (uint32_t i = 0; i < n; ++i) {
        const uint32_t id = ids ? ids[i] : i;
        if (id >= ctx->n_inst || !ctx->inst_alive[id])
            return fail(ctx, TBF_ERR_HANDLE, "instance_stats: unknown instance id");
        f += st[2 * (size_t)id];
        w += st[2 * (size_t)id + 1];
    }
    if (fragments)
        *fragments = f;
    if (active_wheels)
        *active_wheels = w;
    return TBF_OK;
}

int tbf_gpu_instance_count(const tbf_ctx *ctx) { return ctx ? (int)ctx->n_inst : TBF_ERR_ARG; }

int tbf_gpu_instance_reset(tbf_ctx *ctx, uint32_t first_id, uint32_t count)
{
    if (!ctx || (size_t)first_id + count > ctx->n_inst)
        return fail(ctx, TBF_ERR_HANDLE, "instance_reset: range out of bounds");
    if (ctx->in_flight)
        return fail(ctx, TBF_ERR_ARG, "instance_reset: a launch is in flight");
    for (size_t k = first_id; k < (size_t)first_id + count; ++k)
        if (!ctx->inst_alive[k])
            return fail(ctx, TBF_ERR_HANDLE, "instance_reset: instance destroyed");
    CK(cudaSetDevice(ctx->device));
    uint32_t i = first_id;
    const uint32_t end = first_id + count;
    while (i < end) { /* runs of instances sharing both table sets */
        uint32_t j = i + 1;
        while (j < end && ctx->inst_tg[j] == ctx->inst_tg[i] && ctx->inst_wh[j] == ctx->inst_wh[i])
            ++j;
        int rc = replicate(ctx, ctx->inst_tg[i], ctx->inst_wh[i], i, j - i);
        if (rc != TBF_OK)
            return rc;
        i = j;
    }
    return TBF_OK;
}

} /* extern "C" */

template <typename T> static cudaError_t grow(T **p, size_t *cap, size_t need)
{
    if (need <= *cap)
        return cudaSuccess;
    cudaFree(*p);
    *p = nullptr;
    *cap = 0;
    cudaError_t e = cudaMalloc(p, sizeof(T) * need);
    if (e == cudaSuccess)
        *cap = need;
    return e;
}

extern "C" {

/* Validate a batch, upload its events / ids and reset the per-slot event cursors.
 * The prepared batch can then be launched (and re-launched: the instances simply
 * keep playing) without touching the host data again. */
/* Events must come sorted by (slot, fragment); start[s] = index of the first event of a slot >= s
 * (start[n_instances] = n_events).  One compare of a 64-bit key per event; the list is cut into
 * ranges for a few threads, because at millions of events per batch the pass is bound by one core's
 * memory bandwidth and it is host time inside every end-to-end step.  Returns what is wrong, or null. */
static const char *event_csr(const tbf_event *ev, uint64_t n_events, uint32_t n_instances, uint32_t n_fragments,
                             uint64_t *start)
{
    static const char *const unsorted = "render: events not sorted by (slot, fragment)";
    static const char *const range = "render: event slot or fragment out of range";
    struct Part {
        bool sorted = true, bad_slot = false;
        uint32_t frag_max = 0;
    };
    auto scan = [&](uint64_t lo, uint64_t hi, Part *r) {
        uint32_t prev_slot = lo ? ev[lo - 1].instance : 0;
        uint64_t prev_key = lo ? (((uint64_t)ev[lo - 1].instance << 32) | ev[lo - 1].fragment) : 0;
        bool sorted = true;
        uint32_t frag_max = 0;
        for (uint64_t e = lo; e < hi; ++e) {
            const uint32_t slot = ev[e].instance, frag = ev[e].fragment;
            const uint64_t key = ((uint64_t)slot << 32) | frag;
            sorted = sorted && key >= prev_key;
            prev_key = key;
            frag_max = frag > frag_max ? frag : frag_max;
            if (slot != prev_slot) {
                if (slot < prev_slot || slot >= n_instances) { /* nothing is written past the array */
                    r->bad_slot = true;
                    r->sorted = sorted && slot > prev_slot;
                    return;
                }
                for (uint32_t s = prev_slot + 1; s <= slot; ++s)
                    start[s] = e;
                prev_slot = slot;
            }
        }
        r->sorted = sorted;
        r->frag_max = frag_max;
    };
    start[0] = 0;
    const unsigned hw = std::thread::hardware_concurrency();
    const unsigned n_thr = n_events < (1u << 18) ? 1u : std::min(8u, std::max(1u, hw));
    std::vector<Part> parts(n_thr);
    if (n_thr == 1) {
        scan(0, n_events, &parts[0]);
    } else {
        std::vector<std::thread> thr;
        for (unsigned t = 0; t < n_thr; ++t)
            thr.emplace_back(scan, n_events * t / n_thr, n_events * (t + 1) / n_thr, &parts[t]);
        for (auto &t : thr)
            t.join();
    }
    for (const Part &r : parts) { /* the same verdicts, in the same order, as one pass in sequence */
        if (!r.sorted)
            return unsorted;
        if (r.bad_slot || r.frag_max >= n_fragments)
            return range;
    }
    if (n_events && ev[0].instance >= n_instances)
        return range;
    for (uint32_t s = (n_events ? ev[n_events - 1].instance : 0) + 1; s <= n_instances; ++s)
        start[s] = n_events;
    return nullptr;
}

int tbf_gpu_event_index(const tbf_event *events, uint64_t n_events, uint32_t n_instances,
                        uint32_t n_fragments, uint64_t *start, const char **why)
{
    const char *w = nullptr;
    if (!start || !n_instances || !n_fragments || (n_events && !events))
        w = "event_index: null argument or empty batch";
    else
        w = event_csr(events, n_events, n_instances, n_fragments, start);
    if (why)
        *why = w;
    return w ? TBF_ERR_ARG : TBF_OK;
}

int tbf_gpu_batch_prepare(tbf_ctx *ctx, const tbf_batch_desc *d)
{
    if (!ctx || !d)
        return fail(ctx, TBF_ERR_ARG, "render: null argument");
    if (ctx->in_flight)
        return fail(ctx, TBF_ERR_ARG, "render: previous launch not synchronized");
    ctx->has_prepared = false;
    if (int rc = finish_all_pending_synth(ctx)) /* table sets still awaiting their libm check */
        return rc;
    if (!d->n_instances || !d->n_fragments)
        return fail(ctx, TBF_ERR_ARG, "render: empty batch");
    const bool no_out = (d->flags & TBF_RENDER_NO_OUTPUT) != 0;
    const bool dev_out = (d->flags & TBF_RENDER_OUT_DEVICE) != 0;
    const uint64_t frames = (uint64_t)d->n_fragments * TBF_FRAGMENT;
    if (!no_out && (!d->out_l || !d->out_r || d->out_stride < frames))
        return fail(ctx, TBF_ERR_ARG, "render: output buffers missing or stride too small");
    if (frames > 0xFFFFFFFFull)
        return fail(ctx, TBF_ERR_ARG, "render: batch longer than 2^32 samples");
    if (d->flags & TBF_RENDER_WHIRL_INPUT) {
        if (d->flags & TBF_RENDER_WHIRL_BYPASS)
            return fail(ctx, TBF_ERR_ARG, "render: WHIRL_INPUT and WHIRL_BYPASS exclude each other");
        if (!d->in || d->in_stride < frames)
            return fail(ctx, TBF_ERR_ARG, "render: input buffer missing or stride too small");
    }
    if (d->n_events && !d->events)
        return fail(ctx, TBF_ERR_ARG, "render: events missing");
    if ((d->flags & TBF_RENDER_STEMS) && (d->flags & (TBF_RENDER_WHIRL_BYPASS | TBF_RENDER_NO_OUTPUT)))
        return fail(ctx, TBF_ERR_ARG, "render: STEMS needs the Leslie and output buffers");
    if ((d->flags & TBF_RENDER_OUT_CHUNKED) && (no_out || dev_out))
        return fail(ctx, TBF_ERR_ARG, "render: OUT_CHUNKED is a host-output layout");
    if (d->instance_ids) {
        /* an instance may appear once per batch: one CTA owns its state */
        std::vector<uint8_t> seen(ctx->n_inst, 0);
        for (uint32_t i = 0; i < d->n_instances; ++i) {
            const uint32_t id = d->instance_ids[i];
            if (id >= ctx->n_inst || !ctx->inst_alive[id])
                return fail(ctx, TBF_ERR_HANDLE, "render: unknown instance id");
            if (seen[id])
                return fail(ctx, TBF_ERR_ARG, "render: instance listed twice");
            seen[id] = 1;
        }
    } else if (d->n_instances > ctx->n_inst) {
        return fail(ctx, TBF_ERR_HANDLE, "render: more instances than created");
    } else {
        for (uint32_t i = 0; i < d->n_instances; ++i)
            if (!ctx->inst_alive[i])
                return fail(ctx, TBF_ERR_HANDLE, "render: instance destroyed");
    }
    /* stage W's shared-memory delay-line ring: the largest any whirl set OF THIS BATCH needs */
    {
        int ring = 512;
        for (uint32_t i = 0; i < d->n_instances; ++i)
            ring = std::max(ring, ctx->wh[ctx->inst_wh[d->instance_ids ? d->instance_ids[i] : i]].ring_size);
        ctx->prepared_ring = ring;
    }
    /* how many tone-generator table sets the batch's organs share: stage T picks its sample mover by
     * it (many organs per set: the waves stay in L2; one set per organ: they come from HBM) */
    {
        std::vector<uint8_t> used(ctx->tg.size(), 0);
        uint32_t sets = 0;
        for (uint32_t i = 0; i < d->n_instances; ++i) {
            const int h = ctx->inst_tg[d->instance_ids ? d->instance_ids[i] : i];
            if (h >= 0 && (size_t)h < used.size() && !used[h]) {
                used[h] = 1;
                ++sets;
            }
        }
        ctx->prepared_tg_sets = sets;
    }
    /* the events start crossing the link while the host checks them and builds the per-slot CSR */
    CK(cudaSetDevice(ctx->device));
    CK(grow(&ctx->d_events, &ctx->cap_events, std::max<uint64_t>(1, d->n_events)));
    struct DrainOnReturn { /* the caller may free its arrays once we return, on every path */
        cudaStream_t s;
        ~DrainOnReturn() { cudaStreamSynchronize(s); }
    } drain{ctx->stream};
    if (d->n_events)
        CK(cudaMemcpyAsync(ctx->d_events, d->events, sizeof(tbf_event) * d->n_events,
                           cudaMemcpyHostToDevice, ctx->stream));
    std::vector<uint64_t> &evs = ctx->ev_start_host;
    evs.resize((size_t)d->n_instances + 1);
    if (const char *why = event_csr(d->events, d->n_events, d->n_instances, d->n_fragments, evs.data()))
        return fail(ctx, TBF_ERR_ARG, why);
    if ((size_t)d->n_instances + 1 > ctx->cap_slots) {
        size_t c0 = 0, c1 = 0, c2 = 0, c3 = 0;
        CK(grow(&ctx->d_ev_start, &c0, (size_t)d->n_instances + 1));
        CK(grow(&ctx->d_ev_cursor, &c3, 3 * ((size_t)d->n_instances + 1)));
        CK(grow(&ctx->d_ids, &c1, (size_t)d->n_instances + 1));
        CK(grow(&ctx->d_csum, &c2, (size_t)d->n_instances + 1));
        ctx->cap_slots = (size_t)d->n_instances + 1;
    }
    /* host outputs are streamed back in chunks of fragments through two device
     * staging buffers, so the device-to-host copy of chunk i overlaps the render
     * of chunk i+1; device outputs (or none) are one launch */
    uint32_t chunk = d->n_fragments;
    if (!no_out && !dev_out) {
        chunk = (d->flags & TBF_RENDER_OUT_CHUNKED)
                    ? ctx->host_chunk(true) /* the caller's layout is cut by this number */
                    : std::min<uint32_t>(d->n_fragments, ctx->host_chunk(false));
        const size_t planes = (d->flags & TBF_RENDER_STEMS) ? 6 : 2; /* L R (+ HL HR DL DR) */
        const size_t need = (size_t)2 /*buffers*/ * planes * d->n_instances * chunk * TBF_FRAGMENT;
        cudaError_t e = grow(&ctx->d_out, &ctx->cap_out, need);
        if (e != cudaSuccess) {
            cudaGetLastError();
            ctx->err = "render: device output staging buffer does not fit";
            return TBF_ERR_NOMEM;
        }
    }
    CK(cudaMemcpyAsync(ctx->d_ev_start, evs.data(), sizeof(uint64_t) * evs.size(),
                       cudaMemcpyHostToDevice, ctx->stream));
    if (d->instance_ids)
        CK(cudaMemcpyAsync(ctx->d_ids, d->instance_ids, sizeof(uint32_t) * d->n_instances,
                           cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream)); /* the host arrays may be freed after prepare */
    ctx->prepared = *d;
    ctx->prepared.events = nullptr;
    ctx->prepared.instance_ids = nullptr;
    ctx->prepared_has_ids = d->instance_ids != nullptr;
    ctx->prepared_chunk = chunk;
    ctx->has_prepared = true;
    return TBF_OK;
}

/* bytes of stage streams per (slot, fragment): FragCtl + 8 float planes + 2 double planes */
static size_t stream_bytes_per_fragment(bool whirl)
{
    const size_t planes32 = whirl ? 8 : 2, planes64 = whirl ? 2 : 0;
    return sizeof(FragCtl) + (size_t)TBF_FRAGMENT * (4 * planes32 + 8 * planes64);
}

/* Launch the prepared batch on the library's stream (asynchronous).  Every chunk of
 * fragments runs the four stage kernels T -> R -> W -> D (T -> R when the Leslie is
 * bypassed for the whole batch). */
int tbf_gpu_batch_launch(tbf_ctx *ctx)
{
    if (!ctx || !ctx->has_prepared)
        return fail(ctx, TBF_ERR_ARG, "launch: no prepared batch");
    if (ctx->in_flight)
        return fail(ctx, TBF_ERR_ARG, "launch: previous launch not synchronized");
    const tbf_batch_desc *d = &ctx->prepared;
    const bool no_out = (d->flags & TBF_RENDER_NO_OUTPUT) != 0;
    const bool dev_out = (d->flags & TBF_RENDER_OUT_DEVICE) != 0;
    const bool host_out = !no_out && !dev_out;
    const bool whirl = (d->flags & TBF_RENDER_WHIRL_BYPASS) == 0;
    const bool whirl_input = (d->flags & TBF_RENDER_WHIRL_INPUT) != 0;
    const uint32_t chunk = ctx->prepared_chunk;
    const size_t n = d->n_instances;
    CK(cudaSetDevice(ctx->device));
    uint64_t *cur_t = ctx->d_ev_cursor, *cur_r = ctx->d_ev_cursor + ctx->cap_slots,
             *cur_rot = ctx->d_ev_cursor + 2 * ctx->cap_slots;
    CK(cudaMemcpyAsync(cur_t, ctx->d_ev_start, sizeof(uint64_t) * n, cudaMemcpyDeviceToDevice, ctx->stream));
    CK(cudaMemcpyAsync(cur_r, ctx->d_ev_start, sizeof(uint64_t) * n, cudaMemcpyDeviceToDevice, ctx->stream));
    CK(cudaMemcpyAsync(cur_rot, ctx->d_ev_start, sizeof(uint64_t) * n, cudaMemcpyDeviceToDevice, ctx->stream));
    if (d->checksums)
        CK(cudaMemsetAsync(ctx->d_csum, 0, sizeof(uint64_t) * n, ctx->stream));

    /* stage streams, n_bufs copies in rotation (chunk c uses copy c % n_bufs): as many
     * fragments per stage launch as the scratch budget allows.  With two copies the buffer
     * hand-backs close a cycle R(c) -> W(c) -> D(c) -> R(c + 2) that bounds the period of a
     * chunk by (R + W + D) / 2; the third copy breaks it */
    const size_t nb = (size_t)ctx->n_bufs;
    uint32_t sc_frag = (uint32_t)std::min<size_t>(
        ctx->max_stage_chunk,
        std::max<size_t>(1, ctx->scratch_budget / (nb * n * stream_bytes_per_fragment(whirl))));
    sc_frag = std::min(sc_frag, chunk);
    const size_t pitch = (size_t)sc_frag * TBF_FRAGMENT;
    const size_t planes32 = whirl ? 8 : 2;
    /* rows per plane, padded to whole 32-organ tiles: a TMA box never reaches into the next plane */
    const size_t np = (n + 31) / 32 * 32;
    StageScratch &ss = ctx->scratch;
    {
        cudaError_t e = grow(&ss.ctl, &ss.cap_ctl, nb * n * sc_frag);
        if (e == cudaSuccess)
            e = grow(&ss.f32, &ss.cap_f32, nb * np * pitch * planes32);
        if (e == cudaSuccess && whirl)
            e = grow(&ss.f64, &ss.cap_f64, nb * np * pitch * 2);
        if (e != cudaSuccess) {
            cudaGetLastError();
            ctx->err = "render: stage stream buffers do not fit";
            return TBF_ERR_NOMEM;
        }
    }
    /* tensor maps of the scratch as it is laid out for this batch (TBF_NO_TMA_TILES=1: none, the
     * lane = organ kernels then move their tiles with cp.async) */
    static const bool no_tiles = getenv("TBF_NO_TMA_TILES") != nullptr;
    bool tiles = !no_tiles && !(d->flags & 0x8000u);
    const size_t rows64 = whirl ? nb * np * 2 : 0; /* a bypass-only batch has no angle planes */
    if (tiles && !(ss.maps_ok && ss.map_f32 == ss.f32 && ss.map_f64 == (whirl ? ss.f64 : nullptr) &&
                   ss.map_pitch == pitch && ss.map_rows32 == nb * np * planes32 && ss.map_rows64 == rows64)) {
        ss.maps_ok = tile_map_make(&ss.maps.m32, ss.f32, pitch, nb * np * planes32) &&
                     (!whirl || tile_map_make(&ss.maps.m64, ss.f64, 2 * pitch, rows64));
        ss.map_f32 = ss.f32;
        ss.map_f64 = whirl ? ss.f64 : nullptr;
        ss.map_pitch = pitch;
        ss.map_rows32 = nb * np * planes32;
        ss.map_rows64 = rows64;
    }
    tiles = tiles && ss.maps_ok;

    /* instance groups (multiples of 32 slots, so the lane = instance warps stay full) */
    int n_groups = std::min<int>(ctx->n_groups_max, n >= 2048 ? 4 : (n >= 512 ? 2 : 1));
    uint32_t gfirst[TBF_MAX_GROUPS + 1];
    {
        const uint32_t per = (uint32_t)(((n + n_groups - 1) / n_groups + 31) / 32 * 32);
        for (int g = 0; g <= n_groups; ++g)
            gfirst[g] = (uint32_t)std::min<size_t>(n, (size_t)per * g);
        /* experiment: uneven two-way split, group 0 = whole waves of CTAs */
        if (const char *e = getenv("TBF_GROUP0"))
            if (n_groups == 2 && (size_t)atoi(e) < n && atoi(e) > 0)
                gfirst[1] = (uint32_t)(atoi(e) / 32 * 32);
    }
    const int n_stages = whirl ? 4 : 2, last = n_stages - 1;

    RenderParams base_p;
    memset(&base_p, 0, sizeof(base_p));
    base_p.states = ctx->d_states;
    base_p.tg = ctx->d_tg;
    base_p.wh = ctx->d_wh;
    base_p.events = ctx->d_events;
    /* debugging: 0x100 every fragment takes stage T's full active-list pass, 0x200 quiet
     * fragments are summed one at a time instead of in pairs */
    base_p.flags = d->flags | (getenv("TBF_NO_FAST_PASS") ? 0x100u : 0u) | (getenv("TBF_NO_PAIRS") ? 0x200u : 0u);
    base_p.pitch = pitch;
    base_p.one = 1.0f;
    base_p.tg_sets = ctx->prepared_tg_sets;
    ctx->launches = 0;
    const int ring = ctx->prepared_ring;
    CK(cudaEventRecord(ctx->ev0, ctx->stream));
    CK(cudaEventRecord(ctx->fork, ctx->stream));
    for (int g = 0; g < n_groups; ++g) {
        for (int st = 0; st < n_stages; ++st)
            CK(cudaStreamWaitEvent(ctx->sstream[g][st], ctx->fork, 0));
        CK(cudaStreamWaitEvent(ctx->rot_stream[g], ctx->fork, 0));
    }
    const bool stems = (d->flags & TBF_RENDER_STEMS) != 0;
    const bool chunked = (d->flags & TBF_RENDER_OUT_CHUNKED) != 0;
    const size_t n_planes = stems ? 6 : 2;
    const size_t half = n_planes * n * chunk * TBF_FRAGMENT;
    uint32_t cidx[TBF_MAX_GROUPS] = {0, 0, 0, 0}; /* stage chunks launched per group */
    ctx->prof_stage.clear();
    ctx->timeline = !ctx->profile_stages && getenv("TBF_TIMELINE") != nullptr;
    auto prof_mark = [&](int stage, bool begin, cudaStream_t on) -> cudaError_t {
        if (!ctx->profile_stages && !ctx->timeline)
            return cudaSuccess;
        const size_t idx = 2 * ctx->prof_stage.size() + (begin ? 0 : 1);
        while (ctx->prof_events.size() <= idx) {
            cudaEvent_t e;
            cudaError_t rc = cudaEventCreate(&e);
            if (rc != cudaSuccess)
                return rc;
            ctx->prof_events.push_back(e);
        }
        cudaError_t rc = cudaEventRecord(ctx->prof_events[idx], on);
        if (!begin)
            ctx->prof_stage.push_back(stage);
        return rc;
    };
    int k = 0;
    for (uint32_t base = 0; base < d->n_fragments; base += chunk, ++k) {
        const uint32_t nf = std::min<uint32_t>(chunk, d->n_fragments - base);
        const int buf = k & 1;
        float *out_l, *out_r, *out_s[4] = {nullptr, nullptr, nullptr, nullptr};
        uint64_t out_stride;
        if (host_out) {
            /* chunk-major host layout: staging rows keep the full chunk length, so the whole
             * plane of a chunk is one contiguous block on both sides of the link */
            out_stride = (uint64_t)(chunked ? chunk : nf) * TBF_FRAGMENT;
            out_l = ctx->d_out + half * buf;
            out_r = out_l + n * out_stride;
            for (int j = 0; stems && j < 4; ++j)
                out_s[j] = d->stems[j] ? out_l + (size_t)(2 + j) * n * out_stride : nullptr;
        } else {
            out_stride = d->out_stride;
            out_l = no_out ? nullptr : d->out_l + (size_t)base * TBF_FRAGMENT;
            out_r = no_out ? nullptr : d->out_r + (size_t)base * TBF_FRAGMENT;
            for (int j = 0; stems && j < 4; ++j)
                out_s[j] = d->stems[j] ? d->stems[j] + (size_t)base * TBF_FRAGMENT : nullptr;
        }
        for (int g = 0; g < n_groups; ++g) {
            const size_t s0 = gfirst[g];
            const uint32_t ng = gfirst[g + 1] - gfirst[g];
            if (!ng)
                continue;
            cudaStream_t prof_st[4] = {ctx->stream, ctx->stream, ctx->stream, ctx->stream};
            cudaStream_t *st = ctx->profile_stages ? prof_st : ctx->sstream[g];
            /* staging buffer `buf` is free once the copy issued two chunks ago is done */
            if (host_out && k >= 2) {
                CK(cudaStreamWaitEvent(st[last], ctx->copy_done[buf], 0));
                CK(cudaStreamWaitEvent(st[last], ctx->copy_done2[buf], 0));
            }
            RenderParams p = base_p;
            p.n_instances = ng;
            p.ids = ctx->prepared_has_ids ? ctx->d_ids + s0 : nullptr;
            p.id_base = (uint32_t)s0;
            p.ev_start = ctx->d_ev_start + s0;
            p.ev_cursor_t = cur_t + s0;
            p.ev_cursor_r = cur_r + s0;
            p.ev_cursor_rot = cur_rot + s0;
            p.checksums = d->checksums ? ctx->d_csum + s0 : nullptr;
            p.out_stride = out_stride;
            int par = 0;
            for (uint32_t ib = 0; ib < nf; ib += sc_frag) {
                const uint32_t c = cidx[g]++;
                par = (int)(c % (uint32_t)nb);
                float *f32 = ss.f32 + (size_t)par * np * pitch * planes32;
                p.ctl = ss.ctl + ((size_t)par * n + s0) * sc_frag;
                p.u = f32 + s0 * pitch;
                p.php = f32 + (np + s0) * pitch;
                if (whirl) {
                    p.dx0 = f32 + (2 * np + s0) * pitch;
                    p.hx0 = f32 + (3 * np + s0) * pitch;
                    for (int l = 0; l < 4; ++l)
                        p.line[l] = f32 + ((4 + l) * np + s0) * pitch;
                    double *f64 = ss.f64 + (size_t)par * np * pitch * 2;
                    p.hang = f64 + s0 * pitch;
                    p.dang = f64 + (np + s0) * pitch;
                }
                p.tile_maps = tiles ? &ss.maps : nullptr;
                for (int k = 0; k < 8; ++k) /* u php dx0 hx0 line[0..3] */
                    p.tile_row[k] = (uint32_t)(((size_t)par * planes32 + (size_t)k) * np + s0);
                p.tile_row[8] = (uint32_t)(((size_t)par * 2) * np + s0);
                p.tile_row[9] = (uint32_t)(((size_t)par * 2 + 1) * np + s0);
                p.frag_base = base + ib;
                p.n_fragments = std::min<uint32_t>(sc_frag, nf - ib);
                p.out_l = out_l ? out_l + s0 * out_stride + (size_t)ib * TBF_FRAGMENT : nullptr;
                p.out_r = out_r ? out_r + s0 * out_stride + (size_t)ib * TBF_FRAGMENT : nullptr;
                for (int j = 0; j < 4; ++j)
                    p.stems[j] = out_s[j] ? out_s[j] + s0 * out_stride + (size_t)ib * TBF_FRAGMENT : nullptr;
                cudaEvent_t(*ev)[TBF_MAX_STAGE_BUFS] = ctx->sdone[g];
                /* a stage waits for its producer of this chunk and, before it overwrites
                 * copy `par` of the streams, for their consumers of chunk c - n_bufs */
                int rc = 0;
                if (whirl_input) {
                    /* the caller's mono input takes the place of stage T's output plane (stage R
                     * of chunk c - 2, the plane's last reader, precedes on the same stream) */
                    const float *src = d->in + s0 * d->in_stride + (size_t)(base + ib) * TBF_FRAGMENT;
                    CK(cudaMemcpy2DAsync(p.u, sizeof(float) * pitch, src, sizeof(float) * d->in_stride,
                                         sizeof(float) * p.n_fragments * TBF_FRAGMENT, ng,
                                         cudaMemcpyDefault, st[1]));
                } else {
                    if (c >= nb)
                        CK(cudaStreamWaitEvent(st[0], ev[1][par], 0));
                    CK(prof_mark(0, true, st[0]));
                    rc = tbf_launch_stage_t(p, st[0]);
                    CK(prof_mark(0, false, st[0]));
                    CK(cudaEventRecord(ev[0][par], st[0]));
                }
                if (rc == 0) {
                    if (whirl && c >= nb) {
                        CK(cudaStreamWaitEvent(st[1], ev[2][par], 0));
                        CK(cudaStreamWaitEvent(st[1], ev[3][par], 0));
                    }
                    CK(prof_mark(1, true, st[1]));
                    /* the rotors need nothing from stage T: their kernel runs on a stream of its own,
                     * as far ahead of the audio chain as the stage-stream copies allow */
                    const bool rot_apart = whirl && tbf_stage_r_rotors_apart(p);
                    if (rot_apart) {
                        cudaStream_t rs = ctx->profile_stages ? st[1] : ctx->rot_stream[g];
                        if (c >= nb) /* stage W of chunk c - n_bufs has read this copy's angle planes */
                            CK(cudaStreamWaitEvent(rs, ev[2][par], 0));
                        rc = tbf_launch_stage_rot(p, rs);
                        CK(cudaEventRecord(ctx->rot_done[g][par], rs));
                    }
                    if (!whirl_input)
                        CK(cudaStreamWaitEvent(st[1], ev[0][par], 0));
                    if (rc == 0)
                        rc = tbf_launch_stage_r(p, st[1]);
                    CK(prof_mark(1, false, st[1]));
                    CK(cudaEventRecord(ev[1][par], st[1]));
                    if (rot_apart)
                        CK(cudaStreamWaitEvent(st[2], ctx->rot_done[g][par], 0));
                }
                if (rc == 0 && whirl) {
                    CK(cudaStreamWaitEvent(st[2], ev[1][par], 0));
                    if (c >= nb)
                        CK(cudaStreamWaitEvent(st[2], ev[3][par], 0));
                    CK(prof_mark(2, true, st[2]));
                    rc = tbf_launch_stage_w(p, ring, st[2]);
                    CK(prof_mark(2, false, st[2]));
                    CK(cudaEventRecord(ev[2][par], st[2]));
                }
                if (rc == 0 && whirl) {
                    CK(cudaStreamWaitEvent(st[3], ev[2][par], 0));
                    CK(prof_mark(3, true, st[3]));
                    rc = tbf_launch_stage_d(p, st[3]);
                    CK(prof_mark(3, false, st[3]));
                    CK(cudaEventRecord(ev[3][par], st[3]));
                }
                if (rc != 0) {
                    ctx->err = std::string("stage kernel launch: ") + cudaGetErrorString((cudaError_t)rc);
                    return TBF_ERR_CUDA;
                }
                /* stage D is two kernels (D, M); the rotors are one more when they run apart */
                ctx->launches += n_stages + (whirl ? 1 : 0) + (whirl && tbf_stage_r_rotors_apart(p) ? 1 : 0);
            }
            if (host_out) {
                CK(cudaStreamWaitEvent(ctx->copy_stream, ctx->sdone[g][last][par], 0));
                CK(cudaStreamWaitEvent(ctx->copy_stream2, ctx->sdone[g][last][par], 0));
            }
        }
        if (host_out) {
            /* the planes of this chunk cross the link on two copy streams (left / right) */
            const float *src[6] = {out_l, out_r, out_s[0], out_s[1], out_s[2], out_s[3]};
            float *dst[6] = {d->out_l, d->out_r, d->stems[0], d->stems[1], d->stems[2], d->stems[3]};
            const size_t w = sizeof(float) * (size_t)nf * TBF_FRAGMENT;
            for (size_t pl = 0; pl < n_planes; ++pl) {
                if (!src[pl] || !dst[pl])
                    continue;
                cudaStream_t cs = (pl & 1) ? ctx->copy_stream2 : ctx->copy_stream;
                if (chunked) /* chunk k of every instance: one contiguous block */
                    CK(cudaMemcpyAsync(dst[pl] + (size_t)k * n * out_stride, src[pl],
                                       sizeof(float) * n * out_stride, cudaMemcpyDeviceToHost, cs));
                else
                    CK(cudaMemcpy2DAsync(dst[pl] + (size_t)base * TBF_FRAGMENT, sizeof(float) * d->out_stride,
                                         src[pl], w, w, d->n_instances, cudaMemcpyDeviceToHost, cs));
            }
            CK(cudaEventRecord(ctx->copy_done[buf], ctx->copy_stream));
            CK(cudaEventRecord(ctx->copy_done2[buf], ctx->copy_stream2));
        }
    }
    for (int g = 0; g < n_groups; ++g)
        if (cidx[g])
            CK(cudaStreamWaitEvent(ctx->stream, ctx->sdone[g][last][(cidx[g] - 1) % (uint32_t)nb], 0));
    CK(cudaEventRecord(ctx->ev1, ctx->stream));
    if (d->checksums)
        CK(cudaMemcpyAsync(d->checksums, ctx->d_csum, sizeof(uint64_t) * n,
                           cudaMemcpyDeviceToHost, ctx->stream));
    ctx->in_flight = true;
    return TBF_OK;
}

int tbf_gpu_render_batch_launch(tbf_ctx *ctx, const tbf_batch_desc *d)
{
    int rc = tbf_gpu_batch_prepare(ctx, d);
    if (rc != TBF_OK)
        return rc;
    return tbf_gpu_batch_launch(ctx);
}

int tbf_gpu_synchronize(tbf_ctx *ctx)
{
    if (!ctx)
        return TBF_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    ctx->in_flight = false;
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaStreamSynchronize(ctx->copy_stream));
    CK(cudaStreamSynchronize(ctx->copy_stream2));
    /* debugging aid, TBF_TIMELINE=1: when each stage launch of the overlapped schedule became
     * ready (its stream reached it: producers done) and when it finished, ms after the batch
     * started, one line per launch on stderr */
    if (ctx->timeline) {
        for (size_t i = 0; i < ctx->prof_stage.size(); ++i) {
            float t0 = 0.f, t1 = 0.f;
            CK(cudaEventElapsedTime(&t0, ctx->ev0, ctx->prof_events[2 * i]));
            CK(cudaEventElapsedTime(&t1, ctx->ev0, ctx->prof_events[2 * i + 1]));
            fprintf(stderr, "tbf_timeline %c %zu ready %.3f done %.3f\n", "TRWD"[ctx->prof_stage[i]], i,
                    t0, t1);
        }
        ctx->prof_stage.clear();
    }
    for (size_t i = 0; i < ctx->prof_stage.size(); ++i) {
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, ctx->prof_events[2 * i], ctx->prof_events[2 * i + 1]));
        ctx->stage_ms[ctx->prof_stage[i]] += ms;
        ctx->stage_launches[ctx->prof_stage[i]] += 1;
    }
    ctx->prof_stage.clear();
    return TBF_OK;
}

int tbf_gpu_render_batch(tbf_ctx *ctx, const tbf_batch_desc *d)
{
    int rc = tbf_gpu_render_batch_launch(ctx, d);
    if (rc != TBF_OK)
        return rc;
    return tbf_gpu_synchronize(ctx);
}

int tbf_gpu_last_launch_ms(tbf_ctx *ctx, float *ms, int *n_kernel_launches)
{
    if (!ctx || ctx->in_flight)
        return fail(ctx, TBF_ERR_ARG, "last_launch_ms: synchronize first");
    if (ms)
        CK(cudaEventElapsedTime(ms, ctx->ev0, ctx->ev1));
    if (n_kernel_launches)
        *n_kernel_launches = ctx->launches;
    return TBF_OK;
}

/* how the kernels will treat a whirl table set: parallel ordered gather or the serial
 * delay-line path, shared-memory ring size per line, live window in slots */
int tbf_gpu_whirl_tables_info(tbf_ctx *ctx, int handle, int *gather_ok, int *ring_size, int *window)
{
    if (!ctx || handle < 0 || (size_t)handle >= ctx->wh.size())
        return fail(ctx, TBF_ERR_HANDLE, "whirl_tables_info: unknown handle");
    if (gather_ok)
        *gather_ok = ctx->wh[handle].gather_ok;
    if (ring_size)
        *ring_size = ctx->wh[handle].ring_size;
    if (window)
        *window = ctx->wh[handle].window;
    return TBF_OK;
}

/* CUDA-event stopwatch on the library's stream (bench: time K launches on the device) */
int tbf_gpu_timer_start(tbf_ctx *ctx)
{
    if (!ctx)
        return TBF_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    CK(cudaEventRecord(ctx->t0, ctx->stream));
    return TBF_OK;
}

int tbf_gpu_timer_stop(tbf_ctx *ctx, float *ms)
{
    if (!ctx || !ms)
        return TBF_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    CK(cudaEventRecord(ctx->t1, ctx->stream));
    CK(cudaEventSynchronize(ctx->t1));
    CK(cudaEventElapsedTime(ms, ctx->t0, ctx->t1));
    return TBF_OK;
}

/* page-locked host buffers, so host-output batches stream back asynchronously */
void *tbf_gpu_host_alloc(tbf_ctx *ctx, size_t bytes)
{
    void *p = nullptr;
    if (!ctx || cudaSetDevice(ctx->device) != cudaSuccess)
        return nullptr;
    if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return p;
}

void *tbf_gpu_host_alloc2(tbf_ctx *ctx, size_t bytes, unsigned flags)
{
    void *p = nullptr;
    if (!ctx || cudaSetDevice(ctx->device) != cudaSuccess)
        return nullptr;
    const unsigned f = (flags & TBF_HOST_WRITE_COMBINED) ? cudaHostAllocWriteCombined : cudaHostAllocDefault;
    if (cudaHostAlloc(&p, bytes, f | cudaHostAllocPortable) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return p;
}

/* the host link's own ceiling: a plain contiguous device -> page-locked-host copy */
int tbf_gpu_measure_d2h(tbf_ctx *ctx, size_t bytes, int reps, double *gbs)
{
    if (!ctx || !bytes || reps < 1 || !gbs)
        return fail(ctx, TBF_ERR_ARG, "measure_d2h: bad argument");
    if (ctx->in_flight)
        return fail(ctx, TBF_ERR_ARG, "measure_d2h: synchronize first");
    CK(cudaSetDevice(ctx->device));
    void *dev = nullptr, *host = nullptr;
    CK(cudaMalloc(&dev, bytes));
    cudaError_t e = cudaHostAlloc(&host, bytes, cudaHostAllocDefault);
    if (e == cudaSuccess)
        e = cudaMemsetAsync(dev, 0, bytes, ctx->copy_stream);
    float best = 0.f, sum = 0.f;
    for (int r = 0; e == cudaSuccess && r < reps + 1; ++r) { /* first pass untimed */
        e = cudaEventRecord(ctx->t0, ctx->copy_stream);
        if (e == cudaSuccess)
            e = cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, ctx->copy_stream);
        if (e == cudaSuccess)
            e = cudaEventRecord(ctx->t1, ctx->copy_stream);
        if (e == cudaSuccess)
            e = cudaEventSynchronize(ctx->t1);
        float ms = 0.f;
        if (e == cudaSuccess)
            e = cudaEventElapsedTime(&ms, ctx->t0, ctx->t1);
        if (r > 0 && ms > 0.f && (best == 0.f || ms < best))
            best = ms;
        if (r > 0)
            sum += ms;
    }
    cudaFree(dev);
    if (host)
        cudaFreeHost(host);
    if (e != cudaSuccess) {
        ctx->err = std::string("measure_d2h: ") + cudaGetErrorString(e);
        cudaGetLastError();
        return TBF_ERR_CUDA;
    }
    /* the MEAN over the timed passes: with several GPUs copying at once the best pass is the one
     * the neighbours had already finished, which is not the rate a sustained stream gets */
    (void)best;
    *gbs = sum > 0.f ? (double)bytes * reps / (sum * 1e-3) / 1e9 : 0.0;
    return TBF_OK;
}

void tbf_gpu_host_free(tbf_ctx *ctx, void *p)
{
    if (ctx && p)
        cudaFreeHost(p);
}

/* Stage profiling: while on, every stage kernel of a launch runs alone on the library's
 * stream between two CUDA events (no overlap); tbf_gpu_stage_times returns the device time
 * and launch count per stage (T, R, W, D) accumulated since profiling was switched on. */
int tbf_gpu_set_stage_profiling(tbf_ctx *ctx, int on)
{
    if (!ctx || ctx->in_flight)
        return fail(ctx, TBF_ERR_ARG, "set_stage_profiling: synchronize first");
    ctx->profile_stages = on != 0;
    for (int i = 0; i < 4; ++i) {
        ctx->stage_ms[i] = 0.f;
        ctx->stage_launches[i] = 0;
    }
    return TBF_OK;
}

int tbf_gpu_stage_times(tbf_ctx *ctx, float ms[4], int launches[4])
{
    if (!ctx || ctx->in_flight)
        return fail(ctx, TBF_ERR_ARG, "stage_times: synchronize first");
    for (int i = 0; i < 4; ++i) {
        if (ms)
            ms[i] = ctx->stage_ms[i];
        if (launches)
            launches[i] = ctx->stage_launches[i];
    }
    return TBF_OK;
}

int tbf_gpu_set_host_chunk(tbf_ctx *ctx, uint32_t fragments)
{
    if (!ctx || !fragments)
        return TBF_ERR_ARG;
    ctx->host_chunk_fragments = fragments;
    return TBF_OK;
}

uint32_t tbf_gpu_get_host_chunk(const tbf_ctx *ctx) { return ctx ? ctx->host_chunk(true) : 0; }

int tbf_gpu_export_state(tbf_ctx *ctx, uint32_t instance, tbf_state_export *o)
{
    if (!ctx || !o || instance >= ctx->n_inst || !ctx->inst_alive[instance])
        return fail(ctx, TBF_ERR_HANDLE, "export_state: unknown instance");
    if (ctx->in_flight)
        return fail(ctx, TBF_ERR_ARG, "export_state: synchronize first");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream)); /* a create / reset may still be initialising it */
    InstState *s = new InstState();
    cudaError_t e = cudaMemcpy(s, ctx->d_states + instance, sizeof(InstState), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) {
        delete s;
        ctx->err = cudaGetErrorString(e);
        return TBF_ERR_CUDA;
    }
    memset(o, 0, sizeof(*o));
    o->active_osc_l_end = s->tg.activeOscLEnd;
    for (int i = 0; i < s->tg.activeOscLEnd && i <= TBF_NOF_WHEELS; ++i)
        o->active_osc_list[i] = s->activeOscList[i];
    for (int i = 0; i <= TBF_NOF_WHEELS; ++i) {
        o->osc_pos[i] = s->osc[i].pos;
        o->osc_acl_pos[i] = s->osc[i].aclPos;
        o->ref_count[i] = s->osc[i].refCount;
    }
    o->old_routing = s->tg.oldRouting;
    o->new_routing = s->tg.newRouting;
    o->upper_key_count = s->tg.upperKeyCount;
    o->key_down_count = s->tg.keyDownCount;
    o->perc_env_gain = s->rp.percEnvGain;
    o->key_comp_level = s->rp.keyCompLevel;
    o->pz = s->tg.pz;
    o->vib_stator = s->tg.vibStator;
    o->vib_out_pos = s->tg.vibOutPos;
    memcpy(o->vib_buffer, s->vibBuffer, sizeof(o->vib_buffer));
    o->whirl_outpos = s->wl.outpos;
    o->horn_angle = s->rp.hornAngle;
    o->drum_angle = s->rp.drumAngle;
    o->horn_incr = s->rp.hornIncr;
    o->drum_incr = s->rp.drumIncr;
    o->horn_acdc = s->rp.hornAcDc;
    o->drum_acdc = s->rp.drumAcDc;
    memcpy(o->whirl_buf, s->wbuf, sizeof(o->whirl_buf));
    memcpy(o->iir_z[0], s->rp.iirz, sizeof(s->rp.iirz));
    memcpy(o->horn_iir, s->rp.iir, sizeof(s->rp.iir));
    memcpy(o->brake_pos, s->rp.brakePos, sizeof(s->rp.brakePos));
    memcpy(o->ramp_one_minus_l, s->rp.oml, sizeof(s->rp.oml));
    memcpy(o->iir_z[2], s->dr.iirz, sizeof(s->dr.iirz));
    memcpy(o->filt_z, s->wl.z, sizeof(o->filt_z));
    memcpy(o->adx, s->wl.adx, sizeof(o->adx));
    memcpy(o->adi, s->wl.adi, sizeof(o->adi));
    o->error_flags = s->tg.errorFlags | s->rp.errorFlags | s->wl.errorFlags;
    o->stat_fragments = s->tg.statFragments;
    o->stat_active_wheels = s->tg.statActiveWheels;
    delete s;
    return TBF_OK;
}

} /* extern "C" */

/* ------------------------------------------------------------------------------------------
 * tbf_gpu_render_small_noalloc: the render chain for callers that may not allocate (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Everything
 * the stage kernels need between them -- the [slot][sample] stage streams, the per-fragment
 * FragCtl records, the uploaded events with their CSR and cursors, the instance ids and the
 * checksums -- lives in ONE file-scope device array of fixed size instead of the context's
 * grow-on-demand buffers.  Every GPU that loads the library holds these 6.4 MB until exit.
 * All calls (of all contexts on a device) share the array and it keeps the last call's
 * contents: what must start at zero is cleared per call, and calls that overlap on the GPU race
 * -- the entry point therefore runs on the context's stream and returns when it has finished.
 * ---------------------------------------------------------------------------------------- */
#define NA_MAX_INSTANCES TBF_NOALLOC_MAX_INSTANCES
#define NA_MAX_FRAGMENTS TBF_NOALLOC_MAX_FRAGMENTS
#define NA_MAX_EVENTS TBF_NOALLOC_MAX_EVENTS
#define NA_PITCH ((size_t)NA_MAX_FRAGMENTS * TBF_FRAGMENT)
#define NA_UP256(x) (((x) + 255) & ~(size_t)255)
#define NA_OFF_F32 ((size_t)0) /* 8 planes [slot][pitch] of float */
#define NA_OFF_F64 NA_UP256(NA_OFF_F32 + sizeof(float) * 8 * NA_MAX_INSTANCES * NA_PITCH)
#define NA_OFF_CTL NA_UP256(NA_OFF_F64 + sizeof(double) * 2 * NA_MAX_INSTANCES * NA_PITCH)
#define NA_OFF_EVENTS NA_UP256(NA_OFF_CTL + sizeof(FragCtl) * NA_MAX_INSTANCES * NA_MAX_FRAGMENTS)
#define NA_OFF_EVSTART NA_UP256(NA_OFF_EVENTS + sizeof(tbf_event) * NA_MAX_EVENTS)
#define NA_OFF_CURSOR NA_UP256(NA_OFF_EVSTART + sizeof(uint64_t) * (NA_MAX_INSTANCES + 1))
#define NA_OFF_IDS NA_UP256(NA_OFF_CURSOR + sizeof(uint64_t) * 2 * NA_MAX_INSTANCES)
#define NA_OFF_CSUM NA_UP256(NA_OFF_IDS + sizeof(uint32_t) * NA_MAX_INSTANCES)
#define NA_BYTES NA_UP256(NA_OFF_CSUM + sizeof(uint64_t) * NA_MAX_INSTANCES)

static __device__ __align__(256) unsigned char tbf_noalloc_scratch[NA_BYTES];

/* In host code `tbf_noalloc_scratch` is a host-side stand-in, not a device address: the address
 * the kernels get comes from cudaGetSymbolAddress, once per device, with that device current. */
static int noalloc_scratch_address(tbf_ctx *ctx, unsigned char **out)
{
    static unsigned char *addr[64] = {nullptr};
    const int dev = ctx->device;
    unsigned char *p = (dev >= 0 && dev < 64) ? addr[dev] : nullptr;
    if (!p) {
        CK(cudaSetDevice(dev));
        cudaError_t e = cudaGetSymbolAddress((void **)&p, tbf_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            ctx->err = std::string("render_small_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(e);
            cudaGetLastError();
            return TBF_ERR_CUDA;
        }
        if (dev >= 0 && dev < 64)
            addr[dev] = p;
    }
    *out = p;
    return TBF_OK;
}

extern "C" {

unsigned long long tbf_gpu_noalloc_scratch_bytes(void) { return (unsigned long long)NA_BYTES; }

int tbf_gpu_render_small_noalloc(tbf_ctx *ctx, const tbf_batch_desc *d)
{
    if (!ctx || !d)
        return fail(ctx, TBF_ERR_ARG, "render_small_noalloc: null argument");
    if (ctx->in_flight)
        return fail(ctx, TBF_ERR_ARG, "render_small_noalloc: previous launch not synchronized");
    if (int rc = finish_all_pending_synth(ctx))
        return rc;
    if (!d->n_instances || !d->n_fragments)
        return fail(ctx, TBF_ERR_ARG, "render_small_noalloc: empty batch");
    if (d->n_instances > NA_MAX_INSTANCES || d->n_fragments > NA_MAX_FRAGMENTS || d->n_events > NA_MAX_EVENTS) {
        ctx->err = "render_small_noalloc: batch beyond the static scratch (max " +
                   std::to_string(NA_MAX_INSTANCES) + " instances x " + std::to_string(NA_MAX_FRAGMENTS) +
                   " fragments, " + std::to_string(NA_MAX_EVENTS) + " events per call): got " +
                   std::to_string(d->n_instances) + " x " + std::to_string(d->n_fragments) + ", " +
                   std::to_string(d->n_events) + " events; use tbf_gpu_render_batch";
        return TBF_ERR_ARG;
    }
    const uint32_t allowed = TBF_RENDER_WHIRL_BYPASS | TBF_RENDER_WHIRLPROC3 | TBF_RENDER_OUT_DEVICE;
    if ((d->flags & ~allowed) || !(d->flags & TBF_RENDER_OUT_DEVICE))
        return fail(ctx, TBF_ERR_ARG,
                    "render_small_noalloc: flags must be OUT_DEVICE plus WHIRL_BYPASS / WHIRLPROC3 only");
    const uint64_t frames = (uint64_t)d->n_fragments * TBF_FRAGMENT;
    if (!d->out_l || !d->out_r || d->out_stride < frames)
        return fail(ctx, TBF_ERR_ARG, "render_small_noalloc: output buffers missing or stride too small");
    if (d->n_events && !d->events)
        return fail(ctx, TBF_ERR_ARG, "render_small_noalloc: events missing");
    const uint32_t n = d->n_instances;
    uint32_t ids[NA_MAX_INSTANCES];
    int ring = 512;
    for (uint32_t i = 0; i < n; ++i) {
        const uint32_t id = d->instance_ids ? d->instance_ids[i] : i;
        if (id >= ctx->n_inst || !ctx->inst_alive[id])
            return fail(ctx, TBF_ERR_HANDLE, "render_small_noalloc: unknown instance id");
        for (uint32_t j = 0; j < i; ++j)
            if (ids[j] == id)
                return fail(ctx, TBF_ERR_ARG, "render_small_noalloc: instance listed twice");
        ids[i] = id;
        ring = std::max(ring, ctx->wh[ctx->inst_wh[id]].ring_size);
    }
    uint64_t ev_start[NA_MAX_INSTANCES + 1] = {0};
    {
        uint32_t prev_slot = 0, prev_frag = 0;
        for (uint64_t e = 0; e < d->n_events; ++e) {
            const tbf_event &ev = d->events[e];
            if (ev.instance >= n || ev.fragment >= d->n_fragments)
                return fail(ctx, TBF_ERR_ARG, "render_small_noalloc: event slot or fragment out of range");
            if (ev.instance < prev_slot || (ev.instance == prev_slot && ev.fragment < prev_frag))
                return fail(ctx, TBF_ERR_ARG, "render_small_noalloc: events not sorted by (slot, fragment)");
            prev_slot = ev.instance;
            prev_frag = ev.fragment;
            ev_start[ev.instance + 1]++;
        }
        for (uint32_t i = 0; i < n; ++i)
            ev_start[i + 1] += ev_start[i];
    }
    unsigned char *base = nullptr;
    if (int rc = noalloc_scratch_address(ctx, &base))
        return rc;
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    float *f32 = reinterpret_cast<float *>(base + NA_OFF_F32);
    double *f64 = reinterpret_cast<double *>(base + NA_OFF_F64);
    tbf_event *d_events = reinterpret_cast<tbf_event *>(base + NA_OFF_EVENTS);
    uint64_t *d_ev_start = reinterpret_cast<uint64_t *>(base + NA_OFF_EVSTART);
    uint64_t *d_cursor = reinterpret_cast<uint64_t *>(base + NA_OFF_CURSOR);
    uint32_t *d_ids = reinterpret_cast<uint32_t *>(base + NA_OFF_IDS);
    unsigned long long *d_csum = reinterpret_cast<unsigned long long *>(base + NA_OFF_CSUM);

    /* the array keeps the last call's contents: everything read before it is written by this
     * call's own kernels is set here (events, CSR, cursors, ids, checksums).  The stage streams
     * and FragCtl records are fully written by their producer stage before a consumer reads them */
    if (d->n_events)
        CK(cudaMemcpyAsync(d_events, d->events, sizeof(tbf_event) * d->n_events, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(d_ev_start, ev_start, sizeof(uint64_t) * (n + 1), cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(d_cursor, ev_start, sizeof(uint64_t) * n, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(d_cursor + NA_MAX_INSTANCES, ev_start, sizeof(uint64_t) * n, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(d_ids, ids, sizeof(uint32_t) * n, cudaMemcpyHostToDevice, st));
    if (d->checksums)
        CK(cudaMemsetAsync(d_csum, 0, sizeof(uint64_t) * n, st));
    /* the host arrays above are on this function's stack: pageable copies are staged by the
     * runtime before cudaMemcpyAsync returns, and the call synchronizes before it returns */

    const bool whirl = (d->flags & TBF_RENDER_WHIRL_BYPASS) == 0;
    const size_t pitch = (size_t)d->n_fragments * TBF_FRAGMENT;
    RenderParams p;
    memset(&p, 0, sizeof(p));
    p.states = ctx->d_states;
    p.ids = d_ids;
    p.tg = ctx->d_tg;
    p.wh = ctx->d_wh;
    p.events = d_events;
    p.ev_start = d_ev_start;
    p.ev_cursor_t = d_cursor;
    p.ev_cursor_r = d_cursor + NA_MAX_INSTANCES;
    p.out_l = d->out_l;
    p.out_r = d->out_r;
    p.checksums = d->checksums ? d_csum : nullptr;
    p.out_stride = d->out_stride;
    p.n_instances = n;
    p.n_fragments = d->n_fragments;
    p.flags = d->flags;
    p.frag_base = 0;
    p.ctl = reinterpret_cast<FragCtl *>(base + NA_OFF_CTL);
    p.pitch = pitch;
    p.one = 1.0f;
    p.u = f32;
    p.php = f32 + (size_t)n * pitch;
    if (whirl) {
        p.dx0 = f32 + (size_t)2 * n * pitch;
        p.hx0 = f32 + (size_t)3 * n * pitch;
        for (int l = 0; l < 4; ++l)
            p.line[l] = f32 + (size_t)(4 + l) * n * pitch;
        p.hang = f64;
        p.dang = f64 + (size_t)n * pitch;
    }
    CK(cudaEventRecord(ctx->ev0, st));
    int rc = tbf_launch_stage_t(p, st);
    if (rc == 0)
        rc = tbf_launch_stage_r(p, st);
    if (rc == 0 && whirl)
        rc = tbf_launch_stage_w(p, ring, st);
    if (rc == 0 && whirl)
        rc = tbf_launch_stage_d(p, st);
    if (rc != 0) {
        ctx->err = std::string("render_small_noalloc: stage kernel launch: ") + cudaGetErrorString((cudaError_t)rc);
        cudaStreamSynchronize(st);
        return TBF_ERR_CUDA;
    }
    ctx->launches = whirl ? 5 : 2;
    CK(cudaEventRecord(ctx->ev1, st));
    if (d->checksums)
        CK(cudaMemcpyAsync(d->checksums, d_csum, sizeof(uint64_t) * n, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return TBF_OK;
}

unsigned tbf_sizeof_tonegen_tables(void) { return (unsigned)sizeof(tbf_tonegen_tables); }
unsigned tbf_sizeof_whirl_tables(void) { return (unsigned)sizeof(tbf_whirl_tables); }
unsigned tbf_sizeof_state_export(void) { return (unsigned)sizeof(tbf_state_export); }
unsigned tbf_sizeof_batch_desc(void) { return (unsigned)sizeof(tbf_batch_desc); }
unsigned tbf_sizeof_event(void) { return (unsigned)sizeof(tbf_event); }

} /* extern "C" */
