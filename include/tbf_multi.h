/* tbf_multi.h -- the batched renderer over several GPUs of one box from ONE host process.
 *
 * the path shards by organ with no exchange step -- "one host thread + stream per
 * GPU, instances partitioned statically, sorted by table hash, outputs in per-GPU buffers, no
 * collective".  This is that host code in C++ behind a C ABI: a tbf_multi owns one tbf_ctx
 * (include/tbf_gpu.h) and one worker thread per device; a batch lists its slots with the device
 * each organ lives on, and every device renders its slots concurrently into the caller's
 * buffers.  Nothing crosses between GPUs (NVLink is not used: there is nothing to exchange).
 *
 * The reference has no counterpart (it renders one organ per process callback,
 * src/clap.cpp:244-270); the organs a host would spread over its cores are spread over GPUs.
 */
#ifndef TBF_MULTI_H
#define TBF_MULTI_H

#include "tbf_gpu.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct tbf_multi tbf_multi;

/* one context and one worker thread per entry of devices[] (NULL: devices 0 .. n_devices-1).
 * A device may be listed more than once (two contexts on one GPU: tests on a one-GPU box). */
int tbf_multi_create(tbf_multi **out, int n_devices, const int *devices);
void tbf_multi_destroy(tbf_multi *m);
int tbf_multi_device_count(const tbf_multi *m);
/* the context of entry `index`: table uploads, instance creation, timers, export_state ...
 * Different entries' contexts may be used from different host threads at the same time. */
tbf_ctx *tbf_multi_context(tbf_multi *m, int index);
const char *tbf_multi_last_error(const tbf_multi *m);

/* Static block partition of n organs over n_devices, organs SORTED BY TABLE KEY first
 * (SURVEY 8(e)): organs that share wave tables become neighbours and land on one GPU, so each
 * GPU uploads and caches few table sets.  table_key[i] = any hash of organ i's table set(s).
 * order[k] = the organ at position k of the sorted list (ties keep the caller's order);
 * device_of[i] = the entry index organ i is placed on: position k goes to device
 * k / ceil(n / n_devices).  Pure function: no tbf_multi needed. */
int tbf_multi_plan(int n_devices, uint32_t n, const uint64_t *table_key, uint32_t *device_of,
                   uint32_t *order);

/* tbf_batch_desc over several devices.  Slot i is instance instance_ids[i] of entry
 * device_of[i]; events are sorted by (slot, fragment) with `instance` = the slot.  Host
 * outputs: out_l / out_r / stems rows [slot][out_stride] (checksums [slot]).  With
 * TBF_RENDER_OUT_DEVICE the audio stays on the devices: dev_out_l[d] / dev_out_r[d] are device
 * pointers of entry d, rows in the order entry d's slots appear in the batch. */
typedef struct tbf_multi_batch {
    uint32_t n_slots;
    const uint32_t *device_of;
    const uint32_t *instance_ids;
    uint32_t n_fragments;
    const tbf_event *events;
    uint64_t n_events;
    float *out_l, *out_r;
    uint64_t out_stride;
    uint32_t flags; /* TBF_RENDER_*; OUT_CHUNKED is not supported here */
    uint64_t *checksums;
    const float *in; /* TBF_RENDER_WHIRL_INPUT, host, [slot][in_stride] */
    uint64_t in_stride;
    float *stems[4];
    float *const *dev_out_l, *const *dev_out_r; /* [n_devices], TBF_RENDER_OUT_DEVICE only */
} tbf_multi_batch;

/* Render the batch: every device's worker thread runs tbf_gpu_render_batch on its slots (one
 * call per run of consecutive slots on that device: order the slots by device for one call
 * each); returns when all are done, with the first error if any (tbf_multi_last_error).
 * device_ms (optional, [n_devices]) receives each device's summed tbf_gpu_last_launch_ms. */
int tbf_multi_render_batch(tbf_multi *m, const tbf_multi_batch *batch, float *device_ms);

#ifdef __cplusplus
}
#endif
#endif /* TBF_MULTI_H */
