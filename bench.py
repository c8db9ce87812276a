#!/usr/bin/env python
"""bench.py -- realtime-instances benchmark of the tonewheel -> scanner -> Leslie path.

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU engine

A "step" = one pass of the hot path over one batch: `instances` organs x `seconds`
of 48 kHz stereo audio (default: BASELINE config 3 = 4096 instances, tonegen +
vibrato C3 + Leslie fast through whirlProc3, 10 s each).  Multi-GPU: one process per
GPU (torchrun), every rank renders its own `instances` organs (weak scaling, no
collective on the data path); the timed region is bracketed by barriers and the
slowest rank's device time is used.

metric  = organ-seconds of 48 kHz stereo audio rendered per wall-second
value   = device-resident: events uploaded before the timed region, audio kept in HBM
e2e     = through tbf_gpu_render_batch with HOST buffers: events copied host->device and
          all audio copied device->host inside the timed region, every step
"""
import argparse
import json
import os
import sys
import threading
import time

# 2 instance groups x 4 stage streams + copy + control stream: one hardware queue each
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
# stdout carries exactly one JSON line: NCCL's version banner and warnings go to stderr
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np

METRIC = "realtime-instances: 48 kHz stereo organ-seconds rendered per wall-second"
UNIT = "organ-seconds/s"
N_TABLE_SETS = 16  # engine seeds 1 + id mod 16: 16 dither/click realisations
DEFAULT_INSTANCES = {2: 1024, 3: 4096, 4: 0, 5: 16384}  # 4: one organ per accepted scale
WORKLOAD_NAME = {
    2: "config2: tonegen only, drawbars 888888888, 10-note chords re-struck every 0.5 s, "
       "percussion + key click, whirl bypassed",
    3: "config3: tonegen + vibrato C3 + Leslie fast (whirlProc3), drawbars 888888888, 10-note "
       "chords re-struck every 0.5 s, percussion + key click",
    4: "config4: one organ per Scala scale of the reference's archive (own wave tables each, built "
       "by tbf_build_tonegen_tables), fixed 5-note chord re-struck every 0.5 s through config 3's "
       "chain (vibrato C3 + Leslie fast, whirlProc3)",
    5: "config5: config 3's chain and chords over the Cartesian mix of the reference's 10 osc "
       "cfg variants x 3 whirl cfg variants x 16 engine seeds (80 wave-table sets), organs "
       "grouped by table set",
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", type=int, default=3, choices=[2, 3, 4, 5])
    ap.add_argument("--instances", type=int, default=0, help="instances per GPU")
    ap.add_argument("--seconds", type=float, default=0.0,
                    help="audio seconds per instance per step (default 10; config 4: 2)")
    ap.add_argument("--cpu-sample-seconds", type=float, default=15.0,
                    help="target CPU time of the bounded cpu_baseline sample")
    ap.add_argument("--ref-build", default="strict", choices=["strict", "stock"],
                    help="--impl reference: the engine built strict IEEE (the parity oracle) or with "
                         "the project's stock -O3 -ffast-math flags (timing only)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    a = ap.parse_args()
    if a.seconds <= 0:
        a.seconds = 2.0 if a.config == 4 else 10.0
    return a


# --------------------------------------------------------------------------
# CPU arm: the reference's own engine (oracle/_ref), one instance per host core
# --------------------------------------------------------------------------
def _cpu_worker(args):
    """Runs in a forked worker: init its instances, wait for the common start time, render."""
    ids, n_frag, config, start_at = args
    from oracle import ref_engine
    from tunebfree_b200 import workloads
    # a few engine instances per worker, cycled: the render cost does not depend on which
    # organ plays which script, and 3 MB per reference instance adds up on many-core hosts
    engines = []
    for i in ids[:4]:
        if config == 4:  # one tuning per organ: organ i plays scale i of the sweep
            from tunebfree_b200 import tables
            degs = workloads.scala_sweep()[1]
            engines.append(ref_engine.RefInstance(seed=1, freqs=tables.freqs_from_ratios(degs[i % len(degs)])))
            continue
        osc, whirl, seed = workloads.organ_variant(i, config)
        engines.append(ref_engine.RefInstance(seed=seed, cfg=workloads.variant_cfg(osc, whirl)))
    insts = []
    for j, i in enumerate(ids):
        ev = workloads.instance_events(i, n_frag, config)
        insts.append((engines[j % len(engines)], ref_engine.events_array(ev)))
    flags = workloads.render_flags(config)
    outl = np.zeros(n_frag * 128, dtype=np.float32)
    outr = np.zeros(n_frag * 128, dtype=np.float32)
    L = ref_engine.lib()
    import ctypes as C
    while time.time() < start_at:
        time.sleep(0.001)
    t0 = time.time()
    for r, (arr, n) in insts:
        L.ref_render(r._h, n_frag, flags, arr, n, outl.ctypes.data_as(C.c_void_p),
                     outr.ctypes.data_as(C.c_void_p))
    t1 = time.time()
    return t0, t1, float(np.abs(outl).max())


def cpu_reference_pass(config, n_frag, per_core, cores):
    """One bounded sample: cores x per_core instances of n_frag fragments on `cores` worker
    processes (the engine is single-threaded and rand() is process-global).  Returns
    organ-seconds/s over the common render window."""
    import multiprocessing as mp
    ctx = mp.get_context("fork")
    start_at = time.time() + 3.0 + 0.01 * per_core  # every worker finishes its inits first
    jobs = [(list(range(c * per_core, (c + 1) * per_core)), n_frag, config, start_at)
            for c in range(cores)]
    with ctx.Pool(cores) as pool:
        res = pool.map(_cpu_worker, jobs)
    t0 = min(r[0] for r in res)
    t1 = max(r[1] for r in res)
    organ_seconds = cores * per_core * n_frag / 375.0
    return organ_seconds / (t1 - t0), t1 - t0


def cpu_sample_plan(config, seconds, target_cpu_seconds, cores):
    """Instances per core so that one sample is about target_cpu_seconds of wall time
    (the reference runs ~45x realtime per core with the Leslie, ~400x without)."""
    per_core_rate = 45.0 if config >= 3 else 350.0
    per_core = max(1, int(round(target_cpu_seconds * per_core_rate / seconds)))
    return per_core


def run_reference_arm(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # the CPU engine has no use for more ranks: rank 0 alone measures
    if a.ref_build == "stock":
        os.environ["TBF_REF_BUILD"] = "stock"
    from oracle import ref_engine
    ref_engine.lib()
    cores = os.cpu_count() or 1
    n_frag = int(round(a.seconds * 375))
    per_core = cpu_sample_plan(a.config, a.seconds, a.cpu_sample_seconds, cores)
    vals = []
    for _ in range(a.warmup and 1):  # one short warm-up pass: page in the library, spin clocks up
        cpu_reference_pass(a.config, max(8, n_frag // 20), 1, cores)
    t_all = 0.0
    for _ in range(a.steps):
        v, dt = cpu_reference_pass(a.config, n_frag, per_core, cores)
        vals.append(v)
        t_all += dt
    value = float(np.mean(vals))
    sample = "%d instances x %.1f s (%d per core) per step, oracle/_ref = reference engine, %s" % (
        cores * per_core, a.seconds, per_core,
        "stock flags -O3 -ffast-math (not bit-exact with the oracle)" if a.ref_build == "stock"
        else "strict IEEE -O2")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": 1e3 * t_all / max(1, a.steps),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": WORKLOAD_NAME[a.config], "instances_per_step": cores * per_core,
                   "seconds_per_instance": a.seconds, "sample_rate": 48000, "host_cores": cores},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference",
                         "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------
# clocks during the timed region
# --------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    def __init__(self, device_index):
        super().__init__(daemon=True)
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_ev = threading.Event()
        self._h = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self._h = None

    def run(self):
        if self._h is None:
            return
        nv = self._nv
        names = {
            nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
            nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
        }
        while not self._stop_ev.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop_ev.wait(0.1)

    def stop(self):
        self._stop_ev.set()
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def physical_device_index(local_rank):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_rank])
        except Exception:
            return local_rank
    return local_rank


# --------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------
def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return None


def run_gpu_arm(a):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    # CPU baseline first (forks worker processes: must happen before any CUDA context)
    cpu_baseline = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        cores = os.cpu_count() or 1
        n_frag_cpu = int(round(a.seconds * 375))
        per_core = cpu_sample_plan(a.config, a.seconds, a.cpu_sample_seconds, cores)
        cpu_reference_pass(a.config, max(8, n_frag_cpu // 20), 1, cores)
        v, _ = cpu_reference_pass(a.config, n_frag_cpu, per_core, cores)
        cpu_baseline = {"value": v, "unit": UNIT, "cores": cores, "kind": "reference",
                        "sample": "%d instances x %.1f s (%d per core), oracle/_ref = the reference "
                                  "engine compiled strict IEEE -O2, one process per core"
                                  % (cores * per_core, a.seconds, per_core)}

    import torch
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        # NCCL prints its banner on file descriptor 1 when the communicator comes up
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            os.dup2(saved, 1)
            os.close(saved)
    from tunebfree_b200 import cabi, engine, tables, workloads

    def barrier():
        if dist is not None:
            dist.barrier()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    n_inst = a.instances or DEFAULT_INSTANCES[a.config]
    if a.config == 4:
        sweep_names, sweep_degs, _ = workloads.scala_sweep()
        n_inst = n_inst or len(sweep_names)
    n_frag = int(round(a.seconds * 375))
    frames = n_frag * 128
    flags = workloads.render_flags(a.config)
    from tunebfree_b200 import sharding
    # every rank plays different organs; slot i plays global organ first_global + i with
    # table set (global id) mod 16
    variant = lambda g: workloads.organ_variant(g, a.config)
    if a.config == 4:
        # rank r plays scales [r * n_inst, (r + 1) * n_inst) of the sweep (wrapping around)
        global_ids = sharding.shard_plan_sweep(rank, world, n_inst, len(sweep_names))
        groups = {}
    elif a.config == 5:
        global_ids, groups = sharding.shard_plan_grouped(rank, world, n_inst, variant)
    else:
        first_global, per_set = sharding.shard_plan(rank, world, n_inst, N_TABLE_SETS)
        global_ids = first_global + np.arange(n_inst, dtype=np.int64)
        groups = {variant(first_global + slots[0]): slots for slots in per_set if slots}

    ctx = engine.Context(local_rank)
    if os.environ.get("TBF_BENCH_HOST_CHUNK"):  # fragments per device->host staging chunk (sweeps)
        ctx.set_host_chunk(int(os.environ["TBF_BENCH_HOST_CHUNK"]))
    # table sets built by the product's own host-side builder: wave tables per (osc variant,
    # engine seed), Leslie tables per whirl variant
    tg_handles, wh_handles = {}, {}
    for osc, whirl, seed in groups:
        if (osc, seed) not in tg_handles or whirl not in wh_handles:
            spec = tables.InstanceSpec(seed=seed, cfg=workloads.variant_cfg(osc, whirl))
            tt, wt = spec.build()
            if (osc, seed) not in tg_handles:
                tg_handles[(osc, seed)] = ctx.tonegen_tables(tt)
            if whirl not in wh_handles:
                wh_handles[whirl] = ctx.whirl_tables(wt)
            tables.free_tables(tt, wt)
    init = None
    if a.config == 4:
        # config 4's init leg: 128 frequencies -> wave tables, key wiring, envelopes (the
        # product's restatement of initToneGenerator) on the host cores, then the upload
        import ctypes as C
        from concurrent.futures import ThreadPoolExecutor
        threads = os.cpu_count() or 1
        spec0 = tables.InstanceSpec(seed=1)
        tt0, wt0 = spec0.build()
        wh_handles["default"] = ctx.whirl_tables(wt0)
        tables.free_tables(tt0, wt0)

        def build_one(g):
            sp = tables.InstanceSpec(freqs=tables.freqs_from_ratios(sweep_degs[g]), seed=1)
            return sp.build_tonegen()

        ids = np.zeros(n_inst, dtype=np.uint32)
        t_wait = t_upload = 0.0
        t_init0 = time.perf_counter()
        chunks = [[int(g) for g in global_ids[lo:lo + 256]] for lo in range(0, n_inst, 256)]
        with ThreadPoolExecutor(threads) as pool:
            # the host cores build chunk c + 1 while this thread uploads chunk c
            futures = [pool.submit(build_one, g) for g in chunks[0]]
            lo = 0
            for c in range(len(chunks)):
                t0 = time.perf_counter()
                built = [f.result() for f in futures]
                t1 = time.perf_counter()
                futures = [pool.submit(build_one, g) for g in chunks[c + 1]] if c + 1 < len(chunks) else []
                for k, tt in enumerate(built):
                    h = ctx.tonegen_tables(tt)
                    tg_handles[lo + k] = h
                    ids[lo + k] = ctx.create_instances(h, wh_handles["default"], 1)
                    tables.free_tables(tt, None)
                lo += len(built)
                t_wait += t1 - t0
                t_upload += time.perf_counter() - t1
        ctx.synchronize()
        t_init = time.perf_counter() - t_init0
        init = {"organs_initialised_per_s": n_inst / t_init, "wall_s": t_init, "host_threads": threads,
                "upload_s": t_upload, "waiting_for_builders_s": t_wait,
                "what": "128 frequencies -> tone-generator tables on the host cores "
                        "(tbf_build_tonegen_tables), upload, instance create; builders and uploads overlap"}
    else:
        ids = sharding.slot_instance_ids(
            groups, lambda k, count: ctx.create_instances(tg_handles[(k[0], k[2])], wh_handles[k[1]], count))
    ev_np = workloads.batch_events_numpy(0, n_inst, n_frag, a.config, global_ids=global_ids)
    events = engine.events_from_numpy(ev_np)
    h2d_bytes = int(ev_np.shape[0] * 20 + n_inst * 4 + (n_inst + 1) * 8)

    # ---- value: device resident ----
    out = torch.empty((2, n_inst, frames), dtype=torch.float32, device="cuda:%d" % local_rank)
    ctx.prepare(n_inst, n_frag, events, flags | cabi.RENDER_OUT_DEVICE, out[0].data_ptr(),
                out[1].data_ptr(), frames, ids=ids)
    for _ in range(a.warmup):
        ctx.launch()
        ctx.synchronize()
    sampler = ClockSampler(physical_device_index(local_rank))
    torch.cuda.synchronize()
    barrier()
    sampler.start()
    kernel_ms = 0.0
    launches = 0
    ctx.timer_start()
    for _ in range(a.steps):
        ctx.launch()
        ctx.synchronize()
        ms, nl = ctx.last_launch_ms()
        kernel_ms += ms
        launches += nl
    total_ms = ctx.timer_stop()
    torch.cuda.synchronize()
    barrier()
    clocks = sampler.stop()
    total_ms = max_over_ranks(total_ms)
    peak_out = float(out[0, 0].abs().max().item())
    st = ctx.export_state(int(ids[0]))
    n_act = st.stat_active_wheels / max(1, st.stat_fragments)
    # per-stage device times: one extra pass with every stage kernel alone between two CUDA
    # events on its launch stream (outside the timed region: it removes the overlap)
    ctx.set_stage_profiling(True)
    ctx.launch()
    ctx.synchronize()
    stage = ctx.stage_times()
    ctx.set_stage_profiling(False)
    del out
    torch.cuda.empty_cache()

    organ_seconds_step = n_inst * a.seconds
    value = world * organ_seconds_step * a.steps / (total_ms * 1e-3)

    # ---- e2e: host buffers through tbf_gpu_render_batch ----
    e2e = None
    if not a.no_e2e:
        # Host buffers for the whole step are 2 x n_inst x frames floats per rank (15.7 GB in
        # config 3).  If the box cannot hold that for all ranks, the step is rendered in time
        # slices into a smaller buffer: same calls, same bytes over the link, state carried on.
        need = 2 * n_inst * frames * 4
        slices = max(1, int(os.environ.get("TBF_BENCH_E2E_SLICES", "1")))
        try:
            avail = [int(l.split()[1]) * 1024 for l in open("/proc/meminfo") if l.startswith("MemAvailable")][0]
            while slices < 16 and need / slices * world > 0.5 * avail:
                slices *= 2
        except (OSError, IndexError, ValueError):
            pass
        bounds = [int(round(k * n_frag / slices)) for k in range(slices + 1)]
        slice_frames = max(bounds[k + 1] - bounds[k] for k in range(slices)) * 128
        try:
            hl, pl = ctx.host_alloc((n_inst, slice_frames))
            hr, pr = ctx.host_alloc((n_inst, slice_frames))
            pinned = True
        except MemoryError:
            hl = np.empty((n_inst, slice_frames), dtype=np.float32)
            hr = np.empty((n_inst, slice_frames), dtype=np.float32)
            pl = pr = None
            pinned = False
        if slices == 1:
            parts = [(n_frag, events)]
        else:
            parts = []
            for k in range(slices):
                lo, hi = bounds[k], bounds[k + 1]
                sub = ev_np[(ev_np[:, 1] >= lo) & (ev_np[:, 1] < hi)].copy()
                sub[:, 1] -= lo
                parts.append((hi - lo, engine.events_from_numpy(sub)))

        e2e_count = [0]

        def e2e_step():
            for nf_k, ev_k in parts:
                ctx.prepare(n_inst, nf_k, ev_k, flags, hl.ctypes.data, hr.ctypes.data, slice_frames,
                            ids=ids)
                ctx.launch()
                ctx.synchronize()
                e2e_count[0] += ctx.last_launch_ms()[1]

        for _ in range(max(1, min(a.warmup, 2))):
            e2e_step()
        barrier()
        e2e_count[0] = 0
        t0 = time.perf_counter()
        for _ in range(a.steps):
            e2e_step()
        dt = time.perf_counter() - t0
        e2e_launches = e2e_count[0]
        barrier()
        dt = max_over_ranks(dt)
        e2e = {"value": world * organ_seconds_step * a.steps / dt, "unit": UNIT,
               "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": int(2 * n_inst * frames * 4),
               "host_buffers": "page-locked" if pinned else "pageable", "time_slices": slices,
               "peak_abs_sample": float(np.abs(hl[0]).max())}
        if pl:
            ctx.host_free(pl)
            ctx.host_free(pr)

    if rank == 0:
        peaks = measured_peaks()
        hbm_peak = peaks["hbm_gbs"] if peaks else 6650.0
        whirl = 1 if a.config >= 3 else 0
        f_clk = (clocks.get("sm_mhz") or (peaks or {}).get("sm_max_mhz") or 1965.0) * 1e6
        rate = value / world  # organ-seconds per second per GPU
        # dominant kernel = the stage with the largest summed device time in the profiled pass
        names = {"T": "tbf_stage_t", "R": "tbf_stage_r", "W": "tbf_stage_w", "D": "tbf_stage_d"}
        dom = max(stage, key=lambda k: stage[k][0])
        dom_ms, dom_n = stage[dom]
        dur = dom_ms * 1e-3 / max(1, dom_n)                      # average launch duration
        secs_per_dom_launch = organ_seconds_step / max(1, dom_n)  # organ-seconds one launch renders
        # algorithmic HBM-side bytes per organ-second of each stage (DESIGN.md "Work model"):
        # T reads 4 B per active wheel and sample, writes u + php; R reads them, writes dx0, hx0
        # and two double angles; W reads those, writes four line reads; D reads them + hx0, writes L, R
        stage_bytes = {"T": 48000 * (4 * n_act + 8), "R": 48000 * (8 + 24), "W": 48000 * (24 + 16),
                       "D": 48000 * (20 + 8)}
        ach_hbm = stage_bytes[dom] * secs_per_dom_launch / dur / 1e9
        # algorithmic work of the whole path per organ-second
        w32 = 48000 * (6 * n_act + 17 + 253 * whirl)
        bsm = 48000 * (4 * n_act + 70 + 380 * whirl)
        # DRAM bytes of one launch of that kernel: dram__bytes_read.sum + dram__bytes_write.sum of
        # the committed `ncu --set full` capture, scaled by organ-fragments per launch
        traffic = None
        try:
            cap = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))[names[dom]]
            if cap["config"] == a.config:
                frag_now = organ_seconds_step * 375.0 / max(1, dom_n)
                traffic = cap["dram_bytes"] * frag_now / (cap["instances"] * cap["fragments"])
        except (OSError, KeyError, ValueError):
            pass
        roofline = {"bound": "hbm", "achieved": ach_hbm, "peak": hbm_peak, "unit": "GB/s",
                    "frac": ach_hbm / hbm_peak, "traffic": traffic,
                    "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s",
                    "note": "bytes = the dominant stage kernel's algorithmic reads + writes; the path is "
                            "bound by instruction issue, not HBM: see fp32_issue / smem_bw (whole path, "
                            "against the timed region) and profiles/",
                    "kernel": names[dom], "avg_launch_ms": dur * 1e3, "launches_profiled": dom_n,
                    "n_act": n_act,
                    "stage_ms_per_step": {names[k]: stage[k][0] for k in "TRWD"},
                    "fp32_issue": {"achieved": w32 * rate / 1e12,
                                   "peak": 148 * 128 * f_clk / 1e12, "unit": "T lane-ops/s",
                                   "frac": (w32 * rate) / (148 * 128 * f_clk)},
                    "smem_bw": {"achieved": bsm * rate / 1e12,
                                "peak": 148 * 128 * f_clk / 1e12, "unit": "TB/s",
                                "frac": (bsm * rate) / (148 * 128 * f_clk)}}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": total_ms / a.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD_NAME[a.config], "instances_per_gpu": n_inst,
                       "seconds_per_instance": a.seconds, "sample_rate": 48000,
                       "table_sets": len(tg_handles), "whirl_table_sets": len(wh_handles),
                       "tuning": "one Scala scale per organ" if a.config == 4 else "12-TET",
                       "init": init,
                       "l2": "inputs larger than L2 (instance state %.0f MB + %.1f GB of output per step)"
                             % (n_inst * 0.08, 2 * n_inst * frames * 4 / 1e9),
                       "parallelism": "instances sharded over %d GPU(s), no collective" % world},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches + (e2e_launches if e2e else 0),
            "roofline": roofline, "cpu_baseline": cpu_baseline,
            "checks": {"peak_abs_sample": peak_out, "error_flags": int(st.error_flags)},
        }
        print(json.dumps(line), flush=True)
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()


def main():
    a = parse_args()
    if a.impl == "reference":
        run_reference_arm(a)
    else:
        run_gpu_arm(a)


if __name__ == "__main__":
    main()
