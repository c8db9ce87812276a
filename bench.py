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
    ap.add_argument("--no-parity-check", action="store_true",
                    help="skip the sampled bit-for-bit comparison of the timed audio with the oracle")
    a = ap.parse_args()
    if a.seconds <= 0:
        a.seconds = 2.0 if a.config == 4 else 10.0
    return a


# --------------------------------------------------------------------------
# CPU arm: the reference's own engine (oracle/_ref), one instance per host core
# --------------------------------------------------------------------------
_CPU_BARRIER = None  # multiprocessing.Barrier shared with the forked workers


def _ref_instance(g, config):
    """A fresh reference organ for global organ id g of a config: every script gets its own
    engine, so no key of another script is still held when it starts (the GPU arm gives every
    organ its own instance too)."""
    from oracle import ref_engine
    from tunebfree_b200 import workloads
    if config == 4:  # one tuning per organ: organ g plays scale g of the sweep
        from tunebfree_b200 import tables
        degs = workloads.scala_sweep()[1]
        return ref_engine.RefInstance(seed=1, freqs=tables.freqs_from_ratios(degs[g % len(degs)]))
    osc, whirl, seed = workloads.organ_variant(g, config)
    return ref_engine.RefInstance(seed=seed, cfg=workloads.variant_cfg(osc, whirl))


def _cpu_worker(args):
    """Runs in a forked worker: init its instances, meet the other workers, render."""
    ids, n_frag, config = args
    from oracle import ref_engine
    from tunebfree_b200 import workloads
    insts = [(_ref_instance(i, config), ref_engine.events_array(workloads.instance_events(i, n_frag, config)))
             for i in ids]
    flags = workloads.render_flags(config)
    outl = np.zeros(n_frag * 128, dtype=np.float32)
    outr = np.zeros(n_frag * 128, dtype=np.float32)
    L = ref_engine.lib()
    import ctypes as C
    _CPU_BARRIER.wait()  # every worker has finished its inits
    t0 = time.time()
    for r, (arr, n) in insts:
        L.ref_render(r._h, n_frag, flags, arr, n, outl.ctypes.data_as(C.c_void_p),
                     outr.ctypes.data_as(C.c_void_p))
    t1 = time.time()
    frags = wheels = 0
    for r, _ in insts:
        st = r.export_state()
        frags += st.stat_fragments
        wheels += st.stat_active_wheels
        r.close()
    return t0, t1, frags, wheels


def cpu_reference_pass(config, n_frag, per_core, cores):
    """One bounded sample: cores x per_core organs of n_frag fragments on `cores` worker
    processes (the engine is single-threaded and rand() is process-global), one fresh engine
    per organ.  Returns (organ-seconds/s over the common render window, seconds, time-averaged
    active wheels of the sample)."""
    global _CPU_BARRIER
    import multiprocessing as mp
    ctx = mp.get_context("fork")
    _CPU_BARRIER = ctx.Barrier(cores)
    jobs = [(list(range(c * per_core, (c + 1) * per_core)), n_frag, config) for c in range(cores)]
    with ctx.Pool(cores) as pool:
        res = pool.map(_cpu_worker, jobs, chunksize=1)
    t0 = min(r[0] for r in res)
    t1 = max(r[1] for r in res)
    organ_seconds = cores * per_core * n_frag / 375.0
    n_act = sum(r[3] for r in res) / max(1, sum(r[2] for r in res))
    return organ_seconds / (t1 - t0), t1 - t0, n_act


def cpu_sample_plan(config, seconds, target_cpu_seconds, cores):
    """Instances per core so that one sample is about target_cpu_seconds of wall time
    (the reference runs ~45x realtime per core with the Leslie, ~400x without)."""
    per_core_rate = 45.0 if config >= 3 else 350.0
    per_core = max(1, int(round(target_cpu_seconds * per_core_rate / seconds)))
    return per_core


def _parity_worker(args):
    """Forked worker: the oracle's render of ONE organ of the timed batch, replayed over every
    pass bench.py makes (the organs keep their state from pass to pass); returns the audio of the
    passes listed in `keep` for the bit-for-bit comparison with what the GPU arm timed."""
    g, n_frag, config, n_pass, keep, slices = args
    from tunebfree_b200 import workloads
    r = _ref_instance(g, config)
    ev = sorted(workloads.instance_events(g, n_frag, config), key=lambda e: e[0])
    flags = workloads.render_flags(config)
    out = {}
    for k in range(n_pass):
        l, rr = r.render(n_frag, flags, ev)
        if k in keep:
            out[k] = (l, rr)
    r.close()
    return g, out


def parity_reference(global_ids_sampled, n_frag, config, n_pass, keep):
    """{organ: {pass: (L, R)}} from the oracle, computed on forked workers BEFORE the CUDA context
    exists (checker only: nothing of this is timed or shipped)."""
    import multiprocessing as mp
    ctx = mp.get_context("fork")
    jobs = [(int(g), n_frag, config, n_pass, tuple(keep), 1) for g in global_ids_sampled]
    with ctx.Pool(min(len(jobs), os.cpu_count() or 1)) as pool:
        return dict(pool.map(_parity_worker, jobs, chunksize=1))


def bench_config(a, n_inst, world, table_sets=None, whirl_sets=None):
    """The workload description both arms print (the reference arm runs a bounded sample of it)."""
    frames = int(round(a.seconds * 375)) * 128
    return {"workload": WORKLOAD_NAME[a.config], "instances_per_gpu": n_inst,
            "seconds_per_instance": a.seconds, "sample_rate": 48000,
            "tuning": "one Scala scale per organ" if a.config == 4 else "12-TET",
            "l2": "inputs larger than L2 (instance state %.0f MB + %.1f GB of output per step)"
                  % (n_inst * 0.08, 2 * n_inst * frames * 4 / 1e9),
            "parallelism": "instances sharded over %d GPU(s), no collective" % world}


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
    vals, acts = [], []
    for _ in range(a.warmup and 1):  # one short warm-up pass: page in the library, spin clocks up
        cpu_reference_pass(a.config, max(8, n_frag // 20), 1, cores)
    t_all = 0.0
    for _ in range(a.steps):
        v, dt, n_act = cpu_reference_pass(a.config, n_frag, per_core, cores)
        vals.append(v)
        acts.append(n_act)
        t_all += dt
    value = float(np.mean(vals))
    sample = "%d instances x %.1f s (%d per core, a fresh engine each) per step, oracle/_ref = reference engine, %s" % (
        cores * per_core, a.seconds, per_core,
        "stock flags -O3 -ffast-math (not bit-exact with the oracle)" if a.ref_build == "stock"
        else "strict IEEE -O2")
    n_inst = a.instances or DEFAULT_INSTANCES[a.config] or 5176
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": 1e3 * t_all / max(1, a.steps),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": bench_config(a, n_inst, int(os.environ.get("WORLD_SIZE", "1"))),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference",
                         "sample": sample, "instances_per_step": cores * per_core,
                         "n_act": float(np.mean(acts))},
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
    # multi-rank runs: rank 0 measures the CPU baseline alone on the host cores; the other ranks
    # wait at a file gate (no process group yet: the workers are forked before any CUDA / NCCL
    # state exists) and only then start their own oracle workers for the sampled parity check
    gate = os.path.join("/tmp", "tbf_bench_gate_%d_%s" % (os.getppid(), os.environ.get("MASTER_PORT", "0")))
    if world > 1 and rank != 0 and not a.no_cpu_baseline:
        t_gate = time.time()
        while not os.path.exists(gate) and time.time() - t_gate < 900:
            time.sleep(0.2)
    if rank == 0 and not a.no_cpu_baseline:
        cores = os.cpu_count() or 1
        n_frag_cpu = int(round(a.seconds * 375))
        per_core = cpu_sample_plan(a.config, a.seconds, a.cpu_sample_seconds, cores)
        cpu_reference_pass(a.config, max(8, n_frag_cpu // 20), 1, cores)
        v, _, cpu_n_act = cpu_reference_pass(a.config, n_frag_cpu, per_core, cores)
        cpu_baseline = {"value": v, "unit": UNIT, "cores": cores, "kind": "reference",
                        "sample": "%d instances x %.1f s (%d per core, a fresh engine each), oracle/_ref = "
                                  "the reference engine compiled strict IEEE -O2, one process per core"
                                  % (cores * per_core, a.seconds, per_core),
                        "n_act": cpu_n_act}
        if world > 1:
            open(gate, "w").close()

    from tunebfree_b200 import cabi, engine, tables, workloads
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

    # sampled organs of the timed batch, rendered by the oracle over every pass this run makes
    # (forked workers: before this process owns a CUDA context); compared bit for bit with the
    # audio of the last timed pass of both legs, outside the timed regions
    e2e_warm = max(1, min(a.warmup, 2))
    pass_dev = a.warmup + a.steps - 1
    pass_e2e = a.warmup + a.steps + 1 + e2e_warm + a.steps - 1
    parity_ref = None
    if not a.no_parity_check:
        n_samp = 8 if world == 1 else 4
        samp_slots = sorted(set(int(round(k * (n_inst - 1) / max(1, n_samp - 1))) for k in range(n_samp)))
        keep = [pass_dev] + ([] if a.no_e2e else [pass_e2e])
        parity_ref = parity_reference([global_ids[s_] for s_ in samp_slots], n_frag, a.config,
                                      (pass_dev if a.no_e2e else pass_e2e) + 1, keep)


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

    def barrier():
        if dist is not None:
            dist.barrier()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

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

        host_waves = bool(os.environ.get("TBF_BENCH_HOST_WAVES"))  # round 1's path, for comparison
        specs = (cabi.TonegenSpec * n_inst)()
        for k, g in enumerate(global_ids):
            sp = tables.InstanceSpec(freqs=tables.freqs_from_ratios(sweep_degs[int(g)]), seed=1)
            C.memmove(C.byref(specs[k]), C.byref(sp.tg), C.sizeof(cabi.TonegenSpec))
        # tbf_gpu_organs_create (tbf_pipeline.cpp): builder threads on the host cores, the calling
        # thread uploads in order; nothing of the loop runs in Python
        t_init0 = time.perf_counter()
        ids, handles, (t_wall, t_wait, t_upload) = ctx.create_organs(specs, wh_handles["default"], threads,
                                                                     device_waves=not host_waves)
        ctx.synchronize()
        t_init = time.perf_counter() - t_init0
        for k, h in enumerate(handles):
            tg_handles[k] = h
        init = {"organs_initialised_per_s": n_inst / t_init, "wall_s": t_init, "host_threads": threads,
                "upload_s": t_upload, "waiting_for_builders_s": t_wait,
                "wave_synthesis": "host" if host_waves else "device",
                "synth_samples_flagged_differed": None if host_waves else list(ctx.synth_stats()),
                "what": "128 frequencies -> wheel choice, play matrix, loop lengths, envelopes on the host "
                        "cores (tbf_build_tonegen_tables_synth on builder threads), the 256 wavetables of "
                        "every organ synthesised on the device (tbf_gpu_tonegen_tables_create_synth), "
                        "instance create: tbf_gpu_organs_create, builders and uploads overlap"}
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
    parity = None
    if parity_ref is not None:
        bad = 0
        for s_ in samp_slots:
            l_ref, r_ref = parity_ref[int(global_ids[s_])][pass_dev]
            got = out[:, s_].cpu().numpy()
            if not (np.array_equal(got[0].view(np.uint32), l_ref.view(np.uint32))
                    and np.array_equal(got[1].view(np.uint32), r_ref.view(np.uint32))):
                bad += 1
        parity = {"organs": len(samp_slots), "slots": samp_slots, "pass": pass_dev,
                  "device_resident_mismatches": bad,
                  "what": "audio of the last timed pass of the sampled organs, bit patterns, against the "
                          "oracle replaying every pass; checked outside the timed region"}
    st = ctx.export_state(int(ids[0]))
    stat_f, stat_w = ctx.instance_stats(ids)  # every organ of the batch, every pass so far
    n_act = stat_w / max(1, stat_f)
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
        slice_frag = max(bounds[k + 1] - bounds[k] for k in range(slices))
        slice_frames = slice_frag * 128
        # host layout: chunk-major (TBF_RENDER_OUT_CHUNKED: every staging chunk crosses the link as
        # one contiguous copy per channel) unless TBF_BENCH_E2E_LAYOUT=rows asks for [organ][sample]
        chunked = os.environ.get("TBF_BENCH_E2E_LAYOUT", "chunks") != "rows"
        hchunk = ctx.host_chunk()
        n_hchunks = -(-slice_frag // hchunk)
        shape = (n_hchunks, n_inst, hchunk * 128) if chunked else (n_inst, slice_frames)
        e2e_flags = flags | (cabi.RENDER_OUT_CHUNKED if chunked else 0)
        wc = bool(os.environ.get("TBF_BENCH_E2E_WC"))
        try:
            hl, pl = ctx.host_alloc(shape, write_combined=wc)
            hr, pr = ctx.host_alloc(shape, write_combined=wc)
            pinned = True
        except MemoryError:
            hl = np.empty(shape, dtype=np.float32)
            hr = np.empty(shape, dtype=np.float32)
            pl = pr = None
            pinned = False

        def host_rows(h, slot, n_frames):
            """the last call's audio of one slot as a flat array"""
            if chunked:
                return h[:, slot, :].reshape(-1)[:n_frames]
            return h[slot, :n_frames]

        # the link's own ceiling: a plain contiguous 1 GiB device -> page-locked host copy on
        # every rank at once (outside the timed region)
        barrier()
        link_gbs = ctx.measure_d2h(1 << 30, 8)  # ~8 GiB per rank: long enough to overlap on every rank
        barrier()
        # the step's inputs (the events) wait in page-locked memory, like the outputs
        ev_pinned = []

        def pinned_events(ev_arr):
            try:
                evs, addr = engine.events_from_numpy_pinned(ctx, ev_arr)
                ev_pinned.append(addr)
                return evs
            except MemoryError:
                return engine.events_from_numpy(ev_arr)

        if slices == 1:
            parts = [(n_frag, pinned_events(ev_np))]
        else:
            parts = []
            for k in range(slices):
                lo, hi = bounds[k], bounds[k + 1]
                sub = ev_np[(ev_np[:, 1] >= lo) & (ev_np[:, 1] < hi)].copy()
                sub[:, 1] -= lo
                parts.append((hi - lo, pinned_events(sub)))

        e2e_count = [0]

        e2e_split = [0.0, 0.0, 0.0]  # host seconds in prepare (event upload) | launch | synchronize

        def e2e_step():
            for nf_k, ev_k in parts:
                ta = time.perf_counter()
                ctx.prepare(n_inst, nf_k, ev_k, e2e_flags, hl.ctypes.data, hr.ctypes.data, slice_frames,
                            ids=ids)
                tb = time.perf_counter()
                ctx.launch()
                tc = time.perf_counter()
                ctx.synchronize()
                td = time.perf_counter()
                e2e_split[0] += tb - ta
                e2e_split[1] += tc - tb
                e2e_split[2] += td - tc
                e2e_count[0] += ctx.last_launch_ms()[1]

        for _ in range(e2e_warm):
            e2e_step()
        barrier()
        e2e_count[0] = 0
        e2e_split[:] = [0.0, 0.0, 0.0]
        t0 = time.perf_counter()
        for _ in range(a.steps):
            e2e_step()
        dt = time.perf_counter() - t0
        e2e_launches = e2e_count[0]
        barrier()
        dt = max_over_ranks(dt)
        e2e = {"value": world * organ_seconds_step * a.steps / dt, "unit": UNIT,
               "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": int(2 * n_inst * frames * 4),
               "host_buffers": ("page-locked" if pinned else "pageable") + (", write-combined" if wc else ""),
               "host_layout": "chunk-major [chunk][organ][%d samples] (TBF_RENDER_OUT_CHUNKED)" % (hchunk * 128)
                              if chunked else "[organ][sample]",
               "time_slices": slices,
               "host_ms_per_step": {"prepare": 1e3 * e2e_split[0] / a.steps, "launch": 1e3 * e2e_split[1] / a.steps,
                                    "synchronize": 1e3 * e2e_split[2] / a.steps,
                                    "link_alone": 1e3 * 2 * n_inst * frames * 4 / 1e9 / link_gbs if link_gbs > 0 else None},
               "peak_abs_sample": float(np.abs(host_rows(hl, 0, slice_frames)).max())}
        # audio bytes over the link per second against the link's own contiguous-copy rate
        link_min = -max_over_ranks(-link_gbs)
        e2e["link_peak_gbs"] = link_min
        e2e["link_peak_what"] = ("slowest rank's plain contiguous 1 GiB device -> page-locked host "
                                 "cudaMemcpyAsync, mean of 8 passes, all %d rank(s) copying at once" % world)
        e2e["audio_gbs_per_gpu"] = e2e["d2h_bytes_per_step"] * a.steps / dt / 1e9
        e2e["frac_of_link"] = e2e["audio_gbs_per_gpu"] / link_min if link_min > 0 else None
        if parity is not None:
            bad = 0
            lo = bounds[-2] * 128  # the host buffers hold the last time slice of the last pass
            for s_ in samp_slots:
                l_ref, r_ref = parity_ref[int(global_ids[s_])][pass_e2e]
                nfr = l_ref.shape[0] - lo
                if not (np.array_equal(host_rows(hl, s_, nfr).view(np.uint32), l_ref[lo:].view(np.uint32))
                        and np.array_equal(host_rows(hr, s_, nfr).view(np.uint32), r_ref[lo:].view(np.uint32))):
                    bad += 1
            parity["e2e_mismatches"] = bad
            parity["e2e_pass"] = pass_e2e
        if pl:
            ctx.host_free(pl)
            ctx.host_free(pr)
        parts = None
        for addr in ev_pinned:
            ctx.host_free(addr)

    if parity is not None and dist is not None:  # mismatches of every rank's sampled organs
        keys = ["device_resident_mismatches"] + (["e2e_mismatches"] if "e2e_mismatches" in parity else [])
        t = torch.tensor([parity[k] for k in keys] + [parity["organs"]], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        for k, v in zip(keys + ["organs"], t.tolist()):
            parity[k] = int(v)
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
        # algorithmic work per organ-second: FP32 lane-ops (mul / add / sub = 1 each,
        # the reference allows no FMA) W32 = 48000 (6 N_act + 17 + 253 [whirl]) and shared-memory
        # bytes Bsm = 48000 (4 N_act + 70 + 380 [whirl]); per stage:
        # T = 6 N_act + 11 (wheel sum, scanner, percussion differencer), R = 6 (mixdown) + 18 (two
        # horn biquads), W = 198 (taps: interpolation, FIRs, fractional scatter), D = 18 (two drum
        # biquads) + 19 (levels, leak, whirlProc3 mix)
        w32 = 48000 * (6 * n_act + 17 + 253 * whirl)
        bsm = 48000 * (4 * n_act + 70 + 380 * whirl)
        stage_w32 = {"T": 48000 * (6 * n_act + 11), "R": 48000 * (6 + 18 * whirl), "W": 48000 * 198,
                     "D": 48000 * 37}
        issue_peak = 148 * 128 * f_clk  # lane-ops/s = bytes/s of shared memory at 4 B per lane
        ach_fp32 = stage_w32[dom] * secs_per_dom_launch / dur
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
        roofline = {"bound": "fp32_issue", "achieved": ach_fp32 / 1e12, "peak": issue_peak / 1e12,
                    "unit": "T lane-ops/s", "frac": ach_fp32 / issue_peak, "traffic": traffic,
                    "peak_source": "148 SMs x 128 FP32 lanes x the SM clock sampled under this load "
                                   "(SURVEY 8(d): the path is bound by FP32 issue and shared-memory "
                                   "bandwidth, not by HBM or tensor cores)",
                    "note": "dominant stage kernel: algorithmic FP32 lane-ops of one launch / its average "
                            "launch duration (every stage kernel timed alone between CUDA events on its "
                            "launch stream, one extra pass after the timed region); path_* = the whole "
                            "path's algorithmic work x the timed rate",
                    "kernel": names[dom], "avg_launch_ms": dur * 1e3, "launches_profiled": dom_n,
                    "n_act": n_act,
                    "stage_ms_per_step": {names[k]: stage[k][0] for k in "TRWD"},
                    "path_fp32_issue": {"achieved": w32 * rate / 1e12, "peak": issue_peak / 1e12,
                                        "unit": "T lane-ops/s", "frac": (w32 * rate) / issue_peak},
                    "path_smem_bw": {"achieved": bsm * rate / 1e12, "peak": issue_peak / 1e12,
                                     "unit": "TB/s", "frac": (bsm * rate) / issue_peak},
                    "hbm": {"achieved": ach_hbm, "peak": hbm_peak, "unit": "GB/s",
                            "frac": ach_hbm / hbm_peak,
                            "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s",
                            "note": "the dominant kernel's algorithmic reads + writes over its launch time; "
                                    "not the bound: stage T's reads are the wavetables (4 B per wheel and "
                                    "sample), which stay in L2 -- what reaches DRAM is roofline.traffic"}}
        if cpu_baseline is not None and cpu_baseline.get("n_act"):
            cpu_baseline["n_act_gpu_arm"] = n_act
            cpu_baseline["same_workload"] = bool(abs(cpu_baseline["n_act"] - n_act) <= 0.01 * n_act)
        cfg = bench_config(a, n_inst, world)
        cfg.update({"table_sets": len(tg_handles), "whirl_table_sets": len(wh_handles), "init": init})
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": total_ms / a.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": cfg,
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches + (e2e_launches if e2e else 0),
            "roofline": roofline, "cpu_baseline": cpu_baseline,
            "checks": {"peak_abs_sample": peak_out, "error_flags": int(st.error_flags),
                       "parity_sampled": parity},
        }
        print(json.dumps(line), flush=True)
    ctx.close()
    if dist is not None:
        dist.barrier()
        if rank == 0 and os.path.exists(gate):
            os.remove(gate)
        dist.destroy_process_group()


def main():
    a = parse_args()
    if a.impl == "reference":
        run_reference_arm(a)
    else:
        run_gpu_arm(a)


if __name__ == "__main__":
    main()
