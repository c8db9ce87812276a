"""Synthetic event scripts for BASELINE.json's configurations.

All randomness comes from splitmix64(instance_id xor config_id), so the GPU
renderer, the oracle and the CPU baseline see identical inputs.
"""
import numpy as np

from . import cabi, config5

FRAGMENTS_PER_SECOND = 375  # 48000 / 128
MASK64 = (1 << 64) - 1


def splitmix64(state):
    """One step of splitmix64 -> (new_state, output)."""
    state = (state + 0x9E3779B97F4A7C15) & MASK64
    z = state
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & MASK64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & MASK64
    return state, z ^ (z >> 31)


def chord_keys(instance_id, config_id=0, n_notes=10, lo=36, hi=96):
    """n_notes distinct upper-manual keys in [lo, hi] for this instance."""
    s = (instance_id ^ (config_id << 32)) & MASK64
    keys = []
    while len(keys) < n_notes:
        s, r = splitmix64(s)
        k = lo + r % (hi - lo + 1)
        if k not in keys:
            keys.append(k)
    return keys


def setup_events(config):
    """Fragment-0 control events of a config (before any key)."""
    ev = [(0, cabi.EV_DRAWBAR, b, 8) for b in range(9)]          # 888888888 upper
    ev += [(0, cabi.EV_PERC_ENABLE, 1, 0), (0, cabi.EV_PERC_SOFT, 0, 0),
           (0, cabi.EV_PERC_FAST, 1, 0), (0, cabi.EV_PERC_FIRST, 0, 0)]  # normal / fast / 3rd
    if config >= 3:
        ev += [(0, cabi.EV_VIB_UPPER, 1, 0), (0, cabi.EV_VIB_KNOB, 5, 0),  # C3 on the upper manual
               (0, cabi.EV_REV_OPTION, 8, 0)]                               # horn + drum fast
    return ev


SWEEP_CHORD = (36, 48, 55, 64, 70)  # config 4: the same five upper-manual keys in every tuning


def instance_events(instance_id, n_fragments, config=3, restrike=187, hold=140):
    """[(fragment, type, a, b)] for one instance: setup at fragment 0, then a 10-note
    chord struck every `restrike` fragments (~0.5 s) and released after `hold`.  Config 4
    (one organ per Scala scale) plays a fixed 5-note chord instead, same chain as config 3."""
    ev = setup_events(config)
    keys = list(SWEEP_CHORD) if config == 4 else chord_keys(instance_id, config)
    f = 0
    while f < n_fragments:
        ev += [(f, cabi.EV_KEY_ON, k, 0) for k in keys]
        off = f + hold
        if off < n_fragments:
            ev += [(off, cabi.EV_KEY_OFF, k, 0) for k in keys]
        f += restrike
    return ev


def batch_events_numpy(first_instance, n_instances, n_fragments, config=3, global_ids=None):
    """int32 [n, 5] rows (slot, fragment, type, a, b) sorted by (slot, fragment).  Slot i plays
    organ global_ids[i] (default: first_instance + i)."""
    rows = []
    for slot in range(n_instances):
        g = first_instance + slot if global_ids is None else int(global_ids[slot])
        evs = sorted(instance_events(g, n_fragments, config), key=lambda t: t[0])
        a = np.array(evs, dtype=np.int32).reshape(-1, 4)
        s = np.full((a.shape[0], 1), slot, dtype=np.int32)
        rows.append(np.concatenate([s, a], axis=1))
    return np.concatenate(rows, axis=0)


def organ_variant(global_id, config):
    """(osc variant, whirl variant, engine seed) of organ `global_id`.  Configs 2 and 3 use the
    compiled-in defaults with 16 engine seeds; config 5 walks the Cartesian mix of the
    reference's cfg/osc-*.cfg (and the default) x cfg/whirl-*.cfg."""
    seed = 1 + global_id % 16
    if config != 5:
        return ("default", "default", seed)
    osc = list(config5.OSC_VARIANTS)
    whirl = list(config5.WHIRL_VARIANTS)
    return (osc[global_id % len(osc)], whirl[(global_id // len(osc)) % len(whirl)], seed)


def variant_cfg(osc, whirl):
    """cfg text of a variant pair ('default' = no setting)."""
    return config5.OSC_VARIANTS.get(osc, "") + config5.WHIRL_VARIANTS.get(whirl, "")


def render_flags(config):
    if config == 2:
        return cabi.RENDER_WHIRL_BYPASS
    return cabi.RENDER_WHIRLPROC3


_SWEEP = None


def scala_sweep():
    """Config 4's tunings: (names, [degrees as ratios]) of every scale of the reference's Scala
    archive that both sides accept (tunebfree_b200/data/scala_sweep.npz, made by
    tests/golden/make_scala_sweep.py; 128 note frequencies = tables.freqs_from_ratios(degrees))."""
    global _SWEEP
    if _SWEEP is None:
        import os
        z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "scala_sweep.npz"))
        ok = np.flatnonzero(z["status"] == 0)
        off, deg = z["deg_off"], z["degrees"]
        _SWEEP = ([str(z["names"][i]) for i in ok], [deg[off[i]:off[i + 1]].copy() for i in ok],
                  z["digest"][ok])
    return _SWEEP
