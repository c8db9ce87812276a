"""Golden cases shared by the generator (make_golden.py) and the tests.

Each case: name -> dict(seed, cfg, scl (Scala text or None), flags, n_frag, events).
Small on purpose (the .npz files are committed): a few dozen fragments each."""
from tunebfree_b200 import cabi, workloads

E = cabi

PELOG_SCL = """! pelog-ish 5 tone test scale
5 tone unequal scale
 5
!
 120.0
 270.0
 540.0
 670.0
 2/1
"""


def _config1_events():
    """BASELINE config 1 shape: pgm/default.pgm program 1 ("Jazz 1 all": drawbars 888000000,
    percussion on / soft / fast / 3rd, vibrato C3 on the upper manual, rotary chorale =
    useRevOption(4)), C-major triad on at fragment 0, off at fragment 40."""
    ev = [(0, E.EV_DRAWBAR, b, 8 if b < 3 else 0) for b in range(9)]
    ev += [(0, E.EV_PERC_ENABLE, 1, 0), (0, E.EV_PERC_SOFT, 1, 0), (0, E.EV_PERC_FAST, 1, 0),
           (0, E.EV_PERC_FIRST, 0, 0), (0, E.EV_VIB_UPPER, 1, 0), (0, E.EV_VIB_KNOB, 5, 0),
           (0, E.EV_REV_OPTION, 4, 0)]
    ev += [(0, E.EV_KEY_ON, k, 0) for k in (60, 64, 67)]
    ev += [(40, E.EV_KEY_OFF, k, 0) for k in (60, 64, 67)]
    return ev


CASES = {
    "config1_default_pgm": dict(seed=1, cfg="", scl=None, flags=cabi.RENDER_WHIRLPROC3, n_frag=64,
                                events=_config1_events()),
    "config2_tonegen_only": dict(seed=2, cfg="", scl=None, flags=cabi.RENDER_WHIRL_BYPASS, n_frag=48,
                                 events=workloads.instance_events(5, 48, config=2, restrike=30, hold=20)),
    "config3_full_chain": dict(seed=3, cfg="", scl=None, flags=cabi.RENDER_WHIRLPROC3, n_frag=64,
                               events=workloads.instance_events(9, 64, config=3, restrike=40, hold=25)),
    "config4_microtuned": dict(seed=1, cfg="", scl=PELOG_SCL, flags=0, n_frag=40,
                               events=workloads.setup_events(3) + [(0, E.EV_KEY_ON, k, 0)
                                                                   for k in (48, 55, 60, 63, 70)]),
    "config5_sawtooth_sharp": dict(seed=4, cfg="osc.harmonic.1=1.0\nosc.harmonic.2=0.5\nosc.harmonic.3=0.333333333\n"
                                               "whirl.horn.filter.b.gain=-60\n",
                                   scl=None, flags=cabi.RENDER_WHIRLPROC3, n_frag=40,
                                   events=workloads.instance_events(21, 40, config=3, restrike=25, hold=15)),
}
