"""ctypes mirror of ``include/tbf_gpu.h`` (the C-ABI drop-in boundary).

Field order and types must match the header exactly; ``tests/test_cabi.py``
checks the struct sizes against the values the shared libraries report.
"""
import ctypes as C

FRAGMENT = 128
NOF_WHEELS = 256
NOF_BUSES = 27
NOF_DRAWBARS = 9
MAX_KEYS = 384
NOF_FREQS = 300
VIB_TBL = 2048
VIB_BUF = 1024
WHIRL_DISPLC = 16384
WHIRL_BUF = 2048

# tbf_status
OK = 0
ERR_ARG = -1
ERR_CUDA = -2
ERR_NOMEM = -3
ERR_HANDLE = -4
ERR_REFERENCE_UB = -5
ERR_GEOMETRY = -6

# tbf_event_type
EV_KEY_ON = 1
EV_KEY_OFF = 2
EV_DRAWBAR = 3
EV_PERC_ENABLE = 4
EV_PERC_SOFT = 5
EV_PERC_FAST = 6
EV_PERC_FIRST = 7
EV_VIB_UPPER = 8
EV_VIB_LOWER = 9
EV_VIB_KNOB = 10
EV_REV_OPTION = 11
EV_SWELL = 12
EV_REV_SELECT = 13
EV_WHIRL_BRAKE = 14
EV_WHIRL_RAMP = 15
EV_HORN_FILTER = 16

RENDER_WHIRL_BYPASS = 0x1
RENDER_WHIRLPROC3 = 0x2
RENDER_OUT_DEVICE = 0x4
RENDER_NO_OUTPUT = 0x8
RENDER_WHIRL_INPUT = 0x10
RENDER_STEMS = 0x20
RENDER_OUT_CHUNKED = 0x40
# testing / measurement: stage T variants (include/tbf_gpu.h TBF_RENDER_DEBUG_*)
RENDER_DEBUG_FULL_PASS = 0x100
RENDER_DEBUG_NO_PAIRS = 0x200
RENDER_DEBUG_BUS_ALL = 0x400
RENDER_DEBUG_NO_BUS = 0x800
RENDER_DEBUG_T_TMA = 0x1000
RENDER_DEBUG_T_CPASYNC = 0x2000
RENDER_DEBUG_T_DIRECT = 0x3000
RENDER_DEBUG_ALL_BUSES = 0x4000
RENDER_DEBUG_NO_TMA_TILES = 0x8000
HOST_WRITE_COMBINED = 0x1

IFLAG_COREPGM_OVERFLOW = 0x1
IFLAG_KEYCOMP_OOB = 0x2
IFLAG_BAD_EVENT = 0x4
IFLAG_GATHER_FALLBACK = 0x8

_f32p = C.POINTER(C.c_float)
_u16p = C.POINTER(C.c_uint16)


class TonegenTables(C.Structure):
    _fields_ = [
        ("wave_offset", C.c_uint32 * (NOF_WHEELS + 1)),
        ("wave_length", C.c_uint32 * (NOF_WHEELS + 1)),
        ("wave_data", _f32p),
        ("wave_total", C.c_uint64),
        ("contrib_start", C.c_uint32 * (MAX_KEYS + 1)),
        ("contrib_wheel", _u16p),
        ("contrib_bus", _u16p),
        ("contrib_level", _f32p),
        ("attack_env", (C.c_float * FRAGMENT) * NOF_DRAWBARS),
        ("release_env", (C.c_float * FRAGMENT) * NOF_DRAWBARS),
        ("key_comp_table", C.c_float * 128),
        ("draw_bar_level", (C.c_float * NOF_DRAWBARS) * NOF_BUSES),
        ("perc_env_gain_reset_norm", C.c_float),
        ("perc_env_gain_reset_soft", C.c_float),
        ("perc_env_scaling", C.c_float),
        ("perc_decay_fast_norm", C.c_float),
        ("perc_decay_fast_soft", C.c_float),
        ("perc_decay_slow_norm", C.c_float),
        ("perc_decay_slow_soft", C.c_float),
        ("perc_drawbar_normal_gain", C.c_float),
        ("perc_drawbar_soft_gain", C.c_float),
        ("perc_send_bus_a", C.c_int32),
        ("perc_send_bus_b", C.c_int32),
        ("perc_trigger_bus", C.c_int32),
        ("output_level_trim", C.c_float),
        ("vib_offset_table", (C.c_uint32 * VIB_TBL) * 3),
        ("vib_stator_increment", C.c_uint32),
        ("init_draw_bar_gain", C.c_float * NOF_BUSES),
        ("init_swell_pedal_gain", C.c_float),
        ("init_perc_env_gain_reset", C.c_float),
        ("init_perc_env_gain_decay", C.c_float),
        ("init_perc_drawbar_gain", C.c_float),
        ("init_perc_send_bus", C.c_int32),
        ("init_perc_trig_restore", C.c_int32),
        ("init_perc_is_soft", C.c_int32),
        ("init_perc_is_fast", C.c_int32),
        ("init_perc_enabled", C.c_int32),
        ("init_routing", C.c_uint32),
        ("init_vib_table", C.c_int32),
        ("init_vib_mixed", C.c_int32),
        ("init_draw_bar_change", C.c_uint16),
    ]


class WhirlTables(C.Structure):
    _fields_ = [
        ("sample_rate", C.c_double),
        ("hn_fwd_displ", _f32p),
        ("hn_bwd_displ", _f32p),
        ("dr_fwd_displ", _f32p),
        ("dr_bwd_displ", _f32p),
        ("bfw", _f32p),
        ("bbw", _f32p),
        ("horn_phase", C.c_int32 * 6),
        ("drum_phase", C.c_int32 * 6),
        ("horn_spacing", C.c_float * 6),
        ("drum_spacing", C.c_float * 6),
        ("hafw", C.c_float * 5),
        ("hbfw", C.c_float * 5),
        ("drfL", C.c_float * 5),
        ("drfR", C.c_float * 5),
        ("ha_param", C.c_float * 4),
        ("hb_param", C.c_float * 4),
        ("mic_angle", C.c_double),
        ("horn_acc", C.c_float),
        ("horn_dec", C.c_float),
        ("drum_acc", C.c_float),
        ("drum_dec", C.c_float),
        ("rev_horn_target", C.c_double * 9),
        ("rev_drum_target", C.c_double * 9),
        ("revselects", C.c_int32 * 3),
        ("hn_brake_pos", C.c_double),
        ("dr_brake_pos", C.c_double),
        ("horn_level", C.c_float),
        ("leakage", C.c_float),
        ("mic_hll", C.c_float),
        ("mic_hlr", C.c_float),
        ("mic_hrl", C.c_float),
        ("mic_hrr", C.c_float),
        ("mic_dll", C.c_float),
        ("mic_dlr", C.c_float),
        ("mic_drl", C.c_float),
        ("mic_drr", C.c_float),
        ("bypass", C.c_int32),
        ("init_horn_incr", C.c_double),
        ("init_drum_incr", C.c_double),
        ("init_horn_target", C.c_double),
        ("init_drum_target", C.c_double),
        ("init_horn_angle", C.c_double),
        ("init_drum_angle", C.c_double),
        ("init_horn_acdc", C.c_int32),
        ("init_drum_acdc", C.c_int32),
        ("init_rev_select", C.c_int32),
    ]


class Event(C.Structure):
    _fields_ = [
        ("instance", C.c_uint32),
        ("fragment", C.c_uint32),
        ("type", C.c_int32),
        ("a", C.c_int32),
        ("b", C.c_int32),
    ]


class WaveSynth(C.Structure):
    """tbf_wave_synth (include/tbf_tables.h)"""
    _fields_ = [
        ("sample_rate", C.c_double),
        ("seed", C.c_uint32),
        ("pad_", C.c_uint32),
        ("harmonics", C.c_double * 12),
        ("f1", C.c_double * (NOF_WHEELS + 1)),
        ("attenuation", C.c_double * (NOF_WHEELS + 1)),
    ]


class BatchDesc(C.Structure):
    _fields_ = [
        ("n_instances", C.c_uint32),
        ("instance_ids", C.POINTER(C.c_uint32)),
        ("n_fragments", C.c_uint32),
        ("events", C.POINTER(Event)),
        ("n_events", C.c_uint64),
        ("out_l", C.c_void_p),
        ("out_r", C.c_void_p),
        ("out_stride", C.c_uint64),
        ("flags", C.c_uint32),
        ("checksums", C.POINTER(C.c_uint64)),
        ("in_", C.c_void_p),
        ("in_stride", C.c_uint64),
        ("stems", C.c_void_p * 4),
    ]


class StateExport(C.Structure):
    _fields_ = [
        ("active_osc_l_end", C.c_int32),
        ("active_osc_list", C.c_int32 * (NOF_WHEELS + 1)),
        ("osc_pos", C.c_uint32 * (NOF_WHEELS + 1)),
        ("osc_acl_pos", C.c_int32 * (NOF_WHEELS + 1)),
        ("ref_count", C.c_int32 * (NOF_WHEELS + 1)),
        ("old_routing", C.c_uint32),
        ("new_routing", C.c_uint32),
        ("upper_key_count", C.c_uint32),
        ("key_down_count", C.c_int32),
        ("perc_env_gain", C.c_float),
        ("key_comp_level", C.c_float),
        ("pz", C.c_float),
        ("vib_stator", C.c_uint32),
        ("vib_out_pos", C.c_uint32),
        ("vib_buffer", C.c_float * VIB_BUF),
        ("whirl_outpos", C.c_uint32),
        ("horn_angle", C.c_double),
        ("drum_angle", C.c_double),
        ("horn_incr", C.c_double),
        ("drum_incr", C.c_double),
        ("horn_acdc", C.c_int32),
        ("drum_acdc", C.c_int32),
        ("whirl_buf", (C.c_float * WHIRL_BUF) * 4),
        ("iir_z", (C.c_float * 2) * 4),
        ("horn_iir", (C.c_float * 5) * 2),
        ("brake_pos", C.c_double * 2),
        ("ramp_one_minus_l", C.c_double * 4),
        ("filt_z", C.c_float * 4),
        ("adx", (C.c_float * 8) * 3),
        ("adi", C.c_int32 * 3),
        ("error_flags", C.c_uint32),
        ("pad_", C.c_uint32),
        ("stat_fragments", C.c_uint64),
        ("stat_active_wheels", C.c_uint64),
    ]

MAX_EXTRA_HARMONICS = 64


class TonegenSpec(C.Structure):
    """tbf_tonegen_spec (include/tbf_tables.h)"""
    _fields_ = [
        ("sample_rate", C.c_double),
        ("freqs", C.c_double * 128),
        ("target_ratio", C.c_double * 9),
        ("seed", C.c_uint32),
        ("precision", C.c_double),
        ("eq_macro", C.c_int32),
        ("eq_p1y", C.c_double),
        ("eq_r1y", C.c_double),
        ("eq_p4y", C.c_double),
        ("eq_r4y", C.c_double),
        ("compartment_crosstalk", C.c_double),
        ("transformer_crosstalk", C.c_double),
        ("terminalstrip_crosstalk", C.c_double),
        ("wiring_crosstalk", C.c_double),
        ("contribution_floor", C.c_double),
        ("contribution_min", C.c_double),
        ("n_extra_harmonics", C.c_int32),
        ("extra_harmonic_n", C.c_int32 * MAX_EXTRA_HARMONICS),
        ("extra_harmonic_level", C.c_float * MAX_EXTRA_HARMONICS),
        ("attack_model", C.c_int32),
        ("release_model", C.c_int32),
        ("attack_click_level", C.c_float),
        ("release_click_level", C.c_float),
        ("atk_clk_min_length", C.c_int32),
        ("atk_clk_max_length", C.c_int32),
        ("perc_gain_norm", C.c_float),
        ("perc_gain_soft", C.c_float),
        ("perc_gain_scaling", C.c_float),
        ("perc_bus_a", C.c_int32),
        ("perc_bus_b", C.c_int32),
        ("perc_bus_trig", C.c_int32),
        ("scanner_hz", C.c_double),
        ("scanner_amp", C.c_double * 3),
    ]


class WhirlSpec(C.Structure):
    """tbf_whirl_spec (include/tbf_tables.h)"""
    _fields_ = [
        ("sample_rate", C.c_double),
        ("bypass", C.c_int32),
        ("hn_brake_pos", C.c_double),
        ("dr_brake_pos", C.c_double),
        ("horn_rpm_slow", C.c_float),
        ("horn_rpm_fast", C.c_float),
        ("drum_rpm_slow", C.c_float),
        ("drum_rpm_fast", C.c_float),
        ("horn_acc", C.c_float),
        ("horn_dec", C.c_float),
        ("drum_acc", C.c_float),
        ("drum_dec", C.c_float),
        ("mic_angle", C.c_double),
        ("air_speed", C.c_float),
        ("mic_dist_cm", C.c_float),
        ("horn_x_offset_cm", C.c_float),
        ("horn_z_offset_cm", C.c_float),
        ("horn_radius_cm", C.c_float),
        ("drum_radius_cm", C.c_float),
        ("lp_type", C.c_int32),
        ("lp_f", C.c_double),
        ("lp_q", C.c_double),
        ("lp_g", C.c_double),
        ("ha_type", C.c_int32),
        ("ha_f", C.c_float),
        ("ha_q", C.c_float),
        ("ha_g", C.c_float),
        ("hb_type", C.c_int32),
        ("hb_f", C.c_float),
        ("hb_q", C.c_float),
        ("hb_g", C.c_float),
        ("horn_level", C.c_float),
        ("leak_level", C.c_float),
        ("drum_mic_width", C.c_float),
        ("horn_mic_width", C.c_float),
        ("rev_select", C.c_int32),
    ]
