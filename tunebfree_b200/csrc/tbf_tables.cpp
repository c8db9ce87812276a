/* tbf_tables.cpp -- host-side table builder (see include/tbf_tables.h).
 *
 * From-scratch restatement of the reference's cold init path; each function
 * cites the reference lines it follows.  Arithmetic types (float vs double)
 * and evaluation order follow the reference expression by expression, because
 * every table entry has to come out bit-identical.
 * Compiled with -ffp-contract=off, no fast-math.
 */
#include <algorithm>
#include <chrono>
#include <math.h>
#include <cmath>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <strings.h>

#include <limits>
#include <string>
#include <vector>

#include "tbf_tables.h"

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

/* ------------------------------------------------------------------ */
/* glibc rand(): TYPE_3 additive feedback generator, degree 31, sep 3 */
/* (glibc stdlib/random_r.c; call sites src/tonegen.cpp:342, :1450,   */
/* :2577-2660).  r[k] = r[k-31] + r[k-3], output r[k] >> 1, the first */
/* 310 outputs are discarded by srandom().                            */
/* ------------------------------------------------------------------ */
extern "C" void tbf_srand(tbf_rand_state *st, uint32_t seed)
{
    int32_t word = (int32_t)seed;
    if (word == 0)
        word = 1;
    st->r[0] = word;
    for (int i = 1; i < 31; ++i) {
        const long hi = word / 127773;
        const long lo = word % 127773;
        long w = 16807 * lo - 2836 * hi;
        if (w < 0)
            w += 2147483647;
        word = (int32_t)w;
        st->r[i] = word;
    }
    /* ring of 31 words; idx = position of the "front" pointer (starts at 3) */
    st->idx = 3;
    st->r[31] = st->r[32] = st->r[33] = 0;
    for (int i = 0; i < 310; ++i)
        (void)tbf_rand(st);
}

extern "C" int32_t tbf_rand(tbf_rand_state *st)
{
    const int f = st->idx;
    const int r = (f + 31 - 3) % 31;
    const uint32_t v = (uint32_t)st->r[f] + (uint32_t)st->r[r];
    st->r[f] = (int32_t)v;
    st->idx = (f + 1) % 31;
    return (int32_t)(v >> 1);
}

namespace {

const int32_t kRandMax = 2147483647;

struct Rng {
    tbf_rand_state st;
    explicit Rng(uint32_t seed) { tbf_srand(&st, seed); }
    int32_t next() { return tbf_rand(&st); }
    /* drnd(), src/tonegen.cpp:342 */
    double drnd() { return ((double)next()) / (double)kRandMax; }
    /* n draws thrown away: x[i] = x[i - 31] + x[i - 3] on a linear buffer (no ring arithmetic) */
    void skip(uint64_t n)
    {
        enum { CHUNK = 4096 };
        uint32_t lin[31 + CHUNK];
        for (int j = 0; j < 31; ++j)
            lin[j] = (uint32_t)st.r[(st.idx + j) % 31];
        while (n) {
            const uint32_t m = (uint32_t)(n < CHUNK ? n : CHUNK);
            for (uint32_t i = 31; i < 31 + m; ++i)
                lin[i] = lin[i - 31] + lin[i - 3];
            memmove(lin, lin + m, sizeof(uint32_t) * 31);
            n -= m;
        }
        for (int j = 0; j < 31; ++j)
            st.r[j] = (int32_t)lin[j];
        st.idx = 0;
    }
};

/* dBToGain, src/tonegen.cpp:337 */
double db_to_gain(double dB) { return pow(10.0, (dB / 20.0)); }

} // namespace

/* ------------------------------------------------------------------ */
/* tuning                                                              */
/* ------------------------------------------------------------------ */
extern "C" void tbf_freqs_12tet(double freqs[128])
{
    for (int n = 0; n < 128; ++n)
        freqs[n] = 440.0 * pow(2.0, (n - 69) / 12.0);
}

/* Scala scale file: first non-comment line = description, second = number of
 * degrees, then one pitch per line: with a '.' it is cents, otherwise a ratio
 * n/d or an integer.  Degree 0 (1/1) sits on MIDI note 60 at 261.6255653005986 Hz
 * and the last entry is the period. */
static int parse_scl(const char *text, std::vector<double> &ratio)
{
    if (!text)
        return TBF_ERR_ARG;
    std::vector<std::string> lines;
    const char *p = text;
    while (*p) {
        const char *e = strchr(p, '\n');
        std::string ln = e ? std::string(p, e - p) : std::string(p);
        p = e ? e + 1 : p + ln.size();
        while (!ln.empty() && (ln.back() == '\r' || ln.back() == ' ' || ln.back() == '\t'))
            ln.pop_back();
        size_t a = 0;
        while (a < ln.size() && (ln[a] == ' ' || ln[a] == '\t'))
            ++a;
        if (a < ln.size() && ln[a] == '!')
            continue;
        lines.push_back(ln.substr(a));
    }
    if (lines.size() < 2)
        return TBF_ERR_ARG;
    const int n = atoi(lines[1].c_str());
    if (n < 1 || (size_t)n + 2 > lines.size())
        return TBF_ERR_ARG;
    ratio.assign(n + 1, 1.0);
    for (int i = 0; i < n; ++i) {
        std::string tok = lines[2 + i];
        size_t sp = tok.find_first_of(" \t");
        if (sp != std::string::npos)
            tok = tok.substr(0, sp);
        if (tok.empty())
            return TBF_ERR_ARG;
        double r;
        if (tok.find('.') != std::string::npos) {
            r = pow(2.0, atof(tok.c_str()) / 1200.0);
        } else {
            size_t sl = tok.find('/');
            if (sl != std::string::npos) {
                const double den = atof(tok.substr(sl + 1).c_str());
                if (den == 0.0)
                    return TBF_ERR_ARG;
                r = atof(tok.substr(0, sl).c_str()) / den;
            } else {
                r = atof(tok.c_str());
            }
        }
        if (!(r > 0.0))
            return TBF_ERR_ARG;
        ratio[i + 1] = r;
    }
    return n;
}

extern "C" int tbf_scl_ratios(const char *text, double *degrees, int capacity)
{
    std::vector<double> ratio;
    const int n = parse_scl(text, ratio);
    if (n < 0)
        return n;
    if (!degrees || n > capacity)
        return TBF_ERR_ARG;
    memcpy(degrees, ratio.data() + 1, sizeof(double) * n);
    return n;
}

extern "C" int tbf_freqs_from_scl(const char *text, double freqs[128])
{
    std::vector<double> ratio;
    const int n = parse_scl(text, ratio);
    if (n < 0)
        return n;
    return tbf_freqs_from_ratios(ratio.data() + 1, n, freqs);
}

extern "C" int tbf_freqs_from_ratios(const double *degrees, int n, double freqs[128])
{
    if (!degrees || !freqs || n < 1)
        return TBF_ERR_ARG;
    for (int i = 0; i < n; ++i)
        if (!(degrees[i] > 0.0))
            return TBF_ERR_ARG;
    const double base = 261.6255653005986;
    const double period = degrees[n - 1];
    for (int note = 0; note < 128; ++note) {
        const int d = note - 60;
        int oct = d / n, deg = d % n;
        if (deg < 0) {
            deg += n;
            oct -= 1;
        }
        freqs[note] = base * (deg ? degrees[deg - 1] : 1.0) * pow(period, (double)oct);
    }
    return TBF_OK;
}

namespace {

/* inferScaleSize, src/tuning.cpp:48-106: `period` is a float there */
void infer_scale_size(const double *f, int *scale_size, float *period_out)
{
    for (float period = 2.0f; period < 10.0f; period++) {
        for (int ss = 1; ss < 128; ++ss) {
            bool mismatch = false;
            for (int i = 0; i < 128 - ss; ++i) {
                if (fabs(f[i + ss] / f[i] - (double)period) > 1e-6) {
                    mismatch = true;
                    break;
                }
            }
            if (!mismatch) {
                *scale_size = ss;
                *period_out = period;
                return;
            }
        }
    }
    for (int ss = 1; ss < 128; ++ss) {
        const float period = (float)(f[ss] / f[0]);
        bool mismatch = false;
        for (int i = 0; i < 128 - ss; ++i) {
            if (fabs(f[i + ss] / f[i] - (double)period) > 1e-6) {
                mismatch = true;
                break;
            }
        }
        if (!mismatch) {
            *scale_size = ss;
            *period_out = period;
            return;
        }
    }
    *scale_size = -1;
    *period_out = -1.0f;
}

} // namespace

/* extendFrequencies, src/tuning.cpp:115-135 */
extern "C" int tbf_extend_frequencies(const double freqs128[128], double out[TBF_NOF_FREQS],
                                      int *scale_size, float *period)
{
    int ss;
    float per;
    for (int i = 0; i < 128; ++i)
        out[i] = freqs128[i];
    infer_scale_size(out, &ss, &per);
    if (ss > 0) {
        for (int i = 128; i < TBF_NOF_FREQS; ++i)
            out[i] = (double)per * out[i - ss];
    } else {
        for (int i = 128; i < TBF_NOF_FREQS; ++i)
            out[i] = out[127];
    }
    if (scale_size)
        *scale_size = ss;
    if (period)
        *period = per;
    return TBF_OK;
}

/* ------------------------------------------------------------------ */
/* spec defaults                                                       */
/* ------------------------------------------------------------------ */
extern "C" void tbf_tonegen_spec_default(tbf_tonegen_spec *s, double sample_rate)
{
    memset(s, 0, sizeof(*s));
    s->sample_rate = sample_rate;
    tbf_freqs_12tet(s->freqs);
    const double tr[9] = {0.5, 1.5, 1, 2, 3, 4, 5, 6, 8}; /* src/tonegen.cpp:2959 */
    memcpy(s->target_ratio, tr, sizeof(tr));
    s->seed = 1;
    /* initValues, src/tonegen.cpp:196-300 */
    s->precision = 0.001;
    s->eq_macro = TBF_EQ_SPLINE;
    s->eq_p1y = 1.0;
    s->eq_r1y = 0.0;
    s->eq_p4y = 1.0;
    s->eq_r4y = 0.0;
    s->compartment_crosstalk = 0.01;
    s->transformer_crosstalk = 0.0;
    s->terminalstrip_crosstalk = 0.01;
    s->wiring_crosstalk = 0.01;
    s->contribution_floor = 0.0000158;
    s->contribution_min = 0.0;
    s->attack_model = TBF_ENV_CLICK;
    s->release_model = TBF_ENV_LINEAR;
    s->attack_click_level = 0.50;
    s->release_click_level = 0.25;
    s->atk_clk_min_length = -1;
    s->atk_clk_max_length = -1;
    s->perc_gain_norm = 1.0;
    s->perc_gain_soft = 0.5012;
    s->perc_gain_scaling = 11.0; /* HIPASS_PERCUSSION */
    s->perc_bus_a = 3;
    s->perc_bus_b = 4;
    s->perc_bus_trig = 8;
    /* reset_vibrato, src/vibrato.cpp:288-305 */
    s->scanner_hz = 7.25;
    s->scanner_amp[0] = 3.0;
    s->scanner_amp[1] = 6.0;
    s->scanner_amp[2] = 9.0;
}

extern "C" void tbf_whirl_spec_default(tbf_whirl_spec *s, double sample_rate)
{
    memset(s, 0, sizeof(*s));
    s->sample_rate = sample_rate;
    /* initValues, src/whirl.cpp:40-131 */
    s->horn_rpm_slow = 60.0 * 0.672;
    s->horn_rpm_fast = 60.0 * 7.056;
    s->drum_rpm_slow = 60.0 * 0.600;
    s->drum_rpm_fast = 60.0 * 5.955;
    s->horn_acc = 0.161;
    s->horn_dec = 0.321;
    s->drum_acc = 4.127;
    s->drum_dec = 1.371;
    s->mic_angle = 0;
    s->air_speed = 340.0;
    s->mic_dist_cm = 42.0;
    s->horn_x_offset_cm = 0.0;
    s->horn_z_offset_cm = 0.0;
    s->horn_radius_cm = 19.2;
    s->drum_radius_cm = 22.0;
    s->lp_type = 8; /* EQC_HIGH */
    s->lp_f = 811.9695;
    s->lp_q = 1.6016;
    s->lp_g = -38.9291;
    s->ha_type = 0; /* EQC_LPF */
    s->ha_f = 4500;
    s->ha_q = 2.7456;
    s->ha_g = -30.0;
    s->hb_type = 7; /* EQC_LOW */
    s->hb_f = 300.0;
    s->hb_q = 1.0;
    s->hb_g = -30.0;
    s->horn_level = 0.7;
    s->leak_level = 0.15;
    s->drum_mic_width = 0;
    s->horn_mic_width = 0;
    s->rev_select = 0;
}

/* ------------------------------------------------------------------ */
/* tone generator wiring model                                         */
/* ------------------------------------------------------------------ */
namespace {

/* physical layout tables of the instrument (facts about the hardware,
 * src/tonegen.cpp:205-233, src/tuning.cpp:154-165) */
const short kNorthTransformers[] = {85, 66, 90, 71, 47, 64, 86, 69, 45, 62, 86, 67, 91,
                                    72, 48, 65, 89, 70, 46, 63, 87, 68, 44, 61, 0};
const short kSouthTransformers[] = {78, 54, 83, 59, 76, 52, 81, 57, 74, 50, 79, 55, 84,
                                    60, 77, 53, 82, 58, 75, 51, 80, 56, 73, 49, 0};
const short kTerminalStrip[] = {
    85, 42, 30, 76, 66, 18, 6,  54, 90, 35, 83, 71, 23, 11, 59, 47, 40, 28, 76, 64, 16, 4,  52,
    88, 33, 81, 69, 21, 9,  57, 45, 34, 26, 74, 62, 14, 2,  50, 86, 43, 31, 79, 67, 19, 7,  55,
    91, 36, 84, 72, 24, 12, 60, 48, 41, 29, 77, 65, 17, 5,  53, 89, 34, 82, 70, 22, 10, 58, 46,
    39, 27, 75, 63, 15, 3,  51, 87, 32, 80, 68, 20, 8,  56, 44, 37, 25, 73, 61, 13, 1,  49, 0};

/* wheel -> wheel sharing its compartment; zero where none (wheels 37..41) */
short paired_wheel(short n)
{
    const int q = n / 92, r = n % 92;
    int p;
    if (r == 0)
        p = 0;
    else if (r <= 36)
        p = r + 48;
    else if (r <= 41)
        p = 0;
    else if (r <= 48)
        p = r + 43;
    else if (r <= 84)
        p = r - 48;
    else
        p = r - 43;
    return (short)(q * 92 + p);
}

struct Conn { /* one list element: (terminal or wheel, bus, level) */
    short a;
    short bus;
    float level;
};

/* the default taper of a manual key's contact on bus b in dB
 * (taperingModel, src/tonegen.cpp:501-696: a table of break points) */
double tapering_db(int key, int bus)
{
    const double m3 = -10.0, m2 = -7.0, m1 = -3.5, r0 = 0.0, p1 = 3.5, p2 = 7.0;
    struct Seg {
        int below;
        double db;
    };
    static const Seg t0[] = {{12, m3}, {17, m2}, {24, m1}, {36, r0}, {48, p1}, {1 << 30, p2}};
    static const Seg t1[] = {{15, m1}, {38, r0}, {50, p1}, {1 << 30, p2}};
    static const Seg t2[] = {{17, m2}, {22, m1}, {37, r0}, {49, p1}, {1 << 30, p2}};
    static const Seg t3[] = {{17, m1}, {39, r0}, {1 << 30, m1}};
    static const Seg t4[] = {{14, p2}, {20, p1}, {40, r0}, {50, m1}, {1 << 30, m2}};
    static const Seg t5[] = {{12, p2}, {15, p1}, {41, r0}, {54, m1}, {1 << 30, m2}};
    static const Seg t6[] = {{14, p1}, {42, r0}, {50, m1}, {1 << 30, m2}};
    static const Seg t7[] = {{43, r0}, {48, m1}, {1 << 30, m2}};
    static const Seg t8[] = {{43, r0}, {1 << 30, m2}};
    static const Seg *const tabs[9] = {t0, t1, t2, t3, t4, t5, t6, t7, t8};
    const int k = key - 36; /* the fork shifted the key numbers up by 36 */
    if (bus < 0 || bus > 8)
        return r0;
    for (const Seg *s = tabs[bus];; ++s)
        if (k < s->below)
            return s->db;
}

bool east_west(const short *v, int w, int *east, int *west)
{
    *east = *west = 0;
    for (int i = 0; v[i] > 0; ++i) {
        if (v[i] == (short)w) {
            if (i > 0)
                *east = v[i - 1];
            *west = v[i + 1];
            return true;
        }
    }
    return false;
}

struct Wiring {
    std::vector<Conn> terminal_mix[TBF_NOF_WHEELS + 1];
    std::vector<Conn> key_taper[TBF_MAX_KEYS];
    std::vector<Conn> key_crosstalk[TBF_MAX_KEYS];
};

/* getOscillatorFrequency, src/tonegen.cpp:698-706 */
double osc_frequency(const double *freq300, int wheel)
{
    return fmin(fmax(freq300[wheel - 1], 12.0), 2.5e10);
}

/* applyManualDefaults, src/tonegen.cpp:707-802: per key and bus choose the wheel
 * whose frequency ratio to the key is closest (in cents) to the drawbar's target
 * ratio; float arithmetic for ratio / cent difference as in the reference */
void manual_defaults(const tbf_tonegen_spec *s, const double *freq300, Wiring &w, int key_offset,
                     int bus_offset)
{
    double of[TBF_NOF_WHEELS + 1], log_of[TBF_NOF_WHEELS + 1], log_tr[9];
    for (int i = 1; i <= TBF_NOF_WHEELS; ++i) {
        of[i] = osc_frequency(freq300, i);
        log_of[i] = log2(of[i]);
    }
    for (int b = 0; b < 9; ++b)
        log_tr[b] = log2(s->target_ratio[b]);
    bool logs_finite = true;
    for (int i = 1; i <= TBF_NOF_WHEELS; ++i)
        logs_finite = logs_finite && std::isfinite(log_of[i]);
    /* the reference's distance of wheel t, evaluated exactly as it does */
    auto exact_cent = [&](int t, int k, int b) {
        const float ratio = (float)(of[t] / freq300[k]);
        return (float)(1200 * fabs(log2(s->target_ratio[b] / (double)ratio)));
    };
    for (int k = 0; k < 128; ++k) {
        const int key = k + key_offset;
        if (!w.key_taper[key].empty())
            continue;
        w.key_taper[key].reserve(9);
        const double log_fk = log2(freq300[k]);
        for (int b = 0; b < 9; ++b) {
            /* The reference scans all 256 wheels with one log2 each (300 000 calls per manual).
             * Here: the distance from precomputed logarithms first -- it is within
             * 1e-3 + 1e-6 * cent of the reference's float value (float rounding of the ratio
             * and of the result) -- then the reference's own expression, in the reference's
             * order and with its strict comparison, for the wheels within twice that of the
             * approximate minimum: the wheel the full scan picks is always among them. */
            double approx[TBF_NOF_WHEELS + 1]; /* in octaves: cents / 1200 */
            double a_min = std::numeric_limits<double>::infinity();
            const double centre = log_tr[b] + log_fk;
            const bool finite = logs_finite && std::isfinite(centre);
            for (int t = 1; t <= TBF_NOF_WHEELS; ++t) {
                approx[t] = fabs(centre - log_of[t]);
                a_min = approx[t] < a_min ? approx[t] : a_min;
            }
            const double limit = finite ? a_min + 2.0 * (1e-3 / 1200 + 1e-6 * a_min)
                                        : std::numeric_limits<double>::infinity();
            float smallest = std::numeric_limits<float>::infinity();
            int best = 0;
            for (int t = 1; t <= TBF_NOF_WHEELS; ++t) {
                if (finite && approx[t] > limit)
                    continue;
                const float cent = exact_cent(t, k, b);
                if (cent < smallest) {
                    smallest = cent;
                    best = t;
                }
            }
            /* a hit on the first or last wheel means the search ran off the range */
            if (best != 0 && best != 1 && best != TBF_NOF_WHEELS) {
                Conn c;
                c.a = (short)best;
                c.bus = (short)(b + bus_offset);
                c.level = (float)db_to_gain(tapering_db(k, b));
                w.key_taper[key].push_back(c);
            }
        }
    }
}

/* applyPedalDefaults, src/tonegen.cpp:810-841 */
void pedal_defaults(Wiring &w, int n_pedals)
{
    const int offs[9] = {-12, 7, 0, 12, 19, 24, 28, 31, 36};
    for (int k = 0; k < n_pedals; ++k) {
        const int key = k + 256;
        if (!w.key_taper[key].empty())
            continue;
        for (int b = 0; b < 9; ++b) {
            const int t = (k + 1) + offs[b];
            if (t < 1 || t > TBF_NOF_WHEELS)
                continue;
            Conn c;
            c.a = (short)t;
            c.bus = (short)(b + 18);
            c.level = (float)db_to_gain(0.0);
            w.key_taper[key].push_back(c);
        }
    }
}

/* applyDefaultCrosstalk, src/tonegen.cpp:849-879: each contact also hears the
 * other contacts of its key, attenuated by vertical distance */
void default_crosstalk(const tbf_tonegen_spec *s, Wiring &w, int key_offset, int bus_offset)
{
    for (int k = 0; k < 128; ++k) {
        const int key = k + key_offset;
        if (!w.key_crosstalk[key].empty())
            continue;
        w.key_crosstalk[key].reserve(9 * w.key_taper[key].size());
        for (int b = 0; b < 9; ++b) {
            const int bus = bus_offset + b;
            for (const Conn &c : w.key_taper[key]) {
                if (c.bus == bus)
                    continue;
                Conn n;
                n.a = c.a;
                n.bus = (short)bus;
                n.level = (float)((s->wiring_crosstalk * (double)c.level) / abs(bus - c.bus));
                w.key_crosstalk[key].push_back(n);
            }
        }
    }
}

/* applyDefaultConfiguration, src/tonegen.cpp:941-1023 */
int default_configuration(const tbf_tonegen_spec *s, const double *freq300, Wiring &w)
{
    for (int i = 1; i <= TBF_NOF_WHEELS; ++i) {
        Conn c;
        c.a = (short)i;
        c.bus = 0;
        c.level = (float)(1.0 - s->compartment_crosstalk);
        w.terminal_mix[i].push_back(c);
        if (0.0 < s->compartment_crosstalk) {
            const short p = paired_wheel((short)i);
            if (0 < p && p <= TBF_NOF_WHEELS) {
                Conn d;
                d.a = p;
                d.bus = 0;
                d.level = (float)s->compartment_crosstalk;
                w.terminal_mix[i].push_back(d);
            }
        }
    }
    if (0.0 < s->transformer_crosstalk) {
        for (int i = 44; i <= TBF_NOF_WHEELS; ++i) {
            int east, west;
            if (!east_west(kNorthTransformers, i, &east, &west) &&
                !east_west(kSouthTransformers, i, &east, &west))
                return TBF_ERR_REFERENCE_UB; /* reference: assert(0), src/tonegen.cpp:925-937 */
            const int nb[2] = {east, west};
            for (int j = 0; j < 2; ++j)
                if (0 < nb[j]) {
                    Conn c;
                    c.a = (short)nb[j];
                    c.bus = 0;
                    c.level = (float)s->transformer_crosstalk;
                    w.terminal_mix[i].push_back(c);
                }
        }
    }
    if (0.0 < s->terminalstrip_crosstalk) {
        for (int i = 1; i <= TBF_NOF_WHEELS; ++i) {
            int east, west;
            east_west(kTerminalStrip, i, &east, &west);
            const int nb[2] = {east, west};
            for (int j = 0; j < 2; ++j)
                if (0 < nb[j]) {
                    Conn c;
                    c.a = (short)nb[j];
                    c.bus = 0;
                    c.level = (float)s->terminalstrip_crosstalk;
                    w.terminal_mix[i].push_back(c);
                }
        }
    }
    manual_defaults(s, freq300, w, 0, 0);
    manual_defaults(s, freq300, w, 128, 9);
    pedal_defaults(w, 32);
    default_crosstalk(s, w, 0, 0);
    default_crosstalk(s, w, 128, 9);
    return TBF_OK;
}

/* compilePlayMatrix + cpmInsert, src/tonegen.cpp:1042-1213: resolve every
 * (terminal, bus, level) contact of a key through the terminal mix into
 * (wheel, bus) gains, accumulated in float in list order, then ordered by
 * (wheel, bus) */
void compile_play_matrix(const tbf_tonegen_spec *s, const Wiring &w, std::vector<uint32_t> &start,
                         std::vector<uint16_t> &cw, std::vector<uint16_t> &cb,
                         std::vector<float> &cl)
{
    start.assign(TBF_MAX_KEYS + 1, 0);
    cw.reserve(49152);
    cb.reserve(49152);
    cl.reserve(49152);
    static thread_local float gain[TBF_NOF_WHEELS + 1][TBF_NOF_BUSES];
    /* the cells a key touches (a few hundred of the 257 x 27) as bit masks: which buses of a
     * wheel, which wheels -- so that neither the clearing nor the (wheel, bus)-ordered read-out
     * walks the whole matrix 384 times, and nothing has to be sorted */
    uint32_t bus_mask[TBF_NOF_WHEELS + 1] = {0};
    uint64_t wheel_mask[(TBF_NOF_WHEELS + 64) / 64] = {0};
    for (int k = 0; k < TBF_MAX_KEYS; ++k) {
        start[k] = (uint32_t)cw.size();
        const std::vector<Conn> *lists[2] = {&w.key_taper[k], &w.key_crosstalk[k]};
        for (int l = 0; l < 2; ++l)
            for (const Conn &c : *lists[l]) {
                if (c.a < 0 || c.a > TBF_NOF_WHEELS)
                    continue;
                for (const Conn &t : w.terminal_mix[c.a]) {
                    const float g = t.level * c.level;
                    if (g == 0.0)
                        continue;
                    const uint32_t bit = 1u << c.bus;
                    if (!(bus_mask[t.a] & bit)) {
                        bus_mask[t.a] |= bit;
                        wheel_mask[t.a >> 6] |= 1ull << (t.a & 63);
                        gain[t.a][c.bus] = g;
                    } else {
                        gain[t.a][c.bus] += g;
                    }
                }
            }
        /* wheel-major, bus-minor: the reference's order (src/tonegen.cpp:1183-1201) */
        for (size_t word = 0; word < sizeof(wheel_mask) / sizeof(wheel_mask[0]); ++word) {
            uint64_t wm = wheel_mask[word];
            wheel_mask[word] = 0;
            while (wm) {
                const int wh = (int)(word * 64) + __builtin_ctzll(wm);
                wm &= wm - 1;
                uint32_t bm = bus_mask[wh];
                bus_mask[wh] = 0;
                while (bm) {
                    const int b = __builtin_ctz(bm);
                    bm &= bm - 1;
                    float g = gain[wh][b];
                    if ((double)g < s->contribution_floor)
                        continue;
                    if ((double)g < s->contribution_min)
                        g = (float)s->contribution_min;
                    cw.push_back((uint16_t)wh);
                    cb.push_back((uint16_t)b);
                    cl.push_back(g);
                }
            }
        }
    }
    start[TBF_MAX_KEYS] = (uint32_t)cw.size();
}

/* ------------------------------------------------------------------ */
/* wavetables                                                          */
/* ------------------------------------------------------------------ */

/* fitWave, src/tonegen.cpp:1335-1369: shortest loop of a whole number of
 * periods whose length is within `precision` of an integer */
int fit_wave(double hz, double precision, int min_samples, int max_samples, double rate, size_t *out)
{
    double min_err = 99999.9, min_spn = 0.0;
    const int min_waves = (int)ceil((hz * (double)min_samples) / rate);
    const int max_waves = (int)floor((hz * (double)max_samples) / rate);
    if (!(min_samples < max_samples) || !(min_waves <= max_waves) || !(min_waves > 0))
        return TBF_ERR_REFERENCE_UB; /* reference asserts */
    /* The reference divides once per candidate (up to ~2000 candidates for the top wheels).
     * Here a product with the reciprocal comes first: it is within 1e-8 of the quotient (the
     * quotient is below 2^13, the reciprocal's relative error 2^-52), and only a candidate that
     * could lower the running minimum or pass the precision test by that margin is evaluated
     * with the reference's own division -- same minimum, same first hit, same result. */
    const double inv_hz = 1.0 / hz, margin = 1e-8;
    for (int i = min_waves; i <= max_waves; ++i) {
        const double approx = (rate * i) * inv_hz;
        const double approx_err = fabs(approx - rint(approx));
        if (approx_err >= min_err + margin && approx_err >= precision + margin && approx_err < 0.5 - margin)
            continue;
        const double nws = (rate * i) / hz;
        const double spn = rint(nws);
        const double err = fabs(nws - spn);
        if (err < min_err) {
            min_err = err;
            min_spn = spn;
        }
        if (err < precision)
            break;
    }
    if (!(0.0 < min_spn) || !(min_spn <= max_samples))
        return TBF_ERR_REFERENCE_UB;
    *out = (size_t)min_spn;
    return TBF_OK;
}

/* wheel equalisation, src/tonegen.cpp:1222-1316 */
double damper_curve(int tg, int first, int last, double w, double v, double u)
{
    const double x = ((double)(tg - first)) / ((double)(last - first));
    const double z = (x * (u - v)) - u;
    return 1.0 - w * z * z;
}

void wheel_attenuation(const tbf_tonegen_spec *s, double *att)
{
    const int n = TBF_NOF_WHEELS;
    if (s->eq_macro == TBF_EQ_SPLINE) { /* constrained Hermite spline */
        const double k = n - 1;
        for (int i = 1; i <= n; ++i) {
            const double t = ((double)(i - 1)) / k;
            const double t2 = t * t;
            const double t3 = t2 * t;
            const double r = s->eq_p1y * (2.0 * t3 - 3.0 * t2 + 1.0) +
                             s->eq_p4y * (-2.0 * t3 + 3.0 * t2) + s->eq_r1y * (t3 - 2.0 * t2 + t) +
                             s->eq_r4y * (t3 - t2);
            att[i] = (r < 0.0) ? 0.0 : (1.0 < r) ? 1.0 : r;
        }
    } else if (s->eq_macro == TBF_EQ_PEAK24) {
        for (int i = 1; i <= 43; ++i)
            att[i] = damper_curve(i, 1, 43, 0.2, -0.8, 1.0);
        for (int i = 44; i <= 48; ++i)
            att[i] = damper_curve(i, 44, 48, 1.6, -0.4, -0.3);
        for (int i = 49; i <= n; ++i)
            att[i] = damper_curve(i, 49, n, 0.9, -1.0, -0.7);
    } else {
        for (int i = 1; i <= 43; ++i)
            att[i] = damper_curve(i, 1, 43, 0.3, 0.4, 1.0);
        for (int i = 44; i <= 48; ++i)
            att[i] = damper_curve(i, 44, 48, 0.1, -0.4, 0.4);
        for (int i = 49; i <= n; ++i)
            att[i] = damper_curve(i, 49, n, 0.8, -1.0, -0.3);
    }
}

/* writeSamples, src/tonegen.cpp:1402-1457: up to 12 partials, amplitudes
 * normalised by the sum of magnitudes, partials at or above Nyquist muted,
 * one rand() per sample for a one-LSB (1/32767) dither.  Partials whose
 * amplitude is zero add +-0 to a sum that is never -0, so they are skipped. */
struct WheelSynth {
    double apl[TBF_MAX_PARTIALS], pl_hz[TBF_MAX_PARTIALS], U;
};

void wheel_synth(WheelSynth &ws, const double *ap, double attenuation, double f1, double rate)
{
    double sum = 0.0;
    for (int i = 0; i < TBF_MAX_PARTIALS; ++i) {
        ws.apl[i] = ap[i];
        sum += fabs(ws.apl[i]);
        ws.pl_hz[i] = f1 * ((double)(i + 1));
        if ((rate * 0.5) <= ws.pl_hz[i])
            ws.apl[i] = 0.0;
    }
    ws.U = attenuation / sum;
}

inline float synth_sample(const WheelSynth &ws, double rate, unsigned int i, bool dither)
{
    const double full_circle = 2.0 * M_PI;
    const float lsb = (float)(1.0 / 32767.0);
    double sacc = 0.0;
    for (int j = 0; j < TBF_MAX_PARTIALS; ++j) {
        if (ws.apl[j] == 0.0)
            continue;
        sacc += ws.apl[j] * sin(remainder((ws.pl_hz[j] * full_circle * (double)i) / rate, full_circle));
    }
    const float d = dither ? lsb : 0.0f;
    return (float)((double)d + (ws.U * sacc));
}

void write_samples(float *buf, size_t n, const double *ap, double attenuation, double f1, double rate,
                   Rng &rng)
{
    WheelSynth ws;
    wheel_synth(ws, ap, attenuation, f1, rate);
    for (size_t i = 0; i < n; ++i)
        buf[i] = synth_sample(ws, rate, (unsigned int)i, rng.next() < (kRandMax >> 1));
}

} // namespace

/* ------------------------------------------------------------------ */
/* envelopes, key compression, scanner tables                          */
/* ------------------------------------------------------------------ */
namespace {

/* noise burst + two-point average, src/tonegen.cpp:2577-2607 / :2650-2667 */
void click_envelope(float *env, int start, int burst, float level, Rng &rng)
{
    const int bss = TBF_FRAGMENT;
    int i = 0;
    for (; i < start; ++i)
        env[i] = 0.0f;
    for (; i < (start + burst); ++i)
        env[i] = (float)(1.0 - ((double)level * rng.drnd()));
    for (; i < bss; ++i)
        env[i] = 1.0f;
    env[0] = (float)((double)env[0] / 2.0);
    for (i = 1; i < bss; ++i)
        env[i] = (float)((double)(env[i - 1] + env[i]) / 2.0);
}

void shelf_envelope(float *env, int bound, Rng &rng)
{
    const int bss = TBF_FRAGMENT;
    int start = rng.next() % bound;
    if ((bss - 2) <= start)
        start = bss - 2;
    int i = 0;
    for (; i < start; ++i)
        env[i] = 0.0f;
    env[i + 0] = (float)0.33333333;
    env[i + 1] = (float)0.66666666;
    for (i = i + 2; i < bss; ++i)
        env[i] = 1.0f;
}

/* initEnvelopes, src/tonegen.cpp:2562-2728; the order of rand() draws per
 * bus is: attack click, attack shelf, release shelf, release click */
void init_envelopes(const tbf_tonegen_spec *s, int min_len, int max_len, tbf_tonegen_tables *o,
                    Rng &rng)
{
    const int bss = TBF_FRAGMENT;
    const double T = (double)(TBF_FRAGMENT - 1);
    for (int b = 0; b < 9; ++b) {
        float *att = o->attack_env[b];
        float *rel = o->release_env[b];
        int bound = max_len - min_len;
        if (bound < 1)
            bound = 1;
        if (s->attack_model == TBF_ENV_CLICK) {
            int burst = min_len + (rng.next() % bound);
            if (bss <= burst)
                burst = bss - 1;
            const int start = rng.next() % (bss - burst);
            click_envelope(att, start, burst, s->attack_click_level, rng);
        }
        if (s->attack_model == TBF_ENV_SHELF)
            shelf_envelope(att, bound, rng);
        if (s->release_model == TBF_ENV_SHELF)
            shelf_envelope(rel, bound, rng);
        if (s->release_model == TBF_ENV_CLICK) {
            const int burst = 8 + (rng.next() % 32);
            const int start = rng.next() % (bss - burst);
            click_envelope(rel, start, burst, s->release_click_level, rng);
        }
        if (s->attack_model == TBF_ENV_COSINE)
            for (int i = 0; i < bss; ++i) {
                const int d = bss - (i + 1);
                const double a = (M_PI * (double)d) / T;
                att[i] = (float)(0.5 + (0.5 * cos(a)));
            }
        if (s->release_model == TBF_ENV_COSINE)
            for (int i = 0; i < bss; ++i) {
                const double a = (M_PI * (double)i) / T;
                rel[i] = (float)(0.5 - (0.5 * cos(a)));
            }
        if (s->attack_model == TBF_ENV_LINEAR)
            for (int i = 0; i < bss; ++i)
                att[i] = ((float)i) / (float)bss;
        if (s->release_model == TBF_ENV_LINEAR)
            for (int i = 0; i < bss; ++i)
                rel[i] = ((float)i) / (float)bss;
    }
}

/* initKeyCompTable, src/tonegen.cpp:1939-1965 */
void init_key_comp(float *tab)
{
    const float u = -5.0f, v = -9.0f;
    const float m = (float)(1.0 / (128 - 12));
    tab[0] = tab[1] = 1.0f;
    const double db[9] = {-1.1598, -2.0291, -2.4987, -2.9952, -3.5218,
                          -4.0823, -4.6815, -4.9975, -4.9998};
    for (int i = 0; i < 9; ++i)
        tab[2 + i] = (float)db_to_gain(db[i]);
    for (int i = 11; i < 128; ++i) {
        const float a = (float)(i - 11);
        tab[i] = (float)db_to_gain((double)(u + ((v - u) * a * m)));
    }
}

} // namespace

static int build_tonegen(const tbf_tonegen_spec *s, tbf_tonegen_tables *o, tbf_wave_synth *synth);

extern "C" int tbf_build_tonegen_tables(const tbf_tonegen_spec *s, tbf_tonegen_tables *o)
{
    return build_tonegen(s, o, nullptr);
}

extern "C" int tbf_build_tonegen_tables_synth(const tbf_tonegen_spec *s, tbf_tonegen_tables *o,
                                              tbf_wave_synth *synth)
{
    if (!synth)
        return TBF_ERR_ARG;
    return build_tonegen(s, o, synth);
}

extern "C" float tbf_wave_sample(const tbf_wave_synth *y, int wheel, uint32_t i, int dither_bit)
{
    if (!y || wheel < 1 || wheel > TBF_NOF_WHEELS)
        return 0.0f;
    WheelSynth ws;
    wheel_synth(ws, y->harmonics, y->attenuation[wheel], y->f1[wheel], y->sample_rate);
    return synth_sample(ws, y->sample_rate, i, dither_bit != 0);
}

extern "C" void tbf_dither_bits(uint32_t seed, uint64_t n, uint32_t *bits)
{
    Rng rng(seed);
    for (uint64_t k = 0; k < n; k += 32) {
        uint32_t w = 0;
        for (uint64_t j = k; j < n && j < k + 32; ++j)
            if (rng.next() < (kRandMax >> 1))
                w |= 1u << (j - k);
        bits[k >> 5] = w;
    }
}

static int build_tonegen(const tbf_tonegen_spec *s, tbf_tonegen_tables *o, tbf_wave_synth *synth)
{
    if (!s || !o || !(s->sample_rate > 0))
        return TBF_ERR_ARG;
    memset(o, 0, sizeof(*o));
    const double rate = s->sample_rate;

    /* click burst lengths, src/tonegen.cpp:2940-2956 */
    int min_len = s->atk_clk_min_length, max_len = s->atk_clk_max_length;
    if (min_len < 0)
        min_len = (int)floor(rate * 8.0 / 22050.0);
    if (max_len < 0)
        max_len = (int)ceil(rate * 40.0 / 22050.0);
    if (min_len > TBF_FRAGMENT)
        min_len = TBF_FRAGMENT;
    if (max_len > TBF_FRAGMENT)
        max_len = TBF_FRAGMENT;

    /* TBF_BUILD_PROFILE=1: phase times of one build on stderr */
    static const bool prof = getenv("TBF_BUILD_PROFILE") != nullptr;
    auto t_prev = std::chrono::steady_clock::now();
    auto lap = [&](const char *what) {
        if (prof) {
            const auto t1 = std::chrono::steady_clock::now();
            fprintf(stderr, "tbf build %-22s %7.0f us\n", what,
                    std::chrono::duration<double, std::micro>(t1 - t_prev).count());
            t_prev = t1;
        }
    };
    double freq300[TBF_NOF_FREQS];
    tbf_extend_frequencies(s->freqs, freq300, NULL, NULL);
    lap("extend_frequencies");

    Wiring *w = new Wiring();
    int rc = default_configuration(s, freq300, *w);
    if (rc != TBF_OK) {
        delete w;
        return rc;
    }
    std::vector<uint32_t> start;
    std::vector<uint16_t> cw, cb;
    std::vector<float> cl;
    lap("default_configuration");
    compile_play_matrix(s, *w, start, cw, cb, cl);
    delete w;
    lap("compile_play_matrix");

    /* initOscillators, src/tonegen.cpp:1470-1630 */
    double att[TBF_NOF_WHEELS + 1];
    wheel_attenuation(s, att);
    double harmonics[TBF_MAX_PARTIALS] = {1.0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int i = 0; i < s->n_extra_harmonics && i < TBF_MAX_EXTRA_HARMONICS; ++i) {
        const int h = s->extra_harmonic_n[i] - 1;
        if (h < 0)
            return TBF_ERR_REFERENCE_UB; /* assert(0 <= h) */
        if (h < TBF_MAX_PARTIALS)
            harmonics[h] += s->extra_harmonic_level[i];
    }
    size_t len[TBF_NOF_WHEELS + 1];
    uint64_t total = 0;
    const int max_samples = (int)(ceil(rate / 48000.0) * 4096);
    for (int i = 1; i <= TBF_NOF_WHEELS; ++i) {
        rc = fit_wave(osc_frequency(freq300, i), s->precision, 3 * TBF_FRAGMENT, max_samples, rate,
                      &len[i]);
        if (rc != TBF_OK)
            return rc;
        o->wave_offset[i] = (uint32_t)total;
        o->wave_length[i] = (uint32_t)len[i];
        total += len[i];
    }
    lap("fit_wave x 256");
    float *wave = nullptr;
    Rng rng(s->seed);
    if (synth) {
        /* the device makes the samples; the rand() stream moves on as if they had been made */
        memset(synth, 0, sizeof(*synth));
        synth->sample_rate = rate;
        synth->seed = s->seed;
        memcpy(synth->harmonics, harmonics, sizeof(harmonics));
        for (int i = 1; i <= TBF_NOF_WHEELS; ++i) {
            synth->f1[i] = osc_frequency(freq300, i);
            synth->attenuation[i] = att[i];
        }
        rng.skip(total);
    } else {
        wave = (float *)malloc(sizeof(float) * (size_t)total);
        if (!wave)
            return TBF_ERR_NOMEM;
        for (int i = 1; i <= TBF_NOF_WHEELS; ++i)
            write_samples(wave + o->wave_offset[i], len[i], harmonics, att[i], osc_frequency(freq300, i),
                          rate, rng);
    }
    o->wave_data = wave;
    o->wave_total = total;
    lap("waves / rand() skip");

    const size_t nc = cw.size();
    uint16_t *pw = (uint16_t *)malloc(sizeof(uint16_t) * (nc ? nc : 1));
    uint16_t *pb = (uint16_t *)malloc(sizeof(uint16_t) * (nc ? nc : 1));
    float *pl = (float *)malloc(sizeof(float) * (nc ? nc : 1));
    if (!pw || !pb || !pl) {
        free(wave);
        free(pw);
        free(pb);
        free(pl);
        memset(o, 0, sizeof(*o));
        return TBF_ERR_NOMEM;
    }
    if (nc) {
        memcpy(pw, cw.data(), sizeof(uint16_t) * nc);
        memcpy(pb, cb.data(), sizeof(uint16_t) * nc);
        memcpy(pl, cl.data(), sizeof(float) * nc);
    }
    memcpy(o->contrib_start, start.data(), sizeof(o->contrib_start));
    o->contrib_wheel = pw;
    o->contrib_bus = pb;
    o->contrib_level = pl;

    init_key_comp(o->key_comp_table);
    init_envelopes(s, min_len, max_len, o, rng);
    lap("copies, envelopes");

    for (int i = 0; i < TBF_NOF_BUSES; ++i)
        for (int st = 0; st < 9; ++st)
            o->draw_bar_level[i][st] = (float)((double)(float)st / 8.0);

    o->perc_env_gain_reset_norm = s->perc_gain_norm;
    o->perc_env_gain_reset_soft = s->perc_gain_soft;
    o->perc_env_scaling = s->perc_gain_scaling;
    /* computePercResets() is never reached in the reference (only through the
     * uncalled setFast/SlowPercussionDecay), so the decay multipliers keep their
     * initValues constants (src/tonegen.cpp:260-263) */
    o->perc_decay_fast_norm = 0.9995f;
    o->perc_decay_fast_soft = 0.9995f;
    o->perc_decay_slow_norm = 0.9999f;
    o->perc_decay_slow_soft = 0.9999f;
    o->perc_drawbar_normal_gain = (float)0.60512;
    o->perc_drawbar_soft_gain = 1.0f;
    o->perc_send_bus_a = s->perc_bus_a;
    o->perc_send_bus_b = s->perc_bus_b;
    o->perc_trigger_bus = s->perc_bus_trig;
    o->output_level_trim = (float)0.07;

    /* init_vibrato, src/vibrato.cpp:95-100, :224-256 */
    o->vib_stator_increment =
        (unsigned int)(((s->scanner_hz * TBF_VIB_TBL) / rate) * 65536.0);
    for (int i = 0; i < TBF_VIB_TBL; ++i) {
        const double m = sin((2.0 * M_PI * i) / TBF_VIB_TBL);
        for (int k = 0; k < 3; ++k)
            o->vib_offset_table[k][i] =
                (unsigned int)((1.0 + s->scanner_amp[k] + (m * s->scanner_amp[k])) * 65536.0);
    }

    /* run-time scalars after initToneGenerator's trailing setter calls
     * (src/tonegen.cpp:3006-3021): drawbars 8,8,6 / 8,3,8 / 8,-,6 through
     * setMIDIDrawBar (rint((127-v)*8/127)), percussion second/normal/fast/off */
    const int dbus[8] = {0, 1, 2, 9, 10, 11, 18, 20};
    const int dval[8] = {8, 8, 6, 8, 3, 8, 8, 6};
    int trig_restore = 0;
    for (int i = 0; i < 8; ++i) {
        const int setting = (int)rint((127 - dval[i]) * 8.0 / 127.0);
        if (dbus[i] == s->perc_bus_trig)
            trig_restore = setting; /* percussion is off: the gain is still installed */
        o->init_draw_bar_gain[dbus[i]] = o->draw_bar_level[dbus[i]][setting];
    }
    o->init_draw_bar_change = 1;
    o->init_perc_trig_restore = trig_restore;
    o->init_perc_send_bus = s->perc_bus_b; /* setPercussionFirst(FALSE) */
    o->init_perc_is_soft = 0;
    o->init_perc_env_gain_reset = s->perc_gain_scaling * s->perc_gain_norm;
    o->init_perc_drawbar_gain = o->perc_drawbar_normal_gain;
    o->init_perc_is_fast = 1;
    o->init_perc_env_gain_decay = o->perc_decay_fast_norm;
    o->init_perc_enabled = 0;
    /* setPercussionEnabled(FALSE) re-installs the trigger bus gain */
    if (-1 < s->perc_bus_trig)
        o->init_draw_bar_gain[s->perc_bus_trig] = o->draw_bar_level[s->perc_bus_trig][trig_restore];
    o->init_swell_pedal_gain = (float)0.07;
    o->init_routing = 0;
    o->init_vib_table = 2; /* reset_vibrato: offset3Table; setVibrato(v, 0) leaves it */
    o->init_vib_mixed = 0;
    return TBF_OK;
}

extern "C" void tbf_free_tonegen_tables(tbf_tonegen_tables *t)
{
    if (!t)
        return;
    free((void *)t->wave_data);
    free((void *)t->contrib_wheel);
    free((void *)t->contrib_bus);
    free((void *)t->contrib_level);
    memset(t, 0, sizeof(*t));
}

/* ------------------------------------------------------------------ */
/* whirl tables                                                        */
/* ------------------------------------------------------------------ */
namespace {

/* RBJ audio-EQ-cookbook biquad, evaluated with the reference's expression
 * grouping (eqCompute, src/eqcomp.cpp:109-203).  out = a1 a2 b0 b1 b2 / a0. */
void eq_compute(int type, double hz, double Q, double dbG, double rate, double *o_a1, double *o_a2,
                double *o_b0, double *o_b1, double *o_b2)
{
    const double A = pow(10.0, (dbG / 40.0));
    const double omega = (2.0 * M_PI * hz) / rate;
    const double sn = sin(omega);
    const double cs = cos(omega);
    const double alpha = sn / (2.0 * Q);
    const double beta = sqrt(A) / Q;
    double b0 = 0, b1 = 0, b2 = 0, a0 = 1, a1 = 0, a2 = 0;
    switch (type) {
    case 0: /* low pass */
        b0 = (1.0 - cs) / 2.0;
        b1 = 1.0 - cs;
        b2 = (1.0 - cs) / 2.0;
        a0 = 1.0 + alpha;
        a1 = -2.0 * cs;
        a2 = 1.0 - alpha;
        break;
    case 1: /* high pass */
        b0 = (1.0 + cs) / 2.0;
        b1 = -(1.0 + cs);
        b2 = (1.0 + cs) / 2.0;
        a0 = 1.0 + alpha;
        a1 = -2.0 * cs;
        a2 = 1.0 - alpha;
        break;
    case 2: /* band pass, constant skirt gain */
        b0 = sn / 2.0;
        b1 = 0.0;
        b2 = -sn / 2.0;
        a0 = 1.0 + alpha;
        a1 = -2.0 * cs;
        a2 = 1.0 - alpha;
        break;
    case 3: /* band pass, 0 dB peak */
        b0 = alpha;
        b1 = 0.0;
        b2 = -alpha;
        a0 = 1.0 + alpha;
        a1 = -2.0 * cs;
        a2 = 1.0 - alpha;
        break;
    case 4: /* notch */
        b0 = 1.0;
        b1 = -2.0 * cs;
        b2 = 1.0;
        a0 = 1.0 + alpha;
        a1 = -2.0 * cs;
        a2 = 1.0 - alpha;
        break;
    case 5: /* all pass */
        b0 = 1.0 - alpha;
        b1 = -2.0 * cs;
        b2 = 1.0 + alpha;
        a0 = 1.0 + alpha;
        a1 = -2.0 * cs;
        a2 = 1.0 - alpha;
        break;
    case 6: /* peaking */
        b0 = 1.0 + (alpha * A);
        b1 = -2.0 * cs;
        b2 = 1.0 - (alpha * A);
        a0 = 1.0 + (alpha / A);
        a1 = -2.0 * cs;
        a2 = 1.0 - (alpha / A);
        break;
    case 7: /* low shelf */
        b0 = A * ((A + 1) - ((A - 1) * cs) + (beta * sn));
        b1 = (2.0 * A) * ((A - 1) - ((A + 1) * cs));
        b2 = A * ((A + 1) - ((A - 1) * cs) - (beta * sn));
        a0 = (A + 1) + ((A - 1) * cs) + (beta * sn);
        a1 = -2.0 * ((A - 1) + ((A + 1) * cs));
        a2 = (A + 1) + ((A - 1) * cs) - (beta * sn);
        break;
    case 8: /* high shelf */
        b0 = A * ((A + 1) + ((A - 1) * cs) + (beta * sn));
        b1 = -(2.0 * A) * ((A - 1) + ((A + 1) * cs));
        b2 = A * ((A + 1) + ((A - 1) * cs) - (beta * sn));
        a0 = (A + 1) - ((A - 1) * cs) + (beta * sn);
        a1 = 2.0 * ((A - 1) - ((A + 1) * cs));
        a2 = (A + 1) - ((A - 1) * cs) - (beta * sn);
        break;
    }
    *o_b0 = b0 / a0;
    *o_b1 = b1 / a0;
    *o_b2 = b2 / a0;
    *o_a1 = a1 / a0;
    *o_a2 = a2 / a0;
}

/* setIIRFilter, src/whirl.cpp:147-172: out-of-range parameters leave the
 * (zeroed) coefficients untouched -- whirl-sharp.cfg relies on it */
bool set_iir(float *c5, int type, double F, double Q, double G, double rate)
{
    if (Q <= 0.1 || Q >= 6.00)
        return false;
    if (F / rate <= 0.0002 || F / rate >= 0.4998)
        return false;
    if (G <= -48.0 || G >= 48.0)
        return false;
    if (type < 0 || type > 8)
        return false;
    double a1, a2, b0, b1, b2;
    eq_compute(type, F, Q, G, rate, &a1, &a2, &b0, &b1, &b2);
    c5[0] = (float)a1;
    c5[1] = (float)a2;
    c5[2] = (float)b0;
    c5[3] = (float)b1;
    c5[4] = (float)b2;
    return true;
}

/* Angle-dependent horn impulse response: five coefficient curves over
 * -180..180 degrees given as break points (read off fig. 8 of Smith, Serafin,
 * Abel, Berners, "Doppler simulation and the Leslie", DAFx-02), linearly
 * interpolated onto 16384 angle steps (initTables, src/whirl.cpp:338-517). */
struct Pt {
    double deg, level;
};
const Pt kIr0[] = {{-180.0, 1.052}, {-166.4, .881}, {-150.5, .881}, {-135.3, .881}, {-122.4, .792},
                   {-106.5, .792}, {-91.2, .836},  {-75.8, .881},  {-59.4, .851},  {-44.7, .941},
                   {-30.0, 1.298}, {-14.7, 2.119}, {0.0, 2.820},   {15.6, 2.313},  {30.0, 1.492},
                   {44.7, .926},   {60.0, .836},   {74.7, .866},   {90.6, .792},   {100.0, .777},
                   {105.0, .777},  {120.0, .836},  {135.3, .836},  {150.0, .881},  {164.5, .874},
                   {180.0, 1.052}};
const Pt kIr1[] = {{-180.0, -0.07}, {-150.0, 0.10}, {-135.0, -0.10}, {-122.2, 0.16}, {-105.0, 0.15},
                   {-91.2, 0.37},  {-75.3, 0.32},  {-60.1, 0.39},   {-44.5, 0.70},  {-30.0, 0.53},
                   {-12.0, -0.40}, {0.0, -0.81},   {2.7, -0.77},    {15.0, -0.52},  {33.1, 0.38},
                   {43.7, 0.68},   {57.7, 0.49},   {74.1, 0.19},    {89.4, 0.33},   {105.0, 0.03},
                   {120.0, 0.12},  {134.0, -0.13}, {153.3, 0.08},   {180.0, -0.07}};
const Pt kIr2[] = {{-180.0, 0.40}, {-165.0, 0.20}, {-150.0, 0.48}, {-135.0, 0.27}, {-121.2, 0.22},
                   {-89.2, 0.30},  {-69.2, 0.22},  {-58.0, 0.11},  {-40.2, -0.43}, {-29.0, -0.53},
                   {-15.6, -0.43}, {0.0, 0.00},    {14.3, -0.44},  {30.3, -0.60},  {60.3, 0.11},
                   {74.9, 0.32},   {91.5, 0.23},   {104.9, 0.32},  {121.7, 0.19},  {135.0, 0.27},
                   {150.0, 0.45},  {165.0, 0.20},  {180.0, 0.40}};
const Pt kIr3[] = {{-180.0, -0.08}, {-165.2, -0.19}, {-150.0, 0.00}, {-133.9, -0.20}, {-120.0, -0.15},
                   {-106.0, 0.09},  {-89.3, -0.15},  {-76.3, 0.00},  {-60.3, 0.29},   {-44.6, -0.02},
                   {-15.6, -0.22},  {0.0, 0.24},     {14.5, 0.11},   {30.1, -0.10},   {44.6, 0.17},
                   {60.4, 0.22},    {75.9, 0.16},    {90.4, -0.05},  {104.9, 0.07},   {122.8, -0.07},
                   {136.2, -0.07},  {150.0, 0.08},   {165.0, -0.19}, {180.0, -0.08}};
const Pt kIr4[] = {{-180.0, 0.13}, {-165.2, 0.00}, {-150.0, 0.17}, {-135.2, -0.20}, {-120.5, 0.00},
                   {-105.0, 0.00}, {-90.0, 0.04},  {-75.0, -0.09}, {-60.3, -0.14},  {-45.0, 0.16},
                   {-15.6, 0.00},  {0.0, 0.22},    {15.6, -0.21},  {30.1, -0.09},   {45.0, 0.10},
                   {60.3, -0.07},  {74.8, -0.15},  {90.4, -0.03},  {104.9, -0.14},  {120.5, 0.00},
                   {135.2, -0.26}, {150.0, 0.16},  {165.0, -0.02}, {180.0, 0.13}};

void draw_curve(float *bfw, int partial, const Pt *pts, int n)
{
    const double N = (double)TBF_WHIRL_DISPLC;
    double ipx = pts[0].deg, ipy = pts[0].level;
    for (int p = 1; p < n; ++p) {
        double d = ipx;
        while (d < 0.0)
            d += 360.0;
        const int from = (int)((d * N) / 360.0);
        ipx = pts[p].deg;
        double e = ipx;
        while (e < d)
            e += 360.0;
        const int to = (int)((e * N) / 360.0);
        const double range = (double)(to - from);
        for (int i = from; i <= to; ++i) {
            const double x = (double)(i - from);
            const double v = ipy + ((x / range) * (pts[p].level - ipy));
            bfw[5 * (i & (TBF_WHIRL_DISPLC - 1)) + partial] = (float)v;
        }
        ipy = pts[p].level;
    }
}

} // namespace

extern "C" int tbf_iir_coefficients(float c5[5], int type, double hz, double q, double gain_db,
                                    double sample_rate)
{
    if (!c5)
        return TBF_ERR_ARG;
    return set_iir(c5, type, hz, q, gain_db, sample_rate) ? 1 : 0;
}

extern "C" int tbf_build_whirl_tables(const tbf_whirl_spec *s, tbf_whirl_tables *o)
{
    if (!s || !o || !(s->sample_rate > 0))
        return TBF_ERR_ARG;
    memset(o, 0, sizeof(*o));
    const unsigned N = TBF_WHIRL_DISPLC;
    const double rate = s->sample_rate;
    float *tab = (float *)calloc((size_t)N * 14, sizeof(float));
    if (!tab)
        return TBF_ERR_NOMEM;
    float *hn_fwd = tab, *hn_bwd = tab + N, *dr_fwd = tab + 2 * N, *dr_bwd = tab + 3 * N;
    float *bfw = tab + 4 * N, *bbw = tab + 9 * N;
    o->hn_fwd_displ = hn_fwd;
    o->hn_bwd_displ = hn_bwd;
    o->dr_fwd_displ = dr_fwd;
    o->dr_bwd_displ = dr_bwd;
    o->bfw = bfw;
    o->bbw = bbw;
    o->sample_rate = rate;

    /* initialize(), src/whirl.cpp:626-662 */
    o->leakage = s->leak_level * s->horn_level;
    set_iir(o->drfL, s->lp_type, s->lp_f, s->lp_q, s->lp_g, rate);
    set_iir(o->drfR, s->lp_type, s->lp_f, s->lp_q, s->lp_g, rate);
    set_iir(o->hafw, s->ha_type, s->ha_f, s->ha_q, s->ha_g, rate);
    set_iir(o->hbfw, s->hb_type, s->hb_f, s->hb_q, s->hb_g, rate);
    const float hap[4] = {(float)s->ha_type, s->ha_f, s->ha_q, s->ha_g};
    const float hbp[4] = {(float)s->hb_type, s->hb_f, s->hb_q, s->hb_g};
    memcpy(o->ha_param, hap, sizeof hap);
    memcpy(o->hb_param, hbp, sizeof hbp);

    /* computeOffsets(), src/whirl.cpp:534-624 */
    const float hs0[6] = {12.0, 18.0, 53.0, 50.0, 106.0, 116.0};
    const float ds0[6] = {36.0, 39.0, 79.0, 86.0, 123.0, 116.0};
    const double horn_radius = (s->horn_radius_cm * rate / 100.0) / s->air_speed;
    const double drum_radius = (s->drum_radius_cm * rate / 100.0) / s->air_speed;
    const double mic_dist = (s->mic_dist_cm * rate / 100.0) / s->air_speed;
    const double mic_x = (s->horn_x_offset_cm * rate / 100.0) / s->air_speed;
    const double mic_z = (s->horn_z_offset_cm * rate / 100.0) / s->air_speed;
    double maxhn = 0, maxdr = 0;
    for (unsigned i = 0; i < N; ++i) {
        const double v = (2.0 * M_PI * (double)i) / (double)N;
        double a = mic_dist - (horn_radius * cos(v));
        double b = mic_z + horn_radius * sin(v);
        const double dist = sqrt((a * a) + (b * b));
        hn_fwd[i] = (float)(dist + mic_x);
        hn_bwd[N - (i + 1)] = (float)(dist - mic_x);
        if (maxhn < hn_fwd[i])
            maxhn = hn_fwd[i];
        if (maxhn < hn_bwd[N - (i + 1)])
            maxhn = hn_bwd[N - (i + 1)];
        a = mic_dist - (drum_radius * cos(v));
        b = drum_radius * sin(v);
        dr_fwd[i] = (float)sqrt((a * a) + (b * b));
        dr_bwd[N - (i + 1)] = dr_fwd[i];
        if (maxdr < dr_fwd[i])
            maxdr = dr_fwd[i];
    }
    const int phase[6] = {0, (int)(N >> 1), (int)((N * 2) / 6), (int)((N * 5) / 6),
                          (int)((N * 1) / 6), (int)((N * 4) / 6)};
    for (int i = 0; i < 6; ++i) {
        o->horn_phase[i] = phase[i];
        o->drum_phase[i] = phase[i];
        o->horn_spacing[i] = (float)(hs0[i] * rate / 22100.0 + horn_radius + 1.0);
        o->drum_spacing[i] = (float)(ds0[i] * rate / 22100.0 + drum_radius + 1.0);
        if (!(maxhn + o->horn_spacing[i] < TBF_WHIRL_BUF) ||
            !(maxdr + o->drum_spacing[i] < TBF_WHIRL_BUF)) {
            free(tab);
            memset(o, 0, sizeof(*o));
            return TBF_ERR_GEOMETRY; /* reference asserts, src/whirl.cpp:607, :622 */
        }
    }

    /* initTables(), src/whirl.cpp:338-517 */
    draw_curve(bfw, 0, kIr0, (int)(sizeof(kIr0) / sizeof(Pt)));
    draw_curve(bfw, 1, kIr1, (int)(sizeof(kIr1) / sizeof(Pt)));
    draw_curve(bfw, 2, kIr2, (int)(sizeof(kIr2) / sizeof(Pt)));
    draw_curve(bfw, 3, kIr3, (int)(sizeof(kIr3) / sizeof(Pt)));
    draw_curve(bfw, 4, kIr4, (int)(sizeof(kIr4) / sizeof(Pt)));
    double sum = 0.0;
    for (unsigned i = 0; i < N; ++i) {
        double colsum = 0.0;
        for (int j = 0; j < 5; ++j)
            colsum += fabs(bfw[5 * i + j]);
        if (sum < colsum)
            sum = colsum;
    }
    for (unsigned i = 0; i < N; ++i)
        for (int j = 0; j < 5; ++j) {
            bfw[5 * i + j] = (float)((double)bfw[5 * i + j] * (1.0 / sum));
            bbw[5 * (N - i - 1) + j] = bfw[5 * i + j];
        }

    /* computeRotationSpeeds(), src/whirl.cpp:270-293 */
    const double hfast = s->horn_rpm_fast / (rate * 60.0);
    const double hslow = s->horn_rpm_slow / (rate * 60.0);
    const double dfast = s->drum_rpm_fast / (rate * 60.0);
    const double dslow = s->drum_rpm_slow / (rate * 60.0);
    const double hsel[3] = {0, hslow, hfast}, dsel[3] = {0, dslow, dfast};
    for (int i = 0; i < 9; ++i) {
        o->rev_horn_target[i] = hsel[i / 3];
        o->rev_drum_target[i] = dsel[i % 3];
    }
    o->revselects[0] = 4;
    o->revselects[1] = 0;
    o->revselects[2] = 8;

    o->mic_angle = s->mic_angle;
    o->horn_acc = s->horn_acc;
    o->horn_dec = s->horn_dec;
    o->drum_acc = s->drum_acc;
    o->drum_dec = s->drum_dec;
    o->hn_brake_pos = s->hn_brake_pos;
    o->dr_brake_pos = s->dr_brake_pos;
    o->horn_level = s->horn_level;
    /* fsetDrumMicWidth / fsetHornMicWidth, src/whirl.cpp:917-951 */
    {
        const float dw = s->drum_mic_width;
        const float dwP = dw > 0.f ? (dw > 1.f ? 1.f : dw) : 0.f;
        const float dwN = dw < 0.f ? (dw < -1.f ? 1.f : -dw) : 0.f;
        o->mic_dll = sqrtf(1.f - dwP);
        o->mic_dlr = sqrtf(0.f + dwP);
        o->mic_drl = sqrtf(0.f + dwN);
        o->mic_drr = sqrtf(1.f - dwN);
        const float hw = s->horn_mic_width;
        const float hwP = hw > 0.f ? (hw > 1.f ? 1.f : hw) : 0.f;
        const float hwN = hw < 0.f ? (hw < -1.f ? 1.f : -hw) : 0.f;
        o->mic_hll = sqrtf(1.f - hwP);
        o->mic_hlr = sqrtf(0.f + hwP);
        o->mic_hrl = sqrtf(0.f + hwN);
        o->mic_hrr = sqrtf(1.f - hwN);
    }
    o->bypass = s->bypass;

    /* state after setRevSelect(w, revSelect) at the end of computeRotationSpeeds:
     * targets of the selected preset, acceleration flags against incr = 0 */
    const int sel = ((s->rev_select % 3) + 3) % 3;
    const int opt = o->revselects[sel];
    o->init_rev_select = sel;
    o->init_horn_target = o->rev_horn_target[opt];
    o->init_drum_target = o->rev_drum_target[opt];
    o->init_horn_incr = 0;
    o->init_drum_incr = 0;
    o->init_horn_angle = 0;
    o->init_drum_angle = 0;
    o->init_horn_acdc = (0 < o->init_horn_target) ? 1 : (o->init_horn_target < 0 ? -1 : 0);
    o->init_drum_acdc = (0 < o->init_drum_target) ? 1 : (o->init_drum_target < 0 ? -1 : 0);
    return TBF_OK;
}

extern "C" void tbf_free_whirl_tables(tbf_whirl_tables *t)
{
    if (!t)
        return;
    free((void *)t->hn_fwd_displ); /* one block: see tbf_build_whirl_tables */
    memset(t, 0, sizeof(*t));
}

/* ------------------------------------------------------------------ */
/* cfg "key=value" lines of the osc.*, scanner.* and whirl.* families   */
/* (oscConfig src/tonegen.cpp:2173-2556, scannerConfig src/vibrato.cpp: */
/* 334-357, whirlConfig src/whirl.cpp:992-1164).  Range checks and the  */
/* float/double target types are the reference's.                       */
/* ------------------------------------------------------------------ */
namespace {

bool parse_double(const char *v, double *out)
{
    return v && sscanf(v, "%lf", out) == 1;
}
bool parse_int(const char *v, int *out) { return v && sscanf(v, "%d", out) == 1; }

int env_model(const char *v)
{
    if (!strcasecmp(v, "click"))
        return TBF_ENV_CLICK;
    if (!strcasecmp(v, "cosine"))
        return TBF_ENV_COSINE;
    if (!strcasecmp(v, "linear"))
        return TBF_ENV_LINEAR;
    if (!strcasecmp(v, "shelf"))
        return TBF_ENV_SHELF;
    return -1;
}

} // namespace

extern "C" int tbf_spec_apply_cfg(tbf_tonegen_spec *tg, tbf_whirl_spec *wh, const char *key,
                                  const char *value)
{
    if (!key || !value)
        return TBF_ERR_ARG;
    double d = 0;
    int k = 0;
    const bool has_d = parse_double(value, &d);
    const bool has_i = parse_int(value, &k);
#define KEY(name) (!strcasecmp(key, name))
#define NEED_D()        \
    if (!has_d)         \
        return TBF_ERR_ARG;
#define RANGE_D(lo, hi)         \
    NEED_D()                    \
    if (!((lo) <= d && d <= (hi))) \
        return TBF_ERR_ARG;
#define RANGE_I(lo, hi)                  \
    if (!has_i || k < (lo) || k > (hi)) \
        return TBF_ERR_ARG;

    if (tg) {
        if (KEY("osc.tuning") || KEY("osc.temperament") || !strncasecmp(key, "osc.eqv.", 8))
            return 1; /* parsed by the reference but dead */
        if (KEY("osc.x-precision")) { NEED_D() if (0.0 < d) tg->precision = d; return 1; }
        if (KEY("osc.perc.fast") || KEY("osc.perc.slow"))
            return 1; /* stored, never turned into decay constants by the reference */
        if (KEY("osc.perc.normal")) { NEED_D() tg->perc_gain_norm = (float)d; return 1; }
        if (KEY("osc.perc.soft")) { NEED_D() tg->perc_gain_soft = (float)d; return 1; }
        if (KEY("osc.perc.gain")) { NEED_D() tg->perc_gain_scaling = (float)d; return 1; }
        if (KEY("osc.perc.bus.a")) { RANGE_I(0, 8) tg->perc_bus_a = k; return 1; }
        if (KEY("osc.perc.bus.b")) { RANGE_I(0, 8) tg->perc_bus_b = k; return 1; }
        if (KEY("osc.perc.bus.trig")) { RANGE_I(-1, 8) tg->perc_bus_trig = k; return 1; }
        if (KEY("osc.eq.macro")) {
            if (!strcasecmp(value, "chspline")) tg->eq_macro = TBF_EQ_SPLINE;
            else if (!strcasecmp(value, "peak24")) tg->eq_macro = TBF_EQ_PEAK24;
            else if (!strcasecmp(value, "peak46")) tg->eq_macro = TBF_EQ_PEAK46;
            else return TBF_ERR_ARG;
            return 1;
        }
        if (KEY("osc.eq.p1y")) { NEED_D() tg->eq_p1y = d; return 1; }
        if (KEY("osc.eq.r1y")) { NEED_D() tg->eq_r1y = d; return 1; }
        if (KEY("osc.eq.p4y")) { NEED_D() tg->eq_p4y = d; return 1; }
        if (KEY("osc.eq.r4y")) { NEED_D() tg->eq_r4y = d; return 1; }
        if (!strncasecmp(key, "osc.harmonic.", 13)) {
            int n;
            if (sscanf(key, "osc.harmonic.%d", &n) == 1) {
                NEED_D()
                if (tg->n_extra_harmonics >= TBF_MAX_EXTRA_HARMONICS)
                    return TBF_ERR_NOMEM;
                tg->extra_harmonic_n[tg->n_extra_harmonics] = (short)n;
                tg->extra_harmonic_level[tg->n_extra_harmonics] = (float)d;
                tg->n_extra_harmonics++;
                return 1;
            }
            return TBF_ERR_ARG; /* per-wheel lists (osc.harmonic.w<w>.f<h>) not supported yet */
        }
        if (!strncasecmp(key, "osc.terminal.", 13) || !strncasecmp(key, "osc.taper.", 10) ||
            !strncasecmp(key, "osc.crosstalk.", 14))
            return TBF_ERR_ARG; /* per-terminal / per-key wiring overrides not supported yet */
        if (KEY("osc.compartment-crosstalk")) { RANGE_D(0.0, 1.0) tg->compartment_crosstalk = d; return 1; }
        if (KEY("osc.transformer-crosstalk")) { RANGE_D(0.0, 1.0) tg->transformer_crosstalk = d; return 1; }
        if (KEY("osc.terminalstrip-crosstalk")) { RANGE_D(0.0, 1.0) tg->terminalstrip_crosstalk = d; return 1; }
        if (KEY("osc.wiring-crosstalk")) { RANGE_D(0.0, 1.0) tg->wiring_crosstalk = d; return 1; }
        if (KEY("osc.contribution-floor")) { RANGE_D(0.0, 1.0) tg->contribution_floor = d; return 1; }
        if (KEY("osc.contribution-min")) { RANGE_D(0.0, 1.0) tg->contribution_min = d; return 1; }
        if (KEY("osc.attack.click.level")) { RANGE_D(0.0, 1.0) tg->attack_click_level = (float)d; return 1; }
        if (KEY("osc.release.click.level")) { RANGE_D(0.0, 1.0) tg->release_click_level = (float)d; return 1; }
        if (KEY("osc.attack.click.maxlength")) { RANGE_D(0.0, 1.0) tg->atk_clk_max_length = (int)(128.0 * d); return 1; }
        if (KEY("osc.attack.click.minlength")) { RANGE_D(0.0, 1.0) tg->atk_clk_min_length = (int)(128.0 * d); return 1; }
        if (KEY("osc.attack.model") || KEY("osc.release.model")) {
            const int m = env_model(value);
            if (m >= 0) {
                if (KEY("osc.attack.model")) tg->attack_model = m;
                else tg->release_model = m;
            }
            return 1;
        }
        if (KEY("scanner.hz")) { RANGE_D(4.0, 22.0) tg->scanner_hz = d; return 1; }
        if (KEY("scanner.modulation.v1")) { RANGE_D(0.0, 12.0) tg->scanner_amp[0] = d; return 1; }
        if (KEY("scanner.modulation.v2")) { RANGE_D(0.0, 12.0) tg->scanner_amp[1] = d; return 1; }
        if (KEY("scanner.modulation.v3")) { RANGE_D(0.0, 12.0) tg->scanner_amp[2] = d; return 1; }
    }
    if (wh) {
        if (KEY("whirl.horn.slowrpm")) { NEED_D() wh->horn_rpm_slow = (float)d; return 1; }
        if (KEY("whirl.horn.fastrpm")) { NEED_D() wh->horn_rpm_fast = (float)d; return 1; }
        if (KEY("whirl.horn.acceleration")) { NEED_D() wh->horn_acc = (float)d; return 1; }
        if (KEY("whirl.horn.deceleration")) { NEED_D() wh->horn_dec = (float)d; return 1; }
        if (KEY("whirl.drum.slowrpm")) { NEED_D() wh->drum_rpm_slow = (float)d; return 1; }
        if (KEY("whirl.drum.fastrpm")) { NEED_D() wh->drum_rpm_fast = (float)d; return 1; }
        if (KEY("whirl.drum.acceleration")) { NEED_D() wh->drum_acc = (float)d; return 1; }
        if (KEY("whirl.drum.deceleration")) { NEED_D() wh->drum_dec = (float)d; return 1; }
        if (KEY("whirl.horn.radius")) { NEED_D() wh->horn_radius_cm = (float)d; return 1; }
        if (KEY("whirl.drum.radius")) { NEED_D() wh->drum_radius_cm = (float)d; return 1; }
        if (KEY("whirl.horn.level")) { NEED_D() wh->horn_level = (float)d; return 1; }
        if (KEY("whirl.horn.leak")) { NEED_D() wh->leak_level = (float)d; return 1; }
        if (KEY("whirl.drum.width")) { NEED_D() wh->drum_mic_width = (float)d; return 1; }
        if (KEY("whirl.horn.width")) { NEED_D() wh->horn_mic_width = (float)d; return 1; }
        if (KEY("whirl.mic.distance")) { NEED_D() wh->mic_dist_cm = (float)d; return 1; }
        if (KEY("whirl.horn.offset.x")) { NEED_D() wh->horn_x_offset_cm = (float)d; return 1; }
        if (KEY("whirl.horn.offset.z")) { NEED_D() wh->horn_z_offset_cm = (float)d; return 1; }
        if (KEY("whirl.drum.filter.type")) { RANGE_I(0, 8) wh->lp_type = k; return 1; }
        if (KEY("whirl.drum.filter.q")) { NEED_D() wh->lp_q = d; return 1; }
        if (KEY("whirl.drum.filter.hz")) { NEED_D() wh->lp_f = d; return 1; }
        if (KEY("whirl.drum.filter.gain")) { NEED_D() wh->lp_g = d; return 1; }
        if (KEY("whirl.horn.filter.a.type")) { RANGE_I(0, 8) wh->ha_type = k; return 1; }
        if (KEY("whirl.horn.filter.a.hz")) { NEED_D() wh->ha_f = (float)d; return 1; }
        if (KEY("whirl.horn.filter.a.q")) { NEED_D() wh->ha_q = (float)d; return 1; }
        if (KEY("whirl.horn.filter.a.gain")) { NEED_D() wh->ha_g = (float)d; return 1; }
        if (KEY("whirl.horn.filter.b.type")) { RANGE_I(0, 8) wh->hb_type = k; return 1; }
        if (KEY("whirl.horn.filter.b.hz")) { NEED_D() wh->hb_f = (float)d; return 1; }
        if (KEY("whirl.horn.filter.b.q")) { NEED_D() wh->hb_q = (float)d; return 1; }
        if (KEY("whirl.horn.filter.b.gain")) { NEED_D() wh->hb_g = (float)d; return 1; }
        if (KEY("whirl.horn.comb.a.feedback") || KEY("whirl.horn.comb.a.delay") ||
            KEY("whirl.horn.comb.b.feedback") || KEY("whirl.horn.comb.b.delay"))
            return 1; /* parsed, but HORN_COMB_FILTER is compiled out (src/whirl.cpp:25) */
        if (KEY("whirl.speed-preset")) { if (!has_i) return TBF_ERR_ARG; wh->rev_select = k % 3; return 1; }
        if (KEY("whirl.bypass")) { RANGE_I(0, 1) wh->bypass = k; return 1; }
        if (KEY("whirl.horn.mic.angle")) { RANGE_D(0, 180.0) wh->mic_angle = 1.0 - d / 180.0; return 1; }
        if (KEY("whirl.horn.brakepos") || KEY("whirl.horn.breakpos")) { RANGE_D(0, 1.0) wh->hn_brake_pos = d; return 1; }
        if (KEY("whirl.drum.brakepos") || KEY("whirl.drum.breakpos")) { RANGE_D(0, 1.0) wh->dr_brake_pos = d; return 1; }
    }
#undef KEY
#undef NEED_D
#undef RANGE_D
#undef RANGE_I
    return 0;
}

extern "C" unsigned tbf_sizeof_tonegen_spec(void) { return (unsigned)sizeof(tbf_tonegen_spec); }
extern "C" unsigned tbf_sizeof_whirl_spec(void) { return (unsigned)sizeof(tbf_whirl_spec); }
