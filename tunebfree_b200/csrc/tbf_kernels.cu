/* tbf_kernels.cu -- sm_100a kernels for the tonewheel -> scanner -> Leslie path.
 *
 * The per-fragment chain oscGenerateFragment -> vibratoProc -> whirlProc2/3 is a feed-
 * forward pipeline whose recurrences all stay inside one stage, so it runs as four
 * kernels per chunk of fragments, each with the thread layout its stage wants:
 *
 *   stage T (warp = instance; lane = 4 consecutive samples during the wheel sum, then
 *            lane = sample of a 32-sample block)
 *       control events, active-list pass into per-wheel records kept in shared memory across
 *       quiet fragments, additive wheel sum (16-byte cp.async per lane and wheel from the
 *       aligned wave store, packed FP32), scanner vibrato (ordered two-point scatter),
 *       percussion high-pass                              -> u, php, per-fragment FragCtl
 *   stage R (lane = instance: 32 organs advance in lockstep in one warp)
 *       percussion decay / key-compression chase, mixdown, horn biquads A and B, rotor
 *       speed ramp + angle recurrence in double            -> dx0, hx0, hang, dang
 *   stage W (warp = instance, lane = sample of a 32-sample block)
 *       reflection filters (double islands), tap geometry, 5-tap FIRs, ordered scatter
 *       into the four delay lines held in shared memory    -> raw reads of HL HR DL DR
 *   stage D (lane = instance for the drum biquads, lane = sample for the mix)
 *       drum biquads, horn level + leak, whirlProc / whirlProc3 mix -> out_l, out_r
 *
 * The streams between stages live in HBM as [slot][sample]; lane = instance stages
 * transpose 32 x 32 tiles through shared memory so every global access is coalesced.
 *
 * Numerics contract: bit-exact float/double arithmetic in the reference's
 * source operation order.  This file is compiled with -fmad=false -ftz=false
 * -prec-div=true -prec-sqrt=true; no FMA contraction, no re-association, denormals kept.
 *
 * Reference functions restated here (never copied):
 *   oscKeyOn/oscKeyOff          src/tonegen.cpp:3096-3166
 *   oscGenerateFragment         src/tonegen.cpp:3218-3778
 *   vibratoProc                 src/vibrato.cpp:365-411
 *   whirlProc2 / whirlProc3     src/whirl.cpp:1191-1681
 *   setters                     src/tonegen.cpp:1635-1765, :2738-2750; src/vibrato.cpp:95-142;
 *                               src/whirl.cpp:174-236
 */
#include <cuda_pipeline.h>
#include <cuda.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "tbf_device.h"

#define DISPLC_MASK (TBF_WHIRL_DISPLC - 1)
#define WBUF_MASK (TBF_WHIRL_BUF - 1)
#define VBUF_MASK (TBF_VIB_BUF - 1)
#define VRING_MASK (TBF_VRING - 1)
#define FULL 0xFFFFFFFFu
#define BLK 32 /* samples per scatter block = one per lane */
#define TILE_PITCH 33

#define T_WARPS 4
#define W_WARPS 4

namespace {

__device__ __forceinline__ unsigned long long mix64(unsigned long long x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

/* position-keyed additive hash of one stereo sample (order independent) */
__device__ __forceinline__ unsigned long long sample_hash(uint32_t pos, float l, float r)
{
    return mix64(((unsigned long long)pos << 32) | __float_as_uint(l)) +
           mix64((((unsigned long long)pos << 32) | __float_as_uint(r)) ^ 0x5555555555555555ull);
}

/* x86 semantics of `(unsigned int)f` for the float->unsigned casts of
 * src/whirl.cpp:56-58: gcc converts through a 64-bit signed integer and keeps
 * the low 32 bits (CUDA's direct cast would saturate negatives to 0). */
__device__ __forceinline__ unsigned f2u_x86(float f) { return (unsigned)(long long)f; }

/* ------------------------------------------------------------------ */
/* Ordered two-point scatter of one 32-sample block, lane = sample.     */
/* Sample i adds a[i] to slot n[i] and b[i] to slot n[i]+1 (scanner:    */
/* src/vibrato.cpp:389-390; Leslie taps: src/whirl.cpp:1448-1451).      */
/* Because n[] never decreases inside the block, the samples hitting a  */
/* slot form a contiguous run and the reference's sample loop adds to   */
/* slot S first the b-parts of the run that hit S-1, then the a-parts   */
/* of the run that hits S.  The first sample of each run applies        */
/* exactly that sequence; a slot that only receives b-parts (the next   */
/* run skips it, or the block ends) is completed by the last sample of  */
/* the run before.  Distinct lanes own distinct slots: no atomics.      */
/* ------------------------------------------------------------------ */
/* The pass, from two ballots instead of per-lane neighbour compares.  With n[] never decreasing,
 * everything the pass has to know about its neighbours is in e (bit i: n[i] == n[i-1]) and d (bit i:
 * n[i] == n[i-1] + 1): "n[i-2] hits n[i]-1" is d[i] & e[i-1], "n[i+1] hits n[i]" is e[i+1], "n[i+1]
 * hits n[i]+1" is d[i+1].  Written without branches: a term that does not apply is replaced by -0.0f,
 * whose addition changes no value (x + -0 = x for every x, -0 included), so all lanes run the same
 * dozen instructions instead of the union of the branches.  When no two samples share a slot (e
 * empty: most blocks, a tap moves by less than a slot in 32 samples most of the time) every lane owns
 * slot n[i], which takes b[i-1] if the sample before sits right below, then a[i]; slot n[i]+1 is
 * completed with b[i] where the next sample skips it. */
struct RunBits {
    uint32_t e, d;
};
__device__ __forceinline__ RunBits run_bits(const uint32_t n, const uint32_t smask, const int lane, uint32_t &dn)
{
    const uint32_t n1 = __shfl_up_sync(FULL, n, 1);
    dn = (n - n1) & smask;
    RunBits r;
    r.e = __ballot_sync(FULL, lane >= 1 && dn == 0u);
    r.d = __ballot_sync(FULL, lane >= 1 && dn == 1u);
    return r;
}
/* rmask: the ring actually held (a power of two that divides the delay line's length, so the slot
 * index may be reduced by it alone) */
__device__ __forceinline__ void ordered_pass_bits(const uint32_t n, const RunBits rb, const float a, const float b,
                                                  float *ring, const uint32_t rmask, const int lane)
{
    const float nz = -0.0f;
    const float b1 = __shfl_up_sync(FULL, b, 1);
    float *const p0 = ring + (n & rmask), *const p1 = ring + ((n + 1) & rmask);
    if (rb.e == 0u) { /* warp-uniform: no slot is hit twice, every lane owns its slot n */
        const uint32_t dd = rb.d >> lane;
        float v = *p0;
        float w = *p1; /* stored only where no lane has n+1 as its own slot */
        v = __fadd_rn(v, (dd & 1u) ? b1 : nz);
        v = __fadd_rn(v, a);
        w = __fadd_rn(w, b);
        *p0 = v;
        if (!(dd & 2u))
            *p1 = w;
        return;
    }
    const float b2 = __shfl_up_sync(FULL, b, 2);
    const float ax = __shfl_down_sync(FULL, a, 1);
    const uint32_t ee = rb.e >> lane, ep = (rb.e << 1) >> lane, dd = rb.d >> lane;
    const bool e_i = (ee & 1u) != 0, e_x = (ee & 2u) != 0, e_p = (ep & 1u) != 0;
    const bool d_i = (dd & 1u) != 0, d_x = (dd & 2u) != 0;
    float v = *p0;
    float w = *p1; /* a slot no lane has as its main slot whenever it is stored below */
    v = __fadd_rn(v, (d_i && e_p) ? b2 : nz);
    v = __fadd_rn(v, d_i ? b1 : nz);
    v = __fadd_rn(v, a);
    v = __fadd_rn(v, e_x ? ax : nz);
    w = __fadd_rn(w, e_i ? b1 : nz);
    w = __fadd_rn(w, b);
    if (!e_i) /* first sample of the run hitting slot n */
        *p0 = v;
    if (!e_x && !d_x) /* last sample of its run, and nobody hits n+1 as a main slot in this block */
        *p1 = w;
}

/* ================================================================== */
/* stage T: tone generator + scanner                                   */
/* ================================================================== */

/* One entry of the per-fragment "core program": the reference emits 1-2 CoreIns per
 * active wheel (src/tonegen.h:114-129); with a modular index the wrap split collapses
 * and one record per wheel suffices. */
#define REC_TILE 32    /* active-list entries turned into records and summed at a time */
#define REC_CAP TBF_REC_CAP /* longest active list whose records persist across fragments */
#define T_MAX_SLOTS 8  /* wave staging slots per warp = wheels in flight (bulk copies): 2 when the
                          SMs are full (6 KB per warp, 28 warps per SM), 6 for small batches, where a
                          warp's own copies are all that hides the copy latency                    */
struct TWarp {
    TgScalars sc;
    /* core-program records.  Lists of up to REC_CAP wheels keep their records here from one
     * fragment to the next: a fragment without events only advances the read positions
     * (fast_pass) and never touches the per-wheel state in HBM */
    float4 recG[REC_CAP];   /* gs gv gp, env as int bits (-1 plain, 0..7 attack, 8..15 release) */
    uint32_t recQ[REC_CAP]; /* where the wheel's next 128 (256: pairs) samples start in the wave
                               store, in float4 units                                            */
    uint32_t recL[REC_CAP]; /* wave length | read position << 16 (of the record's last fragment) */
    uint16_t recW[REC_CAP]; /* wheel                                                             */
    /* per bus (swell, scanner, percussion): the records whose gain on that bus is not zero, in list
     * order (bus_lists_build) */
    uint8_t bus_list[3][REC_CAP];
    int bus_n[3];
    /* wave samples in flight: one 512-byte bulk copy (TMA) per wheel and fragment lands here,
     * [slot][fragment A / B][lane] = samples 4*lane .. 4*lane+3.  The single-fragment passes use
     * the A halves only; the B halves then hold the envelope deltas of the current tile
     * (rec_d) and, once the last tile has been summed, the removal list (removed_wheel).
     * The slots lie BEHIND this struct in the warp's shared memory (wslot()): their number is a
     * launch-time choice */
    unsigned long long mbar[T_MAX_SLOTS]; /* "slot filled" barriers of the bulk copies          */
    const float *wave4;               /* the organ's wave store (read by the lane that requests) */
    float vring[TBF_VRING];
    uint8_t removed[TBF_NOF_WHEELS + 4];
    int recompute_routing, draw_bar_change, vib_on, perc_on;
    /* raised by any lane during the active-list pass of a fragment.  Shared-memory flags read
     * after a __syncwarp(), not warp votes: ptxas assumed the warp converged behind the tile
     * loop, dropped the WARPSYNC before VOTE.ANY, and lanes got different answers */
    int any_removed, any_env;
};
static_assert(sizeof(TWarp) % 16 == 0, "the staging slots behind the struct are float4-aligned");
static_assert(sizeof(TWarp) + 2 * 1024 <= 7936, "stage T: 7 CTAs of 4 warps per SM need <= 7936 B of shared memory per warp");
static_assert(REC_TILE * sizeof(float4) <= sizeof(float4) * 32 &&
              TBF_NOF_WHEELS * sizeof(uint16_t) <= sizeof(float4) * 32, "slow-path arrays fit a B half");

/* envelope deltas ds dv dp of the current tile's records (slow path, env >= 0 only) */
/* staging slot `slot`, fragment half `f`: 32 float4 */
__device__ __forceinline__ float4 *wslot(TWarp &sm, int slot, int f)
{
    return reinterpret_cast<float4 *>(&sm + 1) + (slot * 2 + f) * 32;
}
__device__ __forceinline__ float4 *rec_d(TWarp &sm) { return wslot(sm, 0, 1); }
__device__ __forceinline__ uint16_t *removed_wheel(TWarp &sm)
{
    return reinterpret_cast<uint16_t *>(wslot(sm, 1, 1));
}

/* -- key events, one lane (src/tonegen.cpp:3096-3166 + the queue drain :3257-3327): the
 *    fallback when a key's contribution list is not wheel-sorted with unique pairs -- */
__device__ void key_off_serial(TgScalars &sc, InstState *st, const DevTonegenTables *tg, int key)
{
    const uint32_t e0 = tg->contrib_start[key], e1 = tg->contrib_start[key + 1];
    for (uint32_t e = e0; e < e1; ++e) {
        const int w = tg->contrib_wheel[e];
        const int b = tg->contrib_bus[e];
        OscState *o = &st->osc[w];
        float *bl = &st->busLevel[w * TBF_NOF_BUSES + b];
        *bl = __fsub_rn(*bl, tg->contrib_level[e]);
        o->refCount -= 1;
        if (o->refCount == 0)
            o->rflags = OR_REM;
        else
            o->rflags |= ORF_MODIFIED;
    }
}

__device__ void key_on_serial(TgScalars &sc, InstState *st, const DevTonegenTables *tg, int key)
{
    const uint32_t e0 = tg->contrib_start[key], e1 = tg->contrib_start[key + 1];
    for (uint32_t e = e0; e < e1; ++e) {
        const int w = tg->contrib_wheel[e];
        const int b = tg->contrib_bus[e];
        OscState *o = &st->osc[w];
        if (o->refCount == 0) {
            o->rflags = OR_ADD;
            if (o->aclPos == -1) {
                o->aclPos = sc.activeOscLEnd;
                st->activeOscList[sc.activeOscLEnd++] = w;
            }
        } else {
            o->rflags |= ORF_MODIFIED;
        }
        float *bl = &st->busLevel[w * TBF_NOF_BUSES + b];
        *bl = __fadd_rn(*bl, tg->contrib_level[e]);
        o->refCount += 1;
    }
}

/* -- key events, whole warp: lane = contribution for the bus levels (every (wheel, bus)
 *    cell of one key is distinct), lane = wheel run for the reference counts, flags and
 *    the order-preserving append to the active list -- */
__device__ void key_off_warp(TWarp &sm, InstState *st, const DevTonegenTables *tg, int key, int lane)
{
    TgScalars &sc = sm.sc;
    if (lane == 0) {
        st->activeKeys[key] = 0;
        if (key < 128)
            sc.upperKeyCount--;
        sc.keyDownCount--;
    }
    if (!tg->coop_events) {
        if (lane == 0)
            key_off_serial(sc, st, tg, key);
        __syncwarp();
        return;
    }
    const uint32_t e0 = tg->contrib_start[key], e1 = tg->contrib_start[key + 1];
    for (uint32_t e = e0 + lane; e < e1; e += 32) {
        float *bl = &st->busLevel[(int)tg->contrib_wheel[e] * TBF_NOF_BUSES + (int)tg->contrib_bus[e]];
        *bl = __fsub_rn(*bl, tg->contrib_level[e]);
    }
    const uint32_t r0 = tg->run_start[key], r1 = tg->run_start[key + 1];
    for (uint32_t r = r0 + lane; r < r1; r += 32) {
        OscState *o = &st->osc[tg->run_wheel[r]];
        const int rc = o->refCount - (int)tg->run_count[r];
        o->refCount = rc;
        if (rc == 0)
            o->rflags = OR_REM;
        else
            o->rflags |= ORF_MODIFIED;
    }
    __syncwarp();
}

__device__ void key_on_warp(TWarp &sm, InstState *st, const DevTonegenTables *tg, int key, int lane)
{
    TgScalars &sc = sm.sc;
    if (st->activeKeys[key]) /* re-striking a held key: off, then on (src/tonegen.cpp:3139-3142) */
        key_off_warp(sm, st, tg, key, lane);
    if (lane == 0) {
        st->activeKeys[key] = 1;
        if (key < 128)
            sc.upperKeyCount++;
        sc.keyDownCount++;
    }
    if (!tg->coop_events) {
        if (lane == 0)
            key_on_serial(sc, st, tg, key);
        __syncwarp();
        return;
    }
    const uint32_t e0 = tg->contrib_start[key], e1 = tg->contrib_start[key + 1];
    for (uint32_t e = e0 + lane; e < e1; e += 32) {
        float *bl = &st->busLevel[(int)tg->contrib_wheel[e] * TBF_NOF_BUSES + (int)tg->contrib_bus[e]];
        *bl = __fadd_rn(*bl, tg->contrib_level[e]);
    }
    const uint32_t r0 = tg->run_start[key], r1 = tg->run_start[key + 1];
    int l_end = sc.activeOscLEnd; /* uniform */
    for (uint32_t base = r0; base < r1; base += 32) {
        const uint32_t r = base + lane;
        const bool valid = r < r1;
        int w = 0, rc = 1;
        bool append = false;
        if (valid) {
            w = tg->run_wheel[r];
            rc = st->osc[w].refCount;
            append = (rc == 0) && (st->osc[w].aclPos == -1);
        }
        const unsigned am = __ballot_sync(FULL, append);
        if (valid) {
            OscState *o = &st->osc[w];
            if (rc == 0) {
                o->rflags = OR_ADD;
                if (append) {
                    const int pos = l_end + __popc(am & ((1u << lane) - 1u));
                    o->aclPos = pos;
                    st->activeOscList[pos] = w;
                }
            } else {
                o->rflags |= ORF_MODIFIED;
            }
            o->refCount = rc + (int)tg->run_count[r];
        }
        l_end += __popc(am);
    }
    __syncwarp();
    if (lane == 0)
        sc.activeOscLEnd = l_end;
    __syncwarp();
}

__device__ void perc_resets(TgScalars &sc, const DevTonegenTables *tg)
{
    if (sc.percIsFast)
        sc.percEnvGainDecay = sc.percIsSoft ? tg->perc_decay_fast_soft : tg->perc_decay_fast_norm;
    else
        sc.percEnvGainDecay = sc.percIsSoft ? tg->perc_decay_slow_soft : tg->perc_decay_slow_norm;
}

/* setters of the tone generator and the scanner, one lane */
__device__ void apply_scalar_event(TgScalars &sc, const DevTonegenTables *tg, const int type,
                                   const int a, const int b)
{
    switch (type) {
    case TBF_EV_DRAWBAR:
        if (a < 0 || a >= TBF_NOF_BUSES || b < 0 || b >= 9) { /* reference asserts */
            sc.errorFlags |= TBF_IFLAG_BAD_EVENT;
            break;
        }
        sc.drawBarChange = 1;
        if (a == tg->perc_trigger_bus) {
            sc.percTrigRestore = b;
            if (sc.percEnabled)
                break;
        }
        sc.drawBarGain[a] = tg->draw_bar_level[a][b];
        break;
    case TBF_EV_PERC_ENABLE:
        if (a) {
            sc.newRouting |= RT_PERC;
            if (-1 < tg->perc_trigger_bus) {
                sc.drawBarGain[tg->perc_trigger_bus] = 0.0f;
                sc.drawBarChange = 1;
            }
        } else {
            sc.newRouting &= ~RT_PERC;
            if (-1 < tg->perc_trigger_bus) {
                sc.drawBarGain[tg->perc_trigger_bus] =
                    tg->draw_bar_level[tg->perc_trigger_bus][sc.percTrigRestore];
                sc.drawBarChange = 1;
            }
        }
        sc.percEnabled = a;
        break;
    case TBF_EV_PERC_SOFT:
        sc.percIsSoft = a;
        sc.percEnvGainReset = __fmul_rn(
            tg->perc_env_scaling, a ? tg->perc_env_gain_reset_soft : tg->perc_env_gain_reset_norm);
        sc.percDrawbarGain = a ? tg->perc_drawbar_soft_gain : tg->perc_drawbar_normal_gain;
        perc_resets(sc, tg);
        break;
    case TBF_EV_PERC_FAST:
        sc.percIsFast = a;
        perc_resets(sc, tg);
        break;
    case TBF_EV_PERC_FIRST:
        sc.percSendBus = a ? tg->perc_send_bus_a : tg->perc_send_bus_b;
        break;
    case TBF_EV_VIB_UPPER:
        if (a)
            sc.newRouting |= RT_UPPRVIB;
        else
            sc.newRouting &= ~RT_UPPRVIB;
        break;
    case TBF_EV_VIB_LOWER:
        if (a)
            sc.newRouting |= RT_LOWRVIB;
        else
            sc.newRouting &= ~RT_LOWRVIB;
        break;
    case TBF_EV_VIB_KNOB:
        if (a >= 0 && a <= 5) { /* V1 C1 V2 C2 V3 C3 */
            sc.vibTable = a >> 1;
            sc.vibMixed = (a & 1) ? 0x80 : 0;
        }
        break;
    case TBF_EV_SWELL:
        sc.swellPedalGain =
            (float)(((double)tg->output_level_trim * (double)(unsigned char)a) / 127.0);
        break;
    case TBF_EV_REV_OPTION:
    case TBF_EV_REV_SELECT:
    case TBF_EV_WHIRL_BRAKE:
    case TBF_EV_WHIRL_RAMP:
    case TBF_EV_HORN_FILTER:
        break; /* rotor and horn filter events belong to stage R */
    default:
        sc.errorFlags |= TBF_IFLAG_BAD_EVENT;
        break;
    }
}

/* read position one fragment later (src/tonegen.cpp:3376, :3524) */
__device__ __forceinline__ uint32_t next_position(uint32_t pos, uint32_t len)
{
    return (len < pos + TBF_FRAGMENT) ? TBF_FRAGMENT - (len - pos) : pos + TBF_FRAGMENT;
}

/* where the 128 samples from `pos` start in the wave store, in float4 units from its base:
 * copy (pos & 3) of the wheel, 16-byte aligned (tbf_wave_copy_len) */
__device__ __forceinline__ uint32_t wave_q(uint32_t off4, uint32_t len, uint32_t pos)
{
    return (off4 + (pos & 3u) * tbf_wave_copy_len(len) + (pos & ~3u)) >> 2;
}

/* active-list pass (src/tonegen.cpp:3333-3594), entries [base, base + cnt) -> records
 * 0..cnt-1.  Entry i of the active list touches only its own wheel's state, so lane l
 * handles entries base + l, base + l + 32. */
__device__ __forceinline__ void active_list_tile(InstState *st, const DevTonegenTables *tg, TWarp &sm,
                                                 int lane, int base, int cnt, int rec0, int &my_ins)
{
    TgScalars &sc = sm.sc;
    const uint32_t routing = sc.oldRouting; /* already updated by lane 0 */
    const bool recompute_routing = sm.recompute_routing != 0;
    const bool draw_bar_change = sm.draw_bar_change != 0;

    for (int i = base + lane; i < base + cnt; i += 32) {
        const int w = st->activeOscList[i];
#ifdef TBF_DEBUG_CHECKS
        if (w < 1 || w > TBF_NOF_WHEELS) {
            printf("active list bad: slot %d i %d w %d base %d cnt %d rec0 %d lend %d lane %d keys %d\n", (int)(blockIdx.x * T_WARPS + threadIdx.x / 32), i, w, base, cnt, rec0, sc.activeOscLEnd, lane, sc.keyDownCount);
            __trap();
        }
#endif
        OscState *o = &st->osc[w];
        const uint32_t rflags = o->rflags;
        const uint32_t len = tg->wave_length[w];
        const uint32_t pos = o->pos;
        int env;
        float gs, gv, gp, ds, dv, dp;

        if (rflags & ORF_REMOVED) {
            sm.removed[i] = 1;
            sm.any_removed = 1;
            env = 8 + (i & 7);
            gs = o->sumSwell;
            gp = o->sumPercn;
            gv = o->sumScanr;
            ds = __fsub_rn(0.0f, gs);
            dp = __fsub_rn(0.0f, gp);
            dv = __fsub_rn(0.0f, gv);
        } else {
            sm.removed[i] = 0;
            bool reroute = false;
            if (rflags & ORF_ADDED) {
                gs = gp = gv = 0.0f;
            } else {
                gs = o->sumSwell;
                gp = o->sumPercn;
                gv = o->sumScanr;
            }
            if ((rflags & ORF_MODIFIED) || draw_bar_change) {
                const float *bl = &st->busLevel[w * TBF_NOF_BUSES];
                float sum = 0.0f;
                for (int d = 0; d < 9; ++d)
                    sum = __fadd_rn(sum, __fmul_rn(bl[d], sc.drawBarGain[d]));
                o->sumUpper = sum;
                sum = 0.0f;
                for (int d = 9; d < 18; ++d)
                    sum = __fadd_rn(sum, __fmul_rn(bl[d], sc.drawBarGain[d]));
                o->sumLower = sum;
                sum = 0.0f;
                for (int d = 18; d < 27; ++d)
                    sum = __fadd_rn(sum, __fmul_rn(bl[d], sc.drawBarGain[d]));
                o->sumPedal = sum;
                reroute = true;
            }
            if (reroute || recompute_routing) {
                float percn = 0.0f;
                if (routing & RT_PERC)
                    percn = st->busLevel[w * TBF_NOF_BUSES + sc.percSendBus];
                float scanr = 0.0f;
                float swell = o->sumPedal;
                if (routing & RT_UPPRVIB)
                    scanr = __fadd_rn(scanr, o->sumUpper);
                else
                    swell = __fadd_rn(swell, o->sumUpper);
                if (routing & RT_LOWRVIB)
                    scanr = __fadd_rn(scanr, o->sumLower);
                else
                    swell = __fadd_rn(swell, o->sumLower);
                o->sumPercn = percn;
                o->sumScanr = scanr;
                o->sumSwell = swell;
            }
            if (rflags & OR_ADD) {
                env = i & 7;
                ds = __fsub_rn(o->sumSwell, gs);
                dp = __fsub_rn(o->sumPercn, gp);
                dv = __fsub_rn(o->sumScanr, gv);
            } else {
                env = -1;
                ds = dp = dv = 0.0f;
            }
        }
        /* advance the read position; wrap test is `length < pos + 128`
         * (src/tonegen.cpp:3376, :3524), so pos may rest at == length */
        if (len < pos + TBF_FRAGMENT) {
            o->pos = TBF_FRAGMENT - (len - pos);
            my_ins += 2;
        } else {
            o->pos = pos + TBF_FRAGMENT;
            my_ins += 1;
        }
        o->rflags = 0;
        const int rec = rec0 + i - base;
        sm.recQ[rec] = wave_q(tg->wave4_offset[w], len, pos);
        sm.recL[rec] = len | (pos << 16);
        sm.recW[rec] = (uint16_t)w;
        sm.recG[rec] = make_float4(gs, gv, gp, __int_as_float(env));
        if (env >= 0) {
            rec_d(sm)[i - base] = make_float4(ds, dv, dp, 0.0f);
            sm.any_env = 1;
        }
    }
}

/* A fragment without events, routing or drawbar change over a list whose records persist:
 * every wheel plays on with unchanged gains, only its read position advances
 * (src/tonegen.cpp:3521-3555: wrap test `length < pos + 128`, pos may rest at == length).
 * NF = 2: the next TWO fragments are quiet, the records are advanced for both at once. */
template <int NF>
__device__ __forceinline__ void fast_pass(const DevTonegenTables *tg, TWarp &sm, int lane, int n_act,
                                          int &ins_a, int &ins_b)
{
    for (int i = lane; i < n_act; i += 32) {
        const uint32_t l = sm.recL[i];
        const uint32_t len = l & 0xFFFFu;
        const uint32_t off4 = __ldg(&tg->wave4_offset[sm.recW[i]]);
        /* the position this record was last rendered from moves on by one fragment */
        const uint32_t pos_a = next_position(l >> 16, len);
        ins_a += (len < pos_a + TBF_FRAGMENT) ? 2 : 1;
        uint32_t pos_b = pos_a;
        if (NF == 2) { /* fragment B reads on from where A ends, in the same image of the wave */
            pos_b = next_position(pos_a, len);
            ins_b += (len < pos_b + TBF_FRAGMENT) ? 2 : 1;
        }
        sm.recQ[i] = wave_q(off4, len, pos_a);
        sm.recL[i] = len | (pos_b << 16);
    }
}

/* persistent records -> HBM: the position after the fragment last rendered from them */
__device__ __forceinline__ void flush_positions(InstState *st, const TWarp &sm, int lane, int n_act)
{
    for (int i = lane; i < n_act; i += 32) {
        const uint32_t l = sm.recL[i];
        st->osc[sm.recW[i]].pos = next_position(l >> 16, l & 0xFFFFu);
    }
}

/* end of the active-list pass: statistics, reference-UB flag, removal list */
__device__ __forceinline__ void active_list_finish(InstState *st, TWarp &sm, int lane, int n_act,
                                                   int my_ins)
{
    TgScalars &sc = sm.sc;
    __syncwarp(); /* removed[], the flags and the wheel state are visible to lane 0 */
    const int n_ins = __reduce_add_sync(FULL, my_ins);
    const int any_removed = sm.any_removed;

    if (lane == 0) {
        sc.statFragments += 1;
        sc.statActiveWheels += (unsigned long long)n_act;
        sc.drawBarChange = 0;
        if (n_ins > 256)
            sc.errorFlags |= TBF_IFLAG_COREPGM_OVERFLOW;
        if (any_removed) {
            /* removal list in ascending active-list index, swap-with-last
             * (src/tonegen.cpp:3576-3594); changes the list order and with it next
             * fragment's summation order and envelope rows.  Victims are identified by
             * wheel number BEFORE any swap moves entries around: the reference collects
             * removedList[] first */
            int l_end = n_act;
            for (int i = 0; i < n_act; ++i)
                if (sm.removed[i])
                    removed_wheel(sm)[i] = (uint16_t)st->activeOscList[i];
            for (int i = 0; i < n_act; ++i) {
                if (!sm.removed[i])
                    continue;
                const int vic = removed_wheel(sm)[i];
                const int actidx = st->osc[vic].aclPos;
                st->osc[vic].aclPos = -1;
                l_end--;
                if (0 < l_end) {
                    const int mov = st->activeOscList[l_end];
                    if (mov != vic) {
                        st->activeOscList[actidx] = mov;
                        st->osc[mov].aclPos = actidx;
                    }
                }
            }
            sc.activeOscLEnd = l_end;
        }
    }
    __syncwarp();
}

/* ---- TMA bulk copies and their mbarriers (cp.async.bulk -> SASS UBLKCP, SYNCS) ---- */
__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "TBF_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra TBF_DONE;\n"
        "bra TBF_WAIT;\n"
        "TBF_DONE:\n"
        "}" ::"r"(bar), "r"(parity)
        : "memory");
}
/* global -> shared, `bytes` a multiple of 16, both addresses 16-byte aligned; completion is
 * signalled on the mbarrier as `bytes` of its expected transaction count */
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
/* generic-proxy writes to shared memory are ordered before later async-proxy (TMA) accesses */
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

/* core interpreter (src/tonegen.cpp:3607-3687): lane holds samples 4*lane .. 4*lane+3 of the
 * fragment while the wheels are summed.  The device's wave store (tbf_api.cu) has a 16-byte
 * aligned, never-wrapping image of every 128-sample read, so ONE lane requests a wheel's 512
 * bytes with one bulk copy (TMA) into a staging slot; its completion flips the slot's mbarrier
 * and every lane reads its four samples back with one LDS.128. */
/* one lane of the (converged) warp, chosen by the hardware: the canonical TMA issue pattern */
__device__ __forceinline__ bool elect_one()
{
    uint32_t pred;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "elect.sync _|p, 0xffffffff;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}" : "=r"(pred));
    return pred != 0;
}

/* NF = 2: the wave store's images run 256 samples past every start (tbf_wave_copy_len), so the
 * reads of two consecutive fragments are ONE contiguous kilobyte, wraps included */
template <int NF>
__device__ __forceinline__ void wave_request(TWarp &sm, int rec, int slot)
{
    const uint32_t q = sm.recQ[rec];
    const float *wave4 = sm.wave4;
    const uint32_t bar = smem_u32(&sm.mbar[slot]);
    mbar_expect_tx(bar, NF * 512u);
    bulk_g2s(smem_u32(wslot(sm, slot, 0)), wave4 + (size_t)q * 4u, NF * 512u, bar);
}

/* Packed pairs of floats for Blackwell's two-wide FP32 multiply (FMUL2): each half is an IEEE
 * round-to-nearest operation, bit-identical to the scalar one.  The sums are scalar FADDs:
 * ptxas contracts a mul.f32x2 feeding an add.f32x2 into FFMA2 even with --fmad=false, which
 * would break the reference's rounding (one rounding instead of two); a packed product
 * feeding scalar adds cannot be contracted, needs no carried products (round 1 carried them
 * one loop trip: 24 more registers and 8 MOVs per wheel) and tests/test_cabi.py checks that
 * the built kernels contain no fused multiply-add at all. */
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 pack2(float lo, float hi)
{
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float &lo, float &hi)
{
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}

/* a + b as a packed instruction that ptxas cannot fuse with the multiply that made `a`:
 * fma(a, one, b) with `one` = 1.0f in both halves, a value the compiler cannot see (it arrives in
 * RenderParams).  a * 1.0 is exact, so the result is the correctly rounded sum -- bit for bit
 * add.rn(a, b), signed zeros included.  (A plain add.rn.f32x2 fed by mul.rn.f32x2 is contracted
 * into FFMA2 -- one rounding instead of the reference's two -- even under --fmad=false.) */
__device__ __forceinline__ f32x2 add2_via_fma(f32x2 a, f32x2 one2, f32x2 b)
{
    f32x2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(one2), "l"(b));
    return r;
}

/* running sums of the three buffers (swell, scanner, percussion) over the lane's four samples,
 * as packed pairs: [0] = samples 4*lane, 4*lane+1; [1] = 4*lane+2, 4*lane+3 */
struct CoreAcc {
    f32x2 s[2], v[2], p[2];
};

/* The first instruction of a fragment copies, the rest add (src/tonegen.cpp:3640-3680):
 * with the sums started at -0.0f an add IS a copy, bit for bit ((-0) + x = x for every x,
 * (-0) + (-0) = -0). */
__device__ __forceinline__ void core_begin(CoreAcc &a)
{
    const f32x2 nz = 0x8000000080000000ull;
#pragma unroll
    for (int h = 0; h < 2; ++h)
        a.s[h] = a.v[h] = a.p[h] = nz;
}

/* sum += x * g for the lane's four samples x (two packed pairs): 2 FMUL2 + 2 FFMA2 */
__device__ __forceinline__ void acc_plain(f32x2 (&sum)[2], const ulonglong2 x, const f32x2 g2, const f32x2 one2)
{
    sum[0] = add2_via_fma(mul2(x.x, g2), one2, sum[0]);
    sum[1] = add2_via_fma(mul2(x.y, g2), one2, sum[1]);
}

/* the same with a gain per sample (key click), scalar: the rare path */
__device__ __forceinline__ void acc_env(f32x2 (&sum)[2], const float (&xs)[4], const float (&ev)[4], float g, float d)
{
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        float lo, hi;
        unpack2(sum[h], lo, hi);
        lo = __fadd_rn(lo, __fmul_rn(xs[2 * h], __fadd_rn(g, __fmul_rn(ev[2 * h], d))));
        hi = __fadd_rn(hi, __fmul_rn(xs[2 * h + 1], __fadd_rn(g, __fmul_rn(ev[2 * h + 1], d))));
        sum[h] = pack2(lo, hi);
    }
}

/* products of one record with the lane's staged samples x, added in list order.  ENV = false:
 * the caller knows that every record plays with constant gains (no key click in flight). */
template <bool ENV, int MASK = 7>
__device__ __forceinline__ void wheel_accumulate(const DevTonegenTables *tg, TWarp &sm, const float4 g,
                                                 int i_tile, int lane, const ulonglong2 x, CoreAcc &a,
                                                 const f32x2 one2)
{
    const int env = ENV ? __float_as_int(g.w) : -1;
    if (!ENV || env < 0) {
        /* MASK: the buses this fragment needs (quiet fragments only: quiet_mask) */
        if (MASK & 1)
            acc_plain(a.s, x, pack2(g.x, g.x), one2);
        if (MASK & 2)
            acc_plain(a.v, x, pack2(g.y, g.y), one2);
        if (MASK & 4)
            acc_plain(a.p, x, pack2(g.z, g.z), one2);
    } else { /* key click: gain = g + env[n] * (next - g), per sample */
        const float4 d = rec_d(sm)[i_tile]; /* only the pass's tiles (<= REC_TILE records) carry envelopes */
        const float *ep = tg->env[env] + 4 * lane;
        const float ev[4] = {__ldg(ep), __ldg(ep + 1), __ldg(ep + 2), __ldg(ep + 3)};
        float xs[4];
        unpack2(x.x, xs[0], xs[1]);
        unpack2(x.y, xs[2], xs[3]);
        acc_env(a.s, xs, ev, g.x, d.x);
        acc_env(a.v, xs, ev, g.y, d.y);
        acc_env(a.p, xs, ev, g.z, d.z);
    }
}

/* one wheel: wait for its samples in staging slot SLOT, read them and the record's gains,
 * hand the slot back (the wheel SLOTS further is requested into it), accumulate */
template <int NF, bool ENV, int SLOTS, int SLOT>
__device__ __forceinline__ void core_step(const DevTonegenTables *tg, TWarp &sm, int lane, int rec, int i_tile,
                                          bool more, CoreAcc (&acc)[NF], uint32_t &parity, const f32x2 one2)
{
    mbar_wait(smem_u32(&sm.mbar[SLOT]), (parity >> SLOT) & 1u);
    parity ^= 1u << SLOT;
    ulonglong2 x[NF];
#pragma unroll
    for (int f = 0; f < NF; ++f)
        x[f] = *reinterpret_cast<const ulonglong2 *>(wslot(sm, SLOT, f) + lane);
    const float4 g = sm.recG[rec];
    __syncwarp(); /* every lane has read the slot: it may be filled again */
    if (more && elect_one())
        wave_request<NF>(sm, rec + SLOTS, SLOT);
#pragma unroll
    for (int f = 0; f < NF; ++f)
        wheel_accumulate<ENV>(tg, sm, g, i_tile, lane, x[f], acc[f], one2);
}

/* compile-time loop over the staging slots */
template <int NF, bool ENV, int SLOTS, int J>
struct CoreSteps {
    /* a full trip: wheels i .. i + SLOTS - 1, one per slot */
    static __device__ __forceinline__ void trip(const DevTonegenTables *tg, TWarp &sm, int lane, int rec0, int i,
                                                int cnt, CoreAcc (&acc)[NF], uint32_t &parity, const f32x2 one2)
    {
        core_step<NF, ENV, SLOTS, J>(tg, sm, lane, rec0 + i + J, i + J, i + J + SLOTS < cnt, acc, parity, one2);
        CoreSteps<NF, ENV, SLOTS, J + 1>::trip(tg, sm, lane, rec0, i, cnt, acc, parity, one2);
    }
    /* the last, partial trip: nothing is requested any more */
    static __device__ __forceinline__ void tail(const DevTonegenTables *tg, TWarp &sm, int lane, int rec0, int i,
                                                int cnt, CoreAcc (&acc)[NF], uint32_t &parity, const f32x2 one2)
    {
        if (i + J < cnt) {
            core_step<NF, ENV, SLOTS, J>(tg, sm, lane, rec0 + i + J, i + J, false, acc, parity, one2);
            CoreSteps<NF, ENV, SLOTS, J + 1>::tail(tg, sm, lane, rec0, i, cnt, acc, parity, one2);
        }
    }
};
template <int NF, bool ENV, int SLOTS>
struct CoreSteps<NF, ENV, SLOTS, SLOTS> {
    static __device__ __forceinline__ void trip(const DevTonegenTables *, TWarp &, int, int, int, int,
                                                CoreAcc (&)[NF], uint32_t &, const f32x2)
    {
    }
    static __device__ __forceinline__ void tail(const DevTonegenTables *, TWarp &, int, int, int, int,
                                                CoreAcc (&)[NF], uint32_t &, const f32x2)
    {
    }
};

/* records rec0 .. rec0+cnt-1 in order (src/tonegen.cpp:3607-3687), SLOTS wheels in flight and
 * SLOTS wheels per trip, so that slot numbers are compile-time constants; NF = 2 sums the same
 * records into two consecutive fragments (acc[0], acc[1]) from the 256-sample read fast_pass<2>
 * laid out.  parity: bit s = the parity slot s's barrier shows when its next fill has landed. */
template <int NF, bool ENV, int SLOTS>
__device__ __forceinline__ void core_tile_tma(const DevTonegenTables *tg, TWarp &sm, int lane, int rec0,
                                              int cnt, CoreAcc (&acc)[NF], uint32_t &parity, const f32x2 one2)
{
    if (cnt > 0 && elect_one()) {
#pragma unroll
        for (int j = 0; j < SLOTS; ++j)
            if (j < cnt)
                wave_request<NF>(sm, rec0 + j, j);
    }
    int i = 0;
#pragma unroll 1
    for (; i + SLOTS <= cnt; i += SLOTS)
        CoreSteps<NF, ENV, SLOTS, 0>::trip(tg, sm, lane, rec0, i, cnt, acc, parity, one2);
    CoreSteps<NF, ENV, SLOTS, 0>::tail(tg, sm, lane, rec0, i, cnt, acc, parity, one2);
}

/* The same sum with the Ampere-style pipeline: every lane copies its own 16 bytes of a wheel (per
 * fragment) into the staging slot with cp.async (LDGSTS), STAGES wheels ahead, and reads back
 * only what it requested itself -- no barrier object, no elected lane, a third fewer instructions
 * per wheel than the bulk-copy variant and a shorter dependent chain per step, at the price of 32
 * requests through the LSU per wheel instead of one through the TMA unit.  It is the variant for
 * SMALL batches, where each scheduler holds one or two warps and the step latency of a warp is
 * what bounds the kernel (profiles/r02_kernels.md); tbf_launch_stage_t picks. */
template <int NF, bool ENV, int STAGES>
__device__ __forceinline__ void core_tile_cpa(const DevTonegenTables *tg, TWarp &sm, int lane, int rec0,
                                              int cnt, CoreAcc (&acc)[NF], const f32x2 one2)
{
    const float4 *base = reinterpret_cast<const float4 *>(sm.wave4) + lane;
    float4 *stage = wslot(sm, 0, 0) + lane; /* slot s, half f: + (2 s + f) * 32 */
#pragma unroll
    for (int j = 0; j < STAGES; ++j) {
        if (j < cnt) {
            const float4 *q = base + sm.recQ[rec0 + j];
#pragma unroll
            for (int f = 0; f < NF; ++f)
                __pipeline_memcpy_async(stage + (2 * j + f) * 32, q + 32 * f, 16);
        }
        __pipeline_commit();
    }
    int s = 0;
#pragma unroll 1
    for (int i = 0; i < cnt; ++i) {
        __pipeline_wait_prior(STAGES - 1); /* this lane's copies of wheel i have landed */
        ulonglong2 x[NF];
#pragma unroll
        for (int f = 0; f < NF; ++f)
            x[f] = *reinterpret_cast<const ulonglong2 *>(stage + (2 * s + f) * 32);
        const float4 g = sm.recG[rec0 + i];
        if (i + STAGES < cnt) {
            const float4 *q = base + sm.recQ[rec0 + i + STAGES];
#pragma unroll
            for (int f = 0; f < NF; ++f)
                __pipeline_memcpy_async(stage + (2 * s + f) * 32, q + 32 * f, 16);
        }
        __pipeline_commit();
#pragma unroll
        for (int f = 0; f < NF; ++f)
            wheel_accumulate<ENV>(tg, sm, g, i, lane, x[f], acc[f], one2);
        s = (s == STAGES - 1) ? 0 : s + 1;
    }
    __pipeline_wait_prior(0);
}

/* The same sum without staging: every lane loads its own 16 bytes of a wheel (per fragment) straight
 * into registers with LDG.128, DEPTH wheels ahead, the register sets rotating at compile time.  Two
 * instructions per wheel and fragment pair fetch the samples (no request, no barrier, no read-back
 * from shared memory), and the L1 data pipe sees the same four wavefronts per 512 bytes a bulk copy's
 * read-back costs; the price is 8 registers per wheel in flight. */
template <int NF>
__device__ __forceinline__ void wave_load_direct(const char *lane_base, uint32_t q, ulonglong2 (&x)[NF])
{
    const ulonglong2 *src = reinterpret_cast<const ulonglong2 *>(lane_base + (size_t)q * 16u);
#pragma unroll
    for (int f = 0; f < NF; ++f)
        x[f] = __ldg(src + 32 * f);
}
/* One trip = DEPTH records, register set J holding record i + J.  Three stages are in flight: record
 * i + J is summed, the samples of record i + J + DEPTH are requested into the set it leaves (their
 * address has been waiting in qn[J] since the trip before, so the request goes out the moment the
 * set is free -- with the address looked up in the same trip ptxas places the loads at the trip's
 * end, right before their first use, and every trip waits out a full L2 round trip), and the address
 * of record i + J + 2 DEPTH is looked up.  Records past the tile's end are clamped to its last one: a
 * load nobody uses instead of a branch (a predicated load makes ptxas shuffle the register sets). */
template <int NF, bool ENV, int MASK, int DEPTH, int J, bool LOAD>
struct DirectSteps {
    static __device__ __forceinline__ void run(const DevTonegenTables *tg, TWarp &sm, int lane, int rec0, int i,
                                               int cnt, const char *lane_base, ulonglong2 (&x)[DEPTH][NF],
                                               uint32_t (&qn)[DEPTH], CoreAcc (&acc)[NF], const f32x2 one2)
    {
        if (LOAD || i + J < cnt) {
            const float4 g = sm.recG[rec0 + i + J];
#pragma unroll
            for (int f = 0; f < NF; ++f)
                wheel_accumulate<ENV, MASK>(tg, sm, g, i + J, lane, x[J][f], acc[f], one2);
            if (LOAD) {
                wave_load_direct<NF>(lane_base, qn[J], x[J]);
                qn[J] = sm.recQ[rec0 + min(i + J + 2 * DEPTH, cnt - 1)];
            }
            DirectSteps<NF, ENV, MASK, DEPTH, J + 1, LOAD>::run(tg, sm, lane, rec0, i, cnt, lane_base, x, qn, acc, one2);
        }
    }
};
template <int NF, bool ENV, int MASK, int DEPTH, bool LOAD>
struct DirectSteps<NF, ENV, MASK, DEPTH, DEPTH, LOAD> {
    static __device__ __forceinline__ void run(const DevTonegenTables *, TWarp &, int, int, int, int, const char *,
                                               ulonglong2 (&)[DEPTH][NF], uint32_t (&)[DEPTH], CoreAcc (&)[NF],
                                               const f32x2)
    {
    }
};
template <int NF, bool ENV, int DEPTH, int MASK = 7>
__device__ __forceinline__ void core_tile_direct(const DevTonegenTables *tg, TWarp &sm, int lane, int rec0,
                                                 int cnt, CoreAcc (&acc)[NF], const f32x2 one2)
{
    if (cnt <= 0)
        return;
    /* the lane's window into the wave store, as a value ptxas keeps in a register pair instead of
     * re-deriving it from the thread index inside the loop */
    const char *lane_base = reinterpret_cast<const char *>(sm.wave4) + lane * 16;
    asm volatile("" : "+l"(lane_base));
    ulonglong2 x[DEPTH][NF];
    uint32_t qn[DEPTH];
#pragma unroll
    for (int j = 0; j < DEPTH; ++j)
        wave_load_direct<NF>(lane_base, sm.recQ[rec0 + min(j, cnt - 1)], x[j]);
#pragma unroll
    for (int j = 0; j < DEPTH; ++j)
        qn[j] = sm.recQ[rec0 + min(DEPTH + j, cnt - 1)];
    int i = 0;
#pragma unroll 1
    for (; i + DEPTH <= cnt; i += DEPTH)
        DirectSteps<NF, ENV, MASK, DEPTH, 0, true>::run(tg, sm, lane, rec0, i, cnt, lane_base, x, qn, acc, one2);
    DirectSteps<NF, ENV, MASK, DEPTH, 0, false>::run(tg, sm, lane, rec0, i, cnt, lane_base, x, qn, acc, one2);
}

/* ---- quiet fragments, one bus at a time over the records that feed it ----
 *
 * Most gains are zero: a wheel played from a manual that is routed to the scanner has no swell
 * gain, only the wheels of the percussion harmonic have a percussion gain.  The reference
 * multiplies by those zeros and adds the products (src/tonegen.cpp:3640-3680); the sums do not
 * need them.  A product x * (+-0) is +-0, and adding +-0 changes a running sum only while that sum
 * is itself +-0, and then only its sign.  So the sum over the records with a NON-ZERO gain, in list
 * order, equals the reference's sum bit for bit wherever the final value is not zero; where it is
 * zero only the sign can differ, and that case is recomputed over all records (bus_sum<.., true>).
 * The buses are independent sums, so each runs as its own loop over its own list, samples loaded
 * straight into registers DEPTH records ahead. */
__device__ __forceinline__ void bus_lists_build(TWarp &sm, int lane, int n_act)
{
    int cnt[3] = {0, 0, 0};
    for (int base = 0; base < n_act; base += 32) {
        const int i = base + lane;
        float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i < n_act)
            g = sm.recG[i];
        const float gb[3] = {g.x, g.y, g.z};
#pragma unroll
        for (int b = 0; b < 3; ++b) {
            const bool mine = gb[b] != 0.0f; /* -0 is a zero too; NaN counts as a gain */
            const uint32_t m = __ballot_sync(FULL, mine);
            if (mine)
                sm.bus_list[b][cnt[b] + __popc(m & ((1u << lane) - 1u))] = (uint8_t)i;
            cnt[b] += __popc(m);
        }
    }
    if (lane < 3)
        sm.bus_n[lane] = lane == 0 ? cnt[0] : (lane == 1 ? cnt[1] : cnt[2]);
    __syncwarp();
}

/* entry e of a bus loop: the record it names, where its samples start, its gain on the bus */
template <bool ALL>
__device__ __forceinline__ void bus_entry(const TWarp &sm, const uint8_t *list, int bus, int e, uint32_t &q, float &g)
{
    const int r = ALL ? e : (int)list[e];
    q = sm.recQ[r];
    g = reinterpret_cast<const float *>(&sm.recG[r])[bus];
}
/* One trip = DEPTH entries, register set J holding entry e + J.  Three stages are in flight: entry
 * e + J is summed, the samples of entry e + J + DEPTH are requested into the set it leaves (their
 * address has been waiting in qn[J] since the trip before, so the request can go out the moment the
 * set is free -- with the address looked up in the same trip the loads end up at the trip's end, right
 * before their first use), and the address and gain of entry e + J + 2 DEPTH are looked up.  Entries
 * past the list's end are clamped to its last one: a load nobody uses instead of a branch. */
template <int NF, int DEPTH, bool ALL, int J, bool LOAD>
struct BusSteps {
    static __device__ __forceinline__ void run(const TWarp &sm, const uint8_t *list, int bus, int e, int cnt,
                                               const char *lane_base, ulonglong2 (&x)[DEPTH][NF], float (&g)[DEPTH],
                                               uint32_t (&qn)[DEPTH], float (&gn)[DEPTH], f32x2 (&sum)[NF][2],
                                               const f32x2 one2)
    {
        if (LOAD || e + J < cnt) {
            const f32x2 g2 = pack2(g[J], g[J]);
#pragma unroll
            for (int f = 0; f < NF; ++f)
                acc_plain(sum[f], x[J][f], g2, one2);
            if (LOAD) {
                wave_load_direct<NF>(lane_base, qn[J], x[J]);
                g[J] = gn[J];
                bus_entry<ALL>(sm, list, bus, min(e + J + 2 * DEPTH, cnt - 1), qn[J], gn[J]);
            }
            BusSteps<NF, DEPTH, ALL, J + 1, LOAD>::run(sm, list, bus, e, cnt, lane_base, x, g, qn, gn, sum, one2);
        }
    }
};
template <int NF, int DEPTH, bool ALL, bool LOAD>
struct BusSteps<NF, DEPTH, ALL, DEPTH, LOAD> {
    static __device__ __forceinline__ void run(const TWarp &, const uint8_t *, int, int, int, const char *,
                                               ulonglong2 (&)[DEPTH][NF], float (&)[DEPTH], uint32_t (&)[DEPTH],
                                               float (&)[DEPTH], f32x2 (&)[NF][2], const f32x2)
    {
    }
};
/* sum[f] = the bus's samples of fragment f (the lane's four, as two packed pairs) over the `cnt`
 * records of `list` (ALL: over records 0 .. cnt-1), starting from -0 */
template <int NF, int DEPTH, bool ALL>
__device__ __forceinline__ void bus_sum(const TWarp &sm, int lane, const uint8_t *list, int bus, int cnt,
                                        f32x2 (&sum)[NF][2], const f32x2 one2)
{
    const f32x2 nz = 0x8000000080000000ull;
#pragma unroll
    for (int f = 0; f < NF; ++f)
        sum[f][0] = sum[f][1] = nz;
    if (cnt <= 0)
        return;
    const char *lane_base = reinterpret_cast<const char *>(sm.wave4) + lane * 16;
    asm volatile("" : "+l"(lane_base));
    ulonglong2 x[DEPTH][NF];
    float g[DEPTH], gn[DEPTH];
    uint32_t qn[DEPTH];
#pragma unroll
    for (int j = 0; j < DEPTH; ++j) {
        uint32_t q;
        bus_entry<ALL>(sm, list, bus, min(j, cnt - 1), q, g[j]);
        wave_load_direct<NF>(lane_base, q, x[j]);
    }
#pragma unroll
    for (int j = 0; j < DEPTH; ++j)
        bus_entry<ALL>(sm, list, bus, min(DEPTH + j, cnt - 1), qn[j], gn[j]);
    int e = 0;
#pragma unroll 1
    for (; e + DEPTH <= cnt; e += DEPTH)
        BusSteps<NF, DEPTH, ALL, 0, true>::run(sm, list, bus, e, cnt, lane_base, x, g, qn, gn, sum, one2);
    BusSteps<NF, DEPTH, ALL, 0, false>::run(sm, list, bus, e, cnt, lane_base, x, g, qn, gn, sum, one2);
}
/* any of the lane's samples exactly +-0 */
template <int NF>
__device__ __forceinline__ bool bus_has_zero(const f32x2 (&sum)[NF][2])
{
    bool z = false;
#pragma unroll
    for (int f = 0; f < NF; ++f)
#pragma unroll
        for (int h = 0; h < 2; ++h)
            z |= ((uint32_t)sum[f][h] << 1) == 0u || ((uint32_t)(sum[f][h] >> 32) << 1) == 0u;
    return z;
}
/* one bus of NF quiet fragments: over its own list, again over all records where a sample came out
 * as zero (its sign is the reference's only with the zero products added), over all records straight
 * away when no record feeds the bus */
template <int NF, int DEPTH>
__device__ __forceinline__ void bus_render(const TWarp &sm, int lane, int bus, int n_act, f32x2 (&sum)[NF][2],
                                           const f32x2 one2, bool force_all)
{
    const int nz = sm.bus_n[bus];
    if (nz > 0 && nz < n_act) {
        bus_sum<NF, DEPTH, false>(sm, lane, sm.bus_list[bus], bus, nz, sum, one2);
        if (!force_all && !__any_sync(FULL, bus_has_zero<NF>(sum)))
            return;
    }
    bus_sum<NF, DEPTH, true>(sm, lane, nullptr, bus, n_act, sum, one2);
}

/* NF quiet fragments with direct loads.  Only the buses somebody reads are summed: the scanner bus
 * without vibrato routing and the percussion bus without percussion are never looked at
 * (post_fragment), and a swell bus that no record feeds is left at -0 when the scanner runs and
 * its delay line held no -0 when the launch began (no_neg_zero): then it never will, the scanner's
 * output vy is never -0, and swell + vy = vy for a swell bus that is +-0 whatever its sign.  The
 * remaining buses are summed by one merged loop over all records, or bus by bus over their own
 * lists when those are short (about 19 instructions per list entry against 9 + 8 per bus and
 * record). */
template <int NF, int DEPTH, int BUS_DEPTH>
__device__ __forceinline__ void quiet_sum(const DevTonegenTables *tg, TWarp &sm, int lane, int n_act,
                                          CoreAcc (&acc)[NF], const f32x2 one2, bool no_neg_zero, uint32_t flags)
{
    const bool vib_on = sm.vib_on != 0, perc_on = sm.perc_on != 0;
    const int ns = sm.bus_n[0], nv = sm.bus_n[1], np = sm.bus_n[2];
    const bool skip_s = ns == 0 && vib_on && no_neg_zero;
    int mask = (skip_s ? 0 : 1) | (vib_on ? 2 : 0) | (perc_on ? 4 : 0);
    if (flags & 0x4000u) /* debugging: every bus, always */
        mask = 7;
    const int entries = ((mask & 1) ? (ns ? ns : n_act) : 0) + ((mask & 2) ? (nv ? nv : n_act) : 0) +
                        ((mask & 4) ? (np ? np : n_act) : 0);
    if (!(flags & 0x800u) && entries * 19 < n_act * (9 + 8 * __popc(mask))) {
        const bool force_all = (flags & 0x400u) != 0;
        f32x2 sum[NF][2];
        if (mask & 1) {
            bus_render<NF, BUS_DEPTH>(sm, lane, 0, n_act, sum, one2, force_all);
#pragma unroll
            for (int f = 0; f < NF; ++f)
                acc[f].s[0] = sum[f][0], acc[f].s[1] = sum[f][1];
        }
        if (mask & 2) {
            bus_render<NF, BUS_DEPTH>(sm, lane, 1, n_act, sum, one2, force_all);
#pragma unroll
            for (int f = 0; f < NF; ++f)
                acc[f].v[0] = sum[f][0], acc[f].v[1] = sum[f][1];
        }
        if (mask & 4) {
            bus_render<NF, BUS_DEPTH>(sm, lane, 2, n_act, sum, one2, force_all);
#pragma unroll
            for (int f = 0; f < NF; ++f)
                acc[f].p[0] = sum[f][0], acc[f].p[1] = sum[f][1];
        }
        return;
    }
    switch (mask) {
    case 1: core_tile_direct<NF, false, DEPTH, 1>(tg, sm, lane, 0, n_act, acc, one2); break;
    case 2: core_tile_direct<NF, false, DEPTH, 2>(tg, sm, lane, 0, n_act, acc, one2); break;
    case 3: core_tile_direct<NF, false, DEPTH, 3>(tg, sm, lane, 0, n_act, acc, one2); break;
    case 4: core_tile_direct<NF, false, DEPTH, 4>(tg, sm, lane, 0, n_act, acc, one2); break;
    case 5: core_tile_direct<NF, false, DEPTH, 5>(tg, sm, lane, 0, n_act, acc, one2); break;
    case 6: core_tile_direct<NF, false, DEPTH, 6>(tg, sm, lane, 0, n_act, acc, one2); break;
    default: core_tile_direct<NF, false, DEPTH, 7>(tg, sm, lane, 0, n_act, acc, one2); break;
    }
}

/* SLOTS > 0: TMA bulk copies through that many staging slots; -8 <= SLOTS < 0: per-lane cp.async
 * through -SLOTS slots; SLOTS <= -10: direct register loads -SLOTS - 10 wheels ahead */
template <int NF, bool ENV, int SLOTS>
__device__ __forceinline__ void core_tile(const DevTonegenTables *tg, TWarp &sm, int lane, int rec0,
                                          int cnt, CoreAcc (&acc)[NF], uint32_t &parity, const f32x2 one2)
{
    if (SLOTS > 0)
        core_tile_tma<NF, ENV, (SLOTS > 0 ? SLOTS : 2)>(tg, sm, lane, rec0, cnt, acc, parity, one2);
    else if (SLOTS > -10)
        core_tile_cpa<NF, ENV, ((SLOTS < 0 && SLOTS > -10) ? -SLOTS : 2)>(tg, sm, lane, rec0, cnt, acc, one2);
    else
        core_tile_direct<NF, ENV, (SLOTS <= -10 ? -SLOTS - 10 : 2)>(tg, sm, lane, rec0, cnt, acc, one2);
}

/* from lane = 4 consecutive samples to lane = sample of a 32-sample block (s[k] = sample
 * 32k + lane) through the idle wave staging slots; an empty active list leaves the buffers
 * zeroed (src/tonegen.cpp:3607-3612) */
__device__ __forceinline__ void core_end(const CoreAcc &a, TWarp &sm, int n_act, int lane, float (&s)[4],
                                         float (&v)[4], float (&p)[4])
{
    float4 *st = wslot(sm, 0, 0); /* 3 x 512 B of the (at least) 2 KB */
    *reinterpret_cast<ulonglong2 *>(&st[lane]) = make_ulonglong2(a.s[0], a.s[1]);
    *reinterpret_cast<ulonglong2 *>(&st[32 + lane]) = make_ulonglong2(a.v[0], a.v[1]);
    *reinterpret_cast<ulonglong2 *>(&st[64 + lane]) = make_ulonglong2(a.p[0], a.p[1]);
    __syncwarp();
    const float *fs = reinterpret_cast<const float *>(st);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        s[k] = fs[BLK * k + lane];
        v[k] = fs[128 + BLK * k + lane];
        p[k] = fs[256 + BLK * k + lane];
    }
    /* the slots are requested into again by the next fragment: these generic-proxy accesses
     * are ordered before its bulk copies */
    fence_async_smem();
    __syncwarp();
    if (n_act == 0) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
            s[k] = v[k] = p[k] = 0.0f;
    }
}

/* scanner vibrato of one 32-sample block (src/vibrato.cpp:365-411), lane = sample.
 * Everything integer is closed form in the sample index m: stator(m) = stator0 + m*inc,
 * outPos(m) = outPos0 + m.  Returns y (the value read at outPos(m)). */
__device__ __forceinline__ float vibrato_block(const DevTonegenTables *tg, TWarp &sm, int k, int lane,
                                               float x)
{
    const TgScalars &sc = sm.sc;
    const float fnorm = (float)(1.0 / 65536.0);
    const uint32_t *tab = tg->vib_offset_table[sc.vibTable];
    const uint32_t inc = tg->vib_stator_increment;
    const uint32_t m = (uint32_t)(BLK * k + lane);
    const uint32_t stator = (sc.vibStator + m * inc) & 0x07FFFFFFu;
    const uint32_t op = (sc.vibOutPos + m) & VBUF_MASK;
    const uint32_t j = ((op << 16) + __ldg(&tab[stator >> 16])) & 0x03FFFFFFu;
    const uint32_t h = j >> 16;
    const float f = __fmul_rn(fnorm, (float)(j & 0xFFFFu));
    const float g = __fmul_rn(f, x);
    const float va = __fsub_rn(x, g), vb = g;
    /* slots ahead of the block's first read; the parallel pass needs slots that never go
     * backwards, at most two samples per slot, and slots that stay ahead of (or at) their
     * own sample's read position */
    const uint32_t rel = (h - op + (uint32_t)lane) & VBUF_MASK;
    int bad = (rel < (uint32_t)lane) || (rel + 1 >= (uint32_t)TBF_VRING);
    uint32_t dh;
    const RunBits rb = run_bits(h, VBUF_MASK, lane, dh);
    if (lane >= 1)
        bad |= dh >= (TBF_VIB_BUF / 2);
    bad |= ((rb.e >> lane) & ((rb.e << 1) >> lane) & 1u) != 0u; /* three samples on one slot */
    bad = __any_sync(FULL, bad);
    float y;
    if (!bad) {
        ordered_pass_bits(h, rb, va, vb, sm.vring, VRING_MASK, lane);
        __syncwarp();
        y = sm.vring[op & VRING_MASK];
        sm.vring[op & VRING_MASK] = 0.0f;
    } else {
        /* the reference's loop as written, one lane (offset tables with backward steps);
         * the other lanes hand their sample over by shuffle */
        y = 0.0f;
        for (int i = 0; i < BLK; ++i) {
            const uint32_t hh = __shfl_sync(FULL, h, i);
            const float ai = __shfl_sync(FULL, va, i), bi = __shfl_sync(FULL, vb, i);
            float yi = 0.0f;
            if (lane == 0) {
                const uint32_t kk = (hh + 1) & VBUF_MASK;
                const uint32_t oo = (sc.vibOutPos + (uint32_t)(BLK * k + i)) & VBUF_MASK;
                sm.vring[hh & VRING_MASK] = __fadd_rn(sm.vring[hh & VRING_MASK], ai);
                sm.vring[kk & VRING_MASK] = __fadd_rn(sm.vring[kk & VRING_MASK], bi);
                yi = sm.vring[oo & VRING_MASK];
                sm.vring[oo & VRING_MASK] = 0.0f;
            }
            yi = __shfl_sync(FULL, yi, 0);
            if (lane == i)
                y = yi;
        }
    }
    __syncwarp();
    return y;
}

#ifdef TBF_DEBUG_CHECKS
__device__ void check_list(InstState *st, TWarp &sm, int lane, int where, int f, int a)
{
    __syncwarp();
    const int n = sm.sc.activeOscLEnd;
    int bad = 0;
    for (int i = lane; i < n; i += 32) {
        const int w = st->activeOscList[i];
        if (w < 1 || w > TBF_NOF_WHEELS || st->osc[w].aclPos != i) {
            bad = 1;
            if (blockIdx.x == 0 && threadIdx.x < 32)
                printf("  list[%d] = %d aclPos %d rflags %x ref %d\n", i, w, (w >= 0 && w <= TBF_NOF_WHEELS) ? st->osc[w].aclPos : -99,
                       (w >= 0 && w <= TBF_NOF_WHEELS) ? st->osc[w].rflags : 0u, (w >= 0 && w <= TBF_NOF_WHEELS) ? st->osc[w].refCount : 0);
        }
    }
    if (n < 0 || n > TBF_NOF_WHEELS)
        bad = 1;
    const int bad_local = bad;
    bad = __any_sync(FULL, bad);
    if (bad) {
        if (blockIdx.x == 0 && threadIdx.x < 32)
            printf("  lane %d bad_local %d n %d active %08x\n", lane, bad_local, n, __activemask());
        if (lane == 0)
            printf("check_list: slot %d where %d frag %d a %d lend %d\n",
                   (int)(blockIdx.x * T_WARPS + threadIdx.x / 32), where, f, a, n);
        __syncwarp();
        __trap();
    }
}
#endif
/* scanner, percussion high-pass and the hand-over to stage R of one fragment whose three
 * buffers are summed (src/tonegen.cpp:3699-3732) */
__device__ __forceinline__ void post_fragment(const DevTonegenTables *tg, TWarp &sm, int lane, uint32_t f,
                                              float *u_row, float *php_row, const float (&swl)[4],
                                              const float (&vib)[4], const float (&prc)[4])
{
    TgScalars &sc = sm.sc;
    const bool vib_on = sm.vib_on != 0, perc_on = sm.perc_on != 0;
    const float pz = sc.pz;
    float *u_out = u_row + (size_t)f * TBF_FRAGMENT;
    float *php_out = php_row + (size_t)f * TBF_FRAGMENT;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        float uu = swl[k];
        if (vib_on) {
            const float y = vibrato_block(tg, sm, k, lane, vib[k]);
            const float vy = sc.vibMixed ? __fmul_rn(__fadd_rn(vib[k], y), (float)0.7071067811865475) : y;
            uu = __fadd_rn(uu, vy);
        }
        u_out[BLK * k + lane] = uu;
        if (perc_on) { /* in-place differencer, src/tonegen.cpp:3719-3732 */
            float prev = __shfl_up_sync(FULL, prc[k], 1);
            float carry = pz; /* sample -1 of the fragment */
            if (k > 0)
                carry = __shfl_sync(FULL, prc[k > 0 ? k - 1 : 0], 31);
            if (lane == 0)
                prev = carry;
            php_out[BLK * k + lane] = __fsub_rn(prev, prc[k]);
        }
    }
    const float prc_last = __shfl_sync(FULL, prc[3], 31);
    if (lane == 0) {
        if (vib_on) {
            sc.vibStator = (sc.vibStator + TBF_FRAGMENT * tg->vib_stator_increment) & 0x07FFFFFFu;
            sc.vibOutPos = (sc.vibOutPos + TBF_FRAGMENT) & VBUF_MASK;
        }
        if (perc_on) {
            sc.outputGain = __fmul_rn(sc.swellPedalGain, sc.percDrawbarGain);
            sc.pz = prc_last;
        }
    }
    __syncwarp();
}

/* statistics and the reference-UB flag of a quiet fragment */
__device__ __forceinline__ void quiet_stats(TgScalars &sc, int lane, int n_act, int my_ins)
{
    const int n_ins = __reduce_add_sync(FULL, my_ins);
    if (lane == 0) {
        sc.statFragments += 1;
        sc.statActiveWheels += (unsigned long long)n_act;
        if (n_ins > 256)
            sc.errorFlags |= TBF_IFLAG_COREPGM_OVERFLOW;
    }
}

/* 7 CTAs = 28 warps per SM at 72 registers: measured ahead of 8 CTAs at 64 registers, where the
 * pair loop spills two of its 24 running sums (profiles/r02_knob_sweep.md) */
#ifndef T_MIN_CTAS
#define T_MIN_CTAS 7
#endif
/* SLOTS > 0: the wheels' samples arrive by TMA bulk copies through that many shared-memory staging
 * slots; SLOTS < 0: by per-lane cp.async through -SLOTS slots */
/* direct loads: 8 registers per wheel in flight */
#ifndef T_DIRECT_CTAS
#define T_DIRECT_CTAS(depth) ((depth) <= 3 ? 7 : 5)
#endif
template <int SLOTS>
__global__ void __launch_bounds__(T_WARPS * 32, SLOTS <= -10 ? T_DIRECT_CTAS(-SLOTS - 10) : (SLOTS >= -3 && SLOTS <= 3) ? T_MIN_CTAS : 4) tbf_stage_t(const RenderParams p)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t slot = blockIdx.x * T_WARPS + warp;
    if (slot >= p.n_instances)
        return; /* warps are independent: no CTA-wide barrier anywhere in this kernel */
    constexpr bool TMA = SLOTS > 0;
    constexpr int N_SLOTS = SLOTS <= -10 ? 2 : SLOTS < 0 ? -SLOTS : SLOTS; /* direct loads: the two slots the transposes use */
    constexpr size_t per_warp = sizeof(TWarp) + (size_t)N_SLOTS * 1024;
    TWarp &sm = *reinterpret_cast<TWarp *>(smem_raw + per_warp * warp);
    const uint32_t id = p.ids ? p.ids[slot] : p.id_base + slot;
    InstState *st = p.states + id;
    TgScalars &sc = sm.sc;

    for (int i = lane; i < (int)(sizeof(TgScalars) / 4); i += 32)
        reinterpret_cast<uint32_t *>(&sm.sc)[i] = reinterpret_cast<const uint32_t *>(&st->tg)[i];
    if (TMA && lane == 0) {
#pragma unroll
        for (int j = 0; j < SLOTS; ++j)
            mbar_init(smem_u32(&sm.mbar[j]), 1u);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    uint32_t parity = 0u; /* bit s: parity of staging slot s's next completed fill */
    const f32x2 one2 = pack2(p.one, p.one); /* 1.0f the compiler cannot see: add2_via_fma() */
    const DevTonegenTables *tg = p.tg + sc.tgTables;
    if (lane == 0)
        sm.wave4 = tg->wave4;
    bool no_neg_zero = true; /* no -0 in the scanner's delay line (quiet_sum) */
    for (int s = lane; s < TBF_VRING; s += 32) {
        const uint32_t g = (sc.vibOutPos + (uint32_t)s) & VBUF_MASK;
        const float y = st->vibBuffer[g];
        no_neg_zero &= __float_as_uint(y) != 0x80000000u;
        sm.vring[g & VRING_MASK] = y;
    }
    no_neg_zero = __all_sync(FULL, no_neg_zero);
    constexpr bool DIRECT = SLOTS <= -10;
    constexpr int BUS_DEPTH = 3; /* list entries in flight in the bus-by-bus loops */
    constexpr int DIRECT_DEPTH = DIRECT ? -SLOTS - 10 : 2;
    bool lists_valid = false;
    uint64_t ev_cur = p.ev_cursor_t[slot];
    const uint64_t ev_end = p.ev_start[slot + 1];
    float *u_row = p.u + (size_t)slot * p.pitch;
    float *php_row = p.php + (size_t)slot * p.pitch;
    FragCtl *ctl_row = p.ctl + (size_t)slot * p.n_fragments;
    __syncwarp();

    bool cache_valid = false; /* the records in shared memory describe the last fragment */
#ifdef TBF_DEBUG_CHECKS
    int dbg_fast = 0;
#endif
    int n_act_cached = 0;
    /* the records the last launch left behind (InstState::recG ...) */
    if (st->rec_valid && !(p.flags & 0x100u)) {
        n_act_cached = st->rec_count;
        for (int i = lane; i < n_act_cached; i += 32) {
            sm.recG[i] = *reinterpret_cast<const float4 *>(st->recG[i]);
            sm.recL[i] = st->recL[i];
            sm.recW[i] = st->recW[i];
        }
        cache_valid = true;
        __syncwarp();
    }
    uint32_t f = 0;
    while (f < p.n_fragments) {
        /* -- control events that precede this fragment -- */
        bool ev_applied = false;
        uint32_t next_ev_fragment = 0xFFFFFFFFu; /* batch fragment of the first event still to come */
        while (ev_cur < ev_end) {
            const tbf_event *evp = p.events + ev_cur;
            if (evp->fragment > p.frag_base + f) {
                next_ev_fragment = evp->fragment;
                break;
            }
            const int type = evp->type, a = evp->a, b = evp->b;
            ev_applied = true;
            if (type == TBF_EV_KEY_ON || type == TBF_EV_KEY_OFF) {
                if (a >= 0 && a < TBF_MAX_KEYS) {
                    if (type == TBF_EV_KEY_ON)
                        key_on_warp(sm, st, tg, a, lane);
                    else if (st->activeKeys[a])
                        key_off_warp(sm, st, tg, a, lane);
                }
            } else if (lane == 0) {
                apply_scalar_event(sc, tg, type, a, b);
            }
            __syncwarp();
#ifdef TBF_DEBUG_CHECKS
            check_list(st, sm, lane, type, (int)f, a);
#endif
            ++ev_cur;
        }
        /* -- routing, key compression target, what stage R needs -- */
        /* a quiet fragment followed by another one: both are summed in one pass over the records */
        const bool pair_ok = f + 1 < p.n_fragments && next_ev_fragment > p.frag_base + f + 1 &&
                             !(p.flags & 0x200u);
        if (lane == 0) {
            int kdc = sc.keyDownCount;
            if (kdc > 127 || kdc < 0) { /* reference reads past keyCompTable[128] */
                sc.errorFlags |= TBF_IFLAG_KEYCOMP_OOB;
                kdc = kdc < 0 ? 0 : 127;
            }
            sm.recompute_routing = (sc.oldRouting != sc.newRouting);
            if (sm.recompute_routing)
                sc.oldRouting = sc.newRouting;
            sm.draw_bar_change = sc.drawBarChange;
            sm.any_removed = sm.any_env = 0;
            sm.vib_on = (sc.oldRouting & RT_VIB) != 0;
            sm.perc_on = (sc.oldRouting & RT_PERC) != 0;
            FragCtl c;
            c.key_comp = tg->key_comp_table[kdc];
            c.gain = sm.perc_on ? __fmul_rn(sc.swellPedalGain, sc.percDrawbarGain) : sc.swellPedalGain;
            c.decay = sc.percEnvGainDecay;
            c.reset = sc.percEnvGainReset;
            c.flags = (sm.perc_on ? FC_PERC_ON : 0u) | (sc.upperKeyCount == 0 ? FC_PERC_RESET : 0u);
            c.pad_[0] = c.pad_[1] = c.pad_[2] = 0u;
            ctl_row[f] = c;
            /* nothing stage R looks at changes without an event */
            if (pair_ok && cache_valid && !ev_applied && !sm.recompute_routing && !sm.draw_bar_change)
                ctl_row[f + 1] = c;
        }
        __syncwarp();

        /* -- active-list pass and additive synthesis -- */
        float swl[4], vib[4], prc[4];
        const int n_act = sc.activeOscLEnd;
        const bool quiet = cache_valid && !ev_applied && !sm.recompute_routing && !sm.draw_bar_change;
        if (quiet && pair_ok) {
            /* two quiet fragments: the records of the last fragment, one and two fragments further */
#ifdef TBF_DEBUG_CHECKS
            dbg_fast += 2;
#endif
            int ins_a = 0, ins_b = 0;
            CoreAcc acc[2];
            core_begin(acc[0]);
            core_begin(acc[1]);
            fast_pass<2>(tg, sm, lane, n_act, ins_a, ins_b);
            if (DIRECT && !lists_valid) {
                __syncwarp();
                bus_lists_build(sm, lane, n_act);
                lists_valid = true;
            }
            __syncwarp();
            if (DIRECT)
                quiet_sum<2, DIRECT_DEPTH, BUS_DEPTH>(tg, sm, lane, n_act, acc, one2, no_neg_zero, p.flags);
            else
                core_tile<2, false, SLOTS>(tg, sm, lane, 0, n_act, acc, parity, one2);
            __syncwarp();
            quiet_stats(sc, lane, n_act, ins_a);
            quiet_stats(sc, lane, n_act, ins_b);
            core_end(acc[0], sm, n_act, lane, swl, vib, prc);
            __syncwarp();
            post_fragment(tg, sm, lane, f, u_row, php_row, swl, vib, prc);
            core_end(acc[1], sm, n_act, lane, swl, vib, prc);
            __syncwarp();
            post_fragment(tg, sm, lane, f + 1, u_row, php_row, swl, vib, prc);
            f += 2;
            continue;
        }
        {
            int my_ins = 0, unused = 0;
            CoreAcc acc[1];
            core_begin(acc[0]);
            if (quiet) {
                /* quiet fragment: the records of the last fragment, one fragment further */
#ifdef TBF_DEBUG_CHECKS
                ++dbg_fast;
#endif
                fast_pass<1>(tg, sm, lane, n_act, my_ins, unused);
                if (DIRECT && !lists_valid) {
                    __syncwarp();
                    bus_lists_build(sm, lane, n_act);
                    lists_valid = true;
                }
                __syncwarp();
                if (DIRECT)
                    quiet_sum<1, DIRECT_DEPTH, BUS_DEPTH>(tg, sm, lane, n_act, acc, one2, no_neg_zero, p.flags);
                else
                    core_tile<1, false, SLOTS>(tg, sm, lane, 0, n_act, acc, parity, one2);
                __syncwarp();
                quiet_stats(sc, lane, n_act, my_ins);
            } else {
                lists_valid = false; /* new records */
                if (cache_valid) /* the pass below reads the positions from HBM */
                    flush_positions(st, sm, lane, n_act_cached);
                __syncwarp();
                const bool persist = n_act <= REC_CAP;
                for (int base = 0; base < n_act; base += REC_TILE) {
                    const int cnt = min(REC_TILE, n_act - base);
                    const int rec0 = persist ? base : 0;
                    active_list_tile(st, tg, sm, lane, base, cnt, rec0, my_ins);
                    fence_async_smem(); /* rec_d shares the staging slots' B halves */
                    __syncwarp();
                    core_tile<1, true, SLOTS>(tg, sm, lane, rec0, cnt, acc, parity, one2);
                    __syncwarp();
                }
                active_list_finish(st, sm, lane, n_act, my_ins);
#ifdef TBF_DEBUG_CHECKS
                check_list(st, sm, lane, 1000, (int)f, n_act);
#endif
                /* the records can be reused if they are all plain and the list kept its order */
                cache_valid = persist && !sm.any_env && !sm.any_removed && !sm.recompute_routing &&
                              !sm.draw_bar_change; /* (their records still carry the old gains) */
                n_act_cached = n_act;
                if (p.flags & 0x100u) /* debugging: always take the full pass */
                    cache_valid = false;
            }
            core_end(acc[0], sm, n_act, lane, swl, vib, prc);
        }
        __syncwarp();
        post_fragment(tg, sm, lane, f, u_row, php_row, swl, vib, prc);
        f += 1;
    }

    /* ---- write the staged state back ---- */
#ifdef TBF_DEBUG_CHECKS
    if (slot == 5 && lane == 0)
        printf("slot 5: %d of %u fragments on the fast path, lend %d\n", dbg_fast, p.n_fragments, sc.activeOscLEnd);
#endif
    if (cache_valid) {
        flush_positions(st, sm, lane, n_act_cached);
        for (int i = lane; i < n_act_cached; i += 32) {
            *reinterpret_cast<float4 *>(st->recG[i]) = sm.recG[i];
            st->recL[i] = sm.recL[i];
            st->recW[i] = sm.recW[i];
        }
    }
    if (lane == 0) {
        st->rec_valid = cache_valid ? 1 : 0;
        st->rec_count = n_act_cached;
        p.ev_cursor_t[slot] = ev_cur;
    }
    for (int i = lane; i < (int)(sizeof(TgScalars) / 4); i += 32)
        reinterpret_cast<uint32_t *>(&st->tg)[i] = reinterpret_cast<const uint32_t *>(&sm.sc)[i];
    /* outside the live window every ring slot is zero (cleared when read) */
    for (int s = lane; s < TBF_VIB_BUF; s += 32) {
        const uint32_t g = (sc.vibOutPos + (uint32_t)s) & VBUF_MASK;
        st->vibBuffer[g] = (s < TBF_VRING) ? sm.vring[g & VRING_MASK] : 0.0f;
    }
}

/* ================================================================== */
/* stage R: lane = instance                                            */
/* ================================================================== */

__device__ void use_rev_option(RampScalars &rp, const DevWhirlTables *wh, int n, int signals)
{
    const int i = n % 9;
    rp.hornTarget = wh->rev_horn_target[i];
    rp.drumTarget = wh->rev_drum_target[i];
    if (rp.hornIncr < rp.hornTarget)
        rp.hornAcDc = 1;
    else if (rp.hornTarget < rp.hornIncr)
        rp.hornAcDc = -1;
    if (rp.drumIncr < rp.drumTarget)
        rp.drumAcDc = 1;
    else if (rp.drumTarget < rp.drumIncr)
        rp.drumAcDc = -1;
    if (signals & 2) {
        const int hr = (n / 3) % 3;
        rp.revSelect = (hr == 2) ? 2 : (hr == 1) ? 0 : 1; /* WHIRL_FAST / SLOW / STOP */
    }
}

/* per-call speed ramp and brake logic of ONE rotor, double precision
 * (src/whirl.cpp:1219-1374; the horn and drum halves there differ only in the constants
 * passed here).  Returns non-zero when the brake is engaged (incr is zeroed after the
 * sample loop, src/whirl.cpp:1635-1636). */
__device__ int rotor_ramp(const double SR, const bool is_horn, int *acdc, double *incr, double *angle,
                          const double target, const double brake_pos, const double oml_acc,
                          const double oml_dec)
{
    const float hardstop_rpm = is_horn ? 10.f : 8.f;
    const float brake_limit_rpm = is_horn ? 60.f : 100.f;
    const float nsmp = (float)TBF_FRAGMENT;
    const double target_pos =
        is_horn ? fmod(__dsub_rn(1.25, brake_pos), 1.0) : fmod(__dadd_rn(brake_pos, .75), 1.0);

    if (*acdc) {
        int flywheel = 0;
        const float hardstop = (float)((double)hardstop_rpm / ((double)60.f * SR));
        if (brake_pos > 0 && target == 0 && *incr > 0 && *incr < (double)hardstop) {
            if (fabs(__dsub_rn(*angle, target_pos)) < (2.0 / TBF_WHIRL_DISPLC)) {
                *angle = target_pos;
                *incr = 0;
            } else {
                const float minspeed = (float)((double)3.f / ((double)60.f * SR));
                const float diffinc = (float)(
                    fmod(__dsub_rn(__dadd_rn(1., target_pos), *angle), 1.0) / (double)nsmp);
                if (*incr > (double)diffinc)
                    *incr = diffinc;
                else if (*incr < (double)minspeed)
                    *incr = minspeed;
                flywheel = 1;
            }
        }
        if (!flywheel) {
            const double oml = (*acdc > 0) ? oml_acc : oml_dec;
            *incr = __dadd_rn(*incr, __dmul_rn(oml, __dsub_rn(target, *incr)));
        }
        if (fabs(__dsub_rn(target, *incr)) < (.05 / ((double)60.f * SR))) {
            *acdc = 0;
            *incr = target;
        }
    }

    int brake = 0;
    if (brake_pos > 0) {
        if (!*acdc && *incr == 0 && *angle != target_pos) {
            brake = 1;
            if (fabs(__dsub_rn(*angle, target_pos)) < (2.0 / TBF_WHIRL_DISPLC)) {
                *angle = target_pos;
            } else {
                const float limit = (float)((double)brake_limit_rpm / (60. * SR));
                *incr = fmod(__dsub_rn(__dadd_rn(1., target_pos), *angle), 1.0) / (double)nsmp;
                if (*incr > (double)limit)
                    *incr = limit;
            }
        }
    }
    return brake;
}

/* rotor angle recurrence, angle = fmod(angle + incr, 1.0) in double
 * (src/whirl.cpp:1618-1619); for 0 <= t < 2 fmod(t,1) is t or t-1 exactly */
__device__ __forceinline__ double angle_step(double a, double incr)
{
    const double t = __dadd_rn(a, incr);
    if (t >= 0.0 && t < 1.0)
        return t;
    if (t >= 1.0 && t < 2.0)
        return __dsub_rn(t, 1.0);
    return fmod(t, 1.0);
}

/* DF-II biquad, float, source operation order (EQ_IIR, src/whirl.cpp:1479-1485) */
__device__ __forceinline__ float eq_iir(const float (&c)[5], float &z0, float &z1, float x)
{
    const float t = __fsub_rn(__fsub_rn(x, __fmul_rn(c[0], z0)), __fmul_rn(c[1], z1));
    const float y = __fadd_rn(__fadd_rn(__fmul_rn(t, c[2]), __fmul_rn(c[3], z0)), __fmul_rn(c[4], z1));
    z1 = z0;
    z0 = t;
    return y;
}

/* ---- 2-D TMA tiles for the lane = organ kernels (stages R and D) ----
 *
 * The stage streams are [slot][sample] planes; a lane = organ kernel wants the transpose.  A tensor
 * map over the whole scratch (rows = every plane's slots, back to back; tbf_api.cu) lets ONE lane
 * move a 32 organs x 32 samples box with one instruction (cp.async.bulk.tensor, SASS UTMALDG /
 * UTMASTG) instead of 32 lanes x 32 four-byte cp.async and a 32-trip transposed store loop.  The box
 * lands row-major, 128 bytes per organ, with the 128-byte swizzle: 16-byte chunk c of row r sits at
 * chunk c ^ (r & 7), so the eight lanes of a quarter-warp reading "their" row's chunk c with one
 * LDS.128 touch all 32 banks once. */
__device__ __forceinline__ void tma_tile_load(uint32_t dst, const CUtensorMap *map, uint32_t col, uint32_t row, uint32_t bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
                 "l"(map), "r"(col), "r"(row), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void tma_tile_store(const CUtensorMap *map, uint32_t col, uint32_t row, uint32_t src)
{
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%1, %2}], [%3];" ::"l"(map), "r"(col),
                 "r"(row), "r"(src)
                 : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
/* the committed stores have READ their shared-memory source (it may be overwritten) */
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
/* all but the most recently committed store have read their shared-memory source */
__device__ __forceinline__ void tma_store_wait_read_but_one() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
/* byte offset of 16-byte chunk c (4 floats / 2 doubles) of row r inside a 32-row, 128-byte-row tile */
__device__ __forceinline__ uint32_t tile_chunk(int r, int c) { return (uint32_t)(r * 128 + ((c ^ (r & 7)) << 4)); }

#define TILE_BYTES 4096 /* 32 rows x 128 B */
struct RSmem {
    float tin[2][2][BLK * TILE_PITCH]; /* [buffer][u, php] (row = instance, col = sample);
                                          filled by cp.async one block ahead of its use   */
    float tout[2][BLK * TILE_PITCH];  /* dx0 | mono, hx0                                 */
    double tang[2][BLK * TILE_PITCH]; /* horn, drum angle                                */
    unsigned long long csum[BLK][BLK]; /* [row = instance][lane]: checksum partial sums.  LAST:
                                          launched without it unless stage R forms the checksum,
                                          so that the CTA takes one stage T / W slot, not two */
};

/* request one 32 x 32 block of u and php: element (row r, sample lane) by cp.async */
__device__ __forceinline__ void r_tile_fetch(const RenderParams &p, float (*tile)[BLK * TILE_PITCH],
                                             uint32_t slot0, uint32_t rows, size_t col0, int lane)
{
    for (uint32_t r = 0; r < rows; ++r) {
        const size_t g = (size_t)(slot0 + r) * p.pitch + col0 + lane;
        __pipeline_memcpy_async(&tile[0][r * TILE_PITCH + lane], p.u + g, 4);
        __pipeline_memcpy_async(&tile[1][r * TILE_PITCH + lane], p.php + g, 4);
    }
}

/* Two warps per 32 organs, lane = organ in both.  The rotor angles do not depend on the audio,
 * so the warps never meet: role 0 mixes down, runs the horn biquads and writes dx0 / hx0 (or
 * L, R when the Leslie is bypassed); role 1 applies the rotor events, ramps the speeds and writes
 * the two angle streams.  Stage R is one serial chain per organ and sits on the critical path of
 * the stage schedule while it shares its SM with 20 stage T / W warps: halving what one warp has
 * to issue shortens that chain. */
__global__ void __launch_bounds__(64) tbf_stage_r(const RenderParams p)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    RSmem &sm = *reinterpret_cast<RSmem *>(smem_raw);
    const int lane = threadIdx.x & 31;
    const bool rotors = (threadIdx.x >> 5) != 0; /* warp-uniform role */
    const uint32_t slot0 = blockIdx.x * 32;
    const uint32_t slot = slot0 + lane;
    const bool valid = slot < p.n_instances;
    const uint32_t rows = min(32u, p.n_instances - slot0);
    const bool whirl_off = (p.flags & TBF_RENDER_WHIRL_BYPASS) != 0;
    const bool store = !(p.flags & TBF_RENDER_NO_OUTPUT);
    /* Leslie only: plane u holds the caller's mono input (whirlProc's inbuffer) and the
     * tone generator's mixdown state (percussion gain, key compression) stays as it is */
    const bool ext_in = (p.flags & TBF_RENDER_WHIRL_INPUT) != 0;
    if (rotors && whirl_off)
        return;

    InstState *st = nullptr;
    const DevWhirlTables *wh = p.wh;
    RampScalars rp;
    if (valid) {
        const uint32_t id = p.ids ? p.ids[slot] : p.id_base + slot;
        st = p.states + id;
        rp = st->rp;
        wh = p.wh + st->tg.whTables;
    } else {
        rp = RampScalars();
    }
    const bool whirl_run = valid && !whirl_off && !wh->bypass;
    uint32_t new_flags = 0;
    uint64_t ev_cur = 0, ev_end = 0;
    if (valid) {
        ev_cur = p.ev_cursor_r[slot];
        ev_end = p.ev_start[slot + 1];
    }
    const uint32_t n_blocks = p.n_fragments * (TBF_FRAGMENT / BLK);

    if (rotors) {
        /* ---------------- role 1: rotor events, speed ramps, angle streams ---------------- */
        double hang = rp.hornAngle, dang = rp.drumAngle;
        const double SR = wh->sample_rate;
        for (uint32_t f = 0; f < p.n_fragments; ++f) {
            while (ev_cur < ev_end) {
                const tbf_event *evp = p.events + ev_cur;
                if (evp->fragment > p.frag_base + f)
                    break;
                const int type = evp->type, a = evp->a;
                if (type == TBF_EV_REV_OPTION) {
                    if (a < 0)
                        new_flags |= TBF_IFLAG_BAD_EVENT;
                    else
                        use_rev_option(rp, wh, a, 2);
                } else if (type == TBF_EV_REV_SELECT) {
                    if (a < 0) {
                        new_flags |= TBF_IFLAG_BAD_EVENT;
                    } else {
                        rp.revSelect = a % 3;
                        use_rev_option(rp, wh, wh->revselects[rp.revSelect], 1);
                    }
                } else if (type == TBF_EV_WHIRL_BRAKE || type == TBF_EV_WHIRL_RAMP) {
                    const int b = evp->b;
                    if (a < 0 || a > (type == TBF_EV_WHIRL_BRAKE ? 1 : 3) || b < 0 || b > 127) {
                        new_flags |= TBF_IFLAG_BAD_EVENT;
                    } else if (type == TBF_EV_WHIRL_BRAKE) {
                        /* w->hnBrakePos = (double)uc / 127.0, src/whirl.cpp:883-893 */
                        rp.brakePos[a] = __ddiv_rn((double)b, 127.0);
                    } else {
                        rp.oml[a] = wh->ramp_one_minus_l[a][b];
                    }
                }
                ++ev_cur;
            }
            /* -- per-call speed ramps (src/whirl.cpp:1219-1374) -- */
            int hbrake = 0, dbrake = 0;
            double hincr = rp.hornIncr, dincr = rp.drumIncr;
            if (whirl_run) {
                hbrake = rotor_ramp(SR, true, &rp.hornAcDc, &hincr, &hang, rp.hornTarget, rp.brakePos[0],
                                    rp.oml[0], rp.oml[1]);
                dbrake = rotor_ramp(SR, false, &rp.drumAcDc, &dincr, &dang, rp.drumTarget, rp.brakePos[1],
                                    rp.oml[2], rp.oml[3]);
            }
            /* fmod(t, 1.0) of a sum in [0, 2) is t or t - 1, both exact */
            const bool hfast = hang >= 0.0 && hang < 1.0 && hincr >= 0.0 && hincr < 1.0;
            const bool dfast = dang >= 0.0 && dang < 1.0 && dincr >= 0.0 && dincr < 1.0;
            for (int blk = 0; blk < TBF_FRAGMENT / BLK; ++blk) {
                const size_t col0 = ((size_t)f * (TBF_FRAGMENT / BLK) + (size_t)blk) * BLK;
                if (whirl_run) {
#pragma unroll 4
                    for (int s = 0; s < BLK; ++s) {
                        sm.tang[0][lane * TILE_PITCH + s] = hang;
                        sm.tang[1][lane * TILE_PITCH + s] = dang;
                        if (hfast) {
                            const double t = __dadd_rn(hang, hincr);
                            hang = (t >= 1.0) ? __dsub_rn(t, 1.0) : t;
                        } else {
                            hang = angle_step(hang, hincr);
                        }
                        if (dfast) {
                            const double t = __dadd_rn(dang, dincr);
                            dang = (t >= 1.0) ? __dsub_rn(t, 1.0) : t;
                        } else {
                            dang = angle_step(dang, dincr);
                        }
                    }
                }
                __syncwarp();
#pragma unroll 4
                for (uint32_t r = 0; r < rows; ++r) {
                    const size_t g = (size_t)(slot0 + r) * p.pitch + col0 + lane;
                    p.hang[g] = sm.tang[0][r * TILE_PITCH + lane];
                    p.dang[g] = sm.tang[1][r * TILE_PITCH + lane];
                }
                __syncwarp();
            }
            if (whirl_run) {
                rp.hornIncr = hbrake ? 0.0 : hincr;
                rp.drumIncr = dbrake ? 0.0 : dincr;
            }
        }
        if (valid) {
            RampScalars &o = st->rp; /* this role's fields only: role 0 stores the others */
            o.hornAcDc = rp.hornAcDc;
            o.drumAcDc = rp.drumAcDc;
            o.revSelect = rp.revSelect;
            o.hornAngle = hang;
            o.drumAngle = dang;
            o.hornIncr = rp.hornIncr;
            o.drumIncr = rp.drumIncr;
            o.hornTarget = rp.hornTarget;
            o.drumTarget = rp.drumTarget;
            o.brakePos[0] = rp.brakePos[0];
            o.brakePos[1] = rp.brakePos[1];
#pragma unroll
            for (int k = 0; k < 4; ++k)
                o.oml[k] = rp.oml[k];
            if (new_flags)
                atomicOr(&o.errorFlags, new_flags);
        }
        return;
    }

    /* ---------------- role 0: mixdown, horn biquads, dx0 / hx0 ---------------- */
    float cA[5], cB[5];
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        cA[k] = rp.iir[0][k];
        cB[k] = rp.iir[1][k];
    }
    float zA0 = rp.iirz[0][0], zA1 = rp.iirz[0][1], zB0 = rp.iirz[1][0], zB1 = rp.iirz[1][1];
    float pg = rp.percEnvGain, kc = rp.keyCompLevel;
    if (p.checksums && whirl_off) /* csum lies past the launch's shared memory otherwise */
        for (int r = 0; r < BLK; ++r)
            sm.csum[r][lane] = 0ull;
    r_tile_fetch(p, sm.tin[0], slot0, rows, 0, lane);
    __pipeline_commit();
    __syncwarp();

    for (uint32_t f = 0; f < p.n_fragments; ++f) {
        /* -- horn filter events that precede this fragment (the rotor's belong to role 1) -- */
        while (ev_cur < ev_end) {
            const tbf_event *evp = p.events + ev_cur;
            if (evp->fragment > p.frag_base + f)
                break;
            if (evp->type == TBF_EV_HORN_FILTER) {
                const int a = evp->a, k = a & 7;
                if (a < 0 || a > 15 || k > 4) {
                    new_flags |= TBF_IFLAG_BAD_EVENT;
                } else {
                    const float v = __int_as_float(evp->b);
                    rp.iir[a >> 3][k] = v;
#pragma unroll
                    for (int j = 0; j < 5; ++j) {
                        if (a == j)
                            cA[j] = v;
                        if (a == 8 + j)
                            cB[j] = v;
                    }
                }
            }
            ++ev_cur;
        }
        FragCtl c;
        c.key_comp = c.gain = c.decay = c.reset = 0.0f;
        c.flags = 0;
        if (valid && !ext_in)
            c = p.ctl[(size_t)slot * p.n_fragments + f];
        const bool perc_on = (c.flags & FC_PERC_ON) != 0;
        /* (keyComp - level) / 128.f: dividing by 2^7 and multiplying by 2^-7 round identically */
        const float delta = __fmul_rn(__fsub_rn(c.key_comp, kc), 0.0078125f);

        for (int blk = 0; blk < TBF_FRAGMENT / BLK; ++blk) {
            const uint32_t gb = f * (TBF_FRAGMENT / BLK) + (uint32_t)blk;
            const size_t col0 = (size_t)gb * BLK;
            const int cur = (int)(gb & 1u);
            /* tile in: row r = instance slot0 + r, this lane = sample col0 + lane; the
             * next block's tiles are requested before this block's serial loop */
            if (gb + 1 < n_blocks)
                r_tile_fetch(p, sm.tin[cur ^ 1], slot0, rows, (size_t)(gb + 1) * BLK, lane);
            __pipeline_commit();
            __pipeline_wait_prior(1);
            __syncwarp();
            const float *t_u = sm.tin[cur][0], *t_p = sm.tin[cur][1];
#pragma unroll 4
            for (int s = 0; s < BLK; ++s) {
                const float uu = t_u[lane * TILE_PITCH + s];
                float mono;
                if (ext_in) {
                    mono = uu;
                } else {
                    if (perc_on) {
                        const float pterm = __fmul_rn(t_p[lane * TILE_PITCH + s], pg);
                        mono = __fmul_rn(__fmul_rn(c.gain, kc), __fadd_rn(uu, pterm));
                        pg = __fmul_rn(pg, c.decay);
                    } else {
                        mono = __fmul_rn(__fmul_rn(c.gain, kc), uu);
                    }
                    kc = __fadd_rn(kc, delta);
                }
                if (whirl_run) {
                    const float x = (float)__dadd_rn((double)mono, 1e-14); /* DENORMAL_HACK */
                    const float ya = eq_iir(cA, zA0, zA1, x);
                    const float yb = eq_iir(cB, zB0, zB1, ya);
                    sm.tout[0][lane * TILE_PITCH + s] = x;
                    sm.tout[1][lane * TILE_PITCH + s] = yb;
                } else {
                    sm.tout[0][lane * TILE_PITCH + s] = mono;
                    sm.tout[1][lane * TILE_PITCH + s] = mono; /* bypassed Leslie: stage D mixes it */
                }
            }
            __syncwarp();
            /* tile out */
            if (whirl_off) {
                for (uint32_t r = 0; r < rows; ++r) {
                    const float m = sm.tout[0][r * TILE_PITCH + lane];
                    if (store) {
                        const size_t g = (size_t)(slot0 + r) * p.out_stride + col0 + lane;
                        p.out_l[g] = m;
                        p.out_r[g] = m;
                    }
                    if (p.checksums) {
                        const uint32_t pos = (uint32_t)((size_t)p.frag_base * TBF_FRAGMENT + col0 + lane);
                        sm.csum[r][lane] += sample_hash(pos, m, m);
                    }
                }
            } else {
#pragma unroll 4
                for (uint32_t r = 0; r < rows; ++r) {
                    const size_t g = (size_t)(slot0 + r) * p.pitch + col0 + lane;
                    p.dx0[g] = sm.tout[0][r * TILE_PITCH + lane];
                    p.hx0[g] = sm.tout[1][r * TILE_PITCH + lane];
                }
            }
            __syncwarp();
        }

        /* -- end of the fragment -- */
        if (whirl_run) {
            /* NaN scrub after the loop (EQ_IIR_NAN, src/whirl.cpp:1622-1625) */
            zA0 = isnan(zA0) ? 0.0f : zA0;
            zA1 = isnan(zA1) ? 0.0f : zA1;
            zB0 = isnan(zB0) ? 0.0f : zB0;
            zB1 = isnan(zB1) ? 0.0f : zB1;
        }
        if (c.flags & FC_PERC_RESET)
            pg = c.reset;
    }

    if (valid) {
        RampScalars &o = st->rp; /* this role's fields only */
        o.iirz[0][0] = zA0;
        o.iirz[0][1] = zA1;
        o.iirz[1][0] = zB0;
        o.iirz[1][1] = zB1;
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            o.iir[0][k] = rp.iir[0][k];
            o.iir[1][k] = rp.iir[1][k];
        }
        o.percEnvGain = pg;
        o.keyCompLevel = kc;
        if (new_flags)
            atomicOr(&o.errorFlags, new_flags);
        p.ev_cursor_r[slot] = ev_cur;
        if (p.checksums && whirl_off) {
            unsigned long long c = 0ull;
            for (int j = 0; j < BLK; ++j)
                c += sm.csum[lane][(j + lane) & (BLK - 1)];
            p.checksums[slot] += c; /* zeroed by the host; accumulates over chunks */
        }
    }
}

/* Stage R with TMA tiles (Leslie running: the bypass-only batch keeps tbf_stage_r): the same two
 * roles and the same arithmetic, the 32 x 32 tiles moved by one lane with cp.async.bulk.tensor and
 * read / written by every lane as 16-byte chunks of its own row (tile_chunk). */
struct RTmaSmem {
    float tin[2][2][TILE_BYTES / 4]; /* [buffer][u, php], filled by TMA one block ahead */
    float tout[3][TILE_BYTES / 4];   /* dx0, hx0; with the biquads in their own warp: dx0 tile 0, dx0 tile 1, hx0 */
    unsigned long long bar[6];       /* input tiles landed [2]; x tile written [2], read [2] */
    unsigned long long pad_[122];    /* keeps tang 1024-byte aligned */
    /* LAST, and only in the launch's shared memory when needed: the angle tiles when the rotors run
     * as a warp of this kernel, the checksum sums of a bypass-only batch (never both) */
    union {
        double tang[2][2 * TILE_BYTES / 8]; /* horn, drum angle: two boxes of 16 samples each */
        unsigned long long csum[BLK][BLK];
    };
};
static_assert(offsetof(RTmaSmem, tang) % 1024 == 0, "TMA tiles are 1024-byte aligned");
struct RotSmem {
    double tang[2][2 * TILE_BYTES / 8];
};
/* The rotor role of stage R: rotor events, per-fragment speed ramps (src/whirl.cpp:1219-1374), the
 * angle recurrences (:1618-1619) and the two angle streams, written as 2-D TMA boxes out of `tang`.
 * It needs nothing from stage T, so it runs either as a warp of tbf_stage_r_tma or, ahead of the audio
 * chain, as its own small kernel (tbf_stage_rot_tma).  WHICH: 0 both rotors in this warp, 1 the horn
 * only, 2 the drum only -- the two recurrences share nothing but the events, so two warps halve the
 * time the kernel's CTAs stay resident; each warp stores the state fields of its own rotor. */
template <int WHICH>
__device__ __forceinline__ void r_rotor_role(const RenderParams &p, const TileMaps &maps,
                                             double (*tang)[2 * TILE_BYTES / 8], const int lane,
                                             const uint32_t slot0, const bool valid, InstState *st,
                                             const DevWhirlTables *wh, RampScalars &rp, const bool whirl_run,
                                             uint64_t ev_cur, const uint64_t ev_end, uint64_t *cursor_out)
{
    uint32_t new_flags = 0;
    /* ---------------- role 1: rotor events, speed ramps, angle streams ---------------- */
    double hang = rp.hornAngle, dang = rp.drumAngle;
    const double SR = wh->sample_rate;
    for (uint32_t f = 0; f < p.n_fragments; ++f) {
        while (ev_cur < ev_end) {
            const tbf_event *evp = p.events + ev_cur;
            if (evp->fragment > p.frag_base + f)
                break;
            const int type = evp->type, a = evp->a;
            if (type == TBF_EV_REV_OPTION) {
                if (a < 0)
                    new_flags |= TBF_IFLAG_BAD_EVENT;
                else
                    use_rev_option(rp, wh, a, 2);
            } else if (type == TBF_EV_REV_SELECT) {
                if (a < 0) {
                    new_flags |= TBF_IFLAG_BAD_EVENT;
                } else {
                    rp.revSelect = a % 3;
                    use_rev_option(rp, wh, wh->revselects[rp.revSelect], 1);
                }
            } else if (type == TBF_EV_WHIRL_BRAKE || type == TBF_EV_WHIRL_RAMP) {
                const int b = evp->b;
                if (a < 0 || a > (type == TBF_EV_WHIRL_BRAKE ? 1 : 3) || b < 0 || b > 127) {
                    new_flags |= TBF_IFLAG_BAD_EVENT;
                } else if (type == TBF_EV_WHIRL_BRAKE) {
                    /* w->hnBrakePos = (double)uc / 127.0, src/whirl.cpp:883-893 */
                    rp.brakePos[a] = __ddiv_rn((double)b, 127.0);
                } else {
                    rp.oml[a] = wh->ramp_one_minus_l[a][b];
                }
            }
            ++ev_cur;
        }
        /* -- per-call speed ramps (src/whirl.cpp:1219-1374) -- */
        int hbrake = 0, dbrake = 0;
        double hincr = rp.hornIncr, dincr = rp.drumIncr;
        if (whirl_run) {
            if (WHICH != 2)
                hbrake = rotor_ramp(SR, true, &rp.hornAcDc, &hincr, &hang, rp.hornTarget, rp.brakePos[0],
                                    rp.oml[0], rp.oml[1]);
            if (WHICH != 1)
                dbrake = rotor_ramp(SR, false, &rp.drumAcDc, &dincr, &dang, rp.drumTarget, rp.brakePos[1],
                                    rp.oml[2], rp.oml[3]);
        }
        /* fmod(t, 1.0) of a sum in [0, 2) is t or t - 1, both exact */
        const bool hfast = hang >= 0.0 && hang < 1.0 && hincr >= 0.0 && hincr < 1.0;
        const bool dfast = dang >= 0.0 && dang < 1.0 && dincr >= 0.0 && dincr < 1.0;
        for (int blk = 0; blk < TBF_FRAGMENT / BLK; ++blk) {
            const size_t col0 = ((size_t)f * (TBF_FRAGMENT / BLK) + (size_t)blk) * BLK;
            if (lane == 0)
                tma_store_wait_read(); /* the last block's angle tiles have left shared memory */
            __syncwarp();
            if (whirl_run) {
#pragma unroll 4
                for (int s2 = 0; s2 < BLK / 2; ++s2) { /* two samples = one 16-byte chunk */
                    double h[2], dd[2];
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        h[e] = hang;
                        dd[e] = dang;
                        if (WHICH == 2) {
                        } else if (hfast) {
                            const double t = __dadd_rn(hang, hincr);
                            hang = (t >= 1.0) ? __dsub_rn(t, 1.0) : t;
                        } else {
                            hang = angle_step(hang, hincr);
                        }
                        if (WHICH == 1) {
                        } else if (dfast) {
                            const double t = __dadd_rn(dang, dincr);
                            dang = (t >= 1.0) ? __dsub_rn(t, 1.0) : t;
                        } else {
                            dang = angle_step(dang, dincr);
                        }
                    }
                    const uint32_t o = (uint32_t)(s2 >> 3) * TILE_BYTES + tile_chunk(lane, s2 & 7);
                    if (WHICH != 2)
                        *reinterpret_cast<double2 *>(reinterpret_cast<unsigned char *>(tang[0]) + o) = make_double2(h[0], h[1]);
                    if (WHICH != 1)
                        *reinterpret_cast<double2 *>(reinterpret_cast<unsigned char *>(tang[1]) + o) = make_double2(dd[0], dd[1]);
                }
            }
            fence_async_smem();
            __syncwarp();
            if (lane == 0) {
                const uint32_t c2 = 2u * (uint32_t)col0; /* the double planes as pairs of floats */
                if (WHICH != 2) {
                    tma_tile_store(&maps.m64, c2, p.tile_row[8] + slot0, smem_u32(tang[0]));
                    tma_tile_store(&maps.m64, c2 + 32u, p.tile_row[8] + slot0, smem_u32(tang[0]) + TILE_BYTES);
                }
                if (WHICH != 1) {
                    tma_tile_store(&maps.m64, c2, p.tile_row[9] + slot0, smem_u32(tang[1]));
                    tma_tile_store(&maps.m64, c2 + 32u, p.tile_row[9] + slot0, smem_u32(tang[1]) + TILE_BYTES);
                }
                tma_store_commit();
            }
        }
        if (whirl_run) {
            if (WHICH != 2)
                rp.hornIncr = hbrake ? 0.0 : hincr;
            if (WHICH != 1)
                rp.drumIncr = dbrake ? 0.0 : dincr;
        }
    }
    if (valid) {
        RampScalars &o = st->rp; /* this role's fields only: role 0 stores the others */
        if (WHICH != 2) {
            o.hornAcDc = rp.hornAcDc;
            o.revSelect = rp.revSelect;
            o.hornAngle = hang;
            o.hornIncr = rp.hornIncr;
            o.hornTarget = rp.hornTarget;
            o.brakePos[0] = rp.brakePos[0];
            o.oml[0] = rp.oml[0];
            o.oml[1] = rp.oml[1];
            if (new_flags)
                atomicOr(&o.errorFlags, new_flags);
        }
        if (WHICH != 1) {
            o.drumAcDc = rp.drumAcDc;
            o.drumAngle = dang;
            o.drumIncr = rp.drumIncr;
            o.drumTarget = rp.drumTarget;
            o.brakePos[1] = rp.brakePos[1];
            o.oml[2] = rp.oml[2];
            o.oml[3] = rp.oml[3];
        }
    }
    if (cursor_out)
        *cursor_out = ev_cur;
    if (lane == 0)
        tma_store_wait_all();
}

template <bool SPLIT, bool ROT>
__global__ void __launch_bounds__((SPLIT && ROT) ? 96 : 64) tbf_stage_r_tma(const RenderParams p, const __grid_constant__ TileMaps maps)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    RTmaSmem &sm = *reinterpret_cast<RTmaSmem *>(smem_raw + (((raw + 1023u) & ~1023u) - raw));
    const int lane = threadIdx.x & 31;
    /* warp-uniform roles: warp 0 mixes down (and, unless SPLIT, runs the horn biquads), then the rotor
     * warp (ROT; otherwise tbf_stage_rot_tma has run them already), then (SPLIT) the warp of the horn
     * biquads, one block behind warp 0 */
    const bool rotors = ROT && (threadIdx.x >> 5) == 1;
    const bool filters = SPLIT && (threadIdx.x >> 5) == (ROT ? 2 : 1);
    const uint32_t slot0 = blockIdx.x * 32;
    const uint32_t slot = slot0 + lane;
    const bool valid = slot < p.n_instances;
    const bool whirl_off = (p.flags & TBF_RENDER_WHIRL_BYPASS) != 0;
    /* Leslie only: plane u holds the caller's mono input (whirlProc's inbuffer) and the
     * tone generator's mixdown state (percussion gain, key compression) stays as it is */
    const bool ext_in = (p.flags & TBF_RENDER_WHIRL_INPUT) != 0;
    if (rotors && whirl_off)
        return;
    const uint32_t bar0 = smem_u32(&sm.bar[0]);
    if (SPLIT) { /* bar[0..1]: input tiles landed; [2..3]: x tile b written; [4..5]: x tile b read */
        if (threadIdx.x == 0) {
#pragma unroll
            for (int j = 0; j < 6; ++j)
                mbar_init(bar0 + 8u * j, 1u);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
    }

    InstState *st = nullptr;
    const DevWhirlTables *wh = p.wh;
    RampScalars rp;
    if (valid) {
        const uint32_t id = p.ids ? p.ids[slot] : p.id_base + slot;
        st = p.states + id;
        rp = st->rp;
        wh = p.wh + st->tg.whTables;
    } else {
        rp = RampScalars();
    }
    const bool whirl_run = valid && !whirl_off && !wh->bypass;
    uint32_t new_flags = 0;
    uint64_t ev_cur = 0, ev_end = 0;
    if (valid) {
        ev_cur = p.ev_cursor_r[slot];
        ev_end = p.ev_start[slot + 1];
    }
    const uint32_t n_blocks = p.n_fragments * (TBF_FRAGMENT / BLK);

    if (rotors) {
        /* ---------------- role 1: rotor events, speed ramps, angle streams ---------------- */
        r_rotor_role<0>(p, maps, sm.tang, lane, slot0, valid, st, wh, rp, whirl_run, ev_cur, ev_end, nullptr);
        return;
    }

    if (filters) {
        /* ---------------- role 2 (SPLIT): horn biquads A and B, x tile -> hx0 ---------------- */
        float cA[5], cB[5];
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            cA[k] = rp.iir[0][k];
            cB[k] = rp.iir[1][k];
        }
        float zA0 = rp.iirz[0][0], zA1 = rp.iirz[0][1], zB0 = rp.iirz[1][0], zB1 = rp.iirz[1][1];
        uint32_t par_full = 0u;
        unsigned char *o_y = reinterpret_cast<unsigned char *>(sm.tout[2]);
        for (uint32_t f = 0; f < p.n_fragments; ++f) {
            /* horn filter events that precede this fragment (the rotor's belong to role 1) */
            while (ev_cur < ev_end) {
                const tbf_event *evp = p.events + ev_cur;
                if (evp->fragment > p.frag_base + f)
                    break;
                if (evp->type == TBF_EV_HORN_FILTER) {
                    const int a = evp->a, k = a & 7;
                    if (a < 0 || a > 15 || k > 4) {
                        new_flags |= TBF_IFLAG_BAD_EVENT;
                    } else {
                        const float v = __int_as_float(evp->b);
                        rp.iir[a >> 3][k] = v;
#pragma unroll
                        for (int j = 0; j < 5; ++j) {
                            if (a == j)
                                cA[j] = v;
                            if (a == 8 + j)
                                cB[j] = v;
                        }
                    }
                }
                ++ev_cur;
            }
            for (int blk = 0; blk < TBF_FRAGMENT / BLK; ++blk) {
                const uint32_t gb = f * (TBF_FRAGMENT / BLK) + (uint32_t)blk;
                const int cur = (int)(gb & 1u);
                mbar_wait(bar0 + 8u * (uint32_t)(2 + cur), (par_full >> cur) & 1u); /* x tile written */
                par_full ^= 1u << cur;
                if (lane == 0)
                    tma_store_wait_read(); /* the last block's hx0 tile has left shared memory */
                __syncwarp();
                const unsigned char *t_x = reinterpret_cast<const unsigned char *>(sm.tout[cur]);
#pragma unroll 2
                for (int c = 0; c < 8; ++c) {
                    const uint32_t o = tile_chunk(lane, c);
                    const float4 x4 = *reinterpret_cast<const float4 *>(t_x + o);
                    const float xs[4] = {x4.x, x4.y, x4.z, x4.w};
                    float ys[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        if (whirl_run) {
                            const float ya = eq_iir(cA, zA0, zA1, xs[e]);
                            ys[e] = eq_iir(cB, zB0, zB1, ya);
                        } else {
                            ys[e] = xs[e]; /* bypassed Leslie: stage D mixes it */
                        }
                    }
                    *reinterpret_cast<float4 *>(o_y + o) = make_float4(ys[0], ys[1], ys[2], ys[3]);
                }
                fence_async_smem();
                __syncwarp();
                if (lane == 0) {
                    mbar_arrive(bar0 + 8u * (uint32_t)(4 + cur)); /* x tile read: warp 0 may refill it */
                    tma_tile_store(&maps.m32, gb * BLK, p.tile_row[3] + slot0, smem_u32(o_y));
                    tma_store_commit();
                }
            }
            if (whirl_run) { /* NaN scrub after the loop (EQ_IIR_NAN, src/whirl.cpp:1622-1625) */
                zA0 = isnan(zA0) ? 0.0f : zA0;
                zA1 = isnan(zA1) ? 0.0f : zA1;
                zB0 = isnan(zB0) ? 0.0f : zB0;
                zB1 = isnan(zB1) ? 0.0f : zB1;
            }
        }
        if (valid) {
            RampScalars &o = st->rp; /* this role's fields only */
            o.iirz[0][0] = zA0;
            o.iirz[0][1] = zA1;
            o.iirz[1][0] = zB0;
            o.iirz[1][1] = zB1;
#pragma unroll
            for (int k = 0; k < 5; ++k) {
                o.iir[0][k] = rp.iir[0][k];
                o.iir[1][k] = rp.iir[1][k];
            }
            if (new_flags)
                atomicOr(&o.errorFlags, new_flags);
            p.ev_cursor_r[slot] = ev_cur;
        }
        if (lane == 0)
            tma_store_wait_all();
        return;
    }

    /* ---------------- role 0: mixdown, horn biquads, dx0 / hx0 ---------------- */
    float cA[5], cB[5];
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        cA[k] = rp.iir[0][k];
        cB[k] = rp.iir[1][k];
    }
    float zA0 = rp.iirz[0][0], zA1 = rp.iirz[0][1], zB0 = rp.iirz[1][0], zB1 = rp.iirz[1][1];
    float pg = rp.percEnvGain, kc = rp.keyCompLevel;
    const uint32_t rows = min(32u, p.n_instances - slot0);
    const bool store = !(p.flags & TBF_RENDER_NO_OUTPUT);
    const bool out_vec = (p.out_stride % 4) == 0 && (((uintptr_t)p.out_l | (uintptr_t)p.out_r) & 15u) == 0;
    if (p.checksums && whirl_off) /* csum lies past the launch's shared memory otherwise */
        for (int r = 0; r < BLK; ++r)
            sm.csum[r][lane] = 0ull;
    if (lane == 0) {
        if (!SPLIT) {
            mbar_init(bar0, 1u);
            mbar_init(bar0 + 8u, 1u);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        mbar_expect_tx(bar0, 2u * TILE_BYTES);
        tma_tile_load(smem_u32(sm.tin[0][0]), &maps.m32, 0u, p.tile_row[0] + slot0, bar0);
        tma_tile_load(smem_u32(sm.tin[0][1]), &maps.m32, 0u, p.tile_row[1] + slot0, bar0);
    }
    uint32_t parity = 0u, par_empty = 0u;
    __syncwarp();

    for (uint32_t f = 0; f < p.n_fragments; ++f) {
        /* -- horn filter events that precede this fragment (the rotor's belong to role 1; with SPLIT
         *    warp 2 applies them and this warp has nothing to look for) -- */
        while (!SPLIT && ev_cur < ev_end) {
            const tbf_event *evp = p.events + ev_cur;
            if (evp->fragment > p.frag_base + f)
                break;
            if (evp->type == TBF_EV_HORN_FILTER) {
                const int a = evp->a, k = a & 7;
                if (a < 0 || a > 15 || k > 4) {
                    new_flags |= TBF_IFLAG_BAD_EVENT;
                } else {
                    const float v = __int_as_float(evp->b);
                    rp.iir[a >> 3][k] = v;
#pragma unroll
                    for (int j = 0; j < 5; ++j) {
                        if (a == j)
                            cA[j] = v;
                        if (a == 8 + j)
                            cB[j] = v;
                    }
                }
            }
            ++ev_cur;
        }
        FragCtl c_;
        c_.key_comp = c_.gain = c_.decay = c_.reset = 0.0f;
        c_.flags = 0;
        if (valid && !ext_in)
            c_ = p.ctl[(size_t)slot * p.n_fragments + f];
        const bool perc_on = (c_.flags & FC_PERC_ON) != 0;
        /* (keyComp - level) / 128.f: dividing by 2^7 and multiplying by 2^-7 round identically */
        const float delta = __fmul_rn(__fsub_rn(c_.key_comp, kc), 0.0078125f);

        for (int blk = 0; blk < TBF_FRAGMENT / BLK; ++blk) {
            const uint32_t gb = f * (TBF_FRAGMENT / BLK) + (uint32_t)blk;
            const size_t col0 = (size_t)gb * BLK;
            const int cur = (int)(gb & 1u);
            /* the next block's tiles are requested before this block's serial loop */
            if (lane == 0 && gb + 1 < n_blocks) {
                const uint32_t bar = bar0 + 8u * (uint32_t)(cur ^ 1);
                mbar_expect_tx(bar, 2u * TILE_BYTES);
                tma_tile_load(smem_u32(sm.tin[cur ^ 1][0]), &maps.m32, (gb + 1) * BLK, p.tile_row[0] + slot0, bar);
                tma_tile_load(smem_u32(sm.tin[cur ^ 1][1]), &maps.m32, (gb + 1) * BLK, p.tile_row[1] + slot0, bar);
            }
            mbar_wait(bar0 + 8u * (uint32_t)cur, (parity >> cur) & 1u);
            parity ^= 1u << cur;
            if (SPLIT) {
                /* x tile `cur` is free: warp 2 has read block gb - 2 out of it (and this warp's own
                 * store of that block, the one but last committed, has left shared memory) */
                if (gb >= 2) {
                    mbar_wait(bar0 + 8u * (uint32_t)(4 + cur), (par_empty >> cur) & 1u);
                    par_empty ^= 1u << cur;
                }
                if (lane == 0)
                    tma_store_wait_read_but_one();
            } else if (lane == 0) {
                tma_store_wait_read(); /* the last block's output tiles have left shared memory */
            }
            __syncwarp();
            const unsigned char *t_u = reinterpret_cast<const unsigned char *>(sm.tin[cur][0]);
            const unsigned char *t_p = reinterpret_cast<const unsigned char *>(sm.tin[cur][1]);
            unsigned char *o_x = reinterpret_cast<unsigned char *>(sm.tout[SPLIT ? cur : 0]);
            unsigned char *o_y = reinterpret_cast<unsigned char *>(sm.tout[SPLIT ? 2 : 1]);
#pragma unroll 2
            for (int c = 0; c < 8; ++c) { /* four samples = one 16-byte chunk of the lane's row */
                const uint32_t o = tile_chunk(lane, c);
                const float4 u4 = *reinterpret_cast<const float4 *>(t_u + o);
                const float4 p4 = *reinterpret_cast<const float4 *>(t_p + o);
                const float us[4] = {u4.x, u4.y, u4.z, u4.w}, ps[4] = {p4.x, p4.y, p4.z, p4.w};
                float xs[4], ys[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float uu = us[e];
                    float mono;
                    if (ext_in) {
                        mono = uu;
                    } else {
                        if (perc_on) {
                            const float pterm = __fmul_rn(ps[e], pg);
                            mono = __fmul_rn(__fmul_rn(c_.gain, kc), __fadd_rn(uu, pterm));
                            pg = __fmul_rn(pg, c_.decay);
                        } else {
                            mono = __fmul_rn(__fmul_rn(c_.gain, kc), uu);
                        }
                        kc = __fadd_rn(kc, delta);
                    }
                    if (whirl_run) {
                        const float x = (float)__dadd_rn((double)mono, 1e-14); /* DENORMAL_HACK */
                        xs[e] = x;
                        if (!SPLIT) {
                            const float ya = eq_iir(cA, zA0, zA1, x);
                            ys[e] = eq_iir(cB, zB0, zB1, ya);
                        }
                    } else {
                        xs[e] = mono;
                        ys[e] = mono; /* bypassed Leslie: stage D mixes it */
                    }
                }
                *reinterpret_cast<float4 *>(o_x + o) = make_float4(xs[0], xs[1], xs[2], xs[3]);
                if (!SPLIT)
                    *reinterpret_cast<float4 *>(o_y + o) = make_float4(ys[0], ys[1], ys[2], ys[3]);
            }
            if (whirl_off) {
                /* the whole batch bypasses the Leslie: L = R = mono, straight to the output rows;
                 * lane = four samples of a row, eight lanes per row, four rows per trip */
                __syncwarp();
                if (store) {
                    const int c = lane & 7;
                    if (out_vec) {
#pragma unroll 2
                        for (uint32_t r = (uint32_t)lane >> 3; r < rows; r += 4) {
                            const float4 m = *reinterpret_cast<const float4 *>(o_x + tile_chunk((int)r, c));
                            const size_t g = (size_t)(slot0 + r) * p.out_stride + col0 + 4 * c;
                            *reinterpret_cast<float4 *>(p.out_l + g) = m;
                            *reinterpret_cast<float4 *>(p.out_r + g) = m;
                        }
                    } else {
                        for (uint32_t r = 0; r < rows; ++r) {
                            const float m = *reinterpret_cast<const float *>(o_x + tile_chunk((int)r, lane >> 2) + 4 * (lane & 3));
                            const size_t g = (size_t)(slot0 + r) * p.out_stride + col0 + lane;
                            p.out_l[g] = m;
                            p.out_r[g] = m;
                        }
                    }
                }
                if (p.checksums) {
                    for (uint32_t r = 0; r < rows; ++r) {
                        const float m = *reinterpret_cast<const float *>(o_x + tile_chunk((int)r, lane >> 2) + 4 * (lane & 3));
                        const uint32_t pos = (uint32_t)((size_t)p.frag_base * TBF_FRAGMENT + col0 + lane);
                        sm.csum[r][lane] += sample_hash(pos, m, m);
                    }
                }
                __syncwarp();
                continue;
            }
            fence_async_smem();
            __syncwarp();
            if (lane == 0) {
                tma_tile_store(&maps.m32, (uint32_t)col0, p.tile_row[2] + slot0, smem_u32(o_x));
                if (!SPLIT)
                    tma_tile_store(&maps.m32, (uint32_t)col0, p.tile_row[3] + slot0, smem_u32(o_y));
                tma_store_commit();
                if (SPLIT)
                    mbar_arrive(bar0 + 8u * (uint32_t)(2 + cur)); /* x tile written: warp 2 may filter it */
            }
        }

        /* -- end of the fragment -- */
        if (!SPLIT && whirl_run) {
            /* NaN scrub after the loop (EQ_IIR_NAN, src/whirl.cpp:1622-1625) */
            zA0 = isnan(zA0) ? 0.0f : zA0;
            zA1 = isnan(zA1) ? 0.0f : zA1;
            zB0 = isnan(zB0) ? 0.0f : zB0;
            zB1 = isnan(zB1) ? 0.0f : zB1;
        }
        if (c_.flags & FC_PERC_RESET)
            pg = c_.reset;
    }

    if (valid) {
        RampScalars &o = st->rp; /* this role's fields only */
        if (!SPLIT) {
            o.iirz[0][0] = zA0;
            o.iirz[0][1] = zA1;
            o.iirz[1][0] = zB0;
            o.iirz[1][1] = zB1;
#pragma unroll
            for (int k = 0; k < 5; ++k) {
                o.iir[0][k] = rp.iir[0][k];
                o.iir[1][k] = rp.iir[1][k];
            }
            if (new_flags)
                atomicOr(&o.errorFlags, new_flags);
            p.ev_cursor_r[slot] = ev_cur;
        }
        o.percEnvGain = pg;
        o.keyCompLevel = kc;
        if (p.checksums && whirl_off) {
            unsigned long long c = 0ull;
            for (int j = 0; j < BLK; ++j)
                c += sm.csum[lane][(j + lane) & (BLK - 1)];
            p.checksums[slot] += c; /* zeroed by the host; accumulates over chunks */
        }
    }
    if (lane == 0)
        tma_store_wait_all();
}

/* The rotors of 32 organs as a kernel of their own: launched on stage R's stream BEFORE it waits for
 * stage T, so the angle streams are written while stage T still runs and the audio chain's CTAs need
 * neither the angle tiles (16 KB) nor a third warp. */
template <bool TWO>
__global__ void __launch_bounds__(TWO ? 64 : 32) tbf_stage_rot_tma(const RenderParams p, const __grid_constant__ TileMaps maps)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    RotSmem &sm = *reinterpret_cast<RotSmem *>(smem_raw + (((raw + 1023u) & ~1023u) - raw));
    const int lane = threadIdx.x & 31;
    const uint32_t slot0 = blockIdx.x * 32;
    const uint32_t slot = slot0 + lane;
    const bool valid = slot < p.n_instances;
    InstState *st = nullptr;
    const DevWhirlTables *wh = p.wh;
    RampScalars rp;
    uint64_t ev_cur = 0, ev_end = 0;
    if (valid) {
        const uint32_t id = p.ids ? p.ids[slot] : p.id_base + slot;
        st = p.states + id;
        rp = st->rp;
        wh = p.wh + st->tg.whTables;
        ev_cur = p.ev_cursor_rot[slot]; /* its own: this kernel runs ahead of the audio chain's */
        ev_end = p.ev_start[slot + 1];
    } else {
        rp = RampScalars();
    }
    const bool whirl_run = valid && !wh->bypass;
    /* warp 0 the horn (and the event cursor), warp 1 the drum: both read the same events */
    if (!TWO)
        r_rotor_role<0>(p, maps, sm.tang, lane, slot0, valid, st, wh, rp, whirl_run, ev_cur, ev_end,
                        valid ? p.ev_cursor_rot + slot : nullptr);
    else if ((threadIdx.x >> 5) == 0)
        r_rotor_role<1>(p, maps, sm.tang, lane, slot0, valid, st, wh, rp, whirl_run, ev_cur, ev_end,
                        valid ? p.ev_cursor_rot + slot : nullptr);
    else
        r_rotor_role<2>(p, maps, sm.tang, lane, slot0, valid, st, wh, rp, whirl_run, ev_cur, ev_end, nullptr);
}

/* ================================================================== */
/* stage W: Leslie reflections, taps, delay lines; warp = instance     */
/* ================================================================== */

/* One tap as the block loop needs it, in processing order: the three taps of the group's left
 * line, farthest delay first, then the three of the right line (tap_order, tbf_api.cu).
 * Filled once per launch so the loop reads parameters with warp-uniform LDS and indexes its
 * per-tap registers statically. */
struct WTap {
    const float2 *dsp; /* displacement table as pairs {d[k], d[k+1]}                    */
    const float *bw;   /* horn: FIR bank, five planes of 16384 coefficients; drum: unused    */
    double ang_off;    /* horn: microphone angle offset added to the rotor angle         */
    float spacing;
    int phase;
    int stage;         /* reflection stage feeding the tap (tap id >> 1)                  */
    int q;             /* tap id inside its group, 0..5 (serial fallback order)           */
};

/* Stage W runs as two independent halves per organ, one warp each: the HORN half (reflection
 * stages with FIR histories, six horn taps, delay lines HL HR) and the DRUM half (two reflection
 * stages in registers, six drum taps, delay lines DL DR).  Nothing flows between them
 * (src/whirl.cpp:1525-1588: the drum path restarts from the unfiltered input), so each half keeps
 * only its own two delay-line rings in shared memory: 6 KB / 4.9 KB per warp instead of 10.9 KB,
 * i.e. 32 resident warps per SM instead of 20. */
struct WWarp {
    float x[3][8 + TBF_FRAGMENT]; /* horn: stage inputs with 8 samples of history; drum: x[0] only,
                                     the path's input with its history at index 7            */
    float zs[2];                  /* FILTER_C states of the half (z[0..1] / z[2..3]), NaN-scrubbed
                                     per fragment                                            */
    WTap tap[6];                  /* the half's taps in processing order                       */
    /* followed by float ring[2][ring_size]: the half's delay lines (left, right) */
};

/* FILTER_C(0.4, 0.4): the literals are double, so the 2-tap FIR is evaluated in
 * double and rounded to float once (src/whirl.cpp:1472-1477) */
__device__ __forceinline__ float filter_c(float x, float z)
{
    return (float)__dadd_rn(__dmul_rn(0.4, (double)x), __dmul_rn(0.4, (double)z));
}

/* floorf(x) and (unsigned)floorf(x) of 0 <= x < 2^22 without a conversion instruction: adding
 * 2^23 with round-down leaves floor(x) in the mantissa (the sum's ulp is 1), both results exact */
__device__ __forceinline__ float floor_small(float x, uint32_t &as_int)
{
    const float y = __fadd_rd(x, 8388608.0f);
    as_int = __float_as_uint(y) & 0x007FFFFFu;
    return __fsub_rn(y, 8388608.0f);
}

/* tap geometry (HN_MOTION / DR_MOTION, src/whirl.cpp:1432-1469).  FAST: the caller has checked
 * that the rotor angle is in [+0, 1) and the host that phases and delays are small and non-
 * negative (DevWhirlTables::small_index), so h1 and t lie in [+0, 2^22): truncf = floorf,
 * roundf(h1) = floorf(h1) + (fractional part >= 0.5), and the float <-> integer conversions of
 * the reference become mantissa reads -- same values, no XU-pipe instruction. */
template <bool FAST>
__device__ __forceinline__ void tap_index(double ang_grd, int phase, float &hd, uint32_t &hl, uint32_t &kround)
{
    const float h1 = (float)__dadd_rn(__dmul_rn(ang_grd, (double)TBF_WHIRL_DISPLC), (double)phase);
    if (FAST) {
        const float fl = floor_small(h1, hl);
        hd = __fsub_rn(h1, fl);
        kround = hl + (hd >= 0.5f ? 1u : 0u); /* roundf of a non-negative float: halves go up */
    } else {
        hd = __fsub_rn(h1, truncf(h1)); /* fmodf(h1, 1.f), exact */
        hl = f2u_x86(floorf(h1));
        kround = f2u_x86(roundf(h1));
    }
    hl &= DISPLC_MASK;
    kround &= DISPLC_MASK;
}

/* d = {displacement[hl], displacement[hl + 1]} -> fractional delay-line position of the tap */
template <bool FAST>
__device__ __forceinline__ void tap_slot(const float2 d, float hd, float spacing, uint32_t outpos, float &frac,
                                         uint32_t &slot)
{
    const float intp = __fadd_rn(__fmul_rn(d.x, __fsub_rn(1.f, hd)), __fmul_rn(hd, d.y));
    if (FAST) {
        /* (float)outpos, outpos < 2^11, off the mantissa */
        const float t_out = __fsub_rn(__uint_as_float(0x4B000000u | outpos), 8388608.0f);
        const float t = __fadd_rn(__fadd_rn(spacing, intp), t_out);
        const float r = floor_small(t, slot);
        frac = __fsub_rn(t, r);
    } else {
        const float t = __fadd_rn(__fadd_rn(spacing, intp), (float)outpos);
        const float r = floorf(t);
        frac = __fsub_rn(t, r);
        slot = f2u_x86(r);
    }
    slot &= WBUF_MASK;
}

/* What the parallel delay-line update relies on, checked per block and tap: the
 * destination slots advance monotonically with the sample index, at most two samples
 * share a slot, and all stay inside [reader, reader + window). */
__device__ __forceinline__ int tap_bad(const uint32_t n, const uint32_t dn, const RunBits rb,
                                       const uint32_t window, const int lane, const uint32_t outpos_blk)
{
    const uint32_t rel = (n - outpos_blk) & WBUF_MASK;
    int bad = (rel < (uint32_t)lane) | (rel + 1 >= window);
    if (lane >= 1)
        bad |= dn >= (TBF_WHIRL_BUF / 2);
    bad |= ((rb.e >> lane) & ((rb.e << 1) >> lane) & 1u) != 0u; /* three samples on one slot */
    return bad;
}

/* Scatter the six taps of one group (horn: HL, HR; drum: DL, DR) of one 32-sample block
 * into the group's two delay lines and read the block's 32 outputs.  tn/ta/tb are in
 * processing order (WTap): 0..2 feed the left line, 3..5 the right one, farthest delay first. */
__device__ __forceinline__ void group_scatter_read(const WTap *tap, float *ring_l,
                                                   float *ring_r, uint32_t rmask, int lane,
                                                   uint32_t outpos_blk, int bad, int gather_ok,
                                                   int &fell_back, const uint32_t (&tn)[6],
                                                   const RunBits (&trb)[6],
                                                   const float (&ta)[6], const float (&tb)[6],
                                                   float &out_l, float &out_r)
{
    bad = __any_sync(FULL, bad);
    const uint32_t op = (outpos_blk + (uint32_t)lane) & WBUF_MASK;
    if (!bad) {
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            ordered_pass_bits(tn[j], trb[j], ta[j], tb[j], ring_l, rmask, lane);
            ordered_pass_bits(tn[3 + j], trb[3 + j], ta[3 + j], tb[3 + j], ring_r, rmask, lane);
            __syncwarp();
        }
        out_l = ring_l[op & rmask];
        ring_l[op & rmask] = 0.0f;
        out_r = ring_r[op & rmask];
        ring_r[op & rmask] = 0.0f;
    } else {
        /* the reference's loop as written, one lane per delay line (lane 0 left, lane 1
         * right): geometries / rotor speeds the parallel passes do not cover.  The lanes
         * hand sample i of every tap over by shuffle; a line applies its three taps in
         * tap-id order (src/whirl.cpp:1540-1588). */
        fell_back |= gather_ok;
        const int j0 = (lane & 1) * 3; /* my line's first processing index */
        int ord[3];                    /* processing index (0..2 in my line) of tap id lane + 2s */
#pragma unroll
        for (int s = 0; s < 3; ++s) {
            const int q = (lane & 1) + 2 * s;
            ord[s] = tap[j0].q == q ? 0 : (tap[j0 + 1].q == q ? 1 : 2);
        }
        out_l = out_r = 0.0f;
        for (int i = 0; i < BLK; ++i) {
            uint32_t n3[3];
            float a3[3], b3[3];
#pragma unroll
            for (int c = 0; c < 3; ++c) { /* my line's taps in processing order */
                const uint32_t nl = __shfl_sync(FULL, tn[c], i), nr = __shfl_sync(FULL, tn[3 + c], i);
                const float al = __shfl_sync(FULL, ta[c], i), ar = __shfl_sync(FULL, ta[3 + c], i);
                const float bl = __shfl_sync(FULL, tb[c], i), br = __shfl_sync(FULL, tb[3 + c], i);
                n3[c] = (lane & 1) ? nr : nl;
                a3[c] = (lane & 1) ? ar : al;
                b3[c] = (lane & 1) ? br : bl;
            }
            float o_val = 0.0f;
            if (lane < 2) {
                float *buf = lane ? ring_r : ring_l;
#pragma unroll
                for (int s = 0; s < 3; ++s) {
                    const uint32_t n = ord[s] == 0 ? n3[0] : (ord[s] == 1 ? n3[1] : n3[2]);
                    const float a = ord[s] == 0 ? a3[0] : (ord[s] == 1 ? a3[1] : a3[2]);
                    const float b = ord[s] == 0 ? b3[0] : (ord[s] == 1 ? b3[1] : b3[2]);
                    buf[n & rmask] = __fadd_rn(buf[n & rmask], a);
                    const uint32_t n1 = (n + 1) & WBUF_MASK;
                    buf[n1 & rmask] = __fadd_rn(buf[n1 & rmask], b);
                }
                const uint32_t o = (outpos_blk + (uint32_t)i) & WBUF_MASK;
                o_val = buf[o & rmask];
                buf[o & rmask] = 0.0f;
            }
            const float vl = __shfl_sync(FULL, o_val, 0), vr = __shfl_sync(FULL, o_val, 1);
            if (lane == i) {
                out_l = vl;
                out_r = vr;
            }
        }
    }
    __syncwarp();
}

/* the six taps of one half for one 32-sample block: geometry, table reads, (horn) FIRs, the
 * fractional split.  Written in groups of three taps -- index math and loads first, consumers
 * after -- so that the reads of a group are in flight together. */
template <bool DRUM, bool FAST, bool CERT>
__device__ __forceinline__ void block_taps(const WWarp &sm, const int i, const double ang, const uint32_t outpos,
                                           const uint32_t outpos_blk, const uint32_t window, const int lane,
                                           uint32_t (&tn)[6], RunBits (&trb)[6],
                                           float (&ta)[6], float (&tb)[6], int &bad)
{
    float d0 = 0.f, d1 = 0.f, d2 = 0.f;
    if (DRUM) {
        /* the drum path's reflection stages of this sample; stage 1 of sample i-1 is
         * recomputed locally instead of exchanged */
        const float d0m = sm.x[0][7 + i];
        d0 = sm.x[0][8 + i];
        d1 = filter_c(d0, d0m);
        const float d1m = i ? filter_c(d0m, i > 1 ? sm.x[0][6 + i] : sm.zs[0]) : sm.zs[1];
        d2 = filter_c(d1, d1m);
    }
#pragma unroll
    for (int g = 0; g < 6; g += 3) {
        float hd[3], b0[3], b1[3], b2[3], b3[3], b4[3];
        float2 d[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const WTap &t = sm.tap[g + c];
            uint32_t hl, kr;
            tap_index<FAST>(DRUM ? ang : __dadd_rn(ang, t.ang_off), t.phase, hd[c], hl, kr);
            d[c] = __ldg(t.dsp + hl);
            if (!DRUM) {
                const float *b = t.bw + kr;
                b0[c] = __ldg(b);
                b1[c] = __ldg(b + DISPLC_MASK + 1);
                b2[c] = __ldg(b + 2 * (DISPLC_MASK + 1));
                b3[c] = __ldg(b + 3 * (DISPLC_MASK + 1));
                b4[c] = __ldg(b + 4 * (DISPLC_MASK + 1));
            }
        }
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const int j = g + c;
            const WTap &t = sm.tap[j];
            float frac, xa;
            tap_slot<FAST>(d[c], hd[c], t.spacing, outpos, frac, tn[j]);
            if (!DRUM) {
                /* horn taps: 5-tap FIR over the stage's history (src/whirl.cpp:1432-1453) */
                const float *hx = &sm.x[t.stage][8 + i];
                xa = __fmul_rn(b0[c], hx[0]);
                xa = __fadd_rn(xa, __fmul_rn(b1[c], hx[-1]));
                xa = __fadd_rn(xa, __fmul_rn(b2[c], hx[-2]));
                xa = __fadd_rn(xa, __fmul_rn(b3[c], hx[-3]));
                xa = __fadd_rn(xa, __fmul_rn(b4[c], hx[-4]));
            } else { /* drum taps (src/whirl.cpp:1455-1469, :1566-1588) */
                xa = t.stage == 0 ? d0 : (t.stage == 1 ? d1 : d2);
            }
            const float qq = __fmul_rn(xa, frac);
            ta[j] = __fsub_rn(xa, qq);
            tb[j] = qq;
            uint32_t dn;
            trb[j] = run_bits(tn[j], WBUF_MASK, lane, dn);
            if (!CERT) /* CERT: the host bounded the slot motion (DevWhirlTables::taps_certified) */
                bad |= tap_bad(tn[j], dn, trb[j], window, lane, outpos_blk);
        }
    }
}

/* one half of stage W for one organ: DRUM = false the horn half, true the drum half */
template <bool DRUM>
__device__ __forceinline__ void stage_w_half(const RenderParams &p, const uint32_t ring_size, WWarp &sm,
                                             float *ring, const uint32_t slot, const int lane)
{
    const uint32_t rmask = ring_size - 1;
    const uint32_t id = p.ids ? p.ids[slot] : p.id_base + slot;
    InstState *st = p.states + id;
    const DevWhirlTables *wh = p.wh + st->tg.whTables;
    if (wh->bypass)
        return; /* whirlProc2 leaves a bypassed Leslie untouched (src/whirl.cpp:1197-1215) */

    const WhirlScalars &ws = st->wl;
    uint32_t outpos0 = DRUM ? ws.outpos_drum : ws.outpos;
    const int line0 = DRUM ? 2 : 0;
    for (int line = 0; line < 2; ++line)
        for (uint32_t s = lane; s < ring_size; s += 32) {
            const uint32_t g = (outpos0 + s) & WBUF_MASK;
            ring[line * ring_size + (g & rmask)] = st->wbuf[line0 + line][g];
        }
    int adi_l = 0;
    if (!DRUM && lane < 24) {
        const int s = lane >> 3, k = lane & 7;
        adi_l = ws.adi[s];
        sm.x[s][7 - k] = ws.adx[s][(adi_l + k) & 7];
    }
    if (lane < 2)
        sm.zs[lane] = ws.z[(DRUM ? 2 : 0) + lane];
    if (lane < 6) {
        const int g = DRUM ? 1 : 0, jj = lane; /* group, processing index */
        const int q = wh->tap_order[2 * g + jj / 3][jj % 3] - 6 * g;
        /* taps 0,3,4 read the forward displacement table with the backward FIR bank,
         * taps 1,2,5 the other way round */
        const bool fwd = (q == 0 || q == 3 || q == 4);
        WTap t;
        if (!DRUM) {
            t.dsp = fwd ? wh->hn_fwd_displ : wh->hn_bwd_displ;
            t.bw = fwd ? wh->bbw : wh->bfw;
            t.ang_off = (q & 1) ? wh->bw_ang : wh->fw_ang;
            t.spacing = wh->horn_spacing[q];
            t.phase = wh->horn_phase[q];
        } else {
            t.dsp = fwd ? wh->dr_fwd_displ : wh->dr_bwd_displ;
            t.bw = nullptr;
            t.ang_off = 0.0;
            t.spacing = wh->drum_spacing[q];
            t.phase = wh->drum_phase[q];
        }
        t.stage = q >> 1;
        t.q = q;
        sm.tap[lane] = t;
    }
    const bool gather_ok = wh->gather_ok != 0;
    const bool small_index = wh->small_index != 0;
    const bool certified = wh->taps_certified != 0 && wh->gather_ok != 0;
    const double max_ang = DRUM ? 1.0 : wh->fast_max_angle; /* rotor angles below it keep h1 < 2^22 */
    const uint32_t window = (uint32_t)wh->window;
    int fell_back = 0;
    const size_t row = (size_t)slot * p.pitch;
    const double *ang_row = DRUM ? p.dang : p.hang;
    const float *in_row = DRUM ? p.dx0 : p.hx0;
    float *out_l_row = p.line[line0], *out_r_row = p.line[line0 + 1];
    double ang_next = ang_row[row + lane];
    __syncwarp();

    for (uint32_t f = 0; f < p.n_fragments; ++f) {
        const size_t col = row + (size_t)f * TBF_FRAGMENT;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int i = BLK * k + lane;
            sm.x[0][8 + i] = in_row[col + i];
        }
        if (DRUM && lane == 0)
            sm.x[0][7] = sm.zs[0];
        __syncwarp();
        if (!DRUM) {
            /* -- the horn path's two reflection filters (2-tap FIRs in double); stage 1 of
             *    sample i-1 is recomputed locally instead of exchanged -- */
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int i = BLK * k + lane;
                const float h0 = sm.x[0][8 + i], h0m = i ? sm.x[0][7 + i] : sm.zs[0];
                const float h1 = filter_c(h0, h0m);
                const float h1m = i ? filter_c(h0m, i > 1 ? sm.x[0][6 + i] : sm.zs[0]) : sm.zs[1];
                sm.x[1][8 + i] = h1;
                sm.x[2][8 + i] = filter_c(h1, h1m);
            }
            __syncwarp();
        }

#pragma unroll 1
        for (int k = 0; k < 4; ++k) {
            const int i = BLK * k + lane;
            const uint32_t outpos_blk = (outpos0 + (uint32_t)(BLK * k)) & WBUF_MASK;
            const uint32_t outpos = (outpos_blk + (uint32_t)lane) & WBUF_MASK;
            const double ang = ang_next;
            { /* the next block's rotor angles, one block ahead of their use */
                const size_t nx = col + i + BLK;
                if (k < 3 || f + 1 < p.n_fragments)
                    ang_next = ang_row[nx];
            }
            /* the conversion-free index math needs the angle in [+0, limit): sign bit clear */
            const bool fast = small_index &&
                              __all_sync(FULL, __double_as_longlong(ang) >= 0 && ang < max_ang);
            uint32_t tn[6];
            RunBits trb[6];
            float ta[6], tb[6];
            int bad = gather_ok ? 0 : 1;
            /* one warp-uniform branch per block, not per tap: inside a variant the six taps are
             * straight-line code, so their table reads overlap */
            if (fast && certified)
                block_taps<DRUM, true, true>(sm, i, ang, outpos, outpos_blk, window, lane, tn, trb, ta, tb, bad);
            else if (fast)
                block_taps<DRUM, true, false>(sm, i, ang, outpos, outpos_blk, window, lane, tn, trb, ta, tb, bad);
            else
                block_taps<DRUM, false, false>(sm, i, ang, outpos, outpos_blk, window, lane, tn, trb, ta, tb, bad);
            float o_l, o_r;
            group_scatter_read(&sm.tap[0], ring, ring + ring_size, rmask, lane, outpos_blk, bad,
                               gather_ok ? 1 : 0, fell_back, tn, trb, ta, tb, o_l, o_r);
            out_l_row[col + i] = o_l;
            out_r_row[col + i] = o_r;
        }

        /* carry FIR histories / filter states into the next fragment */
        float keep = 0.f;
        if (!DRUM && lane < 24)
            keep = sm.x[lane >> 3][TBF_FRAGMENT + (lane & 7)];
        float zc = 0.f;
        if (lane < 2) {
            const float last = sm.x[0][8 + 127];
            if (!DRUM)
                zc = (lane == 0) ? last : sm.x[1][8 + 127];
            else
                zc = (lane == 0) ? last : filter_c(last, sm.x[0][8 + 126]);
        }
        __syncwarp();
        if (!DRUM && lane < 24)
            sm.x[lane >> 3][lane & 7] = keep;
        if (lane < 2)
            sm.zs[lane] = isnan(zc) ? 0.0f : zc; /* src/whirl.cpp:1627-1630 */
        outpos0 = (outpos0 + TBF_FRAGMENT) & WBUF_MASK;
        __syncwarp();
    }

    /* ---- write the staged state back ---- */
    if (!DRUM && lane < 24) {
        const int s = lane >> 3, k = lane & 7;
        st->wl.adx[s][(adi_l + k) & 7] = sm.x[s][7 - k];
    }
    if (lane < 2)
        st->wl.z[(DRUM ? 2 : 0) + lane] = sm.zs[lane];
    if (lane == 0) {
        if (DRUM) /* both halves advance their read position alike */
            st->wl.outpos_drum = outpos0;
        else
            st->wl.outpos = outpos0;
        if (fell_back)
            atomicOr(&st->wl.errorFlags, TBF_IFLAG_GATHER_FALLBACK);
    }
    /* outside the live windows every ring slot is zero (cleared when read) */
    for (int line = 0; line < 2; ++line)
        for (uint32_t s = lane; s < TBF_WHIRL_BUF; s += 32) {
            const uint32_t g = (outpos0 + s) & WBUF_MASK;
            st->wbuf[line0 + line][g] = (s < ring_size) ? ring[line * ring_size + (g & rmask)] : 0.0f;
        }
}

/* blockIdx.y = 0: horn halves, 1: drum halves; W_WARPS organs per CTA */
#ifndef W_MIN_CTAS
#define W_MIN_CTAS 8
#endif
/* RING: the delay-line ring held in shared memory as a compile-time constant (512: every default
 * geometry; the masks become immediates), 0: taken from the launch (wide geometries) */
template <int RING>
__global__ void __launch_bounds__(W_WARPS * 32, W_MIN_CTAS) tbf_stage_w(const RenderParams p, const uint32_t ring_size_arg)
{
    const uint32_t ring_size = RING ? (uint32_t)RING : ring_size_arg;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t slot = blockIdx.x * W_WARPS + warp;
    if (slot >= p.n_instances)
        return; /* warps are independent: no CTA-wide barrier anywhere in this kernel */
    const size_t per_warp = sizeof(WWarp) + (size_t)2 * ring_size * sizeof(float);
    WWarp &sm = *reinterpret_cast<WWarp *>(smem_raw + per_warp * warp);
    float *ring = reinterpret_cast<float *>(smem_raw + per_warp * warp + sizeof(WWarp));
    if (blockIdx.y == 0)
        stage_w_half<false>(p, ring_size, sm, ring, slot, lane);
    else
        stage_w_half<true>(p, ring_size, sm, ring, slot, lane);
}

/* ================================================================== */
/* stage D: drum biquads (lane = instance) + output mix (lane = sample) */
/* ================================================================== */

struct DSmem {
    float t[2][2][BLK * TILE_PITCH]; /* [buffer][DL DR] (row = instance, col = sample), filled by
                                        cp.async one block ahead of its use                  */
};

/* request one 32 x 32 block of the two drum delay-line reads */
__device__ __forceinline__ void d_tile_fetch(const RenderParams &p, float (*tile)[BLK * TILE_PITCH],
                                             uint32_t slot0, uint32_t rows, size_t col0, int lane)
{
    for (uint32_t r = 0; r < rows; ++r) {
        const size_t g = (size_t)(slot0 + r) * p.pitch + col0 + lane;
        const int o = r * TILE_PITCH + lane;
        __pipeline_memcpy_async(&tile[0][o], p.line[2] + g, 4);
        __pipeline_memcpy_async(&tile[1][o], p.line[3] + g, 4);
    }
}

/* Stage D, first kernel: the drum biquads drfL / drfR (src/whirl.cpp:1607-1612), the one serial
 * recurrence left after stage W; lane = organ, the two line reads are filtered IN PLACE in the
 * stage streams.  Everything per-sample after them is elementwise and runs in tbf_stage_m with
 * lane = sample, so this warp -- a latency chain that takes a stage T / W slot of its SM while it
 * runs -- carries 2 of the 7 streams and a third of the instructions it used to. */
__global__ void __launch_bounds__(32) tbf_stage_d(const RenderParams p)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    DSmem &sm = *reinterpret_cast<DSmem *>(smem_raw);
    const int lane = threadIdx.x;
    const uint32_t slot0 = blockIdx.x * 32;
    const uint32_t slot = slot0 + lane;
    const bool valid = slot < p.n_instances;
    const uint32_t rows = min(32u, p.n_instances - slot0);

    InstState *st = nullptr;
    const DevWhirlTables *wh = p.wh;
    float zl0 = 0.f, zl1 = 0.f, zr0 = 0.f, zr1 = 0.f;
    if (valid) {
        const uint32_t id = p.ids ? p.ids[slot] : p.id_base + slot;
        st = p.states + id;
        wh = p.wh + st->tg.whTables;
        zl0 = st->dr.iirz[0][0];
        zl1 = st->dr.iirz[0][1];
        zr0 = st->dr.iirz[1][0];
        zr1 = st->dr.iirz[1][1];
    }
    const bool run = valid && !wh->bypass;
    if (!__any_sync(FULL, run))
        return; /* every organ of the warp has its Leslie bypassed: nothing to filter */
    float cL[5], cR[5];
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        cL[k] = wh->iir[2][k];
        cR[k] = wh->iir[3][k];
    }

    const uint32_t n_blocks = p.n_fragments * (TBF_FRAGMENT / BLK);
    d_tile_fetch(p, sm.t[0], slot0, rows, 0, lane);
    __pipeline_commit();
    for (uint32_t blk = 0; blk < n_blocks; ++blk) {
        const size_t col0 = (size_t)blk * BLK;
        const int cur = (int)(blk & 1u);
        if (blk + 1 < n_blocks)
            d_tile_fetch(p, sm.t[cur ^ 1], slot0, rows, (size_t)(blk + 1) * BLK, lane);
        __pipeline_commit();
        __pipeline_wait_prior(1);
        __syncwarp();
        float *tl = sm.t[cur][0], *tr = sm.t[cur][1];
        if (run) {
#pragma unroll 4
            for (int s = 0; s < BLK; ++s) {
                tl[lane * TILE_PITCH + s] = eq_iir(cL, zl0, zl1, tl[lane * TILE_PITCH + s]);
                tr[lane * TILE_PITCH + s] = eq_iir(cR, zr0, zr1, tr[lane * TILE_PITCH + s]);
            }
            if ((blk & 3u) == 3u) { /* end of a fragment: NaN scrub (src/whirl.cpp:1624-1625) */
                zl0 = isnan(zl0) ? 0.0f : zl0;
                zl1 = isnan(zl1) ? 0.0f : zl1;
                zr0 = isnan(zr0) ? 0.0f : zr0;
                zr1 = isnan(zr1) ? 0.0f : zr1;
            }
        }
        __syncwarp();
#pragma unroll 4
        for (uint32_t r = 0; r < rows; ++r) {
            const size_t g = (size_t)(slot0 + r) * p.pitch + col0 + lane;
            p.line[2][g] = tl[r * TILE_PITCH + lane];
            p.line[3][g] = tr[r * TILE_PITCH + lane];
        }
        __syncwarp();
    }

    if (run) {
        st->dr.iirz[0][0] = zl0;
        st->dr.iirz[0][1] = zl1;
        st->dr.iirz[1][0] = zr0;
        st->dr.iirz[1][1] = zr1;
    }
}

struct DTmaSmem {
    float t[2][2][TILE_BYTES / 4]; /* [buffer][DL DR], swizzled boxes; 1024-byte aligned */
    unsigned long long bar[2][2];  /* [warp][buffer] */
};

/* One warp's share of stage D: NCH of the two drum line reads (from channel ch0), filtered in place
 * in their TMA tiles.  The two channels' filters share nothing (src/whirl.cpp:1609-1612), so one warp
 * per channel halves the time a CTA of stage D stays on its SM. */
template <int NCH>
__device__ __forceinline__ void d_channel_role(const RenderParams &p, const TileMaps &maps, DTmaSmem &sm,
                                               const int lane, const uint32_t slot0, const bool run,
                                               InstState *st, const DevWhirlTables *wh, const int ch0,
                                               const int warp)
{
    float z[NCH][2], c[NCH][5];
    uint32_t row[NCH];
#pragma unroll
    for (int k = 0; k < NCH; ++k) {
        z[k][0] = run ? st->dr.iirz[ch0 + k][0] : 0.f;
        z[k][1] = run ? st->dr.iirz[ch0 + k][1] : 0.f;
#pragma unroll
        for (int j = 0; j < 5; ++j)
            c[k][j] = wh->iir[2 + ch0 + k][j];
        row[k] = p.tile_row[6 + ch0 + k] + slot0; /* line[2], line[3] */
    }
    const uint32_t bar0 = smem_u32(&sm.bar[warp][0]);
    if (lane == 0) {
        mbar_init(bar0, 1u);
        mbar_init(bar0 + 8u, 1u);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        mbar_expect_tx(bar0, (uint32_t)NCH * TILE_BYTES);
#pragma unroll
        for (int k = 0; k < NCH; ++k)
            tma_tile_load(smem_u32(sm.t[0][ch0 + k]), &maps.m32, 0u, row[k], bar0);
    }
    __syncwarp();
    const uint32_t n_blocks = p.n_fragments * (TBF_FRAGMENT / BLK);
    uint32_t parity = 0u;
    for (uint32_t blk = 0; blk < n_blocks; ++blk) {
        const int cur = (int)(blk & 1u);
        if (lane == 0 && blk + 1 < n_blocks) {
            tma_store_wait_read(); /* the other buffer's store (block blk - 1) has left shared memory */
            const uint32_t bar = bar0 + 8u * (uint32_t)(cur ^ 1);
            mbar_expect_tx(bar, (uint32_t)NCH * TILE_BYTES);
#pragma unroll
            for (int k = 0; k < NCH; ++k)
                tma_tile_load(smem_u32(sm.t[cur ^ 1][ch0 + k]), &maps.m32, (blk + 1) * BLK, row[k], bar);
        }
        mbar_wait(bar0 + 8u * (uint32_t)cur, (parity >> cur) & 1u);
        parity ^= 1u << cur;
        if (run) {
#pragma unroll 2
            for (int ck = 0; ck < 8; ++ck) {
                const uint32_t o = tile_chunk(lane, ck);
                float4 a[NCH];
#pragma unroll
                for (int k = 0; k < NCH; ++k)
                    a[k] = *reinterpret_cast<float4 *>(reinterpret_cast<unsigned char *>(sm.t[cur][ch0 + k]) + o);
#pragma unroll
                for (int k = 0; k < NCH; ++k)
                    a[k].x = eq_iir(c[k], z[k][0], z[k][1], a[k].x);
#pragma unroll
                for (int k = 0; k < NCH; ++k)
                    a[k].y = eq_iir(c[k], z[k][0], z[k][1], a[k].y);
#pragma unroll
                for (int k = 0; k < NCH; ++k)
                    a[k].z = eq_iir(c[k], z[k][0], z[k][1], a[k].z);
#pragma unroll
                for (int k = 0; k < NCH; ++k)
                    a[k].w = eq_iir(c[k], z[k][0], z[k][1], a[k].w);
#pragma unroll
                for (int k = 0; k < NCH; ++k)
                    *reinterpret_cast<float4 *>(reinterpret_cast<unsigned char *>(sm.t[cur][ch0 + k]) + o) = a[k];
            }
            if ((blk & 3u) == 3u) { /* end of a fragment: NaN scrub (src/whirl.cpp:1624-1625) */
#pragma unroll
                for (int k = 0; k < NCH; ++k) {
                    z[k][0] = isnan(z[k][0]) ? 0.0f : z[k][0];
                    z[k][1] = isnan(z[k][1]) ? 0.0f : z[k][1];
                }
            }
        }
        fence_async_smem(); /* this lane's writes, before the TMA store reads them */
        __syncwarp();
        if (lane == 0) {
#pragma unroll
            for (int k = 0; k < NCH; ++k)
                tma_tile_store(&maps.m32, blk * BLK, row[k], smem_u32(sm.t[cur][ch0 + k]));
            tma_store_commit();
        }
    }
    if (lane == 0)
        tma_store_wait_all();
    if (run) {
#pragma unroll
        for (int k = 0; k < NCH; ++k) {
            st->dr.iirz[ch0 + k][0] = z[k][0];
            st->dr.iirz[ch0 + k][1] = z[k][1];
        }
    }
}

/* Stage D with TMA tiles: same arithmetic as tbf_stage_d.  TWO: one warp per channel. */
template <bool TWO>
__global__ void __launch_bounds__(TWO ? 64 : 32) tbf_stage_d_tma(const RenderParams p, const __grid_constant__ TileMaps maps)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    DTmaSmem &sm = *reinterpret_cast<DTmaSmem *>(smem_raw + (((raw + 1023u) & ~1023u) - raw));
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t slot0 = blockIdx.x * 32;
    const uint32_t slot = slot0 + lane;
    const bool valid = slot < p.n_instances;

    InstState *st = nullptr;
    const DevWhirlTables *wh = p.wh;
    if (valid) {
        const uint32_t id = p.ids ? p.ids[slot] : p.id_base + slot;
        st = p.states + id;
        wh = p.wh + st->tg.whTables;
    }
    const bool run = valid && !wh->bypass;
    if (!__any_sync(FULL, run))
        return; /* every organ of the warp has its Leslie bypassed: nothing to filter */
    if (TWO)
        d_channel_role<1>(p, maps, sm, lane, slot0, run, st, wh, warp, warp);
    else
        d_channel_role<2>(p, maps, sm, lane, slot0, run, st, wh, 0, 0);
}

#ifndef M_THREADS
#define M_THREADS 256
#endif

/* V consecutive floats as one access (V = 4: LDG.128 / STG.128) */
template <int V> struct MVec {
    float v[V];
};
template <int V> __device__ __forceinline__ MVec<V> m_load(const float *p)
{
    MVec<V> r;
    if (V == 4) {
        const float4 q = *reinterpret_cast<const float4 *>(p);
        r.v[0] = q.x;
        r.v[V > 1 ? 1 : 0] = q.y;
        r.v[V > 2 ? 2 : 0] = q.z;
        r.v[V > 3 ? 3 : 0] = q.w;
    } else {
        r.v[0] = *p;
    }
    return r;
}
template <int V> __device__ __forceinline__ void m_store(float *p, const MVec<V> &r)
{
    if (V == 4)
        __stcs(reinterpret_cast<float4 *>(p), make_float4(r.v[0], r.v[V > 1 ? 1 : 0], r.v[V > 2 ? 2 : 0],
                                                          r.v[V > 3 ? 3 : 0])); /* streamed: never re-read */
    else
        *p = r.v[0];
}

/* Stage D, second kernel: horn level, leakage and the microphone mix (src/whirl.cpp:1603-1616,
 * whirlProc3 :1676-1680) -- elementwise over the five streams, lane = V consecutive samples (V = 4:
 * float4 loads and stores), one organ per blockIdx.x, gridDim.y CTAs along the samples; the
 * output checksum is formed here.  With TBF_RENDER_STEMS whirlProc2's separate horn and drum
 * outputs (outHL outHR outDL outDR, src/whirl.h:232-238) are stored next to L and R. */
template <int V> __global__ void __launch_bounds__(M_THREADS) tbf_stage_m(const RenderParams p)
{
    const uint32_t slot = blockIdx.x;
    const bool whirl3 = (p.flags & TBF_RENDER_WHIRLPROC3) != 0;
    const bool store = !(p.flags & TBF_RENDER_NO_OUTPUT);
    const uint32_t id = p.ids ? p.ids[slot] : p.id_base + slot;
    const DevWhirlTables *wh = p.wh + p.states[id].tg.whTables;
    const float horn_level = wh->horn_level, leakage = wh->leakage;
    const float hll = wh->mic_hll, hlr = wh->mic_hlr, hrl = wh->mic_hrl, hrr = wh->mic_hrr;
    const float dll = wh->mic_dll, dlr = wh->mic_dlr, drl = wh->mic_drl, drr = wh->mic_drr;
    const bool bypass = wh->bypass != 0;
    const uint32_t n = p.n_fragments * TBF_FRAGMENT;
    const size_t in0 = (size_t)slot * p.pitch, out0 = (size_t)slot * p.out_stride;
    unsigned long long csum = 0ull;
    for (uint32_t i = (blockIdx.y * M_THREADS + threadIdx.x) * V; i < n; i += gridDim.y * M_THREADS * V) {
        const MVec<V> h0 = m_load<V>(p.hx0 + in0 + i);
        MVec<V> o_l, o_r, s_hl, s_hr, s_dl, s_dr;
        if (bypass) {
            /* bypassed Leslie: input copied to both horn outputs, drum silent
             * (src/whirl.cpp:1197-1215), then whirlProc3's mic mix (:1676-1680) */
#pragma unroll
            for (int j = 0; j < V; ++j) {
                const float x = h0.v[j];
                s_hl.v[j] = s_hr.v[j] = x;
                s_dl.v[j] = s_dr.v[j] = 0.0f;
                if (!whirl3) {
                    o_l.v[j] = o_r.v[j] = x;
                } else {
                    o_l.v[j] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(x, hll), __fmul_rn(x, hlr)),
                                                   __fmul_rn(0.0f, dll)), __fmul_rn(0.0f, dlr));
                    o_r.v[j] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(x, hrl), __fmul_rn(x, hrr)),
                                                   __fmul_rn(0.0f, drl)), __fmul_rn(0.0f, drr));
                }
            }
        } else {
            const MVec<V> hl = m_load<V>(p.line[0] + in0 + i), hr = m_load<V>(p.line[1] + in0 + i);
            const MVec<V> yl = m_load<V>(p.line[2] + in0 + i), yr = m_load<V>(p.line[3] + in0 + i);
#pragma unroll
            for (int j = 0; j < V; ++j) {
                const float leak = __fmul_rn(h0.v[j], leakage);
                const float hpl = __fmul_rn(horn_level, hl.v[j]), hpr = __fmul_rn(horn_level, hr.v[j]);
                const float ohl = __fadd_rn(hpl, leak), ohr = __fadd_rn(hpr, leak);
                s_hl.v[j] = ohl;
                s_hr.v[j] = ohr;
                s_dl.v[j] = yl.v[j];
                s_dr.v[j] = yr.v[j];
                if (!whirl3) {
                    o_l.v[j] = __fadd_rn(__fadd_rn(yl.v[j], hpl), leak);
                    o_r.v[j] = __fadd_rn(__fadd_rn(yr.v[j], hpr), leak);
                } else {
                    o_l.v[j] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(ohl, hll), __fmul_rn(ohr, hlr)),
                                                   __fmul_rn(yl.v[j], dll)), __fmul_rn(yr.v[j], dlr));
                    o_r.v[j] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(ohl, hrl), __fmul_rn(ohr, hrr)),
                                                   __fmul_rn(yl.v[j], drl)), __fmul_rn(yr.v[j], drr));
                }
            }
        }
        if (store) {
            m_store<V>(p.out_l + out0 + i, o_l);
            m_store<V>(p.out_r + out0 + i, o_r);
            if (p.flags & TBF_RENDER_STEMS) {
                if (p.stems[0])
                    m_store<V>(p.stems[0] + out0 + i, s_hl);
                if (p.stems[1])
                    m_store<V>(p.stems[1] + out0 + i, s_hr);
                if (p.stems[2])
                    m_store<V>(p.stems[2] + out0 + i, s_dl);
                if (p.stems[3])
                    m_store<V>(p.stems[3] + out0 + i, s_dr);
            }
        }
        if (p.checksums) {
#pragma unroll
            for (int j = 0; j < V; ++j)
                csum += sample_hash((uint32_t)((size_t)p.frag_base * TBF_FRAGMENT + i + j), o_l.v[j], o_r.v[j]);
        }
    }
    if (p.checksums) { /* zeroed by the host; accumulates over chunks and CTAs */
#pragma unroll
        for (int d = 16; d > 0; d >>= 1)
            csum += __shfl_xor_sync(FULL, csum, d);
        if ((threadIdx.x & 31) == 0 && csum)
            atomicAdd(&p.checksums[slot], csum);
    }
}

/* instance (re)initialisation without a host-side template: every block clears one InstState and
 * installs the scalars allocTonegen + initToneGenerator + allocWhirl + initWhirl leave behind
 * (the caller fills InstInit from the table sets, tbf_api.cu) -- no upload, no synchronisation */
__global__ void tbf_init_instances_kernel(InstState *states, uint32_t first, uint32_t count, const InstInit init)
{
    for (uint32_t inst = blockIdx.x; inst < count; inst += gridDim.x) {
        InstState *st = states + first + inst;
        uint4 *dst = reinterpret_cast<uint4 *>(st);
        const uint32_t n16 = sizeof(InstState) / 16;
        for (uint32_t i = threadIdx.x; i < n16; i += blockDim.x)
            dst[i] = make_uint4(0u, 0u, 0u, 0u);
        for (uint32_t i = n16 * 16 / 4 + threadIdx.x; i < sizeof(InstState) / 4; i += blockDim.x)
            reinterpret_cast<uint32_t *>(st)[i] = 0u;
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < sizeof(TgScalars) / 4; i += blockDim.x)
            reinterpret_cast<uint32_t *>(&st->tg)[i] = reinterpret_cast<const uint32_t *>(&init.tg)[i];
        for (uint32_t i = threadIdx.x; i < sizeof(RampScalars) / 4; i += blockDim.x)
            reinterpret_cast<uint32_t *>(&st->rp)[i] = reinterpret_cast<const uint32_t *>(&init.rp)[i];
        for (uint32_t i = threadIdx.x; i <= TBF_NOF_WHEELS; i += blockDim.x)
            st->osc[i].aclPos = -1;
        __syncthreads();
    }
}

/* writeSamples (src/tonegen.cpp:1402-1457) on the device: blockIdx.x = wheel - 1, one thread per
 * sample.  Every double operation is the reference's, in its order, individually rounded; the
 * one thing that can differ from the host is libm: CUDA's sin() and glibc's may disagree in the
 * last bit of the double.  A sample whose double value lies so close to the midpoint of two
 * floats that such a disagreement could change the float it rounds to is FLAGGED (bound:
 * attenuation * 2^-46, thirty times what twelve partials of 2-ulp sines can add up to); the
 * host recomputes flagged samples with its own libm and patches them (tbf_api.cu) -- about one
 * sample in ten table sets. */
__global__ void tbf_wave_synth_kernel(float *__restrict__ raw, const uint32_t *__restrict__ meta,
                                      const double *__restrict__ wheel_f1, const double *__restrict__ wheel_att,
                                      const WaveSynthParams sp, const uint32_t *__restrict__ dither_bits,
                                      uint32_t *flag_count, uint4 *flags, uint32_t flag_cap)
{
    const uint32_t w = blockIdx.x + 1;
    const uint32_t len = meta[(TBF_NOF_WHEELS + 1) + w];
    if (!len)
        return;
    const uint32_t off = meta[w];
    const double full_circle = 6.283185307179586; /* 2.0 * M_PI */
    const double f1 = wheel_f1[w], att = wheel_att[w];
    double apl[TBF_MAX_PARTIALS], pl_hz[TBF_MAX_PARTIALS];
    double sum = 0.0;
#pragma unroll
    for (int j = 0; j < TBF_MAX_PARTIALS; ++j) {
        apl[j] = sp.harmonics[j];
        sum = __dadd_rn(sum, fabs(apl[j]));
        pl_hz[j] = __dmul_rn(f1, (double)(j + 1));
        if (__dmul_rn(sp.rate, 0.5) <= pl_hz[j])
            apl[j] = 0.0;
    }
    bool any_partial = false;
#pragma unroll
    for (int j = 0; j < TBF_MAX_PARTIALS; ++j)
        any_partial = any_partial || apl[j] != 0.0;
    const double U = __ddiv_rn(att, sum);
    const double bound = __dmul_rn(fabs(att), 1.4210854715202004e-14); /* 2^-46 */
    const float lsb = (float)(1.0 / 32767.0);
    for (uint32_t i = threadIdx.x; i < len; i += blockDim.x) {
        double sacc = 0.0;
#pragma unroll
        for (int j = 0; j < TBF_MAX_PARTIALS; ++j) {
            if (apl[j] == 0.0)
                continue;
            const double arg = __ddiv_rn(__dmul_rn(__dmul_rn(pl_hz[j], full_circle), (double)i), sp.rate);
            sacc = __dadd_rn(sacc, __dmul_rn(apl[j], sin(remainder(arg, full_circle))));
        }
        const uint32_t k = off + i; /* the sample's position in the rand() stream */
        const float d = ((dither_bits[k >> 5] >> (k & 31u)) & 1u) ? lsb : 0.0f;
        const double v = __dadd_rn((double)d, __dmul_rn(U, sacc));
        const float f = (float)v;
        raw[k] = f;
        /* distance of v from the two rounding boundaries around f */
        const double up = 0.5 * ((double)f + (double)nextafterf(f, INFINITY));
        const double dn = 0.5 * ((double)f + (double)nextafterf(f, -INFINITY));
        /* nothing to decide where no sine enters: a wheel with every partial muted (at or above
         * Nyquist) is the dither alone, and sample 0 is sin(0) = 0 on either side */
        if (any_partial && i > 0 && (fmin(fabs(up - v), fabs(v - dn)) <= bound || !(v == v))) {
            const uint32_t slot = atomicAdd(flag_count, 1u);
            if (slot < flag_cap)
                flags[slot] = make_uint4(w, i, __float_as_uint(f), 0u);
        }
    }
}

__global__ void tbf_replicate_kernel(InstState *states, uint32_t first, uint32_t count,
                                     const InstState *tmpl)
{
    const uint32_t words = sizeof(InstState) / 4;
    const uint32_t *src = reinterpret_cast<const uint32_t *>(tmpl);
    for (uint32_t inst = blockIdx.x; inst < count; inst += gridDim.x) {
        uint32_t *dst = reinterpret_cast<uint32_t *>(states + first + inst);
        for (uint32_t i = threadIdx.x; i < words; i += blockDim.x)
            dst[i] = src[i];
    }
}

/* table upload: the four rotated, wrap-extended copies of every wave (tbf_wave_copy_len) from
 * the raw waves: copy c of wheel w holds dst[i] = src[(i + c) % len], i < cl.  meta = wave_offset,
 * wave_length, wave4_offset, 257 entries each.  blockIdx.x = wheel, blockIdx.y = copy */
__global__ void tbf_wave_store_kernel(float *__restrict__ wave4, const float *__restrict__ raw,
                                      const uint32_t *__restrict__ meta)
{
    const uint32_t w = blockIdx.x + 1, c = blockIdx.y;
    const uint32_t len = meta[(TBF_NOF_WHEELS + 1) + w];
    if (!len)
        return;
    const float *src = raw + meta[w];
    const uint32_t cl = tbf_wave_copy_len(len);
    float *dst = wave4 + meta[2 * (TBF_NOF_WHEELS + 1) + w] + (size_t)c * cl;
    uint32_t pos = (threadIdx.x + c) % len;
    const uint32_t step = blockDim.x % len;
    for (uint32_t i = threadIdx.x; i < cl; i += blockDim.x) {
        dst[i] = src[pos];
        pos += step;
        if (pos >= len)
            pos -= len;
    }
}

/* cudaFuncAttributeMaxDynamicSharedMemorySize is a per-DEVICE attribute and one process may hold
 * a context per GPU (tbf_multi): the largest size opted into is tracked per device */
#define TBF_MAX_DEVICES 64
template <typename K> int set_smem(K kernel, size_t smem, int (&have)[TBF_MAX_DEVICES])
{
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= TBF_MAX_DEVICES)
        dev = TBF_MAX_DEVICES; /* unknown device: set the attribute every time */
    if (dev == TBF_MAX_DEVICES || have[dev] < (int)smem) {
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess)
            return (int)e;
        if (dev < TBF_MAX_DEVICES)
            have[dev] = (int)smem;
    }
    return 0;
}

} /* namespace */

/* TBF_DEBUG_SYNC=1: wait for every stage kernel and name the one that faulted (the pool has no
 * compute-sanitizer; -DTBF_DEBUG_CHECKS adds device-side invariant checks to stage T) */
static void debug_sync(void *stream, const char *stage)
{
    static const int on = getenv("TBF_DEBUG_SYNC") != nullptr;
    if (!on)
        return;
    const cudaError_t e = cudaStreamSynchronize((cudaStream_t)stream);
    if (e != cudaSuccess)
        fprintf(stderr, "tbf: stage %s: %s\n", stage, cudaGetErrorString(e));
}

int tbf_launch_wave_store(float *wave4, const float *raw, const uint32_t *meta, void *stream)
{
    tbf_wave_store_kernel<<<dim3(TBF_NOF_WHEELS, 4), 256, 0, (cudaStream_t)stream>>>(wave4, raw, meta);
    return (int)cudaGetLastError();
}

size_t tbf_stage_t_smem(void) { return (sizeof(TWarp) + 2 * 1024) * T_WARPS; }
size_t tbf_stage_w_smem(int ring_size)
{
    return (sizeof(WWarp) + (size_t)2 * ring_size * sizeof(float)) * W_WARPS;
}

template <int SLOTS> static int launch_stage_t(const RenderParams &p, cudaStream_t stream)
{
    static int have[TBF_MAX_DEVICES] = {0};
    /* experiment: TBF_T_PAD_KB pads the CTA's shared memory, i.e. caps stage T's CTAs per SM */
    static const size_t pad = getenv("TBF_T_PAD_KB") ? (size_t)atoi(getenv("TBF_T_PAD_KB")) * 1024 : 0;
    const size_t smem = (sizeof(TWarp) + (size_t)(SLOTS <= -10 ? 2 : SLOTS < 0 ? -SLOTS : SLOTS) * 1024) * T_WARPS + pad;
    if (int rc = set_smem(tbf_stage_t<SLOTS>, smem, have))
        return rc;
    const uint32_t grid = (p.n_instances + T_WARPS - 1) / T_WARPS;
    tbf_stage_t<SLOTS><<<grid, T_WARPS * 32, smem, stream>>>(p);
    return 0;
}

int tbf_launch_stage_t(const RenderParams &p, void *stream)
{
    /* How the wheels' samples reach the sums (profiles/r02_kernels.md has the sweeps):
     *   direct  -- LDG.128 straight into registers, a few records ahead; quiet fragments bus by bus
     *              over the records with a non-zero gain or masked to the buses in use (quiet_sum).
     *   tma     -- bulk copies (cp.async.bulk + mbarrier) through 2 / 3 / 6 staging slots per warp
     *   cpasync -- per-lane cp.async through 2 / 3 / 6 slots
     * TBF_T_LOADS = direct | tma | cpasync or the TBF_RENDER_DEBUG_T_* flag bits pick one,
     * TBF_T_SLOTS the depth. */
    static const char *env = getenv("TBF_T_LOADS");
    static const int env_mover = !env ? 0 : !strcmp(env, "tma") ? 1 : !strcmp(env, "cpasync") ? 2 : !strcmp(env, "direct") ? 3 : 0;
    static const int forced = getenv("TBF_T_SLOTS") ? atoi(getenv("TBF_T_SLOTS")) : 0;
    const int flag_mover = (int)((p.flags >> 12) & 3u);
    int mover = flag_mover ? flag_mover : env_mover;
    const uint32_t grid = (p.n_instances + T_WARPS - 1) / T_WARPS;
    const bool small = grid <= 148u * 3u; /* fewer than three CTAs per SM: a warp's own loads are all
                                             that hides their latency */
    /* Organs that share table sets read their waves out of L2 and direct loads win (config 3: 185 k
     * organ-s/s against 172 k with bulk copies).  With a table set per organ (config 4) every wave
     * comes from HBM, and what hides that latency is depth, without registers: six wheels in flight
     * through shared memory (per-lane cp.async 181 k, bulk copies 168 k, direct loads 154 - 164 k). */
    const bool from_hbm = p.tg_sets && (uint64_t)p.tg_sets * 8u > p.n_instances;
    if (!mover)
        mover = from_hbm ? 2 : 3;
    int rc;
    if (mover == 3) {
        const int depth = forced ? forced : (small ? 4 : 2);
        rc = depth >= 4   ? launch_stage_t<-14>(p, (cudaStream_t)stream)
             : depth == 3 ? launch_stage_t<-13>(p, (cudaStream_t)stream)
                          : launch_stage_t<-12>(p, (cudaStream_t)stream);
    } else {
        const int slots = forced ? forced : ((small || from_hbm) ? 6 : 2);
        if (mover == 2)
            rc = slots >= 6   ? launch_stage_t<-6>(p, (cudaStream_t)stream)
                 : slots >= 3 ? launch_stage_t<-3>(p, (cudaStream_t)stream)
                              : launch_stage_t<-2>(p, (cudaStream_t)stream);
        else
            rc = slots >= 6   ? launch_stage_t<6>(p, (cudaStream_t)stream)
                 : slots >= 3 ? launch_stage_t<3>(p, (cudaStream_t)stream)
                              : launch_stage_t<2>(p, (cudaStream_t)stream);
    }
    if (rc)
        return rc;
    debug_sync(stream, "T");
    return (int)cudaGetLastError();
}

/* the rotors run as their own kernel, ahead of the audio chain (Leslie running, TMA tiles available) */
bool tbf_stage_r_rotors_apart(const RenderParams &p)
{
    static const bool split = !(getenv("TBF_R_SPLIT") && atoi(getenv("TBF_R_SPLIT")) == 0);
    return split && p.tile_maps && p.ev_cursor_rot && !(p.flags & TBF_RENDER_WHIRL_BYPASS);
}

int tbf_launch_stage_rot(const RenderParams &p, void *stream)
{
    const uint32_t grid = (p.n_instances + 31) / 32;
    static const bool one_warp = getenv("TBF_ROT_WARPS") && atoi(getenv("TBF_ROT_WARPS")) == 1;
    if (one_warp)
        tbf_stage_rot_tma<false><<<grid, 32, sizeof(RotSmem) + 1024, (cudaStream_t)stream>>>(p, *p.tile_maps);
    else
        tbf_stage_rot_tma<true><<<grid, 64, sizeof(RotSmem) + 1024, (cudaStream_t)stream>>>(p, *p.tile_maps);
    debug_sync(stream, "R (rotors)");
    return (int)cudaGetLastError();
}

int tbf_launch_stage_r(const RenderParams &p, void *stream)
{
    static int have[TBF_MAX_DEVICES] = {0};
    if (int rc = set_smem(tbf_stage_r, sizeof(RSmem), have))
        return rc;
    const uint32_t grid = (p.n_instances + 31) / 32;
    if (p.tile_maps) {
        static int have_tma[TBF_MAX_DEVICES] = {0};
        const bool bypass = (p.flags & TBF_RENDER_WHIRL_BYPASS) != 0;
        const size_t smem = (bypass && !p.checksums) ? offsetof(RTmaSmem, tang) + 1024 : sizeof(RTmaSmem) + 1024;
        /* The audio chain as two warps per 32 organs (mixdown | horn biquads one block behind), the
         * rotors in a kernel of their own that tbf_api.cu launches ahead (tbf_launch_stage_rot).  Stage
         * R sits on the critical path of the stage schedule and what it costs there is its CTAs'
         * residency times the stage T / W CTAs they displace: half the chain per warp and 30 KB
         * instead of 46 KB per CTA.  TBF_R_SPLIT=0: everything in one kernel of two warps. */
        if (tbf_stage_r_rotors_apart(p)) {
            static int have_split[TBF_MAX_DEVICES] = {0};
            const size_t smem_a = offsetof(RTmaSmem, tang) + 1024;
            if (int rc = set_smem(tbf_stage_r_tma<true, false>, smem_a, have_split))
                return rc;
            tbf_stage_r_tma<true, false><<<grid, 64, smem_a, (cudaStream_t)stream>>>(p, *p.tile_maps);
            debug_sync(stream, "R");
            return (int)cudaGetLastError();
        }
        if (int rc = set_smem(tbf_stage_r_tma<false, true>, sizeof(RTmaSmem) + 1024, have_tma))
            return rc;
        /* the rotor warp has nothing to do when the Leslie is bypassed */
        tbf_stage_r_tma<false, true><<<grid, bypass ? 32 : 64, smem, (cudaStream_t)stream>>>(p, *p.tile_maps);
        debug_sync(stream, "R");
        return (int)cudaGetLastError();
    }
    const bool r_csum = p.checksums && (p.flags & TBF_RENDER_WHIRL_BYPASS);
    const size_t smem = sizeof(RSmem) - (r_csum ? 0 : sizeof(RSmem::csum));
    /* the rotor warp has nothing to do when the Leslie is bypassed */
    tbf_stage_r<<<grid, (p.flags & TBF_RENDER_WHIRL_BYPASS) ? 32 : 64, smem, (cudaStream_t)stream>>>(p);
    debug_sync(stream, "R");
    return (int)cudaGetLastError();
}

int tbf_launch_stage_w(const RenderParams &p, int ring_size, void *stream)
{
    static int have[TBF_MAX_DEVICES] = {0};
    /* experiment: TBF_W_PAD_KB pads the CTA's shared memory, i.e. caps stage W's CTAs per SM */
    static const size_t pad = getenv("TBF_W_PAD_KB") ? (size_t)atoi(getenv("TBF_W_PAD_KB")) * 1024 : 0;
    const size_t smem = tbf_stage_w_smem(ring_size) + pad;
    const dim3 grid((p.n_instances + W_WARPS - 1) / W_WARPS, 2); /* y: horn halves, drum halves */
    if (ring_size == 512) {
        if (int rc = set_smem(tbf_stage_w<512>, smem, have))
            return rc;
        tbf_stage_w<512><<<grid, W_WARPS * 32, smem, (cudaStream_t)stream>>>(p, 512u);
    } else {
        static int have_dyn[TBF_MAX_DEVICES] = {0};
        if (int rc = set_smem(tbf_stage_w<0>, smem, have_dyn))
            return rc;
        tbf_stage_w<0><<<grid, W_WARPS * 32, smem, (cudaStream_t)stream>>>(p, (uint32_t)ring_size);
    }
    debug_sync(stream, "W");
    return (int)cudaGetLastError();
}

int tbf_launch_stage_d(const RenderParams &p, void *stream)
{
    static int have[TBF_MAX_DEVICES] = {0};
    if (int rc = set_smem(tbf_stage_d, sizeof(DSmem), have))
        return rc;
    const uint32_t grid = (p.n_instances + 31) / 32;
    static const bool one_warp = getenv("TBF_D_WARPS") && atoi(getenv("TBF_D_WARPS")) == 1;
    if (p.tile_maps && one_warp)
        tbf_stage_d_tma<false><<<grid, 32, sizeof(DTmaSmem) + 1024, (cudaStream_t)stream>>>(p, *p.tile_maps);
    else if (p.tile_maps)
        tbf_stage_d_tma<true><<<grid, 64, sizeof(DTmaSmem) + 1024, (cudaStream_t)stream>>>(p, *p.tile_maps);
    else
        tbf_stage_d<<<grid, 32, sizeof(DSmem), (cudaStream_t)stream>>>(p);
    debug_sync(stream, "D");
    const uint32_t n = p.n_fragments * TBF_FRAGMENT;
    /* float4 accesses when every output row starts 16-byte aligned (the stage streams always do) */
    bool vec = (p.out_stride % 4) == 0;
    const float *outs[6] = {p.out_l, p.out_r, p.stems[0], p.stems[1], p.stems[2], p.stems[3]};
    for (const float *o : outs)
        vec = vec && (((uintptr_t)o) & 15u) == 0;
    if (vec) {
        const dim3 mgrid(p.n_instances, std::max(1u, std::min(4u, n / (4 * M_THREADS))));
        tbf_stage_m<4><<<mgrid, M_THREADS, 0, (cudaStream_t)stream>>>(p);
    } else {
        const dim3 mgrid(p.n_instances, std::max(1u, std::min(8u, n / (4 * M_THREADS))));
        tbf_stage_m<1><<<mgrid, M_THREADS, 0, (cudaStream_t)stream>>>(p);
    }
    debug_sync(stream, "M");
    return (int)cudaGetLastError();
}

int tbf_launch_init_instances(InstState *states, uint32_t first, uint32_t count, const InstInit &init,
                              void *stream)
{
    if (!count)
        return 0;
    const uint32_t grid = count < 1184 ? count : 1184; /* 148 SMs x 8 */
    tbf_init_instances_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(states, first, count, init);
    return (int)cudaGetLastError();
}

int tbf_launch_wave_synth(float *raw, const uint32_t *meta, const double *wheel_f1, const double *wheel_att,
                          const WaveSynthParams &sp, const uint32_t *dither_bits, uint32_t *flag_count,
                          void *flags, uint32_t flag_cap, void *stream)
{
    tbf_wave_synth_kernel<<<TBF_NOF_WHEELS, 256, 0, (cudaStream_t)stream>>>(
        raw, meta, wheel_f1, wheel_att, sp, dither_bits, flag_count, reinterpret_cast<uint4 *>(flags), flag_cap);
    return (int)cudaGetLastError();
}

int tbf_launch_replicate(InstState *states, uint32_t first, uint32_t count, const InstState *dev_tmpl,
                         void *stream)
{
    if (!count)
        return 0;
    const uint32_t grid = count < 1184 ? count : 1184; /* 148 SMs x 8 */
    tbf_replicate_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(states, first, count, dev_tmpl);
    return (int)cudaGetLastError();
}
