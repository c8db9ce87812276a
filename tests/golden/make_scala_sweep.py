#!/usr/bin/env python
"""BASELINE config 4 fixture: every .scl of the reference's Scala archive
(tests/scala_scale_archive/scl, 5176 files) through (1) this repo's parser + table builder
and (2) the reference engine (oracle/_ref) given the same 128 frequencies.

Writes tunebfree_b200/data/scala_sweep.npz (package data: the product's config-4 workload reads it):
  names    [S]        file names
  status   [S] int8   0 ok: tables bit-identical on both sides
                      1 unparsable by tbf_scl_ratios (not a scale this repo accepts)
                      2 reference aborts (assert) on this tuning; `ours_rc` is what the
                        builder returned for it
                      3 MISMATCH (must not occur)
                      4 reference fine, builder refused (`ours_rc`)
  ours_rc  [S] int32  tbf_build_tonegen_tables' return code
  deg_off  [S+1]      CSR offsets into `degrees`
  degrees  [...] f64  the scale's degrees as ratios (tbf_scl_ratios); freqs =
                      tbf_freqs_from_ratios(degrees)
  digest   [S,2] u64  blake2b-128 of the REFERENCE's tone-generator tables (status 0)

Run here (needs /root/reference): python tests/golden/make_scala_sweep.py [--jobs 8]
The reference aborts the process on some tunings, so the work is done by restartable worker
processes that log "begin <i>" before every reference call.
"""
import argparse
import ctypes as C
import glob
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np

ARCHIVE = "/root/reference/tests/scala_scale_archive/scl"
OUT = os.path.join(ROOT, "tunebfree_b200", "data", "scala_sweep.npz")


def files():
    return sorted(os.path.basename(p) for p in glob.glob(os.path.join(ARCHIVE, "*.scl")))


def worker(start, end, log_path):
    from oracle import ref_engine
    from tunebfree_b200 import cabi, tables
    names = files()
    with open(log_path, "a") as log:
        for i in range(start, end):
            text = open(os.path.join(ARCHIVE, names[i]), encoding="latin-1").read()
            rec = {"i": i}
            try:
                deg = tables.scl_ratios(text)
            except ValueError:
                rec["status"] = 1
                log.write(json.dumps(rec) + "\n")
                log.flush()
                continue
            freqs = tables.freqs_from_ratios(deg)
            assert np.array_equal(freqs, tables.freqs_from_scl(text))
            rec["degrees"] = [float(x).hex() for x in deg]
            spec = tables.InstanceSpec(freqs=freqs, seed=1)
            tt = cabi.TonegenTables()
            rc = tables._lib().tbf_build_tonegen_tables(C.byref(spec.tg), C.byref(tt))
            rec["ours_rc"] = rc
            ours = tables.table_digest(tt) if rc == cabi.OK else None
            if rc == cabi.OK:
                tables.free_tables(tt, None)
            log.write("begin %d %s\n" % (i, json.dumps(rec)))
            log.flush()
            r = ref_engine.RefInstance(freqs=freqs, seed=1)
            theirs = tables.table_digest(r.tonegen_tables())
            r.close()
            rec["digest"] = list(theirs)
            rec["status"] = 0 if ours == theirs else (4 if ours is None else 3)
            log.write(json.dumps(rec) + "\n")
            log.flush()


def drive_range(start, end, log_path):
    """Run a worker over [start, end); restart after the file on which the reference aborted."""
    if os.path.exists(log_path):
        os.remove(log_path)
    pos = start
    while pos < end:
        p = subprocess.run([sys.executable, os.path.abspath(__file__), "--worker", str(pos), str(end),
                            log_path], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        done, pending = pos - 1, None
        for line in open(log_path):
            if line.startswith("begin "):
                _, i, js = line.split(" ", 2)
                pending = (int(i), json.loads(js))
            else:
                rec = json.loads(line)
                done = rec["i"]
                pending = None
        if p.returncode == 0:
            break
        if pending is None:
            raise RuntimeError("worker died outside the reference call after file %d" % done)
        rec = pending[1]
        rec["status"] = 2
        with open(log_path, "a") as log:
            log.write(json.dumps(rec) + "\n")
        pos = pending[0] + 1


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--worker", nargs=3)
    ap.add_argument("--range", nargs=3)
    ap.add_argument("--jobs", type=int, default=8)
    a = ap.parse_args()
    if a.worker:
        worker(int(a.worker[0]), int(a.worker[1]), a.worker[2])
        return
    if a.range:
        drive_range(int(a.range[0]), int(a.range[1]), a.range[2])
        return
    names = files()
    S = len(names)
    tmp = os.path.join(ROOT, "_exp", "scala_sweep")
    os.makedirs(tmp, exist_ok=True)
    step = (S + a.jobs - 1) // a.jobs
    procs = []
    for j in range(a.jobs):
        lo, hi = j * step, min(S, (j + 1) * step)
        procs.append(subprocess.Popen([sys.executable, os.path.abspath(__file__), "--range", str(lo), str(hi),
                                       os.path.join(tmp, "part%02d.jsonl" % j)]))
    for p in procs:
        if p.wait() != 0:
            raise RuntimeError("a sweep driver failed")
    status = np.full(S, -1, dtype=np.int8)
    ours_rc = np.zeros(S, dtype=np.int32)
    digest = np.zeros((S, 2), dtype=np.uint64)
    degs = [[] for _ in range(S)]
    for j in range(a.jobs):
        for line in open(os.path.join(tmp, "part%02d.jsonl" % j)):
            if line.startswith("begin "):
                continue
            rec = json.loads(line)
            i = rec["i"]
            status[i] = rec["status"]
            ours_rc[i] = rec.get("ours_rc", 0)
            degs[i] = [float.fromhex(x) for x in rec.get("degrees", [])]
            if "digest" in rec:
                digest[i] = rec["digest"]
    assert (status >= 0).all()
    off = np.zeros(S + 1, dtype=np.int64)
    off[1:] = np.cumsum([len(d) for d in degs])
    np.savez_compressed(OUT, names=np.array(names), status=status, ours_rc=ours_rc, deg_off=off,
                        degrees=np.array([x for d in degs for x in d], dtype=np.float64), digest=digest)
    for s, label in enumerate(["ok", "unparsable", "reference aborts", "MISMATCH", "builder refused"]):
        print("%-18s %d" % (label, int((status == s).sum())))


if __name__ == "__main__":
    main()
