#!/usr/bin/env python
"""Generates the golden fixtures of this directory by running the reference's OWN code, compiled from
/root/reference into oracle/_ref/ (oracle/Makefile):

  repeat.fa.gz        220 kbp repeat-rich two-sequence genome (seeded; tests/conftest.py make_repeat_rich)
  reads.fq.gz         reads of 100-250 bp at 0-8 % error + boundary/garbage/short cases
  expected.sam.gz     stdout of `accaln_ref repeat.fa reads.fq 100000 20 300`   (@RG line dropped: time based)
  index.sha256.json   sha256 of the five index files `accidx_ref repeat.fa` wrote (.mta .cat .cat.sa5 .cat.mfi .cat.lch)
  lookups.npz         lc_aln / fmi_aln results of the reference's object code (libref.so) on 4000 seeds, and
                      histo_add / histo_find_2_max traces on random key streams

FLAG/RNAME/POS/MAPQ/SEQ and every lookup are pinned by the reference's code.  CIGAR and ED come from the
GACT stand-in linked behind gact/gact.h (the lisanhu/GACT submodule is absent from the reference tree):
they pin the *stand-in spec*, not the original aligner.

Run here (needs /root/reference):  python tests/golden/make_golden.py
"""
import ctypes as C
import gzip
import hashlib
import json
import os
import shutil
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import longreadmapper_b200 as lm      # noqa: E402  (host helpers only: seeded generators)
import oracle_lib                      # noqa: E402
from conftest import make_repeat_rich  # noqa: E402


def golden_reads(a, b):
    comp = np.zeros(256, np.uint8)
    comp[list(b"ACGT")] = list(b"TGCA")
    seqs = []
    for gi, g in enumerate((a, b)):
        for k, (lmin, lmax, e) in enumerate([(100, 100, 0.0), (150, 150, 0.02), (250, 250, 0.05), (100, 250, 0.08)]):
            r, l, _, _ = lm.sim_reads(4000 + 10 * gi + k, g, 80, lmin, lmax, e, e, 0.7)
            seqs += [r[i, : l[i]].copy() for i in range(len(l))]
    rng = np.random.default_rng(9)
    seqs += [a[:150], a[-150:], b[:150], b[-150:], np.concatenate([a[-75:], b[:75]]),
             a[1000:1022], a[1000:1041], a[1000:1042],
             np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, 150)], np.full(150, ord("A"), np.uint8),
             np.tile(np.frombuffer(b"ACGTTGCA", np.uint8), 18), comp[a[-150:]][::-1].copy(), comp[a[:150]][::-1].copy()]
    return seqs


def main():
    assert oracle_lib.have_ref(), "build oracle/_ref first: make -C oracle ref"
    a, b = make_repeat_rich(lm, 23)
    seqs = golden_reads(a, b)
    wd = tempfile.mkdtemp(prefix="golden_")
    fa, fq = os.path.join(wd, "repeat.fa"), os.path.join(wd, "reads.fq")
    with open(fa, "wb") as f:
        f.write(b">seqA\n" + a.tobytes() + b"\n>seqB\n" + b.tobytes() + b"\n")
    with open(fq, "wb") as f:
        for i, s in enumerate(seqs):
            f.write(b"@g%d\n" % i + s.tobytes() + b"\n+\n" + b"I" * len(s) + b"\n")
    oracle_lib.run_ref_accidx(fa)                       # the reference's accidx: 5 index files
    sha = {}
    for ext in (".mta", ".cat", ".cat.sa5", ".cat.mfi", ".cat.lch"):
        sha[ext] = hashlib.sha256(open(fa + ext, "rb").read()).hexdigest()
    sam = oracle_lib.run_ref_accaln(fa, fq, batch=100000)
    sam = "\n".join(l for l in sam.splitlines() if not l.startswith("@RG")) + "\n"

    # lookups through the reference's object code
    R = oracle_lib.ref()
    fmi, lch = oracle_lib.RefFmi(), oracle_lib.RefLch()
    R.fmi_read(C.byref(fmi), (fa + ".cat").encode())     # fmi_read appends ".mfi" (fmidx.c:247)
    R.lc_read((fa + ".cat.lch").encode(), C.byref(lch))
    cat = np.frombuffer(open(fa + ".cat", "rb").read(), np.uint8)
    rng = np.random.default_rng(1)
    out = {}
    for qlen in (20, 12, 32, 64):
        pos = rng.integers(0, cat.size - 80, 1000)
        seeds = np.stack([cat[p:p + qlen] for p in pos]).copy()
        mut = rng.integers(0, 2, 1000).astype(bool)
        seeds[mut, rng.integers(0, qlen, 1000)[mut]] = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, int(mut.sum()))]
        res = np.zeros((1000, 6), np.uint64)
        for i in range(1000):
            q = seeds[i].tobytes()
            k, l = C.c_uint64(1), C.c_uint64(int(fmi.length) - 1)
            r = R.lc_aln(q, qlen, C.byref(k), C.byref(l), C.byref(fmi), C.byref(lch))
            k2, l2 = C.c_uint64(1), C.c_uint64(int(fmi.length) - 1)
            r2 = R.fmi_aln(C.byref(fmi), q, qlen, C.byref(k2), C.byref(l2))
            res[i] = (k.value, l.value, r, k2.value, l2.value, r2)
        out[f"seeds{qlen}"] = seeds
        out[f"res{qlen}"] = res
    # histo traces
    keys_all, top_all = [], []
    for t in range(40):
        n = int(rng.integers(1, 400))
        base = rng.integers(0, 1 << 40, 6)
        keys = (base[rng.integers(0, 6, n)] + rng.integers(0, 40, n)).astype(np.uint64)
        hh = R.histo_init(300)
        for kx in keys:
            R.histo_add(hh, int(kx))
        top = (oracle_lib.RefEntry * 2)()
        v = R.histo_find_2_max(hh, top)
        R.histo_destroy(hh)
        keys_all.append(keys)
        top_all.append([top[0].key, top[0].val, top[0].bucket, top[1].key, top[1].val, top[1].bucket, v])
    out["histo_keys"] = np.concatenate(keys_all)
    out["histo_lens"] = np.array([len(k) for k in keys_all], np.int64)
    out["histo_top"] = np.array(top_all, np.uint64)
    np.savez_compressed(os.path.join(HERE, "lookups.npz"), **out)

    for name, src in (("repeat.fa.gz", fa), ("reads.fq.gz", fq)):
        with open(src, "rb") as fi, gzip.GzipFile(os.path.join(HERE, name), "wb", mtime=0) as fo:
            shutil.copyfileobj(fi, fo)
    with gzip.GzipFile(os.path.join(HERE, "expected.sam.gz"), "wb", mtime=0) as fo:
        fo.write(sam.encode())
    json.dump(sha, open(os.path.join(HERE, "index.sha256.json"), "w"), indent=1)
    shutil.rmtree(wd)
    print("golden fixtures written:", sorted(os.listdir(HERE)))


if __name__ == "__main__":
    main()
