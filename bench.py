#!/usr/bin/env python
"""bench.py -- mapped reads/s of the AccSeqV9 per-read mapping hot path on B200 (BASELINE.json's metric).

A step = one pass of the hot path (alnmain.c:330-479: seeding + voting, locus/strand, GACT, result fill) over
one batch of synthetic reads.  N=1 workload = BASELINE.json configs[1]: 4.6 Mbp uniform synthetic genome
(distribution of scripts/dna_txt_gen.py), 1M x 150 bp reads at 2 % error; with N ranks every rank maps its
own 1M-read shard against its own replica of the index (weak scaling, no data-path collective; one
all-reduce of the shard statistics at the end).

  value        whole-job reads/s with the batch resident in HBM (as_map_batch_device), CUDA-event timed
  e2e          the same through as_map_batch with pinned HOST buffers (H2D + kernels + D2H in the timed region)
  roofline     the dominant kernel of the step against its bound (GACT: INT32 ALU; seeding/voting: memory)
  cpu_baseline the reference's CPU code (oracle/_ref/accaln_ref, P single-thread processes on P read
               slices -- the reference's mapping loop is single-threaded) on a bounded sample
  --impl reference   times only that CPU arm.

Only the cpu legs touch oracle/ (to write the index files accaln_ref loads and to run it).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: genome bases, reads per rank, read length range, error range, substitution fraction
    "c1": dict(genome=1_000_000, reads=10_000, len=(150, 150), err=(0.01, 0.01), sub=0.9),
    "c2": dict(genome=4_600_000, reads=1_000_000, len=(150, 150), err=(0.02, 0.02), sub=0.8),
    # c2 with planted repeat families (about 40 % of the genome, the share of a mammalian genome): 12-copy 1.5 kbp units at
    # 1 % divergence, 50- and 250-copy exact 300-bp units.  Reads inside the high-copy families overflow the shared-memory
    # vote table and are redone by k_vote<32,BIG>: this workload times that path.  NOT a BASELINE config.
    "c2r": dict(genome=4_600_000, reads=1_000_000, len=(150, 150), err=(0.02, 0.02), sub=0.8,
                repeats=[(60, 1500, 12, 0.01), (20, 300, 50, 0.0), (6, 300, 250, 0.0)]),
    "c3": dict(genome=100_000_000, reads=100_000, len=(10_000, 10_000), err=(0.10, 0.10), sub=1 / 3),
    # human scale: a 6.2 G-row .cat, 40-bit SA rows, 64-bit backward search; the index (~45 GB) is replicated per GPU
    "c4": dict(genome=3_100_000_000, reads=10_000_000, len=(150, 150), err=(0.02, 0.02), sub=0.8),
    "c5": dict(genome=3_100_000_000, reads=125_000, len=(1_000, 50_000), err=(0.05, 0.15), sub=1 / 3),
}
GENOME_SEED = 20240
READS_SEED = 777


def make_genome(wl_name):
    """the workload's genome: seeded uniform bases (scripts/dna_txt_gen.py's distribution), plus planted repeat families"""
    import longreadmapper_b200 as lm
    wl = WORKLOADS[wl_name]
    g = lm.sim_genome(GENOME_SEED, wl["genome"])
    for k, (fam, unit, copies, div) in enumerate(wl.get("repeats", [])):
        lm.sim_plant_repeats(GENOME_SEED + 1 + k, g, n_families=fam, unit_len=unit, copies=copies, divergence=div)
    return g


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: every rank maps its own batch of the workload's size (default); strong: ONE batch of that size "
                         "cut into contiguous per-rank shards (config 4 as BASELINE.json words it)")
    ap.add_argument("--reads", type=int, default=0, help="override reads per rank (smaller = NOT the named config)")
    ap.add_argument("--cpu-sample", type=int, default=0, help="reads in the CPU baseline sample (0 = auto)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-program", action="store_true", help="skip the accaln_b200 FASTQ -> SAM program leg")
    ap.add_argument("--knob", action="append", default=[], metavar="K=V",
                    help="as_params scheduling knob (vote_ectx, gact_order, gact_blocks_per_sm, chunk_reads, chunk_tail); A/B runs")
    ap.add_argument("--opt", action="append", default=[], metavar="K=V",
                    help="as_open_opts field (lc2_hlen, no_sax, build_part_max, sampled_sa); A/B runs")
    ap.add_argument("--gact-band", type=int, default=None,
                    help="half-width of GACT's banded first pass (default: adaptive; -1 = full matrix only); results do not depend on it")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------------
# clocks: nvidia-smi sampled during the timed region (B200_PROFILING.md)
# ---------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        rows = [l for (t, l) in self.lines if t0 <= t <= t1 + 0.2] or [l for (_, l) in self.lines]
        for l in rows:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------------
# CPU arm: the reference's own code (oracle/_ref/accaln_ref), P processes on P disjoint read slices
# ---------------------------------------------------------------------------------------------------
def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def write_fastq(path, reads, lens, start=0):
    stride = reads.shape[1]
    with open(path, "wb") as f:
        for i in range(reads.shape[0]):
            L = int(lens[i])
            f.write(b"@r%d\n" % (start + i))
            f.write(reads[i, :L].tobytes())
            f.write(b"\n+\n")
            f.write(b"I" * L)
            f.write(b"\n")
    return stride


def parse_mlog_seconds(text, begin_key, end_key):
    """mlog lines look like `[ Verbose     0.25 ] msg` (mlog/logger.c:75); returns t(end) - t(begin)."""
    tb = te = None
    for line in text.splitlines():
        if "]" not in line or "[" not in line:
            continue
        head, msg = line.split("]", 1)
        try:
            t = float(head.replace("[", " ").split()[-1])
        except (ValueError, IndexError):
            continue
        if begin_key in msg and tb is None:
            tb = t
        if end_key in msg:
            te = t
    if tb is None or te is None:
        return None
    return te - tb


def cpu_gcups(reads, lens, g, procs, budget_s=4.0):
    """GACT cells per second of the CPU aligner behind accaln_ref (orc_gact: the stand-in linked behind gact.h), same cell
    definition as the GPU's `gcups` (tile area Tq x Tt of the spec, summed over the tiles): `procs` threads, each aligning reads
    against the windows they were simulated from for about budget_s seconds."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib
    import threading as th
    L = oracle_lib.lib()
    done = [0] * procs
    t_end = time.time() + budget_s

    def work(k):
        import ctypes as C
        cells = C.c_uint64(0)
        nops = C.c_int(0)
        i = k
        ops = np.zeros(2 * int(lens.max()) + 16, np.uint8)
        while time.time() < t_end:
            ln = int(lens[i % len(lens)])
            q = reads[i % len(lens), :ln]
            p0 = (i * 7919) % (g.size - ln)
            t = g[p0:p0 + ln]
            L.orc_gact(q.ctypes.data, ln, t.ctypes.data, ln, 320, 120, ops.ctypes.data, C.byref(nops), C.byref(cells))
            i += procs
        done[k] = int(cells.value)

    t0 = time.time()
    ts = [th.Thread(target=work, args=(k,)) for k in range(procs)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    return sum(done) / (time.time() - t0) / 1e9


def cpu_reference_run(wl_name, sample_reads, procs, workdir):
    """Runs the compiled reference (or, when oracle/_ref is absent, the oracle port) on `sample_reads` reads of the
    workload.  Returns dict(value reads/s, cores, kind, sample, seconds)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib
    import longreadmapper_b200 as lm
    wl = WORKLOADS[wl_name]
    g = make_genome(wl_name)
    reads, lens, _, _ = lm.sim_reads(READS_SEED, g, sample_reads, wl["len"][0], wl["len"][1], wl["err"][0], wl["err"][1],
                                     wl["sub"])
    t0 = time.time()
    ix = oracle_lib.OracleIndex.build(["chr1"], [g.tobytes()])
    t_index = time.time() - t0
    if oracle_lib.have_ref():
        fa = os.path.join(workdir, "g.fa")
        with open(fa, "wb") as f:
            f.write(b">chr1\n")
            f.write(g.tobytes())
            f.write(b"\n")
        ix.write(fa)
        ix.close()
        procs = max(1, min(procs, sample_reads))
        per = (sample_reads + procs - 1) // procs
        fqs = []
        for p in range(procs):
            s, e = p * per, min(sample_reads, (p + 1) * per)
            if s >= e:
                break
            fq = os.path.join(workdir, f"r{p}.fq")
            write_fastq(fq, reads[s:e], lens[s:e], s)
            fqs.append((fq, e - s))
        exe = os.path.join(oracle_lib.REF_DIR, "accaln_ref")
        t0 = time.time()
        ps = [subprocess.Popen([exe, fa, fq, str(n), "20", "300"], stdout=subprocess.DEVNULL, stderr=subprocess.PIPE,
                               text=True) for fq, n in fqs]
        errs = [p.communicate()[1] for p in ps]
        wall = time.time() - t0
        if any(p.returncode != 0 for p in ps):
            raise RuntimeError("accaln_ref failed: " + errs[0][-400:])
        # mapping-only time per process: "Begin processing queries" -> "Done aligning" (alnmain.c:311,540)
        secs = [parse_mlog_seconds(e, "Begin processing", "Done aligning") for e in errs]
        secs = [s if s and s > 0 else wall for s in secs]
        t = max(secs)
        gc = cpu_gcups(reads, lens, g, len(fqs))
        return dict(value=sample_reads / t, unit="reads/s", cores=len(fqs), kind="reference", gcups=gc,
                    gcups_note=f"orc_gact (the GACT stand-in accaln_ref links) on {len(fqs)} threads, reads against unrelated windows of their "
                               "length: every tile is filled in full, cells = tile area as in the GPU's gcups",
                    sample=f"{sample_reads} reads of {wl_name} ({len(fqs)} accaln_ref processes, 1 thread each; index load "
                           f"excluded; GACT = the oracle stand-in linked behind gact.h, the original submodule is absent)",
                    seconds=t, wall_seconds=wall, index_build_seconds=t_index)
    # port: the oracle's restatement, one thread
    stride = reads.shape[1]
    t0 = time.time()
    ix.map_batch(reads, stride, lens, myers=False)
    t = time.time() - t0
    ix.close()
    return dict(value=sample_reads / t, unit="reads/s", cores=1, kind="port",
                sample=f"{sample_reads} reads of {wl_name} (oracle/oracle.c orc_map_batch, 1 thread)", seconds=t)


def write_fastq_fast(path, reads, lens, copies=1):
    """FASTQ of fixed-length reads, vectorised: @r<8 digits> / seq / + / IIII...; `copies` repeats the batch under new names"""
    n, L = reads.shape[0], int(lens[0])
    rec = 2 + 8 + 1 + L + 3 + L + 1
    with open(path, "wb") as f:
        for c in range(copies):
            buf = np.empty((n, rec), np.uint8)
            buf[:, 0], buf[:, 1] = ord("@"), ord("r")
            ids = np.arange(c * n, (c + 1) * n, dtype=np.int64)
            for d in range(8):
                buf[:, 2 + 7 - d] = ord("0") + (ids // 10 ** d) % 10
            buf[:, 10] = ord("\n")
            buf[:, 11:11 + L] = reads[:, :L]
            buf[:, 11 + L:14 + L] = np.frombuffer(b"\n+\n", np.uint8)
            buf[:, 14 + L:14 + 2 * L] = ord("I")
            buf[:, 14 + 2 * L] = ord("\n")
            f.write(buf.tobytes())


def program_e2e(wl_name, g, reads, lens, n_gpus, workdir, copies):
    """The whole program, FASTQ file in -> SAM out (to /dev/null, as the reference arm's): accidx_b200 writes the index files,
    accaln_b200 maps.  Window = the reference arm's: 'Begin processing queries' -> 'Done aligning' (alnmain.c:311,540), which on
    both sides holds FASTQ parsing, mapping and SAM formatting, and excludes loading the index."""
    pkg = os.path.join(ROOT, "longreadmapper_b200")
    fa, fq = os.path.join(workdir, "p.fa"), os.path.join(workdir, "p.fq")
    with open(fa, "wb") as f:
        f.write(b">chr1\n")
        f.write(g.tobytes())
        f.write(b"\n")
    t0 = time.time()
    subprocess.run([os.path.join(pkg, "accidx_b200"), fa], check=True)
    t_idx = time.time() - t0
    if int(lens.min()) == int(lens.max()):
        write_fastq_fast(fq, reads, lens, copies)
    else:
        copies = 1
        write_fastq(fq, reads, lens)
    n_prog = copies * reads.shape[0]
    env = dict(os.environ, ACCSEQ_DEVICES=",".join(str(i) for i in range(n_gpus)), ACCSEQ_RG_ID="1")
    best = None
    for _ in range(2):                                    # the first run also pages the file in
        with open(os.devnull, "wb") as dn:
            p = subprocess.run([os.path.join(pkg, "accaln_b200"), fa, fq, "1000000", "20", "300"], stdout=dn, stderr=subprocess.PIPE,
                               env=env, text=True)
        if p.returncode != 0:
            raise RuntimeError("accaln_b200 failed: " + p.stderr[-400:])
        t = parse_mlog_seconds(p.stderr, "Begin processing", "Done aligning")
        best = t if best is None or (t and t < best) else best
    return {"value": n_prog / best if best else None, "unit": "reads/s", "reads": n_prog, "seconds": best, "n_gpus": n_gpus,
            "index_build_and_write_s": round(t_idx, 2), "host_threads": "ingest + emit: nproc/2 each, one mapping thread per GPU",
            "window": "Begin processing queries -> Done aligning (FASTQ parse + map + SAM format; SAM to /dev/null), as the reference arm's",
            "program": "accaln_b200 ref.fa reads.fq 1000000 20 300"}


def auto_cpu_sample(wl_name, cores, budget_s):
    # measured single-thread rates of the reference build (SURVEY 6 + GACT stand-in): ~6k reads/s at 150 bp, ~8/s at 10 kbp
    rate = {"c1": 6000.0, "c2": 6000.0, "c2r": 3000.0, "c3": 8.0}[wl_name]
    return int(max(cores * 50, min(WORKLOADS[wl_name]["reads"], rate * cores * budget_s)))


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = host_cores()
    wl = args.workload
    sample = args.cpu_sample or auto_cpu_sample(wl, cores, 20.0)
    per_step = []
    with tempfile.TemporaryDirectory(prefix="accseq_ref_") as wd:
        for i in range(args.warmup + args.steps):
            if i < args.warmup and i > 0:
                continue                      # one warm-up pass is enough to page the binaries and files in
            r = cpu_reference_run(wl, sample, cores, wd)
            if i >= args.warmup:
                per_step.append(r)
    t = float(np.mean([r["seconds"] for r in per_step]))
    v = sample / t
    r0 = per_step[0]
    line = {"impl": "reference", "metric": "mapped reads/sec", "value": v, "unit": "reads/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * t, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u64/i32", "data": "synthetic",
            "config": config_of(wl, args),
            "run": {"reads_per_step": sample, "note": "a step is a bounded sample of the workload's reads (the CPU takes seconds per 100 k reads)"},
            "cpu_baseline": {"value": v, "unit": "reads/s", "cores": r0["cores"], "kind": r0["kind"], "sample": r0["sample"],
                             "gcups": r0.get("gcups"), "gcups_note": r0.get("gcups_note")},
            "e2e": {"value": v, "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


def config_of(wl, args):
    """what both arms run: the same dict, key for key, in the B200 line and in the --impl reference line"""
    w = WORKLOADS[wl]
    return {"workload": workload_name(wl, args), "genome_bases": w["genome"], "reads_per_step_per_gpu": args.reads or w["reads"],
            "read_len": list(w["len"]), "error_rate": list(w["err"]), "seed_len": 20, "thres": 300, "gact_T": 320, "gact_O": 120}


def workload_name(wl, args):
    w = WORKLOADS[wl]
    n = args.reads or w["reads"]
    rep = ", repeat families planted " + str(w["repeats"]) if w.get("repeats") else ""
    return (f"{wl}: {w['genome'] / 1e6:g} Mbp uniform synthetic genome (dna_txt_gen.py distribution, seeded){rep}, "
            f"{n} reads/GPU of {w['len'][0]}-{w['len'][1]} bp at {w['err'][0] * 100:g}-{w['err'][1] * 100:g}% error")


# ---------------------------------------------------------------------------------------------------
# B200 arm
# ---------------------------------------------------------------------------------------------------
# bit-parallel GACT (as_gact_bp.cu): SASS count of the inner word step = 24 integer instructions per 32-row word of one
# column (DESIGN.md "GACT kernel"), i.e. 0.75 per DP cell; and 2 bits per cell of traceback planes written to HBM
INT_OPS_PER_CELL = 24.0 / 32.0
PLANE_BYTES_PER_CELL = 0.25


def run_b200(args):
    import torch
    import torch.distributed as dist
    import longreadmapper_b200 as lm

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if lm.device_count() < 1 or not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the hot path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    numa = None
    if world > 1 and not os.environ.get("AS_NO_NUMA_BIND"):
        from longreadmapper_b200.shard import bind_to_gpu_numa
        numa = bind_to_gpu_numa(local)      # before any pinned buffer exists: first touch puts them on the GPU's node
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    wl = WORKLOADS[args.workload]
    n = args.reads or wl["reads"]
    n_job = world * n
    strong = args.scaling == "strong" and world > 1
    if strong:
        # the config as written: ONE batch of the workload's reads, cut into contiguous shards (shard.py:shard_range)
        from longreadmapper_b200.shard import shard_range
        n_job = n
        lo, hi = shard_range(n_job, rank, world)
        n = hi - lo

    # ---- inputs: the index replica of this rank (built on the GPU) and this rank's read shard ----
    g = make_genome(args.workload)
    t0 = time.time()
    oo = {}
    for kv in args.opt:
        k, v = kv.split("=")
        oo[k] = int(v)
    mapper = lm.GpuMapper.build([g], local, opts=lm.open_opts(**oo) if oo else None)
    t_index = time.time() - t0
    dev = torch.device("cuda", local)
    if strong:
        reads, lens, truth_pos, truth_strand = (np.ascontiguousarray(x[lo:hi]) for x in
                                                lm.sim_reads(READS_SEED, g, n_job, wl["len"][0], wl["len"][1], wl["err"][0],
                                                             wl["err"][1], wl["sub"]))
    else:
        reads, lens, truth_pos, truth_strand = lm.sim_reads(READS_SEED + 1000 * rank, g, n, wl["len"][0], wl["len"][1],
                                                            wl["err"][0], wl["err"][1], wl["sub"])
    stride = reads.shape[1]
    max_len = int(lens.max())
    store_stride = 2 * max_len
    knobs = {}
    for kv in args.knob:
        k, v = kv.split("=")
        knobs[k] = int(v)
    params = mapper.params(band=args.gact_band, **knobs)

    # one non-default stream for everything that is timed: the library launches on it (as_gpu_set_stream), the
    # restore copy and the CUDA events are issued on it
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    mapper.set_stream(stream.cuda_stream)
    d_pristine = torch.from_numpy(reads).to(dev)
    d_reads = torch.empty_like(d_pristine)
    d_lens = torch.from_numpy(lens.view(np.int32)).to(dev)
    d_res = torch.empty(n * lm.RESULT_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    d_store = torch.empty(n * store_stride, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()

    def step_device():
        d_reads.copy_(d_pristine)            # strand-1 reads are reverse-complemented in place by every pass
        mapper.map_batch_device(d_reads.data_ptr(), stride, d_lens.data_ptr(), n, max_len, d_res.data_ptr(),
                                d_store.data_ptr(), store_stride, params)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step_device()
    barrier()
    mapper.stats_reset()
    launches0 = mapper.launch_count()
    stage_ms = {}
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.25)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_wall0 = time.time()
    e0.record(stream)
    for _ in range(args.steps):
        step_device()
    e1.record(stream)
    barrier()
    t_wall1 = time.time()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop(t_wall0, t_wall1)
    launches = mapper.launch_count() - launches0 + args.steps       # + the restore copy of every step
    tm = mapper.last_timing()
    st = mapper.stats()
    from longreadmapper_b200 import shard
    ms = shard.max_over_ranks(ms, dev)
    # the one collective of the design: shard statistics (alnmain.c:541 Sensitivity line)
    st = shard.reduce_stats(st, dev)
    ms_per_step = ms / args.steps
    value = n_job / (ms_per_step / 1e3)

    # correctness of what was timed: the results of the last step against the simulated truth
    res = np.frombuffer(d_res.cpu().numpy().tobytes(), dtype=lm.RESULT_DTYPE)
    ok = res["valid"] == 1
    same_strand = (res["strand"] == 1) == (truth_strand == 1)
    exact = float(((res["off"] == truth_pos) & ok & same_strand).mean())
    near = float(((np.abs(res["off"].astype(np.int64) - truth_pos.astype(np.int64)) <= 50) & ok & same_strand).mean())

    # ---- e2e: host buffers through the C ABI ----
    e2e = None
    e2e_full = None
    e2e_ascii = None
    if not args.no_e2e:
        L = mapper.L
        import ctypes as C
        nb_reads, nb_res, nb_store = n * stride, n * lm.RESULT_DTYPE.itemsize, n * store_stride
        # every step gets its own pristine pinned copy of the batch (the full-copy mode reverse-complements strand-1
        # reads in place), so no host-side restore sits inside the timed region
        n_e2e_warm = max(1, min(args.warmup, 2))
        n_bufs = n_e2e_warm + args.steps
        hp = [L.as_host_alloc(x) for x in (4 * n, nb_res, nb_store)] + [L.as_host_alloc(nb_reads) for _ in range(n_bufs)]
        if not all(hp):
            raise SystemExit("bench.py: pinned allocation failed")
        h_lens = np.ctypeslib.as_array(C.cast(hp[0], C.POINTER(C.c_uint32)), shape=(n,))
        h_lens[:] = lens
        dops = d_store.cpu().numpy().reshape(n, store_stride)
        wmax = int(res["n_cigar_op"].max()) if n else 0

        def run_e2e(flags, pipeline):
            ep = mapper.params(band=args.gact_band, flags=flags, **knobs)
            for q in hp[3:]:
                np.ctypeslib.as_array(C.cast(q, C.POINTER(C.c_uint8)), shape=(n, stride))[:] = reads

            def step_e2e(k):
                rc = L.as_map_batch(mapper.h, C.byref(ep), hp[3 + k], stride, hp[0], n, hp[1], hp[2], store_stride)
                if rc != 0:
                    raise SystemExit(f"as_map_batch failed ({rc}): {L.as_last_error().decode()}")

            for k in range(n_e2e_warm):
                step_e2e(k)
            barrier()
            f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            f0.record(stream)
            for k in range(args.steps):
                step_e2e(n_e2e_warm + k)
            f1.record(stream)
            barrier()
            ems = shard.max_over_ranks(f0.elapsed_time(f1), dev)
            e2e_tm = mapper.last_timing()
            h2d_b, d2h_b = mapper.last_transfer()
            hres = np.ctypeslib.as_array(C.cast(hp[1], C.POINTER(C.c_uint8)), shape=(nb_res,)).view(lm.RESULT_DTYPE)
            hops = np.ctypeslib.as_array(C.cast(hp[2], C.POINTER(C.c_uint8)), shape=(n, store_stride))
            if flags & lm.AS_F_RLE_CIGAR:
                # same results; the run bytes that came back cover exactly the device path's columns, read by read
                same = all(np.array_equal(hres[f], res[f]) for f in res.dtype.names if f != "n_cigar_op")
                m = hres["n_cigar_op"].astype(np.int64)
                w = int(m.max()) if n else 0
                cols = (hops[:, :w] >> 2).astype(np.int64) * (np.arange(w)[None, :] < m[:, None])
                same = same and bool(np.array_equal(cols.sum(axis=1), np.maximum(res["n_cigar_op"], 0)))
            else:
                # the CIGAR columns that came back equal the device path's, column by column up to n_cigar_op
                same = bool(np.array_equal(hres, res)) and bool(np.array_equal(hops[:, :wmax], dops[:, :wmax]))
            return {"value": n_job / (ems / args.steps / 1e3), "unit": "reads/s",
                    "h2d_bytes_per_step": h2d_b, "d2h_bytes_per_step": d2h_b,
                    "ms_per_step": ems / args.steps, "same_results_as_device_path": bool(same), "pipeline": pipeline,
                    "stage_ms_last_step": {k: round(v, 3) for k, v in e2e_tm.items()}}

        def run_e2e_packed():
            """the call accaln_b200 makes: reads 2-bit packed by the ingest threads, compact records + dense run bytes back"""
            wst = (max_len + 15) // 16
            packed = lm.pack_reads(reads, lens, wst)
            runs_cap = n * 64 if max_len <= 512 else int(lens.astype(np.int64).sum())
            bufs = [L.as_host_alloc(x) for x in (n * wst * 4, n * 16, runs_cap)]
            if not all(bufs):
                raise SystemExit("bench.py: pinned allocation failed")
            np.ctypeslib.as_array(C.cast(bufs[0], C.POINTER(C.c_uint32)), shape=(n, wst))[:] = packed
            ep = mapper.params(band=args.gact_band, **knobs)
            got = C.c_uint64(0)

            def step():
                rc = L.as_map_batch_packed(mapper.h, C.byref(ep), bufs[0], wst, hp[0], n, bufs[1], bufs[2], runs_cap, C.byref(got))
                if rc != 0:
                    raise SystemExit(f"as_map_batch_packed failed ({rc}): {L.as_last_error().decode()}")

            for _ in range(n_e2e_warm):
                step()
            barrier()
            f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            f0.record(stream)
            for _ in range(args.steps):
                step()
            f1.record(stream)
            barrier()
            ems = shard.max_over_ranks(f0.elapsed_time(f1), dev)
            e2e_tm = mapper.last_timing()
            h2d_b, d2h_b = mapper.last_transfer()
            rec = np.ctypeslib.as_array(C.cast(bufs[1], C.POINTER(C.c_uint8)), shape=(n * 16,)).view(lm.COMPACT_DTYPE).copy()
            dense = np.ctypeslib.as_array(C.cast(bufs[2], C.POINTER(C.c_uint8)), shape=(int(got.value),)).copy()
            # same records as the device-resident path; the dense run bytes add up to its CIGAR columns read by read
            mapped = res["seq_id"] >= 0
            same = bool(np.array_equal(rec["pos"] & ((1 << 40) - 1), res["off"] & ((1 << 40) - 1))
                        and np.array_equal(((rec["pos"] >> 40) & 0xFFFF).astype(np.int64) - 1, res["seq_id"].astype(np.int64))
                        and np.array_equal((rec["pos"] >> 56).astype(np.int32), res["flag"]) and np.array_equal(rec["ed"], res["ed"]))
            nr = (rec["n_runs"] & 0x07FFFFFF).astype(np.int64)
            offs = np.concatenate([[0], np.cumsum(nr)])
            cols = np.add.reduceat((dense >> 2).astype(np.int64), offs[:-1][nr > 0]) if dense.size else np.zeros(0, np.int64)
            same = same and bool(offs[-1] == dense.size) and bool(np.array_equal(cols, res["n_cigar_op"][nr > 0]))
            same = same and bool(np.array_equal(nr > 0, mapped & (res["n_cigar_op"] > 0)))
            same = same and bool(np.array_equal((rec["n_runs"] >> 31).astype(np.int32), res["strand"]))
            for q in bufs:
                L.as_host_free(q)
            return {"value": n_job / (ems / args.steps / 1e3), "unit": "reads/s",
                    "h2d_bytes_per_step": h2d_b, "d2h_bytes_per_step": d2h_b, "bytes_per_read": round((h2d_b + d2h_b) / n, 1),
                    "ms_per_step": ems / args.steps, "same_results_as_device_path": bool(same),
                    "pipeline": "as_map_batch_packed: reads cross PCIe 2-bit packed (the ingest threads pack while they parse), "
                                "16-byte compact records and the run bytes back to back come back; chunks: every kernel enqueued up front, "
                                "H2D(i+1) | kernels(i) | D2H(i-1), the dense run bytes on a stream of their own",
                    "stage_ms_last_step": {k: round(v, 3) for k, v in e2e_tm.items()},
                    "host_us_last_step": {k: round(v, 1) for k, v in mapper.last_host_timing().items()}}

        chunked = "chunked: H2D(i+1) | kernels(i) | D2H(i-1) on three streams; "
        # what the host driver (accaln_b200) calls
        e2e = run_e2e_packed()
        # ASCII reads in (the refactor_reads_seq buffer), CIGAR as rows of run bytes, reads not copied back
        e2e_ascii = run_e2e(lm.AS_F_RLE_CIGAR | lm.AS_F_KEEP_READS,
                            chunked + "AS_F_RLE_CIGAR | AS_F_KEEP_READS: CIGAR crosses PCIe as run bytes, strand-1 reads are "
                            "reverse-complemented by the SAM emitter")
        # the reference's buffer layout to the byte: one op byte per column, strand-1 reads reverse-complemented in place
        e2e_full = run_e2e(0, chunked + "CIGAR columns trimmed to the widest row of a chunk, reads copied back")
        for p in hp:
            L.as_host_free(p)

    # ---- roofline of the dominant kernel (stage times: CUDA events inside the library, on the timing stream) ----
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)"
    steps = args.steps
    per_step = {k: st[k] / steps / world for k in st}
    # algorithmic bytes of seeding+voting per SURVEY 8(d): 16 B lc pair per seed + 49 B per backward step + 5 B per SA hit
    vote_bytes = per_step["seeds"] * 16 + per_step["bs_steps"] * 49 + per_step["sa_hits"] * 5 + n * max_len / 4
    sm_clock = (clocks.get("sm_mhz") or 1965.0) * 1e6
    # LOP3/SHF/IADD3 go to the alu pipe, which takes one warp instruction per 2 cycles per SMSP (B300_MICROARCH.md "Pipe
    # rates"; ncu's sm__inst_executed_pipe_alu agrees): 148 SMs x 4 SMSPs x 16 lanes x max clock, in Tops/s
    int_peak = 148 * 64 * 1965.0e6 / 1e12
    stages = {
        "part1_seed_vote": {"ms": tm["part1"], "bound": "hbm", "achieved": vote_bytes / (tm["part1"] / 1e3) / 1e9 if tm["part1"] else None,
                            "peak": hbm_peak, "unit": "GB/s"},
        "part2_locus": {"ms": tm["part2"]},
        "gact": {"ms": tm["gact"], "bound": "int32",
                 "gcups": per_step["dp_cells"] / (tm["gact"] / 1e3) / 1e9 if tm["gact"] else None,
                 # gcups counts the spec's tile area (Tq x Tt per tile); the banded first pass fills only the words that
                 # meet |i-j| <= band, so the INT32 work and plane bytes actually issued follow dp_cells_eval
                 "gcups_filled": per_step["dp_cells_eval"] / (tm["gact"] / 1e3) / 1e9 if tm["gact"] else None,
                 "achieved": per_step["dp_cells_eval"] * INT_OPS_PER_CELL / (tm["gact"] / 1e3) / 1e12 if tm["gact"] else None,
                 "peak": int_peak, "unit": "Tops/s",
                 "hbm_plane_write_gbs": per_step["dp_cells_eval"] * PLANE_BYTES_PER_CELL / (tm["gact"] / 1e3) / 1e9 if tm["gact"] else None,
                 "redone_full_matrix_frac": per_step["gact_redone"] / n if n else None},
    }
    for s in stages.values():
        if s.get("achieved") is not None:
            s["frac"] = s["achieved"] / s["peak"]
    try:        # pipe utilisation ncu measured for the committed capture of this command (profiles/ncu_traffic.json)
        tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json"))).get(args.workload, {})
        if "k_gact_bp_alu_pipe_pct" in tj:
            stages["gact"]["alu_pipe_pct_ncu"] = tj["k_gact_bp_alu_pipe_pct"]
    except (OSError, ValueError):
        pass
    # DRAM bytes per launch of the dominant kernels, from the committed `ncu --set full` capture of this same command
    # (profiles/ncu_traffic.json: dram__bytes_read.sum + dram__bytes_write.sum); null for workloads it was not taken on
    traffic = {}
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json"))).get(args.workload, {})
    except (OSError, ValueError):
        pass
    dom = max(("part1_seed_vote", "gact"), key=lambda k: stages[k]["ms"])
    d = stages[dom]
    # The ncu capture is stamped with the kernel time it saw (ncu's own gpu__time_duration) and the commit it was taken at: the
    # bytes are reported only while this run's live launch time is within 5 % of it -- a changed kernel makes them null, not stale.
    kname = "k_gact_bp" if dom == "gact" else "k_vote"
    cap = traffic.get(kname + "_capture") or {}
    tr = traffic.get(kname) if n == wl["reads"] else None
    if tr is not None and cap.get("ms"):
        if abs(d["ms"] - cap["ms"]) > 0.05 * cap["ms"]:
            tr = None
    roofline = {"kernel": kname, "bound": d["bound"], "achieved": d["achieved"],
                "peak": d["peak"], "unit": d["unit"], "frac": d.get("frac"),
                "traffic": tr, "traffic_capture": cap or None,
                "physical_frac": (tr / (d["ms"] / 1e3) / 1e9 / hbm_peak) if tr else None,
                "algorithmic": vote_bytes if dom != "gact" else per_step["dp_cells_eval"] * INT_OPS_PER_CELL,
                "peak_source": peak_src if d["bound"] == "hbm" else "alu pipe: 148 SMs x 64 INT32 lanes x clocks.max.sm 1965 MHz",
                "ms_per_launch": d["ms"], "share_of_step": d["ms"] / tm["total"] if tm["total"] else None,
                "stages": stages, "sm_clock_mhz_under_load": sm_clock / 1e6}

    line = {"metric": "mapped reads/sec", "value": value, "unit": "reads/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong" if strong else "weak",
            "vs_baseline": None, "dtype": "u64/i32", "data": "synthetic",
            "config": config_of(args.workload, args),
            "run": {"reads_per_step_whole_job": n_job, "l2": f"inputs exceed L2: reads {n * stride / 1e6:.0f} MB + lc table 268 MB + SA {4 * (2 * wl['genome'] + 1) / 1e6:.0f} MB "
                          f"+ GACT planes per pass > 126 MB, no flush needed",
                    "index": "built on the GPU (as_gpu_build), replicated per rank", "index_build_s": round(t_index, 2),
                    "host_numa_node_rank0": numa, "knobs": knobs, "open_opts": oo},
            "clocks": clocks, "e2e": e2e, "e2e_ascii_rle": e2e_ascii, "e2e_full_copy": e2e_full, "gpu_launches": int(launches),
            "roofline": roofline,
            "gcups": stages["gact"]["gcups"],
            "quality": {"mapped_frac": float(ok.mean()), "exact_pos_strand_frac": exact, "within_50bp_strand_frac": near,
                        "valid_total": [st["valid"], st["total"]]},
            "stats_per_step_per_gpu": {k: per_step[k] for k in ("seeds", "bs_steps", "fm_steps_exec", "text_verified", "sa_hits", "dp_cells",
                                                                   "dp_cells_eval", "gact_redone", "gact_queued", "overflow_reads", "overflow_queued", "vote_deferred")}}

    # the program leg builds its own index on the same GPUs: give the memory back first (human scale: ~120 GB per handle)
    mapper.close()
    del d_pristine, d_reads, d_lens, d_res, d_store
    torch.cuda.empty_cache()
    if world > 1:
        dist.barrier()
    if rank == 0 and not args.no_program and wl["len"][1] <= 512:
        try:
            with tempfile.TemporaryDirectory(prefix="accseq_prog_") as wd:
                line["e2e_program"] = program_e2e(args.workload, g, reads, lens, world, wd, copies=max(1, 4_000_000 // n))
        except Exception as ex:
            line["e2e_program"] = {"value": None, "error": str(ex)[:300]}
    if world > 1:
        dist.barrier()
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = host_cores()
        sample = args.cpu_sample or auto_cpu_sample(args.workload, cores, 12.0)
        try:
            with tempfile.TemporaryDirectory(prefix="accseq_cpu_") as wd:
                cb = cpu_reference_run(args.workload, sample, cores, wd)
            line["cpu_baseline"] = {k: cb.get(k) for k in ("value", "unit", "cores", "kind", "sample", "gcups", "gcups_note")}
        except Exception as ex:          # the CPU leg must not lose the GPU numbers
            line["cpu_baseline"] = {"value": None, "unit": "reads/s", "cores": cores, "kind": "reference", "sample": f"failed: {ex}"}
    elif rank == 0:
        line["cpu_baseline"] = None
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
