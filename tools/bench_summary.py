"""one-screen summary of bench.py JSON lines: python tools/bench_summary.py gpurun_out/r2m_bench_*.json"""
import json
import sys
for f in sys.argv[1:]:
    try:
        d = json.load(open(f))
    except Exception as e:
        print(f, "ERR", e)
        continue
    r = d.get("roofline") or {}
    st = r.get("stages") or {}
    e = d.get("e2e") or {}
    print(f"{f}: value {d['value']:.4g} ms/step {d['ms_per_step']:.3f} | e2e {e.get('value') or 0:.4g} ms {e.get('ms_per_step') or 0:.3f} B/read {e.get('bytes_per_read')}")
    if st:
        print("   part1 %.3f ms (frac %.3f)  part2 %.3f  gact %.3f ms (frac %.3f, redone %.4f)" % (
            st["part1_seed_vote"]["ms"], st["part1_seed_vote"].get("frac") or 0, st["part2_locus"]["ms"], st["gact"]["ms"],
            st["gact"].get("frac") or 0, st["gact"].get("redone_full_matrix_frac") or 0))
    s = d.get("stats_per_step_per_gpu")
    if s:
        print("   ", {k: int(v) for k, v in s.items()})
    if e:
        print("    e2e stages", e.get("stage_ms_last_step"))
    for k in ("e2e_ascii_rle", "e2e_full_copy", "e2e_program", "cpu_baseline"):
        if d.get(k):
            print("   ", k, {a: b for a, b in d[k].items() if a in ("value", "ms_per_step", "seconds", "reads", "gcups", "cores")})
    print("    quality", d.get("quality"))
