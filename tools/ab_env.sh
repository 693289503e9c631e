# device-path step and PART 1 under different env settings: each argument is an env assignment ("-" = defaults);
# WL="c2|c3 --reads 20000|c4" selects the workloads
IFS='|' read -ra WLS <<< "${WL:-c2|c3 --reads 20000}"
for c in "$@"; do
  for w in "${WLS[@]}"; do
  ( [ "$c" = "-" ] || export $c
    python bench.py --no-cpu --no-e2e --steps 3 --workload $w | python -c "import sys,json; d=json.loads(sys.stdin.read()); s=d['stats_per_step_per_gpu']; print('$c', '$w', round(d['ms_per_step'],3), {k: round(v['ms'], 2) for k, v in d['roofline']['stages'].items()}, 'fm', s['fm_steps_exec'], 'txt', s['text_verified'])" )
  done
done
