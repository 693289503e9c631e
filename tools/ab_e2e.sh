# e2e (compact mode) under different settings: each argument is an env assignment list ("-" = defaults),
# e.g.  bash tools/ab_e2e.sh - AS_ONE_STREAM=1 AS_CHUNK_READS=125000
for c in "$@"; do
  ( [ "$c" = "-" ] || export $c
    python bench.py --no-cpu --steps 5 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); e=d['e2e']; f=d['e2e_full_copy']
print('$c', 'compact', round(e['ms_per_step'],3), e['same_results_as_device_path'], e['stage_ms_last_step'], 'full', round(f['ms_per_step'],3), f['same_results_as_device_path'])" )
done
