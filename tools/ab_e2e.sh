# e2e (compact mode) under different chunk plans: AS_CHUNK_READS=<fixed size>, or the built-in ramp
for c in "$@"; do
  if [ "$c" = "auto" ]; then unset AS_CHUNK_READS; else export AS_CHUNK_READS=$c; fi
  python bench.py --no-cpu --steps 5 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); e=d['e2e']; f=d['e2e_full_copy']
print('chunk $c', 'compact', round(e['ms_per_step'],3), e['stage_ms_last_step'], 'full', round(f['ms_per_step'],3))"
done
