# A/B of the GACT band on the GPU box: gact stage time, filled cells and redone fraction per setting (c2, c3)
for b in "$@"; do
  for w in "c2" "c3 --reads 20000"; do
    python bench.py --no-cpu --no-e2e --steps 3 --warmup 4 --workload $w --gact-band $b | python -c "
import sys,json; d=json.loads(sys.stdin.read()); g=d['roofline']['stages']['gact']; s=d['stats_per_step_per_gpu']
print('band $b', '$w', 'step', round(d['ms_per_step'],3), 'gact', round(g['ms'],3), 'filled/spec', round(s['dp_cells_eval']/s['dp_cells'],3), 'redone', g['redone_full_matrix_frac'])"
  done
done
