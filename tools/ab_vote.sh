# A/B builds of k_vote on the GPU box: rebuild as_vote.cu with each flag set, print step and PART 1 times.
# EXTRA="--opt l2_fetch_gran=32" adds bench flags to every run
# WL="c2|c3 --reads 20000|c4" selects the workloads (default c2 and c3/20k)
IFS='|' read -ra WLS <<< "${WL:-c2|c3 --reads 20000}"
for v in "$@"; do
  touch longreadmapper_b200/csrc/as_vote.cu
  make -s -C longreadmapper_b200/csrc cuda EXTRA_NVFLAGS="$v" > /dev/null 2>&1
  for w in "${WLS[@]}"; do
    python bench.py --no-cpu --no-e2e --no-program --steps 5 $EXTRA --workload $w | python -c "import sys,json; d=json.loads(sys.stdin.read()); s=d['stats_per_step_per_gpu']; print('$v', '$EXTRA', '$w', round(d['ms_per_step'],3), round(d['roofline']['stages']['part1_seed_vote']['ms'],3), 'fm', s['fm_steps_exec'], 'txt', s['text_verified'])"
  done
done
