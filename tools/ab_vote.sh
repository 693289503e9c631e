for v in "-DAS_OCC_256=0 -DVOTE_CLAIM=1" "-DAS_OCC_256=1 -DVOTE_CLAIM=1" "-DAS_OCC_256=0 -DVOTE_CLAIM=4" "-DAS_OCC_256=0 -DVOTE_CLAIM=16"; do
  touch longreadmapper_b200/csrc/as_vote.cu
  make -s -C longreadmapper_b200/csrc cuda EXTRA_NVFLAGS="$v" > /dev/null 2>&1
  python bench.py --no-cpu --no-e2e --steps 3 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', d['ms_per_step'], d['roofline']['stages']['part1_seed_vote']['ms'])"
done
