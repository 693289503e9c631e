# Trimmed evidence run (tests, c2 bench, reference arm, c2r, launch list, full ncu capture); see refresh_profiles.sh for the long one.
#   bash tools/refresh_quick.sh r2v
T=${1:-rX}
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -q 2>&1 | tail -4 > $O/${T}_gputests.txt
python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > $O/${T}_bench_reference_arm.json 2> $O/${T}_ref.err
python bench.py --no-cpu --no-program --workload c2r > $O/${T}_bench_c2r_repeat_rich.json 2>> $O/${T}_bench.err
python bench.py --no-cpu --no-program --no-e2e --steps 2 --warmup 1 > $O/${T}_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_launches.csv \
    python bench.py --no-cpu --no-program --no-e2e --steps 2 --warmup 1 > $O/${T}_ncu_list.log 2>&1
python bench.py --no-cpu --no-e2e --no-program --steps 1 --warmup 1 > $O/${T}_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_vote|k_gact_bp' -c 5 -o $O/${T}_full -f \
    python bench.py --no-cpu --no-e2e --no-program --steps 1 --warmup 1 > $O/${T}_ncu_full.log 2>&1
ls -la $O | tail -14
