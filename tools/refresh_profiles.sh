# End-of-change evidence run on the GPU box: tests, bench lines, reference arm, ncu launch list and one full capture of
# the two dominant kernels.  Everything lands in gpurun_out/<tag>_*; copy what should be judged into profiles/.
#   bash tools/refresh_profiles.sh r2q
T=${1:-rX}
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -q 2>&1 | tail -4 > $O/${T}_gputests.txt
python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > $O/${T}_bench_reference_arm.json 2> $O/${T}_ref.err
python bench.py --no-cpu --no-program --workload c2r > $O/${T}_bench_c2r_repeat_rich.json 2>> $O/${T}_bench.err
tools/gather_probe 8 > $O/${T}_gather_probe.txt 2>&1
python bench.py --no-cpu --no-program --workload c3 --reads 20000 > $O/${T}_bench_c3_20k.json 2>> $O/${T}_bench.err
python bench.py --no-cpu --no-program --workload c3 --steps 3 > $O/${T}_bench_c3_full.json 2>> $O/${T}_bench.err
python bench.py --no-cpu --no-program --workload c5 --steps 2 --warmup 1 > $O/${T}_bench_c5_1gpu.json 2>> $O/${T}_bench.err
python bench.py --no-cpu --workload c4 --steps 3 > $O/${T}_bench_c4_human_scale.json 2>> $O/${T}_bench.err
python bench.py --no-cpu --no-program --workload c4 --steps 3 --opt sampled_sa=8 > $O/${T}_bench_c4_sampled_sa8.json 2>> $O/${T}_bench.err
python bench.py --no-cpu --no-program --no-e2e --steps 2 --warmup 1 > $O/${T}_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_launches.csv \
    python bench.py --no-cpu --no-program --no-e2e --steps 2 --warmup 1 > $O/${T}_ncu_list.log 2>&1
python bench.py --no-cpu --no-e2e --no-program --steps 1 --warmup 1 > $O/${T}_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_vote|k_gact_bp' -c 5 -o $O/${T}_full -f \
    python bench.py --no-cpu --no-e2e --no-program --steps 1 --warmup 1 > $O/${T}_ncu_full.log 2>&1
ls -la $O | tail -14
