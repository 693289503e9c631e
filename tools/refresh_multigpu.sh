# Multi-GPU evidence (run under `gpurun --gpus N`): the C host's N-GPU SAM test, then the bench on N ranks -- c2 weak (the driver's
# SCALE command), c2 strong, and optionally c4 strong (one 10 M-read batch cut N ways: config 4 as BASELINE.json words it).
#   bash tools/refresh_multigpu.sh r2x 2 [c4]
T=${1:-rX}; N=${2:-2}; O=gpurun_out
mkdir -p $O
python -m pytest tests/test_driver.py -m gpu -q 2>&1 | tail -3 > $O/${T}_driver_tests_${N}gpu.txt
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --no-cpu --no-program"
$RUN > $O/${T}_bench_c2_${N}gpu_weak.json 2> $O/${T}_mg.err
$RUN --scaling strong > $O/${T}_bench_c2_${N}gpu_strong.json 2>> $O/${T}_mg.err
if [ "$3" = "c4" ]; then
  $RUN --workload c4 --steps 3 --scaling strong > $O/${T}_bench_c4_${N}gpu_strong.json 2>> $O/${T}_mg.err
fi
ls -la $O | tail -8
