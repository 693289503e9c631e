# 4-GPU evidence (run under `gpurun --gpus 4`): c2 weak (the driver's SCALE command) and c4 strong (ONE 10 M-read batch cut four ways)
T=${1:-rX}; N=${2:-4}; O=gpurun_out
mkdir -p $O
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --no-cpu --no-program"
$RUN > $O/${T}_bench_c2_${N}gpu_weak.json 2> $O/${T}_mg${N}.err
$RUN --workload c4 --steps 3 --scaling strong > $O/${T}_bench_c4_${N}gpu_strong.json 2>> $O/${T}_mg${N}.err
ls -la $O | tail -4
