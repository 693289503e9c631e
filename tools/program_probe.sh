#!/bin/bash
# GPU box: accaln_b200 FASTQ -> SAM on c2-shaped input (4.6 Mbp genome, N x 150 bp reads), stage timings from stderr.
# usage: bash tools/program_probe.sh [copies-of-1M-reads] [extra env...]
set -e
cd "$(dirname "$0")/.."
C=${1:-4}; shift || true
W=/tmp/prog_probe; mkdir -p $W
python - <<PY
import sys, numpy as np
sys.path.insert(0, '.')
import longreadmapper_b200 as lm
from bench import write_fastq_fast, GENOME_SEED, READS_SEED
g = lm.sim_genome(GENOME_SEED, 4_600_000)
open('$W/p.fa','wb').write(b'>chr1\n' + g.tobytes() + b'\n')
reads, lens, _, _ = lm.sim_reads(READS_SEED, g, 1_000_000, 150, 150, 0.02, 0.02, 0.8)
write_fastq_fast('$W/p.fq', reads, lens, $C)
PY
./longreadmapper_b200/accidx_b200 $W/p.fa
for i in 1 2; do env ACCSEQ_RG_ID=1 "$@" ./longreadmapper_b200/accaln_b200 $W/p.fa $W/p.fq 1000000 20 300 2>&1 >/dev/null | grep -E "Begin processing|Done aligning|stage busy|Sensitivity"; done
