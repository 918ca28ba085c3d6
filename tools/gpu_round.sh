#!/bin/bash
# One gpurun call of the round: GPU test tier, smoke, bench (both arms), ncu launch list + one full capture.
# Usage (from the repo root): gpurun --timeout 2400 -- 'bash tools/gpu_round.sh <tag>'
TAG=${1:-r01}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_gpu.csv
echo "== pytest -m gpu"; python -m pytest tests -x -q -m gpu > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "rc=$?"; tail -5 $OUT/${TAG}_pytest_gpu.log
echo "== smoke"; python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "rc=$?"; tail -3 $OUT/${TAG}_smoke.log
echo "== bench reference"; python bench.py --impl reference --steps 2 --warmup 1 > $OUT/${TAG}_bench_reference.json 2> $OUT/${TAG}_bench_reference.err; echo "rc=$?"; cut -c1-400 $OUT/${TAG}_bench_reference.json
echo "== bench"; python bench.py --steps 3 --warmup 3 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "rc=$?"; cat $OUT/${TAG}_bench.json; tail -3 $OUT/${TAG}_bench.err
echo "== ncu launch list (bench, reduced steps)"
python bench.py --steps 2 --warmup 3 --n-sys 1480 --cpu-sample 64 > $OUT/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv \
    python bench.py --steps 2 --warmup 3 --n-sys 1480 --cpu-sample 64 > $OUT/${TAG}_ncu_launches.log 2>&1
echo "rc=$?"
echo "== ncu full (one wave of DFN: 296 systems = two teams per SM)"
python -m tools.gpu_case dfn 296 1 > $OUT/${TAG}_case_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:pbgpu_solve_kernel -c 1 -f -o $OUT/${TAG}_dfn_full \
    python -m tools.gpu_case dfn 296 1 > $OUT/${TAG}_ncu_full.log 2>&1
echo "rc=$?"; tail -3 $OUT/${TAG}_ncu_full.log
ls -la $OUT
