#!/bin/bash
# One gpurun call of the round: GPU test tier, smoke, bench (both arms + configs 1, 2, 4, 5), ncu launch list + full captures of the three kernel shapes.
# Usage (from the repo root): gpurun --timeout 2400 -- 'bash tools/gpu_round.sh <tag>'
TAG=${1:-r02}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_gpu.csv
echo "== pytest -m gpu"; python -m pytest tests -x -q -m gpu > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "rc=$?"; tail -3 $OUT/${TAG}_pytest_gpu.log
echo "== smoke"; python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "rc=$?"; tail -2 $OUT/${TAG}_smoke.log
echo "== bench reference"; python bench.py --impl reference --steps 3 --warmup 1 > $OUT/${TAG}_bench_reference.json 2> $OUT/${TAG}_bench_reference.err; echo "rc=$?"; cut -c1-300 $OUT/${TAG}_bench_reference.json
echo "== bench"; python bench.py --steps 5 --warmup 3 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "rc=$?"; cut -c1-400 $OUT/${TAG}_bench.json; tail -2 $OUT/${TAG}_bench.err
for m in spm spme; do
  n=10000; [ $m = spm ] && n=1000
  echo "== bench $m"; python bench.py --model $m --n-sys $n --steps 5 --warmup 3 > $OUT/${TAG}_bench_$m.json 2> $OUT/${TAG}_bench_$m.err; echo "rc=$?"; cut -c1-200 $OUT/${TAG}_bench_$m.json
  python bench.py --impl reference --model $m --n-sys $n --steps 3 --warmup 1 > $OUT/${TAG}_bench_${m}_reference.json 2> /dev/null
done
echo "== bench dfn_thermal40 (config 4)"; python bench.py --model dfn_thermal40 --steps 2 --warmup 3 > $OUT/${TAG}_bench_thermal40.json 2> $OUT/${TAG}_bench_thermal40.err; echo "rc=$?"; cut -c1-200 $OUT/${TAG}_bench_thermal40.json
python bench.py --impl reference --model dfn_thermal40 --steps 2 --warmup 1 > $OUT/${TAG}_bench_thermal40_reference.json 2> /dev/null
echo "== bench dfn_cccv (config 5)"; python bench.py --model dfn_cccv --steps 2 --warmup 1 > $OUT/${TAG}_bench_cccv.json 2> $OUT/${TAG}_bench_cccv.err; echo "rc=$?"; cut -c1-200 $OUT/${TAG}_bench_cccv.json
echo "== ncu launch list (bench, reduced steps)"
python bench.py --steps 2 --warmup 3 --n-sys 1480 --cpu-sample 64 --parity-sample 16 --no-full-state > $OUT/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv \
    python bench.py --steps 2 --warmup 3 --n-sys 1480 --cpu-sample 64 --parity-sample 16 --no-full-state > $OUT/${TAG}_ncu_launches.log 2>&1
echo "rc=$?"
echo "== ncu full (one wave of DFN: 296 systems = two teams per SM)"
python -m tools.gpu_case dfn 296 1 > $OUT/${TAG}_case_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:pbgpu_solve_kernel -c 1 -f -o $OUT/${TAG}_dfn_full \
    python -m tools.gpu_case dfn 296 1 > $OUT/${TAG}_ncu_full.log 2>&1
echo "rc=$?"; tail -2 $OUT/${TAG}_ncu_full.log
# the reports are 20 MB each and gpurun_out/ is capped: export what profiles/ keeps and drop them
export_rep() {   # <report stem> <model>
  ncu -i $OUT/$1.ncu-rep --page raw --csv > $OUT/$1_ncu_raw.csv 2> /dev/null
  python -m tools.ncu_lines $OUT/$1.ncu-rep pybamm_b200/_build/pbgpu_model_$2.so > $OUT/$1_ncu_by_source_line.txt 2>&1
  rm -f $OUT/$1.ncu-rep
}
export_rep ${TAG}_dfn_full dfn
echo "== ncu full, warp-per-system path (SPMe, one wave)"
python -m tools.gpu_case spme 1332 1 > $OUT/${TAG}_case_spme_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:pbgpu_solve_kernel -c 1 -f -o $OUT/${TAG}_spme_full \
    python -m tools.gpu_case spme 1332 1 > $OUT/${TAG}_ncu_spme_full.log 2>&1
echo "rc=$?"
export_rep ${TAG}_spme_full spme
echo "== ncu full, config 4 (one team per SM: 148 systems)"
python -m tools.gpu_case dfn_thermal40 148 1 > $OUT/${TAG}_case_thermal_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:pbgpu_solve_kernel -c 1 -f -o $OUT/${TAG}_thermal40_full \
    python -m tools.gpu_case dfn_thermal40 148 1 > $OUT/${TAG}_ncu_thermal_full.log 2>&1
echo "rc=$?"
export_rep ${TAG}_thermal40_full dfn_thermal40
echo "== phases"; python -m tools.gpu_phases dfn 1480 > $OUT/${TAG}_phases_dfn.txt 2>&1; python -m tools.gpu_cycles dfn 296 >> $OUT/${TAG}_phases_dfn.txt 2>&1; tail -12 $OUT/${TAG}_phases_dfn.txt
ls -la $OUT | tail -30
