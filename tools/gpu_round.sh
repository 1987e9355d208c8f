#!/bin/bash
# One GPU-box visit of round 2: parity tests, smoke, bench (no profiler), reference arm, then the ncu passes
# (launch lists of the same bench commands; --set full captures of the wave, cluster and LDS-peak kernels).
# usage: tools/gpu_round.sh <tag> [noprof]
set -u
tag=${1:-r2}; noprof=${2:-}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 1200 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; rc=$?; echo "bench rc=$rc"
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_ref.json 2> gpurun_out/${tag}_bench_ref.err; echo "ref rc=$?"
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/${tag}_bench.json; cat gpurun_out/${tag}_bench_ref.json; tail -3 gpurun_out/${tag}_bench.err
if [ $rc -eq 0 ] && [ -z "$noprof" ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-extra --no-cpu > gpurun_out/ncu_bench.log 2>&1; echo "ncu c2 rc=$?"
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches_c3.csv \
    python bench.py --workload c3 --steps 2 --warmup 3 --no-extra --no-cpu > gpurun_out/ncu_bench_c3.log 2>&1; echo "ncu c3 rc=$?"
  bash tools/gpu_prof.sh ${tag}_wave 'fold_wave_kernel' 100 23680
  bash tools/gpu_prof.sh ${tag}_cluster 'fold_cluster_kernel' 400 592
  timeout 600 ncu --set full --clock-control none -k regex:lds_stream_kernel -s 2 -c 1 -f -o gpurun_out/prof_${tag}_peaks \
    python bench.py --steps 1 --warmup 3 --no-extra --no-cpu > gpurun_out/prof_${tag}_peaks.log 2>&1; echo "ncu peaks rc=$?"
fi
