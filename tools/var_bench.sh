#!/bin/bash
# Compare variant builds (build_var/*.so) against the in-tree library: n=100 headline + n=400 + mixed, no CPU leg.
mkdir -p gpurun_out
for so in rnagent_b200/librnafold_b200.so build_var/*.so; do
  [ -f "$so" ] || continue
  for w in ${VAR_WORKLOADS:-c2}; do
    RNAFOLD_LIB=$PWD/$so timeout 600 python bench.py --workload $w --steps ${VAR_STEPS:-5} --warmup 3 --no-extra --no-cpu 2>gpurun_out/var.err |
      python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$so $w', round(d['value']), 'folds/s  kernel_ms', round(d['roofline']['kernel_ms'],2), 'checksum', d.get('checksum_mfe_dcal'))" || tail -3 gpurun_out/var.err
  done
done
