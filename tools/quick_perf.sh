#!/bin/bash
# quick regression sweep on the GPU box: parity tests, env stepping rate, headline / n=400 / mixed fold rates
python -m pytest tests -m gpu -q -x 2>&1 | tail -2
timeout 200 python tools/env_speed.py 2>&1 | tail -1
timeout 200 python tools/env_speed.py mixed 2>&1 | tail -1
for w in c2 c3 c5; do
  python bench.py --workload $w --steps 3 --warmup 3 --no-cpu --no-extra | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$w', round(d['value']), 'folds/s kernel_ms', round(d['roofline']['kernel_ms'],2), 'e2e', round(d['e2e']['value']))"
done
