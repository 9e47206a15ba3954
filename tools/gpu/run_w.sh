#!/bin/bash
# Final check of the committed build: full GPU suite, smoke(), default bench.
set -u
O=gpurun_out/w
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/status.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/status.txt
timeout 900 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/status.txt
cat $O/status.txt; tail -3 $O/pytest.log; tail -1 $O/smoke.log
python - <<PY
import json
d=json.loads(open('$O/bench.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','value_full_head','eager_trainer_tokens_per_s','gpu_launches')}, d['e2e']['value'], d['roofline']['achieved'], d['roofline']['frac'], d['cpu_baseline']['value'], d['clocks'])
for o in d['roofline']['other_kernels']: print(o['kernel'][:70], o['us'], o['frac'])
PY
