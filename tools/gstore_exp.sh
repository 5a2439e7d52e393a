#!/bin/bash
for rep in 1 2 3; do
  timeout 400 python bench.py --no-retrieval --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(j['ms_per_step'],3), round(j['e2e']['value']), j['step_ms'])"
done
