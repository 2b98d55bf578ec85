#!/bin/bash
O=gpurun_out; mkdir -p $O
for spec in "$@"; do
  name=${spec%%:*}; envs=${spec#*:}; envs=${envs//,/ }
  env $envs timeout 600 python bench.py --steps 2 --warmup 3 --no-tp --no-profile --cpu-budget-s 25 > $O/par_$name.json 2> $O/par_$name.err; echo "$name exit $?"
  python -c "
import json
d=json.loads(open('gpurun_out/par_$name.json').read().strip().splitlines()[-1])
print('$name', d['parity']['max_rel_err'], d['parity']['steps'], d['parity']['prompt_tokens'], d['parity']['tokens_equal'])"
done
