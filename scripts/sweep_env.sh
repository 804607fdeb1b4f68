#!/bin/bash
# usage: sweep_env.sh "<bench args>" VAR=val[,VAR=val] ...   — prints QPS / kernel ms / roofline frac per setting
args="$1"; shift
for setting in "$@"; do
  envs=$(echo "$setting" | tr ',' ' ')
  out=$(env $envs timeout 600 python bench.py $args --no-cpu 2>/dev/null | tail -1)
  echo "$setting: $(echo "$out" | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('qps %.0f kernel_ms %.3f frac %.3f recall %.4f e2e %.0f' % (j['value'], j['roofline']['kernel_ms'], j['roofline']['frac'], j['config']['recall_at_10'], j['e2e']['value']))")"
done
