#!/bin/bash
# developer tool: decode us/token of a truncated model under environment-knob combinations
# usage: tools/envsweep.sh LAYERS "VAR=1 VAR2=3" "VAR=2" ...
L=$1; shift
for combo in "$@"; do
  out=$(env $combo python bench.py --layers $L --steps 2 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1)
  echo "$combo :: $(echo "$out" | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('decode us/token %.1f  prefill tok/s %.0f' % (1e6/d['decode_tokens_per_s'], d['prefill_tokens_per_s']))")"
done
