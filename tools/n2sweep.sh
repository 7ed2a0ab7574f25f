#!/bin/bash
# developer tool: N=2 bench under environment combinations (decode tok/s, prefill tok/s)
port=29600
for combo in "$@"; do
  port=$((port+1))
  out=$(env $combo timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $port bench.py --gpus 2 --layers 8 --steps 2 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1)
  echo "$combo :: $(echo "$out" | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('decode tok/s %.0f  prefill tok/s %.0f value %.0f' % (d['decode_tokens_per_s'], d['prefill_tokens_per_s'], d['value']))")"
done
