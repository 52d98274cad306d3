#!/bin/bash
# usage: tools/gpu_multi_ab.sh N : peer-store A/B (TMA-staged 128-byte rows vs plain 32-byte tile rows) on N GPUs
N=$1
for st in 1 0; do for s in grid65_1080p grid65_4k bunny_4k; do
  RT_STAGE_STORES=$st python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --gpus $N --scene $s --steps 50 --warmup 5 2>/dev/null | python -c "import sys,json; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('N=$N stage=$st', j['config']['workload'], round(j['value']), round(j['ms_per_step'],4), j['parity']['bit_identical'])"
done; done
