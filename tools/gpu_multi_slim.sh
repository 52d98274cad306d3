#!/bin/bash
# usage: tools/gpu_multi_slim.sh N scene...   : one bench line per scene on N GPUs (default settings)
N=$1; shift
mkdir -p gpurun_out
for s in "$@"; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
    bench.py --gpus $N --scene $s --steps 50 --warmup 5 > gpurun_out/multi${N}_$s.json 2> gpurun_out/multi${N}_$s.err
  python -c "import sys,json; j=json.loads(open('gpurun_out/multi${N}_$s.json').read().strip().splitlines()[-1]); p=j['parity']; print('N=$N', j['config']['workload'], round(j['value']), round(j['ms_per_step'],4), p['bit_identical'], p['mask_mismatch'], p['depth_outliers'])"
done
