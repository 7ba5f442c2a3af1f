#!/bin/bash
# strong-scaling sweep of bench.py on one box: N = 1 2 4 8 (whatever fits $1), one JSON line each into gpurun_out/scale_N.json
MAXN=${1:-8}
for N in 1 2 4 8; do
  if [ $N -gt $MAXN ]; then break; fi
  if [ $N -eq 1 ]; then
    python bench.py --gpus 1 --no-extras --no-cpu > gpurun_out/scale_$N.json 2> gpurun_out/scale_$N.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + N)) bench.py --gpus $N > gpurun_out/scale_$N.json 2> gpurun_out/scale_$N.err
  fi
  echo "N=$N rc=$?"
  tail -2 gpurun_out/scale_$N.err
done
