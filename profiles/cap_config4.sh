#!/bin/bash
# configs[4] evidence on N GPUs of one box:  bash profiles/cap_config4.sh N ROWS   (strong scaling: 1000 x ROWS mesh over N strips)
N=${1:-2}; ROWS=${2:-250}; O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for v in 0 1; do
  MFV2D_CG_SINGLE_REDUCE=$v timeout 600 $TR --master-port $((29520 + v)) profiles/config4_solve.py --gpus $N --rows $ROWS --rtol 1e-11 \
      > $O/config4_${N}gpu_rows${ROWS}_sr$v.json 2> $O/config4_${N}gpu_rows${ROWS}_sr$v.err; echo "sr$v rc $?"
done
MFV2D_CG_SINGLE_REDUCE=1 timeout 600 $TR --master-port 29523 profiles/config4_solve.py --gpus $N --rows $ROWS --rtol 1e-11 --precondition \
      > $O/config4_${N}gpu_rows${ROWS}_sr1_jacobi.json 2> $O/config4_${N}gpu_rows${ROWS}_sr1_jacobi.err; echo "sr1 jacobi rc $?"
cat $O/config4_${N}gpu_rows${ROWS}_sr*.json
