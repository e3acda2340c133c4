#!/bin/bash
# configs[4] evidence on N GPUs of one box:  bash profiles/cap_config4.sh N ROWS [variants]   (strong scaling: 1000 x ROWS mesh over N strips)
# variants: list of <single_reduce><graph> pairs, e.g. "01 11 00 10"
N=${1:-2}; ROWS=${2:-250}; VARS=${3:-"01 11"}; O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
port=29520
for v in $VARS; do
  sr=${v:0:1}; gr=${v:1:1}; port=$((port + 1))
  MFV2D_CG_SINGLE_REDUCE=$sr MFV2D_CG_GRAPH=$gr timeout 600 $TR --master-port $port profiles/config4_solve.py --gpus $N --rows $ROWS --rtol 1e-11 \
      > $O/config4_${N}gpu_rows${ROWS}_sr${sr}_graph${gr}.json 2> $O/config4_${N}gpu_rows${ROWS}_sr${sr}_graph${gr}.err; echo "sr$sr graph$gr rc $?"
  python -c "import json,sys; d=json.load(open('$O/config4_${N}gpu_rows${ROWS}_sr${sr}_graph${gr}.json')); print('$v', d['solve'])"
done
