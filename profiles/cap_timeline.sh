#!/bin/bash
# Timeline of the distributed CG on N GPUs (MFV2D_PEER_DEBUG bit 8):  bash profiles/cap_timeline.sh N ROWS "<precond>:<debug bits> ..."
N=${1:-8}; ROWS=${2:-1000}; VARS=${3:-"multilevel:8 multilevel:10 jacobi:8"}; O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
port=29620
for v in $VARS; do
  pc=${v%%:*}; dbg=${v##*:}; port=$((port + 1))
  MFV2D_PEER_DEBUG=$dbg timeout 400 $TR --master-port $port profiles/config4_solve.py --gpus $N --rows $ROWS --rtol 1e-11 --precondition $pc --agg 20 --mid 5 \
      > $O/timeline_${N}gpu_${pc}_dbg${dbg}.json 2> $O/timeline_${N}gpu_${pc}_dbg${dbg}.err; echo "$v rc $?"
  python - <<PY
import json
try:
    d = json.loads(open('$O/timeline_${N}gpu_${pc}_dbg${dbg}.json').read().strip().splitlines()[-1])
    s = d['solve']; print('$v', s['s'], s['iterations'], round(s['us_per_iteration'], 1), s['saddle_relative_residual'])
    for row in (s.get('timeline') or {}).get('us_per_iteration_by_stamp', []): print('   ', row)
except Exception as e:
    print('$v', 'failed', e)
PY
done
