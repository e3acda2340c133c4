#!/bin/bash
# configs[2] (512 x 512 order-5 advection-diffusion time march), strong scaling over N GPUs:  bash profiles/cap_config2.sh "8 4 2"
O=gpurun_out; mkdir -p $O; port=29720
for N in ${1:-"8 4 2"}; do
  port=$((port + 1))
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port profiles/config2_march.py --gpus $N --steps 3 --precondition ${PC:-multilevel} --agg ${AGG:-8} --mid ${MID:-2,4} ${DEV:+--device} \
      > $O/config2_512_${N}gpu${DEV:+_device}_r02.json 2> $O/config2_512_${N}gpu.err; echo "N=$N rc $?"
  python - <<PY
import json
try:
    d = json.loads(open('$O/config2_512_${N}gpu${DEV:+_device}_r02.json').read().strip().splitlines()[-1])
    print('N=$N', 'setup', round(d['system_setup_s'], 3), 'steps', [round(t, 3) for t in d['step_s']], 'krylov', d['krylov_iterations'], 'saddle', d['last_correction_saddle_relative_residual'], 'norms', d['solution_norms'])
except Exception as e:
    print('N=$N failed', e); print(open('$O/config2_512_${N}gpu.err').read()[-1500:])
PY
done
