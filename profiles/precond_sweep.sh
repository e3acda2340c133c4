#!/bin/bash
# iteration count / time of the multilevel preconditioner against its aggregate sizes (one rank's share of configs[4] on one GPU)
O=gpurun_out; mkdir -p $O
for cfg in "20:5" "16:4" "16:4,8" "20:2,5" "25:5" "12:4" "10:5" "20:4,10" "16:2,4,8"; do
  agg=${cfg%%:*}; mid=${cfg##*:}
  timeout 200 python profiles/config4_solve.py --gpus 1 --rows 125 --precondition multilevel --agg $agg --mid $mid --rtol 1e-11 --repeat 2 > $O/sweep_${agg}_${mid//,/-}.json 2>/dev/null
  python - <<PY
import json
try:
    d = json.loads(open("$O/sweep_${agg}_${mid//,/-}.json").read().strip().splitlines()[-1]); s = d["solve"]
    print("agg $agg mid $mid", "n_coarse", d["n_coarse"], "it", s["iterations"], "s", round(s["s"], 4), "us/it", round(s["us_per_iteration"], 1), "coarse setup", round(d["coarse_setup_s"], 3), "res", s["saddle_relative_residual"])
except Exception as e:
    print("agg $agg mid $mid failed", e)
PY
done
