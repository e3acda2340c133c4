#!/bin/bash
# Round-2 evidence run (one B200 under gpurun):  bash profiles/capture_r2.sh
#   1. the judged commands WITHOUT a profiler: smoke, pytest -m gpu, default bench, reference arm
#   2. the ncu launch list of the bench command
#   3. one `--set full` capture each of the round-2 kernels: packed condensed-block product, CG gather / update, fused apply
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke_r02_final.log 2>&1; echo "smoke rc $?" >> $O/smoke_r02_final.log
(timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -6) > $O/pytest_gpu_r02_final.log
timeout 900 python bench.py > $O/bench_r02_final.json 2> $O/bench_r02_final.err; echo "rc $?" >> $O/bench_r02_final.err
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_r02_final_reference.json 2> $O/bench_r02_final_reference.err; echo "rc $?" >> $O/bench_r02_final_reference.err
NCU="ncu --clock-control none"
timeout 900 $NCU --metrics gpu__time_duration.sum -c 900 --csv --log-file $O/launches_r02_final.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu > $O/ncu_l.log 2>&1
for k in schur_local_packed_kernel cg1_gather_kernel cg1_update_kernel element_vector_fused_kernel; do
  timeout 600 $NCU --set full --import-source on -k regex:$k -s 20 -c 1 -f -o $O/${k}_full \
      python bench.py --steps 2 --warmup 3 --no-cpu --skip-e2e > $O/ncu_$k.log 2>&1
  ncu -i $O/${k}_full.ncu-rep --page raw --csv > $O/ncu_full_${k}_r02_final.csv 2>/dev/null
  rm -f $O/${k}_full.ncu-rep
done
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $O/smi_r02.txt
tail -2 $O/smoke_r02_final.log; cat $O/pytest_gpu_r02_final.log | tail -2; tail -1 $O/bench_r02_final.err; tail -1 $O/bench_r02_final_reference.err
python - <<'PY'
import json
for f in ("bench_r02_final.json", "bench_r02_final_reference.json"):
    try:
        d = json.loads(open("gpurun_out/" + f).read().strip().splitlines()[-1])
        print(f, "value", d.get("value"), "e2e", (d.get("e2e") or {}).get("value"), "frac", (d.get("roofline") or {}).get("frac"),
              "apply", (d.get("apply") or {}).get("value"), ((d.get("apply") or {}).get("roofline") or {}).get("frac"),
              "stored", ((d.get("apply_stored") or {}).get("roofline") or {}).get("frac"),
              "solve", {k: (d.get("solve") or {}).get(k) for k in ("s", "iterations", "us_per_iteration", "saddle_relative_residual", "system_setup_s", "pipeline_s")},
              "vs_cpu", d.get("vs_cpu_baseline"))
    except Exception as e:
        print(f, "failed", e)
PY
