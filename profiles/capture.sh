#!/bin/bash
# Evidence run for profiles/ (one B200 under gpurun):  bash profiles/capture.sh
#   1. the judged commands WITHOUT a profiler (smoke, default bench, reference arm)
#   2. the ncu launch list of the bench command
#   3. one `--set full` capture each of the assembly kernel (Stokes p4, mixed Poisson p6), of the two
#      matrix-free apply kernels (DMMA: Stokes p4, DFMA: mixed Poisson p6) and of the stored-block gemv
# Everything lands in gpurun_out/; the summaries that are judged are copied into profiles/ by hand.
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc $?" >> $O/smoke.log
timeout 600 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "rc $?" >> $O/bench_default.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err; echo "rc $?" >> $O/bench_reference.err
timeout 600 python bench.py --workload poisson6 --steps 10 --warmup 3 --no-cpu > $O/bench_poisson6.json 2> $O/bench_poisson6.err; echo "rc $?" >> $O/bench_poisson6.err
NCU="ncu --clock-control none"
timeout 900 $NCU --metrics gpu__time_duration.sum -c 600 --csv --log-file $O/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu > $O/ncu_l.log 2>&1
timeout 600 $NCU --set full --import-source on -k regex:element_matrix_kernel -s 2 -c 1 -f -o $O/asm4_full \
    python bench.py --steps 2 --warmup 3 --no-cpu --skip-e2e --skip-extra > $O/ncu_a4.log 2>&1
timeout 600 $NCU --set full --import-source on -k regex:element_matrix_kernel -s 2 -c 1 -f -o $O/asm6_full \
    python bench.py --workload poisson6 --steps 2 --warmup 3 --no-cpu --skip-e2e --skip-extra > $O/ncu_a6.log 2>&1
timeout 600 $NCU --set full --import-source on -k regex:element_vector_fused_kernel -s 2 -c 1 -f -o $O/apply6_full \
    python bench.py --workload poisson6 --steps 2 --warmup 3 --no-cpu --skip-e2e > $O/ncu_v6.log 2>&1
timeout 600 $NCU --set full --import-source on -k regex:element_vector_dmma_kernel -s 2 -c 1 -f -o $O/applyd4_full \
    python bench.py --steps 2 --warmup 3 --no-cpu --skip-e2e > $O/ncu_vd4.log 2>&1
timeout 600 $NCU --set full --import-source on -k regex:block_gemv_kernel -s 2 -c 1 -f -o $O/gemv_full \
    python bench.py --steps 2 --warmup 3 --no-cpu --skip-e2e > $O/ncu_g.log 2>&1
for f in asm4 asm6 apply6 applyd4 gemv; do
    ncu -i $O/${f}_full.ncu-rep --page raw --csv > $O/${f}_raw.csv 2>/dev/null
done
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $O/smi.txt
