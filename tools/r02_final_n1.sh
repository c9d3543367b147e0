# round-2 N=1 evidence: same-box A/B of the push against the previous commit's library, full GPU suite, bench lines, launch list,
# full ncu captures of the push (saturated) and the fused Yee step
set -x
mkdir -p gpurun_out
T=${TAG:-r02z}
V=build/variants/libminipic_none.so
if [ -f $V ]; then for k in 1 2; do for r in sat quiet thermal c4; do
  timeout 300 python tools/push_ab.py $r | sed 's/^{/{"build": "this", /' >> gpurun_out/${T}_push_ab.jsonl
  MPIC_LIB=$V timeout 300 python tools/push_ab.py $r | sed 's/^{/{"build": "previous commit", /' >> gpurun_out/${T}_push_ab.jsonl
done; done; cat gpurun_out/${T}_push_ab.jsonl; fi
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/${T}_pytest_gpu.log; cat gpurun_out/${T}_pytest_gpu.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/${T}_bench_n1.json 2> gpurun_out/${T}_bench.err; tail -c 600 gpurun_out/${T}_bench_n1.json; tail -3 gpurun_out/${T}_bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_bench_reference_arm.json 2>&1; tail -c 600 gpurun_out/${T}_bench_reference_arm.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 11900 -c 96 --csv --log-file gpurun_out/${T}_launches_two_stream_512.csv python tools/prof_push.py 1510 0.001 0.2 > gpurun_out/ncu_l.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:push_packed -s 1500 -c 1 -o gpurun_out/${T}_push_sat -f python tools/prof_push.py 1510 0.001 0.2 > gpurun_out/ncu_p.log 2>&1
ncu -i gpurun_out/${T}_push_sat.ncu-rep --page raw --csv > gpurun_out/${T}_ncu_full_push_packed_saturated_raw.csv
timeout 900 ncu --set full --clock-control none -k regex:fused_yee -s 5 -c 1 -o gpurun_out/${T}_yee -f python tools/prof_yee.py 8 > gpurun_out/ncu_y.log 2>&1
ncu -i gpurun_out/${T}_yee.ncu-rep --page raw --csv > gpurun_out/${T}_ncu_full_fused_yee_8192_raw.csv
rm -f gpurun_out/*.ncu-rep
ls -la gpurun_out | tail -12
