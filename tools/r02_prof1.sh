# round 2, GPU call 1: where does the push kernel lose time when particles cross cells?
# (a) thermal u_th=0.05 512^2 step 25, (b) two-stream after 1500 steps (saturated); ncu --set full with source.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
timeout 300 python tools/prof_push.py 30 0.05 0.0 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"push_deposit" -s 25 -c 1 -o gpurun_out/r02a_push_thermal -f python tools/prof_push.py 30 0.05 0.0 > gpurun_out/ncu_t.log 2>&1
ncu -i gpurun_out/r02a_push_thermal.ncu-rep --page raw --csv > gpurun_out/r02a_push_thermal_raw.csv 2>/dev/null
ncu -i gpurun_out/r02a_push_thermal.ncu-rep --page source --csv > gpurun_out/r02a_push_thermal_src.csv 2>/dev/null
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"push_deposit" -s 1500 -c 1 -o gpurun_out/r02a_push_sat -f python tools/prof_push.py 1510 0.001 0.2 > gpurun_out/ncu_s.log 2>&1
ncu -i gpurun_out/r02a_push_sat.ncu-rep --page raw --csv > gpurun_out/r02a_push_sat_raw.csv 2>/dev/null
ncu -i gpurun_out/r02a_push_sat.ncu-rep --page source --csv > gpurun_out/r02a_push_sat_src.csv 2>/dev/null
# launch list of the saturated regime (per-kernel durations)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 13500 -c 60 --csv --log-file gpurun_out/r02a_launches_sat.csv python tools/prof_push.py 1510 0.001 0.2 > gpurun_out/ncu_l.log 2>&1
rm -f gpurun_out/*.ncu-rep.tmp
ls -la gpurun_out | tail -12
