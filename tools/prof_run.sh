# full ncu capture of one steady-state push (thermal plasma, step 25) + raw/source CSV exports
set -x
timeout 300 python tools/prof_push.py 30 0.05 0.0 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"push_deposit" -s 25 -c 1 -o gpurun_out/push_thermal -f python tools/prof_push.py 30 0.05 0.0 > gpurun_out/ncu_t.log 2>&1
ncu -i gpurun_out/push_thermal.ncu-rep --page raw --csv > gpurun_out/push_thermal_raw.csv 2>/dev/null
ncu -i gpurun_out/push_thermal.ncu-rep --page source --csv > gpurun_out/push_thermal_src.csv 2>/dev/null
ls -la gpurun_out | tail -5
