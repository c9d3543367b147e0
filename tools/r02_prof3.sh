set -x
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:push_ -s 1500 -c 1 -o gpurun_out/r02t_push_sat -f python tools/prof_push.py 1510 0.001 0.2 > gpurun_out/ncu_t2.log 2>&1
ncu -i gpurun_out/r02t_push_sat.ncu-rep --page raw --csv > gpurun_out/r02t_push_sat_raw.csv
ncu -i gpurun_out/r02t_push_sat.ncu-rep --page source --csv > gpurun_out/r02t_push_sat_src.csv
rm -f gpurun_out/*.ncu-rep.tmp
