set -x
timeout 900 python -m pytest tests/test_gpu_pic.py -x -q 2>&1 | tail -2
for r in quiet sat thermal c4; do timeout 300 python tools/push_ab.py $r >> gpurun_out/r02u_ab.jsonl; done
cat gpurun_out/r02u_ab.jsonl
timeout 900 ncu --set full --clock-control none --import-source on -k regex:perm_stable -s 1500 -c 1 -o gpurun_out/r02u_perm_sat -f python tools/prof_push.py 1510 0.001 0.2 > gpurun_out/ncu_u.log 2>&1
ncu -i gpurun_out/r02u_perm_sat.ncu-rep --page raw --csv > gpurun_out/r02u_perm_sat_raw.csv
ncu -i gpurun_out/r02u_perm_sat.ncu-rep --page source --csv > gpurun_out/r02u_perm_sat_src.csv
rm -f gpurun_out/*.ncu-rep.tmp
