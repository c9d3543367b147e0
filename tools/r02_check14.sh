mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:perm_warp -s 1495 -c 1 -o gpurun_out/r02ao_perm -f python tools/prof_push.py 1505 0.001 0.2 > gpurun_out/ncu_q.log 2>&1
ncu -i gpurun_out/r02ao_perm.ncu-rep --page raw --csv > gpurun_out/r02ao_ncu_full_perm_warp_saturated_raw.csv
ncu -i gpurun_out/r02ao_perm.ncu-rep --page source --csv > gpurun_out/r02ao_perm_src.csv
rm -f gpurun_out/*.ncu-rep
tail -3 gpurun_out/ncu_q.log
