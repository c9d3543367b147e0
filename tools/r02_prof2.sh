set -x
mkdir -p gpurun_out
for gl in 1 32; do
MPIC_PUSH_GL=$gl timeout 900 ncu --set full --clock-control none --import-source on -k regex:push_ -s 1500 -c 1 -o gpurun_out/r02k_push_sat_gl$gl -f python tools/prof_push.py 1510 0.001 0.2 > gpurun_out/ncu_k$gl.log 2>&1
ncu -i gpurun_out/r02k_push_sat_gl$gl.ncu-rep --page raw --csv > gpurun_out/r02k_push_sat_gl${gl}_raw.csv
ncu -i gpurun_out/r02k_push_sat_gl$gl.ncu-rep --page source --csv > gpurun_out/r02k_push_sat_gl${gl}_src.csv
done
rm -f gpurun_out/*.ncu-rep.tmp
ls -la gpurun_out | tail -8
