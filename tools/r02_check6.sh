# new round-2 pieces on the GPU: device rebalancer, async dumps, driver; then a source-level ncu capture of the packed push
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_slab.py tests/test_gpu_driver.py tests/test_gpu_yee.py -x -q 2>&1 | tail -15 > gpurun_out/r02w_pytest.log; cat gpurun_out/r02w_pytest.log
timeout 600 python -m pytest tests/test_gpu_fullsize.py -x -q -s -k "65536" 2>&1 | tail -8 > gpurun_out/r02w_fullsize.log; cat gpurun_out/r02w_fullsize.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:push_packed -s 1500 -c 1 -o gpurun_out/r02w_push_sat -f python tools/prof_push.py 1510 0.001 0.2 > gpurun_out/ncu_w.log 2>&1
ncu -i gpurun_out/r02w_push_sat.ncu-rep --page raw --csv > gpurun_out/r02w_push_sat_raw.csv
ncu -i gpurun_out/r02w_push_sat.ncu-rep --page source --csv > gpurun_out/r02w_push_sat_src.csv
rm -f gpurun_out/*.ncu-rep.tmp
ls -la gpurun_out | tail
