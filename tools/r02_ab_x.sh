OUT=gpurun_out/$1
timeout 900 python -m pytest tests/test_gpu_pic.py tests/test_gpu_driver.py -x -q 2>&1 | tail -2
for r in quiet sat thermal c4; do timeout 300 python tools/push_ab.py $r >> $OUT; done
MPIC_NO_SORT_OVERLAP=1 timeout 300 python tools/push_ab.py sat | sed 's|^{|{"overlap": "off", |' >> $OUT
cat $OUT
SKIP=13500 bash tools/r02_launches.sh gpurun_out/r02x_launches_sat.csv 1510 0.001 0.2
