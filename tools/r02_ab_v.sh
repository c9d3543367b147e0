OUT=gpurun_out/$1
timeout 900 python -m pytest tests/test_gpu_pic.py tests/test_gpu_slab.py -x -q 2>&1 | tail -2
for r in quiet sat thermal c4; do timeout 300 python tools/push_ab.py $r >> $OUT; done
MPIC_PERM_WALK=group timeout 300 python tools/push_ab.py sat | sed 's|^{|{"perm": "group", |' >> $OUT
cat $OUT
