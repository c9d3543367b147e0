# perm_warp with grouped dirs loads + peer-step sort on the side stream: parity, then timing
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_pic.py tests/test_gpu_slab.py -x -q 2>&1 | tail -4 > gpurun_out/r02aa_pytest.log; cat gpurun_out/r02aa_pytest.log
V=build/variants/libminipic_base.so
for r in sat c4; do
  timeout 300 python tools/push_ab.py $r | sed 's/^{/{"build": "this", /' >> gpurun_out/r02aa_ab.jsonl
  MPIC_LIB=$V timeout 300 python tools/push_ab.py $r | sed 's/^{/{"build": "previous commit", /' >> gpurun_out/r02aa_ab.jsonl
done
cat gpurun_out/r02aa_ab.jsonl
timeout 600 python bench.py --no-yee --no-cpu --no-e2e > gpurun_out/r02aa_bench_n1.json 2> gpurun_out/r02aa_bench.err; tail -c 300 gpurun_out/r02aa_bench_n1.json; tail -3 gpurun_out/r02aa_bench.err
