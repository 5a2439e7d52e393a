set -x
cd /root/repo
python -m pytest tests/test_gpu_loss.py -m gpu -x -q -k "stripes" 2>&1 | tail -15 > gpurun_out/stripes.log
for w in 2 4 8; do
  for rep in 1 2; do
  python tools/stripe_step.py $w 32768 8 >> gpurun_out/stripe_ab.log 2>&1
  BSC_LIB_PATH=/root/repo/tools/ab/libbsc_hot_head.so python tools/stripe_step.py $w 32768 8 2>&1 | sed 's/^/HEAD /' >> gpurun_out/stripe_ab.log
  done
  BSC_BWD_GSTORE=0 python tools/stripe_step.py $w 32768 8 >> gpurun_out/stripe_ab.log 2>&1
done
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
tail -3 gpurun_out/stripes.log; cat gpurun_out/stripe_ab.log; tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/bench_n1.json
