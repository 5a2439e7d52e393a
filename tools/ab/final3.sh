set -x
cd /root/repo
NCCL_DEBUG=VERSION timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_n2b.json 2> gpurun_out/bench_n2b.err
wc -l gpurun_out/bench_n2b.json; head -c 300 gpurun_out/bench_n2b.json; echo
timeout 300 python bench.py > gpurun_out/bench_n1b.json 2> gpurun_out/bench_n1b.err
wc -l gpurun_out/bench_n1b.json; head -c 300 gpurun_out/bench_n1b.json; echo
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
