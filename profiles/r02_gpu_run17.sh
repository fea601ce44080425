set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
O=gpurun_out
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 8 --steps 2 --warmup 3 > $O/r02_bench_n8.json 2> $O/r02_bench_n8.err
tail -c 2500 $O/r02_bench_n8.json
tail -5 $O/r02_bench_n8.err
