set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
O=gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531"
timeout 300 $T profiles/shard_apply_time.py 26 > $O/r02_shard2_n26_remx.txt 2>&1; tail -1 $O/r02_shard2_n26_remx.txt
BQA_NO_REMX=1 timeout 300 $T profiles/shard_apply_time.py 26 > $O/r02_shard2_n26_old.txt 2>&1; tail -1 $O/r02_shard2_n26_old.txt
timeout 300 $T profiles/shard_apply_time.py 28 > $O/r02_shard2_n28_remx.txt 2>&1; tail -1 $O/r02_shard2_n28_remx.txt
BQA_NO_REMX=1 timeout 300 $T profiles/shard_apply_time.py 28 > $O/r02_shard2_n28_old.txt 2>&1; tail -1 $O/r02_shard2_n28_old.txt
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "sharded" > $O/r02_pytest_shard.log 2>&1; tail -2 $O/r02_pytest_shard.log
