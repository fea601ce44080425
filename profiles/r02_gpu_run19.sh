set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
O=gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531"
timeout 300 $T profiles/shard_check.py 16 0.2 > $O/r02_shard_check16.txt 2> $O/r02_shard_check16.err; tail -3 $O/r02_shard_check16.txt; grep -v "^\*\|OMP_NUM\|^$" $O/r02_shard_check16.err | tail -30
timeout 300 $T profiles/shard_apply_time.py 28 > $O/r02_shard2_n28_bal.txt 2>&1; tail -1 $O/r02_shard2_n28_bal.txt
BQA_TILED_UNBALANCED=1 timeout 300 $T profiles/shard_apply_time.py 28 > $O/r02_shard2_n28_unbal.txt 2>&1; tail -1 $O/r02_shard2_n28_unbal.txt
timeout 300 $T profiles/shard_apply_time.py 26 > $O/r02_shard2_n26_bal.txt 2>&1; tail -1 $O/r02_shard2_n26_bal.txt
