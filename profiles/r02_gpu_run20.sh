set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
O=gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531"
timeout 200 $T profiles/shard_apply_time.py 30 > $O/r02_shard8_n30_bal_remx.txt 2>&1; tail -1 $O/r02_shard8_n30_bal_remx.txt
BQA_TILED_UNBALANCED=1 timeout 200 $T profiles/shard_apply_time.py 30 > $O/r02_shard8_n30_unbal_remx.txt 2>&1; tail -1 $O/r02_shard8_n30_unbal_remx.txt
BQA_TILED_UNBALANCED=1 BQA_NO_REMX=1 timeout 200 $T profiles/shard_apply_time.py 30 > $O/r02_shard8_n30_old.txt 2>&1; tail -1 $O/r02_shard8_n30_old.txt
timeout 200 $T profiles/shard_apply_time.py 27 > $O/r02_shard8_n27_bal_remx.txt 2>&1; tail -1 $O/r02_shard8_n27_bal_remx.txt
