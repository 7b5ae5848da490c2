for d in 0 1 3; do
echo "== TSM_DIST_DEBUG=$d"
TSM_DIST_DEBUG=$d PYTHONPATH=. timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2951$d tools/launch_cost.py 2>&1 | grep "dist epoch" | tail -1
done
timeout 300 python -m pytest tests/test_gpu_dist.py -x -q 2>&1 | tail -2
