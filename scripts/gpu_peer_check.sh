# GPU-box check of the in-process multi-GPU path: bash scripts/gpu_peer_check.sh N   (N = visible GPUs)
N=${1:-2}
mkdir -p gpurun_out/peer
nvidia-smi -L
timeout 600 python -m pytest tests/test_gpu_combine.py -x -q > gpurun_out/peer/pytest_x$N.txt 2>&1; tail -3 gpurun_out/peer/pytest_x$N.txt
timeout 300 python scripts/peer_bench.py --gpus $N --shard sample > gpurun_out/peer/c3_x${N}_sample.json 2> gpurun_out/peer/c3_x${N}_sample.err; cat gpurun_out/peer/c3_x${N}_sample.json; tail -3 gpurun_out/peer/c3_x${N}_sample.err
timeout 300 python scripts/peer_bench.py --gpus $N --shard tile --check --steps 2 --warmup 1 > gpurun_out/peer/c3_x${N}_tile.json 2> gpurun_out/peer/c3_x${N}_tile.err; cat gpurun_out/peer/c3_x${N}_tile.json; tail -3 gpurun_out/peer/c3_x${N}_tile.err
timeout 300 python scripts/peer_bench.py --gpus $N --config c4_tree --shard tile+sample --steps 2 --warmup 1 > gpurun_out/peer/c4_x${N}_tile_sample.json 2> gpurun_out/peer/c4_x${N}_tile_sample.err; cat gpurun_out/peer/c4_x${N}_tile_sample.json; tail -3 gpurun_out/peer/c4_x${N}_tile_sample.err
