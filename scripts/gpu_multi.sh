# Multi-GPU bench lines as the driver launches them: bash scripts/gpu_multi.sh N   (under gpurun --gpus N)
N=${1:-2}
mkdir -p gpurun_out/multi
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 200)) bench.py --gpus $N "$@"; }
run --steps 5 --warmup 3 > gpurun_out/multi/bench_${N}gpu.json 2> gpurun_out/multi/bench_${N}gpu.err; cut -c1-200 gpurun_out/multi/bench_${N}gpu.json
if [ "$N" = 8 ]; then
  run --steps 3 --warmup 3 --config c4_tree --shard tile+sample > gpurun_out/multi/bench_c4_tree_8gpu_tile_sample.json 2> gpurun_out/multi/bench_c4_8gpu.err; cut -c1-200 gpurun_out/multi/bench_c4_tree_8gpu_tile_sample.json
fi
if [ -n "$PEER" ]; then bash scripts/gpu_peer_check.sh $N 2>&1 | grep -v "^GPU "; fi
