set -x
mkdir -p gpurun_out/v1
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -5
timeout 400 python bench.py --impl reference > gpurun_out/v1/bench_reference.json 2> gpurun_out/v1/bench_reference.err
timeout 400 python bench.py > gpurun_out/v1/bench.json 2> gpurun_out/v1/bench.err
for c in c1_balls c2_cube c2_cube_plain c4_tree c5_stress; do
  timeout 300 python bench.py --impl reference --config $c --steps 2 --warmup 1 > gpurun_out/v1/ref_$c.json 2> gpurun_out/v1/ref_$c.err
  timeout 400 python bench.py --config $c --steps 3 --warmup 3 > gpurun_out/v1/bench_$c.json 2> gpurun_out/v1/bench_$c.err
done
cat gpurun_out/v1/*.json | cut -c1-200
