# Round-end measurements on one GPU box: GPU tests, smoke, bench lines of every config (both arms), then the ncu
# launch list of the bench command and one full capture of the trace kernel (each after the plain command exited 0).
set -x
mkdir -p gpurun_out/v2
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/v2/pytest_gpu.txt 2>&1; echo "rc=$?" >> gpurun_out/v2/pytest_gpu.txt; tail -3 gpurun_out/v2/pytest_gpu.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/v2/smoke.txt 2>&1; echo "rc=$?" >> gpurun_out/v2/smoke.txt; tail -4 gpurun_out/v2/smoke.txt
timeout 400 python bench.py --impl reference > gpurun_out/v2/bench_reference.json 2> gpurun_out/v2/bench_reference.err
timeout 400 python bench.py > gpurun_out/v2/bench.json 2> gpurun_out/v2/bench.err
for c in c1_balls c2_cube c2_cube_plain c4_tree c5_stress; do
  timeout 300 python bench.py --impl reference --config $c --steps 2 --warmup 1 > gpurun_out/v2/ref_$c.json 2> gpurun_out/v2/ref_$c.err
  timeout 400 python bench.py --config $c --steps 3 --warmup 3 > gpurun_out/v2/bench_$c.json 2> gpurun_out/v2/bench_$c.err
done
cut -c1-150 gpurun_out/v2/*.json
timeout 300 python bench.py --steps 1 --warmup 1 > gpurun_out/v2/bench_s1.json 2> gpurun_out/v2/bench_s1.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/v2/launches_bench_steps1.csv python bench.py --steps 1 --warmup 1 > gpurun_out/v2/ncu_bench.log 2>&1
timeout 300 python scripts/profile_run.py c3_teapot 128 > gpurun_out/v2/profile_run.txt 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:trace_kernel -s 1 -c 1 -f -o gpurun_out/v2/prof_r1_final5 python scripts/profile_run.py c3_teapot 128 > gpurun_out/v2/ncu_full.log 2>&1
cat gpurun_out/v2/profile_run.txt; tail -3 gpurun_out/v2/ncu_full.log
