# Round-2 measurements on one GPU box: GPU tests, smoke, bench lines of every config (both arms), the ncu launch list of
# the bench command and one full capture of the trace kernel (each only after the plain command exited 0).
set -x
O=gpurun_out/r2final; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.txt 2>&1; echo "rc=$?" >> $O/pytest_gpu.txt; tail -3 $O/pytest_gpu.txt
timeout 300 python __graft_entry__.py smoke > $O/smoke.txt 2>&1; echo "rc=$?" >> $O/smoke.txt; tail -4 $O/smoke.txt
timeout 400 python bench.py --impl reference > $O/bench_reference.json 2> $O/bench_reference.err
timeout 400 python bench.py > $O/bench.json 2> $O/bench.err
for c in c1_balls c2_cube c2_cube_plain c4_tree c5_stress; do
  timeout 300 python bench.py --impl reference --config $c --steps 2 --warmup 1 > $O/ref_$c.json 2> $O/ref_$c.err
  timeout 400 python bench.py --config $c --steps 3 --warmup 3 > $O/bench_$c.json 2> $O/bench_$c.err
done
cut -c1-150 $O/*.json
timeout 300 python bench.py --steps 1 --warmup 1 > $O/bench_s1.json 2> $O/bench_s1.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file $O/launches_bench_steps1.csv python bench.py --steps 1 --warmup 1 > $O/ncu_bench.log 2>&1
timeout 300 python scripts/profile_run.py c3_teapot 128 > $O/profile_run.txt 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:trace_kernel -s 1 -c 1 -f -o $O/prof_r2_c3 python scripts/profile_run.py c3_teapot 128 > $O/ncu_full.log 2>&1
cat $O/profile_run.txt; tail -3 $O/ncu_full.log
timeout 300 python scripts/profile_run.py c5_stress 8 > $O/profile_run_c5.txt 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:trace_kernel -s 1 -c 1 -f -o $O/prof_r2_c5 python scripts/profile_run.py c5_stress 8 > $O/ncu_full_c5.log 2>&1
cat $O/profile_run_c5.txt; tail -3 $O/ncu_full_c5.log
timeout 200 python scripts/e2e_probe.py > $O/e2e_probe.txt 2>&1; cat $O/e2e_probe.txt
