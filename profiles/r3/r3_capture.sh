# Round-2 ncu captures (run under gpurun; one GPU).  Outputs: gpurun_out/r3cap/
#   launches.csv     : every launch of `bench.py --steps 2 --warmup 3 --no-cpu` with its device time
#   prof_all.ncu-rep : --set full of one launch of every kernel of the library (prof_driver all 1 1)
#   chol64.ncu-rep   : --set full of the macro-tile Cholesky at N = 512 (prof_chol64.py)
# Each profiled command is first run without the profiler.
mkdir -p gpurun_out/r3cap
python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r3cap/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r3cap/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r3cap/ncu_bench.log 2>&1
python profiles/prof_driver.py all 1 1 > gpurun_out/r3cap/plain_prof.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'refine|iter_proj|prep_rays|match_prep|cast_f16|gn_|chol|track_|pose_retr|edge_|pointmap|render_' -c 40 -o gpurun_out/r3cap/prof_all python profiles/prof_driver.py all 1 1 > gpurun_out/r3cap/ncu_prof.log 2>&1
python profiles/prof_chol64.py > gpurun_out/r3cap/plain_chol64.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:chol64_kernel -c 1 -o gpurun_out/r3cap/chol64 python profiles/prof_chol64.py > gpurun_out/r3cap/ncu_chol64.log 2>&1
ls -la gpurun_out/r3cap; tail -2 gpurun_out/r3cap/plain_prof.log; tail -2 gpurun_out/r3cap/ncu_prof.log; tail -2 gpurun_out/r3cap/ncu_chol64.log
