# Round-1 (session 20) ncu captures after the Cholesky protocol fix (consumed flag, back-substitution wait) (run under gpurun; one GPU).  Outputs: gpurun_out/r2o/
#   launches.csv     : every launch of `bench.py --steps 2 --warmup 3 --no-cpu --no-large` with its device time
#   prof_all.ncu-rep : --set full of one launch of every kernel of the library (prof_driver all 1 1)
mkdir -p gpurun_out/r2o
python bench.py --steps 2 --warmup 3 --no-cpu --no-large > gpurun_out/r2o/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r2o/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-large > gpurun_out/r2o/ncu_bench.log 2>&1
python profiles/prof_driver.py all 1 1 > gpurun_out/r2o/plain_prof.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'refine|iter_proj|prep_rays|cast_f16|gn_|chol_|track_|pose_retr|edge_build|pointmap|render_' -c 30 -o gpurun_out/r2o/prof_all python profiles/prof_driver.py all 1 1 > gpurun_out/r2o/ncu_prof.log 2>&1
ls -la gpurun_out/r2o; tail -3 gpurun_out/r2o/plain_prof.log; tail -3 gpurun_out/r2o/ncu_prof.log
