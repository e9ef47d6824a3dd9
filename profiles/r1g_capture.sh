# ncu captures of round 1 (run under gpurun; one GPU).  Outputs: gpurun_out/r1g/
#   launches.csv : every launch of `bench.py --steps 2 --warmup 3 --no-cpu` with its device time
#   prof_all.ncu-rep : --set full of one launch of every kernel of the library (prof_driver all 1 1)
mkdir -p gpurun_out/r1g
if [ "$1" != "full" ]; then
python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r1g/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r1g/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r1g/ncu_bench.log 2>&1
fi
python profiles/prof_driver.py all 1 1 > gpurun_out/r1g/plain_prof.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'refine|iter_proj|prep_rays|cast_f16|gn_|chol_|track_|pose_retr' -c 12 -o gpurun_out/r1g/prof_all python profiles/prof_driver.py all 1 1 > gpurun_out/r1g/ncu_prof.log 2>&1
ls -la gpurun_out/r1g; tail -3 gpurun_out/r1g/ncu_prof.log
