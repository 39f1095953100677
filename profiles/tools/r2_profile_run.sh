set -x
F="--no-cpu-baseline --no-spline --no-cartesian --no-planner --no-stop --no-tbuf --no-cfg4"
python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_v9.json 2> gpurun_out/r2_bench_v9.err
cp gpurun_out/bench_full_n1.json gpurun_out/r2_bench_v9_full.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_v9_bench_steps1.csv python bench.py --steps 1 --warmup 3 $F > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_optimize|k_integrate|k_joint_sample|k_resample_fill' -c 5 -f -o gpurun_out/r2_full_v9 python bench.py --steps 1 --warmup 3 $F > gpurun_out/ncu_full.log 2>&1
ncu --set full --clock-control none -k regex:k_optimize_pair -s 3 -c 1 -f -o gpurun_out/r2_full_cfg4_sweeps python profiles/tools/batch_scaling.py 65536 > gpurun_out/ncu_cfg4.log 2>&1
ncu --set full --clock-control none -k regex:k_spline_eval_poly -c 1 -f -o gpurun_out/r2_full_spline_poly python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-cartesian --no-planner --no-stop --no-tbuf --no-cfg4 > gpurun_out/ncu_spline.log 2>&1
ls -la gpurun_out/*.ncu-rep
tail -3 gpurun_out/ncu_full.log gpurun_out/ncu_cfg4.log
