cd $GRAFT_REPO_ROOT
B="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-sparse"
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r02_n1.json 2> gpurun_out/bench_r02_n1.err; echo "bench rc=$?"
timeout 300 python tools/bench_shipped.py 24 > gpurun_out/shipped_config_r02.json 2> gpurun_out/shipped.err; echo "shipped rc=$?"
timeout 400 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/ref_arm_check.json 2> gpurun_out/ref_arm_check.err; echo "ref rc=$?"
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_ -c 200 --csv --log-file gpurun_out/launches_r02_n1.csv $B > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_scatter_wc -s 4 -c 4 -o gpurun_out/prof_r02_wc_final $B > gpurun_out/ncu_wc.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_build_probe3 -s 1 -c 1 -o gpurun_out/prof_r02_k4 $B > gpurun_out/ncu_k4.log 2>&1
tail -c 300 gpurun_out/bench_r02_n1.err; tail -c 600 gpurun_out/shipped_config_r02.json; tail -c 400 gpurun_out/ref_arm_check.json; ls -la gpurun_out | tail -8
