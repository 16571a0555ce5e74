set -x
J="--steps 3 --warmup 3 --no-cpu --no-e2e --no-parity"
python bench.py $J > gpurun_out/plain_join.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_ncu_launches_join.csv python bench.py $J > gpurun_out/ncu_l.log 2>&1
S="--steps 1 --warmup 1 --no-cpu --no-e2e --no-parity"
python bench.py $S > gpurun_out/plain_join1.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"fused8_radix_pass_kernel|fused8_join_kernel|fused8_block_hist" -s 5 -c 5 -o gpurun_out/r02_prof_join python bench.py $S > gpurun_out/ncu_j.log 2>&1
python bench.py --config sortjoin $S > gpurun_out/plain_sj.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"fused8x_join_kernel" -s 1 -c 1 -o gpurun_out/r02_prof_sortjoin python bench.py --config sortjoin $S > gpurun_out/ncu_sj.log 2>&1
python bench.py --config groupby $S > gpurun_out/plain_gb.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"groupby_fused8_kernel" -s 1 -c 1 -o gpurun_out/r02_prof_groupby python bench.py --config groupby $S > gpurun_out/ncu_gbf.log 2>&1
ls -la gpurun_out/*.ncu-rep
