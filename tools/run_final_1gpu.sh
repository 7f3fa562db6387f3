# Round-2 single-GPU sequence: GPU test tier, the bench line with every extra, launch list and ncu captures.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 600 python bench.py --steps 200 --warmup 10 > gpurun_out/final_bench_1gpu.json 2> gpurun_out/final_bench_1gpu.err; echo "bench rc=$?"
CMD="python bench.py --steps 3 --warmup 3 --no-extra --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_mlp_r2.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"fused_reduce_adamw|cast_zero" -s 14 -c 4 -o gpurun_out/prof_update_r2 $CMD > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out | tail -8
