# Round-2 multi-GPU sequence (N GPUs, default 8): correctness checks, then the MLP / GPT-2 numbers.
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
mkdir -p gpurun_out
timeout 300 $TR --master-port 29541 tools/gpu_check_multi.py > gpurun_out/r2_multi${N}.log 2>&1; grep -E "ok\]|FAIL|PASS|mismatch|rel err|resume|gave up" gpurun_out/r2_multi${N}.log | cut -c1-230
B200_SKIP_TIMING=1 timeout 400 $TR --master-port 29542 tools/gpu_check_multi_gpt.py > gpurun_out/r2_multi${N}_gpt.log 2>&1; grep -E "ok\]|FAIL|PASS|gave up|rror" gpurun_out/r2_multi${N}_gpt.log | cut -c1-230
timeout 200 $TR --master-port 29543 tools/comm_phase_timeline.py --out gpurun_out/r2_comm_phases.md 2>&1 | tail -1
timeout 200 $TR --master-port 29544 tools/comm_phase_timeline.py --config mlp_fsdp --out gpurun_out/r2_comm_phases.md 2>&1 | tail -1
timeout 500 $TR --master-port 29545 bench.py --gpus $N --steps 300 --warmup 10 2>&1 | tail -1 > gpurun_out/r2_bench_mlp_dp_${N}.json; python - <<PY
import json
d=json.load(open("gpurun_out/r2_bench_mlp_dp_${N}.json"))
print("mlp_dp", d["ms_per_step"], d["value"], "e2e", d.get("e2e",{}).get("value"), "gpt2s_dp", {k:d["extra"]["gpt2s_dp"].get(k) for k in ("ms_per_step","value","model_tflops_per_gpu","error")})
PY
for MODE in 1; do echo -n "mlp_dp publish mode $MODE: "; B200DIST_DEFER_PUBLISH=$MODE timeout 300 $TR --master-port 29546 bench.py --gpus $N --steps 300 --warmup 10 --no-extra --no-e2e 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'])"; done
echo -n "mlp_fsdp: "; timeout 300 $TR --master-port 29547 bench.py --gpus $N --steps 300 --warmup 10 --no-e2e --config mlp_fsdp 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'])"
for CB in 32 64; do echo -n "gpt2s_dp comm_blocks=$CB: "; B200DIST_COMM_BLOCKS=$CB timeout 300 $TR --master-port 29548 bench.py --gpus $N --steps 20 --warmup 4 --no-e2e --config gpt2s_dp 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'])"; done
echo -n "gpt2s_fsdp: "; timeout 300 $TR --master-port 29549 bench.py --gpus $N --steps 20 --warmup 4 --no-e2e --config gpt2s_fsdp 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'])"
