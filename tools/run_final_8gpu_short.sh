set -x
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 $TR --nproc-per-node 8 --master-port 29541 tools/gpu_check_multi.py 2>&1 | grep -E "ok|FAIL|PASS|rror" | cut -c1-220
timeout 400 $TR --nproc-per-node 8 --master-port 29543 bench.py --gpus 8 --steps 200 --warmup 10 2>&1 | tail -1 > gpurun_out/final2_ours_8gpu.json; cut -c1-300 gpurun_out/final2_ours_8gpu.json
timeout 400 $TR --nproc-per-node 4 --master-port 29544 bench.py --gpus 4 --steps 200 --warmup 10 2>&1 | tail -1 > gpurun_out/final2_ours_4gpu.json; cut -c1-300 gpurun_out/final2_ours_4gpu.json
timeout 400 $TR --nproc-per-node 2 --master-port 29545 bench.py --gpus 2 --steps 200 --warmup 10 2>&1 | tail -1 > gpurun_out/final2_ours_2gpu.json; cut -c1-300 gpurun_out/final2_ours_2gpu.json
