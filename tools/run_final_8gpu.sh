set -x
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 $TR --nproc-per-node 8 --master-port 29541 tools/gpu_check_multi.py 2>&1 | tail -7
timeout 400 $TR --nproc-per-node 8 --master-port 29542 tools/gpu_check_multi_gpt.py 2>&1 | tail -8
timeout 400 $TR --nproc-per-node 8 --master-port 29543 bench.py --gpus 8 --steps 100 --warmup 5 2>&1 | tail -1 > gpurun_out/final_ours_8gpu.json; cut -c1-1800 gpurun_out/final_ours_8gpu.json
timeout 400 $TR --nproc-per-node 4 --master-port 29544 bench.py --gpus 4 --steps 100 --warmup 5 2>&1 | tail -1 > gpurun_out/final_ours_4gpu.json; cut -c1-1800 gpurun_out/final_ours_4gpu.json
timeout 300 $TR --nproc-per-node 8 --master-port 29545 bench.py --gpus 8 --config gpt2s_fsdp --steps 10 --warmup 3 --no-extra 2>&1 | tail -1 | cut -c1-900
timeout 300 $TR --nproc-per-node 8 --master-port 29546 bench.py --gpus 8 --config mlp_fsdp --steps 100 --warmup 5 --no-extra 2>&1 | tail -1 | cut -c1-900
timeout 400 $TR --nproc-per-node 8 --master-port 29547 bench.py --gpus 8 --config gpt2m_2d --steps 8 --warmup 3 --no-extra 2>&1 | tail -1 | cut -c1-900
timeout 300 $TR --nproc-per-node 4 --master-port 29548 bench.py --gpus 4 --config gpt2s_pp4 --steps 8 --warmup 3 --no-extra 2>&1 | tail -1 | cut -c1-900
timeout 300 $TR --nproc-per-node 8 --master-port 29549 bench.py --gpus 8 --impl baseline --steps 50 --warmup 5 2>&1 | tail -1 | cut -c1-500
timeout 300 $TR --nproc-per-node 8 --master-port 29550 bench.py --gpus 8 --impl baseline --config gpt2s_dp --steps 8 --warmup 3 2>&1 | tail -1 | cut -c1-500
