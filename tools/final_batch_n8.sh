set -x
cd /root/repo
mkdir -p gpurun_out/final8
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
nvidia-smi -L | head -8
$T bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/final8/bench_c3_n8.json 2> gpurun_out/final8/bench_c3_n8.err
$T bench.py --gpus 8 --workload c5 --steps 2 --warmup 2 --e2e-steps 1 > gpurun_out/final8/bench_c5_n8.json 2> gpurun_out/final8/bench_c5_n8.err
$T bench.py --gpus 8 --workload c4 --steps 2 --warmup 2 --e2e-steps 1 > gpurun_out/final8/bench_c4_n8.json 2> gpurun_out/final8/bench_c4_n8.err
$T bench.py --gpus 8 --workload c1 --steps 2 --warmup 1 --e2e-steps 1 > gpurun_out/final8/bench_c1_n8.json 2> gpurun_out/final8/bench_c1_n8.err
tail -c 400 gpurun_out/final8/*.err
ls -la gpurun_out/final8
