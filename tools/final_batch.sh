set -x
cd /root/repo
mkdir -p gpurun_out/final
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py --steps 20 --warmup 5 > gpurun_out/final/bench_c3_n1.json 2> gpurun_out/final/bench_c3_n1.err
(time python bench.py --impl reference --steps 20 --warmup 5) > gpurun_out/final/ref_c3.json 2> gpurun_out/final/ref_c3.err
for w in c1 c2; do
  python bench.py --workload $w --steps 3 --warmup 3 --e2e-steps 2 > gpurun_out/final/bench_${w}_n1.json 2> gpurun_out/final/bench_${w}_n1.err
  (time python bench.py --impl reference --workload $w --steps 10 --warmup 1 --cpu-budget-s 300) > gpurun_out/final/ref_$w.json 2> gpurun_out/final/ref_$w.err
done
python bench.py --workload c4 --steps 3 --warmup 3 --e2e-steps 1 > gpurun_out/final/bench_c4_n1.json 2> gpurun_out/final/bench_c4_n1.err
python bench.py --workload c5 --steps 2 --warmup 3 --e2e-steps 1 > gpurun_out/final/bench_c5_n1.json 2> gpurun_out/final/bench_c5_n1.err
(time python bench.py --impl reference --workload c4 --steps 2 --warmup 0 --cpu-budget-s 240) > gpurun_out/final/ref_c4.json 2> gpurun_out/final/ref_c4.err
(time python bench.py --impl reference --workload c5 --steps 2 --warmup 0 --cpu-budget-s 240) > gpurun_out/final/ref_c5.json 2> gpurun_out/final/ref_c5.err
ls -la gpurun_out/final
