cd /root/repo
mkdir -p gpurun_out/r2b
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
GENTUN_B200_LIB=/root/repo/build_exp/libprof.so timeout 300 python tools/prep_phases.py c4 c5 2>&1 | grep -v "labels\|non-finite\|tables\|host\|row orders"
python bench.py > gpurun_out/r2b/bench_c3_n1.json 2> gpurun_out/r2b/bench_c3_n1.err
python bench.py --workload c4 --steps 2 --warmup 3 --e2e-steps 1 > gpurun_out/r2b/bench_c4_n1.json 2> gpurun_out/r2b/bench_c4_n1.err
python bench.py --workload c5 --steps 1 --warmup 3 --e2e-steps 1 > gpurun_out/r2b/bench_c5_n1.json 2> gpurun_out/r2b/bench_c5_n1.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2b/*.json')):
    try:
        d=json.loads(open(f).read().strip().split('\n')[-1]); print(f, 'value %.3f e2e %.3f ms/step %.0f frac %s prepare_ms %.0f' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['roofline']['frac'], d.get('prepare_ms',0)))
    except Exception as ex: print(f, 'ERR', ex)
PY
