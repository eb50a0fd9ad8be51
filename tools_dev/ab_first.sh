# A/B of -DSB2_SPLIT_FIRST (front vehicles of a block's lanes grouped in the work list): parity tests and bench for both builds
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/ff_pytest.log 2>&1; tail -2 gpurun_out/ff_pytest.log
SUMOB200_LIB=$PWD/build_variants/first.so timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/ff_pytest_first.log 2>&1; tail -2 gpurun_out/ff_pytest_first.log
timeout 300 python bench.py > gpurun_out/ff_bench.json 2> gpurun_out/ff_bench.err
SUMOB200_LIB=$PWD/build_variants/first.so timeout 300 python bench.py > gpurun_out/ff_bench_first.json 2> gpurun_out/ff_bench_first.err
for f in ff_bench ff_bench_first; do python - <<PY
import json
d=json.loads(open("gpurun_out/$f.json").read().strip().splitlines()[-1])
print("$f", d["value"]/1e6, d["ms_per_step"], d["e2e"]["value"]/1e6, d["roofline"]["achieved"], {k: v for k, v in d.items() if "kernel" in k})
PY
done
