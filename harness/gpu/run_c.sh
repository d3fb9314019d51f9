set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r2c_tests.log 2>&1
echo "tests rc=$?"
tail -15 gpurun_out/r2c_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2c_smoke.log 2>&1; echo "smoke rc=$?"
tail -3 gpurun_out/r2c_smoke.log
python bench.py --steps 40 --warmup 8 --no-cpu > gpurun_out/r2c_bench_ours.json 2> gpurun_out/r2c_bench_ours.err; echo "bench rc=$?"
tail -3 gpurun_out/r2c_bench_ours.err
H3DGS_HOST=ctypes python bench.py --steps 40 --warmup 8 --no-cpu --no-raster-only > gpurun_out/r2c_bench_ctypes.json 2> gpurun_out/r2c_bench_ctypes.err; echo "bench ctypes rc=$?"
cat gpurun_out/r2c_bench_ours.json | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','e2e','gpu_launches') if k in d}); print(d.get('kernel_ms')); print(d.get('roofline'))"
cat gpurun_out/r2c_bench_ctypes.json | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','e2e') if k in d})"
