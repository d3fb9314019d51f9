set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -p no:cacheprovider -x > gpurun_out/r2b_tests.log 2>&1
echo "tests rc=$?"
tail -5 gpurun_out/r2b_tests.log
python bench.py --steps 40 --warmup 8 --no-cpu > gpurun_out/r2b_bench_ours.json 2> gpurun_out/r2b_bench_ours.err; echo "bench rc=$?"
tail -3 gpurun_out/r2b_bench_ours.err
