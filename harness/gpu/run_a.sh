set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r2a_tests.log 2>&1
echo "tests rc=$?"
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2a_smoke.log 2>&1; echo "smoke rc=$?"
python bench.py --steps 20 --warmup 8 > gpurun_out/r2a_bench_ours.json 2> gpurun_out/r2a_bench_ours.err; echo "bench rc=$?"
python bench.py --impl reference --steps 10 --warmup 3 --no-cpu > gpurun_out/r2a_bench_ref.json 2> gpurun_out/r2a_bench_ref.err; echo "ref rc=$?"
tail -c 1500 gpurun_out/r2a_tests.log
