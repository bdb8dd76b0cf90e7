mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests/ -x -q -m gpu ) > gpurun_out/v_pytest.log 2>&1; echo "pytest rc=$?" 
tail -5 gpurun_out/v_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
( time python bench.py ) > gpurun_out/v_bench.log 2>&1; echo "bench rc=$?"; tail -c 600 gpurun_out/v_bench.log
( time python bench.py --impl reference --steps 4 --warmup 1 ) > gpurun_out/v_bench_ref.log 2>&1; echo "ref rc=$?"; tail -c 500 gpurun_out/v_bench_ref.log
