set -x
python profiles/prof_sweep.py > gpurun_out/p1.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:sweep2tc_kernel -c 1 -o gpurun_out/r3h_sweep2tc python profiles/prof_sweep.py > gpurun_out/n1.log 2>&1
python profiles/prof_lin.py 6553600 soa > gpurun_out/p2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:linearize_tile -c 1 -o gpurun_out/r3h_lin_tile python profiles/prof_lin.py 6553600 soa > gpurun_out/n2.log 2>&1
python profiles/prof_hess.py 262144 > gpurun_out/p3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:hessian2 -c 1 -o gpurun_out/r3h_hessian2 python profiles/prof_hess.py 262144 > gpurun_out/n3.log 2>&1
python bench.py > gpurun_out/r3h_bench_n1.json 2> gpurun_out/bench.err
tail -c 600 gpurun_out/r3h_bench_n1.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r3h_launches.csv python bench.py --steps 2 --warmup 1 > gpurun_out/ncu_bench.log 2>&1
ls -la gpurun_out | tail -12
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python -m pytest tests/test_zz_no_alloc.py -x -q 2>&1 | tail -2
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
