O=gpurun_out/r02_o; mkdir -p $O
ls -la data/cfg3 > $O/data_ls.txt 2>&1
( time python -m pytest tests -x -q -m gpu -k "widened or no_alloc or larger_than_l2 or reference_files" ) > $O/gpu_tests.log 2>&1; echo "tests rc=$?"; tail -3 $O/gpu_tests.log
python tools/trace_run.py cfg3 runs 8 > $O/trace_runs.log 2>&1; tail -4 $O/trace_runs.log
python tools/trace_run.py cfg3 ms 6 > $O/trace_ms.log 2>&1; grep -E "trace_run|widened" $O/trace_ms.log | tail -8
python tools/trace_run.py cfg3 mems 6 > $O/trace_mems.log 2>&1; grep -E "trace_run|widened" $O/trace_mems.log | tail -3
( time python bench.py --steps 10 --warmup 3 ) > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"; tail -12 $O/bench.err
MONI_GPU_TRACE=1 python tests/scale/long_run.py --workload cfg3 --reads 200000 --parity-reads 0 --runs-reads 1000 --dynamic 1 --out $O/long_cfg3.json > $O/long.log 2>&1; echo "long rc=$?"
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__thread_inst_executed_per_inst_executed.ratio,sm__warps_active.avg.pct_of_peak_sustained_active,lts__t_sector_hit_rate.pct --clock-control none -k regex:'k_ms_pointers_dyn|k_ms_lengths_dyn' -c 12 --csv --log-file $O/long_launches.csv python tests/scale/long_run.py --workload cfg3 --reads 100000 --parity-reads 0 --runs-reads 1000 --dynamic 1 --out $O/long_cfg3_ncu.json > $O/long_ncu.log 2>&1; echo "long ncu rc=$?"
ls -la $O
