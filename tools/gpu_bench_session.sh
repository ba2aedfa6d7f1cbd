# One GPU-box session: GPU suite, reference arm (builds and caches the cfg3 index on the box), our arm, ncu launch list, ncu --set full
# of the two passes.  Usage: gpurun -- 'bash tools/gpu_bench_session.sh <tag>'   (outputs under gpurun_out/<tag>/)
O=gpurun_out/${1:-session}; mkdir -p $O
( time python -m pytest tests -x -q -m gpu ) > $O/gpu_suite.log 2>&1; echo "suite rc=$?"; tail -3 $O/gpu_suite.log
( time python bench.py --impl reference --steps 2 --warmup 1 ) > $O/bench_reference.json 2> $O/bench_reference.err; echo "ref rc=$?"
( time python bench.py --steps 10 --warmup 3 ) > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
tail -4 $O/bench.err
if [ -z "$NO_NCU" ]; then
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_ms_pointers|k_ms_lengths' -s 6 -c 2 -o $O/full python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/ncu_full.log 2>&1
ncu -i $O/full.ncu-rep --page raw --csv > $O/full_raw.csv 2>/dev/null
ncu -i $O/full.ncu-rep --page details > $O/full_details.txt 2>/dev/null
rm -f $O/full.ncu-rep.tmp
fi
ls -la $O
