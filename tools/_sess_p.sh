O=gpurun_out/r02_p; mkdir -p $O
( time python -m pytest tests -x -q -m gpu -k "dynamic or widened or ragged or bounds or runs or no_alloc" ) > $O/gpu_tests.log 2>&1; echo "tests rc=$?"; tail -3 $O/gpu_tests.log
python tests/scale/long_run.py --workload cfg3 --reads 400000 --parity-reads 2000 --runs-reads 1000 --dynamic 101,102,103,104 --out $O/long_cfg3_rpt.json > $O/long.log 2>&1; echo "long rc=$?"; grep -E "^\[|wall_ms" $O/long.log | cut -c1-400 | tail -12
