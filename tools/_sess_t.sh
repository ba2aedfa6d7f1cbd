O=gpurun_out/r02_t; mkdir -p $O
( time python -m pytest tests -x -q -m gpu -k "dynamic or widened or ragged or chunking or sharded or concurrent or no_alloc or larger_than_l2" ) > $O/gpu_tests.log 2>&1; echo "tests rc=$?"; tail -3 $O/gpu_tests.log
MONI_GPU_TRACE=1 python tests/scale/long_run.py --workload cfg3 --reads 1000000 --parity-reads 20000 --runs-reads 1000 --dynamic L4R3G0,L4R3G1,L6R3G1,L6R2G1 --out $O/long_cfg3_1M_guided.json > $O/long.log 2>&1; echo "long rc=$?"; grep -E "^\[long_run\] L" $O/long.log | cut -c1-200
grep -c identical $O/long_cfg3_1M_guided.json
