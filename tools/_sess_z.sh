O=gpurun_out/r02_z; mkdir -p $O
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/smoke.log
NO_NCU=1 bash tools/gpu_bench_session.sh r02_z
