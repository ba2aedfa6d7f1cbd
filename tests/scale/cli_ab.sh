# A/B timing of the drop-in CLIs on the bench workload (2M reads), three repeats; prints the CLI's own phase lines.
set -e
cd "$(dirname "$0")/../.."
python - <<'PY'
import sys, os
sys.path.insert(0,'tools')
import workloads, synth
pre = workloads.ensure_index('cfg2s')
reads = workloads.reads_of('cfg2s', pairs=1000000)
os.makedirs('/tmp/ab', exist_ok=True)
synth.write_fastq_matrix('/tmp/ab/reads.fq', reads)
PY
for rep in 1 2; do
 for exe in ms mems; do
   s=$(date +%s%N)
   MONI_GPU_VERBOSE=1 moni_align_b200/bin/$exe data/cfg2s/ref.fa -p /tmp/ab/reads.fq -o /tmp/ab/o -t 16 > /tmp/ab/log 2>&1 || true
   e=$(date +%s%N)
   echo "== $rep $exe wall $(( (e - s) / 1000000 )) ms"; grep "Elapsed\|gpu-call\|cli \|index load" /tmp/ab/log
 done
done
