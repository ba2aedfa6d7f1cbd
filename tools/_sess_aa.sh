O=gpurun_out/r02_aa; mkdir -p $O
for t in 4 1 0; do
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --opt widen_threads=$t > $O/bench_wt$t.json 2> $O/bench_wt$t.err; echo "wt=$t rc=$?"
python - <<PY
import json
d=[json.loads(l) for l in open("$O/bench_wt$t.json") if l.strip().startswith("{")][0]
print("wt=$t", "e2e", round(d["e2e"]["value"]/1e6,1), d["e2e"]["d2h_bytes_per_step"], d["e2e"].get("host_widen_threads"), "link" , round(d.get("e2e_mems_u64_link",{}).get("value",0)/1e6,1), "ms e2e", round(d["ms"]["e2e"]["value"]/1e6,1), d["ms"]["e2e"]["d2h_bytes_per_step"], [k for k in d["ms"] if k.startswith("e2e")], all(v.get("same_bits_as_resident_run",True) for k,v in d.items() if k.startswith("e2e") and isinstance(v,dict)))
PY
done
