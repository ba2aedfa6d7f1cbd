import numpy as np, time, sys, os
ROOT=os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0,ROOT); sys.path.insert(0,os.path.join(ROOT,"tools"))
import moni_align_b200 as m, workloads
pre=workloads.index_prefix("cfg2s")
reads=workloads.reads_of("cfg2s", pairs=1000000)
n,L=reads.shape
hs=m.pinned_empty(n*L,np.uint8); hs[:]=reads.reshape(-1)
ho=m.pinned_empty(n+1,np.uint64); ho[:]=np.arange(n+1,dtype=np.uint64)*np.uint64(L)
out=(m.pinned_empty(n+1,np.uint64), m.pinned_empty(n*L,np.uint32))
ix=m.MoniIndex(pre,gpus=[0])
for _ in range(3): ix.maxrimal_exact_matches_packed(hs,ho,out=out)
os.environ["MONI_GPU_TRACE"]="1"
t=time.perf_counter()
ix.maxrimal_exact_matches_packed(hs,ho,out=out)
print("wall ms", (time.perf_counter()-t)*1e3, ix.timing(), file=sys.stderr)
