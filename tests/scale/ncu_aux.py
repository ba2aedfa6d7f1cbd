#!/usr/bin/env python
"""A short run that launches the kernels bench.py does not time (k_ms_untranspose, k_seed_gather, k_occ_count, k_occ_fill) once
each at a useful size, for an ncu capture:  ncu --set full -k regex:'k_occ|k_ms_untranspose|k_seed_gather' python tests/scale/ncu_aux.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import moni_align_b200 as m  # noqa: E402
import workloads  # noqa: E402

prefix = os.path.join(ROOT, "data", "cfg4s_full", "ref.fa")
ix = m.MoniIndex(prefix, gpus=[0])
reads = workloads.reads_of("cfg4s", pairs=int(os.environ.get("PAIRS", "500000")))
flat, offs = m.pack_reads(reads)
rb = ix.resident_batch(flat, offs)
print("ms", rb.run("ms"))
rb.close()
so, ref, qpos, ln = ix.find_mems(flat, offs, 25)
oo, occs, trunc = ix.find_MEM_occs(ref, ln, max_per_direction=4096)
print("seeds", len(ref), "occurrences", len(occs), ix.timing())
