"""Adds the `moni align` front-end fixtures to the committed golden cases, by running UNMODIFIED reference code in THIS
container (oracle/_ref/bin, built from /root/reference by oracle/build_ref.sh):

    python tests/golden/make_golden_seeds.py

For each case already under tests/golden/ (its builder intermediates .bwt.heads/.bwt.len/.ssa/.esa/.thr_pos are kept there):
  <case>.fa.thrbv.full.ms   reference rlebwt_ms_full_build -n   (src/align/rlebwt_ms_full_build.cpp: the align index)
  <case>.seeds0 / .seeds25  reference seed_finder<plain_slp_t, ms_pointers<>>::find_mems with min_len 0 / 25, loaded from
                            that .thrbv.full.ms, printed per read as '(pos,len,idx) ' by oracle/ref_seeds_harness.cpp
  <case>.occs12             the same with min_len 12 and every seed followed by its occurrences (seed_finder::find_MEM_occs,
                            unfiltered): '(pos,len,idx:occ,occ,...) '
"""
import os
import shutil
import subprocess
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
BIN = os.path.join(ROOT, "oracle", "_ref", "bin")


def main():
    for case in ("pan4", "withN", "twoletter"):
        with tempfile.TemporaryDirectory() as td:
            for f in os.listdir(HERE):
                if f.startswith(case + "."):
                    shutil.copy(os.path.join(HERE, f), td)
            fa = os.path.join(td, case + ".fa")
            subprocess.run([os.path.join(BIN, "rlebwt_ms_full_build"), fa, "-n"], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            open(fa + ".ldx", "wb").close()      # the stubbed lift-over index (never read by find_mems)
            for ml in (0, 25):
                subprocess.run([os.path.join(BIN, "ref_seeds"), fa, os.path.join(td, case + ".fq"), str(ml),
                                os.path.join(td, f"{case}.seeds{ml}")], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
                shutil.copy(os.path.join(td, f"{case}.seeds{ml}"), HERE)
            subprocess.run([os.path.join(BIN, "ref_seeds"), fa, os.path.join(td, case + ".fq"), "12", os.path.join(td, f"{case}.occs12"), "occs"],
                           check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            shutil.copy(os.path.join(td, f"{case}.occs12"), HERE)
            shutil.copy(fa + ".thrbv.full.ms", HERE)


if __name__ == "__main__":
    main()
    print(sorted(f for f in os.listdir(HERE) if "seeds" in f or "full" in f or "occs" in f))
