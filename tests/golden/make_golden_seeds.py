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


def lcp_variant():
    """<case>.fa.thrbv.full.lcp.ms: the sampled-LCP align index (reference rlebwt_ms_full_build without -n).  It needs the
    builder's .slcp, which is not kept under tests/golden/, so the text is regenerated from its seed (make_golden.cases), the
    builder is re-run, and its run files are checked against the committed ones.  The reference's LCP-based enumeration
    (seed_finder<plain_slp_t, moni_lcp<>>, ref_seeds_lcp) must print exactly the committed .seeds25 / .occs12."""
    import filecmp
    import sys
    sys.path.insert(0, HERE)
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import make_golden
    import synth
    for case, records, kw in make_golden.cases():
        with tempfile.TemporaryDirectory() as td:
            fa = os.path.join(td, case + ".fa")
            synth.write_fasta(fa, records)
            for cmd in ([os.path.join(BIN, "newscanNT.x"), fa, "-w", "10", "-p", "100", "-f"], [os.path.join(BIN, "pfp_thresholds"), fa, "-w", "10", "-r"]):
                subprocess.run(cmd, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            for ext in (".bwt.heads", ".bwt.len", ".ssa", ".esa", ".thr_pos"):
                assert filecmp.cmp(fa + ext, os.path.join(HERE, case + ".fa" + ext), shallow=False), (case, ext)
            subprocess.run([os.path.join(BIN, "rlebwt_ms_full_build"), fa], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            shutil.copy(os.path.join(HERE, case + ".fa.plain.slp"), td)
            open(fa + ".ldx", "wb").close()
            fq = os.path.join(HERE, case + ".fq")
            subprocess.run([os.path.join(BIN, "ref_seeds_lcp"), fa, fq, "25", os.path.join(td, "s25")], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            subprocess.run([os.path.join(BIN, "ref_seeds_lcp"), fa, fq, "12", os.path.join(td, "o12"), "occs"], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            assert filecmp.cmp(os.path.join(td, "s25"), os.path.join(HERE, case + ".seeds25"), shallow=False), case
            assert filecmp.cmp(os.path.join(td, "o12"), os.path.join(HERE, case + ".occs12"), shallow=False), case
            shutil.copy(fa + ".thrbv.full.lcp.ms", HERE)


def main():
    lcp_variant()
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
            # seed_finder::populate_seeds (half-MEM splitting, seed filtering off) on the same MEMs
            subprocess.run([os.path.join(BIN, "ref_seeds"), fa, os.path.join(td, case + ".fq"), "12", os.path.join(td, f"{case}.pop12"), "pop"],
                           check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            shutil.copy(os.path.join(td, f"{case}.pop12"), HERE)
            shutil.copy(fa + ".thrbv.full.ms", HERE)


if __name__ == "__main__":
    main()
    print(sorted(f for f in os.listdir(HERE) if "seeds" in f or "full" in f or "occs" in f))
