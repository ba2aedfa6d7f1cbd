"""Generates `<case>.fa.slp` (the shaped grammar, SelfShapedSlp<u32, DagcSd, DagcSd, SelSd>) for the golden
cases with the UNMODIFIED reference encoder (oracle/_ref/bin/SlpEncBuild -e SelfShapedSlp_SdSd_Sd, reference
thirdparty/ShapedSlp/SlpEncBuild.cpp:145), and `<case>.text.sha256` = the hash of the text the reference's own
decoder (oracle/_ref/bin/ref_slp_text = SelfShapedSlp::load + expandSubstr) writes for that file.

    python tests/golden/make_golden_shaped.py

Same FASTA and same .R/.C as `<case>.fa.plain.slp`, so both grammars expand to the same text.
"""
import hashlib
import os
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tools"))
sys.path.insert(0, HERE)
import build_index  # noqa: E402
import make_golden  # noqa: E402
import synth  # noqa: E402


def shaped_slp(fasta: str, trivial_grammar: bool = False) -> str:
    """FASTA -> <fasta>.slp via tools/mkgrammar (.R/.C) and the reference encoder."""
    mk = build_index.build_mkgrammar()
    subprocess.check_call([mk, fasta, fasta] + (["--trivial"] if trivial_grammar else []),
                          stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    subprocess.check_call([build_index.ref_tool("SlpEncBuild"), "-i", fasta, "-o", fasta + ".slp",
                           "-e", "SelfShapedSlp_SdSd_Sd", "-f", "Bigrepair"],
                          stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return fasta + ".slp"


def main():
    for name, records, kw in make_golden.cases():
        with tempfile.TemporaryDirectory() as td:
            fa = os.path.join(td, name + ".fa")
            synth.write_fasta(fa, records)
            # (the reference encoder crashes on the left-deep "trivial" grammar of the twoletter case: that case gets the
            # regular grammar here; the text is the same)
            slp = shaped_slp(fa, False)
            text = subprocess.check_output([build_index.ref_tool("ref_slp_text"), slp])
            dst = os.path.join(HERE, name + ".fa.slp")
            with open(slp, "rb") as f, open(dst, "wb") as g:
                g.write(f.read())
            with open(os.path.join(HERE, name + ".text.sha256"), "w") as g:
                g.write(hashlib.sha256(text).hexdigest() + " " + str(len(text)) + "\n")
            print(name, os.path.getsize(dst), "bytes; text", len(text))


if __name__ == "__main__":
    main()
