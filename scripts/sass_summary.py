"""profiles/sass_summary.txt: per-kernel counts of the Blackwell-specific SASS mnemonics in libb200pipe.so."""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "distributed-neural-networks_b200", "libb200pipe.so")
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
rows = []
for f in re.split(r"\n\s*Function : ", txt)[1:]:
    name = f.split("\n", 1)[0].strip()
    c = collections.Counter({op: len(re.findall(r"\b" + op + r"\b", f)) for op in
                             ("UTCHMMA", "LDTM", "STTM", "UTMALDG", "UBLKCP", "UTCBAR", "FFMA2")})
    c["WATERFALL"] = len(re.findall(r"BRA\.U\.ANY", f))
    if c["UTCHMMA"] or c["LDTM"] or c["UTMALDG"] or c["UBLKCP"]:
        rows.append((subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()[:110], c))
out = ["# Per-kernel counts of the Blackwell-specific SASS mnemonics in libb200pipe.so (cuobjdump -sass, sm_100a).",
       "# UTCHMMA = tcgen05.mma (kind::f16 / tf32), LDTM / STTM = tcgen05.ld / st (tensor memory), UTMALDG = TMA tensor load,",
       "# UBLKCP = cp.async.bulk (bulk copies), UTCBAR = tcgen05.commit, FFMA2 = packed fp32 FMA.  Regenerate: python scripts/sass_summary.py",
       "# BRA.U.ANY = ELECT / R2UR.BROADCAST loops around single-thread instructions issued from divergent code (what is left belongs to TMA loads).",
       "%-112s %8s %6s %6s %8s %7s %7s %6s %10s" % ("kernel", "UTCHMMA", "LDTM", "STTM", "UTMALDG", "UBLKCP", "UTCBAR", "FFMA2", "BRA.U.ANY")]
for name, c in sorted(rows, key=lambda r: -r[1]["UTCHMMA"]):
    out.append("%-112s %8d %6d %6d %8d %7d %7d %6d %10d" % (name, c["UTCHMMA"], c["LDTM"], c["STTM"], c["UTMALDG"], c["UBLKCP"], c["UTCBAR"], c["FFMA2"], c["WATERFALL"]))
open(os.path.join(ROOT, "profiles", "sass_summary.txt"), "w").write("\n".join(out) + "\n")
print("\n".join(out))
