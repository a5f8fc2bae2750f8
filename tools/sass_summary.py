"""Per-kernel SASS evidence of the Blackwell instructions in libb200vit.so (run where cuobjdump is installed):

    python tools/sass_summary.py > profiles/r02_sass_summary.txt

Counts, per kernel: UTCHMMA / UTCQMMA (tcgen05.mma), LDTM / STTM (tcgen05.ld / st), UTMALDG / UTMASTG (TMA load /
store), UTCBAR (tcgen05.commit), SYNCS (mbarrier), HMMA (mma.sync), FFMA2 (packed fp32), MUFU, and STL / LDL (spills)."""
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
lib = ROOT / "pytorch-vit-paper-replication_b200" / "b200vit" / "lib" / "libb200vit.so"
sass = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True).stdout
names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True, text=True).stdout.split("\n")
ops = ["UTCHMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTCBAR", "SYNCS", "HMMA", "FFMA2", "MUFU", "STL", "LDL"]
print(f"{lib.name}: {lib.stat().st_size / 1e6:.1f} MB; instruction counts per kernel (static SASS)")
print("kernel".ljust(78) + "".join(o.rjust(8) for o in ops) + "   insts")
blocks = sass.split("Function : ")[1:]
for blk, name in zip(blocks, names):
    body = blk.split("\n", 1)[1]
    insts = re.findall(r"/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_]+)", body)
    cnt = {o: sum(1 for i in insts if i.startswith(o)) for o in ops}
    short = re.sub(r"\(.*", "", name).replace("b200vit::", "").replace("void ", "")
    print(short[:77].ljust(78) + "".join(str(cnt[o]).rjust(8) for o in ops) + str(len(insts)).rjust(8))
