"""Condense an `ncu --page raw --csv` export of tools/profile_kernels.py into the numbers DESIGN.md / bench.py cite.

    python tools/summarize_ncu.py raw.csv [out.txt] [traffic.json]
Writes one block per profiled launch (duration, DRAM bytes, DRAM / tensor-pipe / issue utilisation, registers, smem)
and, for the 14 GEMM launches (in bench.gemm_shapes_per_step order), profiles/r01_gemm_traffic.json.
"""
import csv
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "lts__t_sector_hit_rate.pct", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__grid_size", "launch__block_size"]

rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
hdr, units = rows[hi], rows[hi + 1]
col = {k: hdr.index(k) for k in KEYS if k in hdr}
kn = hdr.index("Kernel Name")
launches = []
for r in rows[hi + 2:]:
    if len(r) <= kn:
        continue
    rec = {"name": r[kn]}
    for k, c in col.items():
        try:
            rec[k] = float(r[c].replace(",", ""))
        except ValueError:
            rec[k] = r[c]
        rec[k + "@unit"] = units[c]
    launches.append(rec)


def to_bytes(v, unit):
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


def to_us(v, unit):
    return v * {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(unit, 1)


out = []
for i, L in enumerate(launches):
    rd = to_bytes(L.get("dram__bytes_read.sum", 0), L.get("dram__bytes_read.sum@unit"))
    wr = to_bytes(L.get("dram__bytes_write.sum", 0), L.get("dram__bytes_write.sum@unit"))
    us = to_us(L.get("gpu__time_duration.sum", 0), L.get("gpu__time_duration.sum@unit"))
    L["_dram"], L["_us"] = rd + wr, us
    out.append(f"[{i}] {L['name'][:110]}\n"
               f"    duration {us:.1f} us | DRAM read {rd / 1e6:.1f} MB write {wr / 1e6:.1f} MB ({(rd + wr) / us / 1e3:.0f} GB/s) | "
               f"dram {L.get('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 0):.1f}% tensor "
               f"{L.get('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 0):.1f}% issue "
               f"{L.get('smsp__issue_active.avg.pct_of_peak_sustained_active', 0):.1f}% | L2 hit "
               f"{L.get('lts__t_sector_hit_rate.pct', 0):.1f}% | regs {L.get('launch__registers_per_thread', 0):.0f} "
               f"smem {L.get('launch__shared_mem_per_block_dynamic', 0):.1f} {L.get('launch__shared_mem_per_block_dynamic@unit', '')} "
               f"grid {L.get('launch__grid_size', 0):.0f} x {L.get('launch__block_size', 0):.0f} | inst {L.get('smsp__inst_executed.sum', 0):.3g}")
text = "\n".join(out)
print(text)
if len(sys.argv) > 2:
    Path(sys.argv[2]).write_text(text + "\n")
if len(sys.argv) > 3:
    import bench
    gemm = [L for L in launches if "gemm_tcgen05" in L["name"]]
    shapes = bench.gemm_shapes_per_step(256)
    assert len(gemm) >= len(shapes), (len(gemm), len(shapes))
    rec = {"batch": 256, "source": "ncu --set full, one launch per GEMM shape (tools/profile_kernels.py); " + Path(sys.argv[2]).name,
           "dram_bytes": {s[0]: round(g["_dram"]) for s, g in zip(shapes, gemm)},
           "duration_us_under_ncu": {s[0]: round(g["_us"], 1) for s, g in zip(shapes, gemm)}}
    Path(sys.argv[3]).write_text(json.dumps(rec, indent=1) + "\n")
