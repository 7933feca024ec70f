"""profiles/gemm_traffic.json from an `ncu --set full` report of one C3 layer's GEMM launches:

    python tools/ncu_traffic.py report.ncu-rep [config ...]        (default configs: C3 C2)

Takes every `gemm*_bf16_tcgen05_kernel` launch in the report, averages dram__bytes_read.sum + dram__bytes_write.sum per
launch and stores it together with the sha256 of the CUDA sources (bench.kernel_source_hash), so that bench.py reports
`roofline.traffic` only for the build the capture was taken from (and null once the kernels change).
"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def to_bytes(v, unit):
    v = float(v.replace(",", ""))
    u = unit.lower()
    return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)


def main():
    rep = sys.argv[1]
    cfgs = sys.argv[2:] or ["C3", "C2"]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h, units = rows[0], rows[1]
    ir, iw, ik = h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum"), h.index("Kernel Name")
    per = []
    for r in rows[2:]:
        if "_bf16_tcgen05_kernel" in r[ik] and "gemm" in r[ik]:
            per.append(to_bytes(r[ir], units[ir]) + to_bytes(r[iw], units[iw]))
    if not per:
        raise SystemExit("no GEMM launches in the report")
    from bench import kernel_source_hash
    d = {"source_hash": kernel_source_hash(), "report": os.path.basename(rep), "launches": len(per),
         "bytes_per_launch": {c: sum(per) / len(per) for c in cfgs}}
    with open(os.path.join(ROOT, "profiles", "gemm_traffic.json"), "w") as f:
        json.dump(d, f, indent=1)
    print(d)


if __name__ == "__main__":
    main()
