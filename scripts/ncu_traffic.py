#!/usr/bin/env python
"""profiles/traffic.json from `ncu --set full` captures: per kernel, the per-launch DRAM traffic
(dram__bytes_read.sum + dram__bytes_write.sum) and L2 traffic of the LAST captured launch of that
kernel (the warm one in scripts/prof_run.py).  bench.py reads it for roofline.traffic.
usage: python scripts/ncu_traffic.py gpurun_out/a.ncu-rep [b.ncu-rep ...]"""
import csv
import io
import json
import os
import re
import subprocess
import sys

SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    out_path = os.path.join(ROOT, "profiles", "traffic.json")
    out = json.load(open(out_path)) if os.path.exists(out_path) else {}
    for path in sys.argv[1:]:
        raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(raw)))
        hdr, units = rows[0], rows[1]

        def val(r, key):
            if key not in hdr:
                return None
            i = hdr.index(key)
            return float(r[i].replace(",", "")) * SCALE.get(units[i], 1.0) if r[i] != "" else None

        for r in rows[2:]:
            name = re.sub(r"\(.*", "", r[hdr.index("Kernel Name")]).replace("<unnamed>::", "").replace("void ", "").strip()
            name = re.sub(r"<.*>", "", name)
            rd, wr = val(r, "dram__bytes_read.sum"), val(r, "dram__bytes_write.sum")
            if rd is None or wr is None:
                continue
            sect = val(r, "lts__t_sectors.sum")
            sect_rd = val(r, "lts__t_sectors_srcunit_tex_op_read.sum")
            out[name] = {"dram_bytes": rd + wr, "dram_read": rd, "dram_write": wr,
                         # L2 traffic of the launch: every sector the L2 slices served (32 B each), the sectors the
                         # SMs' loads asked for, and the bytes the crossbar delivered to the L1s
                         "l2_bytes": sect * 32.0 if sect is not None else None,
                         "l2_read_sectors_from_sm": sect_rd,
                         "l2_to_sm_read_bytes": val(r, "l1tex__m_xbar2l1tex_read_bytes.sum"),
                         "l2_hit_rate_pct": val(r, "lts__t_sector_hit_rate.pct"),
                         "l1_hit_rate_pct": val(r, "l1tex__t_sector_hit_rate.pct"),
                         "duration_us_under_ncu": (val(r, "gpu__time_duration.sum") or 0) * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(units[hdr.index("gpu__time_duration.sum")], 1.0),
                         "source": os.path.basename(path), "grid": r[hdr.index("Grid Size")]}
    with open(out_path, "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print(json.dumps({k: v["dram_bytes"] for k, v in out.items()}))


if __name__ == "__main__":
    main()
