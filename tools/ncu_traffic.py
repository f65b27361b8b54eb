#!/usr/bin/env python3
"""Extract per-launch DRAM traffic (dram__bytes_read.sum + dram__bytes_write.sum) of named kernels from `ncu --set full`
reports and merge it into profiles/ncu_traffic.json, which bench.py reads for `roofline.traffic` (the .ncu-rep files
themselves are too large to commit).

    python tools/ncu_traffic.py gpurun_out/r02_lstm.ncu-rep lstm_bwd=lstm_bwd_ks_kernel lstm_fwd=lstm_fwd_kernel
    python tools/ncu_traffic.py gpurun_out/r02_stft.ncu-rep stft=stft_dft256_kernel

Each key=regex pair picks the LAST launch whose kernel name matches the regex.
"""
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles", "ncu_traffic.json")
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def main():
    rep, pairs = sys.argv[1], [a.split("=", 1) for a in sys.argv[2:]]
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--metrics",
                          "dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum"],
                         stdout=subprocess.PIPE, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units, body = rows[0], rows[1], rows[2:]
    col = {n: i for i, n in enumerate(hdr)}
    data = json.load(open(OUT)) if os.path.exists(OUT) else {}
    for key, rx in pairs:
        hit = [r for r in body if re.search(rx, r[col["Kernel Name"]])]
        if not hit:
            print("no launch matches", rx)
            continue
        r = hit[-1]
        val = lambda m: float(r[col[m]].replace(",", "")) * UNIT.get(units[col[m]], 1.0)
        rd, wr = val("dram__bytes_read.sum"), val("dram__bytes_write.sum")
        data[key] = {"bytes": rd + wr, "read": rd, "write": wr, "kernel": r[col["Kernel Name"]][:80],
                     "ncu_duration": "%s %s" % (r[col["gpu__time_duration.sum"]], units[col["gpu__time_duration.sum"]]),
                     "source": "ncu --set full, one launch, %s (tools/ncu_traffic.py)" % os.path.basename(rep)}
        print(key, data[key])
    json.dump(data, open(OUT, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
