"""Write profiles/ncu_traffic.json from an `ncu --set full` capture (run HERE, no GPU needed):

    python tools/ncu_traffic.py gpurun_out/<capture>.ncu-rep --batch 256 --chans 32 --blocks 10 --dtype bf16

bench.py prints `roofline.traffic` from this file only while its `source_sha16` equals the hash of the kernel sources it is
running and the configuration matches -- a capture of an older build reads as null, never as a stale number.  Per kernel the
LAST captured launch is kept (the earlier ones may be warm-up); values are per launch."""
import argparse
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import source_sha16  # noqa: E402

# bench.py breakdown key <- substring of the kernel name
KEYS = {"fwd_blocks(persistent)": "k_tc_fwd_persist", "bwd_blocks(persistent)": "k_tc_bwd_persist", "stem_fwd": "k_stem_tc_fwd",
        "stem_bwd": "k_stem_tc_bwd", "head": "k_head_fused", "body_reduce_sgd": "k_body_reduce", "sgd_pack": "k_sgd_pack"}
UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1, "ms": 1e3, "ns": 1e-3, "s": 1e6, "usecond": 1, "msecond": 1e3, "nsecond": 1e-3, "second": 1e6}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep")
    ap.add_argument("--batch", type=int, default=256); ap.add_argument("--chans", type=int, default=32)
    ap.add_argument("--blocks", type=int, default=10); ap.add_argument("--dtype", default="bf16")
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "ncu_traffic.json"))
    a = ap.parse_args()
    raw = subprocess.run(["ncu", "-i", a.rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}

    def val(r, name):
        i = col[name]
        return float(r[i].replace(",", "")) * UNIT.get(units[i], 1)

    kernels = {}
    for r in rows[2:]:
        name = r[col["Kernel Name"]]
        for key, sub in KEYS.items():
            if sub in name:
                kernels[key] = {"kernel": name.split("(")[0], "dram_read_bytes": val(r, "dram__bytes_read.sum"), "dram_write_bytes": val(r, "dram__bytes_write.sum"),
                                "duration_us_under_ncu": val(r, "gpu__time_duration.sum"),
                                "tensor_pipe_active_pct": float(r[col["sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"]]),
                                "warps_active_pct": float(r[col["sm__warps_active.avg.pct_of_peak_sustained_active"]]),
                                "lts_throughput_pct": float(r[col["lts__throughput.avg.pct_of_peak_sustained_elapsed"]])}
    out = {"source_sha16": source_sha16(), "config": {"B": a.batch, "C": a.chans, "NB": a.blocks, "dtype": a.dtype},
           "capture": os.path.basename(a.rep), "how": "ncu --set full --clock-control none, one launch per kernel (values per launch)", "kernels": kernels}
    with open(a.out, "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
