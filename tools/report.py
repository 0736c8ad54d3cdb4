#!/usr/bin/env python3
"""Summarise the sections a proxy run printed (this build's binaries or the reference's: same grammar).

    bin/dlnb_launch -n 8 bin/b200/cpp/data_parallel/dp llama3_8b_64_bfloat16 240 . > run.out
    python tools/report.py run.out [--csv out_prefix]

Prints one line of headline numbers per section; with --csv also writes the per-rank / per-run tables the
reference's plots read (plots/parser.py: dp -> one table, fsdp -> runtime and per-unit communication tables)."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from dlnetbench_b200.parser import dp_metrics_df, fsdp_metrics_df, parse_sections, summarize  # noqa: E402


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("files", nargs="+")
    ap.add_argument("--csv", default=None, help="prefix for the per-rank tables (dp / fsdp sections)")
    a = ap.parse_args()
    for f in a.files:
        text = sys.stdin.read() if f == "-" else open(f).read()
        for sid, sec in parse_sections(text).items():
            print(json.dumps({"file": f, **summarize(sid, sec)}))
            if a.csv and sid == "dp":
                dp_metrics_df(sec).to_csv(f"{a.csv}_dp.csv", index=False)
            if a.csv and sid == "fsdp":
                rt, comm = fsdp_metrics_df(sec)
                rt.to_csv(f"{a.csv}_fsdp_runtime.csv", index=False)
                comm.to_csv(f"{a.csv}_fsdp_comm.csv", index=False)


if __name__ == "__main__":
    main()
