#!/usr/bin/env python3
"""Runs bench.py's ingest leg alone (GPU box): device time + HBM roofline of the ingest kernel,
host packing, whole call, literal-mirror CPU baseline.  Prints one JSON object."""
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import bench  # noqa: E402
from k8s_llm_scheduler_b200 import _lib  # noqa: E402

if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    print(json.dumps(bench.run_ingest_extra(torch, _lib.load(), torch.device("cuda:0"), n_pods=n)))
