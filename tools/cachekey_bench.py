#!/usr/bin/env python3
"""Runs bench.py's cache-key leg alone (GPU box)."""
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import bench  # noqa: E402
from k8s_llm_scheduler_b200 import _lib  # noqa: E402

if __name__ == "__main__":
    print(json.dumps(bench.run_cachekey_extra(torch, _lib.load(), torch.device("cuda:0"))))
