"""Timing + full-size parity of the perspective-frustum querier alone (not collected by pytest; run on the GPU box):

    python tests/run_frustum_timing.py > gpurun_out/frustum_timing.json

Prints the ``frustum_querier_wcoord0`` object of the bench line (bench.py: frustum_bench) without the rest of the bench."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402

if __name__ == "__main__":
    print(json.dumps(bench.frustum_bench(torch.device("cuda", 0))))
