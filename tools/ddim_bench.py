"""DDIM-50 throughput only (the bench.py ddim50 leg).  usage: python tools/ddim_bench.py [B]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
dev = torch.device("cuda:0")
model = bench.build_model(dev)
print(B, bench.ddim50(model, dev, B, False, 1))
