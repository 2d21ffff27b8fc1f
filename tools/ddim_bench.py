"""DDIM-50 throughput only (the `ddim50` leg of bench.py).  usage: python tools/ddim_bench.py [batch]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
dev = torch.device("cuda:0")
print(json.dumps(bench.ddim50(bench.build_model(dev), dev, B, False, 1)))
