#!/usr/bin/env python3
"""stage breakdown of BASELINE config 4 (3840x2160, 8000 features, 12 levels, batch 16) and config 3, device-resident"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
print(json.dumps({"c4": bench.bench_config(torch, dev, 3840, 2160, 8000, 12, 16, steps=5),
                  "c3": bench.bench_config(torch, dev, 752, 480, 1200, 8, 256, steps=5)}))
