"""Smallest program that launches every kernel of one cfg2 inference a few times (for ncu):
    python tests/cuda/ncu_run.py [n_inferences] [dtype]"""
import os, sys, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3
m = bench.make_model(sys.argv[2] if len(sys.argv) > 2 else "bf16").to("cuda").eval()
mel = torch.randn(16, 80, 862, device="cuda")
for _ in range(n):
    m.inference(mel)
torch.cuda.synchronize()
print("ok")
