import os, sys, torch
sys.path.insert(0, "/root/repo")
import bench
m = bench.make_model("bf16").to("cuda").eval()
mel = torch.randn(16, 80, 862, device="cuda")
m.inference(mel); torch.cuda.synchronize()
os.environ["TTSB200_STACK_TRACE"] = "/root/repo/gpurun_out/stack_trace.txt"
m.inference(mel); torch.cuda.synchronize()
