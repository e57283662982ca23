"""Debug aid: capture the fused-stack timeline of CTA 0 (needs a build with TTSB200_NVCC_DEFS=-DTTSB_STACK_TRACE)."""
import os, sys, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import bench
m = bench.make_model("bf16").to("cuda").eval()
mel = torch.randn(16, 80, 862, device="cuda")
m.inference(mel); torch.cuda.synchronize()
for dbg in sys.argv[1:] or ["0"]:
    os.environ["TTSB200_STACK_DEBUG"] = dbg
    os.environ["TTSB200_STACK_TRACE"] = f"/root/repo/gpurun_out/stack_trace_{dbg}.txt"
    m.inference(mel); torch.cuda.synchronize()
