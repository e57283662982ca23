"""Debug aid: capture the fused-stack timeline of CTA 0 (needs a build with TTSB200_NVCC_DEFS=-DTTSB_STACK_TRACE).

    python tests/cuda/trace_run.py [debug bits ...]      # one trace file per value under gpurun_out/

The trace file and the knock-out bits are read once, at plan creation, so every run builds its own model."""
import os, sys, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import bench
mel = torch.randn(16, 80, 862, device="cuda")
bench.make_model("bf16").to("cuda").eval().inference(mel)      # warm: attributes, allocator
torch.cuda.synchronize()
for dbg in sys.argv[1:] or ["0"]:
    os.environ["TTSB200_STACK_DEBUG"] = dbg.split(":")[0]
    os.environ["TTSB200_CONV_DEBUG"] = dbg.split(":")[1] if ":" in dbg else "0"
    os.environ["TTSB200_STACK_TRACE"] = os.path.join(bench.ROOT, "gpurun_out", f"stack_trace_{dbg}.txt")
    m = bench.make_model("bf16").to("cuda").eval()
    m.inference(mel); torch.cuda.synchronize()
    del m
