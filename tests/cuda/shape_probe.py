"""Per-kernel times and back-to-back step time of one inference shape (debug aid):
    python tests/cuda/shape_probe.py B T [steps] [generator class]"""
import os, sys, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import bench
from tts_b200.engine import Profile
B, T = int(sys.argv[1]), int(sys.argv[2])
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 50
m = bench.make_model("bf16", *(sys.argv[4:5])).to("cuda").eval(); m.freeze()
mel = torch.randn(B, 80, T, device="cuda")
for _ in range(5): m.inference(mel)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(steps): m.inference(mel)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / steps
prof = Profile()
m.inference_with_taps(mel, [], profile=prof) if False else m._run(mel, m.inference_padding, m._SKIP_PQMF, profile=prof)
torch.cuda.synchronize()
recs = prof.read()
print(f"B={B} T={T}: {ms:.4f} ms per step back to back; profiled sum {sum(r['ms'] for r in recs):.4f} ms")
for r in recs: print(f"   {r['name']}@{r['layer']}: {r['ms']*1e3:.1f} us")
