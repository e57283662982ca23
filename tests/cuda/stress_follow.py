"""Stress for the per-segment kernel hand-over: thousands of inferences enqueued back to back (no host sync), every
result compared bit for bit with a lone call's.  python tests/cuda/stress_follow.py [n]"""
import os, sys, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
cases = [("MultibandMelganGenerator", {}, 3, 37), ("MelganGenerator", {}, 1, 200),
         ("FullbandMelganGenerator", {"upsample_factors": (5, 5, 4, 3)}, 1, 50), ("MultibandMelganGenerator", {}, 16, 862)]
if os.environ.get("STRESS_ONLY_CFG2"):
    cases = cases[3:]
bad = 0
for cls, kw, B, T in cases:
    m = bench.make_model("bf16", cls, **kw).to("cuda").eval(); m.freeze()
    g = torch.Generator().manual_seed(1)
    mels = [torch.randn(B, 80, T, generator=g).cuda() for _ in range(4)]
    lone = []
    for x in mels:
        lone.append(m.inference(x).clone()); torch.cuda.synchronize()
    reps = n if B * T < 2000 else max(40, n // 20)
    outs = []
    for i in range(reps):
        outs.append(m.inference(mels[i % 4]))
        if os.environ.get("STRESS_SYNC"):
            torch.cuda.synchronize()
        if len(outs) == 40:
            torch.cuda.synchronize()
            for j, y in enumerate(outs):
                k = (i - 39 + j) % 4
                if not torch.equal(y, lone[k]):
                    bad += 1
                    print("MISMATCH", cls, B, T, "iteration", i - 39 + j, float((y - lone[k]).abs().max()))
            outs = []
    torch.cuda.synchronize()
    print(cls, B, T, reps, "inferences checked, mismatches so far", bad, flush=True)
print("STRESS", "FAILED" if bad else "PASSED")
sys.exit(1 if bad else 0)
