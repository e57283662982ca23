"""Which launch differs when a back-to-back inference does not equal a lone call: every iteration taps the tensors
the fused kernels write themselves (fusion stays on) and reports the first tensor that differs and where.
python tests/cuda/stress_locate.py [n]"""
import os, sys, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 400
B, T = 16, 862
m = bench.make_model("bf16", "MultibandMelganGenerator").to("cuda").eval(); m.freeze()
idx = [i for i, l in enumerate(m.layers) if isinstance(l, torch.nn.ConvTranspose1d)]
layers = [1] + idx + [len(m.layers) - 1]
print("tapped layers", layers, flush=True)
g = torch.Generator().manual_seed(1)
mels = [torch.randn(B, 80, T, generator=g).cuda() for _ in range(4)]
lone = []
for x in mels:
    y, t = m._run(x, 2, False, tap_layers=layers)
    torch.cuda.synchronize()
    lone.append((y.clone(), {k: v.clone() for k, v in t.items()}))
bad = 0
for i in range(n):
    k = i % 4
    y, t = m._run(mels[k], 2, False, tap_layers=layers)
    if os.environ.get("LOCATE_PLAIN_TOO"):
        m.inference(mels[(k + 1) % 4])
    torch.cuda.synchronize()
    ry, rt = lone[k]
    if torch.equal(y, ry) and all(torch.equal(t[l], rt[l]) for l in layers):
        continue
    bad += 1
    for l in layers + ["out"]:
        a, b = (y, ry) if l == "out" else (t[l], rt[l])
        if torch.equal(a, b):
            continue
        d = (a != b).nonzero()
        print(f"iteration {i}: first differing tensor: layer {l} shape {tuple(a.shape)}: {d.shape[0]} elements, "
              f"utterances {sorted(set(d[:, 0].tolist()))}, channels {int(d[:, 1].min())}..{int(d[:, 1].max())}, "
              f"time {int(d[:, 2].min())}..{int(d[:, 2].max())}, max abs {float((a - b).abs().max()):.3g}", flush=True)
        if d.shape[0] < 4000:
            ts = sorted(set(d[:, 2].tolist()))
            print("   times:", ts[:40], "..." if len(ts) > 40 else "", " channels:", sorted(set(d[:, 1].tolist()))[:48])
        break
print(n, "inferences, mismatching", bad)
