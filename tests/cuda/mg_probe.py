import os, sys, time, torch
sys.path.insert(0, "/root/repo")
import bench
dt = sys.argv[1]
m = bench.make_model(dt, "MelganGenerator").to("cuda").eval(); m.freeze()
mel = torch.randn(1, 80, 200, device="cuda")
for n in (1, 2, 5, 50, 300):
    t0 = time.time()
    for _ in range(n): m.inference(mel)
    torch.cuda.synchronize()
    print(dt, n, "ok", round((time.time() - t0) * 1e3 / n, 4), "ms per call", flush=True)
