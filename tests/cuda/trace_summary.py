"""Summarise a TTSB200_STACK_TRACE file: mean duration between consecutive tags per role."""
import collections, sys
for path in sys.argv[1:]:
    ev = collections.defaultdict(list)
    for l in open(path):
        cp, r, tag, clk = map(int, l.split())
        ev[(cp, r)].append((clk, tag))
    print("==", path)
    for (cp, r), e in sorted(ev.items()):
        e.sort()
        d = collections.defaultdict(list)
        for (c0, t0), (c1, t1) in zip(e, e[1:]):
            d[(t0, t1)].append(c1 - c0)
        print(f"cp={cp} role={r} span={e[-1][0]-e[0][0]} " + " ".join(f"{a}>{b}:{sum(v)//len(v)}x{len(v)}" for (a, b), v in sorted(d.items()) if len(v) > 2))
