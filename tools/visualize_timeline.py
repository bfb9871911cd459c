#!/usr/bin/env python3
"""Per-step worker-arrival timeline + "average number of gradients received per step"
(reference MPI_code/src/python/visualize_timeline.py:50-62). Rows are ``time_ms step is_master``."""
import sys


def summarize(path):
    rows = [l.split() for l in open(path) if l.strip()]
    steps = {}
    for t, s, is_master in rows:
        d = steps.setdefault(int(s), {"start": None, "arrivals": []})
        if int(is_master):
            d["start"] = float(t)
        else:
            d["arrivals"].append(float(t))
    n = max(1, len(steps))
    avg = sum(len(d["arrivals"]) for d in steps.values()) / n
    return steps, avg


def main(path, png="timeline.png"):
    steps, avg = summarize(path)
    print("Average number of gradients received per step: %.3f" % avg)
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
        for s, d in sorted(steps.items()):
            if d["start"] is None:
                continue
            plt.scatter([a - d["start"] for a in d["arrivals"]], [s] * len(d["arrivals"]), s=4)
        plt.xlabel("ms since step start")
        plt.ylabel("step")
        plt.savefig(png)
        print("wrote", png)
    except ImportError:
        for s, d in sorted(steps.items())[:10]:
            if d["start"] is not None and d["arrivals"]:
                print(f"step {s}: {len(d['arrivals'])} arrivals, last at +{max(d['arrivals']) - d['start']:.2f} ms")
    return avg


if __name__ == "__main__":
    main(*sys.argv[1:])
