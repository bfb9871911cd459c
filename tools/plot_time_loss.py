#!/usr/bin/env python3
"""Loss-vs-time overlay of every ``time_loss_out_*`` file (reference MPI_code/src/python/plot_time_loss.py).
Rows are ``step time_ms loss err``. Without matplotlib a text table is printed instead of a PNG."""
import glob
import os
import sys


def load(path):
    rows = [tuple(float(x) for x in l.split()) for l in open(path) if l.strip()]
    return [r for r in rows if len(r) == 4]


def main(out_dir="outfiles", png="time_loss.png"):
    files = sorted(glob.glob(os.path.join(out_dir, "time_loss_out_*")))
    series = {os.path.basename(f)[len("time_loss_out_"):]: load(f) for f in files}
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
        for name, rows in series.items():
            plt.plot([r[1] for r in rows], [r[2] for r in rows], label=name)
        plt.xlabel("time (ms)")
        plt.ylabel("loss")
        plt.legend()
        plt.savefig(png)
        print("wrote", png)
    except ImportError:
        for name, rows in series.items():
            if rows:
                print(f"{name}: {len(rows)} points, loss {rows[0][2]:.4f} -> {rows[-1][2]:.4f}, err {rows[0][3]:.3f} -> {rows[-1][3]:.3f}, "
                      f"{rows[-1][1]:.1f} ms")
    return series


if __name__ == "__main__":
    main(*sys.argv[1:])
