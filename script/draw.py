#!/usr/bin/env python
"""Plot column y against column x of a whitespace-separated log (the reference's script/draw.py); falls back to a
text table when matplotlib is unavailable.   usage: draw.py LOG XCOL YCOL [out.png]"""
import sys


def main():
    path, xc, yc = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    xs, ys = [], []
    for line in open(path):
        f = line.split()
        try:
            xs.append(float(f[xc])); ys.append(float(f[yc]))
        except (IndexError, ValueError):
            continue
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
        plt.plot(xs, ys)
        plt.xlabel(f"col {xc}"); plt.ylabel(f"col {yc}")
        out = sys.argv[4] if len(sys.argv) > 4 else path + ".png"
        plt.savefig(out)
        print("wrote", out)
    except ImportError:
        for x, y in zip(xs, ys):
            print(f"{x}\t{y}")


if __name__ == "__main__":
    main()
