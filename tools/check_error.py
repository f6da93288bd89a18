#!/usr/bin/env python
"""Headless restatement of the reference's frame-file comparison (utils/check_error.py:19-38).

The reference loads two frame files (`index,x,y,z,mass` per body per frame), takes
|file1 - file2| row by row, picks body 0's row of every frame, plots its |dx| against the
frame number and fits a straight line through frames 30..99 (np.polyfit, degree 1).  There
is no display here, so this tool prints the numbers the plot would show (and can write them
as CSV) instead of drawing them; the arithmetic is the same.

    python tools/check_error.py a.csv b.csv [--body 0] [--fit 30 100] [--series out.csv]

Exit status 0; `--max-error E` turns it into a check (exit 1 when any |d| of any body in any
frame exceeds E).
"""
import argparse
import json
import sys

import numpy as np


def load_frames(path):
    """(frames, bodies, 5) array of a frame file; n_bodies = max(index) + 1 like the reference
    (utils/check_error.py:22)."""
    data = np.loadtxt(path, delimiter=",", ndmin=2)
    if data.size == 0:
        return np.zeros((0, 0, 5))
    if data.shape[1] != 5:
        raise ValueError("%s: %d columns, expected index,x,y,z,mass" % (path, data.shape[1]))
    n = int(data[:, 0].max()) + 1
    if data.shape[0] % n:
        raise ValueError("%s: %d rows is not a whole number of %d-body frames" % (path, data.shape[0], n))
    return data.reshape(-1, n, 5)


def compare(a, b, body=0, fit=(30, 100)):
    """The reference's numbers: per-frame |d| of `body` (x, y, z), the least-squares line
    through |dx| over frames fit[0]..fit[1]-1 (utils/check_error.py:25-37), plus the maximum
    over every body and frame."""
    if a.shape != b.shape:
        raise ValueError("frame files differ in shape: %s vs %s" % (a.shape, b.shape))
    diff = np.abs(a - b)                         # :25
    series = diff[:, body, 1:4]                  # :27-29 rows of body 0, columns x y z
    out = {"frames": int(a.shape[0]), "bodies": int(a.shape[1]), "body": int(body),
           "max_abs_error": float(diff[:, :, 1:4].max()) if diff.size else 0.0,
           "body_final_error": [float(v) for v in series[-1]] if len(series) else []}
    lo, hi = fit
    hi = min(hi, len(series))
    if hi - lo >= 2:
        x = np.arange(lo, hi)
        m, q = np.polyfit(x, series[lo:hi, 0], 1)  # :33-35
        out["fit"] = {"from": int(lo), "to": int(hi), "slope": float(m), "intercept": float(q)}
    else:
        out["fit"] = None
    return out, series


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("file1")
    ap.add_argument("file2")
    ap.add_argument("--body", type=int, default=0)
    ap.add_argument("--fit", type=int, nargs=2, default=(30, 100), metavar=("FROM", "TO"))
    ap.add_argument("--series", help="write frame,|dx|,|dy|,|dz| of the body to this CSV")
    ap.add_argument("--max-error", type=float, default=None)
    args = ap.parse_args(argv)
    a, b = load_frames(args.file1), load_frames(args.file2)
    out, series = compare(a, b, args.body, tuple(args.fit))
    if args.series:
        with open(args.series, "w") as f:
            for k, row in enumerate(series):
                f.write("%d,%.17g,%.17g,%.17g\n" % (k, row[0], row[1], row[2]))
    print(json.dumps(out))
    if args.max_error is not None and out["max_abs_error"] > args.max_error:
        return 1
    return 0


if __name__ == "__main__":
    sys.exit(main())
