"""Writes tests/golden/udon_ply_golden.json from the reference's only real 3DGS cloud,
models/udon/point_cloud/iteration_30000/point_cloud4.ply (N = 2,153): per-array shapes, float64 sums and the first
and last record, as parsed by an independent byte-level reader (struct.unpack), for tests/test_ply_io.py.
    python tests/golden/make_ply_golden.py /root/reference"""
import json
import os
import struct
import sys

import numpy as np


def independent_read(path):
    raw = open(path, "rb").read()
    end = raw.index(b"end_header\n") + len(b"end_header\n")
    head = raw[:end].decode("ascii").split("\n")
    n = [int(l.split()[2]) for l in head if l.startswith("element vertex")][0]
    names = [l.split()[2] for l in head if l.startswith("property float")]
    w = len(names)
    vals = struct.unpack("<%df" % (n * w), raw[end:end + 4 * n * w])
    return names, np.array(vals, np.float32).reshape(n, w)


def main():
    ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
    path = os.path.join(ref, "models/udon/point_cloud/iteration_30000/point_cloud4.ply")
    names, a = independent_read(path)
    col = {n: a[:, i] for i, n in enumerate(names)}
    out = {"n": int(a.shape[0]), "n_properties": len(names), "file_bytes": os.path.getsize(path),
           "sum": {k: float(np.sum(col[k].astype(np.float64))) for k in ("x", "y", "z", "f_dc_0", "f_rest_0", "f_rest_44", "opacity",
                                                                        "scale_0", "scale_2", "rot_0", "rot_3")},
           "first": [float(x) for x in a[0]], "last": [float(x) for x in a[-1]], "names": names}
    dst = os.path.join(os.path.dirname(os.path.abspath(__file__)), "udon_ply_golden.json")
    json.dump(out, open(dst, "w"))
    print("wrote", dst, out["n"], out["n_properties"])


if __name__ == "__main__":
    main()
