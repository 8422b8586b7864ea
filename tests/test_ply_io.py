"""3DGS PLY reader/writer (utils/ply_io.py) against gaussian_model.py:191-256's layout: a round trip
on synthetic Gaussians, the reference's one real cloud (models/udon/.../point_cloud4.ply, N = 2,153) against the golden
numbers an independent byte-level reader produced (tests/golden/make_ply_golden.py), and the xyz-only frame dump."""
import json
import os

import numpy as np
import pytest

from utils import ply_io

HERE = os.path.dirname(os.path.abspath(__file__))
UDON = "/root/reference/models/udon/point_cloud/iteration_30000/point_cloud4.ply"


def _synthetic(P=1234, seed=4):
    rng = np.random.default_rng(seed)
    f = lambda *s: rng.standard_normal(s).astype(np.float32)
    return {"xyz": f(P, 3), "features_dc": f(P, 1, 3), "features_rest": f(P, 15, 3), "opacity": f(P, 1), "scaling": f(P, 3),
            "rotation": f(P, 4)}


def test_round_trip_and_layout(tmp_path):
    g = _synthetic()
    path = str(tmp_path / "pc" / "point_cloud.ply")
    ply_io.write_gaussian_ply(path, g)
    P = g["xyz"].shape[0]
    assert os.path.getsize(path) == len(open(path, "rb").read().split(b"end_header\n")[0]) + len(b"end_header\n") + P * 62 * 4
    v = ply_io.read_ply_vertices(path)
    assert list(v.dtype.names) == ply_io.attribute_names() and v.shape == (P,)
    assert not np.any(v["nx"]) and not np.any(v["nz"])                                   # normals are zeros (:195)
    # f_dc / f_rest are stored channel-major: f_rest_k = features_rest[:, k % 15, k // 15] (:196-197 transpose(1,2).flatten)
    assert np.array_equal(v["f_rest_16"], g["features_rest"][:, 1, 1]) and np.array_equal(v["f_dc_2"], g["features_dc"][:, 0, 2])
    h = ply_io.read_gaussian_ply(path)
    for k in g:
        assert h[k].dtype == np.float32 and np.array_equal(h[k], g[k]), k
    with pytest.raises(AssertionError):
        ply_io.read_gaussian_ply(path, max_sh_degree=2)                                   # gaussian_model.py:231
    # empty cloud
    e = {k: a[:0] for k, a in g.items()}
    ply_io.write_gaussian_ply(str(tmp_path / "empty.ply"), e)
    assert ply_io.read_gaussian_ply(str(tmp_path / "empty.ply"))["xyz"].shape == (0, 3)


def test_deformed_dump_changes_only_positions(tmp_path):
    g = _synthetic(500)
    src, dst = str(tmp_path / "a.ply"), str(tmp_path / "out" / "b.ply")
    ply_io.write_gaussian_ply(src, g)
    mask = np.zeros(500, bool); mask[100:350] = True
    new = np.random.default_rng(1).standard_normal((250, 3)).astype(np.float32)
    ply_io.write_deformed_ply(src, dst, mask, new)
    h = ply_io.read_gaussian_ply(dst)
    assert np.array_equal(h["xyz"][mask], new) and np.array_equal(h["xyz"][~mask], g["xyz"][~mask])
    for k in ("features_dc", "features_rest", "opacity", "scaling", "rotation"):
        assert np.array_equal(h[k], g[k])
    with pytest.raises(ValueError):
        ply_io.write_deformed_ply(src, dst, mask, new[:10])


def test_rejects_what_it_cannot_read(tmp_path):
    p = tmp_path / "ascii.ply"
    p.write_bytes(b"ply\nformat ascii 1.0\nelement vertex 1\nproperty float x\nend_header\n0.0\n")
    with pytest.raises(ValueError):
        ply_io.read_ply_vertices(str(p))
    q = tmp_path / "junk.ply"
    q.write_bytes(b"not a ply\nend_header\n")
    with pytest.raises(ValueError):
        ply_io.read_ply_vertices(str(q))


@pytest.mark.skipif(not os.path.exists(UDON), reason="the reference tree is not mounted here")
def test_real_reference_cloud_matches_golden():
    gold = json.load(open(os.path.join(HERE, "golden", "udon_ply_golden.json")))
    assert os.path.getsize(UDON) == gold["file_bytes"]
    v = ply_io.read_ply_vertices(UDON)
    assert v.shape[0] == gold["n"] == 2153 and list(v.dtype.names) == gold["names"] and len(v.dtype.names) == 62
    for k, s in gold["sum"].items():
        assert float(np.sum(v[k].astype(np.float64))) == s, k
    assert [float(v[0][n]) for n in v.dtype.names] == gold["first"]
    assert [float(v[-1][n]) for n in v.dtype.names] == gold["last"]
    g = ply_io.read_gaussian_ply(UDON)
    assert g["features_rest"].shape == (2153, 15, 3) and g["rotation"].shape == (2153, 4)
    assert float(g["features_rest"][0, 1, 2]) == gold["first"][gold["names"].index("f_rest_31")]     # 2 * 15 + 1
