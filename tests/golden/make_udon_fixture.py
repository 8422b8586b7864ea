"""Writes tests/golden/udon_case.npz + udon_case.json from the reference's only real trained-Gaussian cloud and its
config, for the GPU parity test of the full configs/udon.json boundary-condition list (tests/test_gpu_configs.py):

  * models/udon/point_cloud/iteration_30000/point_cloud4.ply (N = 2,153): the three arrays the simulator consumes
    (xyz, log-scaling, rotation quaternion; gaussian_model.py:215-256), float32, read with utils/ply_io.py (itself
    pinned against an independent byte-level reader by tests/golden/make_ply_golden.py);
  * configs/udon.json: the "mpm" block (material, grid, gravity, the eleven boundary conditions incl.
    additional_params and modify_material), copied as data.

/root/reference does not exist on the GPU box, hence the committed fixture.
    python tests/golden/make_udon_fixture.py /root/reference"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "gaussian-splatting-mpm_b200"))

from utils import ply_io  # noqa: E402


def main():
    ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
    g = ply_io.read_gaussian_ply(os.path.join(ref, "models/udon/point_cloud/iteration_30000/point_cloud4.ply"))
    np.savez_compressed(os.path.join(HERE, "udon_case.npz"), xyz=g["xyz"], scaling=g["scaling"], rotation=g["rotation"])
    with open(os.path.join(ref, "configs/udon.json")) as f:
        cfg = json.load(f)
    with open(os.path.join(HERE, "udon_case.json"), "w") as f:
        json.dump({"source": "configs/udon.json", "mpm": cfg["mpm"]}, f, indent=1)
    print("wrote udon_case.npz / udon_case.json:", g["xyz"].shape[0], "Gaussians,", len(cfg["mpm"]["boundary_conditions"]), "boundary conditions")


if __name__ == "__main__":
    main()
