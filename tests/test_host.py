"""CPU-side checks: the C-ABI library loads and exports every symbol include/mpmk.h declares, the
ctypes structs match the header, and the host logic (registries, activity windows, ti shim)."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "mpmk.h")


def _declared():
    src = open(HEADER).read()
    return sorted(set(re.findall(r"MPMK_API\s+[\w\s\*]+?\b(mpmk_\w+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from mpm_solver import _native
    L = _native.lib()
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(L, n), n
    assert L.mpmk_version() == 1
    # and nothing the binding uses is missing from the header
    bound = {n for n in dir(_native.Context) if n.startswith("mpmk_")}
    assert bound <= set(names)


def test_ctypes_structs_match_header_layout(tmp_path):
    """compile a tiny C program against the header and compare sizeof/offsetof with ctypes"""
    from mpm_solver import _native
    prog = tmp_path / "sz.c"
    prog.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "mpmk.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu\\n",'
                    'sizeof(mpmk_particle_set),sizeof(mpmk_bindings),sizeof(mpmk_params),sizeof(mpmk_grid_op),sizeof(mpmk_impulse),'
                    'sizeof(mpmk_substep_ops),sizeof(mpmk_counters),offsetof(mpmk_substep_ops,n_impulses),offsetof(mpmk_bindings,keys));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["/usr/bin/gcc", "-I", os.path.join(ROOT, "include"), str(prog), "-o", str(exe)])
    got = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    want = [C.sizeof(_native.ParticleSet), C.sizeof(_native.Bindings), C.sizeof(_native.Params), C.sizeof(_native.GridOp),
            C.sizeof(_native.Impulse), C.sizeof(_native.SubstepOps), C.sizeof(_native.Counters),
            _native.SubstepOps.n_impulses.offset, _native.Bindings.keys.offset]
    assert got == want


def test_no_cpu_fallback():
    import torch
    from mpm_solver.solver import MPM_Simulator, ti
    from harness import synthetic
    with pytest.raises(RuntimeError):
        ti.init(arch=ti.cpu)
    if not torch.cuda.is_available():
        with pytest.raises(Exception):
            MPM_Simulator(torch.zeros(4, 3), torch.zeros(4, 6), torch.ones(4), synthetic.sim_args())


def test_bc_registry_and_activity_windows():
    """boundary_conditions.py:12-31,97-117 and the time accumulation of solver.py:52"""
    from mpm_solver.boundary_conditions import (BasicBC, ImpulseBC, boundaryConditionTypeCallBacks, preprocess_bc,
                                                postprocess_bc, init_bc)
    from harness import synthetic
    args = synthetic.sim_args(substep_dt=1e-4)
    bc = BasicBC(10, {"id": 3, "type": "fixed_cube", "center": [0, 1, 1], "size": [0.1, 2, 2], "start_time": 0, "num_dt": 1}, args)
    t, active = 0.0, []
    for s in range(4):
        active.append(bc.isActive(t)); t += 1e-4
    assert active == [True, False, False, False]          # num_dt=1 acts on substep 0 only
    imp = ImpulseBC(10, {"id": 2, "type": "impulse", "center": [1, .65, 1.22], "size": [1.4, .18, .27], "force": [-.8, 0, 0],
                         "start_time": 0.8, "num_dt": 10}, args)
    t, n_on = 0.0, 0
    for s in range(8100):
        n_on += imp.isActive(t); t += 1e-4
    assert n_on in (9, 10, 11)                            # window evaluated with accumulated doubles, as the reference
    assert imp.type in preprocess_bc and bc.type in postprocess_bc and "modify_material" in init_bc
    assert set(boundaryConditionTypeCallBacks) == {"fixed_cube", "impulse", "sticky_ground", "additional_params", "modify_material"}
    op = bc.as_grid_op()
    assert op.type == 0 and list(op.b) == pytest.approx([0.1, 2, 2])
    with pytest.raises(KeyError):
        boundaryConditionTypeCallBacks["nope"]


def test_material_table_and_unknown_material():
    from mpm_solver.utils import material_types
    assert material_types["jelly"] == 0 and material_types["metal"] == 1 and material_types["sand"] == 2 and material_types["foam"] == 3
    assert material_types["fluid"] == 4 and material_types["plasticine"] == 5


def test_synthetic_workload_sizes():
    from harness import synthetic
    c = synthetic.make_cloud(64, 2.0, (22, 22, 22), (20, 20, 20))
    assert c["xyz"].shape == (64000, 3) and c["cov"].shape == (64000, 6)
    n_grid, ext, lo, cn, dt = synthetic.WORKLOADS["S-4M"]
    assert cn[0] * cn[1] * cn[2] * 8 == 4096000
    n_grid, ext, lo, cn, dt = synthetic.WORKLOADS["S-32M-bar"]
    assert cn[0] * cn[1] * cn[2] * 8 == 33030144
