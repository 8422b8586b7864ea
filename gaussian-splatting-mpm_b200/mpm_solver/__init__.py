"""B200-native drop-in for the reference's `mpm_solver` package (mpm_solver/solver.py,
model.py, boundary_conditions.py, collider.py, utils.py, constitutive_models.py).

Same import paths and Python surface as the reference; every Taichi kernel is replaced by
hand-written sm_100a CUDA in lib/libmpmk.so (ABI: include/mpmk.h).  No Taichi, no CPU fallback.
"""
