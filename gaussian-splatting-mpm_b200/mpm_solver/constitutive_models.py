"""Constitutive models of the reference's mpm_solver/constitutive_models.py.

The arithmetic is implemented once, in CUDA: csrc/constitutive.cuh (`constitutive_update`), fused
into the g2p kernel's epilogue.  This module only documents the dispatch so that readers of the
reference find the same names.

    material id  return mapping (reference lines)                         Kirchhoff stress
    0 jelly      none                                                      FCR :11-20 (0 under strict Q1, utils.py:37)
    1 metal      von_mises_return_mapping :63-103 (+ hardening)            StVK-Hencky :24-38
    2 sand       sand_return_mapping :106-140                              Drucker-Prager :42-58
    3 foam       viscoplasticity_return_mapping_with_StVK :217-259         StVK-Hencky
    4 fluid      fluid_return_mapping :143-213 (never dispatched in ref)   StVK-Hencky   [unpinned]
    5 plasticine von Mises with softening                  StVK-Hencky   [unpinned]
"""
MODELS = {
    0: ("jelly", None, "kirchoff_stress_FCR"),
    1: ("metal", "von_mises_return_mapping", "kirchoff_stress_StVK"),
    2: ("sand", "sand_return_mapping", "kirchoff_stress_Drucker_Prager"),
    3: ("foam", "viscoplasticity_return_mapping_with_StVK", "kirchoff_stress_StVK"),
    4: ("fluid", "fluid_return_mapping", "kirchoff_stress_StVK"),
    5: ("plasticine", "von_mises_return_mapping_with_softening", "kirchoff_stress_StVK"),
}
