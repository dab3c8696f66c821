"""Self-written CPU oracle for the MSRCR and Zero-DCE hot paths.

THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE, AND NOT CODE FROM
CILAB-IITM/Low_Light_Image_Enhancement.  The mounted reference contains only a
two-line README (/root/reference/README.md:1-2) -- no code, no tests, no golden
vectors -- so there is nothing of the reference's to restate.  What "correct"
means is pinned by BASELINE.json#north_star clauses (a) and (b) and by the
normative spec in SURVEY.md section 9, which this package implements in NumPy.

PARITY UNPINNED BY THE REFERENCE: the reference ships no fixture or
known-answer test for this path.  The oracle is pinned instead by
  * closed-form known-answer tests (tests/test_oracle_kat.py),
  * agreement with independent libraries (SciPy ndimage, OpenCV, torch CPU;
    tests/test_oracle_crosscheck.py),
  * committed golden vectors generated from this oracle (tests/golden/).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this package.  The product path
(low_light_image_enhancement_b200) never does.
"""
from .msrcr import (MSRCR_DEFAULTS, gaussian_kernel_1d, blur_edge_clamp,
                    msrcr, msrcr_window, msrcr_unclipped)
from .zero_dce import (DCE_LAYERS, DCE_NUM_PARAMS, init_weights, pack_weights,
                       unpack_weights, dce_net, apply_curves, zero_dce)
from .synth import synth_lowlight, hard_cases, HARD_CASE_CLASS
from .u8 import from_u8, to_u8, msrcr_u8, zero_dce_u8, from_u16, to_u16, msrcr_u16, zero_dce_u16

__all__ = [
    "MSRCR_DEFAULTS", "gaussian_kernel_1d", "blur_edge_clamp", "msrcr",
    "msrcr_window", "msrcr_unclipped", "DCE_LAYERS", "DCE_NUM_PARAMS", "init_weights",
    "pack_weights", "unpack_weights", "dce_net", "apply_curves", "zero_dce",
    "synth_lowlight", "hard_cases", "HARD_CASE_CLASS", "from_u8", "to_u8", "msrcr_u8", "zero_dce_u8", "from_u16", "to_u16", "msrcr_u16", "zero_dce_u16",
]
