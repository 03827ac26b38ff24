"""codesign-kernels_b200 -- host-side mirror of the reference's Fortran interface for the
`biharmonic_wk` hot path (atmosphere/biharmonic_wk_kernel.F90, cited as F90:<line>), on top of
the C ABI in include/bwk_b200.h (libbwk_b200.so, hand-written CUDA for sm_100a).

The directory name has a hyphen, so import it through the shim at the repo root:

    import codesign_kernels_b200 as bwk

Names follow the reference: `allocate_data`, `initialize_data`, `biharmonic_wk_scalar`,
`save_qtens`, `compute_l2norm` (module kernel_utils F90:7-93, drivers F90:186,317,525).
Arrays are numpy arrays in Fortran (column-major) order with the reference's shapes, so the
bytes handed to the library are exactly what a Fortran caller would pass.

There is no CPU fallback: if libbwk_b200.so is missing, or no B200 is visible, calls raise.
"""
from .binding import (  # noqa: F401
    BwkError,
    ELEMENT_DTYPE,
    DERIVATIVE_DTYPE,
    NP,
    KernelUtils,
    lib,
    lib_path,
    declared_symbols,
    cubed_sphere_gdof,
    partition,
    Plan,
    NOALLOC_MAX_NELEMD,
    biharmonic_wk_scalar_noalloc,
)
