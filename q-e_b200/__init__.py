"""qe_b200 - B200-native H|psi> path of PWscf (host-side mirror of the C ABI).

The product is q-e_b200/libqeb200.so (hand-written sm_100a CUDA behind the C ABI
of include/qeb200.h). This package only loads it and mirrors the reference's
entry points (h_psi, vloc_psi_k/gamma, invfft/fwfft, calbec, add_vuspsi) for
Python callers, tests and the benchmark. There is no CPU fallback: creating a
context without a CUDA device raises.
"""
from .hpsi import (COMM_BGRP, COMM_RG, HPsi, QebError, allgather_columns, divide_all, h_psi_multi, lib_path,  # noqa: F401
                   load_library)
from . import descriptor  # noqa: F401,E402
