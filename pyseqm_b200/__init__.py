"""pyseqm_b200 — B200-native (sm_100a) implementation of PYSEQM's batched SCF hot path.

Mirrors the reference's `seqm` package layout for that path (`basics`, `seqm_functions.*`,
`MolecularDynamics`, `XLBOMD`); every numerical op runs in hand-written CUDA kernels reached through
the C-ABI library `libseqm_b200.so` (see include/seqm_b200.h).  There is no CPU fallback: using an op
without the library, or with CPU tensors, raises.
"""
__version__ = "0.1.0"
