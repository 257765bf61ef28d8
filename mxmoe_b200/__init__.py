"""mxmoe_b200 — B200-native mixed-precision MoE GroupGEMM path (drop-in for the path run_mxmoe_gg.py drives).

Python here is host plumbing only (device memory, streams, torch.distributed); the work is done by the
hand-written sm_100a kernels in mxmoe_b200/csrc behind the C-ABI of include/mxmoe_b200.h.
"""
__version__ = "0.1.0"
