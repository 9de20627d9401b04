"""circuitsimulator_b200 — B200-native batched Newton/MNA hot path of ZyuRao/CircuitSimulator.

`capi` binds the C ABI (include/csim_b200.h) of libcsim_b200.so, whose kernels live in csrc/.
"""
__version__ = "0.1.0"
