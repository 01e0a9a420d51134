"""uda_b200 — a B200-native MapReduce shuffle-and-merge engine with the capabilities, public API and
on-disk formats of Mellanox UDA (chaoyang223/UDA), re-designed for sm_100a + NVLink 5.

Layers (SURVEY §1 ↔ this package):
  L1 common runtime      uda_b200.utils (VInt, IFile, logging, config), csrc/runtime
  L2 transport           uda_b200.parallel (symmetric arena over CUDA IPC, fused push shuffle)
  L3a consumer/NetMerger uda_b200.consumer, uda_b200.ops.merger, csrc/kernels/merge.cu
  L3b provider/MOFSupplier uda_b200.provider, uda_b200.mof
  L4 JNI bridge          csrc/jni (libuda.so), uda_b200.bridge (the same C ABI driven from Python)
  L5 Java plugins        java/
"""
__version__ = "1.0.0"
