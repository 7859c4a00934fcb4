"""vermilion_b200 — B200-native (sm_100a) engine for Vermilion's ray-tracing hot path.

csrc/   CUDA kernels + the C ABI (include/vermilion_b200.h) -> libvermilion_b200.so
host/   C++ host side: procedural scene generators and the Vermilion::Integrator adapter (B200PathTracer)
engine.py / scenes.py   ctypes plumbing over the two libraries for tests and bench.py
"""
