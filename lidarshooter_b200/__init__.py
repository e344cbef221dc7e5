"""lidarshooter_b200 -- B200-native LiDAR ray-casting backend behind lidarshooter's tracer interface.

The product is two shared libraries built in-tree by `make -C lidarshooter_b200`:
  liblsb200.so       hand-written sm_100a CUDA kernels + the C ABI of include/lsb200.h
  liblsb200_host.so  C++ host layer (LidarDevice / B200Tracer / mesh + config readers) over that ABI
This Python package only binds them (ctypes) for tests and bench.py and generates synthetic workloads.
"""
from . import capi, synth  # noqa: F401
