"""paths_b200: B200-native (sm_100a) implementation of the PATHS hierarchical slide-level forward/backward pass behind
the reference's Python API.  CUDA only: importing works anywhere, computing needs libpaths_b200.so and a CUDA device."""
__version__ = "0.1.0"
