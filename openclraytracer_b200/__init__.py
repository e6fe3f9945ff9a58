"""openclraytracer_b200 — B200-native (sm_100a) ray-tracing core behind the plugin interface of kotturtech/OpenCLRayTracer.

Only the hot path lives here: csrc/ (CUDA kernels + the C ABI of include/clrt.h, built into libclrt_cuda.so) and the
host-side mirror of the reference's interface (api.py).  Importing the package does not load the CUDA library;
`openclraytracer_b200.api` does, and fails loudly if it has not been built.
"""
__version__ = "0.1.0"
