"""fp64 peak of this GPU as the Brownian kernels see it: a DFMA chain micro-benchmark (torch is only the launcher of a
tiny inline CUDA kernel compiled at run time with nvcc -> cubin -> cuda-python driver API would need more plumbing, so
the kernel lives in libswarm_b200.so as swarm_microbench_dfma)."""
import ctypes as C, json, sys, pathlib
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import torch
from swarmrl_b200 import _lib
L = _lib.load()
out = (C.c_double * 4)()
_lib.check(L.swarm_microbench_dfma(0, 8, 4096, out))
print(json.dumps({"dfma_per_s": out[0], "fp64_tflops": 2 * out[0] / 1e12, "ms": out[1], "sm_count": out[2], "ilp": 8}))
