"""The C-ABI library loads on a box without a GPU and exports every symbol include/*.h declares
(no compute calls here); the ctypes table of swarmrl_b200._lib covers exactly those symbols."""
import ctypes
import pathlib
import re

import pytest

ROOT = pathlib.Path(__file__).resolve().parent.parent


def _declared():
    names = set()
    for h in (ROOT / "include").glob("*.h"):
        txt = re.sub(r"/\*.*?\*/", "", h.read_text(), flags=re.S)
        names |= set(re.findall(r"\b(swarm_[a-z0-9_]+)\s*\(", txt))
    return names


def test_library_exports_every_declared_symbol():
    import __graft_entry__

    __graft_entry__.build()
    from swarmrl_b200 import _lib

    declared = _declared()
    assert len(declared) >= 25
    lib = ctypes.CDLL(str(_lib.LIB_PATH))
    missing = [n for n in sorted(declared) if not hasattr(lib, n)]
    assert not missing, missing
    assert set(_lib.SYMBOLS) == declared
    L = _lib.load()
    assert L.swarm_abi_version() == 1


def test_philox_host_entry_point_matches_oracle():
    """swarm_philox_host is pure host code: callable without a GPU."""
    import numpy as np

    from oracle import oracle
    from swarmrl_b200 import _lib

    L = _lib.load()
    c = (ctypes.c_uint64 * 4)(5, 6, 7, 8)
    k = (ctypes.c_uint64 * 2)(11, 12)
    o = (ctypes.c_uint64 * 4)()
    assert L.swarm_philox_host(c, k, o) == 0
    assert np.array_equal(np.array(list(o), dtype=np.uint64), oracle.philox4x64([5, 6, 7, 8], [11, 12]))
    # the thermostat's noise words (two Philox4x32-10 blocks): library == oracle, bit for bit
    for seed, pid, step, stream in ((42, 7, 1000, 3), (2**64 - 1, 2**40 + 9, 2**40 + 5, 5), (0, 0, 0, 0)):
        assert L.swarm_noise_words_host(seed, pid, step, stream, o) == 0
        assert np.array_equal(np.array(list(o), dtype=np.uint64), oracle.noise_words(seed, pid, step, stream))


def test_no_cpu_fallback_without_a_device():
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from swarmrl_b200 import _lib
    from swarmrl_b200.device import DeviceSwarm

    with pytest.raises(_lib.SwarmLibraryError):
        DeviceSwarm(2, (1, 1, 1), True, 1e-3, 0.0, 0, [dict(type=0, gamma_t=(1,) * 3, gamma_r=(1,) * 3, mass=1.0,
                                                           rinertia=(1,) * 3)], dict(sigma=[[1.0]], eps=[[1.0]], cut=[[1.0]]))


def test_product_never_imports_the_oracle():
    for f in (ROOT / "swarmrl_b200").rglob("*"):
        if f.suffix in (".py", ".cu", ".cuh", ".h"):
            txt = f.read_text()
            assert "import oracle" not in txt and "from oracle" not in txt and "swarm_oracle" not in txt, f


def test_cluster_kernel_compiles_to_the_hardware_cluster_barrier():
    """k_bd2d_cluster (csrc/swarm_cluster.cuh) must reach the thread-block-cluster hardware, not a software barrier: its SASS
    for sm_100a holds the cluster barrier's arrive / wait pair (UCGABAR_ARV / UCGABAR_WAIT) and the library is built for
    sm_100a only.  Needs cuobjdump (part of the CUDA toolkit the build uses), no GPU."""
    import shutil
    import subprocess

    import pytest

    from swarmrl_b200 import _lib

    tool = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not pathlib.Path(tool).exists():
        pytest.skip("cuobjdump not found")
    elf = subprocess.run([tool, "-lelf", str(_lib.LIB_PATH)], capture_output=True, text=True).stdout
    assert "sm_100a" in elf and "sm_90" not in elf and "sm_80" not in elf
    sass = subprocess.run([tool, "-sass", "-fun", "_Z14k_bd2d_clusterILb1EEv9DevParams9DevArrays7RunArgs", str(_lib.LIB_PATH)],
                          capture_output=True, text=True).stdout
    assert sass.count("UCGABAR_ARV") >= 2 and sass.count("UCGABAR_WAIT") >= 2
