"""pytest configuration: the `gpu` marker, in-tree builds, shared scene fixtures.

`-m "not gpu"` covers the oracles (port vs the unmodified reference / committed golden vectors), the host scene
compiler and the C-ABI surface; `-m gpu` are the parity tests proper, all through the C ABI (libb200pt.so).
Nothing here reads /root/reference at run time: the reference harness is the prebuilt oracle/_ref/ref_driver.
"""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

SCENE_FILES = ["three_orbs", "cornell_box", "cornell_box_small_lights", "cornell_box_with_volume",
               "cornell_box_with_volume2", "light_test", "book2_final", "cornell_box_image_textures"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "slow: long statistical comparison")


def _build_once():
    lib = os.path.join(ROOT, "pathtrace_b200", "libb200pt.so")
    port = os.path.join(ROOT, "oracle", "_build", "libpt_oracle.so")
    if os.path.exists(lib) and os.path.exists(port):
        return
    import __graft_entry__ as g
    g.build()


@pytest.fixture(scope="session", autouse=True)
def built():
    _build_once()


def scene_path(name):
    return os.path.join(ROOT, "scenes", name + ".json")


@pytest.fixture(scope="session")
def has_gpu():
    try:
        out = subprocess.run(["nvidia-smi", "-L"], capture_output=True, text=True, timeout=30)
        return out.returncode == 0 and "GPU" in out.stdout
    except Exception:
        return False
