import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, REPO)


def _ensure_built():
    """Built artefacts are git-ignored: (re)build what is missing before collection.
    oracle/_ref needs /root/reference and is only built where that exists."""
    env = dict(os.environ, PATH="/usr/local/cuda/bin:" + os.environ.get("PATH", ""))
    if not os.path.exists(os.path.join(REPO, "oracle", "liboracle.so")):
        subprocess.check_call(["make", "-C", os.path.join(REPO, "oracle"), "liboracle.so"], env=env)
    if not os.path.exists(os.path.join(REPO, "oracle", "_ref", "libref_runtime.so")) and os.path.isdir("/root/reference/taichi"):
        subprocess.check_call(["make", "-C", os.path.join(REPO, "oracle"), "ref"], env=env)
    if not os.path.exists(os.path.join(REPO, "taichi_b200", "lib", "libb200mpm.so")):
        subprocess.check_call(["make", "-C", os.path.join(REPO, "taichi_b200", "csrc"), "-j8"], env=env)


_ensure_built()


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
