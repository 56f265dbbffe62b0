"""Build libst2dec.so in-tree: ``python styletts2_b200/build.py`` (nvcc, sm_100a, -lineinfo)."""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))


def build(verbose: bool = False, jobs: int = 8) -> str:
    csrc = os.path.join(HERE, "csrc")
    env = dict(os.environ)
    env["PATH"] = "/usr/local/cuda/bin:" + env.get("PATH", "")
    r = subprocess.run(["make", "-C", csrc, f"-j{jobs}"], env=env, capture_output=True, text=True)
    if verbose or r.returncode:
        sys.stdout.write(r.stdout)
        sys.stderr.write(r.stderr)
    if r.returncode:
        raise RuntimeError("building libst2dec.so failed")
    out = os.path.join(HERE, "_lib", "libst2dec.so")
    assert os.path.exists(out)
    return out


if __name__ == "__main__":
    print(build(verbose=True))
