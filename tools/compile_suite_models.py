"""Compiles the suite's MJCF files into FlatModel caches (dm_control_b200/suite/models).

Run in the build container, where /root/reference is mounted; the GPU box has no
reference tree, so the product and the tests load these compiled models instead
(the analogue of shipping MJB files, wrapper/core.py:510-519).
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dm_control_b200 import mjcf_compiler as C   # noqa: E402

SUITE = "/root/reference/dm_control/suite"
DOMAINS = ["cartpole", "acrobot", "pendulum", "cheetah", "walker", "humanoid",
           "hopper", "reacher", "ball_in_cup", "finger", "swimmer"]
OUT = os.path.join(ROOT, "dm_control_b200", "suite", "models")


def main():
  os.makedirs(OUT, exist_ok=True)
  for d in DOMAINS:
    m = C.compile_file(os.path.join(SUITE, d + ".xml"))
    path = os.path.join(OUT, d + ".npz")
    m.save(path)
    print(d, "nq", m.nq, "nv", m.nv, "nbody", m.nbody, "->", os.path.getsize(path), "B")


if __name__ == "__main__":
  main()
