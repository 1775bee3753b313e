#!/usr/bin/env python
"""Generates tests/golden/yaw_ref.npz from the REFERENCE's own Trajectory1D::AngleController
(oracle/_ref/libyaw_ref.so = /root/reference/.../src/trajectory1d.cpp compiled where it lies; `make -C oracle ref`).
Run in the build container (the reference tree does not exist on the GPU box); the fixture is committed."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import yaw_binding as yb  # noqa: E402


def main():
    assert yb.have_ref(), "build oracle/_ref first: make -C oracle ref"
    rng = np.random.default_rng(20261104)
    w, dw, yaw = yb.random_queries(rng, 512)
    # the reference's own unit-test pairs (test/unit_test.cpp:12-13,40-41), at rest
    ut_from = np.array([np.pi + 0.1, np.pi, 0.0, 1.32, -3.0, 1.7, 2.3, -2.3])
    ut_to = np.array([np.pi, np.pi + 0.1, 1.32, 0.0, 1.7, -3.0, -1.4, 1.4])
    w = np.concatenate([ut_from, w])
    dw = np.concatenate([np.zeros(8), dw])
    yaw = np.concatenate([ut_to, yaw])
    out = {"w": w, "dw": dw, "yaw": yaw}
    for tag, (vm, am, n) in {"a": (1.0, 1.0, 25), "b": (2.0, 0.7, 20), "c": (0.5, 3.0, 0)}.items():
        r = yb.ref_profiles(w, dw, yaw, vm, am, n)
        out[f"{tag}_params"] = np.array([vm, am, n])
        for k, v in r.items():
            out[f"{tag}_{k}"] = v
    np.savez_compressed(os.path.join(HERE, "yaw_ref.npz"), **out)
    print("wrote yaw_ref.npz:", {k: v.shape for k, v in out.items() if k.startswith("a_")})


if __name__ == "__main__":
    main()
