"""Shared helpers of the parity tests."""
import numpy as np

from stereo_matching_b200 import synth


def case_inputs(c):
    """Rebuilds the inputs of a tests/golden/kat_hashes.json case (same recipe as make_golden.py)."""
    L, R, _ = synth.random_dot_pair(c["h"], c["w"], seed=c["seed"], bg=c["bg"], fg=c["fg"], noise=c["noise"])
    sky = synth.sky_masks(c["h"] // c["scale"], c["w"] // c["scale"]) if c["sky"] else (None, None)
    return L, R, sky


def bits_equal(a, b):
    a = np.ascontiguousarray(a)
    b = np.ascontiguousarray(b)
    return a.shape == b.shape and a.dtype == b.dtype and bool((a.view(np.uint8) == b.view(np.uint8)).all())


def n_diff(a, b):
    a = np.asarray(a)
    b = np.asarray(b)
    if a.dtype.kind == "f":
        return int((a.view(np.uint32) != b.view(np.uint32)).sum())
    return int((a != b).sum())
