"""Regenerates tests/golden/oracle_r01.npz from the CPU oracle:  python tests/golden/make_fixtures.py

The reference itself (Rust + NASM) cannot be built in this image, so these vectors are
oracle outputs, frozen so that neither the oracle nor the CUDA path can drift silently; the
vectors that come from the reference's own tests live in ref_dist_kat.json (src/dist.rs:379-402,
437-460) and are what pins the oracle."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)

import cases  # noqa: E402
from oracle import r1o  # noqa: E402

if __name__ == "__main__":
    v = cases.oracle_vectors(r1o)
    np.savez_compressed(os.path.join(HERE, "oracle_r01.npz"), **v)
    print("wrote", len(v), "vectors,", os.path.getsize(os.path.join(HERE, "oracle_r01.npz")), "bytes")
