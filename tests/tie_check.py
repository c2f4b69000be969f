"""Parity tooling (tests only): decides whether a pixel that differs from the oracle differs because of an EPSILON-TIE.

The reference accepts a candidate hit when t lies in the epsilon-closed range [tmin, best] (Range::inRange,
include/Util/Range.cuh:33-43, quirk Q1): a candidate within 1e-6 of the current best REPLACES it even when it is farther.
Which of two tied primitives (typically two triangles sharing an edge) is kept therefore depends on the order in which the
traversal meets them — the reference's order follows its own median-split BVH, ours follows the wide LBVH.  Both apply the
same rule to the same candidates; only the order differs, and no order-free rule can reproduce the reference's choice.
`is_epsilon_tie` replays the pixel with the CPU twin (the product's own cores, bit-equal to the CUDA path), traces every
path segment with the oracle as well, and checks that at the FIRST segment where the two disagree they hit the same
instance at distances less than 1e-6 apart — i.e. that the disagreement is a tie and nothing else."""
import ctypes as C

import numpy as np

import emul_ctypes as ec

_twins = {}


def is_epsilon_tie(sc, oracle_handle, x: int, y: int, seed: int, verbose: bool = True) -> bool:
    key = id(sc)
    if key not in _twins:
        _twins.clear()
        _twins[key] = ec.Emul(sc)
    e = _twins[key]
    cap = int(sc.sample_count) * max(1, int(sc.ray_trace_depth)) + 8
    out = np.zeros((cap, 12), np.float32)
    e.lib.emul_pixel_paths.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_void_p, C.c_int]
    e.lib.emul_pixel_paths.restype = C.c_int
    n = e.lib.emul_pixel_paths(e.h, seed, x, y, out.ctypes.data, cap)
    out = out[:min(n, cap)]
    t, inst, prim, _ = oracle_handle.trace(out[:, 2:8])
    ot = np.where(inst >= 0, t, np.inf).astype(np.float32)
    differs = (ot != out[:, 8]) | (inst != out[:, 9].astype(np.int32)) | ((inst >= 0) & ((prim & 0x0FFFFFFF) != out[:, 10].astype(np.int64)))
    idx = np.nonzero(differs)[0]
    if len(idx) == 0:
        return False                      # every segment agrees: the difference is not in the traversal
    k = int(idx[0])
    tie = bool(inst[k] >= 0 and out[k, 9] == inst[k] and abs(float(ot[k]) - float(out[k, 8])) < 1e-6)
    if verbose:
        print(f"pixel ({x},{y}) segment {k} (sample {int(out[k, 0])}, depth {int(out[k, 1])}): ours t={float(out[k, 8])!r} prim {int(out[k, 10])}, "
              f"reference t={float(ot[k])!r} prim {int(prim[k]) & 0x0FFFFFFF}: {'epsilon-tie' if tie else 'NOT a tie'}")
    return tie
