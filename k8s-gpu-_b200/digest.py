"""Result digests: one sha256 over a placement (or reconcile) result, so that a whole result at
BASELINE.json's full sizes can be pinned by a 64-character string (tests/golden/digests.json) and
printed in every bench line (G-invariance shows up as equal digests at 1/2/4/8 GPUs).

The byte string hashed is the little-endian contents of the arrays, in the fixed order below,
each preceded by its element count as a little-endian u64.  Pure hashlib: nothing here computes a
placement, it only hashes one.
"""
from __future__ import annotations

import hashlib

import numpy as np

PLACE_FIELDS = (("assign", np.int32), ("free_gpu", np.uint32), ("free_mcpu", np.uint32), ("free_mib", np.uint32))
RECONCILE_FIELDS = (("pool_of", np.int32), ("current", np.int32), ("delta", np.int32), ("ready", np.int32),
                    ("create_count", np.int32), ("victim_off", np.int32), ("victim_vm", np.uint32),
                    ("member_off", np.int32), ("member_vm", np.uint32), ("create_off", np.int32),
                    ("create_ordinal", np.uint32))


def _digest(result: dict, fields) -> str:
    h = hashlib.sha256()
    for name, dt in fields:
        a = np.ascontiguousarray(np.asarray(result[name]), dtype=np.dtype(dt).newbyteorder("<"))
        h.update(np.uint64(a.size).astype("<u8").tobytes())
        h.update(a.tobytes())
    return h.hexdigest()


def place_digest(result: dict) -> str:
    """sha256 of assign[m] + the three final free columns [n] (k8sgpu_place_out)."""
    return _digest(result, PLACE_FIELDS)


def reconcile_digest(result: dict) -> str:
    """sha256 of every array of k8sgpu_reconcile_out (member and create lists included)."""
    return _digest(result, RECONCILE_FIELDS)
