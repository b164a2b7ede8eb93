"""Synthetic cluster snapshots for BASELINE.json configs 2-5.

Every column is a splitmix64 stream: element i of a column is the i-th output of the stream
seeded with 0xB200_0000_0000_0000 + config_id * 2**32 + column_id, so any language can
regenerate the same tables.  Nothing here is taken from the reference: it holds no data
beyond one sample CR (README.md:267-286) and one vcjob (GPU调度平台搭建.md:642-673), whose
shapes (Standard_NC4as_T4_v3; nvidia.com/gpu: 1, gpu-1x-16c-32g-1gpu) seed the SKU classes.
"""
from __future__ import annotations

import numpy as np

GAMMA = np.uint64(0x9E3779B97F4A7C15)
SEED_BASE = 0xB200_0000_0000_0000
NONE_U32 = np.uint32(0xFFFFFFFF)

NODE_COLS = ("free_gpu", "free_mcpu", "free_mib", "cap_gpu", "sku_id", "zone_id", "taint_mask", "flags")
POD_COLS = ("req_gpu", "req_mcpu", "req_mib", "sku_mask", "zone_mask", "tol_mask", "pref_zone", "priority")

# SKU classes: shape of Standard_NC4as_T4_v3 (README.md:100,276) and of an
# ND H100 v5 VM (external Azure facts, not in the reference).
SKU_T4 = dict(sku_id=0, cap_gpu=1, cap_mcpu=4000, cap_mib=28672)
SKU_H100 = dict(sku_id=1, cap_gpu=8, cap_mcpu=96000, cap_mib=1945600)


def mix64(z: np.ndarray) -> np.ndarray:
    z = z.astype(np.uint64, copy=True)
    with np.errstate(over="ignore"):
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def stream(seed: int, n: int) -> np.ndarray:
    """First n outputs of splitmix64 seeded with `seed`."""
    with np.errstate(over="ignore"):
        ctr = (np.arange(1, n + 1, dtype=np.uint64)) * GAMMA + np.uint64(seed & 0xFFFFFFFFFFFFFFFF)
    return mix64(ctr)


def col_seed(config_id: int, column_id: int) -> int:
    return (SEED_BASE + (config_id << 32) + column_id) & 0xFFFFFFFFFFFFFFFF


def _pct(config_id, col, n):
    return (stream(col_seed(config_id, col), n) % np.uint64(100)).astype(np.uint32)


def gen_nodes(n: int, config_id: int = 3, h100_pct: int = 30, taint_pct: int = 5,
              cordon_pct: int = 1, zones: int = 3) -> dict:
    """Config 3/4/5 node table: h100_pct % H100-like, rest T4-like; U{0..cap} GPUs pre-used,
    cpu/memory pro rata; taint bit 0 on taint_pct %; cordon_pct % not schedulable."""
    is_h = _pct(config_id, 0, n) < h100_pct
    cap_gpu = np.where(is_h, SKU_H100["cap_gpu"], SKU_T4["cap_gpu"]).astype(np.uint32)
    cap_mcpu = np.where(is_h, SKU_H100["cap_mcpu"], SKU_T4["cap_mcpu"]).astype(np.uint64)
    cap_mib = np.where(is_h, SKU_H100["cap_mib"], SKU_T4["cap_mib"]).astype(np.uint64)
    zone = (stream(col_seed(config_id, 1), n) % np.uint64(zones)).astype(np.uint32)
    used = (stream(col_seed(config_id, 2), n) % (cap_gpu.astype(np.uint64) + np.uint64(1))).astype(np.uint32)
    free_gpu = cap_gpu - used
    free_mcpu = (cap_mcpu * free_gpu.astype(np.uint64) // cap_gpu.astype(np.uint64)).astype(np.uint32)
    free_mib = (cap_mib * free_gpu.astype(np.uint64) // cap_gpu.astype(np.uint64)).astype(np.uint32)
    taint = (_pct(config_id, 3, n) < taint_pct).astype(np.uint32)
    flags = (_pct(config_id, 4, n) >= cordon_pct).astype(np.uint32)
    return {
        "free_gpu": free_gpu, "free_mcpu": free_mcpu, "free_mib": free_mib, "cap_gpu": cap_gpu,
        "sku_id": np.where(is_h, SKU_H100["sku_id"], SKU_T4["sku_id"]).astype(np.uint32),
        "zone_id": zone, "taint_mask": taint, "flags": flags,
    }


def gen_pods(m: int, config_id: int = 3, zones: int = 3) -> dict:
    """Config 3/4/5 pod table: req_gpu in {1:70,2:15,4:10,8:5}%; cpu/mem = 4000 mcpu and
    28672 MiB per GPU; 20 % H100-only; 50 % prefer a zone; 10 % single-zone; 50 % tolerate taint 0."""
    r = _pct(config_id, 8, m)
    g = np.where(r < 70, 1, np.where(r < 85, 2, np.where(r < 95, 4, 8))).astype(np.uint32)
    sku_mask = np.where(_pct(config_id, 9, m) < 20, np.uint32(1 << SKU_H100["sku_id"]), NONE_U32).astype(np.uint32)
    has_pref = _pct(config_id, 10, m) < 50
    pz = (stream(col_seed(config_id, 11), m) % np.uint64(zones)).astype(np.uint32)
    pref_zone = np.where(has_pref, pz, NONE_U32).astype(np.uint32)
    single = _pct(config_id, 12, m) < 10
    sz = (stream(col_seed(config_id, 13), m) % np.uint64(zones)).astype(np.uint32)
    zone_mask = np.where(single, np.uint32(1) << sz, NONE_U32).astype(np.uint32)
    tol = (_pct(config_id, 14, m) < 50).astype(np.uint32)
    return {
        "req_gpu": g, "req_mcpu": (g * np.uint32(4000)).astype(np.uint32),
        "req_mib": (g * np.uint32(28672)).astype(np.uint32), "sku_mask": sku_mask,
        "zone_mask": zone_mask, "tol_mask": tol, "pref_zone": pref_zone,
        "priority": np.zeros(m, np.uint32),
    }


def gen_identical(n: int, m: int, free_gpu: int = 8, req_gpu: int = 1) -> tuple[dict, dict]:
    """Config 5 adversarial ties: all nodes identical, all pods identical (K12)."""
    nodes = {
        "free_gpu": np.full(n, free_gpu, np.uint32), "free_mcpu": np.full(n, 12000 * free_gpu, np.uint32),
        "free_mib": np.full(n, 243200 * free_gpu, np.uint32), "cap_gpu": np.full(n, free_gpu, np.uint32),
        "sku_id": np.ones(n, np.uint32), "zone_id": np.zeros(n, np.uint32),
        "taint_mask": np.zeros(n, np.uint32), "flags": np.ones(n, np.uint32),
    }
    pods = {
        "req_gpu": np.full(m, req_gpu, np.uint32), "req_mcpu": np.full(m, 4000 * req_gpu, np.uint32),
        "req_mib": np.full(m, 28672 * req_gpu, np.uint32), "sku_mask": np.full(m, NONE_U32, np.uint32),
        "zone_mask": np.full(m, NONE_U32, np.uint32), "tol_mask": np.zeros(m, np.uint32),
        "pref_zone": np.full(m, NONE_U32, np.uint32), "priority": np.zeros(m, np.uint32),
    }
    return nodes, pods


def gen_reconcile(p: int, v: int, config_id: int = 2) -> dict:
    """Config 2: 90 % of VMs owned by a uniform pool, 5 % foreign owner id, 5 % managed-by
    clear; rows interleaved (generation order); desired = max(0, current + U{-20..20}).
    Pool p's owner id is P-1-p, so the owner->pool map is not the identity."""
    r = _pct(config_id, 0, v)
    u = (stream(col_seed(config_id, 1), v) % np.uint64(max(p, 1))).astype(np.uint32)
    vm_owner = np.where((r >= 90) & (r < 95), u + np.uint32(p), u).astype(np.uint32)
    vm_flags = np.where(r >= 95, 0, 1).astype(np.uint32)
    pool_owner = (np.uint32(p) - np.uint32(1) - np.arange(p, dtype=np.uint32)).astype(np.uint32)
    managed = (vm_flags == 1) & (vm_owner < p)
    current = np.bincount((p - 1 - vm_owner[managed]).astype(np.int64), minlength=p).astype(np.int64)
    jitter = (stream(col_seed(config_id, 2), p) % np.uint64(41)).astype(np.int64) - 20
    desired = np.maximum(0, current + jitter).astype(np.int32)
    span = max(4 * v // max(p, 1), 8)
    ordinal = (stream(col_seed(config_id, 3), v) % np.uint64(span)).astype(np.uint32)
    ordinal = np.where(_pct(config_id, 4, v) < 5, NONE_U32, ordinal).astype(np.uint32)
    return {"desired": desired, "pool_owner_id": pool_owner, "vm_owner_id": vm_owner, "vm_flags": vm_flags,
            "vm_ordinal": ordinal, "n_owner_ids": 2 * p}
