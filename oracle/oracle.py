"""ctypes front end of the CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY.

Importers allowed: tests/, __graft_entry__.smoke(), bench.py's cpu_baseline / --impl reference
legs.  Nothing in the product package imports this module.  PARITY UNPINNED: the reference
(/root/reference) is Markdown only; see oracle/k8s_oracle.h.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

NODE_COLS = ("free_gpu", "free_mcpu", "free_mib", "cap_gpu", "sku_id", "zone_id", "taint_mask", "flags")
POD_COLS = ("req_gpu", "req_mcpu", "req_mib", "sku_mask", "zone_mask", "tol_mask", "pref_zone", "priority")
NONE_U32 = 0xFFFFFFFF
ASSIGN_NONE = -1
ASSIGN_PENDING = -2
ASSIGN_GANG = -3
POLICY_ROT_ZERO = 1

u32p = C.POINTER(C.c_uint32)
i32p = C.POINTER(C.c_int32)


class NodeTable(C.Structure):
    _fields_ = [("n", C.c_uint32)] + [(c, u32p) for c in NODE_COLS]


class PodTable(C.Structure):
    _fields_ = [("m", C.c_uint32)] + [(c, u32p) for c in POD_COLS]


class PlaceStats(C.Structure):
    _fields_ = [("rounds", C.c_uint32), ("placed", C.c_uint32), ("evals", C.c_uint64)]


class PoolTable(C.Structure):
    _fields_ = [("p", C.c_uint32), ("desired", i32p), ("owner_id", u32p)]


class VmTable(C.Structure):
    _fields_ = [("v", C.c_uint32), ("owner_id", u32p), ("flags", u32p), ("ordinal", u32p)]


class ReconcileOut(C.Structure):
    _fields_ = [("pool_of", i32p), ("current", i32p), ("delta", i32p), ("ready", i32p),
                ("create_count", i32p), ("victim_off", i32p), ("victim_vm", u32p),
                ("member_off", i32p), ("member_vm", u32p), ("create_off", i32p),
                ("create_ordinal", u32p)]


class Result(C.Structure):
    _fields_ = [("requeue_after_s", C.c_int64), ("err", C.c_int)]


def build(force: bool = False) -> str:
    """Compile oracle/*.c into oracle/liboracle.so with the committed Makefile."""
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("k8s_oracle.c", "reconcile_sm.c", "k8s_oracle.h")]
    stale = (not os.path.exists(so)) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs)
    if force or stale:
        subprocess.run(["make", "-s", "-C", _HERE, "liboracle.so"], check=True)
    return so


def lib() -> C.CDLL:
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        L.orc_mix64.restype = C.c_uint64
        L.orc_mix64.argtypes = [C.c_uint64]
        L.orc_splitmix64_at.restype = C.c_uint64
        L.orc_splitmix64_at.argtypes = [C.c_uint64, C.c_uint64]
        L.orc_rot.restype = C.c_uint32
        L.orc_rot.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32]
        L.orc_validate_tables.argtypes = [C.POINTER(NodeTable), C.POINTER(PodTable)]
        L.orc_score_key.restype = C.c_uint64
        L.orc_score_key.argtypes = [C.POINTER(NodeTable), u32p, u32p, u32p, C.POINTER(PodTable),
                                    C.c_uint32, C.c_uint32, C.c_uint32]
        L.orc_score_round.restype = None
        L.orc_score_round.argtypes = [C.POINTER(NodeTable), u32p, u32p, u32p, C.POINTER(PodTable), u32p,
                                      C.c_uint32, C.c_uint32, C.c_int, i32p]
        L.orc_admit_round.restype = C.c_uint32
        L.orc_admit_round.argtypes = [C.c_uint32, u32p, u32p, u32p, C.POINTER(PodTable), u32p, C.c_uint32,
                                      i32p, i32p, u32p]
        L.orc_place.argtypes = [C.POINTER(NodeTable), C.POINTER(PodTable), C.c_uint32, C.c_uint32, C.c_int,
                                i32p, u32p, u32p, u32p, C.POINTER(PlaceStats)]
        L.orc_place_gang.argtypes = [C.POINTER(NodeTable), C.POINTER(PodTable), u32p, u32p, C.c_uint32, C.c_uint32,
                                     C.c_uint32, C.c_int, i32p, u32p, u32p, u32p, i32p, C.POINTER(PlaceStats), u32p]
        L.orc_reconcile_batch.argtypes = [C.POINTER(PoolTable), C.POINTER(VmTable), C.c_uint32, C.c_int,
                                          C.POINTER(ReconcileOut)]
        for f in ("orc_arm_new", "orc_cluster_new"):
            getattr(L, f).restype = C.c_void_p
        L.orc_arm_free.argtypes = [C.c_void_p]
        L.orc_cluster_free.argtypes = [C.c_void_p]
        L.orc_arm_add_vm.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p]
        L.orc_arm_vm_count.argtypes = [C.c_void_p]
        L.orc_arm_vm_name.argtypes = [C.c_void_p, C.c_int]
        L.orc_arm_vm_name.restype = C.c_char_p
        for f in ("orc_arm_fail_list_at", "orc_arm_fail_create_at", "orc_arm_fail_delete_at",
                  "orc_arm_fail_client"):
            getattr(L, f).argtypes = [C.c_void_p, C.c_int]
            getattr(L, f).restype = None
        for f in ("orc_arm_calls_list", "orc_arm_calls_create", "orc_arm_calls_delete"):
            getattr(L, f).argtypes = [C.c_void_p]
        L.orc_cluster_put_pool.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int32]
        L.orc_cluster_set_replicas.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int32]
        L.orc_cluster_delete_pool.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p]
        L.orc_cluster_fail_status_update.argtypes = [C.c_void_p, C.c_int]
        L.orc_cluster_fail_status_update.restype = None
        L.orc_cluster_ready_replicas.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p]
        L.orc_cluster_ready_replicas.restype = C.c_int32
        L.orc_cluster_status_vm_count.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p]
        L.orc_cluster_status_vm.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int]
        L.orc_cluster_status_vm.restype = C.c_char_p
        L.orc_cluster_status_writes.argtypes = [C.c_void_p]
        L.orc_reconcile_one.argtypes = [C.c_void_p, C.c_void_p, C.c_char_p, C.c_char_p]
        L.orc_reconcile_one.restype = Result
        _LIB = L
    return _LIB


def _u32(a):
    a = np.ascontiguousarray(a, dtype=np.uint32)
    return a, a.ctypes.data_as(u32p)


def _i32(a):
    a = np.ascontiguousarray(a, dtype=np.int32)
    return a, a.ctypes.data_as(i32p)


def _node_table(nodes):
    keep = []
    t = NodeTable()
    t.n = len(nodes[NODE_COLS[0]])
    for c in NODE_COLS:
        a, p = _u32(nodes[c])
        keep.append(a)
        setattr(t, c, p)
    return t, keep


def _pod_table(pods):
    keep = []
    t = PodTable()
    t.m = len(pods[POD_COLS[0]])
    for c in POD_COLS:
        a, p = _u32(pods[c])
        keep.append(a)
        setattr(t, c, p)
    return t, keep


def rot(pod_index: int, n_nodes: int, flags: int = 0) -> int:
    return int(lib().orc_rot(pod_index, n_nodes, flags))


def score_key(nodes, pods, p: int, n: int, flags: int = 0) -> int:
    nt, k1 = _node_table(nodes)
    pt, k2 = _pod_table(pods)
    r = rot(p, nt.n, flags)
    return int(lib().orc_score_key(C.byref(nt), nt.free_gpu, nt.free_mcpu, nt.free_mib, C.byref(pt), p, n, r))


def validate(nodes, pods) -> int:
    nt, k1 = _node_table(nodes)
    pt, k2 = _pod_table(pods)
    return int(lib().orc_validate_tables(C.byref(nt), C.byref(pt)))


def score_round(nodes, free, pods, active, flags: int = 0, nthreads: int = 1) -> np.ndarray:
    """free = (free_gpu, free_mcpu, free_mib) current columns."""
    nt, k1 = _node_table(nodes)
    pt, k2 = _pod_table(pods)
    fg, fgp = _u32(free[0]); fc, fcp = _u32(free[1]); fm, fmp = _u32(free[2])
    act, actp = _u32(active)
    best = np.empty(len(act), dtype=np.int32)
    lib().orc_score_round(C.byref(nt), fgp, fcp, fmp, C.byref(pt), actp, len(act), flags, nthreads,
                          best.ctypes.data_as(i32p))
    return best


def admit_round(n_nodes, free, pods, active, best, assign):
    """Mutates free (3 uint32 arrays) and assign (int32) in place; returns next active list."""
    pt, k2 = _pod_table(pods)
    for a in free:
        assert a.dtype == np.uint32 and a.flags.c_contiguous
    assert assign.dtype == np.int32 and assign.flags.c_contiguous
    act, actp = _u32(active)
    b, bp = _i32(best)
    nxt = np.empty(len(act) + 1, dtype=np.uint32)
    k = lib().orc_admit_round(n_nodes, free[0].ctypes.data_as(u32p), free[1].ctypes.data_as(u32p),
                              free[2].ctypes.data_as(u32p), C.byref(pt), actp, len(act), bp,
                              assign.ctypes.data_as(i32p), nxt.ctypes.data_as(u32p))
    return nxt[:k].copy()


def place(nodes, pods, flags: int = 0, max_rounds: int = 0, nthreads: int = 1):
    nt, k1 = _node_table(nodes)
    pt, k2 = _pod_table(pods)
    assign = np.empty(pt.m, dtype=np.int32)
    fg = np.empty(nt.n, dtype=np.uint32); fc = np.empty(nt.n, dtype=np.uint32); fm = np.empty(nt.n, dtype=np.uint32)
    st = PlaceStats()
    rc = lib().orc_place(C.byref(nt), C.byref(pt), flags, max_rounds, nthreads, assign.ctypes.data_as(i32p),
                         fg.ctypes.data_as(u32p), fc.ctypes.data_as(u32p), fm.ctypes.data_as(u32p), C.byref(st))
    if rc != 0:
        raise ValueError("oracle: tables outside the input domain")
    return {"assign": assign, "free_gpu": fg, "free_mcpu": fc, "free_mib": fm,
            "rounds": int(st.rounds), "placed": int(st.placed), "evals": int(st.evals)}


def place_gang(nodes, pods, gang_id, min_available, flags: int = 0, max_rounds: int = 0, nthreads: int = 1):
    nt, k1 = _node_table(nodes)
    pt, k2 = _pod_table(pods)
    gid, gidp = _u32(gang_id)
    mn, mnp = _u32(min_available)
    assign = np.empty(pt.m, dtype=np.int32)
    fg = np.empty(nt.n, dtype=np.uint32); fc = np.empty(nt.n, dtype=np.uint32); fm = np.empty(nt.n, dtype=np.uint32)
    ok = np.empty(len(mn), dtype=np.int32)
    st, it = PlaceStats(), C.c_uint32()
    rc = lib().orc_place_gang(C.byref(nt), C.byref(pt), gidp, mnp, len(mn), flags, max_rounds, nthreads,
                              assign.ctypes.data_as(i32p), fg.ctypes.data_as(u32p), fc.ctypes.data_as(u32p),
                              fm.ctypes.data_as(u32p), ok.ctypes.data_as(i32p), C.byref(st), C.byref(it))
    if rc != 0:
        raise ValueError("oracle: tables outside the input domain or bad gang id")
    return {"assign": assign, "free_gpu": fg, "free_mcpu": fc, "free_mib": fm, "gang_ok": ok,
            "rounds": int(st.rounds), "placed": int(st.placed), "evals": int(st.evals), "iterations": int(it.value)}


def reconcile_batch(desired, pool_owner_id, vm_owner_id, vm_flags, n_owner_ids, vm_ordinal=None,
                    nthreads: int = 1):
    d, dp = _i32(desired)
    po, pop = _u32(pool_owner_id)
    vo, vop = _u32(vm_owner_id)
    vf, vfp = _u32(vm_flags)
    P, V = len(d), len(vo)
    pools = PoolTable(P, dp, pop)
    if vm_ordinal is not None:
        vr, vrp = _u32(vm_ordinal)
    else:
        vr, vrp = None, u32p()
    vms = VmTable(V, vop, vfp, vrp)
    o = {
        "pool_of": np.empty(V, np.int32), "current": np.empty(P, np.int32), "delta": np.empty(P, np.int32),
        "ready": np.empty(P, np.int32), "create_count": np.empty(P, np.int32),
        "victim_off": np.empty(P + 1, np.int32), "victim_vm": np.zeros(V + 1, np.uint32),
        "member_off": np.empty(P + 1, np.int32), "member_vm": np.zeros(V + 1, np.uint32),
        "create_off": np.empty(P + 1, np.int32),
    }
    # create_ordinal capacity = sum(max(desired - current, 0)) <= sum(desired)
    cap = int(np.maximum(d.astype(np.int64), 0).sum()) + 1
    o["create_ordinal"] = np.zeros(cap, np.uint32)
    out = ReconcileOut()
    for k, v in o.items():
        setattr(out, k, v.ctypes.data_as(i32p if v.dtype == np.int32 else u32p))
    rc = lib().orc_reconcile_batch(C.byref(pools), C.byref(vms), n_owner_ids, nthreads, C.byref(out))
    if rc != 0:
        raise ValueError("oracle: invalid pool table")
    o["victim_vm"] = o["victim_vm"][: o["victim_off"][P]].copy()
    o["member_vm"] = o["member_vm"][: o["member_off"][P]].copy()
    o["create_ordinal"] = o["create_ordinal"][: o["create_off"][P]].copy()
    return o


def config1_ns_per_reconcile(iterations: int = 20000) -> float:
    """BASELINE.json config 1 on the CPU state machine: nanoseconds per Reconcile call (K1..K5 loop in C)."""
    import time

    L = lib()
    L.orc_config1_loop.argtypes = [C.c_int]
    L.orc_config1_loop.restype = C.c_long
    if L.orc_config1_loop(100) != 600:
        raise RuntimeError("config 1 sequence did not end as K1..K5 say")
    t0 = time.perf_counter()
    calls = L.orc_config1_loop(iterations)
    dt = time.perf_counter() - t0
    if calls != 6 * iterations:
        raise RuntimeError("config 1 sequence did not end as K1..K5 say")
    return 1e9 * dt / calls


class FakeArm:
    """Fake Azure resource group (insertion-ordered VM list with tags + failure injection)."""

    def __init__(self):
        self._h = lib().orc_arm_new()

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_arm_free(self._h)
            self._h = None

    def add_vm(self, name, managed_by="azurevmpool-operator", owner=None):
        enc = lambda s: None if s is None else s.encode()
        return lib().orc_arm_add_vm(self._h, name.encode(), enc(managed_by), enc(owner))

    def names(self):
        L = lib()
        return [L.orc_arm_vm_name(self._h, i).decode() for i in range(L.orc_arm_vm_count(self._h))]

    def fail_list_at(self, k): lib().orc_arm_fail_list_at(self._h, k)
    def fail_create_at(self, k): lib().orc_arm_fail_create_at(self._h, k)
    def fail_delete_at(self, k): lib().orc_arm_fail_delete_at(self._h, k)
    def fail_client(self, on=True): lib().orc_arm_fail_client(self._h, int(on))

    @property
    def calls(self):
        L = lib()
        return {"list": L.orc_arm_calls_list(self._h), "create": L.orc_arm_calls_create(self._h),
                "delete": L.orc_arm_calls_delete(self._h)}


class FakeCluster:
    """Fake API server holding AzureVmPool objects (spec.replicas + status)."""

    def __init__(self):
        self._h = lib().orc_cluster_new()

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_cluster_free(self._h)
            self._h = None

    def put_pool(self, ns, name, replicas): return lib().orc_cluster_put_pool(self._h, ns.encode(), name.encode(), replicas)
    def set_replicas(self, ns, name, replicas): return lib().orc_cluster_set_replicas(self._h, ns.encode(), name.encode(), replicas)
    def delete_pool(self, ns, name): return lib().orc_cluster_delete_pool(self._h, ns.encode(), name.encode())
    def fail_status_update(self, on=True): lib().orc_cluster_fail_status_update(self._h, int(on))
    def ready_replicas(self, ns, name): return lib().orc_cluster_ready_replicas(self._h, ns.encode(), name.encode())

    def status_vms(self, ns, name):
        L = lib()
        n = L.orc_cluster_status_vm_count(self._h, ns.encode(), name.encode())
        return [L.orc_cluster_status_vm(self._h, ns.encode(), name.encode(), i).decode() for i in range(max(n, 0))]

    @property
    def status_writes(self): return lib().orc_cluster_status_writes(self._h)


def reconcile_one(cluster: FakeCluster, arm: FakeArm, ns: str, name: str):
    r = lib().orc_reconcile_one(cluster._h, arm._h, ns.encode(), name.encode())
    return int(r.requeue_after_s), int(r.err)
