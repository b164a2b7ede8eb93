"""ctypes front end of libk8sgpu_host.so: the C++ mirror of the reference's reconciler
(k8s-gpu-_b200/host/).  `World` bundles a fake API server, a fake ARM client, the single-pool
`AzureVmPoolReconciler` (README.md:170-235) and the GPU-backed `BatchReconciler`."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libk8sgpu_host.so")
ERR = {"None": 0, "Client": 1, "List": 2, "Create": 3, "Delete": 4, "Status": 5, "Gpu": 6, "Conflict": 7}
FAIL = {"client": 0, "list": 1, "create": 2, "delete": 3, "status": 4}
_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: run __graft_entry__.build()")
        L = C.CDLL(LIB_PATH)
        vp, cp = C.c_void_p, C.c_char_p
        L.khost_new.restype = vp
        L.khost_free.argtypes = [vp]
        L.khost_put_pool.argtypes = [vp, cp, cp, C.c_int32, cp]
        L.khost_delete_pool.argtypes = [vp, cp, cp]
        L.khost_add_vm.argtypes = [vp, cp, cp, cp, cp]
        L.khost_fail.argtypes = [vp, C.c_int, C.c_int]
        L.khost_fail.restype = None
        L.khost_calls.argtypes = [vp, C.c_int]
        L.khost_status_writes.argtypes = [vp]
        L.khost_vm_count.argtypes = [vp, cp]
        L.khost_vm_name.argtypes = [vp, cp, C.c_int]
        L.khost_vm_name.restype = cp
        L.khost_status_ready.argtypes = [vp, cp, cp]
        L.khost_status_vm_count.argtypes = [vp, cp, cp]
        L.khost_status_vm.argtypes = [vp, cp, cp, C.c_int]
        L.khost_status_vm.restype = cp
        L.khost_status_condition.argtypes = [vp, cp, cp]
        L.khost_status_condition.restype = cp
        L.khost_status_conditions.argtypes = [vp, cp, cp]
        L.khost_status_conditions.restype = cp
        L.khost_reconcile.argtypes = [vp, cp, cp, C.POINTER(C.c_int64), C.POINTER(C.c_int)]
        L.khost_reconcile_batch.argtypes = [vp, C.c_int, C.c_int, C.POINTER(cp), C.POINTER(cp),
                                            C.POINTER(C.c_int64), C.POINTER(C.c_int)]
        L.khost_last_gpu_launches.argtypes = [vp]
        L.khost_last_gpu_launches.restype = C.c_uint64
        _lib = L
    return _lib


class World:
    def __init__(self):
        self._L = load()
        self._h = self._L.khost_new()

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.khost_free(self._h)
            self._h = None

    def put_pool(self, ns, name, replicas, rg="rg"):
        return self._L.khost_put_pool(self._h, ns.encode(), name.encode(), replicas, rg.encode())

    def delete_pool(self, ns, name):
        return self._L.khost_delete_pool(self._h, ns.encode(), name.encode())

    def add_vm(self, name, managed_by="azurevmpool-operator", owner=None, rg="rg"):
        enc = lambda s: None if s is None else s.encode()
        return self._L.khost_add_vm(self._h, rg.encode(), name.encode(), enc(managed_by), enc(owner))

    def fail(self, kind, k=1):
        self._L.khost_fail(self._h, FAIL[kind], k)

    def calls(self, kind):
        return self._L.khost_calls(self._h, FAIL[kind])

    @property
    def status_writes(self):
        return self._L.khost_status_writes(self._h)

    def vm_names(self, rg="rg"):
        n = self._L.khost_vm_count(self._h, rg.encode())
        return [self._L.khost_vm_name(self._h, rg.encode(), i).decode() for i in range(n)]

    def ready_replicas(self, ns, name):
        return self._L.khost_status_ready(self._h, ns.encode(), name.encode())

    def status_vms(self, ns, name):
        n = self._L.khost_status_vm_count(self._h, ns.encode(), name.encode())
        return [self._L.khost_status_vm(self._h, ns.encode(), name.encode(), i).decode() for i in range(max(n, 0))]

    def condition(self, ns, name):
        c = self._L.khost_status_condition(self._h, ns.encode(), name.encode())
        return c.decode() if c else None

    def conditions(self, ns, name):
        """Every condition of the pool's status: {"Ready": ("True", "Reconciled"), ...}."""
        c = self._L.khost_status_conditions(self._h, ns.encode(), name.encode())
        if not c:
            return None
        out = {}
        for item in c.decode().split(";"):
            if item:
                t, rest = item.split("=", 1)
                st, reason = rest.split(":", 1)
                out[t] = (st, reason)
        return out

    def reconcile(self, ns, name):
        """AzureVmPoolReconciler.Reconcile for one pool -> (requeue_after_seconds, error kind)."""
        rq, err = C.c_int64(), C.c_int()
        self._L.khost_reconcile(self._h, ns.encode(), name.encode(), C.byref(rq), C.byref(err))
        return rq.value, err.value

    def reconcile_batch(self, keys, device=0):
        """BatchReconciler.ReconcileAll over [(ns, name), ...] through the GPU."""
        n = len(keys)
        ns = (C.c_char_p * n)(*[k[0].encode() for k in keys])
        nm = (C.c_char_p * n)(*[k[1].encode() for k in keys])
        rq, err = (C.c_int64 * n)(), (C.c_int * n)()
        rc = self._L.khost_reconcile_batch(self._h, device, n, ns, nm, rq, err)
        if rc != 0:
            raise RuntimeError(f"k8sgpu error {rc}: no CUDA device or library failure (there is no CPU fallback)")
        return [(rq[i], err[i]) for i in range(n)]

    @property
    def last_gpu_launches(self):
        return int(self._L.khost_last_gpu_launches(self._h))


NODE_COLS = ("free_gpu", "free_mcpu", "free_mib", "cap_gpu", "sku_id", "zone_id", "taint_mask", "flags")
POD_COLS = ("req_gpu", "req_mcpu", "req_mib", "sku_mask", "zone_mask", "tol_mask", "pref_zone", "priority")


class Snapshot:
    """host/snapshot.hpp: AzureVmPool manifests (YAML) + ARM VM list (JSON) + Volcano jobs (YAML)
    -> the u32 tables of include/k8sgpu.h, as numpy columns."""

    def __init__(self):
        import numpy as np

        self.np = np
        L = self._L = load()
        vp, cp = C.c_void_p, C.c_char_p
        L.khost_snap_new.restype = vp
        L.khost_snap_free.argtypes = [vp]
        L.khost_snap_error.argtypes = [vp]
        L.khost_snap_error.restype = cp
        for f in ("khost_snap_add_pools_yaml", "khost_snap_add_vms_json", "khost_snap_add_jobs_yaml"):
            getattr(L, f).argtypes = [vp, cp]
        L.khost_snap_register_sku.argtypes = [vp, cp, C.c_uint32, C.c_uint32, C.c_uint32]
        L.khost_snap_register_sku.restype = None
        L.khost_snap_build.argtypes = [vp]
        L.khost_snap_build.restype = None
        L.khost_snap_rows.argtypes = [vp, C.c_int]
        L.khost_snap_rows.restype = C.c_uint32
        L.khost_snap_n_owner_ids.argtypes = [vp]
        L.khost_snap_n_owner_ids.restype = C.c_uint32
        L.khost_snap_col.argtypes = [vp, C.c_int, C.c_int]
        L.khost_snap_col.restype = C.POINTER(C.c_uint32)
        L.khost_snap_name.argtypes = [vp, C.c_int, C.c_uint32]
        L.khost_snap_name.restype = cp
        L.khost_snap_pool_field.argtypes = [vp, C.c_uint32, cp]
        L.khost_snap_pool_field.restype = cp
        L.khost_snap_conflict.argtypes = [vp, C.c_uint32]
        L.khost_snap_conflict.restype = cp
        self._h = L.khost_snap_new()

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.khost_snap_free(self._h)
            self._h = None

    def _add(self, fn, text):
        if getattr(self._L, fn)(self._h, text.encode()) != 0:
            raise ValueError(self._L.khost_snap_error(self._h).decode())

    def add_pools_yaml(self, text): self._add("khost_snap_add_pools_yaml", text)
    def add_vms_json(self, text): self._add("khost_snap_add_vms_json", text)
    def add_jobs_yaml(self, text): self._add("khost_snap_add_jobs_yaml", text)

    def register_sku(self, vm_size, gpu, mcpu, mib):
        self._L.khost_snap_register_sku(self._h, vm_size.encode(), gpu, mcpu, mib)

    def pool_field(self, i, field):
        v = self._L.khost_snap_pool_field(self._h, i, field.encode())
        return None if v is None else v.decode()

    def conflicts(self):
        """Pools left out of the reconcile tables (after build()): [(dropped "ns/name", kept "ns/name")]."""
        out, i = [], 0
        while True:
            v = self._L.khost_snap_conflict(self._h, i)
            if v is None:
                return out
            out.append(tuple(v.decode().split("|", 1)))
            i += 1

    def _col(self, table, col, n, dtype=None):
        np = self.np
        if n == 0:
            return np.zeros(0, dtype or np.uint32)
        p = self._L.khost_snap_col(self._h, table, col)
        a = np.ctypeslib.as_array(p, shape=(n,)).copy()
        return a.view(dtype) if dtype is not None else a

    def _names(self, table, n):
        return [self._L.khost_snap_name(self._h, table, i).decode() for i in range(n)]

    def build(self) -> dict:
        """-> {"reconcile": kwargs of Context.reconcile_batch, "nodes": {...}, "pods": {...}, names}"""
        np, L = self.np, self._L
        L.khost_snap_build(self._h)
        P, V, N, M = (L.khost_snap_rows(self._h, t) for t in range(4))
        rec = dict(desired=self._col(0, 0, P, np.int32), pool_owner_id=self._col(0, 1, P),
                   vm_owner_id=self._col(1, 0, V), vm_flags=self._col(1, 1, V), vm_ordinal=self._col(1, 2, V),
                   n_owner_ids=int(L.khost_snap_n_owner_ids(self._h)))
        nodes = {c: self._col(2, i, N) for i, c in enumerate(NODE_COLS)}
        pods = {c: self._col(3, i, M) for i, c in enumerate(POD_COLS)}
        return dict(reconcile=rec, nodes=nodes, pods=pods, pool_keys=self._names(0, P), vm_names=self._names(1, V),
                    node_names=self._names(2, N), pod_names=self._names(3, M))
