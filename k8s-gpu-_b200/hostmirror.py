"""ctypes front end of libk8sgpu_host.so: the C++ mirror of the reference's reconciler
(k8s-gpu-_b200/host/).  `World` bundles a fake API server, a fake ARM client, the single-pool
`AzureVmPoolReconciler` (README.md:170-235) and the GPU-backed `BatchReconciler`."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libk8sgpu_host.so")
ERR = {"None": 0, "Client": 1, "List": 2, "Create": 3, "Delete": 4, "Status": 5, "Gpu": 6}
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
