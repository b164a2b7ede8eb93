"""ctypes binding of libk8sgpu.so (include/k8sgpu.h).  Every call goes to the CUDA library;
there is no Python or CPU fallback: a missing library or a missing GPU raises."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libk8sgpu.so")

OK, E_ARG, E_CUDA, E_NCCL, E_CAPACITY = 0, -1, -2, -3, -4
NONE_U32 = 0xFFFFFFFF
ASSIGN_NONE, ASSIGN_PENDING = -1, -2
POLICY_ROT_ZERO = 1

NODE_COLS = ("free_gpu", "free_mcpu", "free_mib", "cap_gpu", "sku_id", "zone_id", "taint_mask", "flags")
POD_COLS = ("req_gpu", "req_mcpu", "req_mib", "sku_mask", "zone_mask", "tol_mask", "pref_zone", "priority")

# every symbol include/k8sgpu.h declares (tests check the library exports all of them)
EXPORTS = (
    "k8sgpu_abi_version", "k8sgpu_create", "k8sgpu_destroy", "k8sgpu_last_error", "k8sgpu_get_stats",
    "k8sgpu_reconcile_batch", "k8sgpu_reconcile_upload", "k8sgpu_reconcile_run", "k8sgpu_reconcile_download",
    "k8sgpu_place", "k8sgpu_place_upload", "k8sgpu_place_run", "k8sgpu_place_download", "k8sgpu_score_once",
    "k8sgpu_place_round_begin", "k8sgpu_place_round_score", "k8sgpu_place_round_admit", "k8sgpu_place_round_end",
    "k8sgpu_debug_sort_pairs", "k8sgpu_debug_exclusive_scan", "k8sgpu_exclusive_scan_noalloc",
    "k8sgpu_place_gang", "k8sgpu_peer_export", "k8sgpu_peer_import", "k8sgpu_place_round_score_p2p",
    "k8sgpu_place_gang_begin", "k8sgpu_place_gang_verdicts", "k8sgpu_place_gang_end",
    "k8sgpu_create_ex", "k8sgpu_set_tuning", "k8sgpu_reconcile_out_size", "k8sgpu_comm_unique_id",
    "k8sgpu_comm_init_rank", "k8sgpu_group_info", "k8sgpu_peer_close", "k8sgpu_host_alloc", "k8sgpu_host_free",
)
UNIQUE_ID_BYTES = 128
TUNING_KEYS = ("res", "ppt", "splits", "coop", "coop_grid", "rec_path", "rec_grid", "rec_segments", "coop_prof",
               "rec_prof", "shard_admit", "exchange", "order", "force_cls", "shard_upload", "slot_admit")
ASSIGN_GANG = -3
SCAN_NOALLOC_MAX_N = 1 << 24

u32p = C.POINTER(C.c_uint32)
i32p = C.POINTER(C.c_int32)


class PoolTable(C.Structure):
    _fields_ = [("p", C.c_uint32), ("desired", i32p), ("owner_id", u32p)]


class VmTable(C.Structure):
    _fields_ = [("v", C.c_uint32), ("owner_id", u32p), ("flags", u32p), ("ordinal", u32p)]


class ReconcileOut(C.Structure):
    _fields_ = [("pool_of", i32p), ("current", i32p), ("delta", i32p), ("ready", i32p), ("create_count", i32p),
                ("victim_off", i32p), ("victim_vm", u32p), ("member_off", i32p), ("member_vm", u32p),
                ("create_off", i32p), ("create_ordinal", u32p), ("create_capacity", C.c_uint64)]


class ReconcileLayout(C.Structure):
    _fields_ = [(k, C.c_uint64) for k in ("pool_of", "current", "ready", "delta", "create_count", "victim_off",
                                          "member_off", "create_off", "victim_vm", "member_vm", "create_ordinal",
                                          "total")]


class NodeTable(C.Structure):
    _fields_ = [("n", C.c_uint32)] + [(c, u32p) for c in NODE_COLS]


class PodTable(C.Structure):
    _fields_ = [("m", C.c_uint32)] + [(c, u32p) for c in POD_COLS]


class Policy(C.Structure):
    _fields_ = [("flags", C.c_uint32), ("max_rounds", C.c_uint32)]


class PlaceOut(C.Structure):
    _fields_ = [("assign", i32p), ("free_gpu", u32p), ("free_mcpu", u32p), ("free_mib", u32p)]


class Stats(C.Structure):
    _fields_ = [("rounds", C.c_uint32), ("placed", C.c_uint32), ("evals", C.c_uint64),
                ("kernel_launches", C.c_uint64), ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64),
                ("device_ms", C.c_float), ("score_ms", C.c_float), ("score_evals", C.c_uint64),
                ("score_launches", C.c_uint32), ("gang_iterations", C.c_uint32)]

    def as_dict(self):
        return {f: getattr(self, f) for f, _ in self._fields_}


class GangTable(C.Structure):
    _fields_ = [("n_gangs", C.c_uint32), ("gang_id", u32p), ("min_available", u32p)]


class K8sGpuError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"k8sgpu error {code}: {msg}")
        self.code = code


_lib = None


def load() -> C.CDLL:
    """dlopen libk8sgpu.so (built in-tree by __graft_entry__.build / csrc/Makefile)."""
    global _lib
    if _lib is None:
        path = os.environ.get("K8SGPU_LIB", LIB_PATH)  # developer override: a differently tuned build
        if not os.path.exists(path):
            raise ImportError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(there is no fallback path)")
        L = C.CDLL(path)
        vp = C.c_void_p
        L.k8sgpu_abi_version.restype = C.c_int
        L.k8sgpu_create.argtypes = [C.POINTER(C.c_int), C.c_int, C.POINTER(vp)]
        L.k8sgpu_destroy.argtypes = [vp]
        L.k8sgpu_last_error.argtypes = [vp]
        L.k8sgpu_last_error.restype = C.c_char_p
        L.k8sgpu_get_stats.argtypes = [vp, C.POINTER(Stats)]
        L.k8sgpu_reconcile_batch.argtypes = [vp, C.POINTER(PoolTable), C.POINTER(VmTable), C.c_uint32,
                                             C.POINTER(ReconcileOut)]
        L.k8sgpu_reconcile_upload.argtypes = [vp, C.POINTER(PoolTable), C.POINTER(VmTable), C.c_uint32]
        L.k8sgpu_reconcile_run.argtypes = [vp, C.c_int, C.c_int]
        L.k8sgpu_reconcile_download.argtypes = [vp, C.POINTER(ReconcileOut)]
        L.k8sgpu_place.argtypes = [vp, C.POINTER(NodeTable), C.POINTER(PodTable), C.POINTER(Policy),
                                   C.POINTER(PlaceOut)]
        L.k8sgpu_place_gang.argtypes = [vp, C.POINTER(NodeTable), C.POINTER(PodTable), C.POINTER(GangTable),
                                        C.POINTER(Policy), C.POINTER(PlaceOut), i32p]
        L.k8sgpu_place_upload.argtypes = [vp, C.POINTER(NodeTable), C.POINTER(PodTable)]
        L.k8sgpu_place_run.argtypes = [vp, C.POINTER(Policy)]
        L.k8sgpu_place_download.argtypes = [vp, C.POINTER(PlaceOut)]
        L.k8sgpu_score_once.argtypes = [vp, C.POINTER(Policy), i32p]
        L.k8sgpu_place_round_begin.argtypes = [vp, C.POINTER(Policy), vp, u32p]
        L.k8sgpu_place_round_score.argtypes = [vp, C.c_uint32, C.c_uint32, vp, vp]
        L.k8sgpu_place_round_admit.argtypes = [vp, vp, vp, u32p]
        L.k8sgpu_place_round_end.argtypes = [vp, vp]
        L.k8sgpu_peer_export.argtypes = [vp, C.c_uint32, vp, C.POINTER(vp)]
        L.k8sgpu_peer_import.argtypes = [vp, C.c_int, C.c_int, vp]
        L.k8sgpu_peer_close.argtypes = [vp]
        L.k8sgpu_host_alloc.argtypes = [C.c_size_t, C.POINTER(vp)]
        L.k8sgpu_host_free.argtypes = [vp]
        L.k8sgpu_place_round_score_p2p.argtypes = [vp, C.c_uint32, C.c_uint32, vp]
        L.k8sgpu_place_gang_begin.argtypes = [vp, C.POINTER(GangTable), C.POINTER(Policy), vp, u32p]
        L.k8sgpu_place_gang_verdicts.argtypes = [vp, vp, u32p]
        L.k8sgpu_place_gang_end.argtypes = [vp, vp, i32p]
        L.k8sgpu_debug_sort_pairs.argtypes = [vp, u32p, u32p, C.c_uint32, C.c_int, u32p, u32p]
        L.k8sgpu_debug_exclusive_scan.argtypes = [vp, u32p, C.c_uint32, u32p, u32p]
        L.k8sgpu_exclusive_scan_noalloc.argtypes = [vp, vp, C.c_uint32, vp, vp, vp]
        L.k8sgpu_create_ex.argtypes = [C.POINTER(C.c_int), C.c_int, C.POINTER(vp), C.c_char_p, C.c_size_t]
        L.k8sgpu_set_tuning.argtypes = [vp, C.c_char_p, C.c_int]
        L.k8sgpu_reconcile_out_size.argtypes = [C.c_uint32, C.c_uint32, C.c_uint64, C.POINTER(ReconcileLayout)]
        L.k8sgpu_comm_unique_id.argtypes = [vp]
        L.k8sgpu_comm_init_rank.argtypes = [vp, C.c_int, C.c_int, vp]
        L.k8sgpu_group_info.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                        C.POINTER(C.c_int)]
        for name in EXPORTS:
            if name != "k8sgpu_last_error":
                getattr(L, name).restype = C.c_int
        _lib = L
    return _lib


def _u32(a):
    a = np.ascontiguousarray(a, dtype=np.uint32)
    return a, a.ctypes.data_as(u32p)


def _i32(a):
    a = np.ascontiguousarray(a, dtype=np.int32)
    return a, a.ctypes.data_as(i32p)


def node_table(nodes: dict):
    t, keep = NodeTable(), []
    t.n = len(nodes[NODE_COLS[0]])
    for c in NODE_COLS:
        a, p = _u32(nodes[c])
        keep.append(a)
        setattr(t, c, p)
    return t, keep


def pod_table(pods: dict):
    t, keep = PodTable(), []
    t.m = len(pods[POD_COLS[0]])
    for c in POD_COLS:
        if c == "priority" and pods.get(c) is None:  # reserved column: may be left out
            continue
        a, p = _u32(pods[c])
        keep.append(a)
        setattr(t, c, p)
    return t, keep


class PinnedBlock:
    """`n` uint32 elements of page-locked host memory from k8sgpu_host_alloc, as a numpy array
    (`.array`); freed when the object is closed or collected."""

    def __init__(self, n: int):
        self._L = load()
        self._p = C.c_void_p()
        rc = self._L.k8sgpu_host_alloc(max(int(n), 1) * 4, C.byref(self._p))
        if rc != OK:
            raise K8sGpuError(rc, "k8sgpu_host_alloc failed")
        self.array = np.ctypeslib.as_array(C.cast(self._p, u32p), shape=(max(int(n), 1),))[: int(n)]

    def close(self):
        if getattr(self, "_p", None) and self._p.value:
            self.array = None
            self._L.k8sgpu_host_free(self._p)
            self._p = C.c_void_p()

    __del__ = close


def unique_id() -> bytes:
    """A fresh ncclUniqueId (rank 0 makes it and hands it to the other ranks)."""
    buf = (C.c_ubyte * UNIQUE_ID_BYTES)()
    rc = load().k8sgpu_comm_unique_id(C.cast(buf, C.c_void_p))
    if rc != OK:
        raise K8sGpuError(rc, "k8sgpu_comm_unique_id failed (is NCCL loadable?)")
    return bytes(buf)


def reconcile_out_size(p: int, v: int, create_capacity: int) -> dict:
    """Element offsets of every output array inside one block (k8sgpu_reconcile_out_size)."""
    lay = ReconcileLayout()
    rc = load().k8sgpu_reconcile_out_size(p, v, create_capacity, C.byref(lay))
    if rc != OK:
        raise K8sGpuError(rc, "k8sgpu_reconcile_out_size")
    return {k: int(getattr(lay, k)) for k, _ in ReconcileLayout._fields_}


class Context:
    """One k8sgpu_ctx: one GPU (device: int) or a placement group of several GPUs of this
    process (device: a list of ints).  Not re-entrant: one call at a time."""

    def __init__(self, device=0):
        self._L = load()
        self._h = C.c_void_p()
        devs = [int(device)] if isinstance(device, (int, np.integer)) else [int(d) for d in device]
        dev = (C.c_int * len(devs))(*devs)
        err = C.create_string_buffer(512)
        rc = self._L.k8sgpu_create_ex(dev, len(devs), C.byref(self._h), err, len(err))
        if rc != OK:
            raise K8sGpuError(rc, err.value.decode())
        self.device = devs[0]
        self.devices = devs

    def set_tuning(self, **knobs):
        """Developer / test knobs (k8sgpu_set_tuning); None or -1 = automatic."""
        for k, v in knobs.items():
            self._check(self._L.k8sgpu_set_tuning(self._h, k.encode(), -1 if v is None else int(v)))

    def reset_tuning(self):
        self.set_tuning(**{k: (0 if k.endswith("_prof") else -1) for k in TUNING_KEYS})

    def comm_init_rank(self, rank: int, world: int, uid: bytes):
        """Join `world` one-GPU contexts (one per process) into a placement group; collective."""
        buf = (C.c_ubyte * UNIQUE_ID_BYTES).from_buffer_copy(uid)
        self._check(self._L.k8sgpu_comm_init_rank(self._h, rank, world, C.cast(buf, C.c_void_p)))

    def group_info(self) -> dict:
        w, r, nl, v = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        self._check(self._L.k8sgpu_group_info(self._h, C.byref(w), C.byref(r), C.byref(nl), C.byref(v)))
        return {"world": w.value, "rank": r.value, "nlocal": nl.value, "nccl_version": v.value}

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.k8sgpu_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc):
        if rc != OK:
            raise K8sGpuError(rc, self._L.k8sgpu_last_error(self._h).decode())

    def stats(self) -> dict:
        s = Stats()
        self._check(self._L.k8sgpu_get_stats(self._h, C.byref(s)))
        return s.as_dict()

    # ------------------------------------------------------------ placement
    def place(self, nodes: dict, pods: dict, flags: int = 0, max_rounds: int = 0, out: dict = None) -> dict:
        """Host tables in, host results out (the drop-in call).  `out` may hold caller-owned result
        arrays (assign i32[m], free_gpu / free_mcpu / free_mib u32[n]), e.g. pinned ones."""
        nt, k1 = node_table(nodes)
        pt, k2 = pod_table(pods)
        out, res = self._place_out(nt.n, pt.m, out)
        pol = Policy(flags, max_rounds)
        self._check(self._L.k8sgpu_place(self._h, C.byref(nt), C.byref(pt), C.byref(pol), C.byref(out)))
        res.update(self.stats())
        return res

    def place_gang(self, nodes: dict, pods: dict, gang_id, min_available, flags: int = 0, max_rounds: int = 0) -> dict:
        """All-or-nothing gangs on top of the placement rounds (k8sgpu_place_gang)."""
        nt, k1 = node_table(nodes)
        pt, k2 = pod_table(pods)
        gid, gidp = _u32(gang_id)
        mn, mnp = _u32(min_available)
        gt = GangTable(len(mn), gidp, mnp)
        out, res = self._place_out(nt.n, pt.m)
        ok = np.zeros(len(mn), np.int32)
        pol = Policy(flags, max_rounds)
        self._check(self._L.k8sgpu_place_gang(self._h, C.byref(nt), C.byref(pt), C.byref(gt), C.byref(pol),
                                              C.byref(out), ok.ctypes.data_as(i32p)))
        res["gang_ok"] = ok
        res.update(self.stats())
        return res

    def _place_out(self, n, m, given=None):
        if given is not None:
            res = {k: given[k] for k in ("assign", "free_gpu", "free_mcpu", "free_mib")}
            ok = res["assign"].dtype == np.int32 and len(res["assign"]) == m and res["assign"].flags.c_contiguous
            ok = ok and all(res[k].dtype == np.uint32 and len(res[k]) == n and res[k].flags.c_contiguous
                            for k in ("free_gpu", "free_mcpu", "free_mib"))
            if not ok:
                raise ValueError("out: need contiguous assign int32[m] and free_* uint32[n]")
        else:
            res = {"assign": np.empty(m, np.int32), "free_gpu": np.empty(n, np.uint32),
                   "free_mcpu": np.empty(n, np.uint32), "free_mib": np.empty(n, np.uint32)}
        out = PlaceOut(res["assign"].ctypes.data_as(i32p), res["free_gpu"].ctypes.data_as(u32p),
                       res["free_mcpu"].ctypes.data_as(u32p), res["free_mib"].ctypes.data_as(u32p))
        return out, res

    def place_upload(self, nodes: dict, pods: dict):
        nt, k1 = node_table(nodes)
        pt, k2 = pod_table(pods)
        self._n, self._m = nt.n, pt.m
        self._check(self._L.k8sgpu_place_upload(self._h, C.byref(nt), C.byref(pt)))

    def place_run(self, flags: int = 0, max_rounds: int = 0) -> dict:
        pol = Policy(flags, max_rounds)
        self._check(self._L.k8sgpu_place_run(self._h, C.byref(pol)))
        return self.stats()

    def place_download(self) -> dict:
        out, res = self._place_out(self._n, self._m)
        self._check(self._L.k8sgpu_place_download(self._h, C.byref(out)))
        return res

    def score_once(self, flags: int = 0) -> np.ndarray:
        best = np.empty(self._m, np.int32)
        pol = Policy(flags, 0)
        self._check(self._L.k8sgpu_score_once(self._h, C.byref(pol), best.ctypes.data_as(i32p)))
        return best

    # round-at-a-time form (device pointers; see placement.ShardedPlacer)
    def round_begin(self, flags: int = 0, max_rounds: int = 0, stream: int = 0) -> int:
        pol = Policy(flags, max_rounds)
        na = C.c_uint32()
        self._check(self._L.k8sgpu_place_round_begin(self._h, C.byref(pol), C.c_void_p(stream), C.byref(na)))
        return na.value

    def round_score(self, lo: int, hi: int, d_best_ptr: int, stream: int = 0):
        self._check(self._L.k8sgpu_place_round_score(self._h, lo, hi, C.c_void_p(d_best_ptr), C.c_void_p(stream)))

    def round_admit(self, d_best_all_ptr: int, stream: int = 0) -> int:
        na = C.c_uint32()
        self._check(self._L.k8sgpu_place_round_admit(self._h, C.c_void_p(d_best_all_ptr), C.c_void_p(stream),
                                                     C.byref(na)))
        return na.value

    # gang placement one step at a time (see placement.ShardedPlacer.run_gang)
    def gang_begin(self, gang_id, min_available, flags: int = 0, max_rounds: int = 0, stream: int = 0) -> int:
        gid, gidp = _u32(gang_id)
        mn, mnp = _u32(min_available)
        self._n_gangs = len(mn)
        gt = GangTable(len(mn), gidp, mnp)
        pol = Policy(flags, max_rounds)
        na = C.c_uint32()
        self._check(self._L.k8sgpu_place_gang_begin(self._h, C.byref(gt), C.byref(pol), C.c_void_p(stream), C.byref(na)))
        return na.value

    def gang_verdicts(self, stream: int = 0) -> int:
        na = C.c_uint32()
        self._check(self._L.k8sgpu_place_gang_verdicts(self._h, C.c_void_p(stream), C.byref(na)))
        return na.value

    def gang_end(self, stream: int = 0) -> dict:
        ok = np.zeros(self._n_gangs, np.int32)
        self._check(self._L.k8sgpu_place_gang_end(self._h, C.c_void_p(stream), ok.ctypes.data_as(i32p)))
        st = self.stats()
        st["gang_ok"] = ok
        return st

    def peer_export(self, m: int):
        """-> (64-byte IPC handle of this rank's exchange buffer, its device pointer)."""
        h = (C.c_ubyte * 64)()
        p = C.c_void_p()
        self._check(self._L.k8sgpu_peer_export(self._h, m, C.cast(h, C.c_void_p), C.byref(p)))
        return bytes(h), p.value

    def peer_close(self):
        self._check(self._L.k8sgpu_peer_close(self._h))

    def peer_import(self, rank: int, world: int, handles: bytes):
        buf = (C.c_ubyte * len(handles)).from_buffer_copy(handles)
        self._check(self._L.k8sgpu_peer_import(self._h, rank, world, C.cast(buf, C.c_void_p)))

    def round_score_p2p(self, lo: int, hi: int, stream: int = 0):
        self._check(self._L.k8sgpu_place_round_score_p2p(self._h, lo, hi, C.c_void_p(stream)))

    def round_end(self, stream: int = 0) -> dict:
        self._check(self._L.k8sgpu_place_round_end(self._h, C.c_void_p(stream)))
        return self.stats()

    # ------------------------------------------------------------ reconcile
    @staticmethod
    def _rec_tables(desired, pool_owner_id, vm_owner_id, vm_flags, vm_ordinal):
        d, dp = _i32(desired)
        po, pop = _u32(pool_owner_id)
        vo, vop = _u32(vm_owner_id)
        vf, vfp = _u32(vm_flags)
        if vm_ordinal is not None:
            vr, vrp = _u32(vm_ordinal)
        else:
            vr, vrp = None, u32p()
        return PoolTable(len(d), dp, pop), VmTable(len(vo), vop, vfp, vrp), (d, po, vo, vf, vr)

    @staticmethod
    def _rec_out(P, V, create_cap, members=True, ordinals=True, alloc=None):
        """alloc(n, dtype) -> array: lets a caller hand in pinned or reused host buffers."""
        mk = alloc or (lambda n, dt: np.zeros(n, dt))
        o = {"pool_of": mk(V, np.int32), "current": mk(P, np.int32), "delta": mk(P, np.int32),
             "ready": mk(P, np.int32), "create_count": mk(P, np.int32),
             "victim_off": mk(P + 1, np.int32), "victim_vm": mk(V + 1, np.uint32)}
        if members:
            o["member_off"] = mk(P + 1, np.int32)
            o["member_vm"] = mk(V + 1, np.uint32)
        if ordinals:
            o["create_off"] = mk(P + 1, np.int32)
            o["create_ordinal"] = mk(create_cap + 1, np.uint32)
        out = ReconcileOut()
        for k, v in o.items():
            setattr(out, k, v.ctypes.data_as(i32p if v.dtype == np.int32 else u32p))
        out.create_capacity = create_cap + 1 if ordinals else 0
        return out, o

    @staticmethod
    def _rec_trim(o, P):
        o["victim_vm"] = o["victim_vm"][: o["victim_off"][P]].copy()
        if "member_vm" in o:
            o["member_vm"] = o["member_vm"][: o["member_off"][P]].copy()
        if "create_ordinal" in o:
            o["create_ordinal"] = o["create_ordinal"][: o["create_off"][P]].copy()
        return o

    def reconcile_batch(self, desired, pool_owner_id, vm_owner_id, vm_flags, n_owner_ids, vm_ordinal=None,
                        members=True, ordinals=True, create_capacity=None, alloc=None, trim=True) -> dict:
        pools, vms, keep = self._rec_tables(desired, pool_owner_id, vm_owner_id, vm_flags, vm_ordinal)
        cap = int(np.maximum(keep[0].astype(np.int64), 0).sum()) if create_capacity is None else create_capacity
        out, o = self._rec_out(pools.p, vms.v, cap, members, ordinals, alloc)
        self._check(self._L.k8sgpu_reconcile_batch(self._h, C.byref(pools), C.byref(vms), n_owner_ids, C.byref(out)))
        if trim:
            o = self._rec_trim(o, pools.p)
        o.update(self.stats())
        return o

    def reconcile_upload(self, desired, pool_owner_id, vm_owner_id, vm_flags, n_owner_ids, vm_ordinal=None):
        pools, vms, keep = self._rec_tables(desired, pool_owner_id, vm_owner_id, vm_flags, vm_ordinal)
        self._rp, self._rv = pools.p, vms.v
        self._rcap = int(np.maximum(keep[0].astype(np.int64), 0).sum())
        self._check(self._L.k8sgpu_reconcile_upload(self._h, C.byref(pools), C.byref(vms), n_owner_ids))

    def reconcile_run(self, members=True, ordinals=True) -> dict:
        self._check(self._L.k8sgpu_reconcile_run(self._h, int(members), int(ordinals)))
        return self.stats()

    def reconcile_download(self, members=True, ordinals=True) -> dict:
        out, o = self._rec_out(self._rp, self._rv, self._rcap, members, ordinals)
        self._check(self._L.k8sgpu_reconcile_download(self._h, C.byref(out)))
        return self._rec_trim(o, self._rp)

    # ------------------------------------------------------------ primitives
    def sort_pairs(self, keys, vals, bits: int):
        k, kp = _u32(keys)
        v, vp_ = _u32(vals)
        ok, ov = np.empty_like(k), np.empty_like(v)
        self._check(self._L.k8sgpu_debug_sort_pairs(self._h, kp, vp_, len(k), bits, ok.ctypes.data_as(u32p),
                                                    ov.ctypes.data_as(u32p)))
        return ok, ov

    def exclusive_scan(self, a):
        x, xp = _u32(a)
        out = np.empty_like(x)
        tot = C.c_uint32()
        self._check(self._L.k8sgpu_debug_exclusive_scan(self._h, xp, len(x), out.ctypes.data_as(u32p), C.byref(tot)))
        return out, tot.value

    def exclusive_scan_noalloc(self, d_in_ptr: int, n: int, d_out_ptr: int, d_total_ptr: int = 0, stream: int = 0):
        """Device pointers in and out; enqueues on `stream` and returns (k8sgpu.h)."""
        self._check(self._L.k8sgpu_exclusive_scan_noalloc(self._h, C.c_void_p(d_in_ptr), n, C.c_void_p(d_out_ptr),
                                                          C.c_void_p(d_total_ptr), C.c_void_p(stream)))
