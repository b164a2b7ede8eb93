"""The integer identities the scoring kernel's PTX rests on (k8s-gpu-_b200/csrc/place.cu: eval_pair4,
eval_pair4_u2, eval_pair4_r, resmin_note), restated with numpy uint32 arithmetic and checked over
their whole input domain (include/k8sgpu.h: free_gpu <= 0x7FFF, req_* < 2^31, ids < 32).  CPU only:
the GPU parity tests check the kernels themselves against the oracle; these show that the field
layouts and constants the kernels use (restated here) make each shortcut an identity, not an
approximation."""
import numpy as np

U32 = np.uint32
rng = np.random.default_rng(20261107)


def u32(x):
    return np.asarray(x, dtype=np.uint64).astype(np.uint64) & np.uint64(0xFFFFFFFF)


def sign(x):  # bit 31 of a value held mod 2^32
    return (u32(x) >> np.uint64(31)).astype(bool)


def grid(a, b):
    a, b = np.meshgrid(np.asarray(a, np.uint64), np.asarray(b, np.uint64), indexing="ij")
    return a.ravel(), b.ravel()


FREE = np.unique(np.concatenate([np.arange(0, 70), [127, 128, 255, 256, 0x3FFF, 0x7FFE, 0x7FFF], rng.integers(0, 0x8000, 200)]))
REQ = np.unique(np.concatenate([np.arange(0, 70), [0x7FFF, 0x8000, 0x8001, 0xFFFF, 0x10000, 0x7FFFFFFF], rng.integers(0, 0x8000, 200),
                                rng.integers(0, 2 ** 31, 200)]))


def test_free_gpu_test_from_the_key_word():
    # khi = 1<<31 | (0x7FFF - free) << 16 | low 16 bits (cap field, index bits): khi + (1<<31 + min(req, 0x8000) << 16)
    # has bit 31 set exactly when free < req
    fg, rq = grid(FREE, REQ)
    low = rng.integers(0, 1 << 16, fg.size).astype(np.uint64)
    khi = np.uint64(0x80000000) | ((np.uint64(0x7FFF) - fg) << np.uint64(16)) | low
    c = np.uint64(0x80000000) + (np.minimum(rq, 0x8000) << np.uint64(16))
    assert np.array_equal(sign(khi + c), fg < rq)
    # as the kernel issues it: rq_reg = -c, d1 = rq_reg * (-1) + khi
    rq_reg = u32(np.uint64(1 << 32) - u32(c))
    assert np.array_equal(sign(rq_reg * np.uint64(0xFFFFFFFF) + khi), fg < rq)
    # an unschedulable or padding record is never feasible through free_mcpu = -1, whatever the CPU request
    rc = np.concatenate([[0, 1, 2 ** 31 - 1], rng.integers(0, 2 ** 31, 100)]).astype(np.uint64)
    assert sign(np.uint64(0xFFFFFFFF) + rc * np.uint64(0xFFFFFFFF)).all()


def zrot_word(zone, free, sched, tainted):
    sfree = np.where(sched, np.minimum(free, 127), 0xFF).astype(np.uint64) & np.uint64(0xFF)
    tfree = np.where(sched & ~tainted, free, 0xFFFFFFFF).astype(np.uint64)
    return u32(((zone.astype(np.uint64) - np.uint64(7)) & np.uint64(31)) | (sfree << np.uint64(5)) | u32(tfree << np.uint64(13)))


def test_free_gpu_fields_of_the_zrot_word():
    fg, rq = grid(FREE, REQ)
    n = fg.size
    zone = rng.integers(0, 32, n)
    sched = rng.random(n) > 0.2
    tainted = rng.random(n) < 0.3
    w = zrot_word(zone, fg, sched, tainted)
    # TT = 1: bits 13..31 against min(req, 0x8000) << 13 -- negative iff tainted / unschedulable / too few GPUs
    d1 = w + u32(np.uint64(1 << 32) - (np.minimum(rq, 0x8000) << np.uint64(13)))
    assert np.array_equal(sign(d1), ~sched | tainted | (fg < rq))
    # R, TT = 0: bits 5..12 brought to the top by a multiply by 2^19, against min(req, 64) << 24
    # (this loop only runs when every schedulable node has fewer than 64 free GPUs)
    small = fg < 64
    d1 = u32(w * np.uint64(1 << 19)) + u32(np.uint64(1 << 32) - (np.minimum(rq, 64) << np.uint64(24)))
    assert np.array_equal(sign(d1)[small], (~sched | (fg < rq))[small])
    # the funnel shift of the preference clause reads only bits 0..4
    assert np.array_equal(w & np.uint64(31), (zone.astype(np.uint64) - np.uint64(7)) & np.uint64(31))


def brev(x):
    x = u32(x)
    out = np.zeros_like(x)
    for i in range(32):
        out |= ((x >> np.uint64(i)) & np.uint64(1)) << np.uint64(31 - i)
    return out


def test_allow_masks_by_multiplication_and_by_and():
    mask = np.concatenate([rng.integers(0, 2 ** 32, 500, dtype=np.uint64), [0, 0xFFFFFFFF, 1, 1 << 31]]).astype(np.uint64)
    ids = np.arange(32, dtype=np.uint64)
    m, i = grid(mask, ids)
    allowed = ((m >> i) & np.uint64(1)).astype(bool)
    not_mask = u32(~m)
    # SKU: pod word brev(~mask), node word 1 << id: the product's bit 31 is ~mask's bit id
    assert np.array_equal(sign(brev(not_mask) * (np.uint64(1) << i)), ~allowed)
    # zone: pod word ~mask, node word 1 << (31 - id)
    assert np.array_equal(sign(not_mask * (np.uint64(1) << (np.uint64(31) - i))), ~allowed)
    # eval_pair4_u2 / _ur2: AND tests on the un-mirrored words (brev of the stored ones)
    assert np.array_equal((brev(brev(not_mask)) & (np.uint64(1) << i)) != 0, ~allowed)
    assert np.array_equal((brev(not_mask) & (np.uint64(1) << (np.uint64(31) - i))) != 0, ~allowed)


def test_zone_preference_bit_by_rotation():
    pz, zone = grid(np.arange(32), np.arange(32))
    pzb = np.uint64(1) << pz
    r = (zone - np.uint64(7)) & np.uint64(31)
    rot = u32((pzb >> r) | (pzb << ((np.uint64(32) - r) & np.uint64(31))))
    rot = np.where(r == 0, pzb, rot)
    assert np.array_equal((rot & np.uint64(0x80)) != 0, pz == zone)
    assert not ((np.uint64(0) >> r) & np.uint64(0x80)).any()  # no preference: the one-hot word is 0


def test_taint_classes():
    taints = rng.choice(np.array([0, 1, 2, 3, 0x80000000, 0x40000001], np.uint64), 300)
    tunion = np.bitwise_or.reduce(taints)
    for tol in (0, 1, 2, 3, 0x80000000, 0xC0000001, 0xFFFFFFFF, 0x40000000):
        ntol = np.uint64(~tol & 0xFFFFFFFF)
        ok = (taints & ntol) == 0  # the clause
        x = ntol & tunion
        if x == 0:  # TT = 0: always true
            assert ok.all()
        elif x == tunion:  # TT = 1: true exactly on untainted nodes
            assert np.array_equal(ok, taints == 0)


def test_resource_implication_from_the_round_summary():
    # resmin_note + the suffix minima of score_kernel: a pod with req_mcpu <= SC[req_gpu] and req_mib <= SM[req_gpu]
    # has enough CPU and memory at every schedulable node with enough GPUs; checked by brute force
    for case in range(20):
        n, m = 400, 300
        fg = rng.integers(0, rng.choice([2, 9, 64]), n)
        per_c = rng.choice([1500, 4000, 12000], n) if case % 2 else np.full(n, 4000)
        per_m = rng.choice([9000, 28672, 243200], n) if case % 3 else np.full(n, 28672)
        fc, fm = fg * per_c + rng.integers(0, 3, n), fg * per_m
        sched = rng.random(n) > 0.1
        tab_c = np.full(64, 0xFFFFFFFF, np.uint64)
        tab_m = np.full(64, 0xFFFFFFFF, np.uint64)
        for g, c, mm, s in zip(fg, fc, fm, sched):
            if s:
                tab_c[g] = min(tab_c[g], c)
                tab_m[g] = min(tab_m[g], mm)
        sc = np.append(np.minimum.accumulate(tab_c[::-1])[::-1], 0xFFFFFFFF)
        sm = np.append(np.minimum.accumulate(tab_m[::-1])[::-1], 0xFFFFFFFF)
        rg = rng.choice([0, 1, 2, 4, 8, 63, 64, 100], m)
        rc = rg * rng.choice([1000, 4000, 13000], m)
        rm = rg * rng.choice([1024, 28672, 300000], m)
        g = np.minimum(rg, 64)
        implied = (rc <= sc[g]) & (rm <= sm[g])
        enough_gpu = sched[None, :] & (fg[None, :] >= rg[:, None])
        enough_rest = (fc[None, :] >= rc[:, None]) & (fm[None, :] >= rm[:, None])
        assert not (enough_gpu & ~enough_rest)[implied].any()
        assert implied.any() or case % 2  # (the uniform tables qualify some pods)


def test_sort_key_fields_do_not_overlap():
    # pod_prep_kernel: (15 - class) << 20 | hash (hash_bits) << (20 - hash_bits) | where (20 - hash_bits bits)
    for hb in (0, 4, 12):
        rb = 20 - hb
        where_max = (1 << rb) - 1
        h_max = (1 << hb) - 1
        key = (15 << 20) | (h_max << rb) | where_max
        assert key < (1 << 24) and (h_max << rb) & where_max == 0
