"""cm_gpu_align_batch_noalloc (the entry point that allocates nothing: fixed 4 MiB device array)
against cm_gpu_align_batch on the same problems, bit for bit (integer DP results), two calls in a row per
kind -- the second on other inputs and a smaller batch, so stale contents of the shared array would show -- and the size
limit.  Collected last (file name) so that a failure here hides no other test under `pytest -x`."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _problems(seed, n, stride, kind):
    rng = np.random.default_rng(seed)
    acgt = np.frombuffer(b"ACGT", np.uint8)
    s = np.zeros((n, stride), np.uint8)
    t = np.zeros((n, stride), np.uint8)
    ns = np.zeros(n, np.uint16)
    nt = np.zeros(n, np.uint16)
    aux = np.zeros(n, np.int32)
    for i in range(n):
        L = int(rng.integers(8, 110))
        a = acgt[rng.integers(0, 4, size=L + 3)]
        b = a[:L].copy()
        for _ in range(int(rng.integers(0, 4))):          # substitutions
            b[int(rng.integers(0, L))] = acgt[rng.integers(0, 4)]
        if rng.random() < 0.3 and L > 12:                 # one deletion in the read piece
            p = int(rng.integers(2, L - 2))
            b = np.delete(b, p)
        if rng.random() < 0.05:
            b[int(rng.integers(0, len(b)))] = ord("N")
        if kind in (1, 3):
            a, b = a[::-1], b[::-1]
        if kind in (2, 3):                                # read piece first, exon text second
            a, b = b, a[:L]
        if kind == 4:                                     # one-side banded: |t| = |s| + w
            w = int(rng.integers(0, 4))
            a, b = a[:L], acgt[rng.integers(0, 4, size=L + w)]
            b[:L] = a
            aux[i] = w
        s[i, :len(a)] = a
        t[i, :len(b)] = b
        ns[i], nt[i] = len(a), len(b)
    return s, ns, t, nt, aux


@pytest.mark.parametrize("kind", [0, 1, 2, 3, 4])
def test_noalloc_entry_equals_the_allocating_one(kind):
    from circminer_b200.api import Mapper, default_params
    m = Mapper(default_params())
    for call, (seed, n) in enumerate(((100 + kind, 3000), (200 + kind, 1700))):
        s, ns, t, nt, aux = _problems(seed, n, 128, kind)
        want = m.align(kind, s, ns, t, nt, aux)
        got = m.align_noalloc(kind, s, ns, t, nt, aux)
        for k in ("ed", "sclen", "indel", "score"):
            bad = np.nonzero(got[k] != want[k])[0]
            assert len(bad) == 0, f"kind {kind} call {call} {k}: {len(bad)} differ, first {bad[:5]}: {got[k][bad[:5]]} vs {want[k][bad[:5]]}"
        assert len(np.unique(want["ed"])) > 2, "the problems exercise the DP"
    m.close()


def test_noalloc_entry_refuses_what_does_not_fit():
    from circminer_b200.api import CmGpuError, Mapper, default_params
    m = Mapper(default_params())
    m.lib.cm_gpu_noalloc_scratch_bytes.restype = C.c_size_t
    cap = m.lib.cm_gpu_noalloc_scratch_bytes()
    assert cap == 4 << 20
    n = cap // (2 * 128) + 1                               # the two text regions alone exceed the array
    s, ns, t, nt, aux = _problems(1, 8, 128, 0)
    big = lambda a: np.ascontiguousarray(np.resize(a, (n,) + a.shape[1:]))
    with pytest.raises(CmGpuError, match="scratch"):
        m.align_noalloc(0, big(s), big(ns), big(t), big(nt), big(aux))
    # and the context still works afterwards
    got = m.align_noalloc(0, s, ns, t, nt, aux)
    want = m.align(0, s, ns, t, nt, aux)
    assert all((got[k] == want[k]).all() for k in got)
    m.close()
