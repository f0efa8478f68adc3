"""v2m_cuda_ingest_genotypes_no_alloc (K5 with its scratch in a static device array; no counterpart in the reference)
against the existing entry point v2m_cuda_paths_from_genotypes, through the C-ABI.
Collected last on purpose (the file name sorts after every other test file)."""
import ctypes as C

import numpy as np
import pytest

from tests import randgraph as rg

pytestmark = pytest.mark.gpu

MAX_COPIES = 65536          # V2M_NO_ALLOC_MAX_COPIES in include/v2m_cuda.h


@pytest.fixture(scope="module")
def v2m():
    import vcf2multialign_b200 as v
    v._lib.check(v.lib().v2m_cuda_init(0))
    return v


class _DeviceArrays:
    """Device copies of the caller's arrays, made with the library's own allocator BEFORE the call under test."""

    def __init__(self, v2m):
        self.v2m, self.L, self.ptrs = v2m, v2m.lib(), []

    def put(self, a):
        a = np.ascontiguousarray(a)
        d = self.L.v2m_cuda_device_alloc(max(a.nbytes, 16))
        assert d
        self.ptrs.append(d)
        if a.nbytes:
            self.v2m._lib.check(self.L.v2m_cuda_memcpy_h2d(d, a.ctypes.data_as(C.c_void_p), a.nbytes))
        return d

    def empty(self, n_bytes, fill=0xA5):
        return self.put(np.full(max(n_bytes, 16), fill, np.uint8))       # stale contents: the call has to clear what it uses

    def get(self, d, dtype, count):
        out = np.zeros(count, dtype)
        if out.nbytes:
            self.v2m._lib.check(self.L.v2m_cuda_memcpy_d2h(out.ctypes.data_as(C.c_void_p), d, out.nbytes))
        return out

    def free(self):
        for d in self.ptrs:
            self.L.v2m_cuda_device_free(d)
        self.ptrs = []


def _no_alloc(v2m, gt, alt_offsets, edges_by_alt, h_pad, e_pad, ref_pos=None, edge_target_pos=None, overlap_cap=0):
    """Returns (status, path words, sorted overlap pairs or None, needed overlap count)."""
    gt = np.ascontiguousarray(gt, np.uint16)
    R, H = gt.shape
    want = ref_pos is not None
    dev = _DeviceArrays(v2m)
    try:
        d_gt = dev.put(gt)
        d_off = dev.put(np.asarray(alt_offsets, np.uint64))
        d_eba = dev.put(np.asarray(edges_by_alt, np.uint64))
        d_ref = dev.put(np.asarray(ref_pos, np.uint64)) if want else None
        d_tgt = dev.put(np.asarray(edge_target_pos, np.uint64)) if want else None
        n_edges = len(edge_target_pos) if want else 0
        d_paths = dev.empty(h_pad * e_pad // 8)
        d_ov = dev.empty(16 * overlap_cap) if want else None
        n = C.c_uint64(0)
        st = dev.L.v2m_cuda_ingest_genotypes_no_alloc(d_gt, R, H, d_off, d_eba, d_ref, d_tgt, n_edges, h_pad, e_pad, d_paths,
                                                      d_ov, overlap_cap if want else 0, C.byref(n) if want else None, None)
        words = dev.get(d_paths, np.uint64, h_pad * e_pad // 64)
        pairs = None
        if want:
            k = min(int(n.value), overlap_cap)
            pairs = sorted(map(tuple, dev.get(d_ov, np.uint64, 2 * k).reshape(-1, 2).tolist()))
        return st, words, pairs, int(n.value)
    finally:
        dev.free()


def _case(tmp_path, seed, **kw):
    og = rg.random_oracle_graph(str(tmp_path), seed, **kw)
    return og.arrays(), og.genotype_trace()


def test_no_alloc_ingest_equals_the_allocating_entry_point_twice_in_a_row(v2m, tmp_path):
    """Two calls in a row on DIFFERENT inputs (the second one smaller, so that state the first call left in the shared scratch
    would show), then the first input again: path matrix bit-exact, overlap pairs equal as sorted lists."""
    cases = [
        _case(tmp_path, 4101, ref_len=1500, n_records=260, n_samples=700, ploidy=2, overlap=True, sv=True, alt_freq=0.2, many_alts_prob=0.1),
        _case(tmp_path, 4102, ref_len=900, n_records=90, n_samples=37, ploidy=3, overlap=True, sv=False, alt_freq=0.3, many_alts_prob=0.2),
    ]
    overlaps_seen = 0
    for a, t in (cases[0], cases[1], cases[0]):
        ref_words, ref_ov = v2m.paths_from_genotypes(t["gt"], t["alt_offsets"], t["edges_by_alt"], a["h_pad"], a["e_pad"], t["ref_pos"], t["edge_target_pos"])
        st, words, pairs, n = _no_alloc(v2m, t["gt"], t["alt_offsets"], t["edges_by_alt"], a["h_pad"], a["e_pad"], t["ref_pos"], t["edge_target_pos"],
                                        overlap_cap=max(16, len(ref_ov)))
        assert st == 0, v2m.lib().v2m_cuda_last_error()
        assert np.array_equal(words, ref_words)                       # integer work: bit-exact
        assert n == len(ref_ov) and pairs == sorted(map(tuple, ref_ov.tolist()))
        overlaps_seen += n
        # without overlap reports
        st, words, pairs, _ = _no_alloc(v2m, t["gt"], t["alt_offsets"], t["edges_by_alt"], a["h_pad"], a["e_pad"])
        assert st == 0 and pairs is None and np.array_equal(words, ref_words)
    assert overlaps_seen > 0


def test_no_alloc_ingest_limits_and_errors(v2m, tmp_path):
    L = v2m.lib()
    # beyond the static scratch: refused before anything is launched
    st = L.v2m_cuda_ingest_genotypes_no_alloc(None, 0, MAX_COPIES + 1, None, None, None, None, 0, MAX_COPIES + 64, 64, C.c_void_p(16), None, 0, None, None)
    assert st != 0 and b"V2M_NO_ALLOC_MAX_COPIES" in L.v2m_cuda_last_error()
    # a genotype naming an ALT its record does not have (variant_graph.cc:399), then a good call: the flag does not stick
    bad = np.array([[0, 2, 0, 1]], np.uint16)
    st, _, _, _ = _no_alloc(v2m, bad, [0, 1], [0], 64, 64)
    assert st != 0 and b"ALT" in L.v2m_cuda_last_error()
    good = np.array([[0, 1, v2m.GT_NULL, 1]], np.uint16)
    st, words, _, _ = _no_alloc(v2m, good, [0, 1], [0], 64, 64)
    ref_words, _ = v2m.paths_from_genotypes(good, [0, 1], [0], 64, 64)
    assert st == 0 and np.array_equal(words, ref_words) and int(words[0]) == 0b1010
    # overlap capacity too small: needed count reported, first pairs valid, a retry with the count succeeds
    a, t = _case(tmp_path, 4103, ref_len=800, n_records=120, n_samples=40, ploidy=2, overlap=True, sv=False, alt_freq=0.3)
    _, ref_ov = v2m.paths_from_genotypes(t["gt"], t["alt_offsets"], t["edges_by_alt"], a["h_pad"], a["e_pad"], t["ref_pos"], t["edge_target_pos"])
    assert len(ref_ov) > 1
    st, _, pairs, n = _no_alloc(v2m, t["gt"], t["alt_offsets"], t["edges_by_alt"], a["h_pad"], a["e_pad"], t["ref_pos"], t["edge_target_pos"], overlap_cap=1)
    assert st != 0 and n == len(ref_ov) and set(pairs) <= set(map(tuple, ref_ov.tolist()))
    st, _, pairs, n = _no_alloc(v2m, t["gt"], t["alt_offsets"], t["edges_by_alt"], a["h_pad"], a["e_pad"], t["ref_pos"], t["edge_target_pos"], overlap_cap=n)
    assert st == 0 and pairs == sorted(map(tuple, ref_ov.tolist()))
