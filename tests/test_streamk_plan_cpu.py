"""The stream-K decomposition of the tensor-core GEMM (csrc/tc_gemm_streamk.inl), checked on the host: the kernel's own
segment function, compiled for the host too, must hand every (tile, K-block) unit to exactly one CTA, in ranges that
are contiguous and balanced to within one unit.  (What needs a GPU — the pipeline — is tests/test_tc_streamk_gpu.py.)"""
import ctypes as C

import pytest

from torch_distlearn_b200 import _lib

BM, BK = 128, 32


def _plan(M, N, K, sms):
    lib = _lib.cuda()
    grid = C.c_int()
    buf = (C.c_int * (4 * 4096))()
    n0 = lib.dl_tc_streamk_plan(M, N, K, sms, 0, C.byref(grid), buf, 4096)
    assert n0 >= 0, lib.dl_last_error()
    ctas = []
    for cta in range(grid.value):
        n = lib.dl_tc_streamk_plan(M, N, K, sms, cta, C.byref(grid), buf, 4096)
        assert 0 <= n <= 4096
        ctas.append([tuple(buf[4 * i:4 * i + 4]) for i in range(n)])
    return grid.value, ctas


@pytest.mark.parametrize("M,N,K", [(32768, 128, 1600), (8192, 256, 3200), (2048, 512, 6400), (131072, 64, 96), (128, 128, 32),
                                   (200, 72, 76), (512, 6400, 2048), (128, 10, 2048), (8192, 8192, 8192), (1000, 300, 4100)])
@pytest.mark.parametrize("sms", [148, 7])
def test_every_unit_is_done_exactly_once(M, N, K, sms):
    bn = 64 if N <= 64 else 128
    m_tiles, n_tiles, kbt = -(-M // BM), -(-N // bn), -(-K // BK)
    units = m_tiles * n_tiles * kbt
    grid, ctas = _plan(M, N, K, sms)
    assert 1 <= grid <= sms and (grid == sms or grid == max(units // 4, 1))
    seen = {}
    sizes = []
    for cta, segs in enumerate(ctas):
        total = 0
        prev_end = None
        for m0, n0, kb0, nkb in segs:
            assert m0 % BM == 0 and n0 % bn == 0 and 0 <= m0 < m_tiles * BM and 0 <= n0 < n_tiles * bn
            assert nkb >= 1 and 0 <= kb0 and kb0 + nkb <= kbt
            tile = (m0 // BM) * n_tiles + n0 // bn
            start = tile * kbt + kb0
            if prev_end is not None:
                assert start == prev_end and kb0 == 0          # contiguous: a later segment starts a new tile
            prev_end = start + nkb
            for kb in range(kb0, kb0 + nkb):
                assert (tile, kb) not in seen
                seen[(tile, kb)] = cta
            total += nkb
        sizes.append(total)
    assert len(seen) == units
    assert max(sizes) - min(sizes) <= 1                        # balanced to one K block
    # a tile is either whole in one CTA (plain store) or split between CTAs (all red.add): never mixed
    for segs in ctas:
        for m0, n0, kb0, nkb in segs:
            tile = (m0 // BM) * n_tiles + n0 // bn
            owners = {seen[(tile, kb)] for kb in range(kbt)}
            assert (nkb == kbt) == (len(owners) == 1)


def test_layer2_forward_costs_what_the_design_note_says():
    # 256 tiles x 50 K blocks on 148 SMs: 86 or 87 K-block times per CTA instead of two waves of 50
    grid, ctas = _plan(32768, 128, 1600, 148)
    assert grid == 148
    assert {sum(s[3] for s in segs) for segs in ctas} == {86, 87}
