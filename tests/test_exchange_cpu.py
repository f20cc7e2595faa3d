"""CPU test of the peer-memory exchange plan of the h-sharded transforms (makani_b200/distributed: _PolarExchange).

The row descriptors that ``sfno_rows_scatter`` executes on the GPU (include/makani_b200.h: sfno_copy_desc) are pure
index arithmetic over the split shapes, so they are checked here without GPUs: "peer memory" is a dict of numpy arrays,
the scatter kernel is a ten-line numpy loop following the header's definition, and the h ranks run one after the other
in this process.  Every direction is compared with the transposition it must realise on a global tensor, for even and
uneven latitude / channel / degree splits (the reference's compute_split_shapes, e.g. 721/4 -> 181,181,181,178)."""
import types

import numpy as np
import pytest
import torch

import makani_b200.distributed as thd
from makani_b200.distributed import _PolarExchange, _rup4, _rup8, compute_split_shapes, p2p


class FakeBuffers:
    """Two alternating receive buffers per rank; a 'device pointer' is the key of the numpy array."""
    mem = {}

    def __init__(self, key, nfloats, h, rank):
        self.local, self.peer_ptrs, self._use = [], [], 0
        for b in range(2):
            ptrs = []
            for r in range(h):
                k = hash((key, nfloats, r, b)) & 0x7FFFFFFFFFFF
                if k not in FakeBuffers.mem:
                    FakeBuffers.mem[k] = np.zeros(nfloats, dtype=np.float32)
                ptrs.append(k)
            self.peer_ptrs.append(ptrs)
            self.local.append(torch.from_numpy(FakeBuffers.mem[ptrs[rank]]))

    def next(self):
        b = self._use & 1
        self._use += 1
        return b


def fake_scatter(src, arr, boxed=False):
    """sfno_rows_scatter as the header defines it."""
    s = src.numpy().reshape(-1)
    for d in arr:
        dst = FakeBuffers.mem[d.dst]
        for i0 in range(d.n0):
            for i1 in range(d.n1):
                if d.tri_l0 >= 0 and i1 >= ((d.tri_l0 + i0 + 16) & ~15):
                    continue
                for i2 in range(d.n2):
                    so = d.src_off + i0 * d.ss0 + i1 * d.ss1 + i2 * d.ss2
                    do = d.dst_off + i0 * d.ds0 + i1 * d.ds1 + i2 * d.ds2
                    dst[do:do + d.len] = s[so:so + d.len]


@pytest.fixture
def fake_peer_memory(monkeypatch):
    FakeBuffers.mem = {}
    state = {"rank": 0, "h": 1}
    monkeypatch.setattr(p2p, "buffers", lambda key, n, group, device: FakeBuffers(key, n, state["h"], state["rank"]))
    monkeypatch.setattr(p2p, "scatter", fake_scatter)
    class NoFence:                       # one process plays the ranks in turn: nothing to order
        def __call__(self):
            pass

        def pre_write(self):
            pass

    monkeypatch.setattr(p2p, "fence", lambda group, device: NoFence())
    monkeypatch.setattr(thd, "polar_group", lambda: None)
    return state


def make_mod(rank, h, nlat, L, M):
    return types.SimpleNamespace(comm_rank_polar=rank, comm_size_polar=h, mmax=M, lmax=L, nlat=nlat,
                                 lat_shapes=compute_split_shapes(nlat, h), l_shapes=compute_split_shapes(L, h))


CASES = [(2, 37, 20, 21, 1, 6), (3, 23, 17, 9, 2, 7), (4, 41, 33, 34, 1, 10), (4, 16, 16, 17, 1, 8)]


@pytest.mark.parametrize("h,nlat,L,M,B,C", CASES)
def test_exchange_descriptors_realise_the_transpositions(fake_peer_memory, h, nlat, L, M, B, C):
    check_exchange(fake_peer_memory, h, nlat, L, M, B, C, torch.device("cpu"))


def check_exchange(state, h, nlat, L, M, B, C, device):
    """Run the four exchanges for h ranks (one after the other in this process; `state` selects the acting rank for the
    patched p2p.buffers) and compare with the transpositions of a global tensor.  Shared with the single-GPU kernel
    test (tests/test_gpu_kernels.py), where the buffers are device memory and the scatter is the CUDA kernel."""
    rng = np.random.default_rng(0)
    J = 2 * B
    lat, cs, ls = compute_split_shapes(nlat, h), compute_split_shapes(C, h), compute_split_shapes(L, h)
    lat0, c0, l0 = np.cumsum([0] + lat), np.cumsum([0] + cs), np.cumsum([0] + ls)
    latp0 = np.cumsum([0] + [_rup8(k) for k in lat])
    Kpg = int(latp0[-1])
    Cp = _rup4(C)
    state["h"] = h
    xm = rng.standard_normal((M, J, C, nlat)).astype(np.float32)          # global XM[m][j][c][k]
    sp = rng.standard_normal((L, M, J, C)).astype(np.float32)             # global SP[l][m][j][c]
    exs = []
    for r in range(h):
        state["rank"] = r
        exs.append(_PolarExchange(make_mod(r, h, nlat, L, M), B, C, device))

    def run_all(method, make_src):
        outs = []
        for r in range(h):                    # all ranks scatter ...
            state["rank"] = r
            outs.append(getattr(exs[r], method)(torch.from_numpy(np.ascontiguousarray(make_src(r))).to(device)))
        return [o.cpu().numpy().copy() for o in outs]   # ... then everybody reads its receive buffer

    # ---- lat -> chan: rank r holds XM_loc [M][J][Cp][Kl], ends with XM_ch [M][J][Cp_r][Kpg] (padded latitude order)
    def xm_loc(r):
        Kl = _rup4(lat[r])
        a = np.zeros((M, J, Cp, Kl), dtype=np.float32)
        a[:, :, :C, :lat[r]] = xm[..., lat0[r]:lat0[r + 1]]
        return a.reshape(-1)

    got = run_all("lat_to_chan", xm_loc)
    xm_ch = []
    for r in range(h):
        a = got[r].reshape(M, J, _rup4(cs[r]), Kpg)
        for p in range(h):
            np.testing.assert_array_equal(a[:, :, :cs[r], latp0[p]:latp0[p] + lat[p]],
                                          xm[:, :, c0[r]:c0[r + 1], lat0[p]:lat0[p + 1]])
            assert not a[:, :, :, latp0[p] + lat[p]:latp0[p + 1]].any()     # pad latitudes stay zero
        assert not a[:, :, cs[r]:, :].any()                                  # pad channels stay zero
        xm_ch.append(a.reshape(-1))

    # ---- chan -> lat is its inverse
    got = run_all("chan_to_lat", lambda r: xm_ch[r])
    for r in range(h):
        a = got[r].reshape(M, J, Cp, _rup4(lat[r]))
        np.testing.assert_array_equal(a[:, :, :C, :lat[r]], xm[..., lat0[r]:lat0[r + 1]])

    # ---- chan -> l: rank r holds SP_ch [L][M][J][Cp_r], ends with SP_l [L_r][M][J][Cp]; only m < round_up(l+1,16) moves
    def sp_ch(r):
        a = np.zeros((L, M, J, _rup4(cs[r])), dtype=np.float32)
        a[..., :cs[r]] = sp[..., c0[r]:c0[r + 1]]
        return a.reshape(-1)

    defined = np.array([[m < ((l + 16) & ~15) for m in range(M)] for l in range(L)])
    got = run_all("chan_to_l", sp_ch)
    sp_l = []
    for r in range(h):
        a = got[r].reshape(-1)[: ls[r] * M * J * Cp].reshape(ls[r], M, J, Cp)
        d = defined[l0[r]:l0[r + 1]]
        np.testing.assert_array_equal(a[d][..., :C], sp[l0[r]:l0[r + 1]][d])
        assert not a[~d].any()                                               # undefined entries are never written
        sp_l.append(a.reshape(-1))

    # ---- l -> chan is its inverse on the defined entries
    got = run_all("l_to_chan", lambda r: sp_l[r])
    for r in range(h):
        a = got[r].reshape(L, M, J, _rup4(cs[r]))
        np.testing.assert_array_equal(a[defined][..., :cs[r]], sp[defined][..., c0[r]:c0[r + 1]])


def test_copy_desc_matches_the_c_struct():
    """ctypes mirror of sfno_copy_desc: 96 bytes, field order of include/makani_b200.h."""
    import ctypes
    from makani_b200 import _ext
    assert ctypes.sizeof(_ext.CopyDesc) == 96
    assert [f[0] for f in _ext.CopyDesc._fields_] == ["dst", "src_off", "dst_off", "n0", "n1", "n2", "len",
                                                       "ss0", "ss1", "ss2", "ds0", "ds1", "ds2", "tri_l0"]
