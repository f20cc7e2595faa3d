"""CPU tests of makani_b200.comm (the makani.utils.comm API over torch.distributed):
group-layout arithmetic, single-process defaults (makani/utils/comm.py:30-94) and a world-size-4 gloo run that
uses the groups for what the spectral path uses them for (thd.init + distributed transposes, sfnonet.py:528-531)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_plan_groups_layout():
    from makani_b200 import comm
    g = comm.plan_groups(8, dict(h=2, w=2, fin=1, fout=1, data=2))
    assert g["world"] == [list(range(8))]
    assert g["model"] == [[0, 1, 2, 3], [4, 5, 6, 7]]
    assert g["data"] == [[0, 4], [1, 5], [2, 6], [3, 7]]
    assert g["spatial"] == [[0, 1, 2, 3], [4, 5, 6, 7]]
    assert g["h"] == [[0, 1], [2, 3], [4, 5], [6, 7]]               # h varies fastest: neighbouring ranks
    assert g["w"] == [[0, 2], [1, 3], [4, 6], [5, 7]]
    assert g["matmul"] == [[r] for r in range(8)] and g["fin"] == g["fout"] == g["matmul"]
    # every rank is in exactly one group of every name, and group sizes multiply to the world size
    for ws, sizes in [(8, dict(h=4, w=1, fin=1, fout=1, data=2)), (16, dict(h=2, w=2, fin=2, fout=1, data=2)),
                      (12, dict(h=3, w=1, fin=1, fout=2, data=2))]:
        g = comm.plan_groups(ws, sizes)
        for name, groups in g.items():
            flat = sorted(r for grp in groups for r in grp)
            assert flat == list(range(ws)), name
        for r in range(ws):
            coords = {n: next(grp.index(r) for grp in g[n] if r in grp) for n in ("h", "w", "fin", "fout", "data")}
            lin = (((coords["data"] * sizes["fout"] + coords["fout"]) * sizes["fin"] + coords["fin"]) * sizes["w"]
                   + coords["w"]) * sizes["h"] + coords["h"]
            assert lin == r
    with pytest.raises(ValueError):
        comm.plan_groups(6, dict(h=4, w=1, fin=1, fout=1, data=2))


def test_single_process_defaults():
    from makani_b200 import comm
    comm.cleanup()
    assert comm.get_size("h") == 1 and comm.get_rank("w") == 0 and comm.get_group("h") is None
    assert comm.get_world_size() == 1 and comm.get_world_rank() == 0 and comm.get_local_rank() == 0
    assert comm.get_names() == [] and not comm.is_distributed("spatial") and comm.get_root("model") == 0
    env = {k: os.environ.pop(k) for k in ("WORLD_SIZE", "RANK") if k in os.environ}
    try:
        assert comm.init([1, 1, 1, 1], ["h", "w", "fin", "fout"]) == 1
        assert comm.get_size("spatial") == 1 and comm.get_rank("data") == 0 and comm.is_distributed("h")
        assert set(comm.get_names()) == {"world", "model", "spatial", "matmul", "h", "w", "fin", "fout", "data"}
        with pytest.raises(ValueError):
            comm.init([2, 1, 1, 1])
    finally:
        os.environ.update(env)
        comm.cleanup()


def _worker(rank, world, port, h, w):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    from makani_b200 import comm
    import makani_b200.distributed as thd
    try:
        msize = comm.init(model_parallel_sizes=[h, w, 1, 1], model_parallel_names=["h", "w", "fin", "fout"])
        assert dist.is_initialized() and dist.get_backend() == "gloo"
        assert msize == h * w == comm.get_size("model") == comm.get_size("spatial")
        assert comm.get_size("data") == world // (h * w) and comm.get_size("matmul") == 1
        assert comm.get_world_size() == world and comm.get_world_rank() == rank
        rh, rw, rd = rank % h, (rank // h) % w, rank // (h * w)
        assert (comm.get_rank("h"), comm.get_rank("w"), comm.get_rank("data")) == (rh, rw, rd)
        assert comm.get_root("h") == rank - rh and comm.get_root("data") == rank % (h * w)
        # the groups really connect the ranks the plan says: sum of world ranks over each group
        for name in ("h", "w", "data", "model", "world"):
            t = torch.tensor([float(rank)])
            dist.all_reduce(t, group=comm.get_group(name))
            mine = next(g for g in comm.group_ranks(name) if rank in g)
            assert t.item() == float(sum(mine)), (name, t.item(), mine)

        # what sfnonet.py:528-531 does with them
        polar = None if comm.get_size("h") == 1 else comm.get_group("h")
        azimuth = None if comm.get_size("w") == 1 else comm.get_group("w")
        thd.init(polar, azimuth)
        assert thd.polar_group_size() == h and thd.azimuth_group_size() == w
        assert thd.polar_group_rank() == rh and thd.azimuth_group_rank() == rw
        if h > 1:
            # channel <-> latitude transpose inside the h group, uneven latitude split; every data replica on its own
            torch.manual_seed(5 + rd)
            C, K, M = 2 * h, 7, 3
            full = torch.randn(1, C, K, M, dtype=torch.complex64)
            lat = thd.compute_split_shapes(K, h)
            ch = thd.compute_split_shapes(C, h)
            x = torch.split(full, lat, dim=2)[rh].contiguous()
            y = thd.distributed_transpose_polar.apply(x, (1, 2), lat)
            ref = torch.split(full, ch, dim=1)[rh]
            assert torch.equal(y, ref)
        dist.barrier()
    finally:
        comm.cleanup()
        if dist.is_initialized():
            dist.destroy_process_group()


@pytest.mark.parametrize("h,w", [(2, 1), (2, 2)])
def test_comm_init_world4_gloo(h, w):
    port = _free_port()
    mp.spawn(_worker, args=(4, port, h, w), nprocs=4, join=True)
