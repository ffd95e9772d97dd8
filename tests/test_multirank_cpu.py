"""CPU, world_size 2 over gloo: the rank -> brick logic bench.py uses at N>1 (process grid, local
bounds, lattice ownership) partitions the global lattice exactly, and agrees with the C host layer's
initDecomposition arithmetic (decomposition.c:18-50)."""
import os
import socket

import numpy as np
import pytest

import pycomd
from pycomd import potentials as P


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    return port


def _worker(rank, world, port, grid, ncells, out):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pot = P.load("lj")
    lat = pot["lat"]
    gext = [ncells[a] * lat for a in range(3)]
    coord = (rank % grid[0], (rank // grid[0]) % grid[1], rank // (grid[0] * grid[1]))
    geom = pycomd.make_geometry(gext, pot["cutoff"], grid, coord)
    lmin = [geom.localMin[a] for a in range(3)]; lmax = [geom.localMax[a] for a in range(3)]
    gid, r = P.fcc_block(ncells, lat, lmin, lmax)
    n = torch.tensor([len(gid)], dtype=torch.int64)
    dist.all_reduce(n)                                   # addIntParallel (parallel.c:120)
    sizes = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(sizes, torch.tensor([len(gid)], dtype=torch.int64))
    pad = int(max(s.item() for s in sizes))
    mine = torch.full((pad,), -1, dtype=torch.int64); mine[:len(gid)] = torch.from_numpy(gid.astype(np.int64))
    allg = [torch.zeros(pad, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(allg, mine)
    if rank == 0:
        union = np.concatenate([a.numpy()[a.numpy() >= 0] for a in allg])
        np.savez(out, total=int(n.item()), union=union, lmin=np.array(lmin), lmax=np.array(lmax), grid=np.array(geom.gridSize[:]))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("grid,ncells", [((2, 1, 1), (8, 6, 6)), ((1, 2, 1), (6, 10, 6))])
def test_two_ranks_partition_lattice_gloo(tmp_path, grid, ncells):
    import torch.multiprocessing as mp
    out = str(tmp_path / "r0.npz")
    mp.spawn(_worker, args=(2, _free_port(), grid, ncells, out), nprocs=2, join=True)
    d = np.load(out)
    n_global = 4 * ncells[0] * ncells[1] * ncells[2]
    assert int(d["total"]) == n_global
    assert np.array_equal(np.sort(d["union"]), np.arange(n_global))          # every gid on exactly one rank
    # rank 0's brick: [0, L/2) on the split axis, whole box on the others
    ax = grid.index(2)
    L = ncells[ax] * 3.615
    assert d["lmin"][ax] == 0.0 and d["lmax"][ax] == 0.0 + 1 * (L / 2)
    assert int(d["grid"][ax]) == int((L / 2) / (2.5 * 2.315))
